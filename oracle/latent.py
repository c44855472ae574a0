"""
oracle/latent.py -- TEST INFRASTRUCTURE (see oracle/__init__.py).

CPU restatement of the latent-distance adjacency model's HMC target and its gradient
(adjacency.py:209-219, 244-275, 356-421) and of hips-lib's HMC step (not in /root/reference;
version unpinned; restated from its published code -- "recollection").
The reference obtains the gradient from autograd (adjacency.py:375-379); here it is hand-derived
 and pinned by a finite-difference test (tests/test_oracle_latent.py).
"""
import numpy as np


def logit_matrix(L, mu_0, mu_self):
    """adjacency.py:209-214 / 252-258:  -||l_i - l_j||^2 + mu_0 + mu_self * I."""
    N = L.shape[0]
    D = -((L[:, None, :] - L[None, :, :]) ** 2).sum(2)
    return D + mu_0 + mu_self * np.eye(N)


def hmc_log_probability(L, mu_0, mu_self, A, sigma):
    """adjacency.py:244-275, literally (P = 1/(1+exp(-eta)); ll = sum A log P + (1-A) log(1-P))."""
    eta = logit_matrix(L, mu_0, mu_self)
    P = 1.0 / (1 + np.exp(-eta))
    ll = np.sum(A * np.log(P) + (1 - A) * np.log(1 - P))
    lp = -0.5 * np.sum(L * L / sigma)
    lp += -0.5 * mu_0 ** 2
    lp += -0.5 * mu_self ** 2
    return ll + lp


def log_probability_stable(L, mu_0, mu_self, A, sigma):
    """Same value through softplus: A log s(eta) + (1-A) log(1-s(eta)) = A eta - softplus(eta)."""
    eta = logit_matrix(L, mu_0, mu_self)
    ll = np.sum(A * eta - np.logaddexp(0.0, eta))
    return ll - 0.5 * np.sum(L * L) / sigma - 0.5 * mu_0 ** 2 - 0.5 * mu_self ** 2


def log_probability_and_grad(L, mu_0, mu_self, A, sigma):
    """Value and hand-derived gradients (what autograd returns at adjacency.py:375-379,398,415).

    g_ij = A_ij - s(eta_ij);  d/dl_n = -2 sum_j (l_n - l_j)(g_nj + g_jn) - l_n/sigma;
    d/dmu_0 = sum g - mu_0;  d/dmu_self = sum_i g_ii - mu_self.
    """
    eta = logit_matrix(L, mu_0, mu_self)
    sig = 1.0 / (1.0 + np.exp(-eta))
    g = A - sig
    gs = g + g.T
    lp = np.sum(A * eta - np.logaddexp(0.0, eta)) - 0.5 * np.sum(L * L) / sigma - 0.5 * mu_0 ** 2 - 0.5 * mu_self ** 2
    gL = -2.0 * (gs.sum(1)[:, None] * L - gs.dot(L)) - L / sigma
    gmu0 = g.sum() - mu_0
    gmuself = np.trace(g) - mu_self
    return lp, gL, gmu0, gmuself


def hmc(U, grad_U, step_sz, n_steps, q_curr, draws, negative_log_prob=True,
        avg_accept_rate=0.9, adaptive_step_sz=False, tgt_accept_rate=0.9,
        avg_accept_time_const=0.95, min_step_sz=0.00005, max_step_sz=1.0):
    """hips.inference.hmc.hmc (recollection), Neal's leapfrog HMC.

    ``draws`` = (p0, log_u): injected momentum (standard normal, shape of q) and log-uniform for the
    accept test.  With negative_log_prob=False, U is a log-probability (the reference's call,
    adjacency.py:382-385) and is negated internally.
    Returns q or (q, new_step_sz, new_accept_rate) when adaptive_step_sz.
    """
    p0, log_u = draws
    if negative_log_prob:
        Uf, dUf = U, grad_U
    else:
        Uf, dUf = (lambda q: -1.0 * U(q)), (lambda q: -1.0 * grad_U(q))
    q = np.array(q_curr, dtype=np.float64, copy=True)
    p = np.array(p0, dtype=np.float64, copy=True)
    p_curr = p.copy()
    p = p - step_sz * dUf(q) / 2.0
    for i in range(n_steps):
        q = q + step_sz * p
        if i < n_steps - 1:
            p = p - step_sz * dUf(q)
    p = p - step_sz * dUf(q) / 2.0
    p = -p
    U_curr = Uf(q_curr); K_curr = np.sum(p_curr ** 2) / 2.0
    U_prop = Uf(q); K_prop = np.sum(p ** 2) / 2.0
    accept = log_u < (U_curr - U_prop + K_curr - K_prop)
    q_next = q if accept else np.array(q_curr, dtype=np.float64, copy=True)
    if adaptive_step_sz:
        new_rate = avg_accept_time_const * avg_accept_rate + (1.0 - avg_accept_time_const) * float(accept)
        new_step = step_sz * 1.02 if new_rate > tgt_accept_rate else step_sz * 0.98
        new_step = min(max(new_step, min_step_sz), max_step_sz)
        return q_next, new_step, new_rate
    return q_next


def resample(L, mu_0, mu_self, A, sigma, step_sz, accept_rate, draws):
    """adjacency.py:356-421: HMC on L (adaptive, 10 steps), then mu_0 and mu_self (step 0.005, 10 steps).
    draws = dict(L=(p0[N,D], log_u), mu_0=(p0, log_u), mu_self=(p0, log_u))."""
    lpL = lambda Lq: log_probability_and_grad(Lq, mu_0, mu_self, A, sigma)[0]
    dlpL = lambda Lq: log_probability_and_grad(Lq, mu_0, mu_self, A, sigma)[1]
    L, step_sz, accept_rate = hmc(lpL, dlpL, step_sz, 10, L.copy(), draws["L"], negative_log_prob=False,
                                  avg_accept_rate=accept_rate, adaptive_step_sz=True)
    lp0 = lambda m: log_probability_and_grad(L, float(m), mu_self, A, sigma)[0]
    dlp0 = lambda m: np.array(log_probability_and_grad(L, float(m), mu_self, A, sigma)[2])
    mu_0 = float(hmc(lp0, dlp0, 0.005, 10, np.array(mu_0), draws["mu_0"], negative_log_prob=False))
    lps = lambda m: log_probability_and_grad(L, mu_0, float(m), A, sigma)[0]
    dlps = lambda m: np.array(log_probability_and_grad(L, mu_0, float(m), A, sigma)[3])
    mu_self = float(hmc(lps, dlps, 0.005, 10, np.array(mu_self), draws["mu_self"], negative_log_prob=False))
    return L, mu_0, mu_self, step_sz, accept_rate


def rows_value_and_grad(L, mu_0, mu_self, A_rows, At_rows, r0, sigma):
    """The same value / gradient restricted to the row block [r0, r0 + R): the terms of
    ``log_probability_and_grad`` that involve rows r0.. of A (``A_rows`` = A[r0:r0+R, :]) and, for the gradient of
    those nodes, the matching columns (``At_rows`` = A[:, r0:r0+R].T).  Summing ``ll`` / ``gsum`` over a partition of
    the rows and stacking ``gL`` reproduces the full evaluation (tests/test_oracle_latent.py); the bench's CPU
    baseline times this on a bounded block of rows.  Vectorised NumPy, fp64, O(R N dim) memory."""
    R = A_rows.shape[0]
    Lr = L[r0:r0 + R]
    diff = Lr[:, None, :] - L[None, :, :]
    eta = -(diff ** 2).sum(2) + mu_0
    eta[np.arange(R), r0 + np.arange(R)] += mu_self
    sig = 1.0 / (1.0 + np.exp(-eta))
    ll = np.sum(A_rows * eta - np.logaddexp(0.0, eta))
    g = A_rows - sig
    gs = g + (At_rows - sig)                 # g_nj + g_jn (eta is symmetric)
    gL = -2.0 * np.einsum("rj,rjd->rd", gs, diff) - Lr / sigma
    return ll, gL, g.sum(), g[np.arange(R), r0 + np.arange(R)].sum()


# ------------------------------------------------------------------------------------------------------------------
# LatentDistanceGaussianWeightDistribution (weights.py:613-901, B = 1)
# ------------------------------------------------------------------------------------------------------------------
def weight_log_probability_and_grad(L, b, A, W, sigma_w, eta):
    """weights.py:697-731 (_hmc_log_probability) and the gradient autograd derives from it.

    Mu = -||l_i - l_j||^2 + b;  X = (W - Mu) A_offdiag;  lp = -1/2 sum X^2 / sigma_w - sum L^2 / (2 eta) - b^2 / 2.
    With r = W + d^2 - b on the off-diagonal edges:  d lp / d l_n = -2/sigma_w sum_j (A_nj r_nj + A_jn r_jn)(l_n - l_j) - l_n / eta,
    d lp / d b = sum A r / sigma_w - b.   W: N x N (the single weight dimension)."""
    N = L.shape[0]
    D2 = ((L[:, None, :] - L[None, :, :]) ** 2).sum(2)
    Aoff = np.asarray(A, dtype=np.float64) * (1.0 - np.eye(N))
    R = (W + D2 - b) * Aoff
    lam = 1.0 / sigma_w
    lp = np.sum(-0.5 * R ** 2 * lam) - 0.5 * np.sum(L * L / eta) - 0.5 * b ** 2
    Rs = R + R.T
    gL = -2.0 * lam * (Rs.sum(1)[:, None] * L - Rs.dot(L)) - L / eta
    gb = lam * R.sum() - b
    return lp, gL, gb


def weight_stats(L, b, A, W):
    """(n, sum r, sum r^2) over the off-diagonal edges: what weights.py:838-867 reduce W to."""
    N = L.shape[0]
    D2 = ((L[:, None, :] - L[None, :, :]) ** 2).sum(2)
    e = (np.asarray(A) != 0) & ~np.eye(N, dtype=bool)
    r = (W + D2 - b)[e]
    return float(r.size), float(r.sum()), float((r * r).sum())


def weight_resample_L(L, b, A, W, sigma_w, eta, draws, step=0.005, nsteps=10):
    """weights.py:765-780: one hips HMC transition on L, fixed step 0.005, 10 leapfrog steps."""
    lp = lambda Lq: weight_log_probability_and_grad(Lq, b, A, W, sigma_w, eta)[0]
    dlp = lambda Lq: weight_log_probability_and_grad(Lq, b, A, W, sigma_w, eta)[1]
    return hmc(lp, dlp, step, nsteps, L.copy(), draws, negative_log_prob=False)


def weight_b_posterior(n, sum_x, sigma_w):
    """weights.py:838-860: X = W + D ~ N(b, sigma_w) on the off-diagonal edges, b ~ N(0, 1)  ->  (mean, variance)."""
    sigma_post = 1.0 / (1.0 + n / sigma_w)
    return sigma_post * (sum_x / sigma_w), sigma_post
