"""
CPU oracle (TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
leg, never by the product path) for the mean-field SBM updates, SURVEY.md row a18.

Restates, in Python 3 / fp64:
  * graphistician/stochastic_block_model.py:358-364  mf_update_m
  * graphistician/stochastic_block_model.py:366-389  mf_update_p
  * graphistician/stochastic_block_model.py:392-445  mf_update_c   (`mf_update_c_loops` is the literal loop nest)
  * graphistician/stochastic_block_model.py:340-344 + internals/weights.py:105-132,158-171
                                                     weighted NIW natural-parameter update
  * graphistician/internals/weights.py:200-218       mf_expected_* of the weight model
  * graphistician/internals/distributions.py:75-103 (Bernoulli.negentropy), :278-308 (Gaussian.negentropy),
    :358-364 (Beta.expected_log_p / expected_log_notp)

Pinning: the three internals/distributions.py functions import unmodified under Python 3; tests/golden/
make_golden.py runs `mf_update_c_loops` WITH THE REFERENCE'S OWN Beta / Bernoulli / Gaussian classes plugged in
and commits the result (tests/golden/mf_sbm_small.npz) -- so the per-edge arithmetic is pinned by the reference.

Reference quirks kept:
  * `_expected_log_likelihood_W` (stochastic_block_model.py:631-644) calls `expected_log_likelihood` with an
    (E_W, E_WWT) tuple, which the GaussianWeights override (internals/weights.py:174-198, takes a *network*) cannot
    accept -- the reference would raise.  The per-pair quantity that override sums is
    `Gaussian().negentropy(E_x, E_xxT, E_mu, E_mumuT, E_Sigma_inv, E_logdet_Sigma)`; that is what is used here.
  * Gaussian.negentropy's constant is -0.5*log(2*pi), not -D/2*log(2*pi) (distributions.py:303).
  * mf_expected_mumuT ignores the 1/kappa term (internals/weights.py:207-211).
  * the forward (n1 -> n2) term includes n2 == n1, the reverse term does not (stochastic_block_model.py:410-435).
  * the reverse-edge branch `if n2 != n1:` (:425) is indented one level too shallow: it runs once per n2 after the
    `for cn2` loop, so only cn2 == C-1 (weight mf_m[n2, C-1]) contributes to the reverse term.  `literal_reverse=True`
    (default, what the reference computes) keeps that; `literal_reverse=False` is the evidently intended sum over cn2.
pybasicbayes' natural-parameter NIW (`_standard_to_natural`, `_natural_to_standard`, `_loglmbdatilde`) is a
recollection of public code that is not in the tree.
"""
import numpy as np
from scipy.special import psi, logsumexp


# ---- internals/distributions.py restated ------------------------------------------------------------
def beta_expected_logs(tau1, tau0):
    """Beta(tau1, tau0).expected_log_p(), .expected_log_notp()  (distributions.py:324-325, 358-364)."""
    tau1, tau0 = np.asarray(tau1, float), np.asarray(tau0, float)
    tot = psi(tau1 + tau0)
    return psi(tau1) - tot, psi(tau0) - tot


def bernoulli_negentropy(E_x, E_notx, E_ln_p, E_ln_notp):
    """distributions.py:75-103 (float branch)."""
    return E_x * E_ln_p + E_notx * E_ln_notp


def gaussian_negentropy(E_x, E_xxT, E_mu, E_mumuT, E_Sigma_inv, E_logdet_Sigma):
    """distributions.py:278-308."""
    H = -0.5 * np.log(2 * np.pi)
    H += -0.5 * E_logdet_Sigma
    H += -0.5 * np.trace(E_Sigma_inv.dot(E_xxT))
    H += E_x.T.dot(E_Sigma_inv).dot(E_mu)
    H += -0.5 * np.trace(E_Sigma_inv.dot(E_mumuT))
    return H


# ---- pybasicbayes natural-parameter NIW (recollection) --------------------------------------------------
def standard_to_natural(mu, sigma, kappa, nu):
    D = sigma.shape[0]
    out = np.zeros((D + 2, D + 2))
    out[:D, :D] = sigma + kappa * np.outer(mu, mu)
    out[:D, -2] = out[-2, :D] = kappa * mu
    out[-2, -2] = kappa
    out[-1, -1] = nu + 2 + D
    return out


def natural_to_standard(nat):
    D = nat.shape[0] - 2
    A, b, c, d = nat[:D, :D], nat[:D, -2], nat[-2, -2], nat[-1, -1]
    return b / c, A - np.outer(b, b) / c, c, d - 2 - D


def weighted_statistics(E_x, E_xxT, weights):
    """internals/weights.py:158-171 (`_get_weighted_statistics`); E_x [M,D], E_xxT [M,D,D], weights [M]."""
    D = E_x.shape[1]
    out = np.zeros((D + 2, D + 2))
    out[:D, :D] = (weights[:, None, None] * E_xxT).sum(0)
    out[-2, :D] = out[:D, -2] = (weights[:, None] * E_x).sum(0)
    out[-2, -2] = out[-1, -1] = weights.sum()
    return out


def statistics_matrix(n, sum_x, sum_xxT):
    D = len(sum_x)
    out = np.zeros((D + 2, D + 2))
    out[:D, :D] = sum_xxT
    out[-2, :D] = out[:D, -2] = sum_x
    out[-2, -2] = out[-1, -1] = n
    return out


def loglmbdatilde(sigma_mf, nu_mf):
    """pybasicbayes Gaussian._loglmbdatilde: E_q[log |Sigma^-1|]."""
    D = sigma_mf.shape[0]
    chol = np.linalg.cholesky(sigma_mf)
    return -2 * np.log(np.diag(chol)).sum() + D * np.log(2) + psi((nu_mf - np.arange(D)) / 2.0).sum()


def weight_expectations(mu_mf, sigma_mf, kappa_mf, nu_mf):
    """internals/weights.py:200-218: E_mu, E_mumuT, E_Sigma_inv, E_logdet_Sigma of one cell's q(mu, Sigma)."""
    D = len(mu_mf)
    E_Sigma = sigma_mf / (nu_mf - D - 1)
    return (mu_mf, E_Sigma + np.outer(mu_mf, mu_mf), nu_mf * np.linalg.inv(sigma_mf),
            -loglmbdatilde(sigma_mf, nu_mf))


# ---- the updates ------------------------------------------------------------------------------------
def mf_update_m(mf_pi, pi, mf_m, stepsize=1.0):
    """stochastic_block_model.py:358-364."""
    return (1.0 - stepsize) * mf_pi + stepsize * (pi + mf_m.sum(axis=0))


def mf_update_p_loops(E_A, mf_m, tau1, tau0, mf_tau1, mf_tau0, stepsize=1.0):
    """stochastic_block_model.py:366-389 (both branches of allow_self_connections are the same code)."""
    C = mf_m.shape[1]
    E_notA = 1 - E_A
    t1, t0 = mf_tau1.copy(), mf_tau0.copy()
    for c1 in range(C):
        for c2 in range(C):
            pc1c2 = mf_m[:, c1][:, None] * mf_m[:, c2][None, :]
            tau1_hat = tau1 + (pc1c2 * E_A).sum()
            tau0_hat = tau0 + (pc1c2 * E_notA).sum()
            t1[c1, c2] = (1.0 - stepsize) * t1[c1, c2] + stepsize * tau1_hat
            t0[c1, c2] = (1.0 - stepsize) * t0[c1, c2] + stepsize * tau0_hat
    return t1, t0


def soft_block_stats(E_A, E_W, mf_m):
    """mf_m^T (E_A o x) mf_m for x = 1, W_b, W_b W_b' over ALL pairs (diagonal included): the quantities
    stochastic_block_model.py:340-344,366-389 and internals/weights.py:158-171 reduce to when E_WWT = W W^T."""
    n = mf_m.T.dot(E_A).dot(mf_m)
    sw = np.einsum("ik,ijb,jl->klb", mf_m, E_A[:, :, None] * E_W, mf_m)
    sww = np.einsum("ik,ijbc,jl->klbc", mf_m, E_A[:, :, None, None] * E_W[:, :, :, None] * E_W[:, :, None, :], mf_m)
    return n, sw, sww


def mf_update_weight_models_loops(E_A, E_W, E_WWT, mf_m, nat_prior):
    """stochastic_block_model.py:340-344 -> internals/weights.py:105-132: per cell, mf_natural_hypparam =
    natural_hypparam + weighted statistics with weights E_A * pc1c2.  nat_prior: [C][C] natural hyperparameters."""
    N, C = mf_m.shape
    D = E_W.shape[2]
    out = [[None] * C for _ in range(C)]
    for c1 in range(C):
        for c2 in range(C):
            pc1c2 = mf_m[:, c1][:, None] * mf_m[:, c2][None, :]
            w = (E_A * pc1c2).reshape(N * N)
            out[c1][c2] = nat_prior[c1][c2] + weighted_statistics(E_W.reshape(N * N, D), E_WWT.reshape(N * N, D, D), w)
    return out


class _OwnDists(object):
    """the restated distributions, behind the call shapes the reference's loop uses"""
    @staticmethod
    def beta_logs(t1, t0):
        return beta_expected_logs(t1, t0)

    @staticmethod
    def bern_negentropy(E_x, E_notx, E_ln_p, E_ln_notp):
        return bernoulli_negentropy(E_x, E_notx, E_ln_p, E_ln_notp)

    @staticmethod
    def gauss_negentropy(**kw):
        return gaussian_negentropy(**kw)


def mf_update_c_loops(E_A, E_W, E_WWT, mf_m, mf_pi, mf_tau1, mf_tau0, wexp, stepsize=1.0, dists=_OwnDists, nodes=None,
                      literal_reverse=True):
    """Literal transcription of stochastic_block_model.py:392-445.  wexp[c1][c2] = (E_mu, E_mumuT, E_Sigma_inv,
    E_logdet_Sigma) of the cell's weight model, or None for an unweighted model.  Returns (mf_m, lp[N, C])."""
    N, C = mf_m.shape
    mf_m = mf_m.copy()
    E_notA = 1 - E_A
    lps = np.zeros((N, C))
    E_log_m = psi(mf_pi) - psi(mf_pi.sum())

    def ell_w(n1, cn1, n2, cn2):
        if wexp is None:
            return 0.0
        E_mu, E_mumuT, E_Sinv, E_ld = wexp[cn1][cn2]
        return dists.gauss_negentropy(E_x=E_W[n1, n2, :], E_xxT=E_WWT[n1, n2, :, :], E_mu=E_mu, E_mumuT=E_mumuT,
                                      E_Sigma_inv=E_Sinv, E_logdet_Sigma=E_ld)

    def reverse(n1, cn1, n2, cn2, pcn2):
        E_ln_p, E_ln_notp = dists.beta_logs(mf_tau1[cn2, cn1], mf_tau0[cn2, cn1])
        return (pcn2 * dists.bern_negentropy(E_A[n2, n1], E_notA[n2, n1], E_ln_p, E_ln_notp)
                + E_A[n2, n1] * pcn2 * ell_w(n2, cn2, n1, cn1))

    for n1 in (range(N) if nodes is None else nodes):
        lp = np.zeros(C)
        lp += E_log_m
        for cn1 in range(C):
            for n2 in range(N):
                for cn2 in range(C):
                    pcn2 = mf_m[n2, cn2]
                    E_ln_p, E_ln_notp = dists.beta_logs(mf_tau1[cn1, cn2], mf_tau0[cn1, cn2])
                    lp[cn1] += pcn2 * dists.bern_negentropy(E_A[n1, n2], E_notA[n1, n2], E_ln_p, E_ln_notp)
                    lp[cn1] += E_A[n1, n2] * pcn2 * ell_w(n1, cn1, n2, cn2)
                    if not literal_reverse and n2 != n1:
                        lp[cn1] += reverse(n1, cn1, n2, cn2, pcn2)
                # REFERENCE QUIRK: `if n2 != n1:` (:425) is indented at the level of `for cn2` (:410), i.e. it runs
                # once per n2 AFTER the cn2 loop, with cn2 == C-1 and pcn2 == mf_m[n2, C-1] left over from it.
                if literal_reverse and n2 != n1:
                    lp[cn1] += reverse(n1, cn1, n2, cn2, pcn2)
        Z = logsumexp(lp)
        mk_hat = np.exp(lp - Z)
        lps[n1] = lp
        mf_m[n1, :] = (1.0 - stepsize) * mf_m[n1, :] + stepsize * mk_hat
    return mf_m, lps


# ---- vectorised form (T1): per-cell coefficient tables + soft per-node statistics ---------------------------
def cell_tables(mf_tau1, mf_tau0, wexp, B, literal_reverse=True):
    """Coefficients of the per-pair term in lp[k] for a partner of class k':
         forward  (n1 -> n2):  base_out[k,k'] + a12 * (c0_out + c1_out . w12 + w12^T C2_out w12)      cell (k , k')
         reverse  (n2 -> n1):  base_in [k,k'] + a21 * (c0_in  + c1_in  . w21 + w21^T C2_in  w21)      cell (k', k )
       with base = E[ln(1-p)], c0 = E[ln p] - E[ln(1-p)] + the constant of Gaussian.negentropy, c1 = E[Sigma^-1] E[mu],
       C2 = -E[Sigma^-1]/2.  literal_reverse masks the reverse tables to k' == C-1 (the reference's indentation quirk)."""
    C = mf_tau1.shape[0]
    Elp, Elnp = beta_expected_logs(mf_tau1, mf_tau0)
    base = Elnp.copy()
    c0 = Elp - Elnp
    c1 = np.zeros((C, C, B)); C2 = np.zeros((C, C, B, B))
    if wexp is not None:
        for a in range(C):
            for b in range(C):
                E_mu, E_mumuT, E_Sinv, E_ld = wexp[a][b]
                c0[a, b] += -0.5 * np.log(2 * np.pi) - 0.5 * E_ld - 0.5 * np.trace(E_Sinv.dot(E_mumuT))
                c1[a, b] = E_Sinv.dot(E_mu)          # x^T E_Sinv E_mu
                C2[a, b] = -0.5 * E_Sinv
    mask = np.ones(C)
    if literal_reverse:
        mask[:] = 0.0; mask[C - 1] = 1.0
    t = dict(base_out=base, c0_out=c0, c1_out=c1, C2_out=C2,
             base_in=base.T * mask[None, :], c0_in=c0.T * mask[None, :],
             c1_in=np.transpose(c1, (1, 0, 2)) * mask[None, :, None],
             C2_in=np.transpose(C2, (1, 0, 2, 3)) * mask[None, :, None, None])
    return t


def node_lp(n1, A, W, mf_m, E_log_m, t):
    """lp[k] of node n1 from the current mf_m (one row and one column of A / W)."""
    N, C = mf_m.shape
    a_out = A[n1, :].astype(float); a_in = A[:, n1].astype(float).copy(); a_in[n1] = 0.0
    notself = np.ones(N); notself[n1] = 0.0
    lp = E_log_m + t["base_out"].dot(mf_m.sum(0)) + t["base_in"].dot((mf_m * notself[:, None]).sum(0))
    n_out = (a_out[:, None] * mf_m).sum(0)                       # [k']
    n_in = (a_in[:, None] * mf_m).sum(0)
    lp = lp + t["c0_out"].dot(n_out) + t["c0_in"].dot(n_in)
    if W is not None and W.shape[-1] > 0:
        w_out, w_in = W[n1, :, :], W[:, n1, :]
        s_out = np.einsum("j,jb,jk->kb", a_out, w_out, mf_m); s_in = np.einsum("j,jb,jk->kb", a_in, w_in, mf_m)
        ss_out = np.einsum("j,jb,jc,jk->kbc", a_out, w_out, w_out, mf_m)
        ss_in = np.einsum("j,jb,jc,jk->kbc", a_in, w_in, w_in, mf_m)
        lp = lp + np.einsum("klb,lb->k", t["c1_out"], s_out) + np.einsum("klb,lb->k", t["c1_in"], s_in)
        lp = lp + np.einsum("klbc,lbc->k", t["C2_out"], ss_out) + np.einsum("klbc,lbc->k", t["C2_in"], ss_in)
    return lp


def mf_update_c(A, W, mf_m, mf_pi, mf_tau1, mf_tau0, wexp, stepsize=1.0, literal_reverse=True, nodes=None):
    """Same sequential update as mf_update_c_loops for a fixed network (E_A = A, E_W = W, E_WWT = W W^T)."""
    N, C = mf_m.shape
    B = 0 if W is None else W.shape[-1]
    mf_m = mf_m.copy()
    E_log_m = psi(mf_pi) - psi(mf_pi.sum())
    t = cell_tables(mf_tau1, mf_tau0, wexp, B, literal_reverse)
    lps = np.zeros((N, C))
    for n1 in (range(N) if nodes is None else nodes):
        lp = node_lp(n1, A, W, mf_m, E_log_m, t)
        lps[n1] = lp
        mk_hat = np.exp(lp - logsumexp(lp))
        mf_m[n1, :] = (1.0 - stepsize) * mf_m[n1, :] + stepsize * mk_hat
    return mf_m, lps


def mf_update_p(A, mf_m, tau1, tau0, mf_tau1, mf_tau0, stepsize=1.0):
    """stochastic_block_model.py:366-389 as two small matrix products."""
    n1 = mf_m.T.dot(A.astype(float)).dot(mf_m)
    ms = mf_m.sum(0)
    n0 = np.outer(ms, ms) - n1
    return ((1.0 - stepsize) * mf_tau1 + stepsize * (tau1 + n1), (1.0 - stepsize) * mf_tau0 + stepsize * (tau0 + n0))


# ---- variational lower bound (stochastic_block_model.py:536-568) ---------------------------------------------------
def discrete_negentropy(p, E_x=None, E_ln_p=None):
    """internals/distributions.py:36-50 (nan_to_num on the SUM: a vector with an exact zero gives 0)."""
    E_x = p if E_x is None else E_x
    with np.errstate(divide="ignore", invalid="ignore"):
        E_ln_p = np.log(p) if E_ln_p is None else E_ln_p
        return np.nan_to_num(E_x.dot(E_ln_p))


def dirichlet_negentropy(gamma, E_ln_g=None):
    """internals/distributions.py:327-340"""
    from scipy.special import gammaln
    gamma = np.asarray(gamma, dtype=np.float64)
    if E_ln_g is None:
        E_ln_g = psi(gamma) - psi(gamma.sum(axis=-1, keepdims=True))
    return gammaln(gamma.sum(axis=-1, keepdims=True)).sum() - gammaln(gamma).sum() + ((gamma - 1) * E_ln_g).sum()


def beta_negentropy(tau1, tau0, E_ln_p=None, E_ln_notp=None):
    """internals/distributions.py:343-372 (gamma = [tau1, tau0] on a new last axis, after atleast_1d)."""
    gamma = np.concatenate((np.atleast_1d(tau1)[..., None], np.atleast_1d(tau0)[..., None]), axis=-1)
    E_ln_g = None
    if E_ln_p is not None and E_ln_notp is not None:
        E_ln_g = np.concatenate((E_ln_p[..., None], E_ln_notp[..., None]), axis=-1)
    return dirichlet_negentropy(gamma, E_ln_g)


def niw_vlb(mu_0, sigma_0, kappa_0, nu_0, mu_mf, sigma_mf, kappa_mf, nu_mf):
    """pybasicbayes ``Gaussian.get_vlb`` (dependency absent from /root/reference: restated from Bishop 10.74 / 10.77,
    parity unpinned by the reference; pinned instead by the identity  get_vlb = -KL(q || prior)  in tests/test_oracle_mf.py:
    zero at q = prior and equal to a Monte Carlo estimate of E_q[ln p - ln q])."""
    from scipy.special import gammaln
    D = len(mu_0)

    def logZ(sigma, nu):
        hl = np.log(np.diag(np.linalg.cholesky(sigma))).sum()
        return -(nu * hl - (nu * D / 2.0 * np.log(2) + D * (D - 1) / 4.0 * np.log(np.pi) + gammaln((nu - np.arange(D)) / 2.0).sum()))

    hl = np.log(np.diag(np.linalg.cholesky(sigma_mf))).sum()
    E_logdet_lmbda = psi((nu_mf - np.arange(D)) / 2.0).sum() + D * np.log(2) - 2 * hl
    iw_entropy = logZ(sigma_mf, nu_mf) - (nu_mf - D - 1) / 2.0 * E_logdet_lmbda + nu_mf * D / 2.0
    llt = loglmbdatilde(sigma_mf, nu_mf)
    q_entropy = -0.5 * (llt + D * (np.log(kappa_mf / (2 * np.pi)) - 1)) + iw_entropy
    d = mu_mf - mu_0
    p_avg = 0.5 * (D * np.log(kappa_0 / (2 * np.pi)) + llt - D * kappa_0 / kappa_mf
                   - kappa_0 * nu_mf * d.dot(np.linalg.solve(sigma_mf, d))) \
        - logZ(sigma_0, nu_0) + (nu_0 - D - 1) / 2.0 * llt - 0.5 * nu_mf * np.trace(np.linalg.solve(sigma_mf, sigma_0))
    return p_avg + q_entropy


def sbm_vlb_loops(mf_m, pi, mf_pi, tau1, tau0, mf_tau1, mf_tau0, niw_terms):
    """stochastic_block_model.py:536-568, statement by statement; niw_terms[c1][c2] = weight_models[c1][c2].get_vlb()."""
    N, C = mf_m.shape
    vlb = 0.0
    E_ln_m = psi(mf_pi) - psi(mf_pi.sum())
    for n in range(N):
        vlb += discrete_negentropy(None, E_x=mf_m[n, :], E_ln_p=E_ln_m)
        vlb -= discrete_negentropy(mf_m[n, :])
    vlb += beta_negentropy(tau1, tau0, E_ln_p=psi(mf_tau1) - psi(mf_tau0 + mf_tau1),
                           E_ln_notp=psi(mf_tau0) - psi(mf_tau0 + mf_tau1)).sum()
    vlb -= beta_negentropy(mf_tau1, mf_tau0).sum()
    vlb += dirichlet_negentropy(pi, E_ln_g=E_ln_m)
    vlb -= dirichlet_negentropy(mf_pi)
    for c1 in range(C):
        for c2 in range(C):
            vlb += niw_terms[c1][c2]
    return vlb
