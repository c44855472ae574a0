"""
oracle/eigenmodel.py -- TEST INFRASTRUCTURE (see oracle/__init__.py).

CPU restatement of the (logistic / probit) eigenmodel Gibbs hot path, eigenmodel.py:75-83,
227-236, 268-379, 803-965.  ``*_loops`` = literal transcription (T0); un-suffixed = contraction
form (T1) with identical arithmetic up to summation order.

Draws are injected: Omega (Polya-gamma) or uniforms (probit Z), and standard normals for the
Gaussian conditionals.  The reference samples with ``np.random.multivariate_normal`` (SVD map,
eigenmodel.py:908); parity is therefore on the *conditionals* (precision, mean-dot-precision) and,
for samples, on the fixed map  f = mean + L^-T z  with  prec = L L^T  (documented in DESIGN.md).
"""
import numpy as np
import scipy.linalg
from scipy.special import erfc, erfcinv


def mu_matrix(F, mu_0, lmbda):
    """eigenmodel.py:75-83: Mu = mu_0 + F diag(lmbda) F^T with the diagonal forced to mu_0."""
    Mu = mu_0 + (F * lmbda[None, :]).dot(F.T)
    np.fill_diagonal(Mu, mu_0)
    return Mu


def logistic_P(F, mu_0, lmbda):
    """eigenmodel.py:803-811."""
    P = 1.0 / (1.0 + np.exp(-mu_matrix(F, mu_0, lmbda)))
    return np.clip(P, 1e-16, 1 - 1e-16)


def probit_P(F, mu_0, lmbda):
    """eigenmodel.py:227-236 (0.5*(1+erf(Mu/sqrt2)) == 0.5*erfc(-Mu/sqrt2))."""
    from scipy.special import erf
    P = 0.5 * (1 + erf(mu_matrix(F, mu_0, lmbda) / np.sqrt(2.0)))
    return np.clip(P, 1e-16, 1 - 1e-16)


def log_likelihood(A, P):
    """eigenmodel.py:110-129: sum nan_to_num(A log P + (1-A) log(1-P))."""
    ll = A * np.log(P) + (1 - A) * np.log(1 - P)
    return np.nan_to_num(ll).sum()


def log_prior(F, mu_0, lmbda, sigma_F, mu_mu_0, sigma_mu0, sigma_lmbda=None):
    """eigenmodel.py:92-108."""
    lp = -0.5 * (F * F / sigma_F).sum()
    lp += -0.5 * (mu_0 - mu_mu_0) ** 2 / sigma_mu0
    if sigma_lmbda is not None:
        lp += -0.5 * (lmbda * lmbda / sigma_lmbda).sum()
    return lp


def gaussian_from_prec(prec, h, z):
    """Draw from N(prec^-1 h, prec^-1) with the fixed map x = mean + L^-T z, prec = L L^T."""
    L = np.linalg.cholesky(prec)
    y = scipy.linalg.solve_triangular(L, h, lower=True)
    mean = scipy.linalg.solve_triangular(L.T, y, lower=False)
    return mean + scipy.linalg.solve_triangular(L.T, z, lower=False), mean


# ----------------------------------------------------------------------------------------------
# Logistic (Polya-gamma) Gibbs
# ----------------------------------------------------------------------------------------------
def logistic_resample_F_loops(A, O, F, mu_0, lmbda, sigma_F, z):
    """eigenmodel.py:876-908, loop for loop, single graph.  z[N,D] = injected standard normals.
    Returns (F_new, prec[N,D,D], h[N,D]) -- the conditionals at the time each node is drawn."""
    N, D = F.shape
    F = F.copy()
    precs = np.zeros((N, D, D)); hs = np.zeros((N, D))
    llT = np.outer(lmbda, lmbda)
    for n in range(N):
        post_prec = 1.0 / sigma_F * np.eye(D)
        post_mean_dot_prec = np.zeros(D)
        zcent = (A - 0.5) / O - mu_0
        for nn in range(N):
            if nn == n:
                continue
            post_prec += np.outer(F[nn, :], F[nn, :]) * llT * O[n, nn]
            post_prec += np.outer(F[nn, :], F[nn, :]) * llT * O[nn, n]
            post_mean_dot_prec += O[n, nn] * zcent[n, nn] * F[nn, :] * lmbda
            post_mean_dot_prec += O[nn, n] * zcent[nn, n] * F[nn, :] * lmbda
        precs[n] = post_prec; hs[n] = post_mean_dot_prec
        F[n, :], _ = gaussian_from_prec(post_prec, post_mean_dot_prec, z[n])
    return F, precs, hs


def resample_F(S, Ksym_minus_mu0S, F, lmbda, sigma_F, z, base_scale=None, nodes=None):
    """Contraction form of the sequential F sweep.

    S = Omega + Omega^T (logistic) or 2 (probit), zero diagonal irrelevant (skipped);
    R = (kappa + kappa^T) - mu_0 S  (logistic)  or  zc + zc^T  (probit).
      prec_n = I/sigma_F + reshape(S[n,:] . (F (x) F)) o (lmbda lmbda^T)
      h_n    = (R[n,:] . F) o lmbda          both over m != n, with the *current* F.
    """
    N, D = F.shape
    F = F.copy()
    precs = np.zeros((N, D, D)); hs = np.zeros((N, D))
    llT = np.outer(lmbda, lmbda)
    # nodes: bounded sample for the timed CPU baseline -- a count (the first `nodes` nodes) or a (lo, hi) range
    lo, hi = (0, N) if nodes is None else ((0, nodes) if np.isscalar(nodes) else nodes)
    for n in range(lo, hi):
        s = S[n].copy(); s[n] = 0.0
        r = Ksym_minus_mu0S[n].copy(); r[n] = 0.0
        prec = np.eye(D) / sigma_F + (F.T * s).dot(F) * llT
        h = r.dot(F) * lmbda
        precs[n] = prec; hs[n] = h
        F[n], _ = gaussian_from_prec(prec, h, z[n])
    return F, precs, hs


def logistic_sym(A, O, mu_0):
    """S = O + O^T and R = (kappa + kappa^T) - mu_0 S with kappa = A - 1/2 (eigenmodel.py:813-820,892,898-901)."""
    S = O + O.T
    kap = A - 0.5
    R = kap + kap.T - mu_0 * S
    return S, R


def logistic_mu0_conditional(A, O, F, lmbda, mu_mu_0, sigma_mu0):
    """eigenmodel.py:910-932 -> (post_prec, post_mean_dot_prec)."""
    post_prec = 1.0 / sigma_mu0 + O.sum()
    FdotF = (F * lmbda[None, :]).dot(F.T)
    np.fill_diagonal(FdotF, 0)
    h = mu_mu_0 / sigma_mu0 + ((A - 0.5) - O * FdotF).sum()
    return post_prec, h


def logistic_lmbda_conditional_loops(A, O, F, mu_0, mu_lmbda, sigma_lmbda):
    """eigenmodel.py:934-961, loop for loop -> (post_prec[D,D], post_mean_dot_prec[D])."""
    N, D = F.shape
    post_prec = 1.0 / sigma_lmbda * np.eye(D)
    h = mu_lmbda / sigma_lmbda * np.ones(D)
    zcent = (A - 0.5) / O - mu_0
    for n1 in range(N):
        for n2 in range(n1 + 1, N):
            post_prec += (O[n2, n1] + O[n1, n2]) * np.outer(F[n1], F[n1]) * np.outer(F[n2], F[n2])
            Fn1n2 = F[n1] * F[n2]
            h += O[n1, n2] * zcent[n1, n2] * Fn1n2
            h += O[n2, n1] * zcent[n2, n1] * Fn1n2
    return post_prec, h


def lmbda_conditional(S, R, F, mu_lmbda, sigma_lmbda):
    """Contraction form of a13: prec = I/s + 1/2 sum_n (f_n f_n^T) o (S Phi)_n ; h = mu/s + 1/2 sum_n f_n o (R F)_n
    (S, R with zeroed diagonals)."""
    N, D = F.shape
    S0 = S.copy(); np.fill_diagonal(S0, 0.0)
    R0 = R.copy(); np.fill_diagonal(R0, 0.0)
    FF = np.einsum("na,nb->nab", F, F)
    SPhi = np.einsum("nm,mab->nab", S0, FF)
    prec = np.eye(D) / sigma_lmbda + 0.5 * np.einsum("nab,nab->ab", FF, SPhi)
    h = mu_lmbda / sigma_lmbda * np.ones(D) + 0.5 * np.einsum("na,na->a", F, R0.dot(F))
    return prec, h


def logistic_gibbs_sweep(A, O, state, hyp, zF, zmu, zlmb=None):
    """One ``_GibbsLogisticEigenmodel.resample`` (eigenmodel.py:836-857) with injected Omega and normals.
    state = dict(F, mu_0, lmbda); hyp = dict(sigma_F, mu_mu_0, sigma_mu0, lmbda_given, mu_lmbda, sigma_lmbda)."""
    S, R = logistic_sym(A, O, state["mu_0"])
    F, precs, hs = resample_F(S, R, state["F"], state["lmbda"], hyp["sigma_F"], zF)
    pp, ph = logistic_mu0_conditional(A, O, F, state["lmbda"], hyp["mu_mu_0"], hyp["sigma_mu0"])
    mu_0 = ph / pp + np.sqrt(1.0 / pp) * zmu
    lmbda = state["lmbda"]
    lam_cond = None
    if not hyp.get("lmbda_given", True):
        S2, R2 = logistic_sym(A, O, mu_0)
        lam_cond = lmbda_conditional(S2, R2, F, hyp["mu_lmbda"], hyp["sigma_lmbda"])
        lmbda, _ = gaussian_from_prec(lam_cond[0], lam_cond[1], zlmb)
    return dict(F=F, mu_0=mu_0, lmbda=lmbda), dict(prec=precs, h=hs, mu0=(pp, ph), lmbda=lam_cond)


# ----------------------------------------------------------------------------------------------
# Probit Gibbs
# ----------------------------------------------------------------------------------------------
def normal_cdf(x, mu=0.0, sigma=1.0):
    """internals/utils.py:15-17."""
    return 0.5 * erfc(-((x - mu) / sigma) / np.sqrt(2))


def sample_truncnorm(mu, lb, ub, u, sigma=1.0):
    """internals/utils.py:53-76 with the uniforms injected (``np.random.rand(*shp)`` at :66)."""
    cdflb = normal_cdf(lb, mu, sigma)
    cdfub = normal_cdf(ub, mu, sigma)
    cdfsamples = cdflb + u * (cdfub - cdflb)
    cdfsamples = np.clip(cdfsamples, 1e-15, 1 - 1e-15)
    zs = -np.sqrt(2) * erfcinv(2 * cdfsamples)
    return sigma * zs + mu


def probit_resample_Z(A, F, mu_0, lmbda, u):
    """eigenmodel.py:268-287."""
    N = A.shape[0]
    Mu = mu_matrix(F, mu_0, lmbda)
    lb = -np.inf * np.ones((N, N)); lb[A == 1] = 0
    ub = np.inf * np.ones((N, N)); ub[A == 0] = 0
    return sample_truncnorm(Mu, lb, ub, u)


def probit_resample_F_loops(Z, F, mu_0, lmbda, sigma_F, z):
    """eigenmodel.py:289-321, loop for loop."""
    N, D = F.shape
    F = F.copy()
    precs = np.zeros((N, D, D)); hs = np.zeros((N, D))
    llT = np.outer(lmbda, lmbda)
    for n in range(N):
        post_prec = 1.0 / sigma_F * np.eye(D)
        h = np.zeros(D)
        zcent = Z - mu_0
        for nn in range(N):
            if nn == n:
                continue
            post_prec += 2 * np.outer(F[nn], F[nn]) * llT
            h += zcent[nn, n] * F[nn] * lmbda
            h += zcent[n, nn] * F[nn] * lmbda
        precs[n] = post_prec; hs[n] = h
        F[n], _ = gaussian_from_prec(post_prec, h, z[n])
    return F, precs, hs


def probit_sym(Z, mu_0):
    """S == 2, R = zc + zc^T."""
    zc = Z - mu_0
    return 2.0 * np.ones_like(Z), zc + zc.T


def probit_mu0_conditional(Z, F, lmbda, mu_mu_0, sigma_mu0):
    """eigenmodel.py:323-347."""
    N = Z.shape[0]
    post_prec = 1.0 / sigma_mu0 + N ** 2 - N
    Zcent = Z - (F * lmbda[None, :]).dot(F.T)
    np.fill_diagonal(Zcent, 0.0)
    return post_prec, mu_mu_0 / sigma_mu0 + Zcent.sum()
