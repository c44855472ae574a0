"""
CPU oracle (TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg,
never by the product path) for the mean-field logistic eigenmodel, SURVEY.md row a19.

Restates, in Python 3 / fp64, graphistician/eigenmodel.py:
  :1177-1189  mf_expected_mu           :1191-1224  mf_expected_musq / mf_variance_mu / mf_expected_mu0sq
  :1226-1240  mf_expected_ffT / llT    :1245-1248  mf_sample_mus
  :1055-1064  _meanfieldupdate_Omega   (Monte Carlo over N_samples = 1000 activations PER ELEMENT)
  :1066-1088  _meanfieldupdate_mu0     :1090-1124  _meanfieldupdate_F  (`*_loops` = the literal loop nest)
  :1126-1157  _meanfieldupdate_lmbda   :998-1022   meanfieldupdate (order: Omega, mu0, F, lmbda)

Parity unpinned by the reference: eigenmodel.py cannot be imported (Python 2 syntax at :122; undefined base classes at
:1319), and it holds no golden vector for this path; the loop transcriptions are cross-checked against the vectorised
forms and frozen in tests/golden/mf_eigen_small.npz.

Reference quirks kept:
  * _meanfieldupdate_F (:1113-1114) builds h_n from `self.F` -- the GIBBS sample of the features -- not from mf_mu_F.
    `F_lin` is that array; pass mf_mu_F (and track=True) for the evidently intended update.
  * zcent = kappa(E_A)/E_O - E_mu0 is only ever used multiplied by E_O, i.e. as kappa - E_O*E_mu0 (no division needed).
  * kappa(E_A) = E_A - 1/2 on every entry, diagonal included (:813-820), and the mu0 update sums it over the diagonal too.
"""
import numpy as np


def expected_ffT(mf_mu_F, mf_Sigma_F):
    """:1226-1232 for every n: [N, D, D]."""
    return mf_Sigma_F + mf_mu_F[:, :, None] * mf_mu_F[:, None, :]


def expected_llT(mf_mu_lmbda, mf_Sigma_lmbda):
    """:1234-1240"""
    return mf_Sigma_lmbda + np.outer(mf_mu_lmbda, mf_mu_lmbda)


def expected_mu(mf_mu_mu0, mf_mu_F, mf_mu_lmbda):
    """:1177-1189"""
    E_mu = mf_mu_mu0 + (mf_mu_F * mf_mu_lmbda[None, :]).dot(mf_mu_F.T)
    np.fill_diagonal(E_mu, mf_mu_mu0)
    return E_mu


def expected_musq(mf_mu_mu0, mf_sigma_mu0, mf_mu_F, mf_Sigma_F, mf_mu_lmbda, mf_Sigma_lmbda):
    """:1191-1218 (the N^2 loop of :1209-1213 as one [N, D^2] x [D^2, N] product)."""
    N, D = mf_mu_F.shape
    mu0sq = mf_sigma_mu0 + mf_mu_mu0 ** 2
    E = np.zeros((N, N)) + mu0sq
    E += 2 * mf_mu_mu0 * (mf_mu_F * mf_mu_lmbda[None, :]).dot(mf_mu_F.T)
    ffT = expected_ffT(mf_mu_F, mf_Sigma_F).reshape(N, D * D)
    llT = expected_llT(mf_mu_lmbda, mf_Sigma_lmbda).reshape(D * D)
    E += (ffT * llT[None, :]).dot(ffT.T)
    np.fill_diagonal(E, mu0sq)
    return E


def variance_mu(*args):
    """:1220-1221"""
    return expected_musq(*args) - expected_mu(args[0], args[2], args[4]) ** 2


def expected_omega_mc(E_mu, Var_mu, normals):
    """:1055-1064 with the standard normals of mf_sample_mus (:1245-1248) supplied: normals [S, N, N]."""
    Zs = E_mu[None] + np.sqrt(Var_mu)[None] * normals
    return 0.5 * (np.tanh(Zs / 2.0) / Zs).mean(axis=0)


def expected_omega_quad(E_mu, Var_mu, order=80):
    """The expectation that Monte Carlo average estimates, by Gauss-Hermite quadrature (test reference)."""
    x, w = np.polynomial.hermite_e.hermegauss(order)
    w = w / w.sum()
    Z = E_mu[..., None] + np.sqrt(np.maximum(Var_mu, 0.0))[..., None] * x
    with np.errstate(invalid="ignore", divide="ignore"):
        g = np.where(np.abs(Z) < 1e-6, 0.25 - Z ** 2 / 48.0, np.tanh(Z / 2.0) / (2.0 * Z))
    return (g * w).sum(-1)


def kappa(E_A):
    return E_A - 0.5


def meanfieldupdate_mu0(E_A, E_O, mf_mu_F, mf_mu_lmbda, mu_mu_0, sigma_mu0, mf_mu_mu0, mf_sigma_mu0, stepsize=1.0):
    """:1066-1088.  Returns (mf_mu_mu0, mf_sigma_mu0, post_prec, post_mean_dot_prec)."""
    post_prec = 1.0 / sigma_mu0 + E_O.sum()
    E_FdotF = (mf_mu_F * mf_mu_lmbda[None, :]).dot(mf_mu_F.T)
    np.fill_diagonal(E_FdotF, 0)
    h = mu_mu_0 / sigma_mu0 + (kappa(E_A) - E_O * E_FdotF).sum()
    new_sigma = (1 - stepsize) * mf_sigma_mu0 + stepsize * 1.0 / post_prec
    new_mu = (1 - stepsize) * mf_mu_mu0 + stepsize * new_sigma * h
    return new_mu, new_sigma, post_prec, h


def meanfieldupdate_F_loops(E_A, E_O, F_lin, mf_mu_F, mf_Sigma_F, E_mu0, E_lmbda, E_llT, sigma_F, stepsize=1.0, track=False):
    """Literal transcription of :1090-1124.  F_lin = `self.F` (the reference's quirk); track=True makes the h term follow
    mf_mu_F as it is updated (the intended form).  Returns (mf_mu_F, mf_Sigma_F, prec [N,D,D], h [N,D])."""
    N, D = mf_mu_F.shape
    mf_mu_F, mf_Sigma_F = mf_mu_F.copy(), mf_Sigma_F.copy()
    precs, hs = np.zeros((N, D, D)), np.zeros((N, D))
    with np.errstate(divide="ignore", invalid="ignore"):
        zcent = kappa(E_A) / E_O - E_mu0
    for n in range(N):
        post_prec = 1.0 / sigma_F * np.eye(D)
        post_mean_dot_prec = np.zeros(D)
        for nn in range(N):
            if nn == n:
                continue
            E_ffT = mf_Sigma_F[nn] + np.outer(mf_mu_F[nn], mf_mu_F[nn])
            post_prec += E_ffT * E_llT * (E_O[n, nn] + E_O[nn, n])
            Fl = mf_mu_F[nn] if track else F_lin[nn]
            post_mean_dot_prec += E_O[n, nn] * zcent[n, nn] * Fl * E_lmbda
            post_mean_dot_prec += E_O[nn, n] * zcent[nn, n] * Fl * E_lmbda
        precs[n], hs[n] = post_prec, post_mean_dot_prec
        mf_Sigma_F[n] = (1 - stepsize) * mf_Sigma_F[n] + stepsize * np.linalg.inv(post_prec)
        mf_mu_F[n] = (1 - stepsize) * mf_mu_F[n] + stepsize * mf_Sigma_F[n].dot(post_mean_dot_prec)
    return mf_mu_F, mf_Sigma_F, precs, hs


def meanfieldupdate_F(E_A, E_O, F_lin, mf_mu_F, mf_Sigma_F, E_mu0, E_lmbda, E_llT, sigma_F, stepsize=1.0, track=False):
    """Same sequential update, one matrix-vector contraction per node (S = E_O + E_O^T, R = kappa + kappa^T - E_mu0 S)."""
    N, D = mf_mu_F.shape
    mf_mu_F, mf_Sigma_F = mf_mu_F.copy(), mf_Sigma_F.copy()
    S = E_O + E_O.T
    R = kappa(E_A) + kappa(E_A).T - E_mu0 * S
    np.fill_diagonal(S, 0.0); np.fill_diagonal(R, 0.0)
    precs, hs = np.zeros((N, D, D)), np.zeros((N, D))
    ffT = expected_ffT(mf_mu_F, mf_Sigma_F)
    for n in range(N):
        prec = np.eye(D) / sigma_F + np.einsum("m,mab->ab", S[n], ffT) * E_llT
        h = R[n].dot(mf_mu_F if track else F_lin) * E_lmbda
        precs[n], hs[n] = prec, h
        mf_Sigma_F[n] = (1 - stepsize) * mf_Sigma_F[n] + stepsize * np.linalg.inv(prec)
        mf_mu_F[n] = (1 - stepsize) * mf_mu_F[n] + stepsize * mf_Sigma_F[n].dot(h)
        ffT[n] = mf_Sigma_F[n] + np.outer(mf_mu_F[n], mf_mu_F[n])
    return mf_mu_F, mf_Sigma_F, precs, hs


def meanfieldupdate_lmbda_loops(E_A, E_O, mf_mu_F, mf_Sigma_F, E_mu0, mu_lmbda, sigma_lmbda, mf_mu_lmbda, mf_Sigma_lmbda,
                                stepsize=1.0):
    """Literal transcription of :1126-1157.  Returns (mf_mu_lmbda, mf_Sigma_lmbda, post_prec, post_mean_dot_prec)."""
    N, D = mf_mu_F.shape
    post_prec = 1.0 / sigma_lmbda * np.eye(D)
    post_mean_dot_prec = mu_lmbda / sigma_lmbda * np.ones(D)
    with np.errstate(divide="ignore", invalid="ignore"):
        zcent = kappa(E_A) / E_O - E_mu0
    ffT = expected_ffT(mf_mu_F, mf_Sigma_F)
    for n1 in range(N):
        for n2 in range(n1 + 1, N):
            post_prec = post_prec + (E_O[n2, n1] + E_O[n1, n2]) * ffT[n1] * ffT[n2]
            E_f1f2 = mf_mu_F[n1] * mf_mu_F[n2]
            post_mean_dot_prec = post_mean_dot_prec + E_O[n1, n2] * zcent[n1, n2] * E_f1f2
            post_mean_dot_prec = post_mean_dot_prec + E_O[n2, n1] * zcent[n2, n1] * E_f1f2
    new_Sigma = (1.0 - stepsize) * mf_Sigma_lmbda + stepsize * np.linalg.inv(post_prec)
    new_mu = (1.0 - stepsize) * mf_mu_lmbda + stepsize * new_Sigma.dot(post_mean_dot_prec)
    return new_mu, new_Sigma, post_prec, post_mean_dot_prec


def pair_stats(E_A, E_O, mf_mu_F, mf_Sigma_F):
    """Pair sums over n1 < n2 that the lmbda (and mu0) updates need:
         Pq[a,b] = sum S12 ffT1[a,b] ffT2[a,b],  hS[a] = sum S12 f1a f2a,  hK[a] = sum (kappa12 + kappa21) f1a f2a."""
    S = E_O + E_O.T
    K = kappa(E_A) + kappa(E_A).T
    iu = np.triu(np.ones_like(S), 1)
    ffT = expected_ffT(mf_mu_F, mf_Sigma_F)
    Pq = np.einsum("nm,nab,mab->ab", S * iu, ffT, ffT)
    hS = np.einsum("nm,na,ma->a", S * iu, mf_mu_F, mf_mu_F)
    hK = np.einsum("nm,na,ma->a", K * iu, mf_mu_F, mf_mu_F)
    return Pq, hS, hK


def meanfieldupdate_lmbda(E_A, E_O, mf_mu_F, mf_Sigma_F, E_mu0, mu_lmbda, sigma_lmbda, mf_mu_lmbda, mf_Sigma_lmbda,
                          stepsize=1.0):
    D = mf_mu_F.shape[1]
    Pq, hS, hK = pair_stats(E_A, E_O, mf_mu_F, mf_Sigma_F)
    post_prec = np.eye(D) / sigma_lmbda + Pq
    h = mu_lmbda / sigma_lmbda * np.ones(D) + hK - E_mu0 * hS
    new_Sigma = (1.0 - stepsize) * mf_Sigma_lmbda + stepsize * np.linalg.inv(post_prec)
    new_mu = (1.0 - stepsize) * mf_mu_lmbda + stepsize * new_Sigma.dot(h)
    return new_mu, new_Sigma, post_prec, h


# ---- variational lower bound (eigenmodel.py:1262-1298) ---------------------------------------------------------------
def scalar_gaussian_negentropy(mu, sigmasq, E_x=None, E_xsq=None):
    """internals/distributions.py:130-158 with the distribution's own parameters for mu / sigmasq."""
    E_x = mu if E_x is None else E_x
    E_xsq = sigmasq + mu ** 2 if E_xsq is None else E_xsq
    H = -0.5 * np.log(2 * np.pi)
    H += -0.5 * np.log(sigmasq)
    H += -0.5 / sigmasq * E_xsq
    H += 1.0 / sigmasq * E_x * mu
    H += -0.5 / sigmasq * mu ** 2
    return H


def gaussian_negentropy(mu, Sigma, E_x=None, E_xxT=None):
    """internals/distributions.py:278-308 with the distribution's own parameters (note -0.5 log 2 pi, not D times it)."""
    E_x = mu if E_x is None else E_x
    E_xxT = Sigma + np.outer(mu, mu) if E_xxT is None else E_xxT
    Si = np.linalg.inv(Sigma)
    H = -0.5 * np.log(2 * np.pi)
    H += -0.5 * np.linalg.slogdet(Sigma)[1]
    H += -0.5 * np.trace(Si.dot(E_xxT))
    H += E_x.T.dot(Si).dot(mu)
    H += -0.5 * np.trace(Si.dot(np.outer(mu, mu)))
    return H


def get_vlb_loops(mu_mu_0, sigma_mu0, sigma_F, mu_lmbda, sigma_lmbda, mf_mu_mu0, mf_sigma_mu0, mf_mu_F, mf_Sigma_F,
                  mf_mu_lmbda, mf_Sigma_lmbda, lmbda_given=False):
    """eigenmodel.py:1262-1298, statement by statement (prior / entropy terms only: the reference leaves out the
    likelihood because the entropy of q(omega) is unknown)."""
    N, D = mf_mu_F.shape
    vlb = 0.0
    vlb += scalar_gaussian_negentropy(mu_mu_0, sigma_mu0, E_x=mf_mu_mu0, E_xsq=mf_sigma_mu0 + mf_mu_mu0 ** 2)
    vlb -= scalar_gaussian_negentropy(mf_mu_mu0, mf_sigma_mu0)
    for n in range(N):
        vlb += gaussian_negentropy(np.zeros(D), sigma_F * np.eye(D), E_x=mf_mu_F[n, :], E_xxT=mf_Sigma_F[n] + np.outer(mf_mu_F[n], mf_mu_F[n]))      # mf_expected_ffT(n), :1226-1232
        vlb -= gaussian_negentropy(mf_mu_F[n, :], mf_Sigma_F[n, :, :])
    if not lmbda_given:
        vlb += gaussian_negentropy(mu_lmbda, sigma_lmbda * np.eye(D), E_x=mf_mu_lmbda, E_xxT=expected_llT(mf_mu_lmbda, mf_Sigma_lmbda))
        vlb -= gaussian_negentropy(mf_mu_lmbda, mf_Sigma_lmbda)
    return vlb
