"""
Generate the committed golden fixtures under tests/golden/.

Run in the BUILD container only (needs /root/reference):   python tests/golden/make_golden.py

1. ref_internals.npz -- outputs of the reference's OWN code: graphistician/internals/utils.py and
   internals/distributions.py are imported unmodified by file path (they parse under Python 3) with a
   3-alias NumPy shim (np.Inf / np.int / np.float / np.bool were removed from NumPy 2).  These pin the
   device functions (logistic, normal_cdf, sample_truncnorm, expected_truncnorm, truncated-normal
   moments, Beta/Bernoulli expectations) against the reference itself.
2. sbm_*.npz / eigen_*.npz / latent_*.npz -- outputs of the literal loop transcriptions (T0) in oracle/,
   fixed seeds, small N.  The reference cannot run these paths (Python 2 + missing packages), so they
   are "parity unpinned by the reference"; the fixtures freeze the oracle so later edits cannot drift.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/graphistician"
sys.path.insert(0, ROOT)


def _import_reference_internals():
    for name, val in (("Inf", np.inf), ("int", int), ("float", float), ("bool", bool)):
        if not hasattr(np, name):
            setattr(np, name, val)
    pkg = types.ModuleType("graphistician"); pkg.__path__ = [REF]
    sub = types.ModuleType("graphistician.internals"); sub.__path__ = [os.path.join(REF, "internals")]
    sys.modules["graphistician"] = pkg
    sys.modules["graphistician.internals"] = sub
    mods = {}
    for name in ("utils", "distributions"):
        spec = importlib.util.spec_from_file_location(
            "graphistician.internals." + name, os.path.join(REF, "internals", name + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = m
        spec.loader.exec_module(m)
        mods[name] = m
    return mods["utils"], mods["distributions"]


def make_ref_internals():
    utils, dist = _import_reference_internals()
    rng = np.random.RandomState(1234)
    x = np.concatenate([np.linspace(-12, 12, 97), rng.randn(64) * 3])
    out = dict(x=x, logistic=utils.logistic(x), normal_cdf=utils.normal_cdf(x), normal_pdf=utils.normal_pdf(x))
    # sample_truncnorm with np.random seeded: record the uniforms it will consume
    N = 24
    mu = rng.randn(N, N) * 1.5
    A = (rng.rand(N, N) < 0.3).astype(float)
    lb = -np.inf * np.ones((N, N)); lb[A == 1] = 0
    ub = np.inf * np.ones((N, N)); ub[A == 0] = 0
    np.random.seed(99)
    u = np.random.rand(N, N)
    np.random.seed(99)
    z = utils.sample_truncnorm(mu=mu, lb=lb, ub=ub)
    out.update(tn_mu=mu, tn_A=A, tn_u=u, tn_z=z,
               tn_expected=utils.expected_truncnorm(mu=mu, lb=lb, ub=ub))
    # TruncatedScalarGaussian known-answer values of test/test_distributions.py
    t0 = dist.TruncatedScalarGaussian(lb=0)
    t1 = dist.TruncatedScalarGaussian(mu=3, sigmasq=np.sqrt(1), lb=0, ub=10)
    out.update(tsg0_mean=t0.expected_x(), tsg0_var=t0.variance_x(),
               tsg1_mean=t1.expected_x(), tsg1_var=t1.variance_x(), tsg1_negent=t1.negentropy())
    tg = dist.TruncatedScalarGaussian(mu=mu, sigmasq=1.0, lb=lb, ub=ub)
    out.update(tsg_mean=tg.expected_x(), tsg_var=tg.variance_x())
    # Beta / Bernoulli expectations used by the mean-field SBM
    t1_, t0_ = rng.gamma(2.0, 2.0, (5, 5)) + 0.1, rng.gamma(2.0, 2.0, (5, 5)) + 0.1
    b = dist.Beta(t1_, t0_)
    out.update(beta_t1=t1_, beta_t0=t0_, beta_Elogp=b.expected_log_p(), beta_Elognotp=b.expected_log_notp(),
               beta_Ep=b.expected_p())
    p = rng.rand(7)
    xb = (rng.rand(7) < 0.5).astype(float)
    out.update(bern_p=p, bern_x=xb, bern_lp=dist.Bernoulli(p).log_probability(xb))
    np.savez(os.path.join(HERE, "ref_internals.npz"), **out)
    print("ref_internals.npz", sorted(out))


def _sbm_problem(rng, N, K, B):
    c = rng.randint(0, K, N)
    m = rng.dirichlet(np.ones(K) * 3)
    p = rng.beta(1, 1, (K, K))
    mus = rng.randn(K, K, B)
    sig = np.zeros((K, K, B, B))
    for a in range(K):
        for b in range(K):
            X = rng.randn(B, B)
            sig[a, b] = X.dot(X.T) + 0.5 * np.eye(B)
    A = (rng.rand(N, N) < p[np.ix_(c, c)]).astype(np.uint8)
    W = rng.randn(N, N, B) + mus[np.ix_(c, c)]
    return A, W, c, m, p, mus, sig


def make_sbm():
    from oracle import sbm
    rng = np.random.RandomState(4321)
    N, K, B = 20, 4, 2
    A, W, c, m, p, mus, sig = _sbm_problem(rng, N, K, B)
    u = rng.rand(N)
    Af = A.astype(float)
    out = dict(A=A, W=W, c=c, m=m, p=p, mus=mus, sigmas=sig, u=u)
    st = sbm.block_pair_stats_loops(Af, W, c, K)
    out.update({"st_" + k: v for k, v in st.items()})
    c_old, lp_old = sbm.gibbs_sbm_resample_c_loops(Af, W, c, m, p, mus, sig, u)
    out.update(c_old=c_old, lp_old=lp_old)
    for sp in (True, False):
        ca, la = sbm.sbm_adjacency_resample_c_loops(Af, c, m, p, sp, u)
        cw, lw = sbm.sbm_weights_resample_c_loops(Af, W, c, m, mus, sig, sp, u)
        out.update({"c_adj_%d" % sp: ca, "lp_adj_%d" % sp: la, "c_w_%d" % sp: cw, "lp_w_%d" % sp: lw})
    cz, lz = sbm.ais_naive_sample_f_loops(Af, c, p, m, 0.37, u)
    out.update(c_ais=cz, lp_ais=lz, ais_beta=0.37)
    np.savez(os.path.join(HERE, "sbm_small.npz"), **out)
    print("sbm_small.npz")


def make_eigen():
    from oracle import eigenmodel as eg
    rng = np.random.RandomState(777)
    N, D = 12, 3
    F = rng.randn(N, D); lm = rng.randn(D); mu0 = -0.8
    A = (rng.rand(N, N) < eg.logistic_P(F, mu0, lm)).astype(np.uint8)
    O = rng.gamma(2.0, 0.1, (N, N)) + 0.02
    zF = rng.randn(N, D)
    Af = A.astype(float)
    F1, prec, h = eg.logistic_resample_F_loops(Af, O, F, mu0, lm, 1.3, zF)
    mu0c = eg.logistic_mu0_conditional(Af, O, F1, lm, -1.2, 0.9)
    lamc = eg.logistic_lmbda_conditional_loops(Af, O, F1, mu0, 0.0, 2.0)
    u = rng.rand(N, N)
    Z = eg.probit_resample_Z(Af, F, mu0, lm, u)
    F2, prec2, h2 = eg.probit_resample_F_loops(Z, F, mu0, lm, 1.3, zF)
    pm0 = eg.probit_mu0_conditional(Z, F2, lm, -1.2, 0.9)
    np.savez(os.path.join(HERE, "eigen_small.npz"), A=A, O=O, F=F, lmbda=lm, mu_0=mu0, zF=zF, sigma_F=1.3,
             F_logistic=F1, prec_logistic=prec, h_logistic=h, mu0_prec=mu0c[0], mu0_h=mu0c[1],
             lam_prec=lamc[0], lam_h=lamc[1], u=u, Z=Z, F_probit=F2, prec_probit=prec2, h_probit=h2,
             pmu0_prec=pm0[0], pmu0_h=pm0[1], ll_logistic=eg.log_likelihood(Af, eg.logistic_P(F, mu0, lm)),
             ll_probit=eg.log_likelihood(Af, eg.probit_P(F, mu0, lm)))
    print("eigen_small.npz")


def make_latent():
    from oracle import latent as lt
    rng = np.random.RandomState(555)
    N, D = 16, 2
    L = np.sqrt(2.0) * rng.randn(N, D); mu0, ms, sig = 0.4, -0.3, 2.0
    A = (rng.rand(N, N) < 1.0 / (1 + np.exp(-lt.logit_matrix(L, mu0, ms)))).astype(np.uint8)
    Af = A.astype(float)
    lp_lit = lt.hmc_log_probability(L, mu0, ms, Af, sig)
    lp, gL, g0, gs = lt.log_probability_and_grad(L, mu0, ms, Af, sig)
    dr = dict(L=(rng.randn(N, D), np.log(rng.rand())), mu_0=(rng.randn(), np.log(rng.rand())),
              mu_self=(rng.randn(), np.log(rng.rand())))
    L1, m1, s1, st, ar = lt.resample(L, mu0, ms, Af, sig, 0.01, 0.9, dr)
    np.savez(os.path.join(HERE, "latent_small.npz"), A=A, L=L, mu_0=mu0, mu_self=ms, sigma=sig,
             lp_literal=lp_lit, lp=lp, gL=gL, gmu0=g0, gmuself=gs,
             pL=dr["L"][0], luL=dr["L"][1], p0=dr["mu_0"][0], lu0=dr["mu_0"][1],
             ps=dr["mu_self"][0], lus=dr["mu_self"][1], L_new=L1, mu_0_new=m1, mu_self_new=s1,
             step=st, rate=ar)
    print("latent_small.npz")


def make_mf_sbm():
    """mf_sbm_small.npz: the literal mf_update_c loop nest (stochastic_block_model.py:392-445) run with the REFERENCE'S
    OWN Beta / Bernoulli / Gaussian classes (internals/distributions.py, imported unmodified) doing the per-pair
    arithmetic -- this fixture pins oracle/mf_sbm.py's restatement of them."""
    from oracle import mf_sbm as M
    _, dist = _import_reference_internals()

    class RefDists(object):
        @staticmethod
        def beta_logs(t1, t0):
            b = dist.Beta(t1, t0)
            return float(b.expected_log_p()[0]), float(b.expected_log_notp()[0])

        @staticmethod
        def bern_negentropy(E_x, E_notx, E_ln_p, E_ln_notp):
            return dist.Bernoulli().negentropy(E_x=np.float64(E_x), E_notx=np.float64(E_notx),
                                               E_ln_p=E_ln_p, E_ln_notp=E_ln_notp)

        @staticmethod
        def gauss_negentropy(**kw):
            return dist.Gaussian().negentropy(**kw)

    rng = np.random.RandomState(2468)
    N, C, B = 10, 3, 2
    A = (rng.rand(N, N) < 0.45).astype(np.uint8)
    W = rng.randn(N, N, B)
    E_WWT = W[:, :, :, None] * W[:, :, None, :]
    mf_m = rng.dirichlet(2 * np.ones(C), size=N)
    mf_pi = rng.gamma(2.0, 1.0, C) + 0.5
    t1 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    t0 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    niw = np.zeros((C, C, 2 + B + B * B))          # (kappa, nu, mu[B], sigma[B*B]) of each cell's q
    wexp = [[None] * C for _ in range(C)]
    for a in range(C):
        for b in range(C):
            X = rng.randn(B, B)
            mu, sig, kap, nu = rng.randn(B), X.dot(X.T) + np.eye(B), 1.0 + rng.rand(), B + 2.0 + 3 * rng.rand()
            niw[a, b] = np.concatenate([[kap, nu], mu, sig.ravel()])
            wexp[a][b] = M.weight_expectations(mu, sig, kap, nu)
    Af = A.astype(float)
    out = dict(A=A, W=W, mf_m=mf_m, mf_pi=mf_pi, mf_tau1=t1, mf_tau0=t0, niw=niw)
    for lit in (1, 0):
        for tag, st in (("full", 1.0), ("step", 0.35)):
            m_ref, lp_ref = M.mf_update_c_loops(Af, W, E_WWT, mf_m, mf_pi, t1, t0, wexp, st, dists=RefDists,
                                                literal_reverse=bool(lit))
            out["m_%s_lit%d" % (tag, lit)] = m_ref
            out["lp_%s_lit%d" % (tag, lit)] = lp_ref
    m_noW, lp_noW = M.mf_update_c_loops(Af, W, E_WWT, mf_m, mf_pi, t1, t0, None, 1.0, dists=RefDists)
    out.update(m_noW=m_noW, lp_noW=lp_noW)
    nt1, nt0 = M.mf_update_p_loops(Af, mf_m, 1.0, 1.0, t1, t0, 0.6)
    out.update(p_tau1=nt1, p_tau0=nt0)
    nat0 = [[M.standard_to_natural(np.zeros(B), np.eye(B), 1.0, B + 2.0) for _ in range(C)] for _ in range(C)]
    nat = M.mf_update_weight_models_loops(Af, W, E_WWT, mf_m, nat0)
    out["w_nat"] = np.array(nat)
    # get_vlb (stochastic_block_model.py:536-563) without the weight models' terms, by the reference's OWN Discrete /
    # Beta / Dirichlet classes, statement by statement (scalar tau1 = tau0 = 1 and pi = 1, the class defaults)
    pi0 = 1.0 * np.ones(C)
    mf_m0 = mf_m.copy(); mf_m0[3] = np.array([0.0, 0.4, 0.6])       # an exact zero: Discrete.negentropy -> nan_to_num -> 0
    for tag, mm in (("vlb_nogauss", mf_m), ("vlb_nogauss_zero", mf_m0)):
        from scipy.special import psi as _psi
        vlb = 0
        E_ln_m = _psi(mf_pi) - _psi(mf_pi.sum())
        for n in range(N):
            vlb += dist.Discrete().negentropy(E_x=mm[n, :], E_ln_p=E_ln_m)
            with np.errstate(divide="ignore", invalid="ignore"):
                vlb -= dist.Discrete(mm[n, :]).negentropy()
        vlb += dist.Beta(1.0, 1.0).negentropy(E_ln_p=(_psi(t1) - _psi(t0 + t1)), E_ln_notp=(_psi(t0) - _psi(t0 + t1))).sum()
        vlb -= dist.Beta(t1, t0).negentropy().sum()
        vlb += dist.Dirichlet(pi0).negentropy(E_ln_g=E_ln_m)
        vlb -= dist.Dirichlet(mf_pi).negentropy()
        out[tag] = float(vlb)
    out["mf_m_zero"] = mf_m0
    np.savez(os.path.join(HERE, "mf_sbm_small.npz"), **out)
    print("mf_sbm_small.npz")


def make_mf_eigen():
    """mf_eigen_small.npz: the literal loop nests of eigenmodel.py:1090-1157 (oracle/mf_eigen.py `*_loops`), small N.
    Parity unpinned by the reference (eigenmodel.py cannot be imported); frozen so the oracle cannot drift."""
    from oracle import mf_eigen as M
    rng = np.random.RandomState(1357)
    N, D = 14, 3
    A = (rng.rand(N, N) < 0.3).astype(np.uint8)
    mu = rng.randn(N, D) * 0.7
    X = rng.randn(N, D, D) * 0.3
    Sg = X @ X.transpose(0, 2, 1) + 0.1 * np.eye(D)
    lam = rng.randn(D)
    Y = rng.randn(D, D) * 0.2
    SL = Y @ Y.T + 0.05 * np.eye(D)
    mu0, s0 = -0.5, 0.2
    Fg = rng.randn(N, D)
    Em = M.expected_mu(mu0, mu, lam)
    V = M.variance_mu(mu0, s0, mu, Sg, lam, SL)
    z = rng.randn(50, N, N)
    EO = M.expected_omega_mc(Em, V, z)
    llT = M.expected_llT(lam, SL)
    Af = A.astype(float)
    out = dict(A=A, mu=mu, Sig=Sg, lam=lam, SigL=SL, mu0=mu0, s0=s0, Fg=Fg, E_mu=Em, Var_mu=V, z=z, E_O=EO,
               E_O_quad=M.expected_omega_quad(Em, V))
    m0 = M.meanfieldupdate_mu0(Af, EO, mu, lam, -1.2, 0.9, mu0, s0, 0.8)
    out.update(mu0_new=m0[0], s0_new=m0[1], mu0_prec=m0[2], mu0_h=m0[3])
    for tr in (0, 1):
        for tag, st in (("full", 1.0), ("step", 0.4)):
            r = M.meanfieldupdate_F_loops(Af, EO, Fg, mu, Sg, mu0, lam, llT, 1.5, st, bool(tr))
            for nm, v in zip(("mu", "Sig", "prec", "h"), r):
                out["F_%s_%s_tr%d" % (nm, tag, tr)] = v
    r = M.meanfieldupdate_lmbda_loops(Af, EO, mu, Sg, mu0, 0.0, 2.0, lam, SL, 0.7)
    out.update(lam_new=r[0], SigL_new=r[1], lam_prec=r[2], lam_h=r[3])
    # get_vlb (eigenmodel.py:1262-1298) by the reference's OWN ScalarGaussian / Gaussian classes, statement by statement
    _, dist = _import_reference_internals()
    mu_mu_0, sigma_mu0, sigma_F, mu_lmbda, sigma_lmbda = -1.2, 0.9, 1.5, np.zeros(D), 2.0
    vlb = 0
    vlb += dist.ScalarGaussian(mu=mu_mu_0, sigmasq=sigma_mu0).negentropy(E_x=mu0, E_xsq=s0 + mu0 ** 2)
    vlb -= dist.ScalarGaussian(mu=mu0, sigmasq=s0).negentropy()
    for n in range(N):
        vlb += dist.Gaussian(mu=np.zeros(D), Sigma=sigma_F * np.eye(D)).negentropy(E_x=mu[n, :], E_xxT=Sg[n] + np.outer(mu[n], mu[n]))
        vlb -= dist.Gaussian(mu=mu[n, :], Sigma=Sg[n, :, :]).negentropy()
    out["vlb_lmbda_given"] = float(vlb)
    vlb += dist.Gaussian(mu=mu_lmbda, Sigma=sigma_lmbda * np.eye(D)).negentropy(E_x=lam, E_xxT=M.expected_llT(lam, SL))
    vlb -= dist.Gaussian(mu=lam, Sigma=SL).negentropy()
    out["vlb"] = float(vlb)
    np.savez(os.path.join(HERE, "mf_eigen_small.npz"), **out)
    print("mf_eigen_small.npz")


if __name__ == "__main__":
    make_ref_internals()
    make_sbm()
    make_eigen()
    make_latent()
    make_mf_sbm()
    make_mf_eigen()
