"""
make_ref_golden.py -- fixtures produced by EXECUTING THE REFERENCE'S OWN METHOD BODIES (tests/golden/ref_shim.py
loads /root/reference/graphistician/*.py under Python 3).  Build container only:

    python tests/golden/make_ref_golden.py            # writes tests/golden/ref_*.npz

Each fixture stores the inputs, the injected random draws and what the reference's code computed from them
(conditionals recorded at the third-party call sites + the resulting state).  tests/test_oracle_refpin.py asserts
oracle/ == these; the `-m gpu` tests compare the CUDA path with the same files.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_shim                                              # noqa: E402
from ref_shim import REC, patched_numpy                      # noqa: E402


def _save(name, out):
    np.savez(os.path.join(HERE, name), **out)
    print(name, len(out), "arrays")


def _spd(rng, B):
    X = rng.randn(B, B)
    return X.dot(X.T) + 0.5 * np.eye(B)


def _sbm_problem(rng, N, C, B):
    c = rng.randint(0, C, N)
    m = rng.dirichlet(3.0 * np.ones(C))
    p = rng.beta(1, 1, (C, C))
    mus = rng.randn(C, C, B)
    sig = np.array([[_spd(rng, B) for _ in range(C)] for _ in range(C)])
    A = (rng.rand(N, N) < p[np.ix_(c, c)]).astype(np.float64)
    W = rng.randn(N, N, B) + mus[np.ix_(c, c)]
    return A, W, c, m, p, mus, sig


def _stats_of(data, B):
    d = np.asarray(data, dtype=np.float64).reshape(-1, B)
    return d.shape[0], d.sum(0), d.T.dot(d)


# --------------------------------------------------------------------------------------------------
# old API: GaussianStochasticBlockModel (stochastic_block_model.py)
# --------------------------------------------------------------------------------------------------
def make_ref_sbm(R):
    rng = np.random.RandomState(20240)
    N, C, B = 14, 3, 2
    A, W, c, m, p, mus, sig = _sbm_problem(rng, N, C, B)
    net = R.Network(A, W)
    out = dict(A=A.astype(np.uint8), W=W, c=c, m=m, p=p, mus=mus, sigmas=sig)

    def fresh():
        np.random.seed(1)
        mdl = R.sbm.GaussianStochasticBlockModel(N, B, C=C, c=c.astype(int), m=m.copy(), p=p.copy())
        for a in range(C):
            for b in range(C):
                mdl.weight_models[a][b].mu = mus[a, b].copy()
                mdl.weight_models[a][b].sigma = sig[a, b].copy()
        return mdl

    # resample_c (stochastic_block_model.py:252-302) with injected uniforms: lp handed to sample_discrete_from_log, c
    u = rng.rand(N)
    mdl = fresh()
    REC.reset(); REC.uniforms = list(u)
    mdl.resample_c(net)
    out.update(u=u, c_new=mdl.c.copy(), lp=np.array(REC.lp))
    # resample_p (:232-250): np.random seeded -> p
    mdl = fresh()
    np.random.seed(77)
    mdl.resample_p([A])
    out.update(p_seed=77, p_new=mdl.p.copy())
    # the weight-model loop of resample (:221-226): the data the reference's masks hand to Gaussian.resample
    mdl = fresh()
    REC.reset()
    np.random.seed(78)
    for c1 in range(C):
        for c2 in range(C):
            mask = (mdl.c == c1)[:, None] * (mdl.c == c2)[None, :]
            mdl.weight_models[c1][c2].resample(R.Network(A * mask, W))
    st = [_stats_of(d, B) for d in REC.gauss_data]
    out.update(cell_n=np.array([s[0] for s in st]).reshape(C, C),
               cell_sum=np.array([s[1] for s in st]).reshape(C, C, B),
               cell_sumsq=np.array([s[2] for s in st]).reshape(C, C, B, B),
               w_seed=78, mus_new=np.array([[mdl.weight_models[a][b].mu for b in range(C)] for a in range(C)]),
               sigmas_new=np.array([[mdl.weight_models[a][b].sigma for b in range(C)] for a in range(C)]))
    # three full resample() sweeps (:206-230), every draw from np.random (seed 4242): the host draw ORDER
    mdl = fresh()
    REC.reset()
    np.random.seed(4242)
    tr_c, tr_m, tr_p, tr_mu, tr_sig = [], [], [], [], []
    for _ in range(3):
        mdl.resample(net)
        tr_c.append(mdl.c.copy()); tr_m.append(mdl.m.copy()); tr_p.append(mdl.p.copy())
        tr_mu.append(np.array([[mdl.weight_models[a][b].mu for b in range(C)] for a in range(C)]))
        tr_sig.append(np.array([[mdl.weight_models[a][b].sigma for b in range(C)] for a in range(C)]))
    out.update(sweep_seed=4242, sweep_c=np.array(tr_c), sweep_m=np.array(tr_m), sweep_p=np.array(tr_p),
               sweep_mus=np.array(tr_mu), sweep_sigmas=np.array(tr_sig), sweep_lp=np.array(REC.lp).reshape(3, N, C))
    out["log_prior"] = fresh().log_prior()
    _save("ref_sbm.npz", out)


# --------------------------------------------------------------------------------------------------
# new API: adjacency.py (BetaBernoulli, SBMAdjacency, LatentDistance)
# --------------------------------------------------------------------------------------------------
def make_ref_adjacency(R):
    rng = np.random.RandomState(20241)
    N, C = 15, 3
    A, _, c, m, p, _, _ = _sbm_problem(rng, N, C, 1)
    out = dict(A=A.astype(np.uint8), c=c, m=m, p=p)
    u = rng.rand(N)
    out["u"] = u
    for sp in (1, 0):
        np.random.seed(2)
        mdl = R.adjacency.SBMAdjacencyDistribution(N, C=C, c=c.astype(int), m=m.copy(), p=p.copy(),
                                                   special_case_self_conns=bool(sp))
        REC.reset(); REC.uniforms = list(u)
        mdl.resample_c(A)
        out["c_new_%d" % sp] = mdl.c.copy(); out["lp_%d" % sp] = np.array(REC.lp)
        mdl.c = c.astype(int).copy()
        np.random.seed(31 + sp)
        mdl.resample_p([A])
        out["p_new_%d" % sp] = mdl.p.copy()
        if sp:
            out["p_self_new"] = mdl.p_self
        out["ll_%d" % sp] = mdl.log_likelihood(A)
    # BetaBernoulli (adjacency.py:153-181)
    for sc in (0, 1):
        np.random.seed(5)
        kw = dict(tau1_self=1.5, tau0_self=0.5) if sc else {}
        bb = R.adjacency.BetaBernoulliAdjacencyDistribution(N, tau1=2.0, tau0=3.0, **kw)
        np.random.seed(41 + sc)
        bb.resample(A)
        out["bb_p_%d" % sc] = bb.p
        if sc:
            out["bb_p_self"] = bb.p_self
    _save("ref_adjacency.npz", out)


def make_ref_latent(R):
    rng = np.random.RandomState(20242)
    N, dim = 13, 2
    sigma = 2.0
    np.random.seed(3)
    mdl = R.adjacency.LatentDistanceAdjacencyDistribution(N, dim=dim, sigma=sigma, mu0=0.3, mu_self=-0.4)
    L0 = np.sqrt(sigma) * rng.randn(N, dim)
    mdl.L = L0.copy()
    A = (rng.rand(N, N) < mdl.P).astype(np.float64)
    out = dict(A=A.astype(np.uint8), L=L0, mu_0=0.3, mu_self=-0.4, sigma=sigma)
    # _hmc_log_probability (adjacency.py:244-275) and its derivative (complex step through the reference's own code)
    Ls = [L0, L0 + 0.3 * rng.randn(N, dim), 0.1 * rng.randn(N, dim)]
    pars = [(0.3, -0.4), (-1.1, 0.8), (0.0, 0.0)]
    lps, gLs, g0s, gss = [], [], [], []
    for Lq, (m0, ms) in zip(Ls, pars):
        lps.append(float(mdl._hmc_log_probability(Lq, m0, ms, A)))
        gLs.append(ref_shim._grad(lambda x: mdl._hmc_log_probability(x, m0, ms, A))(Lq))
        g0s.append(float(ref_shim._grad(lambda x: mdl._hmc_log_probability(Lq, x, ms, A))(np.array(m0))))
        gss.append(float(ref_shim._grad(lambda x: mdl._hmc_log_probability(Lq, m0, x, A))(np.array(ms))))
    out.update(eval_L=np.array(Ls), eval_mu0=np.array([q[0] for q in pars]), eval_muself=np.array([q[1] for q in pars]),
               eval_lp=np.array(lps), eval_gL=np.array(gLs), eval_gmu0=np.array(g0s), eval_gmuself=np.array(gss))
    out["ll"] = mdl.log_likelihood(A); out["log_prior"] = mdl.log_prior()
    # four full resample(A) sweeps (adjacency.py:356-421) with the HMC draws injected
    nsw = 4
    pL = rng.randn(nsw, N, dim); p0 = rng.randn(nsw); ps = rng.randn(nsw)
    lu = np.log(rng.rand(nsw, 3))
    REC.reset()
    REC.hmc_draws = []
    for s in range(nsw):
        REC.hmc_draws += [(pL[s], lu[s, 0]), (np.array(p0[s]), lu[s, 1]), (np.array(ps[s]), lu[s, 2])]
    tr = []
    for s in range(nsw):
        mdl.resample(A)
        tr.append((mdl.L.copy(), mdl.mu_0, mdl.mu_self, mdl._L_step_sz, mdl._L_accept_rate))
    out.update(hmc_pL=pL, hmc_p0=p0, hmc_ps=ps, hmc_logu=lu, sweep_L=np.array([t[0] for t in tr]),
               sweep_mu0=np.array([t[1] for t in tr]), sweep_muself=np.array([t[2] for t in tr]),
               sweep_step=np.array([t[3] for t in tr]), sweep_rate=np.array([t[4] for t in tr]))
    _save("ref_latent.npz", out)


# --------------------------------------------------------------------------------------------------
# new API: weights.py (NIW, SBM Gaussian weights)
# --------------------------------------------------------------------------------------------------
def make_ref_weights(R):
    rng = np.random.RandomState(20243)
    N, C, B = 13, 3, 2
    A, W, c, m, p, mus, sig = _sbm_problem(rng, N, C, B)
    out = dict(A=A.astype(np.uint8), W=W, c=c, m=m, mus=mus, sigmas=sig)
    # NIWGaussianWeightDistribution.resample (weights.py:154-162)
    np.random.seed(6)
    niw = R.weights.NIWGaussianWeightDistribution(N, B=B)
    REC.reset()
    np.random.seed(61)
    niw.resample((A, W))
    s_off, s_on = _stats_of(REC.gauss_data[0], B), _stats_of(REC.gauss_data[1], B)
    out.update(niw_seed=61, niw_off_n=s_off[0], niw_off_sum=s_off[1], niw_off_sumsq=s_off[2], niw_on_n=s_on[0],
               niw_on_sum=s_on[1], niw_on_sumsq=s_on[2], niw_mu=niw._gaussian.mu, niw_sigma=niw._gaussian.sigma,
               niw_self_mu=niw._self_gaussian.mu, niw_self_sigma=niw._self_gaussian.sigma,
               niw_ll=niw.log_likelihood((A, W)))
    u = rng.rand(N)
    out["u"] = u
    for sp in (1, 0):
        np.random.seed(7)
        mdl = R.weights.SBMGaussianWeightDistribution(N, B=B, C=C, special_case_self_conns=bool(sp))
        mdl.c = c.astype(int).copy(); mdl.m = m.copy()
        for a in range(C):
            for b in range(C):
                mdl._gaussians[a][b].mu = mus[a, b].copy(); mdl._gaussians[a][b].sigma = sig[a, b].copy()
        if sp:
            mdl._self_gaussian.mu = mus[0, 1].copy(); mdl._self_gaussian.sigma = sig[1, 0].copy()
        out["ll_%d" % sp] = mdl.log_likelihood((A, W))
        REC.reset(); REC.uniforms = list(u)
        mdl.resample_c(A, W)
        out["c_new_%d" % sp] = mdl.c.copy(); out["lp_%d" % sp] = np.array(REC.lp)
        mdl.c = c.astype(int).copy()
        REC.reset()
        np.random.seed(71 + sp)
        mdl.resample_mu_and_Sig(A, W)
        st = [_stats_of(d, B) for d in REC.gauss_data]
        out["cell_n_%d" % sp] = np.array([s[0] for s in st[:C * C]]).reshape(C, C)
        out["cell_sum_%d" % sp] = np.array([s[1] for s in st[:C * C]]).reshape(C, C, B)
        out["cell_sumsq_%d" % sp] = np.array([s[2] for s in st[:C * C]]).reshape(C, C, B, B)
        out["mus_new_%d" % sp] = mdl._Mu.copy(); out["sigmas_new_%d" % sp] = mdl._Sigma.copy()
        if sp:
            out["self_n"], out["self_sum"], out["self_sumsq"] = st[C * C]
            out["self_mu_new"] = mdl._self_gaussian.mu; out["self_sigma_new"] = mdl._self_gaussian.sigma
    _save("ref_weights.npz", out)


# --------------------------------------------------------------------------------------------------
# eigenmodel.py: logistic / probit Gibbs
# --------------------------------------------------------------------------------------------------
def make_ref_eigen(R):
    rng = np.random.RandomState(20244)
    N, D = 11, 3
    E = R.eigenmodel
    out = {}
    for link in ("logistic", "probit"):
        for given in (0, 1):
            tag = "%s_g%d" % (link, given)
            F0 = rng.randn(N, D); lm0 = 0.7 * rng.randn(D) + 0.5; mu00 = -0.6
            REC.reset()
            REC.omegas = np.ones(N * N)                      # the constructor's own resample() call needs a value
            np.random.seed(8)
            cls = E.LogisticEigenmodel if link == "logistic" else E.ProbitEigenmodel
            kw = dict(lmbda=lm0.copy()) if given else dict(mu_lmbda=0.2, sigma_lmbda=1.7)
            mdl = cls(N, D, p=0.3, sigma_mu0=0.9, sigma_F=1.3, **kw)
            mdl.F = F0.copy(); mdl.mu_0 = mu00; mdl.lmbda = lm0.copy()
            A = (rng.rand(N, N) < mdl.P).astype(np.float64)
            out.update({tag + "_A": A.astype(np.uint8), tag + "_F": F0, tag + "_lmbda": lm0, tag + "_mu_0": mu00,
                        tag + "_mu_mu_0": mdl.mu_mu_0, tag + "_ll": mdl.log_likelihood(A), tag + "_log_prior": mdl.log_prior()})
            zF = rng.randn(N, D); zmu = rng.randn(); zl = rng.randn(D)
            out.update({tag + "_zF": zF, tag + "_zmu": zmu, tag + "_zl": zl})
            REC.reset()
            REC.mvn_z = [z for z in zF] + ([] if given else [zl])
            REC.normal_z = [zmu]
            if link == "logistic":
                O = rng.gamma(2.0, 0.1, (N, N)) + 0.02
                REC.omegas = O.ravel()
                out[tag + "_O"] = O
                with patched_numpy():
                    mdl.resample(R.Network(A, None))
            else:
                np.random.seed(81)
                u = np.random.rand(N, N)                     # the uniforms sample_truncnorm (internals/utils.py:66) will draw
                np.random.seed(81)
                Zs = mdl._resample_Z([A])
                out[tag + "_u"] = u; out[tag + "_Z"] = Zs[0]
                with patched_numpy():
                    mdl._resample_F(Zs)
                    mdl._resample_mu_0(Zs)
                    if not given:
                        mdl._resample_lmbda(Zs)
            out[tag + "_prec"] = np.array(REC.inv_in[:N])
            out[tag + "_mean"] = np.array([mv[0] for mv in REC.mvn[:N]])
            out[tag + "_F_new"] = mdl.F.copy()
            out[tag + "_mu0_mean"], out[tag + "_mu0_std"] = REC.normal[0]
            out[tag + "_mu_0_new"] = mdl.mu_0
            if not given:
                out[tag + "_lam_prec"] = REC.inv_in[N]
                out[tag + "_lam_mean"] = REC.mvn[N][0]
                out[tag + "_lmbda_new"] = mdl.lmbda.copy()
    _save("ref_eigen.npz", out)


# --------------------------------------------------------------------------------------------------
# internals/graph_models.py: tempered ER / SBM steps used by AIS
# --------------------------------------------------------------------------------------------------
def make_ref_ais(R):
    rng = np.random.RandomState(20245)
    G = R.graph_models
    N, Rb = 12, 3
    z = rng.randint(0, Rb, N)
    Bm = rng.beta(1, 1, (Rb, Rb)); pi = rng.dirichlet(3.0 * np.ones(Rb))
    A = (rng.rand(N, N) < Bm[np.ix_(z, z)]).astype(np.float64)
    out = dict(A=A.astype(np.uint8), z=z, B=Bm, pi=pi)
    sbm = G.StochasticBlockModel(Rb, 1.0, 1.0, 10.0 * np.ones(Rb))
    er = G.ErdosRenyiNetwork(x=(1.0, 1.0))
    betas = np.array([0.0, 0.37, 1.0])
    out["betas"] = betas
    u = rng.rand(N)
    out["u"] = u
    for bi, beta in enumerate(betas):
        f = list(z)
        REC.reset(); REC.uniforms = list(u)
        for n in range(N):
            f[n] = sbm.sample_f((Bm, pi), (n, A, f), beta=beta)
        out["z_new_%d" % bi] = np.array(f); out["lp_%d" % bi] = np.array(REC.lp)
        np.random.seed(90 + bi)
        Bn, pin = sbm.sample_theta((A, list(z)), beta=beta)
        out["B_new_%d" % bi] = np.asarray(Bn); out["pi_new_%d" % bi] = pin
        np.random.seed(95 + bi)
        out["rho_new_%d" % bi] = er.sample_theta((A, None), beta=beta)
        out["sbm_logpr_%d" % bi] = sbm.logpr(A, list(z), (Bm, pi), beta=beta)
        out["er_logpr_%d" % bi] = er.logpr(A, [None] * N, 0.31, beta=beta)
    # collapsed_sample_f (:299-335; unused by the reference's drivers): it draws inside its r loop, so the recorder sees R
    # vectors per call -- the last one is the complete ln_pi_post, and the label returned is the draw made from it
    cn = np.array([0, 5, N - 1])
    uc = rng.rand(len(cn), Rb)
    out["coll_nodes"], out["coll_u"] = cn, uc
    for k, n in enumerate(cn):
        REC.reset(); REC.uniforms = list(uc[k])
        zn = sbm.collapsed_sample_f((Bm, pi), int(n), A, list(z))
        out["coll_lp_%d" % k] = np.array(REC.lp[-1]); out["coll_z_%d" % k] = int(zn)
        assert len(REC.lp) == Rb
    _save("ref_ais.npz", out)


# --------------------------------------------------------------------------------------------------
# mean-field SBM: the WHOLE loop nests of stochastic_block_model.py:330-445 on the reference's own class
# --------------------------------------------------------------------------------------------------
def make_ref_mf_sbm(R):
    rng = np.random.RandomState(20246)
    N, C, B = 9, 3, 2
    A = (rng.rand(N, N) < 0.45).astype(np.float64)
    W = rng.randn(N, N, B)
    net = R.Network(A, W)
    Gn = R.distributions.Gaussian

    class Model(R.sbm.GaussianStochasticBlockModel):
        def _expected_log_likelihood_W(self, network, n1, cn1, n2, cn2):
            # stochastic_block_model.py:631-644 hands a TUPLE to GaussianWeights.expected_log_likelihood, which reads
            # network.E_W (internals/weights.py:177): TypeError in the reference.  The per-pair quantity that method
            # sums (internals/weights.py:190-196) is evaluated here with the reference's own Gaussian().negentropy.
            wm = self.weight_models[cn1][cn2]
            return Gn().negentropy(E_x=network.E_W[n1, n2, :], E_xxT=network.E_WWT[n1, n2, :, :],
                                   E_mu=wm.mf_expected_mu(), E_mumuT=wm.mf_expected_mumuT(),
                                   E_Sigma_inv=wm.mf_expected_Sigma_inv(),
                                   E_logdet_Sigma=wm.mf_expected_logdet_Sigma()).sum()

    np.random.seed(9)
    mdl = Model(N, B, C=C, pi=1.0, tau0=1.0, tau1=1.0)
    mdl.mf_m = rng.dirichlet(2.0 * np.ones(C), size=N)
    mdl.mf_pi = rng.gamma(2.0, 1.0, C) + 0.5
    mdl.mf_tau1 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    mdl.mf_tau0 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    niw = np.zeros((C, C, 2 + B + B * B))
    for a in range(C):
        for b in range(C):
            wm = mdl.weight_models[a][b]
            wm.mu_mf, wm.sigma_mf, wm.kappa_mf, wm.nu_mf = rng.randn(B), _spd(rng, B) + 0.5 * np.eye(B), 1.0 + rng.rand(), B + 2.0 + 3 * rng.rand()
            niw[a, b] = np.concatenate([[wm.kappa_mf, wm.nu_mf], wm.mu_mf, wm.sigma_mf.ravel()])
    out = dict(A=A.astype(np.uint8), W=W, mf_m=mdl.mf_m.copy(), mf_pi=mdl.mf_pi.copy(), mf_tau1=mdl.mf_tau1.copy(),
               mf_tau0=mdl.mf_tau0.copy(), niw=niw)
    import copy
    for tag, st in (("full", 1.0), ("step", 0.35)):
        m2 = copy.deepcopy(mdl)
        m2.mf_update_c(net, stepsize=st)
        out["c_m_" + tag] = m2.mf_m.copy()
        m2 = copy.deepcopy(mdl)
        m2.mf_update_p(net, stepsize=st)
        out["p_tau1_" + tag] = m2.mf_tau1.copy(); out["p_tau0_" + tag] = m2.mf_tau0.copy()
        m2 = copy.deepcopy(mdl)
        m2.mf_update_m(stepsize=st)
        out["m_pi_" + tag] = m2.mf_pi.copy()
    # one whole meanfieldupdate (:330-344): p, c, m, then the weight models' natural parameters
    m2 = copy.deepcopy(mdl)
    m2.meanfieldupdate(net)
    out.update(mfu_m=m2.mf_m.copy(), mfu_pi=m2.mf_pi.copy(), mfu_tau1=m2.mf_tau1.copy(), mfu_tau0=m2.mf_tau0.copy(),
               mfu_nat=np.array([[m2.weight_models[a][b].mf_natural_hypparam for b in range(C)] for a in range(C)]))
    out["vlb"] = m2.get_vlb()
    _save("ref_mf_sbm.npz", out)


# --------------------------------------------------------------------------------------------------
# mean-field eigenmodels (logistic :969-1311, probit :382-770)
# --------------------------------------------------------------------------------------------------
def make_ref_mf_eigen(R):
    rng = np.random.RandomState(20247)
    N, D = 10, 3
    E = R.eigenmodel
    out = {}
    for link in ("logistic", "probit"):
        for given in (0, 1):
            tag = "%s_g%d" % (link, given)
            REC.reset(); REC.omegas = np.ones(N * N)
            np.random.seed(10)
            cls = E.LogisticEigenmodel if link == "logistic" else E.ProbitEigenmodel
            lm0 = 0.7 * rng.randn(D) + 0.5
            kw = dict(lmbda=lm0.copy()) if given else dict(mu_lmbda=0.2, sigma_lmbda=1.7)
            mdl = cls(N, D, p=0.3, sigma_mu0=0.9, sigma_F=1.3, **kw)
            A = (rng.rand(N, N) < 0.35).astype(np.float64)
            mdl.F = rng.randn(N, D)
            mdl.mf_mu_F = 0.7 * rng.randn(N, D)
            mdl.mf_Sigma_F = np.array([_spd(rng, D) * 0.3 for _ in range(N)])
            mdl.mf_mu_mu0, mdl.mf_sigma_mu0 = -0.5, 0.2
            if not given:
                mdl.mf_mu_lmbda = lm0.copy(); mdl.mf_Sigma_lmbda = _spd(rng, D) * 0.1
            st0 = dict(A=A.astype(np.uint8), F=mdl.F.copy(), mf_mu_F=mdl.mf_mu_F.copy(), mf_Sigma_F=mdl.mf_Sigma_F.copy(),
                       mf_mu_mu0=mdl.mf_mu_mu0, mf_sigma_mu0=mdl.mf_sigma_mu0, lmbda=lm0, mu_mu_0=mdl.mu_mu_0)
            if not given:
                st0.update(mf_mu_lmbda=mdl.mf_mu_lmbda.copy(), mf_Sigma_lmbda=mdl.mf_Sigma_lmbda.copy())
            out.update({tag + "_" + k: v for k, v in st0.items()})
            out[tag + "_E_mu"] = mdl.mf_expected_mu(); out[tag + "_E_musq"] = mdl.mf_expected_musq()
            net = R.Network(A, None)
            for stag, st in (("full", 1.0), ("step", 0.4)):
                import copy
                m2 = copy.deepcopy(mdl)
                REC.reset()
                np.random.seed(101)
                with patched_numpy():
                    if link == "logistic":
                        E_O = m2._meanfieldupdate_Omega()
                        out[tag + "_E_O"] = E_O
                        m2._meanfieldupdate_mu0(A, E_O, stepsize=st)
                        out["%s_%s_mu0" % (tag, stag)] = np.array([m2.mf_mu_mu0, m2.mf_sigma_mu0])
                        m2._meanfieldupdate_F(A, E_O, stepsize=st)
                        out["%s_%s_F_prec" % (tag, stag)] = np.array(REC.inv_in[:N])
                        out["%s_%s_mu_F" % (tag, stag)] = m2.mf_mu_F.copy()
                        out["%s_%s_Sigma_F" % (tag, stag)] = m2.mf_Sigma_F.copy()
                        if not given:
                            m2._meanfieldupdate_lmbda(A, E_O, stepsize=st)
                            out["%s_%s_lam_prec" % (tag, stag)] = REC.inv_in[N]
                            out["%s_%s_mu_lmbda" % (tag, stag)] = m2.mf_mu_lmbda.copy()
                            out["%s_%s_Sigma_lmbda" % (tag, stag)] = m2.mf_Sigma_lmbda.copy()
                    else:
                        if stag == "step":                   # the probit updates take no step size (eigenmodel.py:570-674)
                            continue
                        m2.mf_A = A
                        m2._meanfieldupdate_Z()
                        out[tag + "_mu_Z"] = m2.mf_mu_Z.copy()
                        out[tag + "_E_Z"] = m2.mf_expected_Z()
                        out[tag + "_E_Zsq"] = m2.mf_expected_Zsq()
                        m2._meanfieldupdate_mu0()
                        out["%s_%s_mu0" % (tag, stag)] = np.array([m2.mf_mu_mu0, m2.mf_sigma_mu0])
                        m2._meanfieldupdate_F()
                        out["%s_%s_F_prec" % (tag, stag)] = np.array(REC.inv_in[:N])
                        out["%s_%s_mu_F" % (tag, stag)] = m2.mf_mu_F.copy()
                        out["%s_%s_Sigma_F" % (tag, stag)] = m2.mf_Sigma_F.copy()
                        if not given:
                            m2._meanfieldupdate_lmbda()
                            out["%s_%s_lam_prec" % (tag, stag)] = REC.inv_in[N]
                            out["%s_%s_mu_lmbda" % (tag, stag)] = m2.mf_mu_lmbda.copy()
                            out["%s_%s_Sigma_lmbda" % (tag, stag)] = m2.mf_Sigma_lmbda.copy()
                        out[tag + "_ell"] = m2.expected_log_likelihood(A)
                try:
                    out["%s_%s_vlb" % (tag, stag)] = float(m2.get_vlb())
                except Exception as e:                       # recorded, not hidden
                    out["%s_%s_vlb_error" % (tag, stag)] = np.array(repr(e))
    _save("ref_mf_eigen.npz", out)


# --------------------------------------------------------------------------------------------------
# predictive path (abstractions.py:141-182, 289-357; adjacency.py:102,183,291,562; weights.py:59,144,375)
# --------------------------------------------------------------------------------------------------
def make_ref_predictive(R):
    rng = np.random.RandomState(20249)
    N, C, B, M = 12, 3, 2, 7
    A, W, c, m, p, mus, sig = _sbm_problem(rng, N, C, B)
    Arow = (rng.rand(N + 1) < 0.4); Acol = (rng.rand(N + 1) < 0.4); Acol[-1] = Arow[-1]
    Wrow = rng.randn(N + 1, B); Wcol = rng.randn(N + 1, B); Wcol[-1] = Wrow[-1]
    out = dict(c=c, m=m, p=p, mus=mus, sigmas=sig, Arow=Arow, Acol=Acol, Wrow=Wrow, Wcol=Wcol, M=M)

    def adj_record(tag, mdl):
        np.random.seed(101)
        pr, pc = mdl.sample_predictive_parameters()
        out[tag + "_Prow"], out[tag + "_Pcol"] = np.array(pr, dtype=float), np.array(pc, dtype=float)
        np.random.seed(102)
        ar, ac = mdl.sample_predictive_distribution()
        out[tag + "_Arow"], out[tag + "_Acol"] = np.array(ar), np.array(ac)
        np.random.seed(103)
        out[tag + "_pll"] = mdl.approx_predictive_ll(Arow, Acol, M=M)

    def w_record(tag, mdl):
        np.random.seed(111)
        mr, mc, lr, lc = mdl.sample_predictive_parameters()
        out[tag + "_Murow"], out[tag + "_Mucol"], out[tag + "_Lrow"], out[tag + "_Lcol"] = \
            [np.array(x, dtype=float) for x in (mr, mc, lr, lc)]
        np.random.seed(112)
        try:                                             # abstractions.py:297-298 asserts on arrays: ValueError for B > 1
            wr, wc = mdl.sample_predictive_distribution()
            out[tag + "_Wrow"], out[tag + "_Wcol"] = wr, wc
        except ValueError as e:
            out[tag + "_dist_error"] = np.array(repr(e))
        np.random.seed(113)
        out[tag + "_pll"] = mdl.approx_predictive_ll(Arow, Acol, Wrow, Wcol, M=M)

    adj_record("bern", R.adjacency.BernoulliAdjacencyDistribution(N, p=0.3))
    for sc in (0, 1):
        np.random.seed(5)
        kw = dict(tau1_self=1.5, tau0_self=0.5) if sc else {}
        bb = R.adjacency.BetaBernoulliAdjacencyDistribution(N, tau1=2.0, tau0=3.0, **kw)
        bb.p = 0.35
        if sc:
            bb.p_self = 0.8
        adj_record("bb%d" % sc, bb)
    for sp in (1, 0):
        np.random.seed(2)
        mdl = R.adjacency.SBMAdjacencyDistribution(N, C=C, c=c.astype(int), m=m.copy(), p=p.copy(),
                                                   special_case_self_conns=bool(sp))
        adj_record("sbm%d" % sp, mdl)
    np.random.seed(3)
    ld = R.adjacency.LatentDistanceAdjacencyDistribution(N, dim=2, sigma=1.5, mu0=0.3, mu_self=-0.4)
    ld.L = rng.randn(N, 2)
    out["ld_L"] = ld.L.copy(); out["ld_sigma"] = ld.sigma; out["ld_mu0"] = ld.mu_0; out["ld_mu_self"] = ld.mu_self
    adj_record("ld", ld)

    np.random.seed(6)
    fx = R.weights.FixedGaussianWeightDistribution(N, B, mus[0, 1].copy(), sig[0, 1].copy())
    w_record("fixed", fx)
    np.random.seed(6)
    niw = R.weights.NIWGaussianWeightDistribution(N, B=B)
    niw._gaussian.mu = mus[0, 0].copy(); niw._gaussian.sigma = sig[0, 0].copy()
    niw._self_gaussian.mu = mus[1, 1].copy(); niw._self_gaussian.sigma = sig[1, 1].copy()
    w_record("niw", niw)
    for sp in (1, 0):
        np.random.seed(7)
        mdl = R.weights.SBMGaussianWeightDistribution(N, B=B, C=C, special_case_self_conns=bool(sp))
        mdl.c = c.astype(int).copy(); mdl.m = m.copy()
        for a in range(C):
            for b in range(C):
                mdl._gaussians[a][b].mu = mus[a, b].copy(); mdl._gaussians[a][b].sigma = sig[a, b].copy()
        if sp:
            mdl._self_gaussian.mu = mus[0, 1].copy(); mdl._self_gaussian.sigma = sig[1, 0].copy()
        w_record("sbmw%d" % sp, mdl)
    # B = 1: the only case in which the reference's sample_predictive_distribution runs
    np.random.seed(8)
    mdl = R.weights.SBMGaussianWeightDistribution(N, B=1, C=C, special_case_self_conns=True)
    mdl.c = c.astype(int).copy(); mdl.m = m.copy()
    for a in range(C):
        for b in range(C):
            mdl._gaussians[a][b].mu = mus[a, b, :1].copy(); mdl._gaussians[a][b].sigma = sig[a, b, :1, :1].copy()
    mdl._self_gaussian.mu = mus[0, 1, :1].copy(); mdl._self_gaussian.sigma = sig[1, 0, :1, :1].copy()
    np.random.seed(112)
    wr, wc = mdl.sample_predictive_distribution()
    out["sbmw_b1_Wrow"], out["sbmw_b1_Wcol"] = wr, wc
    _save("ref_predictive.npz", out)


# --------------------------------------------------------------------------------------------------
# weights.py:501-901: shared-covariance SBM weights, latent-distance Gaussian weights
# --------------------------------------------------------------------------------------------------
def make_ref_latent_weights(R):
    rng = np.random.RandomState(20250)
    N, dim = 12, 2
    np.random.seed(9)
    mdl = R.weights.LatentDistanceGaussianWeightDistribution(N, B=1, dim=dim, b=-1.5, eta=0.8)
    L0 = np.sqrt(0.8) * rng.randn(N, dim)
    mdl.L = L0.copy()
    mdl.cov.sigma = np.array([[0.6]])
    mdl._self_gaussian.mu = np.array([0.4]); mdl._self_gaussian.sigma = np.array([[1.3]])
    A = (rng.rand(N, N) < 0.6).astype(np.float64)
    W = mdl.Mu + np.sqrt(0.6) * rng.randn(N, N, 1)
    out = dict(A=A.astype(np.uint8), W=W, L=L0, b=-1.5, eta=0.8, sigma_w=0.6, self_mu=0.4, self_sigma=1.3)
    Ls = [L0, L0 + 0.3 * rng.randn(N, dim), 0.1 * rng.randn(N, dim)]
    bs = [-1.5, 0.7, 0.0]
    lps, gLs, gbs = [], [], []
    for Lq, bq in zip(Ls, bs):
        lps.append(float(mdl._hmc_log_probability(Lq, bq, A, W)))
        gLs.append(ref_shim._grad(lambda x: mdl._hmc_log_probability(x, bq, A, W))(Lq))
        gbs.append(float(ref_shim._grad(lambda x: mdl._hmc_log_probability(Lq, x, A, W))(np.array(bq))))
    out.update(eval_L=np.array(Ls), eval_b=np.array(bs), eval_lp=np.array(lps), eval_gL=np.array(gLs), eval_gb=np.array(gbs))
    out["ll"] = mdl.log_likelihood((A, W)); out["log_prior"] = mdl.log_prior()
    # _resample_L (:765-780) three times with the HMC draws injected, then _resample_b (:838-860) with np.random seeded
    nsw = 3
    pL = rng.randn(nsw, N, dim); lu = np.log(rng.rand(nsw))
    REC.reset(); REC.hmc_draws = [(pL[s], lu[s]) for s in range(nsw)]
    tr = []
    for s in range(nsw):
        mdl._resample_L(A, W)
        tr.append(mdl.L.copy())
    out.update(hmc_pL=pL, hmc_logu=lu, hmc_L=np.array(tr))
    np.random.seed(123)
    mdl._resample_b(A, W)
    out["b_seed"] = 123; out["b_new"] = mdl.b
    # _resample_cov (:863-867): the residuals handed to GaussianFixedMean.resample
    mask = (~np.eye(N, dtype=bool)) & A.astype(bool)
    resid = (W - mdl.Mu)[mask]
    out["cov_n"] = resid.shape[0]; out["cov_S"] = float((resid ** 2).sum())
    Mr, Mc, Lr, Lc = None, None, None, None
    np.random.seed(321)
    Mr, Mc, Lr, Lc = mdl.sample_predictive_parameters()
    out.update(pred_seed=321, pred_Murow=Mr, pred_Mucol=Mc, pred_Lrow=Lr, pred_Lcol=Lc)
    _save("ref_latent_weights.npz", out)


def make_ref_shared_cov(R):
    rng = np.random.RandomState(20251)
    N, C, B = 13, 3, 2
    A, W, c, m, p, mus, sig = _sbm_problem(rng, N, C, B)
    out = dict(A=A.astype(np.uint8), W=W, c=c, m=m, mus=mus)
    shared = _spd(rng, B)
    out["shared"] = shared
    for sp in (1, 0):
        np.random.seed(7)
        mdl = R.weights.SBMGaussianWeightSharedCov(N, B=B, C=C, special_case_self_conns=bool(sp))
        mdl.c = c.astype(int).copy(); mdl.m = m.copy()
        mdl._cov_model.sigma = shared.copy()
        for a in range(C):
            for b in range(C):
                mdl._gaussians[a][b].mu = mus[a, b].copy(); mdl._gaussians[a][b].sigma = shared.copy()
        out["ll_%d" % sp] = mdl.log_likelihood((A, W))
        out["log_prior_%d" % sp] = mdl.log_prior()
        with patched_numpy() as rec:
            REC.reset()
            np.random.seed(71 + sp)
            mdl.resample_mu_and_Sig(A, W)
        out["mvn_mean_%d" % sp] = np.array([q[0] for q in REC.mvn[:C * C]])
        out["mvn_cov_%d" % sp] = np.array([q[1] for q in REC.mvn[:C * C]])
        out["mus_new_%d" % sp] = np.array([[mdl._gaussians[a][b].mu for b in range(C)] for a in range(C)])
        # what the covariance update saw: residuals of the off-diagonal edges around the new means
        Aoff = A.astype(bool) & ~np.eye(N, dtype=bool)
        resid = (W - mdl.Mu)[Aoff]
        out["cov_n_%d" % sp] = resid.shape[0]; out["cov_S_%d" % sp] = resid.T.dot(resid)
    _save("ref_shared_cov.npz", out)


# --------------------------------------------------------------------------------------------------
# internals/weights.py:174-198: GaussianWeights.expected_log_likelihood, the N^2 negentropy loop, on the reference's class
# --------------------------------------------------------------------------------------------------
def make_ref_gw_ell(R):
    rng = np.random.RandomState(20251)
    N, B = 6, 2
    A = (rng.rand(N, N) < 0.5).astype(np.float64)
    W = rng.randn(N, N, B)
    net = R.Network(A, W)
    out = dict(A=A.astype(np.uint8), W=W)
    X = rng.randn(B, B)
    hyp = [dict(), dict(mu_0=np.array([0.4, -1.1]), Sigma_0=X.dot(X.T) + 0.7 * np.eye(B), nu_0=6.5, kappa_0=2.5)]
    for k, h in enumerate(hyp):
        np.random.seed(70 + k)                 # pybasicbayes' Gaussian draws its state (and the shim's stand-in its q) at construction
        gw = R.iweights.GaussianWeights(B, **h)
        if k == 1:                             # a q away from the prior: one mean-field update on the network
            gw.meanfieldupdate(net, np.ones((N, N)))
        out["ell_%d" % k] = float(gw.expected_log_likelihood(net))
        out["q_%d_mu" % k], out["q_%d_sigma" % k] = np.array(gw.mu_mf), np.array(gw.sigma_mf)
        out["q_%d_kappa" % k], out["q_%d_nu" % k] = float(gw.kappa_mf), float(gw.nu_mf)
        for name, v in h.items():
            out["hyp_%d_%s" % (k, name)] = np.asarray(v)
    _save("ref_gw_ell.npz", out)


if __name__ == "__main__":
    R = ref_shim.load_reference()
    which = sys.argv[1:] or ["sbm", "adjacency", "latent", "weights", "eigen", "ais", "mf_sbm", "mf_eigen", "predictive", "latent_weights", "shared_cov", "gw_ell"]
    for w in which:
        globals()["make_ref_" + w](R)
