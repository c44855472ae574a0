"""oracle/ == the REFERENCE'S OWN CODE.  tests/golden/ref_*.npz were written by tests/golden/make_ref_golden.py, which
loads /root/reference/graphistician/*.py under Python 3 (tests/golden/ref_shim.py) and executes the method bodies
themselves on small fixed-seed problems, recording the conditionals they compute.  These tests pin the oracle's literal
transcriptions (T0) and vectorised forms (T1) against those outputs."""
import os

import numpy as np
import pytest

from oracle import eigenmodel as eg
from oracle import latent as lt
from oracle import mf_eigen, mf_sbm, niw, sbm

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


# ---------------------------------------------------------------------------------------------- old-API SBM
def test_gibbs_sbm_resample_c_is_the_references():
    """stochastic_block_model.py:252-302 executed by the reference itself: lp of every node, labels."""
    g = load("ref_sbm.npz")
    A, W = g["A"].astype(float), g["W"]
    for fn in (sbm.gibbs_sbm_resample_c_loops,
               lambda *a: sbm.sbm_assign_sweep(*a, self_mode=sbm.SELF_OLD)):
        c, lp = fn(A, W, g["c"], g["m"], g["p"], g["mus"], g["sigmas"], g["u"])
        assert np.array_equal(c, g["c_new"])
        assert np.allclose(lp, g["lp"], rtol=1e-11, atol=1e-10)


def test_gibbs_sbm_cell_statistics_and_draws_are_the_references():
    """The data the reference's masks hand to Gaussian.resample (:221-226) and resample_p's Beta draws (:232-250)."""
    g = load("ref_sbm.npz")
    A, W, c = g["A"].astype(float), g["W"], g["c"]
    K = len(g["m"])
    for st in (sbm.block_pair_stats_loops(A, W, c, K), sbm.block_pair_stats(A, W, c, K)):
        n_edge, n_noedge, sum_w, sum_ww = sbm.cell_stats_with_diag(st, include_diag=True)
        assert np.array_equal(n_edge, g["cell_n"])
        assert np.allclose(sum_w, g["cell_sum"], rtol=1e-12, atol=1e-12)
        assert np.allclose(sum_ww, g["cell_sumsq"], rtol=1e-12, atol=1e-12)
    rs = np.random.RandomState(int(g["p_seed"]))
    assert np.allclose(sbm.resample_p(n_edge, n_noedge, 1.0, 1.0, rng=rs), g["p_new"], rtol=1e-13)
    rs = np.random.RandomState(int(g["w_seed"]))
    B = W.shape[2]
    mus, sig = sbm.resample_gaussians(n_edge, sum_w, sum_ww, (np.zeros(B), np.eye(B), 1.0, B + 2), rng=rs)
    assert np.allclose(mus, g["mus_new"], rtol=1e-10) and np.allclose(sig, g["sigmas_new"], rtol=1e-10)


def test_gibbs_sbm_full_sweeps_follow_the_references_draw_order():
    """Three whole resample() calls of the reference (:206-230) from np.random.seed: the oracle's sweep, fed the same
    legacy stream, reproduces every state (weight models -> p -> c -> m, one uniform per node)."""
    g = load("ref_sbm.npz")
    A, W = g["A"].astype(float), g["W"]
    B, K, N = W.shape[2], len(g["m"]), A.shape[0]
    state = dict(c=g["c"].copy(), m=g["m"].copy(), p=g["p"].copy(), mus=g["mus"].copy(), sigmas=g["sigmas"].copy())
    hyp = dict(pi=np.ones(K), tau0=1.0, tau1=1.0, niw=(np.zeros(B), np.eye(B), 1.0, B + 2))
    rs = np.random.RandomState(int(g["sweep_seed"]))

    for s in range(3):
        st = sbm.block_pair_stats(A, W, state["c"], K)
        n_edge, n_noedge, sum_w, sum_ww = sbm.cell_stats_with_diag(st, True)
        mus, sig = sbm.resample_gaussians(n_edge, sum_w, sum_ww, hyp["niw"], rs)
        p = sbm.resample_p(n_edge, n_noedge, 1.0, 1.0, rng=rs)
        u = rs.random_sample(N)
        c, lp = sbm.sbm_assign_sweep(A, W, state["c"], state["m"], p, mus, sig, u, sbm.SELF_OLD)
        m = sbm.resample_m(c, hyp["pi"], rs)
        state = dict(c=c, m=m, p=p, mus=mus, sigmas=sig)
        assert np.array_equal(c, g["sweep_c"][s]), s
        assert np.allclose(lp, g["sweep_lp"][s], rtol=1e-9, atol=1e-9)
        assert np.allclose(p, g["sweep_p"][s]) and np.allclose(m, g["sweep_m"][s])
        assert np.allclose(mus, g["sweep_mus"][s]) and np.allclose(sig, g["sweep_sigmas"][s])


# ---------------------------------------------------------------------------------------------- new-API adjacency / weights
@pytest.mark.parametrize("sp", [1, 0])
def test_sbm_adjacency_resample_is_the_references(sp):
    g = load("ref_adjacency.npz")
    A = g["A"].astype(float)
    mode = sbm.SELF_SKIP if sp else sbm.SELF_NEW
    c, lp = sbm.sbm_adjacency_resample_c_loops(A, g["c"], g["m"], g["p"], bool(sp), g["u"])
    assert np.array_equal(c, g["c_new_%d" % sp]) and np.allclose(lp, g["lp_%d" % sp], rtol=1e-11, atol=1e-10)
    c, lp = sbm.sbm_assign_sweep(A, None, g["c"], g["m"], g["p"], None, None, g["u"], mode)
    assert np.array_equal(c, g["c_new_%d" % sp]) and np.allclose(lp, g["lp_%d" % sp], rtol=1e-11, atol=1e-10)
    K = len(g["m"])
    st = sbm.block_pair_stats(A, None, g["c"], K)
    n_edge, n_noedge, _, _ = sbm.cell_stats_with_diag(st, include_diag=not sp)
    rs = np.random.RandomState(31 + sp)
    assert np.allclose(sbm.resample_p(n_edge, n_noedge, 1.0, 1.0, rng=rs), g["p_new_%d" % sp], rtol=1e-13)
    if sp:
        d1 = st["d_edge"].sum()
        assert np.isclose(rs.beta(1.0 + d1, 1.0 + A.shape[0] - d1), g["p_self_new"], rtol=1e-13)


def test_beta_bernoulli_resample_is_the_references():
    """adjacency.py:153-181; note :172 reuses tau1 / tau0 (not tau1_self / tau0_self) for the diagonal."""
    g = load("ref_adjacency.npz")
    A = g["A"].astype(float)
    N = A.shape[0]
    n1, d1 = A.sum(), np.trace(A)
    rs = np.random.RandomState(41)
    assert np.isclose(rs.beta(2.0 + n1, 3.0 + N * N - n1), g["bb_p_0"], rtol=1e-13)
    rs = np.random.RandomState(42)
    assert np.isclose(rs.beta(2.0 + n1 - d1, 3.0 + N * N - N - (n1 - d1)), g["bb_p_1"], rtol=1e-13)
    assert np.isclose(rs.beta(2.0 + d1, 3.0 + N - d1), g["bb_p_self"], rtol=1e-13)


@pytest.mark.parametrize("sp", [1, 0])
def test_sbm_weights_resample_is_the_references(sp):
    g = load("ref_weights.npz")
    A, W = g["A"].astype(float), g["W"]
    mode = sbm.SELF_SKIP if sp else sbm.SELF_NEW
    c, lp = sbm.sbm_weights_resample_c_loops(A, W, g["c"], g["m"], g["mus"], g["sigmas"], bool(sp), g["u"])
    assert np.array_equal(c, g["c_new_%d" % sp]) and np.allclose(lp, g["lp_%d" % sp], rtol=1e-11, atol=1e-10)
    c, lp = sbm.sbm_assign_sweep(A, W, g["c"], g["m"], None, g["mus"], g["sigmas"], g["u"], mode)
    assert np.array_equal(c, g["c_new_%d" % sp]) and np.allclose(lp, g["lp_%d" % sp], rtol=1e-11, atol=1e-10)
    K = len(g["m"])
    st = sbm.block_pair_stats(A, W, g["c"], K)
    n_edge, _, sum_w, sum_ww = sbm.cell_stats_with_diag(st, include_diag=not sp)
    assert np.array_equal(n_edge, g["cell_n_%d" % sp])
    assert np.allclose(sum_w, g["cell_sum_%d" % sp]) and np.allclose(sum_ww, g["cell_sumsq_%d" % sp])
    if sp:
        assert st["d_edge"].sum() == g["self_n"] and np.allclose(st["d_sum_w"].sum(0), g["self_sum"])
        assert np.allclose(st["d_sum_ww"].sum(0), g["self_sumsq"])


def test_niw_weights_statistics_are_the_references():
    """weights.py:154-162: off-diagonal and diagonal edge weights handed to the two Gaussians, and the draws."""
    g = load("ref_weights.npz")
    A, W = g["A"].astype(float), g["W"]
    B = W.shape[2]
    st = sbm.block_pair_stats(A, W, np.zeros(A.shape[0], int), 1)
    assert st["n_edge"][0, 0] == g["niw_off_n"] and np.allclose(st["sum_w"][0, 0], g["niw_off_sum"])
    assert np.allclose(st["sum_ww"][0, 0], g["niw_off_sumsq"])
    assert st["d_edge"][0] == g["niw_on_n"] and np.allclose(st["d_sum_w"][0], g["niw_on_sum"])
    rs = np.random.RandomState(int(g["niw_seed"]))
    hyp = (np.zeros(B), np.eye(B), 1.0, B + 2)
    mu, sig = niw.niw_resample(*hyp, stats=(st["n_edge"][0, 0], st["sum_w"][0, 0], st["sum_ww"][0, 0]), rng=rs)
    assert np.allclose(mu, g["niw_mu"]) and np.allclose(sig, g["niw_sigma"])
    mu, sig = niw.niw_resample(*hyp, stats=(st["d_edge"][0], st["d_sum_w"][0], st["d_sum_ww"][0]), rng=rs)
    assert np.allclose(mu, g["niw_self_mu"]) and np.allclose(sig, g["niw_self_sigma"])


# ---------------------------------------------------------------------------------------------- latent distance
def test_latent_lp_and_gradient_are_the_references():
    """adjacency.py:244-275 evaluated by the reference; gradients = complex-step derivative THROUGH the reference's code."""
    g = load("ref_latent.npz")
    A = g["A"].astype(float)
    for i in range(len(g["eval_lp"])):
        L, m0, ms = g["eval_L"][i], float(g["eval_mu0"][i]), float(g["eval_muself"][i])
        assert np.isclose(lt.hmc_log_probability(L, m0, ms, A, float(g["sigma"])), g["eval_lp"][i], rtol=1e-12)
        lp, gL, g0, gs = lt.log_probability_and_grad(L, m0, ms, A, float(g["sigma"]))
        assert np.isclose(lp, g["eval_lp"][i], rtol=1e-11)
        assert np.allclose(gL, g["eval_gL"][i], rtol=1e-9, atol=1e-10)
        assert np.isclose(g0, g["eval_gmu0"][i], rtol=1e-9) and np.isclose(gs, g["eval_gmuself"][i], rtol=1e-9)


def test_latent_resample_chain_is_the_references():
    """Four whole resample(A) calls (adjacency.py:356-421) with the HMC momenta / log-uniforms injected."""
    g = load("ref_latent.npz")
    A = g["A"].astype(float)
    L, m0, ms, step, rate = g["L"].copy(), float(g["mu_0"]), float(g["mu_self"]), 0.01, 0.9
    for s in range(len(g["sweep_mu0"])):
        lu = g["hmc_logu"][s]
        dr = dict(L=(g["hmc_pL"][s], lu[0]), mu_0=(g["hmc_p0"][s], lu[1]), mu_self=(g["hmc_ps"][s], lu[2]))
        L, m0, ms, step, rate = lt.resample(L, m0, ms, A, float(g["sigma"]), step, rate, dr)
        assert np.allclose(L, g["sweep_L"][s], rtol=1e-9, atol=1e-11), s
        assert np.isclose(m0, g["sweep_mu0"][s], rtol=1e-9) and np.isclose(ms, g["sweep_muself"][s], rtol=1e-9)
        assert np.isclose(step, g["sweep_step"][s]) and np.isclose(rate, g["sweep_rate"][s])


def test_latent_weight_lp_gradient_and_moves_are_the_references():
    """weights.py:697-731 evaluated by the reference (gradients = complex step THROUGH its code), three _resample_L
    transitions (:765-780) with injected draws, the b posterior (:838-860) and the residual statistics of :863-867."""
    g = load("ref_latent_weights.npz")
    A, W = g["A"].astype(float), g["W"][:, :, 0]
    sw, eta = float(g["sigma_w"]), float(g["eta"])
    for i in range(len(g["eval_lp"])):
        lp, gL, gb = lt.weight_log_probability_and_grad(g["eval_L"][i], float(g["eval_b"][i]), A, W, sw, eta)
        assert np.isclose(lp, g["eval_lp"][i], rtol=1e-12)
        assert np.allclose(gL, g["eval_gL"][i], rtol=1e-9, atol=1e-10) and np.isclose(gb, g["eval_gb"][i], rtol=1e-9)
    L = g["L"].copy()
    for s in range(len(g["hmc_logu"])):
        L = lt.weight_resample_L(L, float(g["b"]), A, W, sw, eta, (g["hmc_pL"][s], float(g["hmc_logu"][s])))
        assert np.allclose(L, g["hmc_L"][s], rtol=1e-9, atol=1e-11), s
    n, sr, srr = lt.weight_stats(L, float(g["b"]), A, W)
    mean, var = lt.weight_b_posterior(n, sr + n * float(g["b"]), sw)
    np.random.seed(int(g["b_seed"]))
    b_new = mean + np.sqrt(var) * np.random.randn()
    assert np.isclose(b_new, float(g["b_new"]), rtol=1e-12)
    n2, sr2, srr2 = lt.weight_stats(L, b_new, A, W)
    assert n2 == int(g["cov_n"]) and np.isclose(srr2, float(g["cov_S"]), rtol=1e-12)
    d = b_new - float(g["b"])
    assert np.isclose(srr - 2 * d * sr + n * d * d, srr2, rtol=1e-10)          # the identity the host class uses


# ---------------------------------------------------------------------------------------------- eigenmodel Gibbs
@pytest.mark.parametrize("given", [0, 1])
def test_logistic_eigenmodel_sweep_is_the_references(given):
    """eigenmodel.py:836-965 executed by the reference with Omega and the normals injected: per-node post_prec (the
    matrix it inverts), post_mean, the resulting F, the mu_0 conditional, the lmbda conditional."""
    g = load("ref_eigen.npz")
    t = "logistic_g%d_" % given
    A, O, F, lm, mu0 = g[t + "A"].astype(float), g[t + "O"], g[t + "F"], g[t + "lmbda"], float(g[t + "mu_0"])
    F1, prec, h = eg.logistic_resample_F_loops(A, O, F, mu0, lm, 1.3, g[t + "zF"])
    assert np.allclose(prec, g[t + "prec"], rtol=1e-11)
    assert np.allclose(np.linalg.solve(prec, h[:, :, None])[:, :, 0], g[t + "mean"], rtol=1e-9, atol=1e-12)
    assert np.allclose(F1, g[t + "F_new"], rtol=1e-9, atol=1e-11)
    S, R = eg.logistic_sym(A, O, mu0)
    F2, prec2, h2 = eg.resample_F(S, R, F, lm, 1.3, g[t + "zF"])
    assert np.allclose(prec2, g[t + "prec"], rtol=1e-10) and np.allclose(F2, g[t + "F_new"], rtol=1e-8, atol=1e-10)
    pp, hh = eg.logistic_mu0_conditional(A, O, F1, lm, float(g[t + "mu_mu_0"]), 0.9)
    assert np.isclose(hh / pp, g[t + "mu0_mean"], rtol=1e-10) and np.isclose(np.sqrt(1 / pp), g[t + "mu0_std"], rtol=1e-12)
    mu0n = float(g[t + "mu_0_new"])
    assert np.isclose(hh / pp + np.sqrt(1 / pp) * float(g[t + "zmu"]), mu0n, rtol=1e-10)
    if not given:
        lp_, lh_ = eg.logistic_lmbda_conditional_loops(A, O, F1, mu0n, 0.2 * np.ones(len(lm)), 1.7)
        assert np.allclose(lp_, g[t + "lam_prec"], rtol=1e-11)
        assert np.allclose(np.linalg.solve(lp_, lh_), g[t + "lam_mean"], rtol=1e-9)
        S, R = eg.logistic_sym(A, O, mu0n)
        lp2, lh2 = eg.lmbda_conditional(S, R, F1, 0.2 * np.ones(len(lm)), 1.7)
        assert np.allclose(lp2, g[t + "lam_prec"], rtol=1e-10) and np.allclose(lh2, lh_, rtol=1e-9)
        assert np.allclose(eg.gaussian_from_prec(lp_, lh_, g[t + "zl"])[0], g[t + "lmbda_new"], rtol=1e-9)
    assert np.isclose(eg.log_likelihood(A, eg.logistic_P(F, mu0, lm)), g[t + "ll"], rtol=1e-12)


@pytest.mark.parametrize("given", [0, 1])
def test_probit_eigenmodel_sweep_is_the_references(given):
    """eigenmodel.py:268-379 executed by the reference (Z through its own sample_truncnorm on a seeded stream)."""
    g = load("ref_eigen.npz")
    t = "probit_g%d_" % given
    A, F, lm, mu0 = g[t + "A"].astype(float), g[t + "F"], g[t + "lmbda"], float(g[t + "mu_0"])
    Z = eg.probit_resample_Z(A, F, mu0, lm, g[t + "u"])
    assert np.allclose(Z, g[t + "Z"], rtol=1e-12, atol=1e-13)
    F1, prec, h = eg.probit_resample_F_loops(Z, F, mu0, lm, 1.3, g[t + "zF"])
    assert np.allclose(prec, g[t + "prec"], rtol=1e-11) and np.allclose(F1, g[t + "F_new"], rtol=1e-9, atol=1e-11)
    S, R = eg.probit_sym(Z, mu0)
    F2, prec2, _ = eg.resample_F(S, R, F, lm, 1.3, g[t + "zF"])
    assert np.allclose(prec2, g[t + "prec"], rtol=1e-10) and np.allclose(F2, g[t + "F_new"], rtol=1e-8, atol=1e-10)
    pp, hh = eg.probit_mu0_conditional(Z, F1, lm, float(g[t + "mu_mu_0"]), 0.9)
    assert np.isclose(hh / pp, g[t + "mu0_mean"], rtol=1e-10) and np.isclose(np.sqrt(1 / pp), g[t + "mu0_std"], rtol=1e-12)
    if not given:
        mu0n = float(g[t + "mu_0_new"])
        S, R = eg.probit_sym(Z, mu0n)
        lp2, lh2 = eg.lmbda_conditional(S, R, F1, 0.2 * np.ones(len(lm)), 1.7)
        assert np.allclose(lp2, g[t + "lam_prec"], rtol=1e-10)
        assert np.allclose(np.linalg.solve(lp2, lh2), g[t + "lam_mean"], rtol=1e-9)
    assert np.isclose(eg.log_likelihood(A, eg.probit_P(F, mu0, lm)), g[t + "ll"], rtol=1e-12)


# ---------------------------------------------------------------------------------------------- AIS steps
def test_ais_tempered_steps_are_the_references():
    """internals/graph_models.py:261-297 (naive_sample_f), :337-371 and :168-182 (sample_theta), :42-64 (logpr)."""
    g = load("ref_ais.npz")
    A, z, Bm, pi = g["A"].astype(float), g["z"], g["B"], g["pi"]
    N, R = A.shape[0], len(pi)
    for bi, beta in enumerate(g["betas"]):
        zn, lp = sbm.ais_naive_sample_f_loops(A, z, Bm, pi, float(beta), g["u"])
        assert np.array_equal(zn, g["z_new_%d" % bi]) and np.allclose(lp, g["lp_%d" % bi], rtol=1e-11, atol=1e-10)
        zn, lp = sbm.sbm_assign_sweep(A, None, z, pi, Bm, None, None, g["u"], sbm.SELF_AIS, beta=float(beta), sampler="lse")
        assert np.array_equal(zn, g["z_new_%d" % bi]) and np.allclose(lp, g["lp_%d" % bi], rtol=1e-11, atol=1e-10)
        # sample_theta (:337-371): Dirichlet first, then Beta per cell in row-major order, stored in a float32 matrix
        rs = np.random.RandomState(90 + bi)
        pin = rs.dirichlet(10.0 * np.ones(R) + np.bincount(z, minlength=R))
        st = sbm.block_pair_stats(A, None, z, R)
        n_edge, n_noedge, _, _ = sbm.cell_stats_with_diag(st, include_diag=True)
        Bn = np.zeros((R, R), np.float32)
        for r1 in range(R):
            for r2 in range(R):
                Bn[r1, r2] = rs.beta(1.0 + beta * n_edge[r1, r2], 1.0 + beta * n_noedge[r1, r2])
        assert np.allclose(pin, g["pi_new_%d" % bi], rtol=1e-13) and np.array_equal(Bn, g["B_new_%d" % bi])
        rs = np.random.RandomState(95 + bi)
        assert np.isclose(rs.beta(1.0 + beta * A.sum(), 1.0 + beta * (N * N - A.sum())), g["rho_new_%d" % bi], rtol=1e-13)


def test_ais_collapsed_sampler_is_the_references():
    """internals/graph_models.py:299-335 (collapsed_sample_f) executed on the reference's own class: the complete log
    posterior vector and the label drawn from it (the R-th uniform: the reference draws once per candidate, :331-333)."""
    g = load("ref_ais.npz")
    A, z, pi = g["A"].astype(float), g["z"], g["pi"]
    for k, n in enumerate(g["coll_nodes"]):
        lp = sbm.ais_collapsed_ln_post_loops(A, z, int(n), pi, 1.0, 1.0)
        assert np.allclose(lp, g["coll_lp_%d" % k], rtol=1e-12, atol=1e-10)
        assert sbm.log_sum_exp_sample(lp, g["coll_u"][k][-1]) == int(g["coll_z_%d" % k])


# ---------------------------------------------------------------------------------------------- mean-field SBM
def _wexp_from(niw_arr, B):
    C = niw_arr.shape[0]
    return [[mf_sbm.weight_expectations(niw_arr[a, b, 2:2 + B], niw_arr[a, b, 2 + B:].reshape(B, B), niw_arr[a, b, 0],
                                        niw_arr[a, b, 1]) for b in range(C)] for a in range(C)]


@pytest.mark.parametrize("tag,st", [("full", 1.0), ("step", 0.35)])
def test_mf_sbm_updates_are_the_references(tag, st):
    """The WHOLE loop nest of mf_update_c (stochastic_block_model.py:392-445, incl. the reverse-edge indentation of
    :425), mf_update_p (:366-389) and mf_update_m (:358-364) executed on the reference's own class."""
    g = load("ref_mf_sbm.npz")
    A, W = g["A"].astype(float), g["W"]
    B = W.shape[2]
    E_WWT = W[:, :, :, None] * W[:, :, None, :]
    wexp = _wexp_from(g["niw"], B)
    m1, _ = mf_sbm.mf_update_c_loops(A, W, E_WWT, g["mf_m"], g["mf_pi"], g["mf_tau1"], g["mf_tau0"], wexp, st,
                                     literal_reverse=True)
    assert np.allclose(m1, g["c_m_" + tag], rtol=1e-9, atol=1e-12)
    m2, _ = mf_sbm.mf_update_c(A, W, g["mf_m"], g["mf_pi"], g["mf_tau1"], g["mf_tau0"], wexp, st, literal_reverse=True)
    assert np.allclose(m2, g["c_m_" + tag], rtol=1e-8, atol=1e-11)
    t1, t0 = mf_sbm.mf_update_p_loops(A, g["mf_m"], 1.0, 1.0, g["mf_tau1"], g["mf_tau0"], st)
    assert np.allclose(t1, g["p_tau1_" + tag], rtol=1e-12) and np.allclose(t0, g["p_tau0_" + tag], rtol=1e-12)
    assert np.allclose(mf_sbm.mf_update_m(g["mf_pi"], np.ones(len(g["mf_pi"])), g["mf_m"], st), g["m_pi_" + tag], rtol=1e-12)


def test_mf_sbm_whole_meanfieldupdate_is_the_references():
    """meanfieldupdate (:330-344): p, then c, then m, then every cell's natural NIW parameters."""
    g = load("ref_mf_sbm.npz")
    A, W = g["A"].astype(float), g["W"]
    B, C = W.shape[2], g["mf_m"].shape[1]
    E_WWT = W[:, :, :, None] * W[:, :, None, :]
    t1, t0 = mf_sbm.mf_update_p_loops(A, g["mf_m"], 1.0, 1.0, g["mf_tau1"], g["mf_tau0"], 1.0)
    m1, _ = mf_sbm.mf_update_c_loops(A, W, E_WWT, g["mf_m"], g["mf_pi"], t1, t0, _wexp_from(g["niw"], B), 1.0,
                                     literal_reverse=True)
    pi1 = mf_sbm.mf_update_m(g["mf_pi"], np.ones(C), m1, 1.0)
    assert np.allclose(t1, g["mfu_tau1"]) and np.allclose(m1, g["mfu_m"], rtol=1e-9, atol=1e-12) and np.allclose(pi1, g["mfu_pi"])
    nat0 = [[mf_sbm.standard_to_natural(np.zeros(B), np.eye(B), 1.0, B + 2.0) for _ in range(C)] for _ in range(C)]
    nat = np.array(mf_sbm.mf_update_weight_models_loops(A, W, E_WWT, m1, nat0))
    assert np.allclose(nat, g["mfu_nat"], rtol=1e-10, atol=1e-12)


# ---------------------------------------------------------------------------------------------- mean-field eigenmodel
@pytest.mark.parametrize("given", [0, 1])
@pytest.mark.parametrize("tag,st", [("full", 1.0), ("step", 0.4)])
def test_mf_logistic_eigenmodel_updates_are_the_references(given, tag, st):
    """eigenmodel.py:1055-1157 executed by the reference (E_O by its own 1000-sample Monte Carlo on a seeded stream)."""
    g = load("ref_mf_eigen.npz")
    t = "logistic_g%d_" % given
    A, EO = g[t + "A"].astype(float), g[t + "E_O"]
    mu, Sg, Fg, lam = g[t + "mf_mu_F"], g[t + "mf_Sigma_F"], g[t + "F"], g[t + "lmbda"]
    D = mu.shape[1]
    SL = g[t + "mf_Sigma_lmbda"] if not given else 1e-16 * np.eye(D)
    mu0, s0 = float(g[t + "mf_mu_mu0"]), float(g[t + "mf_sigma_mu0"])
    assert np.allclose(mf_eigen.expected_mu(mu0, mu, lam), g[t + "E_mu"], rtol=1e-12)
    assert np.allclose(mf_eigen.expected_musq(mu0, s0, mu, Sg, lam, SL), g[t + "E_musq"], rtol=1e-11)
    # E_O itself: same seeded normals through the oracle's restatement of :1055-1064 / :1245-1248
    rs = np.random.RandomState(101)
    z = rs.randn(1000, A.shape[0], A.shape[0])
    EO2 = mf_eigen.expected_omega_mc(mf_eigen.expected_mu(mu0, mu, lam), mf_eigen.variance_mu(mu0, s0, mu, Sg, lam, SL), z)
    assert np.allclose(EO2, EO, rtol=1e-10)
    r0 = mf_eigen.meanfieldupdate_mu0(A, EO, mu, lam, float(g[t + "mu_mu_0"]), 0.9, mu0, s0, st)
    assert np.allclose([r0[0], r0[1]], g["%s%s_mu0" % (t, tag)], rtol=1e-10)
    llT = mf_eigen.expected_llT(lam, SL)
    for fn in (mf_eigen.meanfieldupdate_F_loops, mf_eigen.meanfieldupdate_F):
        mu1, Sg1, prec, _ = fn(A, EO, Fg, mu, Sg, r0[0], lam, llT, 1.3, st, False)
        assert np.allclose(prec, g["%s%s_F_prec" % (t, tag)], rtol=1e-9)
        assert np.allclose(mu1, g["%s%s_mu_F" % (t, tag)], rtol=1e-8, atol=1e-11)
        assert np.allclose(Sg1, g["%s%s_Sigma_F" % (t, tag)], rtol=1e-8, atol=1e-12)
    if not given:
        r = mf_eigen.meanfieldupdate_lmbda_loops(A, EO, mu1, Sg1, r0[0], 0.2 * np.ones(D), 1.7, lam, SL, st)
        assert np.allclose(r[2], g["%s%s_lam_prec" % (t, tag)], rtol=1e-9)
        assert np.allclose(r[0], g["%s%s_mu_lmbda" % (t, tag)], rtol=1e-8) and np.allclose(r[1], g["%s%s_Sigma_lmbda" % (t, tag)], rtol=1e-8)
