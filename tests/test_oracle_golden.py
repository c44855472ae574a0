"""Oracle vs the reference's own code (ref_internals.npz was produced by importing
/root/reference/graphistician/internals/{utils,distributions}.py) and vs the frozen T0 fixtures."""
import os

import numpy as np
from scipy.stats import truncnorm

from oracle import eigenmodel as eg, latent as lt, sbm


def _g(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_reference_known_answers(golden_dir):
    # test/test_distributions.py:14,28,32,36 -- the reference's only asserts
    g = _g(golden_dir, "ref_internals.npz")
    assert np.allclose(g["tsg0_mean"], truncnorm(0, np.inf).mean())
    tn = truncnorm(-3.0, 7.0, loc=3)
    assert np.allclose(g["tsg1_mean"], tn.mean())
    assert np.allclose(g["tsg1_var"], tn.var())
    assert np.allclose(-g["tsg1_negent"], tn.entropy())
    assert np.allclose([g["tsg1_mean"], g["tsg1_var"]], [3.00443783903, 0.98666678839])


def test_oracle_truncnorm_matches_reference(golden_dir):
    g = _g(golden_dir, "ref_internals.npz")
    A = g["tn_A"]; N = A.shape[0]
    lb = -np.inf * np.ones((N, N)); lb[A == 1] = 0
    ub = np.inf * np.ones((N, N)); ub[A == 0] = 0
    z = eg.sample_truncnorm(g["tn_mu"], lb, ub, g["tn_u"])
    assert np.array_equal(z, g["tn_z"])          # same arithmetic, same uniforms -> bit-identical
    assert np.allclose(eg.normal_cdf(g["x"]), g["normal_cdf"], rtol=0, atol=0)
    assert np.allclose(1.0 / (1.0 + np.exp(-g["x"])), g["logistic"], rtol=0, atol=0)


def test_oracle_bernoulli_matches_reference(golden_dir):
    g = _g(golden_dir, "ref_internals.npz")
    assert np.array_equal(sbm._bern_logp(g["bern_p"], g["bern_x"]), g["bern_lp"])


def test_sbm_fixture(golden_dir):
    g = _g(golden_dir, "sbm_small.npz")
    A = g["A"].astype(float); W = g["W"]; c = g["c"]; K = len(g["m"])
    st = sbm.block_pair_stats(A, W, c, K)
    for k, v in st.items():
        assert np.allclose(v, g["st_" + k], rtol=1e-12, atol=1e-12), k
    assert np.array_equal(st["n_edge"], g["st_n_edge"]) and np.array_equal(st["n_pair"], g["st_n_pair"])
    c1, lp1 = sbm.sbm_assign_sweep(A, W, c, g["m"], g["p"], g["mus"], g["sigmas"], g["u"], sbm.SELF_OLD)
    assert np.array_equal(c1, g["c_old"]) and np.allclose(lp1, g["lp_old"], rtol=1e-10)
    for sp, mode in ((1, sbm.SELF_SKIP), (0, sbm.SELF_NEW)):
        ca, la = sbm.sbm_assign_sweep(A, None, c, g["m"], g["p"], None, None, g["u"], mode)
        assert np.array_equal(ca, g["c_adj_%d" % sp]) and np.allclose(la, g["lp_adj_%d" % sp], rtol=1e-10)
        cw, lw = sbm.sbm_assign_sweep(A, W, c, g["m"], None, g["mus"], g["sigmas"], g["u"], mode)
        assert np.array_equal(cw, g["c_w_%d" % sp]) and np.allclose(lw, g["lp_w_%d" % sp], rtol=1e-10)
    cz, lz = sbm.sbm_assign_sweep(A, None, c, g["m"], g["p"], None, None, g["u"], sbm.SELF_AIS,
                                  beta=float(g["ais_beta"]), sampler="lse")
    assert np.array_equal(cz, g["c_ais"]) and np.allclose(lz, g["lp_ais"], rtol=1e-10)


def test_eigen_fixture(golden_dir):
    g = _g(golden_dir, "eigen_small.npz")
    A = g["A"].astype(float); O = g["O"]; F = g["F"]; lm = g["lmbda"]; mu0 = float(g["mu_0"])
    S, R = eg.logistic_sym(A, O, mu0)
    F1, prec, h = eg.resample_F(S, R, F, lm, float(g["sigma_F"]), g["zF"])
    assert np.allclose(F1, g["F_logistic"], rtol=1e-9) and np.allclose(prec, g["prec_logistic"], rtol=1e-10)
    assert np.allclose(h, g["h_logistic"], rtol=1e-9, atol=1e-12)
    lp, lh = eg.lmbda_conditional(S, R, g["F_logistic"], 0.0, 2.0)
    assert np.allclose(lp, g["lam_prec"], rtol=1e-10) and np.allclose(lh, g["lam_h"], rtol=1e-9)
    pp, ph = eg.logistic_mu0_conditional(A, O, g["F_logistic"], lm, -1.2, 0.9)
    assert np.allclose([pp, ph], [g["mu0_prec"], g["mu0_h"]])
    Z = eg.probit_resample_Z(A, F, mu0, lm, g["u"])
    assert np.array_equal(Z, g["Z"])
    S2, R2 = eg.probit_sym(Z, mu0)
    F2, prec2, h2 = eg.resample_F(S2, R2, F, lm, float(g["sigma_F"]), g["zF"])
    assert np.allclose(F2, g["F_probit"], rtol=1e-9) and np.allclose(prec2, g["prec_probit"], rtol=1e-10)


def test_latent_fixture(golden_dir):
    g = _g(golden_dir, "latent_small.npz")
    A = g["A"].astype(float)
    lp, gL, g0, gs = lt.log_probability_and_grad(g["L"], float(g["mu_0"]), float(g["mu_self"]), A, float(g["sigma"]))
    assert np.allclose(lp, g["lp_literal"], rtol=1e-12) and np.allclose(gL, g["gL"])
    eps = 1e-6
    L = g["L"].copy(); L[3, 1] += eps
    lp2 = lt.hmc_log_probability(L, float(g["mu_0"]), float(g["mu_self"]), A, float(g["sigma"]))
    assert abs((lp2 - g["lp_literal"]) / eps - gL[3, 1]) < 1e-3
    dr = dict(L=(g["pL"], float(g["luL"])), mu_0=(float(g["p0"]), float(g["lu0"])),
              mu_self=(float(g["ps"]), float(g["lus"])))
    L1, m1, s1, st, ar = lt.resample(g["L"], float(g["mu_0"]), float(g["mu_self"]), A, float(g["sigma"]), 0.01, 0.9, dr)
    assert np.allclose(L1, g["L_new"]) and np.allclose([m1, s1, st, ar], [g["mu_0_new"], g["mu_self_new"], g["step"], g["rate"]])
