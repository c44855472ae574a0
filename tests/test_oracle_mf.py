"""CPU: the mean-field SBM oracle (oracle/mf_sbm.py) against the golden fixture produced with the reference's own
internals/distributions.py classes doing the per-pair arithmetic (tests/golden/make_golden.py:make_mf_sbm)."""
import os

import numpy as np
import pytest

from oracle import mf_sbm as M

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "mf_sbm_small.npz"))


def _wexp():
    niw = G["niw"]
    C, B = niw.shape[0], G["W"].shape[2]
    return [[M.weight_expectations(niw[a, b, 2:2 + B], niw[a, b, 2 + B:].reshape(B, B), niw[a, b, 0], niw[a, b, 1])
             for b in range(C)] for a in range(C)]


@pytest.mark.parametrize("lit", [1, 0])
@pytest.mark.parametrize("tag,step", [("full", 1.0), ("step", 0.35)])
def test_update_c_loops_and_vectorised_match_reference_arithmetic(lit, tag, step):
    A, W = G["A"].astype(float), G["W"]
    E_WWT = W[:, :, :, None] * W[:, :, None, :]
    args = (G["mf_m"], G["mf_pi"], G["mf_tau1"], G["mf_tau0"], _wexp(), step)
    m_loops, lp_loops = M.mf_update_c_loops(A, W, E_WWT, *args, literal_reverse=bool(lit))
    m_vec, lp_vec = M.mf_update_c(A, W, *args, literal_reverse=bool(lit))
    for m, lp in ((m_loops, lp_loops), (m_vec, lp_vec)):
        np.testing.assert_allclose(lp, G["lp_%s_lit%d" % (tag, lit)], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(m, G["m_%s_lit%d" % (tag, lit)], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(m.sum(1), 1.0, rtol=1e-12)


def test_reverse_edge_quirk_changes_the_result():
    assert np.abs(G["lp_full_lit1"] - G["lp_full_lit0"]).max() > 1e-3


def test_unweighted_update_c():
    A, W = G["A"].astype(float), G["W"]
    m, lp = M.mf_update_c(A, None, G["mf_m"], G["mf_pi"], G["mf_tau1"], G["mf_tau0"], None, 1.0)
    np.testing.assert_allclose(lp, G["lp_noW"], rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(m, G["m_noW"], rtol=1e-9, atol=1e-12)


def test_update_p_and_weight_statistics():
    A, W = G["A"].astype(float), G["W"]
    t1, t0 = M.mf_update_p(A, G["mf_m"], 1.0, 1.0, G["mf_tau1"], G["mf_tau0"], 0.6)
    np.testing.assert_allclose(t1, G["p_tau1"], rtol=1e-12)
    np.testing.assert_allclose(t0, G["p_tau0"], rtol=1e-12)
    n, sw, sww = M.soft_block_stats(A, W, G["mf_m"])
    C, B = n.shape[0], W.shape[2]
    nat0 = M.standard_to_natural(np.zeros(B), np.eye(B), 1.0, B + 2.0)
    for a in range(C):
        for b in range(C):
            np.testing.assert_allclose(nat0 + M.statistics_matrix(n[a, b], sw[a, b], sww[a, b]), G["w_nat"][a, b],
                                       rtol=1e-11, atol=1e-11)


def test_natural_parameter_round_trip():
    rng = np.random.RandomState(0)
    X = rng.randn(2, 2)
    mu, sig, kap, nu = rng.randn(2), X.dot(X.T) + np.eye(2), 1.7, 5.5
    mu2, sig2, kap2, nu2 = M.natural_to_standard(M.standard_to_natural(mu, sig, kap, nu))
    np.testing.assert_allclose(mu2, mu); np.testing.assert_allclose(sig2, sig)
    assert abs(kap2 - kap) < 1e-12 and abs(nu2 - nu) < 1e-12


# ---- get_vlb (stochastic_block_model.py:536-568) ---------------------------------------------------------------------
@pytest.mark.parametrize("tag,key", [("vlb_nogauss", "mf_m"), ("vlb_nogauss_zero", "mf_m_zero")])
def test_vlb_matches_reference_distribution_classes(tag, key):
    """golden values made by the reference's own Discrete / Beta / Dirichlet classes (make_golden.py), including the
    nan_to_num quirk for a q(c_n) with an exact zero and the once-only Beta normaliser for scalar tau."""
    C = G["mf_pi"].size
    zero = [[0.0] * C for _ in range(C)]
    v = M.sbm_vlb_loops(G[key], np.ones(C), G["mf_pi"], 1.0, 1.0, G["mf_tau1"], G["mf_tau0"], zero)
    assert np.isclose(v, float(G[tag]), rtol=1e-13)


def test_niw_vlb_is_minus_kl():
    """pybasicbayes is absent: the NIW term is pinned by what it must equal, -KL(q || prior)."""
    from scipy.stats import invwishart, multivariate_normal
    mu0, S0, k0, nu0 = np.array([0.3, -0.2]), np.array([[2.0, 0.3], [0.3, 1.0]]), 1.5, 5.0
    assert abs(M.niw_vlb(mu0, S0, k0, nu0, mu0, S0, k0, nu0)) < 1e-12
    mu, S, k, nu = np.array([0.8, 0.1]), np.array([[3.0, -0.5], [-0.5, 2.5]]), 4.0, 9.0
    v = M.niw_vlb(mu0, S0, k0, nu0, mu, S, k, nu)
    assert v < 0
    rng = np.random.default_rng(0)
    n = 6000
    Sg = invwishart.rvs(df=nu, scale=S, size=n, random_state=1)
    vals = np.empty(n)
    for i in range(n):
        m = rng.multivariate_normal(mu, Sg[i] / k)
        vals[i] = (invwishart.logpdf(Sg[i], df=nu0, scale=S0) + multivariate_normal.logpdf(m, mu0, Sg[i] / k0)
                   - invwishart.logpdf(Sg[i], df=nu, scale=S) - multivariate_normal.logpdf(m, mu, Sg[i] / k))
    assert abs(vals.mean() - v) < 4.5 * vals.std() / np.sqrt(n)


def test_model_get_vlb_matches_oracle():
    """host class (no device work in get_vlb, as in the reference)"""
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    rng = np.random.RandomState(5)
    N, B, C = 12, 2, 3
    model = GaussianStochasticBlockModel(N, B, C=C)
    assert abs(sum(w.get_vlb() for row in model.weight_models for w in row)) < 1e-10      # q starts at the prior
    model.mf_m = rng.dirichlet(2 * np.ones(C), size=N)
    model.mf_m[4] = np.array([0.0, 0.25, 0.75])
    model.mf_pi = rng.gamma(2.0, 1.0, C) + 0.5
    model.mf_tau1 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    model.mf_tau0 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    terms = [[0.0] * C for _ in range(C)]
    for a in range(C):
        for b in range(C):
            X = rng.randn(B, B)
            wm = model.weight_models[a][b]
            wm.mf_natural_hypparam = wm._standard_to_natural(rng.randn(B), X.dot(X.T) + np.eye(B), 1.0 + rng.rand(),
                                                             B + 2.0 + 3 * rng.rand())
            terms[a][b] = M.niw_vlb(wm.mu_0, wm.sigma_0, wm.kappa_0, wm.nu_0, wm.mu_mf, wm.sigma_mf, wm.kappa_mf, wm.nu_mf)
            assert np.isclose(wm.get_vlb(), terms[a][b], rtol=1e-10)
    ref = M.sbm_vlb_loops(model.mf_m, model.pi, model.mf_pi, model.tau1, model.tau0, model.mf_tau1, model.mf_tau0, terms)
    assert np.isclose(model.get_vlb(), ref, rtol=1e-10)
