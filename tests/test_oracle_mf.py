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
