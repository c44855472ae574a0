"""Mean-field SBM on the GPU through the C ABI, against oracle/mf_sbm.py and the golden fixture made
with the reference's own distribution classes.  fp conditionals: 1e-5 relative with fp64 device storage, 1e-3 with fp32
(north_star tolerances)."""
import os

import numpy as np
import pytest

from oracle import mf_sbm as M

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "mf_sbm_small.npz"))


def _model(N, B, C, dtype, rng, literal=True):
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    model = GaussianStochasticBlockModel(N, B, C=C, dtype=dtype)
    model.literal_reverse_edge = literal
    model.mf_m = rng.dirichlet(2 * np.ones(C), size=N)
    model.mf_pi = rng.gamma(2.0, 1.0, C) + 0.5
    model.mf_tau1 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    model.mf_tau0 = rng.gamma(2.0, 2.0, (C, C)) + 0.3
    for a in range(C):
        for b in range(C):
            X = rng.randn(B, B)
            wm = model.weight_models[a][b]
            wm.mf_natural_hypparam = wm._standard_to_natural(rng.randn(B), X.dot(X.T) + np.eye(B), 1.0 + rng.rand(),
                                                             B + 2.0 + 3 * rng.rand())
    return model


def _wexp(model):
    C = model.C
    if model.B == 0:
        return None
    return [[M.weight_expectations(w.mu_mf, w.sigma_mf, w.kappa_mf, w.nu_mf) for w in row] for row in model.weight_models]


def _golden_model(lit):
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    A, W, niw = G["A"], G["W"], G["niw"]
    N, C, B = A.shape[0], niw.shape[0], W.shape[2]
    np.random.seed(0)
    model = GaussianStochasticBlockModel(N, B, C=C, dtype="f64")
    model.literal_reverse_edge = bool(lit)
    model.mf_m, model.mf_pi = G["mf_m"].copy(), G["mf_pi"].copy()
    model.mf_tau1, model.mf_tau0 = G["mf_tau1"].copy(), G["mf_tau0"].copy()
    for a in range(C):
        for b in range(C):
            wm = model.weight_models[a][b]
            wm.mf_natural_hypparam = wm._standard_to_natural(niw[a, b, 2:2 + B], niw[a, b, 2 + B:].reshape(B, B),
                                                             niw[a, b, 0], niw[a, b, 1])
    return model, FixedGaussianNetwork(A, W)


@pytest.mark.parametrize("lit", [1, 0])
@pytest.mark.parametrize("tag,step", [("full", 1.0), ("step", 0.35)])
def test_update_c_matches_reference_golden(lit, tag, step):
    model, net = _golden_model(lit)
    lp = model.mf_update_c(net, stepsize=step, want_lp=True)
    ref_lp = G["lp_%s_lit%d" % (tag, lit)]
    np.testing.assert_allclose(lp, ref_lp, rtol=1e-9, atol=1e-9 * np.abs(ref_lp).max())
    np.testing.assert_allclose(model.mf_m, G["m_%s_lit%d" % (tag, lit)], rtol=1e-7, atol=1e-12)


def test_update_p_and_weights_match_reference_golden():
    model, net = _golden_model(1)
    t1, t0 = model.mf_tau1.copy(), model.mf_tau0.copy()
    model.mf_update_p(net, stepsize=0.6)
    np.testing.assert_allclose(model.mf_tau1, G["p_tau1"], rtol=1e-11)
    np.testing.assert_allclose(model.mf_tau0, G["p_tau0"], rtol=1e-11)
    C = model.C
    for a in range(C):
        for b in range(C):
            model.weight_models[a][b]._mf_nat = None          # q back at the prior
    model._mf_update_weight_models(net)
    for a in range(C):
        for b in range(C):
            np.testing.assert_allclose(model.weight_models[a][b].mf_natural_hypparam, G["w_nat"][a, b], rtol=1e-10,
                                       atol=1e-10)


@pytest.mark.parametrize("N,C,B,dtype,lit", [(70, 3, 2, "f64", True), (333, 16, 2, "f64", False), (520, 5, 1, "f64", True),
                                             (257, 7, 1, "f64", True), (600, 16, 2, "f32", True), (300, 32, 1, "f32", False)])
def test_update_c_matches_oracle(N, C, B, dtype, lit):
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    rng = np.random.RandomState(N + C)
    np.random.seed(1)
    A = (rng.rand(N, N) < 0.3).astype(np.uint8)
    W = rng.randn(N, N, max(B, 1))[:, :, :B]
    if dtype == "f32":
        W = W.astype(np.float32)
    model = _model(N, B, C, dtype, rng, lit)
    net = FixedGaussianNetwork(A, W if B else None)
    Wd = W.astype(np.float64) if B else None
    for step in (1.0, 0.4):
        m0 = model.mf_m.copy()
        m_ref, lp_ref = M.mf_update_c(A.astype(float), Wd, m0, model.mf_pi, model.mf_tau1, model.mf_tau0, _wexp(model),
                                      step, literal_reverse=lit)
        lp = model.mf_update_c(net, stepsize=step, want_lp=True)
        tol = 1e-5 if dtype == "f64" else 1e-3
        scale = np.abs(lp_ref).max()
        assert np.abs(lp - lp_ref).max() <= tol * scale, (np.abs(lp - lp_ref).max(), scale)
        # class probabilities: compare where the oracle's softmax is not saturated away
        np.testing.assert_allclose(model.mf_m, m_ref, rtol=0, atol=(1e-6 if dtype == "f64" else 2e-2))
        np.testing.assert_allclose(model.mf_m.sum(1), 1.0, rtol=1e-9)


@pytest.mark.parametrize("N,C,B,dtype", [(200, 4, 2, "f64"), (700, 16, 2, "f32"), (130, 9, 1, "f64")])
def test_soft_stats_match_oracle(N, C, B, dtype):
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    rng = np.random.RandomState(5)
    np.random.seed(1)
    A = (rng.rand(N, N) < 0.4).astype(np.uint8)
    W = rng.randn(N, N, B)
    if dtype == "f32":
        W = W.astype(np.float32)
    model = _model(N, B, C, dtype, rng)
    n, sw, sww = model._soft_stats(FixedGaussianNetwork(A, W))
    n_ref, sw_ref, sww_ref = M.soft_block_stats(A.astype(float), W.astype(np.float64), model.mf_m)
    tol = 1e-10 if dtype == "f64" else 2e-5
    np.testing.assert_allclose(n, n_ref, rtol=tol)
    np.testing.assert_allclose(sw, sw_ref, rtol=0, atol=tol * np.abs(sww_ref).max())
    np.testing.assert_allclose(sww, sww_ref, rtol=0, atol=tol * np.abs(sww_ref).max())


def test_meanfieldupdate_sequence_matches_oracle():
    """model.meanfieldupdate(network) (:327-344) against the oracle functions composed in the same order."""
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    rng = np.random.RandomState(11)
    np.random.seed(2)
    N, C, B = 150, 4, 2
    A = (rng.rand(N, N) < 0.35).astype(np.uint8)
    W = rng.randn(N, N, B)
    model = _model(N, B, C, "f64", rng)
    net = FixedGaussianNetwork(A, W)
    Af = A.astype(float)
    mf_m, mf_pi, t1, t0 = model.mf_m.copy(), model.mf_pi.copy(), model.mf_tau1.copy(), model.mf_tau0.copy()
    nat = [[w.mf_natural_hypparam.copy() for w in row] for row in model.weight_models]
    prior = model.weight_models[0][0].natural_hypparam
    for it in range(2):
        model.meanfieldupdate(net)
        t1, t0 = M.mf_update_p(Af, mf_m, model.tau1, model.tau0, t1, t0)
        wexp = [[M.weight_expectations(*M.natural_to_standard(nat[a][b])) for b in range(C)] for a in range(C)]
        mf_m, _ = M.mf_update_c(Af, W, mf_m, mf_pi, t1, t0, wexp, 1.0)
        mf_pi = M.mf_update_m(mf_pi, model.pi, mf_m)
        n, sw, sww = M.soft_block_stats(Af, W, mf_m)
        nat = [[prior + M.statistics_matrix(n[a, b], sw[a, b], sww[a, b]) for b in range(C)] for a in range(C)]
        np.testing.assert_allclose(model.mf_tau1, t1, rtol=1e-9)
        np.testing.assert_allclose(model.mf_m, mf_m, rtol=0, atol=1e-7)
        np.testing.assert_allclose(model.mf_pi, mf_pi, rtol=1e-8)
        for a in range(C):
            for b in range(C):
                np.testing.assert_allclose(model.weight_models[a][b].mf_natural_hypparam, nat[a][b], rtol=1e-7, atol=1e-7)
    model.resample_from_mf()
    assert model.c.shape == (N,) and model.p.shape == (C, C)
    E = model.mf_expected_p()
    assert E.shape == (N, N) and np.all((E >= 0) & (E <= 1))
    assert model.mf_expected_mu().shape == (N, N, B) and model.mf_expected_Sigma_inv().shape == (N, N, B, B)
