"""GaussianStochasticBlockModel.resample(network) on the GPU against the oracle's full sweep
(stochastic_block_model.py:206-230), same np.random stream: c bit-exact, (mu, Sigma, p, m) to 1e-5 (fp64 data)."""
import numpy as np
import pytest

from oracle import sbm

pytestmark = pytest.mark.gpu


def _oracle_niw_cells(n_edge, sw, sww, niw):
    """Per-cell NIW posterior + injected-draw sample (oracle/niw.py), consuming np.random in the engine's order."""
    from oracle import niw as oniw
    C, B = n_edge.shape[0], sw.shape[2]
    post = [[oniw.niw_posterior(*niw, n_edge[a, b], sw[a, b], sww[a, b]) for b in range(C)] for a in range(C)]
    nu = np.array([[post[a][b][3] for b in range(C)] for a in range(C)])
    chi2 = np.random.chisquare(nu[..., None] - np.arange(B))
    z_off = np.random.standard_normal((C, C, B * (B - 1) // 2))
    z_mu = np.random.standard_normal((C, C, B))
    mus = np.zeros((C, C, B)); sigmas = np.zeros((C, C, B, B))
    for a in range(C):
        for b in range(C):
            mus[a, b], sigmas[a, b] = oniw.sample_niw_injected(*post[a][b], chi2[a, b], z_off[a, b], z_mu[a, b])
    return mus, sigmas


@pytest.mark.parametrize("N,C,B,dtype", [(50, 5, 2, "f64"), (160, 4, 1, "f64"), (300, 16, 2, "f32"), (120, 4, 3, "f64"), (200, 6, 4, "f64")])
def test_resample_matches_oracle(N, C, B, dtype):
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    np.random.seed(3)
    truth = GaussianStochasticBlockModel(N, B, C=C, dtype=dtype)
    net = truth.rvs()
    if dtype == "f32":
        net.W = net.W.astype(np.float32)
    model = GaussianStochasticBlockModel(N, B, C=C, dtype=dtype)
    state = dict(c=model.c.copy(), m=model.m.copy(), p=model.p.copy())
    hyp = dict(pi=model.pi, tau0=model.tau0, tau1=model.tau1,
               niw=(np.zeros(B), np.eye(B), 1.0, B + 2))
    A = net.A.astype(float); W = net.W.astype(np.float64)
    for sweep in range(3):
        np.random.seed(100 + sweep)
        model.resample(net)
        np.random.seed(100 + sweep)
        st = sbm.block_pair_stats(A, W, state["c"], C)
        n_edge, n_noedge, sw, sww = sbm.cell_stats_with_diag(st, include_diag=True)
        mus, sigmas = _oracle_niw_cells(n_edge, sw, sww, hyp["niw"])
        p = sbm.resample_p(n_edge, n_noedge, hyp["tau1"], hyp["tau0"])
        u = np.random.random(N)
        c, _ = sbm.sbm_assign_sweep(A, W, state["c"], state["m"], p, mus, sigmas, u, sbm.SELF_OLD)
        m = sbm.resample_m(c, hyp["pi"])
        state = dict(c=c, m=m, p=p)
        tol = 1e-5 if dtype == "f64" else 1e-3
        assert np.allclose(model.mus, mus, rtol=tol, atol=tol)
        assert np.allclose(model.sigmas, sigmas, rtol=tol, atol=tol)
        assert np.allclose(model.p, p, rtol=1e-12)
        assert np.array_equal(model.c, c)
        assert np.allclose(model.m, m, rtol=1e-12)
    ll = model.log_likelihood(net)
    Ps = np.clip(model.P, 1e-300, 1 - 1e-16)
    ll_ref = np.sum(A * np.log(Ps) + (1 - A) * np.log1p(-Ps)) + sbm.gaussian_edge_log_likelihood_loops(A, W, model.Mu, model.Sigma)
    assert np.isclose(ll, ll_ref, rtol=1e-5 if dtype == "f64" else 1e-3)
    assert np.isfinite(model.log_probability(net))
