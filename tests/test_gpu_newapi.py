"""New-API classes (adjacency.py / weights.py / networks.py mirrors) on the GPU against the oracle.

Same np.random stream on both sides: integer statistics and block assignments bit-exact, the sampled Beta / NIW
parameters equal to round-off (fp64 device data) because they are functions of exact counts / fp64 sums."""
import numpy as np
import pytest

from oracle import sbm, niw as oniw

pytestmark = pytest.mark.gpu


def _er_data(N, B, p, seed, self_edges=True):
    rng = np.random.RandomState(seed)
    A = (rng.rand(N, N) < p).astype(np.uint8)
    if not self_edges:
        np.fill_diagonal(A, 0)
    W = rng.randn(N, N, B) * 0.7 + 0.3
    return A, W


@pytest.mark.parametrize("N", [10, 77, 300])
@pytest.mark.parametrize("self_conn", [False, True])
def test_beta_bernoulli_resample_and_ll(N, self_conn):
    """Config 1's adjacency factor (adjacency.py:153-181): counts bit-exact, the Beta draw identical."""
    from graphistician_b200.adjacency import BetaBernoulliAdjacencyDistribution
    A, _ = _er_data(N, 1, 0.3, N)
    np.random.seed(1)
    kw = dict(tau1_self=2.0, tau0_self=3.0) if self_conn else {}
    model = BetaBernoulliAdjacencyDistribution(N, tau1=1.5, tau0=2.5, **kw)
    np.random.seed(7)
    model.resample(A)
    np.random.seed(7)
    Af = A.astype(float)
    if self_conn:
        off = Af.sum() - np.trace(Af)
        p = np.random.beta(1.5 + off, 2.5 + N * (N - 1) - off)
        ps = np.random.beta(1.5 + np.trace(Af), 2.5 + N - np.trace(Af))      # the reference's (tau1, tau0) quirk
        assert model.p == p and model.p_self == ps
    else:
        p = np.random.beta(1.5 + Af.sum(), 2.5 + N * N - Af.sum())
        assert model.p == p
    P = model.P
    ref = np.sum(Af * np.log(P) + (1 - Af) * np.log(1 - P))
    assert np.isclose(model.log_likelihood(A), ref, rtol=1e-12)


@pytest.mark.parametrize("N,B", [(10, 2), (130, 1), (257, 2)])
def test_niw_weights_resample_and_ll(N, B):
    """Config 1's weight factor (weights.py:154-162): edge count exact, fp64 sums, same NIW draw as the oracle."""
    from graphistician_b200.weights import NIWGaussianWeightDistribution
    A, W = _er_data(N, B, 0.4, 3 * N)
    np.random.seed(2)
    model = NIWGaussianWeightDistribution(N, B=B, dtype="f64")
    np.random.seed(11)
    model.resample((A, W))
    np.random.seed(11)
    off = A.astype(bool) & ~np.eye(N, dtype=bool)
    on = A.astype(bool) & np.eye(N, dtype=bool)
    (n_off, sw, sww), (n_on, dsw, dsww) = model._last_stats
    assert n_off == off.sum() and n_on == on.sum()
    assert np.allclose(sw, W[off].sum(0), rtol=1e-10) and np.allclose(sww, W[off].T @ W[off], rtol=1e-10)
    mu1, sig1 = oniw.niw_resample(np.zeros(B), np.eye(B), 1.0, B + 2, data=W[off])
    mu2, sig2 = oniw.niw_resample(np.zeros(B), np.eye(B), 1.0, B + 2, data=W[on])
    assert np.allclose(model._gaussian.mu, mu1, rtol=1e-8, atol=1e-10)
    assert np.allclose(model._gaussian.sigma, sig1, rtol=1e-8, atol=1e-10)
    assert np.allclose(model._self_gaussian.mu, mu2, rtol=1e-8, atol=1e-10)
    assert np.allclose(model._self_gaussian.sigma, sig2, rtol=1e-8, atol=1e-10)
    ref = sbm.gaussian_edge_log_likelihood_loops(A.astype(float), W, model.Mu, model.Sigma) if N <= 130 else None
    if ref is not None:
        assert np.isclose(model.log_likelihood((A, W)), ref, rtol=1e-9)


@pytest.mark.parametrize("special", [True, False])
@pytest.mark.parametrize("N,C", [(60, 3), (200, 8)])
def test_sbm_adjacency_sweep_matches_reference_loops(N, C, special):
    """adjacency.py:620-695 literally (the oracle's loop transcription) vs the device path."""
    from graphistician_b200.adjacency import SBMAdjacencyDistribution
    np.random.seed(5)
    truth = SBMAdjacencyDistribution(N, C=C, special_case_self_conns=special)
    A = truth.rvs().astype(np.uint8)
    model = SBMAdjacencyDistribution(N, C=C, special_case_self_conns=special)
    c0, m0 = model.c.copy(), model.m.copy()
    np.random.seed(9)
    model.resample_p([A])
    u = np.random.random(N)
    lp = model.resample_c(A, uniforms=u, want_lp=True)
    # reference: masks + loops
    np.random.seed(9)
    Af = A.astype(float)
    p = np.zeros((C, C))
    for c1 in range(C):
        for c2 in range(C):
            mask = (c0 == c1)[:, None] * (c0 == c2)[None, :]
            if special:
                mask &= ~np.eye(N, dtype=bool)
            p[c1, c2] = np.random.beta(1.0 + Af[mask].sum(), 1.0 + (1 - Af[mask]).sum())
    assert np.array_equal(model._last_counts[0], model._last_counts[0].astype(np.int64))
    assert np.allclose(model.p, p, rtol=1e-12)
    if special:
        d = np.diag(Af)
        assert model.p_self == np.random.beta(1.0 + d.sum(), 1.0 + (1 - d).sum())
    c_ref, _ = sbm.sbm_adjacency_resample_c_loops(Af, c0, m0, model.p, special, u)
    assert np.array_equal(model.c, c_ref)
    P = model.P
    ref = np.sum(Af * np.log(P) + (1 - Af) * np.log(1 - np.clip(P, 1e-64, 1 - 1e-64)))
    assert np.isclose(model.log_likelihood(A), ref, rtol=1e-10)


@pytest.mark.parametrize("special", [True, False])
@pytest.mark.parametrize("N,C,B", [(50, 3, 1), (120, 4, 2)])
def test_sbm_weights_sweep_matches_reference_loops(N, C, B, special):
    """weights.py:399-488 literally vs the device path (fp64 device data)."""
    from graphistician_b200.weights import SBMGaussianWeightDistribution
    np.random.seed(6)
    truth = SBMGaussianWeightDistribution(N, B=B, C=C, special_case_self_conns=special)
    W = truth.rvs()
    A = (np.random.rand(N, N) < 0.35).astype(np.uint8)
    model = SBMGaussianWeightDistribution(N, B=B, C=C, special_case_self_conns=special, dtype="f64")
    c0, m0 = model.c.copy(), model.m.copy()
    np.random.seed(13)
    model.resample_mu_and_Sig(A, W)
    n, sw, sww, d = model._last_stats
    for c1 in range(C):
        for c2 in range(C):
            mask = (c0 == c1)[:, None] * (c0 == c2)[None, :] & A.astype(bool)
            if special:
                mask &= ~np.eye(N, dtype=bool)
            assert n[c1, c2] == mask.sum()
            assert np.allclose(sw[c1, c2], W[mask].sum(0), rtol=1e-9, atol=1e-9)
            assert np.allclose(sww[c1, c2], W[mask].T @ W[mask], rtol=1e-9, atol=1e-9)
    u = np.random.random(N)
    model.resample_c(A, W, uniforms=u)
    c_ref, _ = sbm.sbm_weights_resample_c_loops(A.astype(float), W, c0, m0, model._Mu, model._Sigma, special, u)
    assert np.array_equal(model.c, c_ref)
    ref = sbm.gaussian_edge_log_likelihood_loops(A.astype(float), W, model.Mu, model.Sigma)
    assert np.isclose(model.log_likelihood((A, W)), ref, rtol=1e-9)
    assert np.isfinite(model.log_prior())


def test_factorized_network_config1_chain():
    """configs[0] (examples/demo_gaussian_er_gibbs.py: N=10, B=2, seed 1234, 100 sweeps): the device chain and the
    reference arithmetic (masks + NIW draws, same np.random stream) stay identical for the whole run."""
    from graphistician_b200.networks import FactorizedNetworkDistribution
    from graphistician_b200.adjacency import BetaBernoulliAdjacencyDistribution
    from graphistician_b200.weights import NIWGaussianWeightDistribution
    N, B = 10, 2
    np.random.seed(1234)
    true_model = FactorizedNetworkDistribution(N, BetaBernoulliAdjacencyDistribution, {},
                                               NIWGaussianWeightDistribution, {"B": B, "dtype": "f64"})
    A, W = true_model.rvs()
    A = A.astype(np.uint8)
    test_model = FactorizedNetworkDistribution(N, BetaBernoulliAdjacencyDistribution, {},
                                               NIWGaussianWeightDistribution, {"B": B, "dtype": "f64"})
    off = A.astype(bool) & ~np.eye(N, dtype=bool)
    on = A.astype(bool) & np.eye(N, dtype=bool)
    lps = []
    for it in range(100):
        st = np.random.get_state()
        test_model.resample((A, W))
        after = np.random.get_state()
        np.random.set_state(st)
        p = np.random.beta(1.0 + A.sum(), 1.0 + N * N - A.sum())
        mu1, sig1 = oniw.niw_resample(np.zeros(B), np.eye(B), 1.0, B + 2, data=W[off])
        mu2, sig2 = oniw.niw_resample(np.zeros(B), np.eye(B), 1.0, B + 2, data=W[on])
        assert test_model.adjacency.p == p
        assert np.allclose(test_model.weights._gaussian.mu, mu1, rtol=1e-9, atol=1e-12)
        assert np.allclose(test_model.weights._gaussian.sigma, sig1, rtol=1e-9, atol=1e-12)
        assert np.allclose(test_model.weights._self_gaussian.sigma, sig2, rtol=1e-9, atol=1e-12)
        assert all(np.array_equal(a, b) for a, b in zip(after, np.random.get_state()) if isinstance(a, np.ndarray))
        lps.append(test_model.log_probability((A, W)))
    assert np.all(np.isfinite(lps))
