"""The factorised weighted models of the old API on the GPU: GaussianErdosRenyiFixedSparsity (BASELINE config 1: the model
examples/demo_gaussian_er_gibbs.py builds, N = 10, B = 2, seed 1234, 100 sweeps) and GaussianWeightedEigenmodel
(eigenmodel.py:1319-1416, the model of examples/demo_logistic_eigenmodel_gibbs.py)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _host_edge_stats(A, W):
    e = np.asarray(A) != 0
    w = np.asarray(W, dtype=np.float64)[e]
    return int(e.sum()), w.sum(0), w.T.dot(w)


def test_config1_demo_gaussian_er_gibbs():
    import graphistician_b200 as g
    from oracle import niw
    np.random.seed(1234)
    N, B = 10, 2
    true_model = g.GaussianErdosRenyiFixedSparsity(N, B)
    network = true_model.rvs()
    test_model = g.GaussianErdosRenyiFixedSparsity(N, B, dtype="f64")
    n, sw, sww = _host_edge_stats(network.A, network.W)
    lps = [test_model.log_probability(network)]
    mus = []
    for _ in range(100):
        test_model.resample(network)
        lps.append(test_model.log_probability(network))
        mus.append(test_model.weight_dist.mu.copy())
    assert np.all(np.isfinite(lps))
    # log_probability == Bernoulli(p = 0.5) term + Gaussian term over the edges, computed on the host
    wd = test_model.weight_dist
    ll_ref = N * N * np.log(0.5) + niw.gaussian_log_likelihood(np.asarray(network.W)[np.asarray(network.A) != 0], wd.mu, wd.sigma).sum()
    assert np.isclose(lps[-1], ll_ref + test_model.log_prior(), rtol=1e-10)
    # the NIW posterior the chain targets has the oracle's hyper-parameters (device statistics == host statistics)
    from graphistician_b200.weighted_models import _edge_stats
    dn, dsw, dsww = _edge_stats(network, "f64")
    assert dn == n and np.allclose(dsw, sw, rtol=1e-12) and np.allclose(dsww, sww, rtol=1e-12)
    mu_n, sig_n, kap_n, nu_n = niw.niw_posterior(wd.mu_0, wd.sigma_0, wd.kappa_0, wd.nu_0, n, sw, sww)
    post_mean = np.mean(mus, 0)
    se = np.sqrt(np.diag(sig_n) / max(nu_n - B - 1, 1) / kap_n / len(mus))
    assert np.all(np.abs(post_mean - mu_n) < 6 * se + 1e-3)
    # mean-field half of the same class (demo_gaussian_er_mf.py): one update lands on the exact NIW posterior
    test_model.meanfieldupdate(network)
    assert np.allclose(wd.mu_mf, mu_n, rtol=1e-10) and np.isclose(wd.kappa_mf, kap_n) and np.isclose(wd.nu_mf, nu_n)
    test_model.resample_from_mf()
    assert np.isfinite(test_model.get_vlb())


def test_gaussian_weighted_eigenmodel():
    import graphistician_b200 as g
    from oracle import niw
    np.random.seed(7)
    N, B, D = 48, 2, 2
    true_model = g.GaussianWeightedEigenmodel(N, B, D=D, p=0.2, sigma_F=3.0, lmbda=np.ones(D))
    network = true_model.rvs()
    model = g.GaussianWeightedEigenmodel(N, B, D=D, p=0.2, sigma_F=3.0, lmbda=np.ones(D), dtype="f64")
    assert model.F.shape == (N, D) and model.Mu.shape == (N, N, B) and model.Sigma.shape == (N, N, B, B)
    lp0 = model.log_probability(network)
    for _ in range(30):
        model.resample(network)
    lp1 = model.log_probability(network)
    assert np.isfinite(lp0) and np.isfinite(lp1) and lp1 > lp0
    wd = model.weight_dist
    e = np.asarray(network.A) != 0
    ll_w = niw.gaussian_log_likelihood(np.asarray(network.W)[e], wd.mu, wd.sigma).sum()
    ll_a = model.adjacency_dist.log_likelihood(network)
    assert np.isclose(model.log_likelihood(network), ll_a + ll_w, rtol=1e-10)
    model.meanfieldupdate(network)
    model.resample_from_mf()
    assert np.all(np.isfinite(model.F)) and np.isfinite(model.get_vlb())
