"""SURVEY.md 8f rank 4 variants of the weight models on the GPU, through the C ABI:

* LatentDistanceGaussianWeightDistribution (weights.py:613-901): k_latw_rows value / gradient / statistics against the
  reference-executed fixture tests/golden/ref_latent_weights.npz and the oracle; HMC transitions with injected draws;
  the conjugate b / Sigma moves; a chain on planted data.
* SBMGaussianWeightSharedCov (weights.py:501-611): the conditionals handed to the third-party draws and the residual
  statistics of the covariance update against tests/golden/ref_shared_cov.npz."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _ldw(g, dtype, L=None, b=None):
    from graphistician_b200.weights import LatentDistanceGaussianWeightDistribution
    N, dim = g["L"].shape
    np.random.seed(9)
    m = LatentDistanceGaussianWeightDistribution(N, B=1, dim=dim, b=float(g["b"]) if b is None else b, eta=float(g["eta"]),
                                                 dtype=dtype)
    m.L = (g["L"] if L is None else L).copy()
    m.cov.sigma = np.array([[float(g["sigma_w"])]])
    m._self_gaussian.mu = np.array([float(g["self_mu"])]); m._self_gaussian.sigma = np.array([[float(g["self_sigma"])]])
    return m


@pytest.mark.parametrize("dtype,tol", [("f64", 1e-10), ("f32", 2e-4)])
def test_latent_weight_lp_and_gradient(dtype, tol):
    g = np.load(os.path.join(GOLD, "ref_latent_weights.npz"))
    A, W = g["A"], g["W"]
    for i in range(len(g["eval_lp"])):
        m = _ldw(g, dtype, L=g["eval_L"][i], b=float(g["eval_b"][i]))
        lp, gL, gb = m.hmc_log_probability_and_grad(A, W)
        assert np.isclose(lp, g["eval_lp"][i], rtol=tol), (i, lp, g["eval_lp"][i])
        scale = np.abs(g["eval_gL"][i]).max()
        assert np.allclose(gL, g["eval_gL"][i], rtol=tol, atol=tol * scale), i
        assert np.isclose(gb, g["eval_gb"][i], rtol=tol, atol=tol * 10)
    m = _ldw(g, dtype)
    assert np.isclose(m.log_likelihood((A, W)), float(g["ll"]), rtol=max(tol, 1e-9))
    assert m.log_prior() == float(g["log_prior"])          # -inf in the reference too: invgamma.logpdf(eta, 1, 1) is (a, loc)


def test_latent_weight_moves_match_the_reference():
    g = np.load(os.path.join(GOLD, "ref_latent_weights.npz"))
    A, W = g["A"], g["W"]
    m = _ldw(g, "f64")
    dev = m._device_state(A, W)
    for s in range(len(g["hmc_logu"])):
        m._resample_L(dev, (g["hmc_pL"][s], float(g["hmc_logu"][s])))
        assert np.allclose(m.L, g["hmc_L"][s], rtol=1e-8, atol=1e-10), s
    stats = dev.weight_stats()
    np.random.seed(int(g["b_seed"]))
    delta = m._resample_b(stats)
    assert np.isclose(m.b, float(g["b_new"]), rtol=1e-10)
    n, sr, srr = stats
    assert int(round(n)) == int(g["cov_n"])
    assert np.isclose(srr - 2 * delta * sr + n * delta * delta, float(g["cov_S"]), rtol=1e-9)
    np.random.seed(int(g["pred_seed"]))
    Mr, Mc, Lr, Lc = m.sample_predictive_parameters()
    assert np.allclose(Mr, g["pred_Murow"], rtol=1e-10) and np.allclose(Lc, g["pred_Lcol"], rtol=1e-10)


@pytest.mark.parametrize("N,dim", [(200, 2), (1000, 3), (130, 1)])
def test_latent_weight_kernel_vs_oracle_ragged_sizes(N, dim):
    from oracle import latent as lt
    from graphistician_b200.weights import LatentDistanceGaussianWeightDistribution
    rng = np.random.RandomState(N)
    L = rng.randn(N, dim)
    A = (rng.rand(N, N) < 0.3).astype(np.uint8)
    D2 = ((L[:, None, :] - L[None, :, :]) ** 2).sum(2)
    W = (-D2 + 0.5 + 0.7 * rng.randn(N, N))[:, :, None]
    for dtype, tol in (("f64", 1e-10), ("f32", 5e-4)):
        np.random.seed(1)
        m = LatentDistanceGaussianWeightDistribution(N, B=1, dim=dim, b=0.3, eta=1.5, dtype=dtype)
        m.L = L.copy(); m.cov.sigma = np.array([[0.49]])
        lp, gL, gb = m.hmc_log_probability_and_grad(A, W if dtype == "f64" else W.astype(np.float32))
        lp_ref, gL_ref, gb_ref = lt.weight_log_probability_and_grad(L, 0.3, A, W[:, :, 0], 0.49, 1.5)
        assert np.isclose(lp, lp_ref, rtol=tol)
        assert np.allclose(gL, gL_ref, rtol=tol, atol=tol * np.abs(gL_ref).max())
        assert np.isclose(gb, gb_ref, rtol=tol, atol=tol * 10)
        assert np.allclose(m._device_state(A, W).weight_stats(), lt.weight_stats(L, 0.3, A, W[:, :, 0]), rtol=tol)


def test_latent_weight_chain_recovers_the_offset_and_noise():
    """examples/demo_gaussian_distance_model.py in small: complete graph, weights from the model; the chain's b and
    Sigma approach the planted values and the log probability rises."""
    from graphistician_b200.adjacency import CompleteAdjacencyDistribution
    from graphistician_b200.weights import LatentDistanceGaussianWeightDistribution
    from graphistician_b200.networks import FactorizedNetworkDistribution
    np.random.seed(1234)
    N = 30
    kw = dict(dim=2, b=-2.0)
    true = FactorizedNetworkDistribution(N, CompleteAdjacencyDistribution, {}, LatentDistanceGaussianWeightDistribution, kw)
    r = 1.5 + np.arange(N) // (N / 2.)
    th = np.linspace(0, 4 * np.pi, N, endpoint=False)
    true.weights.L = np.hstack(((r * np.cos(th))[:, None], (r * np.sin(th))[:, None]))
    true.weights.cov.sigma = np.array([[0.25]])
    A, W = true.rvs()
    test = FactorizedNetworkDistribution(N, CompleteAdjacencyDistribution, {}, LatentDistanceGaussianWeightDistribution,
                                         dict(dim=2, b=-2.0, dtype="f64"))
    lps = [test.weights.log_likelihood((A, W))]
    for it in range(600):
        test.resample((A, W))
        lps.append(test.weights.log_likelihood((A, W)))
    assert np.all(np.isfinite(lps)) and lps[-1] > lps[0] + 100
    d = lambda X: ((X[:, None, :] - X[None, :, :]) ** 2).sum(2)
    err0 = np.sqrt(np.mean((d(np.zeros((N, 2))) - d(true.weights.L)) ** 2))
    err = np.sqrt(np.mean((d(test.weights.L) - d(true.weights.L)) ** 2))
    assert err < 0.5 * err0, (err, err0)


@pytest.mark.parametrize("sp", [1, 0])
def test_shared_cov_conditionals_match_the_reference(sp):
    from graphistician_b200.weights import SBMGaussianWeightSharedCov
    g = np.load(os.path.join(GOLD, "ref_shared_cov.npz"))
    A, W, c, C = g["A"], g["W"], g["c"], 3
    N, B = A.shape[0], W.shape[2]
    np.random.seed(7)
    m = SBMGaussianWeightSharedCov(N, B=B, C=C, special_case_self_conns=bool(sp), dtype="f64")
    m.c = c.astype(int).copy(); m.m = g["m"].copy()
    m._cov_model.sigma = g["shared"].copy()
    for a in range(C):
        for b in range(C):
            m._gaussians[a][b].mu = g["mus"][a, b].copy(); m._gaussians[a][b].sigma = g["shared"].copy()
    assert np.isclose(m.log_likelihood((A, W)), float(g["ll_%d" % sp]), rtol=1e-10)
    assert np.isclose(m.log_prior(), float(g["log_prior_%d" % sp]), rtol=1e-10)
    # conditionals of the cell means: what the reference handed to np.random.multivariate_normal
    rec = []
    orig = np.random.multivariate_normal

    def spy(mean, cov, *a, **k):
        rec.append((np.array(mean), np.array(cov)))
        return orig(mean, cov, *a, **k)

    np.random.multivariate_normal = spy
    try:
        np.random.seed(71 + sp)
        m.resample_mu_and_Sig(A, W)
    finally:
        np.random.multivariate_normal = orig
    assert np.allclose([r[0] for r in rec[:C * C]], g["mvn_mean_%d" % sp], rtol=1e-9, atol=1e-12)
    assert np.allclose([r[1] for r in rec[:C * C]], g["mvn_cov_%d" % sp], rtol=1e-9, atol=1e-12)
    # the means are drawn in the same order from the same stream -> identical; so are the residual statistics
    mus_new = np.array([[m._gaussians[a][b].mu for b in range(C)] for a in range(C)])
    assert np.allclose(mus_new, g["mus_new_%d" % sp], rtol=1e-9)
    Aoff = A.astype(bool) & ~np.eye(N, dtype=bool)
    resid = (W - m.Mu)[Aoff]
    assert resid.shape[0] == int(g["cov_n_%d" % sp]) and np.allclose(resid.T.dot(resid), g["cov_S_%d" % sp], rtol=1e-9)
    assert np.allclose(m._Sigma, m.shared_sigma[None, None])


def test_shared_cov_chain_recovers_planted_covariance():
    from graphistician_b200.weights import SBMGaussianWeightSharedCov
    rng = np.random.RandomState(5)
    N, C, B = 96, 3, 2
    c = rng.randint(0, C, N)
    mus = 2.0 * rng.randn(C, C, B)
    Sig = np.array([[0.5, 0.2], [0.2, 0.3]])
    A = (rng.rand(N, N) < 0.5).astype(np.uint8)
    W = mus[np.ix_(c, c)] + rng.multivariate_normal(np.zeros(B), Sig, size=(N, N))
    np.random.seed(3)
    m = SBMGaussianWeightSharedCov(N, B=B, C=C, dtype="f64")
    m.c = c.copy()                                     # labels known: the test is about the (mu, Sigma) conditionals
    sig = []
    for it in range(200):
        m.resample_mu_and_Sig(A, W)
        m.resample_m()
        if it >= 50:
            sig.append(m.shared_sigma.copy())
    assert np.allclose(np.mean(sig, 0), Sig, atol=0.03)
    assert np.allclose(m._Mu, mus, atol=0.15)
    m.resample((A, W))                                 # the whole sweep (labels included) runs
    assert m.c.shape == (N,)
