"""north_star correctness check 3: posterior moments of the GPU chains match the reference (oracle) chain within Monte
Carlo error.  For each model a long GPU chain (``model.resample(data)`` through the C ABI) and a long oracle chain
(oracle/*.py, CPU) are run on the same data from independent random streams; permutation / rotation invariant
summaries are compared with a two-sample z statistic whose standard errors come from non-overlapping batch means.

Chains mirrored: examples/demo_gaussian_sbm_gibbs.py:47-52 (SBM), examples/demo_logistic_eigenmodel_gibbs.py:53-57 and
test/test_logistic_eigenmodel_geweke.py:37-47 (eigenmodel), examples/demo_distance_model_gibbs.py (latent distance)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ZMAX = 4.0


def _bm(x, nb=20):
    x = np.asarray(x, dtype=float)
    n = len(x) // nb * nb
    b = x[:n].reshape(nb, -1).mean(1)
    return b.mean(), b.std(ddof=1) / np.sqrt(nb)


def _compare(name, xs_gpu, xs_ref):
    m1, s1 = _bm(xs_gpu)
    m2, s2 = _bm(xs_ref)
    z = (m1 - m2) / max(np.hypot(s1, s2), 1e-12)
    assert abs(z) < ZMAX, "%s: GPU chain %.4f +- %.4f vs oracle chain %.4f +- %.4f (z = %.2f)" % (name, m1, s1, m2, s2, z)


def test_sbm_posterior_moments():
    from oracle import sbm
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    N, C, B, T, burn = 64, 3, 1, 2000, 300
    rng = np.random.RandomState(1234)
    c_true = rng.randint(0, C, N)
    p_true = np.array([[0.55, 0.30, 0.25], [0.20, 0.50, 0.35], [0.30, 0.25, 0.60]])
    mu_true = 0.8 * rng.randn(C, C, B)
    A = (rng.rand(N, N) < p_true[np.ix_(c_true, c_true)]).astype(np.uint8)
    W = mu_true[np.ix_(c_true, c_true)] + rng.randn(N, N, B)
    same_true = c_true[:, None] == c_true[None, :]

    def summaries(c, p, mus):
        same = c[:, None] == c[None, :]
        sz = np.bincount(c, minlength=C)
        return [float((sz ** 2).sum()), float(sz.max()), float((same == same_true).mean()),
                float(p[np.ix_(c, c)].mean()), float(mus[np.ix_(c, c)][:, :, 0].mean())]

    names = ["sum of squared block sizes", "largest block", "pair agreement with the truth", "mean P_ij", "mean Mu_ij"]
    # GPU chain
    np.random.seed(11)
    model = GaussianStochasticBlockModel(N, B, C=C, dtype="f64")
    model.c = c_true.copy()                  # both chains start in the main mode (the merged-block states are metastable)
    net = FixedGaussianNetwork(A, W)
    gpu = []
    for it in range(T + burn):
        model.resample(net)
        if it >= burn:
            gpu.append(summaries(model.c, model.p, model.mus))
    # oracle chain (stochastic_block_model.py:206-230 restated), same hyperparameters, independent stream
    wm = model.weight_models[0][0]
    hyp = dict(pi=model.pi, tau0=model.tau0, tau1=model.tau1, niw=(wm.mu_0, wm.sigma_0, wm.kappa_0, wm.nu_0))
    rs = np.random.RandomState(22)
    m0 = rs.dirichlet(np.ones(C))
    state = dict(c=c_true.copy(), m=m0, p=rs.beta(1, 1, (C, C)), mus=np.zeros((C, C, B)),
                 sigmas=np.tile(np.eye(B), (C, C, 1, 1)))
    Af = A.astype(float)
    ref = []
    for it in range(T + burn):
        state, _ = sbm.gaussian_sbm_gibbs_sweep(Af, W, state, hyp, rs.rand(N), rng=rs)
        if it >= burn:
            ref.append(summaries(state["c"], state["p"], state["mus"]))
    gpu, ref = np.array(gpu), np.array(ref)
    for k, nm in enumerate(names):
        _compare("SBM " + nm, gpu[:, k], ref[:, k])


def test_logistic_eigenmodel_posterior_moments():
    from oracle import eigenmodel as eg, pg
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    N, D, T, burn = 32, 2, 3000, 300
    rng = np.random.RandomState(1234)
    lm = np.ones(D)
    F_true = rng.randn(N, D)
    A = (rng.rand(N, N) < eg.logistic_P(F_true, -0.5, lm)).astype(np.uint8)
    pairs = [(0, 1), (2, 5), (7, 30)]

    def summaries(F, mu0):
        G = F.dot(F.T)
        return [float(mu0), float((F * F).sum()), float(np.abs(G).mean())] + [float(G[i, j]) for i, j in pairs]

    names = ["mu_0", "||F||^2", "mean |F F^T|"] + ["(F F^T)[%d,%d]" % p for p in pairs]
    np.random.seed(5)
    model = LogisticEigenmodel(N, D, p=0.3, sigma_F=1.0, sigma_mu0=1.0, lmbda=lm.copy(), dtype="f64")
    gpu = []
    for it in range(T + burn):
        model.resample(A)
        if it >= burn:
            gpu.append(summaries(model.F, model.mu_0))
    hyp = dict(sigma_F=1.0, mu_mu_0=model.mu_mu_0, sigma_mu0=1.0, lmbda_given=True)
    rs = np.random.RandomState(6)
    state = dict(F=rs.randn(N, D), mu_0=model.mu_mu_0 + rs.randn(), lmbda=lm.copy())
    Af = A.astype(float)
    ref = []
    for it in range(T + burn):
        O = pg.pgdraw(eg.mu_matrix(state["F"], state["mu_0"], state["lmbda"]), seed=1000 + it)     # eigenmodel.py:859-874
        state, _ = eg.logistic_gibbs_sweep(Af, O, state, hyp, rs.randn(N, D), rs.randn())
        if it >= burn:
            ref.append(summaries(state["F"], state["mu_0"]))
    gpu, ref = np.array(gpu), np.array(ref)
    for k, nm in enumerate(names):
        _compare("eigenmodel " + nm, gpu[:, k], ref[:, k])


def test_latent_distance_posterior_moments():
    from oracle import latent
    from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution
    N, dim, T, burn = 32, 2, 4000, 1000
    rng = np.random.RandomState(1234)
    L_true = np.sqrt(2.0) * rng.randn(N, dim)
    P = 1.0 / (1.0 + np.exp(-latent.logit_matrix(L_true, 0.5, 0.0)))
    A = (rng.rand(N, N) < P).astype(np.uint8)
    pairs = [(0, 1), (3, 17), (8, 25)]

    def summaries(L, mu0, mus):
        d2 = ((L[:, None, :] - L[None, :, :]) ** 2).sum(2)
        return [float(mu0), float(mus), float((L * L).sum()), float(d2.mean())] + [float(d2[i, j]) for i, j in pairs]

    names = ["mu_0", "mu_self", "||L||^2", "mean squared distance"] + ["d^2[%d,%d]" % p for p in pairs]
    np.random.seed(3)
    model = LatentDistanceAdjacencyDistribution(N, dim=dim, sigma=2.0, dtype="f64")
    gpu = []
    for it in range(T + burn):
        model.resample(A)
        if it >= burn:
            gpu.append(summaries(model.L, model.mu_0, model.mu_self))
    rs = np.random.RandomState(4)
    L, mu0, mus, step, rate = np.sqrt(2.0) * rs.randn(N, dim), 0.0, 0.0, 0.01, 0.9
    Af = A.astype(float)
    ref = []
    for it in range(T + burn):
        draws = dict(L=(rs.randn(N, dim), np.log(rs.rand())), mu_0=(rs.randn(), np.log(rs.rand())),
                     mu_self=(rs.randn(), np.log(rs.rand())))
        L, mu0, mus, step, rate = latent.resample(L, mu0, mus, Af, 2.0, step, rate, draws)
        if it >= burn:
            ref.append(summaries(L, mu0, mus))
    gpu, ref = np.array(gpu), np.array(ref)
    for k, nm in enumerate(names):
        _compare("latent " + nm, gpu[:, k], ref[:, k])
