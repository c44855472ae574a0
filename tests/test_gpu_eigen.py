"""GPU parity of the logistic eigenmodel path (K3-K6, K9) against the oracle, through the C ABI.

Protocol (north_star): inject Omega and the standard normals; the conditionals (prec_n, h_n, the mu_0 and lmbda
posteriors) must agree to 1e-5 relative with fp64 device storage and 1e-3 with fp32; the Polya-gamma variates are
compared distributionally."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _problem(N, D, seed, p_scale=0.6):
    from oracle import eigenmodel as eg
    rng = np.random.RandomState(seed)
    F = p_scale * rng.randn(N, D)
    lm = rng.randn(D)
    mu0 = -0.7
    A = (rng.rand(N, N) < eg.logistic_P(F, mu0, lm)).astype(np.uint8)
    O = rng.gamma(2.0, 0.1, (N, N)) + 0.02
    z = rng.randn(N, D)
    return A, O, F, lm, mu0, z


def _run(A, O, F, lm, mu0, z, sigma_F, dtype):
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    net = DeviceNetwork(A, None, dtype=dtype)
    eig = DeviceEigenmodel(net, F.shape[1])
    eig.set_params(F, mu0, lm)
    eig.set_S(O + O.T)
    out = eig.resample_F(sigma_F, z, want_conditionals=True)
    return net, eig, out


@pytest.mark.parametrize("N,D", [(12, 3), (200, 2), (333, 8), (130, 1), (64, 5)])
def test_feature_sweep_matches_oracle_f64(N, D):
    from oracle import eigenmodel as eg
    A, O, F, lm, mu0, z = _problem(N, D, 11 + N)
    S, R = eg.logistic_sym(A.astype(float), O, mu0)
    F_ref, prec_ref, h_ref = eg.resample_F(S, R, F, lm, 1.3, z)
    net, eig, (F_gpu, st, prec, h) = _run(A, O, F, lm, mu0, z, 1.3, "f64")
    scale_p = np.abs(prec_ref).max(); scale_h = np.abs(h_ref).max()
    assert np.allclose(prec, prec_ref, rtol=1e-5, atol=1e-9 * scale_p)
    assert np.allclose(h, h_ref, rtol=1e-5, atol=1e-9 * scale_h)
    assert np.allclose(F_gpu, F_ref, rtol=1e-6, atol=1e-8)
    # K6 pair statistics -> mu_0 and lmbda conditionals (eigenmodel.py:910-965)
    pp, ph = eg.logistic_mu0_conditional(A.astype(float), O, F_ref, lm, -1.2, 0.9)
    ph_gpu = -1.2 / 0.9 + (A.sum() - 0.5 * N * N) - lm.dot(st["hS"])
    assert np.isclose(ph_gpu, ph, rtol=1e-8, atol=1e-8 * abs(A.sum()))
    S2, R2 = eg.logistic_sym(A.astype(float), O, 0.37)
    lp_ref, lh_ref = eg.lmbda_conditional(S2, R2, F_ref, np.zeros(D), 2.0)
    lp_gpu = np.eye(D) / 2.0 + st["Pq"]
    lh_gpu = st["hK"] - 0.37 * st["hS"]
    assert np.allclose(lp_gpu, lp_ref, rtol=1e-8, atol=1e-9 * np.abs(lp_ref).max())
    assert np.allclose(lh_gpu, lh_ref, rtol=1e-8, atol=1e-9 * max(1.0, np.abs(lh_ref).max()))


def test_feature_sweep_matches_golden_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "eigen_small.npz"))
    net, eig, (F_gpu, st, prec, h) = _run(g["A"], g["O"], g["F"], g["lmbda"], float(g["mu_0"]), g["zF"],
                                          float(g["sigma_F"]), "f64")
    assert np.allclose(prec, g["prec_logistic"], rtol=1e-9)
    assert np.allclose(h, g["h_logistic"], rtol=1e-9, atol=1e-12)
    assert np.allclose(F_gpu, g["F_logistic"], rtol=1e-8, atol=1e-10)
    N = g["A"].shape[0]
    ph = -1.2 / 0.9 + (g["A"].sum() - 0.5 * N * N) - g["lmbda"].dot(st["hS"])
    assert np.isclose(ph, float(g["mu0_h"]), rtol=1e-9)
    lam_prec = np.eye(3) / 2.0 + st["Pq"]
    lam_h = st["hK"] - float(g["mu_0"]) * st["hS"]
    assert np.allclose(lam_prec, g["lam_prec"], rtol=1e-9)
    assert np.allclose(lam_h, g["lam_h"], rtol=1e-9, atol=1e-12)
    eig.set_params(g["F"], float(g["mu_0"]), g["lmbda"])
    assert np.isclose(eig.log_likelihood(0), float(g["ll_logistic"]), rtol=1e-12)
    assert np.isclose(eig.log_likelihood(1), float(g["ll_probit"]), rtol=1e-12)


@pytest.mark.parametrize("N,D,chol64", [(300, 8, 0), (257, 2, 0), (300, 8, 1)])
def test_feature_sweep_f32_storage(N, D, chol64):
    from oracle import eigenmodel as eg
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    A, O, F, lm, mu0, z = _problem(N, D, 5 + N)
    S, R = eg.logistic_sym(A.astype(float), O, mu0)
    F_ref, prec_ref, h_ref = eg.resample_F(S, R, F, lm, 1.0, z)
    net = DeviceNetwork(A, None, dtype="f32")
    eig = DeviceEigenmodel(net, D)
    eig.set_option(0, chol64)                     # GK_EIG_OPT_CHOL_F64
    eig.set_params(F, mu0, lm)
    eig.set_S(O + O.T)
    F_gpu, st, prec, h = eig.resample_F(1.0, z, want_conditionals=True)
    assert np.allclose(prec, prec_ref, rtol=1e-3, atol=1e-4 * np.abs(prec_ref).max())
    assert np.allclose(h, h_ref, rtol=1e-3, atol=1e-4 * np.abs(h_ref).max())
    assert np.allclose(F_gpu, F_ref, rtol=1e-2, atol=2e-3)


@pytest.mark.parametrize("dtype,N,D", [("f64", 450, 8), ("f32", 700, 3), ("f64", 129, 2)])
def test_feature_sweep_push_overlap_is_equivalent(dtype, N, D):
    """GK_EIG_OPT_OVERLAP_PUSH (second-stream push under the next block's serial kernel) and
    GK_EIG_OPT_FAST_SWEEP = 0 (slab form) give the sweep of the plain single-stream running-G form."""
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    A, O, F, lm, mu0, z = _problem(N, D, 77 + N)
    outs = []
    for fast, overlap in ((1, 1), (1, 0), (0, 0)):
        net = DeviceNetwork(A, None, dtype=dtype)
        eig = DeviceEigenmodel(net, D)
        eig.set_option(1, fast)                    # GK_EIG_OPT_FAST_SWEEP
        eig.set_option(2, overlap)                 # GK_EIG_OPT_OVERLAP_PUSH
        eig.set_params(F, mu0, lm)
        eig.set_S(O + O.T)
        for _ in range(2):                         # two sweeps: the second starts from the first one's F
            F_gpu, st, prec, h = eig.resample_F(1.1, z, want_conditionals=True)
        outs.append((F_gpu, st["Pq"], st["hS"], st["hK"], prec, h))
    tol = dict(rtol=1e-8, atol=1e-9) if dtype == "f64" else dict(rtol=2e-2, atol=5e-3)
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            assert np.allclose(a, b, **tol)


def test_pg_draw_symmetric_and_moments():
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    from oracle import eigenmodel as eg, pg
    N, D = 700, 4
    rng = np.random.RandomState(3)
    F = rng.randn(N, D); lm = np.array([1.5, -1.0, 0.5, 2.0]); mu0 = -1.0
    A = (rng.rand(N, N) < 0.1).astype(np.uint8)
    net = DeviceNetwork(A, None, dtype="f32")
    eig = DeviceEigenmodel(net, D)
    eig.set_params(F, mu0, lm)
    tot = eig.pg_draw(1234)
    S = eig.get_S()
    assert np.array_equal(S, S.T) and np.all(np.diag(S) == 0) and np.all(S[~np.eye(N, dtype=bool)] > 0)
    Mu = eg.mu_matrix(F, mu0, lm)
    iu = np.triu_indices(N, 1)
    z = Mu[iu]; s = S[iu]
    mean, var = 2 * pg.pg1_mean(z), 2 * pg.pg1_var(z)
    # standardised residuals: mean 0, variance 1
    r = (s - mean) / np.sqrt(var)
    n = r.size
    assert abs(r.mean()) < 5.0 / np.sqrt(n)
    assert abs(r.var() - 1.0) < 0.05
    # binned by |z|: the mean must hold across the whole range of logits, not only on average
    az = np.abs(z)
    for lo, hi in [(0, 0.5), (0.5, 1.5), (1.5, 3), (3, 6), (6, 40)]:
        sel = (az >= lo) & (az < hi)
        if sel.sum() > 500:
            assert abs(r[sel].mean()) < 5.0 / np.sqrt(sel.sum()), (lo, hi, r[sel].mean())
    # sum over all N^2 draws = pair sums + N diagonal draws PG(1, mu_0)
    diag_sum = tot - s.sum()
    assert abs(diag_sum - N * pg.pg1_mean(mu0)) < 6 * np.sqrt(N * pg.pg1_var(mu0))
    # a second draw advances the counter
    eig.pg_draw(1234)
    assert not np.array_equal(eig.get_S(), S)


def test_pg_draw_ks_against_c_sampler():
    """Distribution of the device PG(1,z)+PG(1,z) variates vs the plain-C Devroye sampler at fixed z."""
    from scipy.stats import ks_2samp
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    from oracle import pg
    N = 512
    A = np.zeros((N, N), np.uint8)
    net = DeviceNetwork(A, None, dtype="f32")
    eig = DeviceEigenmodel(net, 1)
    for zval in [0.0, 0.8, 2.5, 7.0]:
        eig.set_params(np.zeros((N, 1)), zval, np.ones(1))       # F = 0 => Mu == mu_0 everywhere
        eig.pg_draw(77)
        S = eig.get_S()
        s = S[np.triu_indices(N, 1)][:60000]
        ref = pg.pgdraw(np.full(60000, zval), seed=5) + pg.pgdraw(np.full(60000, zval), seed=6)
        stat, pval = ks_2samp(s, ref)
        assert pval > 1e-4, (zval, stat, pval)


def test_model_resample_matches_oracle_sweep():
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    from oracle import eigenmodel as eg
    N, D = 150, 3
    A, O, F, lm, mu0, z = _problem(N, D, 99)
    np.random.seed(4)
    model = LogisticEigenmodel(N, D, p=0.2, sigma_F=1.5, sigma_mu0=0.8, dtype="f64")      # lmbda inferred
    model.F, model.mu_0, model.lmbda = F.copy(), mu0, lm.copy()
    np.random.seed(10)
    model.resample(A, omega=O, normals=z, want_conditionals=True)
    hyp = dict(sigma_F=1.5, mu_mu_0=model.mu_mu_0, sigma_mu0=0.8, lmbda_given=False, mu_lmbda=np.zeros(D),
               sigma_lmbda=1.0)
    # replay the host draws the model made after the F sweep: np.random.normal for mu_0 then multivariate_normal
    st, cond = eg.logistic_gibbs_sweep(A.astype(float), O, dict(F=F, mu_0=mu0, lmbda=lm), hyp, z, 0.0,
                                       np.zeros(D))
    lc = model._last_conditionals
    assert np.allclose(lc["prec"], cond["prec"], rtol=1e-5)
    assert np.allclose(model.F, st["F"], rtol=1e-6, atol=1e-8)
    assert np.allclose(lc["mu0"], cond["mu0"], rtol=1e-8)
    np.random.seed(10)
    mu0_new = np.random.normal(cond["mu0"][1] / cond["mu0"][0], np.sqrt(1.0 / cond["mu0"][0]))
    assert np.isclose(model.mu_0, mu0_new, rtol=1e-9)
    S2, R2 = eg.logistic_sym(A.astype(float), O, mu0_new)
    lp_ref, lh_ref = eg.lmbda_conditional(S2, R2, st["F"], np.zeros(D), 1.0)
    assert np.allclose(lc["lmbda"][0], lp_ref, rtol=1e-7)
    assert np.allclose(lc["lmbda"][1], lh_ref, rtol=1e-7, atol=1e-9)
    # log likelihood through the model API
    ll = model.log_likelihood(A)
    assert np.isclose(ll, eg.log_likelihood(A.astype(float), eg.logistic_P(model.F, model.mu_0, model.lmbda)), rtol=1e-10)


def test_model_runs_with_device_rng():
    """Full sweep with device-side Polya-gamma draws: finite state, reproducible under np.random.seed."""
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    from oracle import eigenmodel as eg
    N, D = 260, 2
    rng = np.random.RandomState(0)
    Ft = rng.randn(N, D)
    A = (rng.rand(N, N) < eg.logistic_P(Ft, -1.0, np.ones(D))).astype(np.uint8)
    outs = []
    for rep in range(2):
        np.random.seed(42)
        model = LogisticEigenmodel(N, D, p=0.3, lmbda=np.ones(D))
        for _ in range(5):
            model.resample(A)
        assert np.all(np.isfinite(model.F)) and np.isfinite(model.mu_0)
        outs.append((model.F.copy(), model.mu_0))
    # the slab partials are combined with fp64 atomics: summation order may differ in the last bits
    assert np.allclose(outs[0][0], outs[1][0], rtol=1e-6, atol=1e-9) and np.isclose(outs[0][1], outs[1][1], rtol=1e-9)
    ll0 = LogisticEigenmodel(N, D, p=0.3, lmbda=np.ones(D)).log_likelihood(A)
    assert model.log_likelihood(A) > ll0 - 50.0


def test_invalidate_refills_the_device_copy_in_place():
    """network.invalidate() after an in-place edit of A: same device handle, same allocations, new contents -- including
    the derived A + A^T the eigenmodel kernels read (bench.py's e2e leg relies on this)."""
    from graphistician_b200.device import DeviceEigenmodel, DeviceNetwork
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    N, D = 300, 4
    A, O, F, lm, mu0, z = _problem(N, D, 123)
    A2 = _problem(N, D, 456)[0]
    net = FixedGaussianNetwork(A.copy(), None)
    dev = net.device("f64")
    eig = DeviceEigenmodel(dev, D)
    eig.set_params(F, mu0, lm); eig.set_S(O + O.T)
    first = eig.resample_F(1.0, z)[0]
    net.A[...] = A2
    net.invalidate()
    assert net.device("f64") is dev
    eig.set_params(F, mu0, lm); eig.set_S(O + O.T)
    again = eig.resample_F(1.0, z)[0]
    fresh_net = DeviceNetwork(A2, None, dtype="f64")
    fresh = DeviceEigenmodel(fresh_net, D)
    fresh.set_params(F, mu0, lm); fresh.set_S(O + O.T)
    ref = fresh.resample_F(1.0, z)[0]
    assert np.array_equal(again, ref) and not np.allclose(first, ref)
    c = np.random.RandomState(0).randint(0, 3, N)
    assert np.array_equal(dev.block_stats(c, 3)[0], fresh_net.block_stats(c, 3)[0])
