"""GPU parity of the latent-distance path (K8 pair pass + device HMC) against the oracle, through the C ABI.

Protocol (north_star): the fp conditionals -- log probability and its gradients -- must agree to 1e-5 relative in
fp64 and 1e-3 in fp32; with injected momenta and log-uniforms an HMC transition must take the same accept decision
and land on the same state."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _problem(N, dim, seed, diag=True):
    from oracle import latent
    rng = np.random.RandomState(seed)
    L = np.sqrt(2.0) * rng.randn(N, dim)
    mu0, mus = 0.4, -0.3
    P = 1.0 / (1.0 + np.exp(-latent.logit_matrix(L, mu0, mus)))
    A = (rng.rand(N, N) < P).astype(np.uint8)
    if not diag:
        np.fill_diagonal(A, 0)
    return A, L, mu0, mus


def _state(A, L, mu0, mus, sigma, dtype):
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceLatent
    net = DeviceNetwork(A, None, dtype=dtype)
    lat = DeviceLatent(net, L.shape[1])
    lat.set_params(L, mu0, mus, sigma)
    return net, lat


def _relerr(a, b):
    return np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300)


@pytest.mark.parametrize("N,dim", [(5, 2), (37, 1), (128, 2), (129, 3), (300, 2), (513, 2), (256, 3)])
@pytest.mark.parametrize("dtype,tol", [("f64", 1e-9), ("f32", 1e-3)])
def test_lp_and_gradients_match_oracle(N, dim, dtype, tol):
    from oracle import latent
    A, L, mu0, mus = _problem(N, dim, 3 + N)
    net, lat = _state(A, L, mu0, mus, 2.0, dtype)
    ll, lp, gL, g0, gs = lat.lp_grad()
    rlp, rgL, rg0, rgs = latent.log_probability_and_grad(L, mu0, mus, A.astype(float), 2.0)
    rll = rlp + 0.5 * np.sum(L * L) / 2.0 + 0.5 * mu0 ** 2 + 0.5 * mus ** 2
    assert abs(lp - rlp) <= tol * abs(rlp)
    assert abs(ll - rll) <= tol * abs(rll)
    assert _relerr(gL, rgL) <= tol
    assert abs(g0 - rg0) <= tol * max(abs(rg0), np.abs(A - 0.5).sum() * 1e-3)
    assert abs(gs - rgs) <= tol * max(abs(rgs), 1.0)
    # the literal (unstable) form of the reference agrees wherever it is finite
    lit = latent.hmc_log_probability(L, mu0, mus, A.astype(float), 2.0)
    if np.isfinite(lit):
        assert abs(lp - lit) <= max(tol, 1e-9) * abs(lit)


@pytest.mark.parametrize("dtype,tol", [("f64", 1e-9), ("f32", 2e-3)])
@pytest.mark.parametrize("target", [0, 1])
def test_hmc_transition_matches_oracle(dtype, tol, target):
    from oracle import latent
    N, dim, sigma = 200, 2, 2.0
    A, L, mu0, mus = _problem(N, dim, 77)
    rng = np.random.RandomState(5)
    for trial, log_u in enumerate([-1e9, -0.05, 0.5]):            # always accept / data dependent / always reject
        net, lat = _state(A, L, mu0, mus, sigma, dtype)
        Af = A.astype(float)
        if target == 0:
            p0 = rng.randn(N, dim)
            U = lambda q: latent.log_probability_and_grad(q, mu0, mus, Af, sigma)[0]
            dU = lambda q: latent.log_probability_and_grad(q, mu0, mus, Af, sigma)[1]
            q_ref = latent.hmc(U, dU, 0.01, 10, L.copy(), (p0, log_u), negative_log_prob=False)
            acc, ratio = lat.hmc(0, 0.01, 10, p0, log_u)
            q_gpu = lat.get_params()[0]
            moved_ref = not np.array_equal(q_ref, L)
        else:
            p0 = rng.randn()
            U = lambda m: latent.log_probability_and_grad(L, float(m), mus, Af, sigma)[0]
            dU = lambda m: np.array(latent.log_probability_and_grad(L, float(m), mus, Af, sigma)[2])
            q_ref = latent.hmc(U, dU, 0.005, 10, np.array(mu0), (np.array(p0), log_u), negative_log_prob=False)
            acc, ratio = lat.hmc(1, 0.005, 10, p0, log_u)
            q_gpu = lat.get_params()[1]
            moved_ref = float(q_ref) != mu0
        assert np.isfinite(ratio)
        if trial != 1:
            assert acc == moved_ref == (trial == 0)
        if acc == moved_ref:
            assert _relerr(q_gpu, q_ref) <= tol
        net.close()


def test_sharded_passes_sum_to_the_unsharded_one():
    """Row shards: the per-rank accumulators add up to the single-rank one."""
    import ctypes
    import torch
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceLatent
    N, dim = 384, 2
    A, L, mu0, mus = _problem(N, dim, 9)
    At = np.ascontiguousarray(A.T)
    net, lat = _state(A, L, mu0, mus, 2.0, "f64")
    full = lat.lp_grad()
    n = lat.acc_len()
    total = torch.zeros(n, dtype=torch.float64, device="cuda")
    diag = np.ascontiguousarray(np.diag(A))
    for r0, r1 in [(0, 128), (128, 384)]:
        snet = DeviceNetwork(A[r0:r1], None, dtype="f64", row0=r0, nrows=r1 - r0, At_rows=At[r0:r1], diagA=diag)
        slat = DeviceLatent(snet, dim)
        slat.set_params(L, mu0, mus, 2.0)
        acc = torch.zeros(n, dtype=torch.float64, device="cuda")
        slat.eval_pass(ctypes.c_void_p(acc.data_ptr()))
        snet.ctx.synchronize()
        total += acc
        last = slat
    torch.cuda.synchronize()
    out = last.eval_finish(ctypes.c_void_p(total.data_ptr()))
    assert abs(out[1] - full[1]) <= 1e-10 * abs(full[1])
    assert _relerr(out[2], full[2]) <= 1e-10


def test_model_class_resample_matches_oracle_chain():
    """LatentDistanceAdjacencyDistribution.resample(A) with injected draws follows the oracle's chain (fp64)."""
    from oracle import latent
    from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution
    N, dim = 150, 2
    A, L, mu0, mus = _problem(N, dim, 21)
    np.random.seed(0)
    model = LatentDistanceAdjacencyDistribution(N, dim=dim, sigma=2.0, dtype="f64")
    model.L = L.copy(); model.mu_0 = mu0; model.mu_self = mus
    rng = np.random.RandomState(4)
    rL, rmu0, rmus, rstep, rrate = L.copy(), mu0, mus, 0.01, 0.9
    Af = A.astype(float)
    for it in range(4):
        draws = dict(L=(rng.randn(N, dim), np.log(rng.rand())), mu_0=(np.array(rng.randn()), np.log(rng.rand())),
                     mu_self=(np.array(rng.randn()), np.log(rng.rand())))
        rL, rmu0, rmus, rstep, rrate = latent.resample(rL, rmu0, rmus, Af, 2.0, rstep, rrate, draws)
        model.resample(A, draws=draws)
        assert _relerr(model.L, rL) <= 1e-8
        assert abs(model.mu_0 - rmu0) <= 1e-8 and abs(model.mu_self - rmus) <= 1e-8
        assert abs(model._L_step_sz - rstep) <= 1e-15 and abs(model._L_accept_rate - rrate) <= 1e-15
    ll = model.log_likelihood(A)
    ref = latent.log_probability_stable(model.L, model.mu_0, model.mu_self, Af, 2.0) \
        + 0.5 * np.sum(model.L ** 2) / 2.0 + 0.5 * model.mu_0 ** 2 + 0.5 * model.mu_self ** 2
    assert abs(ll - ref) <= 1e-9 * abs(ref)
