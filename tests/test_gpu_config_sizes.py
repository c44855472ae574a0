"""GPU parity at the BASELINE config sizes, through the C ABI, against the vectorised oracle.

* config 2 (Gaussian SBM, N = 4096, K = 16, B = 2): one full assignment sweep from the prior draw of the labels,
  every node's conditional and label against ``oracle.sbm.sbm_assign_sweep`` -- fp64 storage (labels bit-exact,
  lp 1e-5) and fp32 storage (lp 1e-3; labels equal unless the draw sat on a cdf boundary); K1 counts bit-exact.
* config 3 (logistic eigenmodel, D = 8) at N = 4096 with fp64 storage: every node's (prec_n, h_n) against
  ``oracle.eigenmodel.resample_F``.
* N = 16384 (the size every reported number is quoted on), fp32 storage, sampled: the K1 count table (exact) and
  weight sums, the conditionals of 512 random nodes of the SBM sweep, and (prec_n, h_n) of 256 random nodes of the
  eigenmodel sweep, each rebuilt by the oracle from the state the node saw when it was drawn (new values before it,
  old values after it).
Data generators are bench.py's."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))

pytestmark = pytest.mark.gpu


def _prior_state(N, K, seed=5678):
    rs = np.random.RandomState(seed)
    m = rs.dirichlet(np.ones(K))
    c = rs.choice(K, p=m, size=N)
    return m, c


def _boundary_flip(lp, u, tol):
    """True when the oracle's draw from lp with uniform u sits within tol of a cdf boundary (pybasicbayes sampler)."""
    cum = np.cumsum(np.exp(lp - lp.max()))
    return np.min(np.abs(u * cum[-1] - cum)) <= tol * cum[-1]


@pytest.mark.parametrize("dtype", ["f64", "f32"])
def test_sbm_config2_full_sweep(dtype):
    import bench
    from oracle import sbm
    from graphistician_b200 import DeviceNetwork
    N, K, B = 4096, 16, 2
    d = bench.make_sbm_data(N, K, B, pin=False)
    m, c = _prior_state(N, K)
    u = np.random.RandomState(3).rand(N)
    W = d["W"] if dtype == "f32" else d["W"].astype(np.float64)
    net = DeviceNetwork(d["A"], W, dtype=dtype)
    # K1 at the config size: integer counts bit-exact
    n_edge, sw, sww = net.block_stats(c, K)
    ref = sbm.block_pair_stats(d["A"].astype(float), d["W"].astype(np.float64), c, K)
    assert np.array_equal(n_edge, ref["n_edge"])
    tol_s = 1e-10 if dtype == "f64" else 2e-4
    assert np.allclose(sww, ref["sum_ww"], rtol=tol_s, atol=tol_s * np.abs(ref["sum_ww"]).max())
    # K2: the whole sweep
    c_gpu, lp_gpu, nch = net.sbm_sweep(c, np.log(m), d["p"], d["mus"], d["sigmas"], u, want_lp=True)
    c_ref, lp_ref = sbm.sbm_assign_sweep(d["A"].astype(float), d["W"].astype(np.float64), c, m, d["p"], d["mus"],
                                         d["sigmas"], u, sbm.SELF_OLD)
    tol = 1e-5 if dtype == "f64" else 1e-3
    if dtype == "f64":
        assert np.array_equal(c_gpu, c_ref)
        upto = N
    else:
        diff = np.flatnonzero(c_gpu != c_ref)
        upto = N if diff.size == 0 else int(diff[0])
        if diff.size:                                       # the first difference must be a boundary case of the draw
            assert _boundary_flip(lp_ref[upto], u[upto], 1e-3), "label of node %d differs away from a cdf boundary" % upto
            assert upto > N // 2
    scale = np.abs(lp_ref[:upto]).max()
    assert np.allclose(lp_gpu[:upto], lp_ref[:upto], rtol=tol, atol=tol * 1e-2 * scale)


def test_eigen_config3_conditionals_f64():
    import bench
    from oracle import eigenmodel as eg
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    N, D = 4096, 8
    d = bench.make_eigen_data(N, D, pin=False)
    rng = np.random.RandomState(5)
    F = rng.randn(N, D); lm = np.ones(D); mu0 = float(d["mu0"])
    O = rng.gamma(2.0, 0.05, (N, N)) + 0.01
    z = rng.randn(N, D)
    net = DeviceNetwork(d["A"], None, dtype="f64")
    eig = DeviceEigenmodel(net, D)
    eig.set_params(F, mu0, lm)
    eig.set_S(O + O.T)
    F_gpu, st, prec, h = eig.resample_F(1.0, z, want_conditionals=True)
    S, R = eg.logistic_sym(d["A"].astype(float), O, mu0)
    F_ref, prec_ref, h_ref = eg.resample_F(S, R, F, lm, 1.0, z)
    assert np.allclose(prec, prec_ref, rtol=1e-5, atol=1e-9 * np.abs(prec_ref).max())
    assert np.allclose(h, h_ref, rtol=1e-5, atol=1e-7 * np.abs(h_ref).max())
    assert np.allclose(F_gpu, F_ref, rtol=1e-5, atol=1e-7)


def test_sbm_n16384_f32_sampled():
    import bench
    from oracle import sbm
    from graphistician_b200 import DeviceNetwork
    N, K, B = 16384, 16, 2
    d = bench.make_sbm_data(N, K, B, pin=False)
    A, W = d["A"], d["W"]
    m, c = _prior_state(N, K)
    u = np.random.RandomState(3).rand(N)
    net = DeviceNetwork(A, W, dtype="f32")
    # K1: exact counts by a chunked one-hot contraction on the host (fp32 partial sums of 0/1 values below 2^24 are exact)
    n_edge, sw, sww = net.block_stats(c, K)
    oh = np.zeros((N, K), np.float32); oh[np.arange(N), c] = 1.0
    cnt = np.zeros((K, K), np.int64); s0 = np.zeros((K, K))
    for r0 in range(0, N, 1024):
        blk = A[r0:r0 + 1024].astype(np.float32)
        blk[np.arange(blk.shape[0]), np.arange(r0, r0 + blk.shape[0])] = 0.0       # off-diagonal pairs only
        cnt += np.rint(oh[r0:r0 + 1024].T.dot(blk.dot(oh))).astype(np.int64)
        s0 += oh[r0:r0 + 1024].T.astype(np.float64).dot((blk * W[r0:r0 + 1024, :, 0]).astype(np.float64).dot(oh.astype(np.float64)))
    assert np.array_equal(n_edge, cnt)
    assert np.allclose(sw[:, :, 0], s0, rtol=2e-4, atol=2e-4 * np.abs(s0).max())
    # K2: conditionals of 512 random nodes, rebuilt from the labels each node saw (new before it, old after it)
    c_gpu, lp_gpu, nch = net.sbm_sweep(c, np.log(m), d["p"], d["mus"], d["sigmas"], u, want_lp=True)
    gc = sbm.gaussian_coefficients(d["mus"], d["sigmas"])
    logm = np.log(m)
    nodes = np.sort(np.random.RandomState(9).choice(N, 512, replace=False))
    worst = 0.0
    for n in nodes:
        c_seen = np.concatenate([c_gpu[:n], c[n:]])
        lp = sbm.node_conditional(int(n), A, W, c_seen, logm, d["p"], gc, sbm.SELF_OLD)
        scale = np.abs(lp).max()
        worst = max(worst, np.max(np.abs(lp_gpu[n] - lp)) / scale)
        new = sbm.sample_discrete_from_log(lp, u[n])
        assert new == c_gpu[n] or _boundary_flip(lp, u[n], 1e-3), "node %d: label differs from the oracle's draw" % n
    assert worst <= 1e-3, worst


def test_eigen_n16384_f32_sampled():
    import bench
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    N, D = 16384, 8
    d = bench.make_eigen_data(N, D, pin=False)
    A = d["A"]
    rng = np.random.RandomState(5)
    F = rng.randn(N, D); lm = np.ones(D); mu0 = float(d["mu0"])
    z = rng.randn(N, D)
    net = DeviceNetwork(A, None, dtype="f32")
    eig = DeviceEigenmodel(net, D)
    eig.set_params(F, mu0, lm)
    eig.pg_draw(424242)                                   # S = Omega + Omega^T on the device (fp32 storage)
    S = eig.get_S()
    F_gpu, st, prec, h = eig.resample_F(1.0, z, want_conditionals=True)
    nodes = np.sort(np.random.RandomState(11).choice(N, 256, replace=False))
    llT = np.outer(lm, lm)
    for n in nodes:
        Fs = np.concatenate([F_gpu[:n], F[n:]])            # features node n saw: new before it, old after it
        s = S[n].copy(); s[n] = 0.0
        r = A[n].astype(np.float64) + A[:, n] - 1.0 - mu0 * s; r[n] = 0.0
        prec_ref = np.eye(D) + (Fs.T * s).dot(Fs) * llT    # eigenmodel.py:884-901 in contraction form (oracle.eigenmodel.resample_F)
        h_ref = r.dot(Fs) * lm
        assert np.allclose(prec[n], prec_ref, rtol=1e-3, atol=1e-3 * np.abs(prec_ref).max()), n
        assert np.allclose(h[n], h_ref, rtol=1e-3, atol=2e-3 * max(1.0, np.abs(h_ref).max())), n
