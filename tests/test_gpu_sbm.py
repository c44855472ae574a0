"""GPU parity (through the C ABI) of the SBM path against the oracle: K1 block-pair statistics and the
K2 sequential assignment sweep.  Integer counts and assignments bit-exact; conditionals 1e-5 relative in the
fp64 data mode, 1e-3 in fp32 (BASELINE.json north_star)."""
import os

import numpy as np
import pytest

from oracle import sbm

pytestmark = pytest.mark.gpu


def _problem(seed, N, K, B, dens=None):
    rng = np.random.RandomState(seed)
    c = rng.randint(0, K, N)
    m = rng.dirichlet(np.ones(K) * 3) if K > 1 else np.ones(1)
    p = rng.beta(1, 1, (K, K)) if dens is None else np.full((K, K), dens)
    mus = rng.randn(K, K, max(B, 1))
    sig = np.zeros((K, K, max(B, 1), max(B, 1)))
    for a in range(K):
        for b in range(K):
            X = rng.randn(max(B, 1), max(B, 1))
            sig[a, b] = X.dot(X.T) + 0.5 * np.eye(max(B, 1))
    A = (rng.rand(N, N) < p[np.ix_(c, c)]).astype(np.uint8)
    W = (rng.randn(N, N, B) + mus[np.ix_(c, c)][:, :, :B]) if B else None
    return A, W, c, m, p, mus[:, :, :B], sig[:, :, :B, :B], rng.rand(N)


def _net(A, W, dtype):
    from graphistician_b200 import DeviceNetwork
    Wd = None if W is None else (W.astype(np.float32) if dtype == "f32" else W)
    return DeviceNetwork(A, Wd, dtype=dtype), Wd


@pytest.mark.parametrize("N,K,B,dtype", [(50, 5, 2, "f64"), (200, 16, 2, "f32"), (333, 7, 1, "f64"), (210, 6, 3, "f64"), (300, 16, 4, "f32"), (190, 9, 4, "f64"),
                                         (129, 1, 2, "f32"), (64, 32, 0, "f32"), (700, 20, 2, "f64")])
def test_block_stats(N, K, B, dtype):
    A, W, c, m, p, mus, sig, u = _problem(N + K, N, K, B)
    net, Wd = _net(A, W, dtype)
    n_edge, sw, sww = net.block_stats(c, K)
    ref = sbm.block_pair_stats(A.astype(float), None if Wd is None else Wd.astype(np.float64), c, K)
    assert np.array_equal(n_edge, ref["n_edge"])                       # integer counts: bit-exact
    if B:
        tol = 1e-10 if dtype == "f64" else 2e-4
        assert np.allclose(sw, ref["sum_w"], rtol=tol, atol=tol * np.abs(ref["sum_w"]).max())
        assert np.allclose(sww, ref["sum_ww"], rtol=tol, atol=tol * np.abs(ref["sum_ww"]).max())
    d = net.diag_stats(c, K)
    for k in ("d_cnt", "d_edge", "d_sum_w", "d_sum_ww"):
        assert np.allclose(d[k], ref[k])


def _check_sweep(A, W, c, m, p, mus, sig, u, dtype, mode, gmode, beta=1.0, sampler="cumsum", block_nodes=0):
    from graphistician_b200 import _lib
    net, Wd = _net(A, W, dtype)
    W64 = None if Wd is None else Wd.astype(np.float64)
    c_ref, lp_ref = sbm.sbm_assign_sweep(A.astype(float), W64, c, m, p, mus, sig, u, mode, beta=beta, sampler=sampler)
    c_gpu, lp_gpu, nch = net.sbm_sweep(c, np.log(m), p, mus, sig, u, self_mode=gmode,
                                      sampler=_lib.SAMPLER_CUMSUM if sampler == "cumsum" else _lib.SAMPLER_LSE,
                                      beta=beta, block_nodes=block_nodes, want_lp=True)
    tol = 1e-5 if dtype == "f64" else 1e-3
    # conditionals are defined up to the state at draw time, so compare them first, then the assignments
    scale = np.abs(lp_ref).max()
    assert np.allclose(lp_gpu, lp_ref, rtol=tol, atol=tol * 1e-2 * scale), np.abs(lp_gpu - lp_ref).max()
    assert np.array_equal(c_gpu, c_ref)
    assert nch == int(np.sum(_count_changes(A, W64, c, m, p, mus, sig, u, mode, beta, sampler)))


def _count_changes(A, W, c, m, p, mus, sig, u, mode, beta, sampler):
    # number of nodes whose draw differed from their label at draw time
    c = np.array(c).copy()
    logm = np.log(m)
    gc = sbm.gaussian_coefficients(mus, sig) if mus is not None else None
    ch = []
    for n in range(A.shape[0]):
        lp = sbm.node_conditional(n, A.astype(float), W, c, logm, p, gc, mode, beta)
        new = sbm.sample_discrete_from_log(lp, u[n]) if sampler == "cumsum" else sbm.log_sum_exp_sample(lp, u[n])
        ch.append(new != c[n]); c[n] = new
    return ch


@pytest.mark.parametrize("N,K,B,dtype,bn", [(50, 5, 2, "f64", 0), (200, 16, 2, "f64", 64), (333, 7, 1, "f64", 128),
                                            (200, 16, 2, "f32", 0), (130, 3, 2, "f64", 64), (257, 32, 2, "f64", 0),
                                            (300, 4, 1, "f32", 64), (500, 16, 2, "f32", 128), (512, 8, 1, "f64", 64),
                                            (260, 6, 3, "f64", 64), (400, 16, 4, "f32", 128), (330, 5, 4, "f64", 0), (1300, 8, 3, "f32", 0)])
def test_sweep_old_api_shared_c(N, K, B, dtype, bn):
    from graphistician_b200 import _lib
    A, W, c, m, p, mus, sig, u = _problem(7 * N + K, N, K, B)
    _check_sweep(A, W, c, m, p, mus, sig, u, dtype, sbm.SELF_OLD, _lib.SELF_ONCE, block_nodes=bn)


@pytest.mark.parametrize("special", [True, False])
def test_sweep_new_api(special):
    from graphistician_b200 import _lib
    A, W, c, m, p, mus, sig, u = _problem(11, 150, 6, 2)
    mode, gmode = (sbm.SELF_SKIP, _lib.SELF_SKIP) if special else (sbm.SELF_NEW, _lib.SELF_ONCE)
    _check_sweep(A, None, c, m, p, None, None, u, "f64", mode, gmode)            # adjacency.py:650-695
    _check_sweep(A, W, c, m, None, mus, sig, u, "f64", mode, gmode)              # weights.py:415-488


def test_sweep_ais_tempered():
    from graphistician_b200 import _lib
    A, W, c, m, p, mus, sig, u = _problem(13, 120, 3, 0)
    _check_sweep(A, None, c, m, p, None, None, u, "f64", sbm.SELF_AIS, _lib.SELF_TWICE, beta=0.37, sampler="lse")


def test_golden_fixture(golden_dir):
    from graphistician_b200 import _lib, DeviceNetwork
    g = np.load(os.path.join(golden_dir, "sbm_small.npz"))
    net = DeviceNetwork(g["A"], g["W"], dtype="f64")
    c1, lp1, _ = net.sbm_sweep(g["c"], np.log(g["m"]), g["p"], g["mus"], g["sigmas"], g["u"], want_lp=True)
    assert np.array_equal(c1, g["c_old"]) and np.allclose(lp1, g["lp_old"], rtol=1e-9)
    n_edge, sw, sww = net.block_stats(g["c"], len(g["m"]))
    assert np.array_equal(n_edge, g["st_n_edge"]) and np.allclose(sww, g["st_sum_ww"], rtol=1e-12, atol=1e-12)


def test_peaked_posterior_stationary():
    # strong block structure: after convergence almost nothing moves; exercises the long speculative windows
    from graphistician_b200 import _lib
    A, W, c, m, p, mus, sig, u = _problem(21, 1500, 4, 2)
    net, Wd = _net(A, W, "f64")
    c2, _, nch = net.sbm_sweep(c, np.log(m), p, mus, sig, u, want_lp=False)
    ref, _ = sbm.sbm_assign_sweep(A.astype(float), W, c, m, p, mus, sig, u, sbm.SELF_OLD)
    assert np.array_equal(c2, ref) and nch < 50


@pytest.mark.parametrize("N,K,B,dtype,bn", [(700, 5, 2, "f64", 128), (1500, 4, 2, "f64", 0), (900, 16, 1, "f32", 256), (800, 6, 4, "f64", 128), (1200, 5, 3, "f32", 0),
                                            (640, 7, 0, "f32", 64), (2600, 8, 2, "f64", 0)])
def test_sweep_leaves_the_next_k1_table(N, K, B, dtype, bn):
    """gk_sbm_sweep_stats: the K1 statistics of the labels a sweep ends on, folded from the sweep's own per-node statistics plus
    the movers' rows, equal a fresh gk_block_stats pass (counts bit-exact) -- for a first sweep from random labels (many movers),
    the sweep after it (few), and with planted labels (none)."""
    A, W, c, m, p, mus, sig, u = _problem(3 * N + K, N, K, B)
    net, Wd = _net(A, W, dtype)
    assert net.sbm_sweep_stats(K) is None                                # no sweep yet
    rng = np.random.RandomState(N)
    got_valid = 0
    for it in range(3):
        c, _, nch = net.sbm_sweep(c, np.log(m), p, mus if B else None, sig if B else None, rng.rand(N), block_nodes=bn)
        st = net.sbm_sweep_stats(K)
        if nch > 1024:
            assert st is None                                            # too many movers: the host takes a fresh pass
            continue
        assert st is not None
        got_valid += 1
        n_edge, sw, sww = net.block_stats(c, K)
        assert np.array_equal(st[0], n_edge)
        if B:
            tol = 1e-10 if dtype == "f64" else 2e-4
            assert np.allclose(st[1], sw, rtol=tol, atol=tol * np.abs(sww).max())
            assert np.allclose(st[2], sww, rtol=tol, atol=tol * np.abs(sww).max())
    assert got_valid >= 1
    assert net.sbm_sweep_stats(K + 1) is None                            # another K: nothing to offer


@pytest.mark.parametrize("N,K,B,dtype,bn", [(1500, 4, 2, "f64", 256), (2600, 8, 2, "f64", 0), (900, 16, 1, "f32", 128), (640, 7, 0, "f32", 64),
                                            (1100, 5, 4, "f64", 256), (1400, 6, 3, "f32", 0)])
def test_overlapped_sweep_equals_the_serial_order(N, K, B, dtype, bn):
    """GK_CTX_OPT_SBM_OVERLAP = 2: the strip pass of node block b+1 runs beside the serial kernel of block b on a label
    snapshot and k_stats_fix moves the terms of block b's movers.  From random labels (hundreds of movers per block), the
    sweep after it and a third one: labels identical to the serial order (mode 0) and to the oracle, conditionals to
    round-off, and the K1 table the sweep leaves behind equal to a fresh pass (counts bit-exact)."""
    from graphistician_b200 import _lib
    A, W, c_true, m, p, mus, sig, u = _problem(5 * N + K, N, K, B)
    net, Wd = _net(A, W, dtype)
    rng = np.random.RandomState(N + 1)
    c0 = rng.randint(0, K, N)                       # not the generating labels: most nodes move in the first sweep
    us = [rng.rand(N) for _ in range(3)]
    out = {}
    try:
        for mode in (0, 2):
            net.ctx.set_option(_lib.CTX_OPT_SBM_OVERLAP, mode)
            c, res = c0.copy(), []
            for it in range(3):
                c, lp, nch = net.sbm_sweep(c, np.log(m), p, mus if B else None, sig if B else None, us[it], block_nodes=bn, want_lp=True)
                res.append((c.copy(), lp.copy(), nch, net.sbm_sweep_stats(K)))
            out[mode] = res
    finally:
        net.ctx.set_option(_lib.CTX_OPT_SBM_OVERLAP, 3)
    tol = 1e-9 if dtype == "f64" else 1e-3
    for it in range(3):
        (ca, lpa, na, sa), (cb, lpb, nb_, sb) = out[0][it], out[2][it]
        assert np.array_equal(ca, cb) and na == nb_
        assert np.allclose(lpa, lpb, rtol=tol, atol=tol * 1e-2 * np.abs(lpa).max())
        if sb is not None:
            n_edge, sw, sww = net.block_stats(cb, K)
            assert np.array_equal(sb[0], n_edge)
    c_ref, _ = sbm.sbm_assign_sweep(A.astype(float), Wd, c0, m, p, mus if B else None, sig if B else None, us[0], sbm.SELF_OLD)
    assert np.array_equal(out[2][0][0], c_ref)
    assert out[2][0][2] > N // 4, "the first sweep must exercise the correction (many movers)"


@pytest.mark.parametrize("N,K,B,dtype,bn", [(2300, 4, 2, "f64", 256), (1500, 6, 1, "f32", 128), (1200, 5, 0, "f32", 64), (1100, 4, 3, "f64", 256)])
def test_speculated_sweep_equals_the_serial_order(N, K, B, dtype, bn):
    """GK_CTX_OPT_SBM_OVERLAP = 3: a sweep that moved few nodes leaves the strip passes of every block of the NEXT sweep running in
    the background from its final labels; the next sweep corrects them block by block for the nodes that move meanwhile
    (k_append_movers / k_stats_fix_list).  Six sweeps from lightly perturbed labels with the block probabilities shaken every
    sweep (so that nodes keep moving while the speculation is in use), a call in between that claims the scratch buffers, and a
    re-upload: labels, move counts, conditionals and the K1 table equal the serial order (mode 0) throughout."""
    from graphistician_b200 import _lib
    A, W, c_true, m, p, mus, sig, u = _problem(9 * N + K, N, K, B)
    net, Wd = _net(A, W, dtype)
    rng = np.random.RandomState(N + 2)
    c0 = c_true.copy()
    flip = rng.choice(N, 40, replace=False)
    c0[flip] = rng.randint(0, K, 40)
    us = [rng.rand(N) for _ in range(6)]

    def run(mode, beta):
        net.ctx.set_option(_lib.CTX_OPT_SBM_OVERLAP, mode)
        c, res = c0.copy(), []
        for it in range(6):
            c, lp, nch = net.sbm_sweep(c, np.log(m), p, mus if B else None, sig if B else None, us[it], block_nodes=bn, want_lp=True, beta=beta)
            res.append((c.copy(), lp.copy(), nch, net.sbm_sweep_stats(K)))
            if it == 2:
                net.block_stats(c, K)                      # another user of the scratch buffers: the speculation must give way
            if it == 4:
                net.upload(A, Wd)                          # new data behind the same handle: the speculation must be dropped
        return res

    out = {}
    try:
        # a temperature at which a few nodes (but never more than 256, the regime that arms the speculation) move in every sweep
        for beta in (0.05, 0.02, 0.01, 0.006, 0.004, 0.003, 0.002, 0.0015, 0.001, 0.0007, 0.0005):
            out[0] = run(0, beta)
            if all(1 <= r[2] <= 256 for r in out[0]):
                break
        else:
            pytest.skip("no temperature with 1..256 movers per sweep for this problem")
        out[3] = run(3, beta)
    finally:
        net.ctx.set_option(_lib.CTX_OPT_SBM_OVERLAP, 3)
    tol = 1e-9 if dtype == "f64" else 1e-3
    moved_under_speculation = 0
    for it in range(6):
        (ca, lpa, na, sa), (cb, lpb, nb_, sb) = out[0][it], out[3][it]
        assert np.array_equal(ca, cb) and na == nb_, it
        assert np.allclose(lpa, lpb, rtol=tol, atol=tol * 1e-2 * np.abs(lpa).max()), it
        if sb is not None:
            n_edge, sw, sww = net.block_stats(cb, K)
            assert np.array_equal(sb[0], n_edge), it
        if it in (1, 2, 4):
            moved_under_speculation += nb_
    assert moved_under_speculation > 0, "nodes must move while a speculation is in use"
    assert max(r[2] for r in out[0]) <= 256, "every sweep must stay in the regime that arms the speculation"


def test_sweep_with_exact_zero_and_one_probabilities():
    """p == 0 / p == 1 are legal (stochastic_block_model.py:58-60).  The reference's Bernoulli.log_probability turns an
    impossible event into a hard exclusion (nan_to_num(-inf) = -1.8e308) and 0 * log 0 into 0 (internals/distributions.py:59-67):
    the device tables must do the same, not soften it.  Labels against the literal loops."""
    rng = np.random.RandomState(4)
    N, K, B = 48, 3, 1
    c = rng.randint(0, K, N)
    m = np.array([0.3, 0.3, 0.4])
    p = rng.beta(2, 2, (K, K))
    p[0, 1] = 0.0
    p[2, 2] = 1.0
    mus = rng.randn(K, K, B); sig = np.tile(np.eye(B), (K, K, 1, 1)) * 1.3
    A = (rng.rand(N, N) < p[np.ix_(c, c)]).astype(np.uint8)          # consistent with p under the starting labels ...
    W = rng.randn(N, N, B) + mus[np.ix_(c, c)]
    u = rng.rand(N)
    c_ref, lp_ref = sbm.gibbs_sbm_resample_c_loops(A.astype(float), W, c, m, p, mus, sig, u)
    net, _ = _net(A, W, "f64")
    c_gpu, lp_gpu, _ = net.sbm_sweep(c, np.log(m), p, mus, sig, u, want_lp=True)
    assert np.array_equal(c_gpu, c_ref)
    excluded = lp_ref < -1e100                                         # ... so other labels hit impossible events
    assert excluded.any() and (~excluded).any()
    assert np.all(lp_gpu[excluded] < -1e9)                             # hard exclusions stay hard
    # (a label that WAS excluded and is freed by a move comes back through -1e10 + 1e10: good to the ulp of 1e10, 2e-6)
    assert np.allclose(lp_gpu[~excluded], lp_ref[~excluded], rtol=1e-9, atol=1e-4)


def test_weight_dimension_limits_are_reported():
    """B in {0..4} is compiled (template instantiations); beyond it, or when a (K, B, dtype) combination needs more shared
    memory than an SM has, the C ABI says so (GK_EUNSUPPORTED) instead of launching."""
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200._lib import GKError
    rng = np.random.RandomState(0)
    A = (rng.rand(64, 64) < 0.3).astype(np.uint8)
    with pytest.raises(GKError):
        DeviceNetwork(A, rng.randn(64, 64, 5), dtype="f32")
    net = DeviceNetwork(np.tile(A, (4, 4)), rng.randn(256, 256, 4), dtype="f64")
    with pytest.raises(GKError):
        net.block_stats(rng.randint(0, 32, 256), 32)             # 32 x 15 x 65 fp64 slots: more than 160 KB
    net.close()
