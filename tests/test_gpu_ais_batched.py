"""GPU parity of the BATCHED AIS path (gk_ais_* over a bit-packed adjacency matrix, P particles
co-resident) through the C ABI.

Integer quantities (labels, R x R edge counts, class sizes) must be bit-identical to the oracle's transcription of
internals/graph_models.py:261-297, 337-371 for the same uniforms; conditionals within 1e-9 relative (fp64 arithmetic);
the whole AIS log weight of every particle must equal the reference driver's arithmetic (examples/demo_ais.py:10-62) run
with that particle's random stream; gk_ais_run's log Z must agree with an oracle-driven AIS run within Monte Carlo error."""
import numpy as np
import pytest
from scipy.special import betaln, logsumexp

from oracle import sbm

pytestmark = pytest.mark.gpu


def _data(N, R, seed, p_self=None):
    rng = np.random.RandomState(seed)
    z_true = rng.randint(0, R, N)
    Bt = rng.beta(1, 1, (R, R))
    A = (rng.rand(N, N) < Bt[np.ix_(z_true, z_true)]).astype(np.uint8)
    return A, z_true, rng


def _particles(rng, P, N, R):
    z0 = rng.randint(0, R, (P, N))
    Bm = rng.beta(1.5, 1.5, (P, R, R)).astype(np.float32).astype(np.float64)
    pi = rng.dirichlet(5 * np.ones(R), P)
    beta = rng.rand(P)
    beta[0] = 1.0
    if P > 1:
        beta[1] = 0.0
    u = rng.rand(P, N)
    return z0, Bm, pi, beta, u


def _ref_counts(A, z, R):
    n_edge = np.zeros((R, R), np.int64)
    for r1 in range(R):
        for r2 in range(R):
            n_edge[r1, r2] = A[np.ix_(z == r1, z == r2)].sum()
    return n_edge, np.bincount(z, minlength=R)


@pytest.mark.parametrize("N,R,P", [(40, 3, 4), (130, 2, 3), (1100, 3, 5), (2100, 5, 3), (257, 8, 2), (1024, 4, 33)])
def test_batched_sweep_matches_reference_loop(N, R, P):
    from graphistician_b200.device import DeviceAIS
    A, _, rng = _data(N, R, 7 + N)
    z0, Bm, pi, beta, u = _particles(rng, P, N, R)
    dev = DeviceAIS(A, P, R)
    dev.set_labels(z0)
    n_edge, sizes = dev.counts()
    for p in range(P):
        e, s = _ref_counts(A, z0[p], R)
        assert np.array_equal(n_edge[p], e) and np.array_equal(sizes[p], s)
    nchg, lp = dev.sweep(np.log(pi), Bm, beta, uniforms=u, want_lp=True)
    z = dev.get_labels()
    n_edge, sizes = dev.counts()
    n_edge2, sizes2 = dev.counts(recompute=True)
    assert np.array_equal(n_edge, n_edge2) and np.array_equal(sizes, sizes2), "incremental cell table drifted"
    Af = A.astype(float)
    for p in range(P):
        if N <= 300:
            z_ref, lp_ref = sbm.ais_naive_sample_f_loops(Af, z0[p], Bm[p], pi[p], beta[p], u[p])
        else:
            z_ref, lp_ref = sbm.sbm_assign_sweep(Af, None, z0[p], pi[p], Bm[p], None, None, u[p], sbm.SELF_AIS, beta[p], sampler="lse")
        assert np.array_equal(z[p], z_ref), "particle %d: labels differ at %s" % (p, np.nonzero(z[p] != z_ref)[0][:5])
        assert np.allclose(lp[p], lp_ref, rtol=1e-9, atol=1e-9)
        assert nchg[p] == np.count_nonzero(z_ref != z0[p])
        e, s = _ref_counts(A, z_ref, R)
        assert np.array_equal(n_edge[p], e) and np.array_equal(sizes[p], s)
    dev.close()


def test_second_sweep_continues_from_the_first_and_matches_single_particle_path():
    """Two sweeps in a row (masks, sizes and cell table carried on the device) against the existing one-particle
    path (gk_sbm_sweep with GK_SELF_TWICE / GK_SAMPLER_LSE on the uint8 network)."""
    from graphistician_b200 import DeviceNetwork, _lib
    from graphistician_b200.device import DeviceAIS
    N, R, P = 1500, 3, 6
    A, _, rng = _data(N, R, 3)
    z0, Bm, pi, beta, u = _particles(rng, P, N, R)
    u2 = rng.rand(P, N)
    net = DeviceNetwork(A, None, dtype="f64")
    dev = DeviceAIS(net, P, R)                                  # packs from the network's device copy
    dev.set_labels(z0)
    dev.sweep(np.log(pi), Bm, beta, uniforms=u)
    dev.sweep(np.log(pi), Bm, beta, uniforms=u2)
    z = dev.get_labels()
    for p in range(P):
        c1, _, _ = net.sbm_sweep(z0[p], np.log(pi[p]), Bm[p], None, None, u[p], self_mode=_lib.SELF_TWICE,
                                 sampler=_lib.SAMPLER_LSE, beta=float(beta[p]))
        c2, _, _ = net.sbm_sweep(c1, np.log(pi[p]), Bm[p], None, None, u2[p], self_mode=_lib.SELF_TWICE,
                                 sampler=_lib.SAMPLER_LSE, beta=float(beta[p]))
        assert np.array_equal(z[p], c2)
    assert dev.nnz() == int(A.sum())
    dev.close(); net.close()


def test_row_chunk_upload_equals_whole_upload_and_exact_zero_one_probabilities():
    from graphistician_b200.device import DeviceAIS
    N, R, P = 200, 3, 2
    A, _, rng = _data(N, R, 5)
    A[:, :] = 0
    zt = rng.randint(0, R, N)
    Bt = np.array([[1.0, 0.0, 0.3], [0.2, 1.0, 0.0], [0.0, 0.5, 0.6]])           # exact 0 / 1 cells
    A = (rng.rand(N, N) < Bt[np.ix_(zt, zt)]).astype(np.uint8)
    dev = DeviceAIS(None, P, R, N=N)
    for r0 in range(0, N, 64):
        dev.upload_rows(A[r0:r0 + 64], r0)
    assert dev.nnz() == int(A.sum())
    z0 = np.stack([zt, zt])                 # any other start makes every candidate impossible (the reference then divides 0 by 0)
    u = rng.rand(P, N)
    pi = np.ones((P, R)) / R
    dev.set_labels(z0)
    dev.sweep(np.log(pi), Bt, np.array([1.0, 0.4]), uniforms=u)
    z = dev.get_labels()
    with np.errstate(divide="ignore", invalid="ignore"):
        for p, beta in enumerate((1.0, 0.4)):
            z_ref, _ = sbm.ais_naive_sample_f_loops(A.astype(float), z0[p], Bt, pi[p], beta, u[p])
            assert np.array_equal(z[p], z_ref)
    assert np.array_equal(z[0], zt) and np.array_equal(z[1], zt), "hard exclusions: the generating labels are the only ones with non-zero probability"
    dev.close()


@pytest.mark.parametrize("N,R,P", [(1500, 3, 5), (2300, 4, 18), (700, 2, 3)])
def test_incremental_counts_equal_the_counting_pass(N, R, P):
    """GK_AIS_OPT_INCREMENTAL: the class counts of every node are kept on the device and moved with the labels (k_ais_apply)
    instead of being recounted per node block.  Five sweeps from random labels (many movers, then few) in mode 2 (always)
    and in mode 1 (switches itself on once a sweep moved few nodes) against mode 0: identical labels, move counts,
    conditionals and R x R tables; and the tables equal a fresh count."""
    from graphistician_b200 import _lib
    from graphistician_b200.device import DeviceAIS
    A, zt, rng = _data(N, R, 41 + N)
    z0 = rng.randint(0, R, (P, N))
    pi = rng.dirichlet(5 * np.ones(R), P)
    Bt = rng.beta(1, 1, (R, R))
    A = (rng.rand(N, N) < Bt[np.ix_(zt, zt)]).astype(np.uint8)
    Bm = np.tile(np.clip(Bt, 0.02, 0.98), (P, 1, 1))
    beta = np.linspace(1.0, 0.3, P)
    us = [rng.rand(P, N) for _ in range(5)]
    out = {}
    for mode in (0, 2, 1):
        dev = DeviceAIS(A, P, R)
        dev.set_option(_lib.AIS_OPT_INCREMENTAL, mode)
        dev.set_labels(z0)
        res = []
        for it in range(5):
            nchg, lp = dev.sweep(np.log(pi), Bm, beta, uniforms=us[it], want_lp=True)
            res.append((dev.get_labels(), nchg.copy(), lp.copy(), dev.counts()))
        e1, s1 = dev.counts(); e2, s2 = dev.counts(recompute=True)
        assert np.array_equal(e1, e2) and np.array_equal(s1, s2)
        out[mode] = res
        dev.close()
    for mode in (2, 1):
        for it in range(5):
            z_a, n_a, lp_a, (e_a, s_a) = out[0][it]
            z_b, n_b, lp_b, (e_b, s_b) = out[mode][it]
            assert np.array_equal(z_a, z_b) and np.array_equal(n_a, n_b), (mode, it)
            assert np.array_equal(lp_a, lp_b) and np.array_equal(e_a, e_b) and np.array_equal(s_a, s_b), (mode, it)
    assert out[0][0][1].min() > N // 4 and out[0][4][1].max() <= 256, "the sweeps must cover both regimes"


def _ref_sample_theta(rs, A, z, R, b0, b1, alpha0, beta):
    """internals/graph_models.py:337-371 literally, drawing from the particle's stream."""
    alpha_post = np.array([alpha0[r] + np.sum(z == r) for r in range(R)], dtype=float)
    pi = rs.dirichlet(alpha_post)
    B = np.zeros((R, R), dtype=np.float32)
    for r1 in range(R):
        for r2 in range(R):
            b0p, b1p = b0, b1
            Ar = A[np.ix_(z == r1, z == r2)]
            if np.size(Ar) > 0:
                b0p += beta * np.sum(1 - Ar)
                b1p += beta * np.sum(Ar)
            B[r1, r2] = rs.beta(b1p, b0p)
    return B.astype(np.float64), pi


def _ref_driver(A, f0, R, NB, steps, seeds, b0=1.0, b1=1.0, alpha=10.0):
    """examples/demo_ais.py:10-62 on the CPU with the oracle's loops, particle m drawing from RandomState(seeds[m])."""
    Af = A.astype(float)
    N = A.shape[0]
    betas = np.linspace(0, 1, NB)
    w = np.zeros(len(seeds))
    for m, sd in enumerate(seeds):
        rs = np.random.RandomState(int(sd))
        f = np.array(f0)
        for b in range(1, NB):
            for s in range(steps):
                theta = _ref_sample_theta(rs, Af, f, R, b0, b1, alpha * np.ones(R), betas[b])
                u = rs.rand(N)
                f, _ = sbm.sbm_assign_sweep(Af, None, f, theta[1], theta[0], None, None, u, sbm.SELF_AIS, betas[b], sampler="lse")
            P = theta[0][np.ix_(f, f)]
            with np.errstate(divide="ignore", invalid="ignore"):
                ll = np.sum(np.where(Af > 0, np.log(P), np.log(1 - P)))
            w[m] += (betas[b] - betas[b - 1]) * ll
    return w


def test_batched_ais_log_weights_match_reference_driver():
    from graphistician_b200.internals.graph_models import StochasticBlockModel, estimate_marginal_likelihood_batched
    N, R, NB, steps, M = 72, 3, 9, 2, 5
    A, zt, rng = _data(N, R, 21)
    f0 = rng.randint(0, R, N)
    model = StochasticBlockModel(R=R, b0=1.0, b1=1.0, alpha0=10 * np.ones(R))
    seeds = np.arange(M) + 100
    logZ, w = estimate_marginal_likelihood_batched(A, f0, None, model, N_samples=M, B=NB, steps_per_B=steps, seeds=seeds,
                                                   return_weights=True)
    w_ref = _ref_driver(A, f0, R, NB, steps, seeds)
    assert np.allclose(w, w_ref, rtol=1e-10)
    assert np.isclose(logZ, -np.log(M) + logsumexp(w_ref), rtol=1e-10)
    # two batches of 3 + 2 particles give the same weights (every particle owns its stream)
    _, w2 = estimate_marginal_likelihood_batched(A, f0, None, model, N_samples=M, B=NB, steps_per_B=steps, seeds=seeds,
                                                 particles=3, return_weights=True)
    assert np.allclose(w2, w_ref, rtol=1e-10)


def test_gk_ais_run_agrees_with_an_oracle_driven_run_at_N64():
    """log Z agreement with the oracle at N = 64.  gk_ais_run draws theta from its own host stream
    and the sweep uniforms from the device stream, so agreement is distributional: the mean log weight within 4 standard
    errors and log Z within 1.5 nats of an oracle-driven run of the same ladder."""
    from graphistician_b200 import _lib
    from graphistician_b200.device import DeviceAIS
    N, R, NB, steps = 64, 3, 30, 2
    rng = np.random.RandomState(8)
    zt = rng.randint(0, R, N)
    Bt = np.array([[0.8, 0.05, 0.05], [0.05, 0.7, 0.1], [0.1, 0.05, 0.9]])
    A = (rng.rand(N, N) < Bt[np.ix_(zt, zt)]).astype(np.uint8)
    f0 = rng.randint(0, R, N)
    P = 96
    dev = DeviceAIS(A, P, R)
    w = dev.run(1, 1.0, 1.0, 10 * np.ones(R), NB, steps, f0=f0, seed=1234)
    dev.close()
    M = 24
    w_ref = _ref_driver(A, f0, R, NB, steps, np.arange(M) + 5)
    se = np.sqrt(w.var(ddof=1) / P + w_ref.var(ddof=1) / M)
    assert abs(w.mean() - w_ref.mean()) < 4 * se, (w.mean(), w_ref.mean(), se)
    logZ, logZ_ref = -np.log(P) + logsumexp(w), -np.log(M) + logsumexp(w_ref)
    assert abs(logZ - logZ_ref) < 1.5 + 4 * se, (logZ, logZ_ref)
    er = DeviceAIS(A, 64, 1)                                   # Erdos-Renyi: the exact answer is a Beta function
    w_er = er.run(0, 1.0, 1.0, None, 400, 1, seed=5)
    er.close()
    nnz = float(A.sum())
    exact = betaln(1.0 + nnz, 1.0 + N * N - nnz) - betaln(1.0, 1.0)
    assert abs(-np.log(64) + logsumexp(w_er) - exact) < 1.5
    assert logZ > -np.log(64) + logsumexp(w_er) + 50, "block data: the SBM must beat the Erdos-Renyi model clearly"
