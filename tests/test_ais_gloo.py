"""World-size-2 ``gloo`` test (CPU) of the particle partition of the batched AIS driver
(graphistician_b200.internals.graph_models.estimate_marginal_likelihood_batched, SURVEY.md 8e: AIS particles partitioned over
the GPUs, no data-path collective, one gather of the log weights).

There is no GPU here, so the device state is stood in for by ``OracleAIS`` -- a NumPy object with the interface of
``device.DeviceAIS`` (set_labels / counts / sweep / nnz) built on the oracle's transcription of
internals/graph_models.py:261-297.  What is under test is the product's host code: which particles a rank runs, each
particle's private random stream, the tempered Dirichlet / Beta draws from the counts, the log-weight arithmetic and the
final all-reduce.  The two-rank result must equal the one-process result and the per-particle reference driver."""
import os
import socket

import numpy as np
import pytest
from scipy.special import logsumexp

from oracle import sbm


class OracleAIS(object):
    def __init__(self, A, P, R):
        self.A = np.asarray(A, dtype=np.float64)
        self.N, self.P, self.R = self.A.shape[0], int(P), int(R)
        self.z = np.zeros((self.P, self.N), np.int64)

    def set_labels(self, z):
        z = np.asarray(z)
        self.z[:] = z if z.ndim == 2 else z[None, :]

    def counts(self, recompute=False):
        n_edge = np.zeros((self.P, self.R, self.R), np.int64)
        sizes = np.zeros((self.P, self.R), np.int64)
        for p in range(self.P):
            for r1 in range(self.R):
                for r2 in range(self.R):
                    n_edge[p, r1, r2] = int(self.A[np.ix_(self.z[p] == r1, self.z[p] == r2)].sum())
            sizes[p] = np.bincount(self.z[p], minlength=self.R)
        return n_edge, sizes

    def sweep(self, log_pi, B, beta, uniforms=None, seed=0, want_lp=False):
        assert uniforms is not None, "the CPU stand-in has no device stream: pass device_uniforms=False"
        beta = np.broadcast_to(np.asarray(beta, dtype=np.float64), (self.P,))
        nchg = np.zeros(self.P, np.int64)
        for p in range(self.P):
            z_new, _ = sbm.sbm_assign_sweep(self.A, None, self.z[p], np.exp(log_pi[p]), B[p], None, None, uniforms[p],
                                            sbm.SELF_AIS, float(beta[p]), sampler="lse")
            nchg[p] = np.count_nonzero(z_new != self.z[p])
            self.z[p] = z_new
        return nchg, None

    def nnz(self):
        return int(self.A.sum())


def _problem():
    rng = np.random.RandomState(17)
    N, R = 40, 3
    zt = rng.randint(0, R, N)
    Bt = np.array([[0.8, 0.1, 0.05], [0.05, 0.7, 0.1], [0.1, 0.05, 0.9]])
    A = (rng.rand(N, N) < Bt[np.ix_(zt, zt)]).astype(np.uint8)
    return A, rng.randint(0, R, N), R


def _ref_weights(A, f0, R, NB, steps, seeds):
    """examples/demo_ais.py:10-62 per particle with the literal sample_theta of internals/graph_models.py:337-371."""
    Af = A.astype(float)
    N = A.shape[0]
    betas = np.linspace(0, 1, NB)
    w = np.zeros(len(seeds))
    for m, sd in enumerate(seeds):
        rs = np.random.RandomState(int(sd))
        f = np.array(f0)
        for b in range(1, NB):
            for s in range(steps):
                pi = rs.dirichlet(np.array([10.0 + np.sum(f == r) for r in range(R)]))
                Bm = np.zeros((R, R), np.float32)
                for r1 in range(R):
                    for r2 in range(R):
                        Ar = Af[np.ix_(f == r1, f == r2)]
                        Bm[r1, r2] = rs.beta(1.0 + betas[b] * np.sum(Ar), 1.0 + betas[b] * np.sum(1 - Ar))
                Bm = Bm.astype(np.float64)
                f, _ = sbm.sbm_assign_sweep(Af, None, f, pi, Bm, None, None, rs.rand(N), sbm.SELF_AIS, betas[b], sampler="lse")
            P = Bm[np.ix_(f, f)]
            w[m] += (betas[b] - betas[b - 1]) * np.sum(np.where(Af > 0, np.log(P), np.log(1 - P)))
    return w


def _run(comm, M=6, NB=7, steps=2):
    from graphistician_b200.internals.graph_models import StochasticBlockModel, estimate_marginal_likelihood_batched
    A, f0, R = _problem()
    model = StochasticBlockModel(R=R, b0=1.0, b1=1.0, alpha0=10 * np.ones(R))
    seeds = 500 + np.arange(M)
    rank, world = (comm.rank, comm.world) if comm is not None else (0, 1)
    mine = len(range(rank, M, world))
    return estimate_marginal_likelihood_batched(A, f0, None, model, N_samples=M, B=NB, steps_per_B=steps, seeds=seeds, comm=comm,
                                                return_weights=True, device_uniforms=False, dev=OracleAIS(A, mine, R))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from graphistician_b200.sharding import Comm
        comm = Comm()
        assert comm.world == world and comm.device == "cpu"
        logZ, w = _run(comm)
        if rank == 0:
            out.put((float(logZ), np.asarray(w)))
    finally:
        dist.destroy_process_group()


def test_particles_partitioned_over_two_ranks_give_the_single_process_weights():
    import torch.multiprocessing as mp
    logZ1, w1 = _run(None)
    A, f0, R = _problem()
    w_ref = _ref_weights(A, f0, R, 7, 2, 500 + np.arange(6))
    assert np.allclose(w1, w_ref, rtol=1e-10)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    logZ2, w2 = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.allclose(w2, w1, rtol=1e-12) and np.isclose(logZ2, logZ1, rtol=1e-12)
    assert np.isclose(logZ1, -np.log(6) + logsumexp(w_ref), rtol=1e-10)
