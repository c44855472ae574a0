"""World-size-2 ``gloo`` tests (CPU) of the row-sharded host logic in graphistician_b200/sharding.py.

There is no GPU here, so the per-rank kernels are stood in for by ``OracleEngine`` -- a NumPy engine with the same
interface as ``sharding.DeviceEngine`` (partial statistics over the rank's own rows, owner-side sequential draw with
in-block corrections), built on the oracle.  What is under test is the product's host code: the row partition, the
all-reduce of statistics, the owner/broadcast protocol of the sequential sweep, and the K1 reduction -- the sharded
result must equal the oracle's single-process result exactly (labels, integer counts) / to round-off (fp sums)."""
import os
import socket

import numpy as np
import pytest

from oracle import sbm

SELF = {0: sbm.SELF_SKIP, 1: sbm.SELF_OLD, 2: sbm.SELF_AIS}


class OracleEngine(object):
    def __init__(self, A, W, r0, r1):
        A = np.asarray(A)
        self.N, self.r0, self.r1 = A.shape[0], r0, r1
        self.B = 0 if W is None else W.shape[2]
        self.A_rows = (A[r0:r1] != 0)
        self.At_rows = np.ascontiguousarray((A[:, r0:r1] != 0).T)
        self.W_rows = None if W is None else np.asarray(W, dtype=np.float64)[r0:r1]
        self.Wt_rows = None if W is None else np.ascontiguousarray(np.asarray(W, dtype=np.float64)[:, r0:r1].transpose(1, 0, 2))
        idx = np.arange(self.N)
        self.diagA = (A[idx, idx] != 0)
        self.diagW = None if W is None else np.asarray(W, dtype=np.float64)[idx, idx]

    # K1 over this rank's rows, off-diagonal pairs only
    def block_stats(self, c, K):
        c = np.asarray(c, dtype=np.int64)
        B = self.B
        E = self.A_rows.copy()
        E[np.arange(self.r1 - self.r0), np.arange(self.r0, self.r1)] = False
        cell = (c[self.r0:self.r1, None] * K + c[None, :])[E]
        n = np.bincount(cell, minlength=K * K).reshape(K, K).astype(np.int64)
        sw = np.zeros((K, K, B)); sww = np.zeros((K, K, B, B))
        if B:
            We = self.W_rows[E]
            for b in range(B):
                sw[:, :, b] = np.bincount(cell, weights=We[:, b], minlength=K * K).reshape(K, K)
                for b2 in range(B):
                    sww[:, :, b, b2] = np.bincount(cell, weights=We[:, b] * We[:, b2], minlength=K * K).reshape(K, K)
        return n, sw, sww

    # K2, block form
    def sweep_begin(self, c, log_m, p, mu, sigma, u, self_mode, sampler, beta):
        self.c = np.array(c, dtype=np.int64)
        self.K = len(log_m)
        self.logm, self.p, self.u, self.beta = np.asarray(log_m, float), p, np.asarray(u, float), beta
        self.gc = sbm.gaussian_coefficients(mu, sigma) if mu is not None else None
        self.self_mode, self.sampler = SELF[self_mode], sampler
        self.S = 1 + self.B + self.B * self.B
        self.nchanged = 0

    def partial_len(self, nb):
        return nb * 2 * self.K * self.S

    def _add(self, st, k, w, sign=1.0):
        st[k, 0] += sign
        if self.B:
            st[k, 1:1 + self.B] += sign * w
            st[k, 1 + self.B:] += sign * np.outer(w, w).ravel()

    def block_partials(self, n0, nb, buf):
        K, S = self.K, self.S
        st = np.zeros((nb, 2, K, S))
        cl = self.c[self.r0:self.r1]
        for i in range(nb):
            n = n0 + i
            for d, (E, Wd) in enumerate(((self.At_rows, self.Wt_rows), (self.A_rows, self.W_rows))):
                e = E[:, n].copy()                       # d = 0: A[n, r] (out-edges of n); d = 1: A[r, n] (in-edges)
                if self.r0 <= n < self.r1:
                    e[n - self.r0] = False
                for r in np.nonzero(e)[0]:
                    self._add(st[i, d], cl[r], Wd[r, n] if self.B else None)
        buf.numpy()[:st.size] = st.ravel()

    def block_sample(self, n0, nb, buf):
        K, S, B = self.K, self.S, self.B
        assert self.r0 <= n0 and n0 + nb <= self.r1
        st = buf.numpy()[:nb * 2 * K * S].reshape(nb, 2, K, S).copy()
        old = self.c[n0:n0 + nb].copy()
        sizes = np.bincount(self.c, minlength=K).astype(np.float64)
        for i in range(nb):
            n = n0 + i
            s = st[i]
            for j in range(i):                            # in-block nodes that already moved
                m = n0 + j
                if self.c[m] == old[j]:
                    continue
                if self.A_rows[n - self.r0, m]:
                    w = self.W_rows[n - self.r0, m] if B else None
                    self._add(s[0], old[j], w, -1.0); self._add(s[0], self.c[m], w, 1.0)
                if self.At_rows[n - self.r0, m]:
                    w = self.Wt_rows[n - self.r0, m] if B else None
                    self._add(s[1], old[j], w, -1.0); self._add(s[1], self.c[m], w, 1.0)
            sz = sizes.copy(); sz[self.c[n]] -= 1
            lp = self.logm.copy()
            if self.p is not None:
                logp, l1 = np.log(self.p), np.log1p(-self.p)
                lp += self.beta * (logp.dot(s[0, :, 0]) + l1.dot(sz - s[0, :, 0]) + logp.T.dot(s[1, :, 0]) + l1.T.dot(sz - s[1, :, 0]))
            if self.gc is not None:
                c0, c1, C2 = self.gc
                so, si = s[0, :, 1:1 + B], s[1, :, 1:1 + B]
                soo, sii = s[0, :, 1 + B:].reshape(K, B, B), s[1, :, 1 + B:].reshape(K, B, B)
                lp += self.beta * (c0.dot(s[0, :, 0]) + np.einsum("kjb,jb->k", c1, so) + np.einsum("kjab,jab->k", C2, soo))
                lp += self.beta * (c0.T.dot(s[1, :, 0]) + np.einsum("jkb,jb->k", c1, si) + np.einsum("jkab,jab->k", C2, sii))
            k = np.arange(K)
            a11 = float(self.diagA[n])
            if self.self_mode in (sbm.SELF_OLD, sbm.SELF_NEW):
                if self.p is not None:
                    lp += self.beta * (a11 * np.log(self.p[k, k]) + (1 - a11) * np.log1p(-self.p[k, k]))
                if self.gc is not None and a11:
                    w = self.diagW[n]
                    lp += self.beta * (self.gc[0][k, k] + self.gc[1][k, k].dot(w) + np.einsum("kab,a,b->k", self.gc[2][k, k], w, w))
            new = sbm.sample_discrete_from_log(lp, self.u[n])
            if new != self.c[n]:
                sizes[self.c[n]] -= 1; sizes[new] += 1
                self.c[n] = new
                self.nchanged += 1

    def get_block(self, n0, nb):
        return self.c[n0:n0 + nb].astype(np.int32)

    def set_block(self, n0, nb, lab):
        self.c[n0:n0 + nb] = lab

    def sweep_end(self):
        return self.c.copy(), self.nchanged

    def close(self):
        pass


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _worker(rank, world, port, N, K, B, seed, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from graphistician_b200.sharding import Comm, sharded_device_network, row_range
        rng = np.random.RandomState(seed)
        c = rng.randint(0, K, N); m = rng.dirichlet(np.ones(K) * 3); p = rng.beta(1, 1, (K, K))
        A = (rng.rand(N, N) < p[np.ix_(c, c)]).astype(np.uint8)
        mus = rng.randn(K, K, B); sig = np.tile(np.eye(B), (K, K, 1, 1)) * 1.3
        W = rng.randn(N, N, B) + mus[np.ix_(c, c)]
        c0 = rng.randint(0, K, N); u = rng.rand(N)
        comm = Comm()
        assert comm.world == world and comm.device == "cpu"
        net = sharded_device_network(A, W, comm, engine_factory=OracleEngine)
        r0, r1 = row_range(N, rank, world)
        assert (net.row0, net.nrows) == (r0, r1 - r0)
        n, sw, sww = net.block_stats(c0, K)
        c_new, _, nch = net.sbm_sweep(c0, np.log(m), p, mus, sig, u, self_mode=1, sampler=0, block_nodes=128)
        if rank == 0:
            ref = sbm.block_pair_stats(A.astype(float), W, c0, K)
            c_ref, _ = sbm.sbm_assign_sweep(A.astype(float), W, c0, m, p, mus, sig, u, sbm.SELF_OLD)
            out.put(dict(n_ok=bool(np.array_equal(n, ref["n_edge"])), sw_ok=bool(np.allclose(sw, ref["sum_w"], rtol=1e-10)),
                         sww_ok=bool(np.allclose(sww, ref["sum_ww"], rtol=1e-10)), c_ok=bool(np.array_equal(c_new, c_ref)),
                         nch=nch, nch_ref=int(np.sum(c_ref != c0)), moved=int(np.sum(c_ref != c0))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("N,K,B", [(256, 3, 1), (384, 4, 2)])
def test_sharded_sbm_paths_match_single_process_oracle(N, K, B):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, N, K, B, 17 + N, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res["n_ok"] and res["sw_ok"] and res["sww_ok"], res
    assert res["c_ok"], res
    assert res["moved"] > 0 and res["nch"] >= res["nch_ref"]      # nch counts label flips, including back-and-forth


def test_row_ranges_partition_and_align():
    from graphistician_b200.sharding import row_range, owner_of
    for N in (128, 300, 1000, 16384, 65536):
        for world in (1, 2, 4, 8):
            if (N + 127) // 128 < world:
                continue
            cuts = [row_range(N, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == N
            for (a0, a1), (b0, b1) in zip(cuts[:-1], cuts[1:]):
                assert a1 == b0 and a1 % 128 == 0 and a1 > a0
            for r, (a0, a1) in enumerate(cuts):
                assert owner_of(N, world, a0) == r and owner_of(N, world, a1 - 1) == r
