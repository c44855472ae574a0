"""Row-sharded execution over several GPUs of one node: one process per GPU, rows of A / W (and of
their transposes) split contiguously on 128-row boundaries, small statistics all-reduced with torch.distributed
(NCCL on the GPUs; gloo in the CPU tests, which drive the same host logic through an oracle-backed engine).

The host model state (c, m, p, F, L, ...) is replicated: every rank runs the same Python, seeds ``np.random``
identically and therefore makes the same O(K^2) host draws; only the O(N^2) passes are partitioned.

    comm = Comm()                                    # wraps the default process group (or a single process)
    network = FixedGaussianNetwork(A, W).shard(comm) # this rank uploads rows [r0, r1) only
    model.resample(network)                          # unchanged model code
"""
import ctypes

import numpy as np

ALIGN = 128


def row_range(N, rank, world, align=ALIGN):
    """Contiguous rows [r0, r1) of rank ``rank``: whole blocks of ``align`` rows, spread as evenly as possible."""
    nblk = (N + align - 1) // align
    assert nblk >= world, "N=%d is too small for %d ranks (needs at least %d rows per rank)" % (N, world, align)
    b0, b1 = (nblk * rank) // world, (nblk * (rank + 1)) // world
    return b0 * align, min(N, b1 * align)


def owner_of(N, world, n, align=ALIGN):
    nblk = (N + align - 1) // align
    b = n // align
    return min(world - 1, ((b + 1) * world - 1) // nblk)


class Comm(object):
    """The collectives the sharded paths need, over torch.distributed (no-ops in a single process)."""

    def __init__(self, group=None, device=None):
        self.group = group
        try:
            import torch.distributed as dist
            self.dist = dist if dist.is_available() and dist.is_initialized() else None
        except Exception:
            self.dist = None
        self.rank = self.dist.get_rank(group) if self.dist else 0
        self.world = self.dist.get_world_size(group) if self.dist else 1
        if device is None:
            device = "cuda" if (self.dist and self.dist.get_backend(group) == "nccl") else "cpu"
        self.device = device

    def allreduce_(self, tensor):
        """In-place SUM over the ranks (torch tensor on ``self.device``)."""
        if self.dist and self.world > 1:
            self.dist.all_reduce(tensor, op=self.dist.ReduceOp.SUM, group=self.group)
        return tensor

    def allreduce_np(self, arr):
        """SUM of a host array over the ranks (int64 stays exact)."""
        if not (self.dist and self.world > 1):
            return arr
        import torch
        t = torch.from_numpy(np.ascontiguousarray(arr)).to(self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy()

    def broadcast_np(self, arr, src):
        if not (self.dist and self.world > 1):
            return arr
        import torch
        t = torch.from_numpy(np.ascontiguousarray(arr)).to(self.device)
        self.dist.broadcast(t, src=self._global(src), group=self.group)
        return t.cpu().numpy()

    def _global(self, group_rank):
        if self.group is None or not hasattr(self.dist, "get_global_rank"):
            return group_rank
        return self.dist.get_global_rank(self.group, group_rank)

    def buffer(self, n):
        """Zeroed float64 exchange buffer on the communication device."""
        import torch
        return torch.zeros(int(n), dtype=torch.float64, device=self.device)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


class DeviceEngine(object):
    """This rank's shard on its GPU: the libgk calls of the sharded paths (include/gk.h 'sharded form')."""

    def __init__(self, A, W, r0, r1, dtype="f32", ctx=None):
        from .device import DeviceNetwork
        A = np.asarray(A)
        N = A.shape[0]
        idx = np.arange(N)
        At_rows = np.ascontiguousarray(A[:, r0:r1].T)
        Wt_rows = None if W is None else np.ascontiguousarray(np.asarray(W)[:, r0:r1].transpose(1, 0, 2))
        self.net = DeviceNetwork(A[r0:r1], None if W is None else np.asarray(W)[r0:r1], dtype=dtype, ctx=ctx,
                                 row0=r0, nrows=r1 - r0, At_rows=At_rows, Wt_rows=Wt_rows, diagA=A[idx, idx],
                                 diagW=None if W is None else np.asarray(W)[idx, idx])
        self.ctx, self.lib = self.net.ctx, self.net.ctx.lib
        self.N, self.B = N, self.net.B

    # K1
    def block_stats(self, c, K):
        return self.net.block_stats(c, K)

    def diag_stats(self, c, K):
        return self.net.diag_stats(c, K)

    # K2, block form
    def sweep_begin(self, c, log_m, p, mu, sigma, u, self_mode, sampler, beta):
        from ._lib import ptr
        self._keep = [np.ascontiguousarray(c, dtype=np.int32), np.ascontiguousarray(log_m, dtype=np.float64),
                      None if p is None else np.ascontiguousarray(p, dtype=np.float64),
                      None if mu is None else np.ascontiguousarray(mu, dtype=np.float64),
                      None if sigma is None else np.ascontiguousarray(sigma, dtype=np.float64),
                      np.ascontiguousarray(u, dtype=np.float64)]
        c32, lm, pp, mm, ss, uu = self._keep
        self.K = len(lm)
        self.ctx.check(self.lib.gk_sbm_sweep_begin(self.ctx.h, self.net.h, ptr(c32), self.K, ptr(lm), ptr(pp), ptr(mm),
                                                   ptr(ss), int(self_mode), int(sampler), float(beta), ptr(uu)))

    def partial_len(self, nb):
        return int(self.lib.gk_sbm_partial_len(self.ctx.h, int(nb)))

    def block_partials(self, n0, nb, buf):
        self.ctx.check(self.lib.gk_sbm_block_partials(self.ctx.h, self.net.h, int(n0), int(nb), _ptr(buf)))

    def block_sample(self, n0, nb, buf):
        self.ctx.check(self.lib.gk_sbm_block_sample(self.ctx.h, self.net.h, int(n0), int(nb), _ptr(buf)))

    def get_block(self, n0, nb):
        from ._lib import ptr
        out = np.empty(nb, np.int32)
        self.ctx.check(self.lib.gk_sbm_get_block(self.ctx.h, int(n0), int(nb), ptr(out)))
        return out

    def set_block(self, n0, nb, lab):
        from ._lib import ptr
        lab = np.ascontiguousarray(lab, dtype=np.int32)
        self.ctx.check(self.lib.gk_sbm_set_block(self.ctx.h, int(n0), int(nb), ptr(lab)))

    def export_block_dev(self, n0, nb, lab_t):
        """labels of the block -> the device tensor ``lab_t`` (int32), in-stream, no host synchronisation"""
        self.ctx.check(self.lib.gk_sbm_export_block_dev(self.ctx.h, int(n0), int(nb), _ptr(lab_t)))

    def import_block_dev(self, n0, nb, lab_t):
        self.ctx.check(self.lib.gk_sbm_import_block_dev(self.ctx.h, int(n0), int(nb), _ptr(lab_t)))

    def sweep_end(self):
        from ._lib import ptr
        c = np.empty(self.N, np.int32)
        nch = ctypes.c_int64(0)
        self.ctx.check(self.lib.gk_sbm_sweep_end(self.ctx.h, self.net.h, ptr(c), None, ctypes.byref(nch)))
        return c.astype(np.int64), int(nch.value)

    def close(self):
        self.net.close()


class ShardedDeviceNetwork(object):
    """Same methods as ``device.DeviceNetwork`` (block_stats, diag_stats, sbm_sweep), collective over ``comm``."""

    def __init__(self, engine, comm, N, r0, r1, diagA, diagW):
        self.engine, self.comm, self.N, self.row0, self.nrows = engine, comm, N, r0, r1 - r0
        self.B = engine.B
        self.diagA, self.diagW = diagA, diagW
        self.local = getattr(engine, "net", None)    # the DeviceNetwork shard (None for the CPU test engine)
        self.ctx = getattr(engine, "ctx", None)

    def native(self):
        """True when this rank's libgk context holds the engine's own communicator over the same ranks."""
        return self.ctx is not None and getattr(self.ctx, "native_world", 1) == self.comm.world and self.comm.world > 1

    def block_stats(self, c, K):
        n, sw, sww = self.engine.block_stats(c, K)
        n = self.comm.allreduce_np(np.asarray(n, dtype=np.int64))
        if self.B:
            packed = self.comm.allreduce_np(np.concatenate([np.ravel(sw), np.ravel(sww)]))
            sw = packed[:sw.size].reshape(sw.shape)
            sww = packed[sw.size:].reshape(sww.shape)
        return n, sw, sww

    def diag_stats(self, c, K):
        """O(N) and replicated (every rank holds the full diagonal)."""
        c = np.asarray(c)
        e = self.diagA != 0
        B = self.B
        out = dict(d_cnt=np.bincount(c, minlength=K).astype(np.int64),
                   d_edge=np.bincount(c[e], minlength=K).astype(np.int64),
                   d_sum_w=np.zeros((K, B)), d_sum_ww=np.zeros((K, B, B)))
        for b in range(B):
            out["d_sum_w"][:, b] = np.bincount(c[e], weights=self.diagW[e, b], minlength=K)
            for b2 in range(B):
                out["d_sum_ww"][:, b, b2] = np.bincount(c[e], weights=self.diagW[e, b] * self.diagW[e, b2], minlength=K)
        return out

    def sbm_sweep(self, c, log_m, p, mu, sigma, uniforms, self_mode=1, sampler=0, beta=1.0, block_nodes=0,
                  want_lp=False):
        """The exact sequential sweep, sharded: per node block, partial statistics on every rank -> all-reduce ->
        the owner of the block's rows draws it -> label broadcast."""
        assert not want_lp, "per-node log conditionals are available on single-rank networks only"
        eng, comm, N = self.engine, self.comm, self.N
        if self.native():                            # one C call: block loop and NCCL collectives inside libgk, in-stream
            return self.local.sbm_sweep(c, log_m, p, mu, sigma, uniforms, self_mode=self_mode, sampler=sampler, beta=beta,
                                        block_nodes=block_nodes)
        eng.sweep_begin(c, log_m, p, mu, sigma, uniforms, self_mode, sampler, beta)
        bn = block_nodes if block_nodes > 0 else 1024
        bn = max(ALIGN, bn // ALIGN * ALIGN)
        on_device = comm.device == "cuda" and hasattr(eng, "export_block_dev")
        if on_device:                                # exchange buffers live on the device and are reused for every block
            import torch
            if getattr(self, "_xbuf", None) is None or self._xbuf.numel() < eng.partial_len(bn):
                self._xbuf = comm.buffer(eng.partial_len(bn))
                self._xlab = torch.zeros(bn, dtype=torch.int32, device=self._xbuf.device)
        for rk in range(comm.world):
            s0, s1 = row_range(N, rk, comm.world)
            for n0 in range(s0, s1, bn):
                nb = min(bn, s1 - n0)
                if on_device:
                    # partial statistics -> NCCL all-reduce -> owner draws -> NCCL broadcast of the labels: every step is
                    # enqueued on the one stream libgk and NCCL share; the host never waits inside the sweep
                    buf = self._xbuf[:eng.partial_len(nb)]
                    lab = self._xlab[:nb]
                    eng.block_partials(n0, nb, buf)
                    comm.allreduce_(buf)
                    if comm.rank == rk:
                        eng.block_sample(n0, nb, buf)
                        eng.export_block_dev(n0, nb, lab)
                    comm.dist.broadcast(lab, src=comm._global(rk), group=comm.group)
                    if comm.rank != rk:
                        eng.import_block_dev(n0, nb, lab)
                    continue
                buf = comm.buffer(eng.partial_len(nb))
                eng.block_partials(n0, nb, buf)
                comm.allreduce_(buf)
                if comm.rank == rk:
                    eng.block_sample(n0, nb, buf)
                    lab = eng.get_block(n0, nb)
                else:
                    lab = np.zeros(nb, np.int32)
                lab = comm.broadcast_np(lab, rk)
                if comm.rank != rk:
                    eng.set_block(n0, nb, lab)
        c_new, nch = eng.sweep_end()
        nch = int(comm.allreduce_np(np.array([nch], dtype=np.int64))[0])
        return c_new, None, nch

    def close(self):
        self.engine.close()


def sharded_device_network(A, W, comm, dtype="f32", ctx=None, engine_factory=None):
    """This rank's shard of (A, W) as a ShardedDeviceNetwork.  The libgk context is put on torch's current CUDA
    stream so kernels and NCCL collectives are ordered on the device without host synchronisation."""
    A = np.asarray(A)
    N = A.shape[0]
    r0, r1 = row_range(N, comm.rank, comm.world)
    idx = np.arange(N)
    diagA = np.ascontiguousarray((A[idx, idx] != 0).astype(np.uint8))
    diagW = None if W is None else np.ascontiguousarray(np.asarray(W)[idx, idx], dtype=np.float64)
    if engine_factory is None:
        from . import _lib
        ctx = ctx or _lib.default_context()
        if comm.device == "cuda":
            import os
            import torch
            ctx.set_stream(torch.cuda.current_stream().cuda_stream)
            if comm.world > 1 and not os.environ.get("GK_NO_NATIVE_COMM"):
                ctx.init_comm(comm)                  # libgk issues the sweeps' collectives itself from here on
        engine = DeviceEngine(A, W, r0, r1, dtype=dtype, ctx=ctx)
    else:
        engine = engine_factory(A, W, r0, r1)
    return ShardedDeviceNetwork(engine, comm, N, r0, r1, diagA, diagW)
