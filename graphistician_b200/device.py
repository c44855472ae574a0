"""Device-resident network data and the thin numpy-facing wrappers over the C ABI.

A ``DeviceNetwork`` is the engine's view of the reference's ``network`` argument -- the ``(A, W)`` tuple of
the new API (networks.py:69-71) or the ``network.A / network.W`` object of the old API
(stochastic_block_model.py:217-218): A (N x N, {0,1}) as uint8, W (N x N x B) as fp32 (default) or fp64
(parity mode), both also held transposed so that in-edges stream as rows.
"""
import ctypes
import weakref

import numpy as np

from . import _lib
from ._lib import GK_F32, GK_F64, ptr


class DeviceNetwork:
    def __init__(self, A, W=None, dtype="f32", ctx=None, row0=0, nrows=None, At_rows=None, Wt_rows=None,
                 diagA=None, diagW=None):
        self.ctx = ctx or _lib.default_context()
        A = np.asarray(A)
        self.N = int(A.shape[1])
        self.row0 = int(row0)
        self.nrows = int(A.shape[0] if nrows is None else nrows)
        assert A.shape == (self.nrows, self.N), "A must be (nrows, N)"
        self.B = 0 if W is None else int(W.shape[2])
        self.dtype = GK_F64 if dtype in ("f64", np.float64, "float64") else GK_F32
        h = ctypes.c_void_p()
        self.ctx.check(self.ctx.lib.gk_network_create(self.ctx.h, self.N, self.B, self.dtype, self.row0, self.nrows,
                                                      ctypes.byref(h)))
        self.h = h
        self._finalizer = weakref.finalize(self, _destroy, self.ctx, h)
        self.upload(A, W, At_rows, Wt_rows, diagA, diagW)

    def upload(self, A, W=None, At_rows=None, Wt_rows=None, diagA=None, diagW=None):
        """Host -> device copy of this rank's shard (and device-side transposes when unsharded)."""
        A8 = _as_u8(A)
        Wc, wdt = _as_w(W)
        At8 = _as_u8(At_rows) if At_rows is not None else None
        Wtc = _as_w(Wt_rows)[0] if Wt_rows is not None else None
        self.ctx.check(self.ctx.lib.gk_network_upload(self.ctx.h, self.h, ptr(A8), ptr(Wc), wdt, ptr(At8), ptr(Wtc)))
        if diagA is None:
            assert self.nrows == self.N, "row-sharded networks must be given the full diagonal"
            idx = np.arange(self.N)
            diagA = A8[idx, idx]
            diagW = None if W is None else np.asarray(W)[idx, idx]
        dA = np.ascontiguousarray(diagA, dtype=np.uint8)
        dW = None if self.B == 0 else np.ascontiguousarray(diagW, dtype=np.float64)
        self.ctx.check(self.ctx.lib.gk_network_set_diag(self.ctx.h, self.h, ptr(dA), ptr(dW)))
        self.diagA, self.diagW = dA, dW
        self._diag_feat = None
        self._edge_count = None

    def edge_count(self):
        """Number of non-zero entries of A over this rank's rows, diagonal included: one K1 pass (K = 1) on the device,
        cached until the next upload (so an in-place edit of A followed by invalidate() is seen)."""
        if self._edge_count is None:
            n_edge, _, _ = self.block_stats(np.zeros(self.N, np.int32), 1)
            lo, hi = self.row0, self.row0 + self.nrows
            self._edge_count = int(n_edge[0, 0]) + int(np.count_nonzero(self.diagA[lo:hi]))
        return self._edge_count

    def device_bytes(self):
        return int(self.ctx.lib.gk_network_device_bytes(self.h))

    # ---- K1 -------------------------------------------------------------------------------------
    def block_stats(self, c, K):
        """Off-diagonal block-pair statistics over this rank's rows (exact int64 counts, fp64 sums)."""
        c32 = np.ascontiguousarray(c, dtype=np.int32)
        B = self.B
        n_edge = np.zeros((K, K), np.int64)
        sum_w = np.zeros((K, K, B))
        sum_ww = np.zeros((K, K, B, B))
        self.ctx.check(self.ctx.lib.gk_block_stats(self.ctx.h, self.h, ptr(c32), K, ptr(n_edge),
                                                   ptr(sum_w) if B else None, ptr(sum_ww) if B else None))
        return n_edge, sum_w, sum_ww

    def diag_stats(self, c, K):
        """O(N) host-side diagonal statistics per block (composed with block_stats by the model classes)."""
        c = np.asarray(c)
        B = self.B
        if self._diag_feat is None:                      # the diagonal is fixed with the network: features made once
            e = np.flatnonzero(self.diagA != 0)
            X = np.ones((e.size, 1 + B + B * B))
            if B:
                w = self.diagW[e]
                X[:, 1:1 + B] = w
                X[:, 1 + B:] = (w[:, :, None] * w[:, None, :]).reshape(e.size, B * B)
            self._diag_feat = (e, X)
        e, X = self._diag_feat
        onehot = np.zeros((K, e.size))
        onehot[c[e], np.arange(e.size)] = 1.0
        S = onehot.dot(X)
        return dict(d_cnt=np.bincount(c, minlength=K).astype(np.int64), d_edge=np.rint(S[:, 0]).astype(np.int64),
                    d_sum_w=S[:, 1:1 + B].copy(), d_sum_ww=S[:, 1 + B:].reshape(K, B, B).copy())

    # ---- K2 -------------------------------------------------------------------------------------
    def sbm_sweep(self, c, log_m, p, mu, sigma, uniforms, self_mode=_lib.SELF_ONCE, sampler=_lib.SAMPLER_CUMSUM,
                  beta=1.0, block_nodes=0, want_lp=False):
        """Exact sequential block-assignment sweep.  Returns (c_new, lp or None, n_changed)."""
        K = len(log_m)
        c32 = np.array(c, dtype=np.int32, copy=True, order="C")
        lm = np.ascontiguousarray(log_m, dtype=np.float64)
        pp = None if p is None else np.ascontiguousarray(p, dtype=np.float64)
        mm = None if mu is None else np.ascontiguousarray(mu, dtype=np.float64)
        ss = None if sigma is None else np.ascontiguousarray(sigma, dtype=np.float64)
        u = np.ascontiguousarray(uniforms, dtype=np.float64)
        assert u.shape == (self.N,)
        lp = np.zeros((self.N, K)) if want_lp else None
        nch = ctypes.c_int64(0)
        self.ctx.check(self.ctx.lib.gk_sbm_sweep(self.ctx.h, self.h, ptr(c32), K, ptr(lm), ptr(pp), ptr(mm), ptr(ss),
                                                 self_mode, sampler, float(beta), ptr(u), int(block_nodes), ptr(lp),
                                                 ctypes.byref(nch)))
        return c32.astype(np.int64), lp, int(nch.value)

    def sbm_sweep_stats(self, K):
        """K1 statistics of the labels the last ``sbm_sweep`` ended on, folded from that sweep's own per-node statistics
        (no pass over A / W); None when the engine has none to offer (see include/gk.h gk_sbm_sweep_stats)."""
        B = self.B
        n_edge = np.zeros((K, K), np.int64)
        sum_w = np.zeros((K, K, B)); sum_ww = np.zeros((K, K, B, B))
        ok = ctypes.c_int(0)
        self.ctx.check(self.ctx.lib.gk_sbm_sweep_stats(self.ctx.h, self.h, int(K), ptr(n_edge), ptr(sum_w) if B else None,
                                                       ptr(sum_ww) if B else None, ctypes.byref(ok)))
        return (n_edge, sum_w, sum_ww) if ok.value else None

    # ---- K10 (mean field) ------------------------------------------------------------------------
    def mf_soft_stats(self, mf_m):
        """Off-diagonal soft block-pair statistics mf_m^T (A o x) mf_m (fp64)."""
        m = np.ascontiguousarray(mf_m, dtype=np.float64)
        K, B = m.shape[1], self.B
        assert m.shape == (self.N, K)
        n_soft = np.zeros((K, K)); sum_w = np.zeros((K, K, B)); sum_ww = np.zeros((K, K, B, B))
        self.ctx.check(self.ctx.lib.gk_mf_sbm_soft_stats(self.ctx.h, self.h, ptr(m), K, ptr(n_soft),
                                                         ptr(sum_w) if B else None, ptr(sum_ww) if B else None))
        return n_soft, sum_w, sum_ww

    def mf_update_c(self, mf_m, E_log_m, base_out, base_in, coef_out, coef_in, stepsize=1.0, block_nodes=0,
                    want_lp=False):
        """Sequential mean-field update of the class probabilities.  Returns (mf_m_new, lp or None)."""
        m = np.array(mf_m, dtype=np.float64, copy=True, order="C")
        K = m.shape[1]
        S = 1 + self.B + self.B * (self.B + 1) // 2
        args = [np.ascontiguousarray(a, dtype=np.float64) for a in (E_log_m, base_out, base_in, coef_out, coef_in)]
        assert args[0].shape == (K,) and args[1].shape == (K, K) and args[2].shape == (K, K)
        assert args[3].shape == (K, K, S) and args[4].shape == (K, K, S)
        lp = np.zeros((self.N, K)) if want_lp else None
        self.ctx.check(self.ctx.lib.gk_mf_sbm_update_c(self.ctx.h, self.h, ptr(m), K, *[ptr(a) for a in args],
                                                       float(stepsize), int(block_nodes), ptr(lp)))
        return m, lp

    def close(self):
        self._finalizer()


def _destroy(ctx, h):
    if ctx.h:
        ctx.lib.gk_network_destroy(ctx.h, h)


def _as_u8(A):
    A = np.asarray(A)
    if A.dtype != np.uint8:
        A = (A != 0).astype(np.uint8)
    return np.ascontiguousarray(A)


def _as_w(W):
    if W is None:
        return None, GK_F32
    W = np.asarray(W)
    if W.dtype == np.float32:
        return np.ascontiguousarray(W), GK_F32
    return np.ascontiguousarray(W, dtype=np.float64), GK_F64


class DeviceEigenmodel:
    """Device state of one eigenmodel over one DeviceNetwork: S = Omega + Omega^T, F, lmbda, mu_0 (gk_eig)."""

    def __init__(self, net, D):
        # a row-sharded network (sharding.ShardedDeviceNetwork) carries its communicator and its local shard
        self.comm = getattr(net, "comm", None)
        local = getattr(net, "local", net)
        self.net, self.local, self.ctx, self.D, self.N = net, local, local.ctx, int(D), net.N
        h = ctypes.c_void_p()
        self.ctx.check(self.ctx.lib.gk_eig_create(self.ctx.h, local.h, self.D, ctypes.byref(h)))
        self.h = h
        self._finalizer = weakref.finalize(self, _destroy_eig, self.ctx, h)
        if self.comm is not None:
            self._acc = self.comm.buffer(int(self.ctx.lib.gk_eig_acc_len(self.h)))
            self._fblk = self.comm.buffer(int(self.ctx.lib.gk_eig_block_nodes(self.h)) * self.D)

    def set_option(self, option, value):
        self.ctx.check(self.ctx.lib.gk_eig_set_option(self.ctx.h, self.h, int(option), int(value)))

    def set_params(self, F, mu_0, lmbda):
        F = np.ascontiguousarray(F, dtype=np.float64)
        lm = np.ascontiguousarray(lmbda, dtype=np.float64)
        assert F.shape == (self.N, self.D) and lm.shape == (self.D,)
        self.ctx.check(self.ctx.lib.gk_eig_set_params(self.ctx.h, self.h, ptr(F), float(mu_0), ptr(lm)))

    def pg_draw(self, seed):
        """K3: S <- Omega + Omega^T with Omega ~ PG(1, Mu); returns sum(Omega) over all N^2 entries."""
        s = ctypes.c_double(0.0)
        self.ctx.check(self.ctx.lib.gk_eig_pg_draw(self.ctx.h, self.h, ctypes.c_uint64(int(seed)), ctypes.byref(s)))
        if self.comm is not None:                   # every unordered pair is drawn by exactly one rank
            return float(self.comm.allreduce_np(np.array([s.value]))[0])
        return s.value

    def set_S(self, S_rows):
        S = np.ascontiguousarray(S_rows, dtype=np.float64)
        assert S.shape == (self.local.nrows, self.N)
        self.ctx.check(self.ctx.lib.gk_eig_set_S(self.ctx.h, self.h, ptr(S)))

    def tn_draw(self, seed):
        """K7 (probit): S buffer <- Z + Z^T with Z ~ truncated N(Mu, 1); returns sum of Z over n != m."""
        s = ctypes.c_double(0.0)
        self.ctx.check(self.ctx.lib.gk_eig_tn_draw(self.ctx.h, self.h, ctypes.c_uint64(int(seed)), ctypes.byref(s)))
        if self.comm is not None:
            return float(self.comm.allreduce_np(np.array([s.value]))[0])
        return s.value

    def set_Zsym(self, Zsym_rows):
        Z = np.ascontiguousarray(Zsym_rows, dtype=np.float64)
        assert Z.shape == (self.local.nrows, self.N)
        self.ctx.check(self.ctx.lib.gk_eig_set_Zsym(self.ctx.h, self.h, ptr(Z)))

    def get_S(self):
        S = np.empty((self.local.nrows, self.N))
        self.ctx.check(self.ctx.lib.gk_eig_get_S(self.ctx.h, self.h, ptr(S)))
        return S

    # ---- K11 (mean field) ----------------------------------------------------------------------------
    def mf_omega(self, mf_mu_F, mf_Sigma_F, E_llT, E_lmbda, mf_mu_mu0, mf_sigma_mu0, n_samples, seed):
        """S <- E_O + E_O^T by Monte Carlo; returns (sum E_O over all N^2 entries, sum_{n != m} E_O * E_FdotF)."""
        assert self.comm is None, "the mean-field eigenmodel path runs on single-rank networks"
        args = [np.ascontiguousarray(a, dtype=np.float64) for a in (mf_mu_F, mf_Sigma_F, E_llT, E_lmbda)]
        assert args[0].shape == (self.N, self.D) and args[1].shape == (self.N, self.D, self.D)
        s0, s1 = ctypes.c_double(0.0), ctypes.c_double(0.0)
        self.ctx.check(self.ctx.lib.gk_eig_mf_omega(self.ctx.h, self.h, *[ptr(a) for a in args], float(mf_mu_mu0),
                                                   float(mf_sigma_mu0), int(n_samples), ctypes.c_uint64(int(seed)),
                                                   ctypes.byref(s0), ctypes.byref(s1)))
        return s0.value, s1.value

    def mf_probit_Z(self, mf_mu_F, E_lmbda, E_mu0, exact_normalizer=False):
        """Probit mean field: S buffer <- E[Z] + E[Z]^T (probit link); returns sum of E[Z] over n != m."""
        assert self.comm is None, "the mean-field eigenmodel path runs on single-rank networks"
        F = np.ascontiguousarray(mf_mu_F, dtype=np.float64); lam = np.ascontiguousarray(E_lmbda, dtype=np.float64)
        assert F.shape == (self.N, self.D) and lam.shape == (self.D,)
        s = ctypes.c_double(0.0)
        self.ctx.check(self.ctx.lib.gk_eig_mf_probit_Z(self.ctx.h, self.h, ptr(F), ptr(lam), float(E_mu0),
                                                      int(bool(exact_normalizer)), ctypes.byref(s)))
        return s.value

    def mf_update_F(self, mf_mu_F, mf_Sigma_F, F_lin, E_llT, E_lmbda, E_mu0, sigma_F, stepsize=1.0, want_conditionals=False):
        """Serial mean-field update of q(F).  Returns (mf_mu_F, mf_Sigma_F, prec or None, h or None)."""
        assert self.comm is None, "the mean-field eigenmodel path runs on single-rank networks"
        mu = np.array(mf_mu_F, dtype=np.float64, copy=True, order="C")
        Sig = np.array(mf_Sigma_F, dtype=np.float64, copy=True, order="C")
        lin = None if F_lin is None else np.ascontiguousarray(F_lin, dtype=np.float64)
        llT = np.ascontiguousarray(E_llT, dtype=np.float64); lam = np.ascontiguousarray(E_lmbda, dtype=np.float64)
        D = self.D
        prec = np.empty((self.N, D, D)) if want_conditionals else None
        h = np.empty((self.N, D)) if want_conditionals else None
        self.ctx.check(self.ctx.lib.gk_eig_mf_update_F(self.ctx.h, self.h, ptr(mu), ptr(Sig), ptr(lin), ptr(llT), ptr(lam),
                                                      float(E_mu0), float(sigma_F), float(stepsize), ptr(prec), ptr(h)))
        return mu, Sig, prec, h

    def mf_pair_stats(self, mf_mu_F, mf_Sigma_F):
        mu = np.ascontiguousarray(mf_mu_F, dtype=np.float64); Sig = np.ascontiguousarray(mf_Sigma_F, dtype=np.float64)
        D = self.D
        st = np.zeros(D * D + 2 * D)
        self.ctx.check(self.ctx.lib.gk_eig_mf_pair_stats(self.ctx.h, self.h, ptr(mu), ptr(Sig), ptr(st)))
        return dict(Pq=st[:D * D].reshape(D, D), hS=st[D * D:D * D + D], hK=st[D * D + D:])

    def resample_F(self, sigma_F, normals, want_conditionals=False):
        """K4+K5.  Returns (F_new, stats dict, prec or None, h or None)."""
        z = np.ascontiguousarray(normals, dtype=np.float64)
        assert z.shape == (self.N, self.D)
        D = self.D
        if self.comm is not None:
            assert not want_conditionals, "per-node conditionals are available on single-rank networks only"
            if not self.net.native():
                return self._resample_F_sharded(sigma_F, z)
            # the engine's own communicator: one C call, collectives in-stream; the statistics come back already summed
        F = np.empty((self.N, D))
        prec = np.empty((self.N, D, D)) if want_conditionals else None
        h = np.empty((self.N, D)) if want_conditionals else None
        st = np.zeros(D * D + 2 * D)
        self.ctx.check(self.ctx.lib.gk_eig_resample_F(self.ctx.h, self.h, float(sigma_F), ptr(z), ptr(F), ptr(prec),
                                                     ptr(h), ptr(st)))
        stats = dict(Pq=st[:D * D].reshape(D, D).copy(), hS=st[D * D:D * D + D].copy(), hK=st[D * D + D:].copy())
        return F, stats, prec, h

    def _resample_F_sharded(self, sigma_F, z):
        """Block form (include/gk.h): slab partials on every rank -> all-reduce -> the owner of the block's rows draws
        it -> feature broadcast; the pair statistics are summed over the ranks at the end."""
        from .sharding import row_range
        lib, c, h, comm, D = self.ctx.lib, self.ctx.h, self.h, self.comm, self.D
        acc, fblk = ctypes.c_void_p(self._acc.data_ptr()), ctypes.c_void_p(self._fblk.data_ptr())
        self.ctx.check(lib.gk_eig_sweep_begin(c, h, float(sigma_F), ptr(z)))
        bn = int(lib.gk_eig_block_nodes(h))
        for rk in range(comm.world):
            s0, s1 = row_range(self.N, rk, comm.world)
            for n0 in range(s0, s1, bn):
                nb = min(bn, s1 - n0)
                self.ctx.check(lib.gk_eig_block_partials(c, h, n0, nb, acc))
                comm.allreduce_(self._acc)
                if comm.rank == rk:
                    self.ctx.check(lib.gk_eig_block_sample(c, h, n0, nb, acc, fblk))
                if comm.dist is not None and comm.world > 1:
                    comm.dist.broadcast(self._fblk, src=comm._global(rk), group=comm.group)
                if comm.rank != rk:
                    self.ctx.check(lib.gk_eig_set_block(c, h, n0, nb, fblk))
        F = np.empty((self.N, D))
        st = np.zeros(D * D + 2 * D)
        self.ctx.check(lib.gk_eig_sweep_end(c, h, ptr(F), ptr(st)))
        st = comm.allreduce_np(st)
        stats = dict(Pq=st[:D * D].reshape(D, D).copy(), hS=st[D * D:D * D + D].copy(), hK=st[D * D + D:].copy())
        return F, stats, None, None

    def log_likelihood(self, link=0):
        v = ctypes.c_double(0.0)
        self.ctx.check(self.ctx.lib.gk_eig_log_likelihood(self.ctx.h, self.h, int(link), ctypes.byref(v)))
        if self.comm is not None:
            return float(self.comm.allreduce_np(np.array([v.value]))[0])
        return v.value

    def close(self):
        self._finalizer()


def _destroy_eig(ctx, h):
    if ctx.h:
        ctx.lib.gk_eig_destroy(ctx.h, h)


class DeviceLatent:
    """Device state of one latent-distance adjacency model over one DeviceNetwork (gk_latent): L, mu_0 on the
    device; every O(N^2) evaluation is one K8 pair pass."""

    def __init__(self, net, dim):
        # a row-sharded network (sharding.ShardedDeviceNetwork) carries its communicator and its local shard
        self.comm = getattr(net, "comm", None)
        local = getattr(net, "local", net)
        self.net, self.ctx, self.dim, self.N = net, local.ctx, int(dim), net.N
        h = ctypes.c_void_p()
        self.ctx.check(self.ctx.lib.gk_latent_create(self.ctx.h, local.h, self.dim, ctypes.byref(h)))
        self.h = h
        self._finalizer = weakref.finalize(self, _destroy_latent, self.ctx, h)
        self._acc = self.comm.buffer(self.acc_len()) if self.comm is not None else None

    def acc_len(self):
        return int(self.ctx.lib.gk_latent_acc_len(self.h))

    def set_params(self, L, mu_0, mu_self, sigma):
        Lc = None if L is None else np.ascontiguousarray(L, dtype=np.float64)
        assert Lc is None or Lc.shape == (self.N, self.dim)
        self.ctx.check(self.ctx.lib.gk_latent_set_params(self.ctx.h, self.h, ptr(Lc), float(mu_0), float(mu_self),
                                                         float(sigma)))

    def get_params(self):
        L = np.empty((self.N, self.dim))
        mu0 = ctypes.c_double(0.0)
        self.ctx.check(self.ctx.lib.gk_latent_get_params(self.ctx.h, self.h, ptr(L), ctypes.byref(mu0)))
        return L, mu0.value

    def lp_grad(self):
        """(ll, lp, dlp/dL, dlp/dmu_0, dlp/dmu_self) at the current parameters."""
        if self.comm is not None:
            acc = ctypes.c_void_p(self._acc.data_ptr())
            self.eval_pass(acc)
            self.comm.allreduce_(self._acc)
            return self.eval_finish(acc)
        ll, lp, g0, gs = (ctypes.c_double(0.0) for _ in range(4))
        gL = np.empty((self.N, self.dim))
        self.ctx.check(self.ctx.lib.gk_latent_lp_grad(self.ctx.h, self.h, ctypes.byref(ll), ctypes.byref(lp), ptr(gL),
                                                      ctypes.byref(g0), ctypes.byref(gs)))
        return ll.value, lp.value, gL, g0.value, gs.value

    def eval_pass(self, acc_ptr=None):
        self.ctx.check(self.ctx.lib.gk_latent_eval_pass(self.ctx.h, self.h, acc_ptr))

    def eval_finish(self, acc_ptr=None):
        ll, lp, g0, gs = (ctypes.c_double(0.0) for _ in range(4))
        gL = np.empty((self.N, self.dim))
        self.ctx.check(self.ctx.lib.gk_latent_eval_finish(self.ctx.h, self.h, acc_ptr, ctypes.byref(ll), ctypes.byref(lp),
                                                          ptr(gL), ctypes.byref(g0), ctypes.byref(gs)))
        return ll.value, lp.value, gL, g0.value, gs.value

    def hmc(self, target, step, nsteps, p0, log_u, allreduce=None, acc_ptr=None):
        """One hips-style HMC transition on the device.  target 0 = L (p0: N x dim), 1 = mu_0 (p0: scalar).
        ``allreduce`` (row-sharded form) is called between every pair pass and the update that consumes it and
        must sum the accumulator at ``acc_ptr`` over the ranks.  Returns (accepted, log_ratio)."""
        p = np.ascontiguousarray(np.atleast_1d(p0), dtype=np.float64)
        if allreduce is None and self.comm is not None and not self.net.native():
            allreduce, acc_ptr = (lambda: self.comm.allreduce_(self._acc)), ctypes.c_void_p(self._acc.data_ptr())
        acc, ratio = ctypes.c_int(0), ctypes.c_double(0.0)
        lib, c, h = self.ctx.lib, self.ctx.h, self.h
        if allreduce is None:
            self.ctx.check(lib.gk_latent_hmc(c, h, int(target), float(step), int(nsteps), ptr(p), float(log_u),
                                             ctypes.byref(acc), ctypes.byref(ratio)))
        else:
            self.ctx.check(lib.gk_latent_hmc_begin(c, h, int(target), float(step), int(nsteps), ptr(p)))
            for _ in range(int(nsteps) + 1):
                self.ctx.check(lib.gk_latent_hmc_pass(c, h, acc_ptr))
                allreduce()
                self.ctx.check(lib.gk_latent_hmc_advance(c, h, acc_ptr))
            self.ctx.check(lib.gk_latent_hmc_end(c, h, float(log_u), ctypes.byref(acc), ctypes.byref(ratio)))
        return bool(acc.value), ratio.value

    def set_weight_mode(self, on, lam=1.0):
        """Switch the passes to the Gaussian-weight target of weights.py:697-731 (set_params' mu_0 = b, sigma = eta)."""
        self.ctx.check(self.ctx.lib.gk_latent_set_weight_mode(self.ctx.h, self.h, int(bool(on)), float(lam)))

    def weight_stats(self):
        """(n, sum r, sum r^2) over the off-diagonal edges, r = W + d^2 - b, at the current (L, b)."""
        if self.comm is not None:
            acc = ctypes.c_void_p(self._acc.data_ptr())
            self.ctx.check(self.ctx.lib.gk_latent_weight_stats_pass(self.ctx.h, self.h, acc))
            self.comm.allreduce_(self._acc)
            ex = self._acc[self.N * self.dim:].cpu().numpy()
            return float(ex[2]), float(ex[0]), float(ex[1])
        out = np.zeros(3)
        self.ctx.check(self.ctx.lib.gk_latent_weight_stats(self.ctx.h, self.h, ptr(out)))
        return float(out[0]), float(out[1]), float(out[2])

    def close(self):
        self._finalizer()


def _destroy_latent(ctx, h):
    if ctx.h:
        ctx.lib.gk_latent_destroy(ctx.h, h)


def truncnorm_draw(mu, edge, u, ctx=None):
    """Device form of the reference's ``sample_truncnorm`` (internals/utils.py:53-76) with the uniforms supplied:
    N(mu, 1) truncated to [0, inf) where ``edge`` else (-inf, 0]."""
    ctx = ctx or _lib.default_context()
    mu = np.ascontiguousarray(mu, dtype=np.float64)
    e8 = np.ascontiguousarray(np.asarray(edge) != 0, dtype=np.uint8)
    uu = np.ascontiguousarray(u, dtype=np.float64)
    assert mu.shape == e8.shape == uu.shape
    out = np.empty_like(mu)
    ctx.check(ctx.lib.gk_truncnorm_draw(ctx.h, ptr(mu), ptr(e8), ptr(uu), ptr(out), mu.size))
    return out


class DeviceAIS:
    """P co-resident AIS particles over one bit-packed adjacency matrix: the device side of
    ``internals.graph_models.estimate_marginal_likelihood_batched`` (examples/demo_ais.py:10-62, internals/graph_models.py
    :261-297, 337-371).  A: N x N {0,1} host array, an unsharded DeviceNetwork, or None (rows are then supplied with
    ``upload_rows``, host arrays or device pointers)."""

    def __init__(self, A, P, R, ctx=None, N=None):
        self.ctx = ctx or _lib.default_context()
        if isinstance(A, DeviceNetwork):
            self.N = A.N
        elif A is not None:
            A = _as_u8(A)
            self.N = int(A.shape[0])
            assert A.shape == (self.N, self.N)
        else:
            self.N = int(N)
        self.P, self.R = int(P), int(R)
        h = ctypes.c_void_p()
        self.ctx.check(self.ctx.lib.gk_ais_create(self.ctx.h, self.N, self.P, self.R, ctypes.byref(h)))
        self.h = h
        self._finalizer = weakref.finalize(self, _destroy_ais, self.ctx, h)
        if isinstance(A, DeviceNetwork):
            self.ctx.check(self.ctx.lib.gk_ais_upload_network(self.ctx.h, self.h, A.h))
        elif A is not None:
            step = max(32, (1 << 28) // self.N // 32 * 32)              # <= 256 MB of rows per staging copy
            for r0 in range(0, self.N, step):
                self.upload_rows(A[r0:r0 + step], r0)

    def upload_rows(self, rows, row0, device_ptr=None, nrows=None):
        """rows [row0, row0 + nrows) of A as a host uint8 array, or ``device_ptr`` (int) to nrows x N uint8 on the device."""
        if device_ptr is not None:
            self.ctx.check(self.ctx.lib.gk_ais_upload_rows(self.ctx.h, self.h, ctypes.c_void_p(device_ptr), int(row0), int(nrows), 1))
            return
        r8 = _as_u8(rows)
        self.ctx.check(self.ctx.lib.gk_ais_upload_rows(self.ctx.h, self.h, ptr(r8), int(row0), int(r8.shape[0]), 0))

    def device_bytes(self):
        return int(self.ctx.lib.gk_ais_device_bytes(self.h))

    def set_option(self, option, value):
        self.ctx.check(self.ctx.lib.gk_ais_set_option(self.ctx.h, self.h, int(option), int(value)))

    def nnz(self):
        v = ctypes.c_int64()
        self.ctx.check(self.ctx.lib.gk_ais_nnz(self.ctx.h, self.h, ctypes.byref(v)))
        return int(v.value)

    def set_labels(self, z):
        z32 = np.ascontiguousarray(z, dtype=np.int32)
        assert z32.shape in ((self.N,), (self.P, self.N))
        self.ctx.check(self.ctx.lib.gk_ais_set_labels(self.ctx.h, self.h, ptr(z32), int(z32.ndim == 1)))

    def get_labels(self):
        z = np.zeros((self.P, self.N), np.int32)
        self.ctx.check(self.ctx.lib.gk_ais_get_labels(self.ctx.h, self.h, ptr(z)))
        return z

    def counts(self, recompute=False):
        """(n_edge[P, R, R], sizes[P, R]) as int64, diagonal included."""
        n_edge = np.zeros((self.P, self.R, self.R), np.int64)
        sizes = np.zeros((self.P, self.R), np.int64)
        self.ctx.check(self.ctx.lib.gk_ais_counts(self.ctx.h, self.h, int(recompute), ptr(n_edge), ptr(sizes)))
        return n_edge, sizes

    def sweep(self, log_pi, B, beta, uniforms=None, seed=0, want_lp=False):
        """One exact sequential feature sweep of every particle.  Returns (n_changed[P], lp[P, N, R] or None)."""
        P, R = self.P, self.R
        lpi = np.ascontiguousarray(np.broadcast_to(np.asarray(log_pi, np.float64), (P, R)))
        Bm = np.ascontiguousarray(np.broadcast_to(np.asarray(B, np.float64), (P, R, R)))
        bt = np.ascontiguousarray(np.broadcast_to(np.asarray(beta, np.float64), (P,)))
        u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float64)
        assert u is None or u.shape == (P, self.N)
        lp = np.zeros((P, self.N, R)) if want_lp else None
        nchg = np.zeros(P, np.int64)
        self.ctx.check(self.ctx.lib.gk_ais_sweep(self.ctx.h, self.h, ptr(lpi), ptr(Bm), ptr(bt), ptr(u),
                                                 ctypes.c_uint64(int(seed) & (2 ** 64 - 1)), ptr(lp), ptr(nchg)))
        return nchg, lp

    def run(self, model, b0, b1, alpha0, n_temps, steps_per_temp, f0=None, seed=0):
        """gk_ais_run: the whole AIS ladder for the P particles; returns log_weights[P]."""
        w = np.zeros(self.P)
        a0 = None if alpha0 is None else np.ascontiguousarray(alpha0, dtype=np.float64)
        f32 = None if f0 is None else np.ascontiguousarray(f0, dtype=np.int32)
        self.ctx.check(self.ctx.lib.gk_ais_run(self.ctx.h, self.h, int(model), float(b0), float(b1), ptr(a0), int(n_temps),
                                               int(steps_per_temp), ptr(f32), ctypes.c_uint64(int(seed) & (2 ** 64 - 1)), ptr(w)))
        return w

    def close(self):
        self._finalizer()


def _destroy_ais(ctx, h):
    if getattr(ctx, "h", None):
        ctx.lib.gk_ais_destroy(ctx.h, h)
