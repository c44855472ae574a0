"""Aldous-Hoover graph models for AIS marginal-likelihood estimation (mirrors ``graphistician.internals.graph_models``
:12-371, 529-544 and the driver of examples/demo_ais.py:10-62).

The O(N^2) work of every AIS step runs in libgk on the device copy of A:

  ErdosRenyiNetwork.sample_theta / log_lkhd   <- one gk_block_stats pass (K = 1), cached: nnz(A) never changes
  StochasticBlockModel.sample_theta           <- gk_block_stats (tempered Beta posterior, :337-371)
  StochasticBlockModel.sample_f for n = 0..N-1 (the driver's loop, demo_ais.py:44-46)
                                               <- gk_sbm_sweep with beta, diagonal counted in both directions
                                                  (GK_SELF_TWICE, :275-293) and hips' log_sum_exp_sample (GK_SAMPLER_LSE)
  log_lkhd                                     <- gk_block_stats (replaces the N^2 Python loop of :42-48)

Independent AIS particles are partitioned over the ranks of a ``sharding.Comm``: no data-path
collective, one gather of the log weights at the end.
"""
import copy

import numpy as np
from scipy.special import betaln, logsumexp

from .. import _lib
from ..old_abstractions import network_for


def _device(A, dtype="f32"):
    return network_for(A).device(dtype)


def _total_edge_stats(A, z, R):
    """(edges, ordered pairs) per block cell, diagonal INCLUDED (the loops of :42-48, :354-367 run over all i, j)."""
    dev = _device(A)
    z = np.asarray(z, dtype=np.int64)
    n_edge = dev.block_stats(z, R)[0]
    d = dev.diag_stats(z, R)
    k = np.arange(R)
    n_edge[k, k] += d["d_edge"]
    sz = d["d_cnt"]
    return n_edge, np.outer(sz, sz)


class AldousHooverNetwork(object):
    """internals/graph_models.py:12-102."""

    def __init__(self, A=None):
        pass

    def pr_A_given_f(self, f1, f2, theta):
        raise Exception("pr_A_given_f is not implemented!")

    def logpr_f(self, f, theta):
        raise Exception("logpr_f is not implemented!")

    def logpr_theta(self, theta):
        raise Exception("logpr_theta is not implemented!")

    def log_lkhd(self, A, f, theta):
        raise Exception("log_lkhd is not implemented!")

    def logpr(self, A, f, theta, beta=1.0):
        """internals/graph_models.py:50-64."""
        lp = self.logpr_theta(theta)
        lp += self.logpr_fs(f, theta)
        lp += beta * self.log_lkhd(A, f, theta)
        return lp

    def logpr_fs(self, f, theta):
        return sum(self.logpr_f(fi, theta) for fi in f)

    def sample_f(self, theta, naf=(None, None, None), beta=1.0):
        raise Exception("sample f is not implemented!")

    def sample_fs(self, theta, A, f, beta=1.0, uniforms=None):
        """The driver's inner loop ``for n in arange(N): f[n] = model.sample_f(theta, (n, A, f), beta)``
        (examples/demo_ais.py:44-46) as one call."""
        for n in range(len(f)):
            f[n] = self.sample_f(theta, (n, A, f), beta=beta)
        return f

    def sample_theta(self, Af=(None, None), beta=1.0):
        raise Exception("sample_theta is not implemented!")

    def sample_A(self, f, theta):
        """internals/graph_models.py:86-96 (row-major uniforms, one per entry)."""
        N = len(f)
        P = self.P_matrix(f, theta)
        return (np.random.rand(N, N) < P).astype(np.float64)

    def sample_Aij(self, fi, fj, theta):
        return np.random.rand() < self.pr_A_given_f(fi, fj, theta)


class ErdosRenyiNetwork(AldousHooverNetwork):
    """internals/graph_models.py:105-182."""

    def __init__(self, rho=None, x=None):
        if rho is None:
            if x is None:
                raise Exception("Either rho or (a,b) must be specified")
            (a, b) = x
        else:
            if x is not None:
                raise Exception("Either rho or (a,b) must be specified")
            a = b = None
        self.rho, self.a, self.b = rho, a, b

    def pr_A_given_f(self, fi, fj, theta):
        return theta

    def P_matrix(self, f, theta):
        return theta * np.ones((len(f), len(f)))

    def logpr_f(self, f, theta):
        return 0.0

    def logpr_fs(self, f, theta):
        return 0.0

    def logpr_theta(self, theta):
        rho = theta
        if self.rho is not None:
            return 0.0 if self.rho == rho else -np.inf
        return (self.a - 1.0) * np.log(rho) + (self.b - 1.0) * np.log(1.0 - rho) - betaln(self.a, self.b)

    def _nnz(self, A):
        """sum(A) over all N^2 entries from the device (cached on the network object: the data never changes)."""
        net = network_for(A)
        if not hasattr(net, "_nnz_total"):
            n_edge, _ = _total_edge_stats(A, np.zeros(net.N, np.int64), 1)
            net._nnz_total = float(n_edge[0, 0])
        return net._nnz_total

    def log_lkhd(self, A, f, theta):
        N = A.shape[0]
        nnz = self._nnz(A)
        return nnz * np.log(theta) + (N ** 2 - nnz) * np.log(1 - theta)

    def sample_f(self, theta, naf=(None, None, None), beta=1.0):
        return None

    def sample_fs(self, theta, A, f, beta=1.0, uniforms=None):
        return f

    def sample_theta(self, Af=(None, None), beta=1.0):
        """internals/graph_models.py:168-182: rho ~ Beta(a + beta nnz, b + beta (N^2 - nnz))."""
        A, f = Af
        if self.rho is not None:
            return self.rho
        if A is None and f is None:
            return np.random.beta(self.a, self.b)
        N = A.shape[0]
        nnz = self._nnz(A)
        return np.random.beta(self.a + beta * nnz, self.b + beta * (N ** 2 - nnz))


class StochasticBlockModel(AldousHooverNetwork):
    """internals/graph_models.py:185-371.  theta = (B, pi): B[r, r'] edge probability between blocks, pi block weights;
    features f = block labels.  (The reference stores B as float32, :353; values here are rounded through float32
    the same way and used in fp64 arithmetic.)"""

    def __init__(self, R, b0, b1, alpha0):
        self.R, self.b0, self.b1, self.alpha0 = R, b0, b1, np.asarray(alpha0, dtype=np.float64)

    def pr_A_given_f(self, fi, fj, theta):
        (B, pi) = theta
        return B[fi, fj]

    def P_matrix(self, f, theta):
        (B, pi) = theta
        z = np.asarray(f, dtype=np.int64)
        return B[np.ix_(z, z)]

    def logpr_f(self, f, theta):
        (B, pi) = theta
        return float(np.sum(np.log(pi[np.array(f).astype(int)])))

    def logpr_fs(self, f, theta):
        return self.logpr_f(f, theta)

    def logpr_theta(self, theta):
        """internals/graph_models.py:222-240, including its quirk of adding the Dirichlet term R times (:237-238)."""
        (B, pi) = theta
        lp = np.sum((self.b1 - 1.0) * np.log(B) + (self.b0 - 1.0) * np.log(1.0 - B) - betaln(self.b1, self.b0))
        lp += self.R * np.sum((self.alpha0 - 1.0) * np.log(pi))
        return float(lp)

    def log_lkhd(self, A, f, theta):
        """internals/graph_models.py:42-48 (an N^2 Python loop) from the block counts of one K1 pass."""
        (B, pi) = theta
        n_edge, n_pair = _total_edge_stats(A, f, self.R)
        B = np.asarray(B, dtype=np.float64)
        with np.errstate(divide="ignore", invalid="ignore"):
            t1 = np.where(n_edge > 0, n_edge * np.log(B), 0.0)
            t0 = np.where(n_pair - n_edge > 0, (n_pair - n_edge) * np.log(1 - B), 0.0)
        return float(np.sum(t1 + t0))

    def sample_f(self, theta, naf=(None, None, None), beta=1.0):
        """One node (:242-259): the prior draw; conditional single-node draws go through ``sample_fs``."""
        (B, pi) = theta
        n, A, f = naf
        if n is None and A is None and f is None:
            return int(np.sum(np.random.rand() > np.cumsum(pi)))           # hips discrete_sample: inverse cdf
        z = np.array(f, dtype=np.int64)
        u = np.full(len(z), 0.5); u[n] = np.random.rand()
        return int(self._sweep(theta, A, z, beta, u, only=n)[n])

    def collapsed_sample_f(self, theta, n, A, f, uniform=None, return_lp=False):
        """internals/graph_models.py:299-335 (unused by the reference's own drivers): Gibbs update of z[n] with the block
        probabilities B integrated out, ln p(z_n = r) = ln pi[r] + sum_{r1, r2} [betaln(m + b1, mhat + b0) - betaln(b1, b0)]
        with m / mhat the edges / non-edges of A[np.ix_(z == r1, z == r2)] under z_n = r.  The reference rebuilds every one of
        the R^2 sub-matrices for every candidate (O(R^3 N^2) per node); here the R x R table comes from ONE device pass
        (gk_block_stats, cached per labelling by the caller through ``table``) and is moved between candidates by node n's own
        row and column counts.  The reference draws inside its r loop (:331-333, an indentation slip that costs R - 1
        uniforms); the label returned is the draw from the complete vector, as there."""
        (B, pi) = theta
        z = np.array(f, dtype=np.int64)
        R = self.R
        Ab = np.asarray(A) != 0
        m, _ = _total_edge_stats(A, z, R)
        m = m.astype(np.float64)
        sz = np.bincount(z, minlength=R).astype(np.float64)
        a = int(z[n]); s = float(Ab[n, n])
        oth = np.ones(len(z), bool); oth[n] = False
        co = np.bincount(z[Ab[n, :] & oth], minlength=R).astype(np.float64)      # n -> class r
        ci = np.bincount(z[Ab[:, n] & oth], minlength=R).astype(np.float64)      # class r -> n
        m0 = m.copy(); m0[a, :] -= co; m0[:, a] -= ci; m0[a, a] -= s
        sz0 = sz.copy(); sz0[a] -= 1
        lp = np.log(np.asarray(pi, dtype=np.float64)).copy()
        for r in range(R):
            mr = m0.copy(); mr[r, :] += co; mr[:, r] += ci; mr[r, r] += s
            szr = sz0.copy(); szr[r] += 1
            pairs = np.outer(szr, szr)
            lp[r] += np.sum(betaln(mr + self.b1, pairs - mr + self.b0) - betaln(self.b1, self.b0))
        if uniform is None:
            for _ in range(R - 1):                   # the draws the reference makes (and discards) inside its loop
                np.random.rand()
            uniform = np.random.rand()
        mx = lp.max(); p = np.exp(lp - mx); p /= p.sum()
        zn = int(min(np.searchsorted(np.cumsum(p), uniform, side="left"), R - 1))
        return (zn, lp) if return_lp else zn

    def _sweep(self, theta, A, z, beta, u, only=None):
        (B, pi) = theta
        dev = _device(A)
        if only is not None:
            raise NotImplementedError("single-node conditional draws: use sample_fs (the driver's full loop)")
        c, _, _ = dev.sbm_sweep(z, np.log(pi), np.asarray(B, dtype=np.float64), None, None, u,
                                self_mode=_lib.SELF_TWICE, sampler=_lib.SAMPLER_LSE, beta=float(beta))
        return c

    def sample_fs(self, theta, A, f, beta=1.0, uniforms=None):
        """``for n in arange(N): f[n] = model.sample_f(theta, (n, A, f), beta)`` (examples/demo_ais.py:44-46) as ONE
        exact sequential device sweep: naive_sample_f (:261-297) for n = 0..N-1 in order, one uniform per node."""
        z = np.array(f, dtype=np.int64)
        u = np.random.rand(len(z)) if uniforms is None else uniforms
        return self._sweep(theta, A, z, beta, u)

    def sample_theta(self, Af=(None, None), beta=1.0):
        """internals/graph_models.py:337-371: pi ~ Dir(alpha0 + counts) (NOT tempered), then
        B[r1, r2] ~ Beta(b1 + beta edges, b0 + beta non-edges) over A[z == r1][:, z == r2], row-major."""
        A, f = Af
        if A is None and f is None:
            B = np.random.beta(self.b1, self.b0, (self.R, self.R))
            pi = np.random.dirichlet(self.alpha0)
            return (B, pi)
        z = np.array(f, dtype=np.int64)
        pi = np.random.dirichlet(self.alpha0 + np.bincount(z, minlength=self.R))
        n_edge, n_pair = _total_edge_stats(A, z, self.R)
        B = np.random.beta(self.b1 + beta * n_edge, self.b0 + beta * (n_pair - n_edge))
        return (B.astype(np.float32).astype(np.float64), pi)


def invariant_sbm_order(f, R):
    """The free function examples/demo_ais.py:87 calls (only the methods adjacency.py:575 /
    stochastic_block_model.py:165 exist in the reference): nodes ordered by block, blocks ordered by size."""
    z = np.asarray(f, dtype=np.int64)
    corder = np.argsort(np.bincount(z, minlength=R))
    newc = np.zeros(len(z))
    for c in range(R):
        newc[z == corder[c]] = c
    return np.argsort(-newc)


def sample_network(model, N):
    """internals/graph_models.py:529-544."""
    theta = model.sample_theta()
    f = [model.sample_f(theta) for _ in range(N)]
    A = model.sample_A(f, theta)
    return (A, f, theta)


def fit_network(A, model, x0=None, N_iter=1000, callback=None, verbose=False):
    """internals/graph_models.py:546-594: Gibbs on (theta, f) given A; the inner ``for n: f[n] = sample_f(...)`` loop
    is one sequential device sweep (``sample_fs``).  Returns (f_trace, theta_trace, lp_trace)."""
    N = A.shape[0]
    if x0 is None:
        theta0 = model.sample_theta()
        f0 = [model.sample_f(theta0) for _ in range(N)]
    else:
        (f0, theta0) = x0
    f, theta = copy.deepcopy(f0), copy.deepcopy(theta0)
    lp_trace = np.zeros(N_iter)
    f_trace, theta_trace = [], []
    for it in range(N_iter):
        lp_trace[it] = model.logpr(A, f, theta)
        if verbose:
            print("Iteration %d. \tlog pr: %f" % (it, lp_trace[it]))
        theta = model.sample_theta((A, f))
        f = model.sample_fs(theta, A, f)
        if callback is not None:
            callback(f, theta)
        f_trace.append(np.array(f))             # the reference appends the list it keeps mutating (:587); copies here
        theta_trace.append(copy.deepcopy(theta))
    return (f_trace, theta_trace, lp_trace)


def geweke_test(N, model, N_iter=1000, callback=None, verbose=False):
    """internals/graph_models.py:597-650: successive-conditional sampler theta | A, f ; f | A, theta ; A | f, theta.
    A changes every iteration, so every iteration uploads a new network (N is small in the reference's use)."""
    theta0 = model.sample_theta()
    f0 = [model.sample_f(theta0) for _ in range(N)]
    A = np.asarray(model.sample_A(f0, theta0))
    f, theta = copy.deepcopy(f0), copy.deepcopy(theta0)
    lp_trace = np.zeros(N_iter)
    f_trace, theta_trace = [], []
    for it in range(N_iter):
        lp_trace[it] = model.logpr(A, f, theta)
        if verbose:
            print("Iteration %d. \tlog pr: %f" % (it, lp_trace[it]))
        theta = model.sample_theta((A, f))
        f = model.sample_fs(theta, A, f)
        A = np.asarray(model.sample_A(f, theta))
        if callback is not None:
            callback(A, f, theta)
        theta_trace.append(copy.deepcopy(theta))
        f_trace.append(np.array(f))
    return (f_trace, theta_trace, lp_trace)


def estimate_marginal_likelihood(A, f0, theta0, model, N_samples=1000, B=100, steps_per_B=11, comm=None,
                                 return_weights=False):
    """AIS estimate of log p(A | model) (examples/demo_ais.py:10-62): B temperatures, ``steps_per_B`` Gibbs steps
    (sample_theta, then the sequential feature sweep) per temperature, log weight = sum_b (beta_b - beta_{b-1}) ll.

    With ``comm`` (sharding.Comm) the particles m = rank, rank + world, ... run on this rank's GPU and the log
    weights are gathered at the end (particles partitioned, no data-path collective); every rank must
    then seed ``np.random`` differently."""
    betas = np.linspace(0, 1, B)
    rank, world = (comm.rank, comm.world) if comm is not None else (0, 1)
    mine = list(range(rank, N_samples, world))
    log_weights = np.zeros(N_samples)
    for m in mine:
        f, theta = copy.deepcopy((f0, theta0))
        ratios = np.zeros(B - 1)
        for b in range(1, B):
            beta = betas[b]
            for s in range(steps_per_B):
                theta = model.sample_theta((A, f), beta=beta)
                f = model.sample_fs(theta, A, f, beta=beta)
            # logpr(beta_b) - logpr(beta_{b-1}) at the same (f, theta): the prior terms cancel exactly
            ratios[b - 1] = (beta - betas[b - 1]) * model.log_lkhd(A, f, theta)
        log_weights[m] = np.sum(ratios)
    if comm is not None and world > 1:
        log_weights = comm.allreduce_np(log_weights)
    log_Z = -np.log(N_samples) + logsumexp(log_weights)
    return (log_Z, log_weights) if return_weights else log_Z


def estimate_marginal_likelihood_batched(A, f0, theta0, model, N_samples=1000, B=100, steps_per_B=11, particles=None,
                                         comm=None, return_weights=False, device_uniforms=None, seeds=None, dev=None):
    """``estimate_marginal_likelihood`` (examples/demo_ais.py:10-62) with the particles CO-RESIDENT on the device
    (SURVEY.md row a17 / config 5): A is bit-packed once (N^2 / 8 bytes, plus its transpose), every streamed row of A
    serves all particles of a batch (``DeviceAIS``: gk_ais_sweep), the serial feature sweeps of the particles run side by
    side, and the R x R edge counts that ``sample_theta`` (internals/graph_models.py:337-371) and ``log_lkhd`` (:42-48) need
    are kept current by the sweep itself.

    Random numbers: the reference consumes ONE global stream particle after particle, which a batch cannot reproduce;
    here particle m owns ``np.random.RandomState(seeds[m])`` (seeds drawn from the global ``np.random`` unless given) and
    consumes it exactly as the reference's loop body does: Dirichlet, R^2 Betas row-major, then the N uniforms of the
    feature sweep.  ``device_uniforms`` (default for N > 8192) replaces those N uniforms per sweep by the device's
    counter-based stream.  With ``comm`` the particles m = rank, rank + world, ... run on this rank (no data-path
    collective, one gather of the log weights)."""
    from ..device import DeviceAIS
    betas = np.linspace(0, 1, B)
    rank, world = (comm.rank, comm.world) if comm is not None else (0, 1)
    if seeds is None:
        seeds = np.random.randint(0, 2 ** 31 - 1, size=N_samples)
    mine = list(range(rank, N_samples, world))
    log_weights = np.zeros(N_samples)
    N = A.shape[0] if A is not None else dev.N
    if device_uniforms is None:
        device_uniforms = N > 8192
    is_er = isinstance(model, ErdosRenyiNetwork)
    R = 1 if is_er else model.R
    P = len(mine) if particles is None else min(int(particles), max(1, len(mine)))
    for lo in range(0, len(mine), max(P, 1)):
        batch = mine[lo:lo + P]
        rs = [np.random.RandomState(int(seeds[m])) for m in batch]
        if dev is None or dev.P != len(batch) or dev.R != R:
            dev = DeviceAIS(A, len(batch), R)
        w = np.zeros(len(batch))
        if is_er:
            nnz = float(dev.nnz())
            for k, r in enumerate(rs):
                for b in range(1, B):
                    for s in range(steps_per_B):
                        rho = model.rho if model.rho is not None else r.beta(model.a + betas[b] * nnz, model.b + betas[b] * (N ** 2 - nnz))
                    w[k] += (betas[b] - betas[b - 1]) * (nnz * np.log(rho) + (N ** 2 - nnz) * np.log(1 - rho))
        else:
            dev.set_labels(np.asarray(f0, dtype=np.int32))
            n_edge, sizes = dev.counts()
            Bm = np.zeros((len(batch), R, R)); lpi = np.zeros((len(batch), R))
            for b in range(1, B):
                beta = betas[b]
                for s in range(steps_per_B):
                    u = None if device_uniforms else np.zeros((len(batch), N))
                    for k, r in enumerate(rs):
                        # sample_theta (:337-371): pi ~ Dir(alpha0 + sizes) untempered; B ~ Beta(b1 + beta e, b0 + beta (pairs - e)) as float32
                        lpi[k] = np.log(r.dirichlet(model.alpha0 + sizes[k]))
                        n_pair = np.outer(sizes[k], sizes[k])
                        Bm[k] = r.beta(model.b1 + beta * n_edge[k], model.b0 + beta * (n_pair - n_edge[k])).astype(np.float32)
                        if u is not None:
                            u[k] = r.rand(N)
                    dev.sweep(lpi, Bm, beta, uniforms=u, seed=int(seeds[batch[0]]))
                    n_edge, sizes = dev.counts()
                for k in range(len(batch)):
                    n_pair = np.outer(sizes[k], sizes[k])
                    with np.errstate(divide="ignore", invalid="ignore"):
                        t1 = np.where(n_edge[k] > 0, n_edge[k] * np.log(Bm[k]), 0.0)
                        t0 = np.where(n_pair - n_edge[k] > 0, (n_pair - n_edge[k]) * np.log(1 - Bm[k]), 0.0)
                    w[k] += (beta - betas[b - 1]) * float(np.sum(t1 + t0))
        log_weights[batch] = w
    if comm is not None and world > 1:
        log_weights = comm.allreduce_np(log_weights)
    log_Z = -np.log(N_samples) + logsumexp(log_weights)
    return (log_Z, log_weights) if return_weights else log_Z
