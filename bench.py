#!/usr/bin/env python
"""bench.py -- Gibbs sweeps/s of the B200 engine on BASELINE.json's metric, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload sbm|eigen|latent]

A "step" is one full ``model.resample(network)`` Gibbs sweep on synthetic data (seed 1234, generators of
SURVEY.md 8d).  Default workload: Gaussian SBM (configs[1] model: K=16, B=2) at the metric's N=16384.
One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "gibbs_sweeps_per_sec"
UNIT = "sweeps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="sbm", choices=["sbm", "eigen", "latent"])
    ap.add_argument("--n", type=int, default=16384, help="nodes (metric: N=16k)")
    ap.add_argument("--k", type=int, default=16, help="SBM blocks / eigenmodel latent dimension (8)")
    ap.add_argument("--b", type=int, default=2, help="weight dimension")
    ap.add_argument("--d", type=int, default=8, help="eigenmodel latent dimension")
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--shard", default="chains", choices=["chains", "rows"],
                    help="N > 1: independent chains per GPU (weak scaling, default) or ONE chain with rows of A/W sharded "
                         "over the GPUs and NCCL all-reduces of the statistics (strong scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-nodes", type=int, default=0, help="nodes of the CPU-baseline sample (0 = auto)")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
# synthetic data: m ~ Dir(10), c ~ Cat(m), p ~ Beta(1,1), (mu,Sigma) ~ NIW(0,I,B+2,1)
# --------------------------------------------------------------------------------------------------
def pinned_empty(shape, dtype):
    """Host buffer in pinned memory when CUDA is present (the e2e leg copies from it), else plain numpy."""
    try:
        import torch
        if torch.cuda.is_available():
            t = torch.empty(tuple(shape), dtype=getattr(torch, np.dtype(dtype).name), pin_memory=True)
            return t.numpy(), t
    except Exception:
        pass
    return np.empty(shape, dtype), None


def make_sbm_data(N, K, B, seed=1234, rank=0):
    rs = np.random.RandomState(seed + 7919 * rank)
    m = rs.dirichlet(10.0 * np.ones(K))
    c = rs.choice(K, p=m, size=N)
    p = rs.beta(1.0, 1.0, size=(K, K))
    mus = np.zeros((K, K, B)); sigmas = np.zeros((K, K, B, B))
    from graphistician_b200.internals.gaussian import invwishart_draw
    for a in range(K):
        for b in range(K):
            sigmas[a, b] = invwishart_draw(np.eye(B), B + 2, rs)
            mus[a, b] = rs.multivariate_normal(np.zeros(B), sigmas[a, b])
    g = np.random.Generator(np.random.PCG64(seed + 7919 * rank))
    A, keepA = pinned_empty((N, N), np.uint8)
    W, keepW = pinned_empty((N, N, B), np.float32)
    chol = np.linalg.cholesky(sigmas)                        # K,K,B,B
    rows = 1024
    for r0 in range(0, N, rows):
        r1 = min(N, r0 + rows)
        cc = np.ix_(c[r0:r1], c)
        A[r0:r1] = g.random((r1 - r0, N), dtype=np.float32) < p[cc].astype(np.float32)
        z = g.standard_normal((r1 - r0, N, B), dtype=np.float32)
        W[r0:r1] = (np.einsum("ijab,ijb->ija", chol[cc].astype(np.float32), z) + mus[cc].astype(np.float32))
    return dict(A=A, W=W, c=c, m=m, p=p, mus=mus, sigmas=sigmas, _keep=(keepA, keepW))


def make_eigen_data(N, D, seed=1234, rank=0, p=0.01):
    """F_true ~ N(0, I), lmbda = 1, mu_0 = Phi^-1(p) (eigenmodel.py:56), A ~ Bern(logistic(Mu))."""
    from scipy.stats import norm
    g = np.random.Generator(np.random.PCG64(seed + 7919 * rank))
    F = g.standard_normal((N, D)).astype(np.float32)
    mu0 = np.float32(norm.ppf(p))
    A, keepA = pinned_empty((N, N), np.uint8)
    rows = 1024
    for r0 in range(0, N, rows):
        r1 = min(N, r0 + rows)
        eta = F[r0:r1] @ F.T + mu0
        eta[np.arange(r1 - r0), np.arange(r0, r1)] = mu0
        A[r0:r1] = g.random((r1 - r0, N), dtype=np.float32) < 1.0 / (1.0 + np.exp(-eta))
    return dict(A=A, W=None, F=F, mu0=float(mu0), _keep=(keepA,))


def make_latent_data(N, dim, B=1, seed=1234, rank=0, sigma=2.0):
    """L_true ~ sqrt(sigma) N(0, I) (adjacency.py:202), mu_0 = mu_self = 0,
    A ~ Bern(logistic(-||l_i - l_j||^2 + mu_0)), W ~ N(0, 1)."""
    g = np.random.Generator(np.random.PCG64(seed + 7919 * rank))
    L = (np.sqrt(sigma) * g.standard_normal((N, dim))).astype(np.float32)
    A, keepA = pinned_empty((N, N), np.uint8)
    W, keepW = pinned_empty((N, N, B), np.float32)
    sq = (L * L).sum(1)
    rows = 1024
    for r0 in range(0, N, rows):
        r1 = min(N, r0 + rows)
        eta = -(sq[r0:r1, None] + sq[None, :] - 2.0 * (L[r0:r1] @ L.T))
        A[r0:r1] = g.random((r1 - r0, N), dtype=np.float32) < 1.0 / (1.0 + np.exp(-eta))
        W[r0:r1] = g.standard_normal((r1 - r0, N, B), dtype=np.float32)
    return dict(A=A, W=W, L=L, _keep=(keepA, keepW))


def build_latent_model(d, N, dim, B, dtype):
    from graphistician_b200.networks import FactorizedNetworkDistribution
    from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution
    from graphistician_b200.weights import NIWGaussianWeightDistribution
    from graphistician_b200.old_abstractions import network_for
    np.random.seed(5678)
    model = FactorizedNetworkDistribution(N, LatentDistanceAdjacencyDistribution, {"dim": dim, "sigma": 2.0, "dtype": dtype},
                                          NIWGaussianWeightDistribution, {"B": B, "dtype": dtype})
    return model, network_for(d["A"], d["W"])


def build_eigen_model(d, N, D, dtype):
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    np.random.seed(5678)
    model = LogisticEigenmodel(N, D, p=0.01, sigma_F=1.0, sigma_mu0=1.0, lmbda=np.ones(D), dtype=dtype)
    return model, FixedGaussianNetwork(d["A"], None)


# --------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region
# --------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index=0, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period, self.samples, self.reasons, self.stop_flag = index, period, [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(self.period)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def dist_setup(args):
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        import torch
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend)
    return rank, world, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(x, world, device):
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def build_sbm_model(d, N, K, B, dtype):
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    np.random.seed(5678)
    model = GaussianStochasticBlockModel(N, B, C=K, dtype=dtype)
    network = FixedGaussianNetwork(d["A"], d["W"])
    return model, network


def setup_workload(args, rank):
    N, B = args.n, args.b
    if args.workload == "sbm":
        K = args.k
        d = make_sbm_data(N, K, B, rank=rank)
        model, network = build_sbm_model(d, N, K, B, args.dtype)
        s_W = 4 if args.dtype == "f32" else 8
        return dict(N=N, data=d, model=model, network=network,
                    algo_bytes=3.0 * N * N * (1 + B * s_W),       # K1 pass + K2 row pass + column pass
                    h2d=N * N * (1 + B * 4) + N * (1 + 8 * B) + N * 12 + 8 * K * K * (2 + B + B * B),
                    d2h=N * 4 + 8 * K * K * (1 + B + B * B) + 8, dominant=(0, "k_strip_stats"),
                    # ncu --set full, dram read+write per launch averaged over the 17 launches of a sweep (16 block passes
                    # at 306.15 MB, profiles/r1_sbm_full_v3.csv; the cell pass at 2441.7 MB, r1_sbm_strip_n16384_full.csv)
                    traffic=(16 * 306.15e6 + 2441.67e6) / 17 if (N, K, B, args.dtype) == (16384, 16, 2, "f32") else None,
                    cpu_baseline=cpu_baseline_sbm,
                    config={"workload": "gaussian_sbm_gibbs_sweep N=%d K=%d B=%d (configs[1] model at the metric's N)" % (N, K, B),
                            "N": N, "K": K, "B": B})
    if args.workload == "eigen":
        D = args.d
        d = make_eigen_data(N, D, rank=rank)
        model, network = build_eigen_model(d, N, D, args.dtype)
        s_O = 4 if args.dtype == "f32" else 8
        return dict(N=N, data=d, model=model, network=network,
                    algo_bytes=1.0 * N * N * (2 * s_O + 1),       # write S + read S + read A
                    h2d=N * N + N * D * 16 + 8 * D + 8, d2h=N * D * 8 + 8 * 68 + 8, dominant=(3, "k_eig_gfull+k_eig_push"),
                    # ncu --set full (profiles/r1_eig_n16384_full_v2.csv, r1_eig_full_v3.csv): gfull 1361.5 MB, 128 pushes at
                    # 17.23 MB (10.5 MB of S/Asym + the G rows read and written), 127 strip pushes at 0.18 MB; per launch
                    traffic=(1361.5e6 + 128 * 17.23e6 + 127 * 0.18e6) / 256 if (N, D, args.dtype) == (16384, 8, "f32") else None,
                    cpu_baseline=cpu_baseline_eigen,
                    config={"workload": "logistic_eigenmodel_pg_gibbs_sweep N=%d D=%d lmbda given (configs[2])" % (N, D),
                            "N": N, "D": D})
    if args.workload == "latent":
        dim, Bw = 2, 1
        d = make_latent_data(N, dim, Bw, rank=rank)
        model, network = build_latent_model(d, N, dim, Bw, args.dtype)
        tup = (d["A"], d["W"])
        return dict(N=N, data=d, model=model, network=network, step=lambda: model.resample(tup),
                    algo_bytes=22.0 * N * N + 1.0 * N * N * (1 + Bw * 4),   # 11 + 11 pair passes + NIW statistics
                    h2d=N * N * (1 + Bw * 4) + N * dim * 8 + N * (1 + 8 * Bw), d2h=N * dim * 8 + 64,
                    dominant=(4, "k_latent_pair"), cpu_baseline=cpu_baseline_latent,
                    traffic=139.6e6 if (N, args.dtype) == (16384, "f32") else None,   # profiles/r1_latent_n16384_full.csv
                    config={"workload": "latent_distance_hmc_gibbs_sweep + NIW weights N=%d dim=%d B=%d (configs[3])" % (N, dim, Bw),
                            "N": N, "dim": dim, "B": Bw})
    raise SystemExit("workload %s: not built yet" % args.workload)


def run_ours(args):
    import torch
    from graphistician_b200 import _lib
    rank, world, local = dist_setup(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this engine has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _lib.set_default_device(local)
    ctx = _lib.default_context()
    stream = torch.cuda.Stream(device=dev)
    ctx.set_stream(stream.cuda_stream)
    rows = args.shard == "rows" and world > 1
    wl = setup_workload(args, 0 if rows else rank)
    if rows:                                                   # one chain, rows of A / W sharded over the ranks
        from graphistician_b200.sharding import Comm
        torch.cuda.set_stream(stream)                          # libgk and NCCL share this stream
        wl["network"].shard(Comm())
    N, model, network, d, algo_bytes = wl["N"], wl["model"], wl["network"], wl["data"], wl["algo_bytes"]
    step = wl.get("step") or (lambda: model.resample(network))

    def timed(nsteps, e2e):
        barrier(world); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(nsteps):
                if e2e:
                    network.invalidate()                        # host buffers -> device inside the timed region
                step()
            e1.record(stream)
        torch.cuda.synchronize(); barrier(world)
        return max_over_ranks(e0.elapsed_time(e1), world, dev)

    timed(args.warmup, False)
    sampler = ClockSampler(local); sampler.start()
    l0 = ctx.launch_count()
    ms = timed(args.steps, False)
    launches = ctx.launch_count() - l0
    clocks = sampler.result()
    # e2e: same call, network re-uploaded from pinned host memory every step, state read back (c, m, p) every step
    e2e_steps = max(3, min(args.steps, 5))
    timed(1, True)
    ms_e2e = timed(e2e_steps, True)
    h2d, d2h = wl["h2d"], wl["d2h"]
    # per-kernel timing pass (CUDA events around every launch of each kernel family, on the ctx stream)
    ctx.kernel_timing(True)
    for _ in range(3):
        step()
    fam = {f: ctx.kernel_timing_read(f) for f in range(6)}
    ctx.kernel_timing(False)
    kms, kn, kbytes = fam[wl["dominant"][0]]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = kbytes / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
    names = {0: "k_strip_stats", 1: "k_sbm_seq", 2: "k_eig_pg_tile", 3: "k_eig_gfull+k_eig_push", 4: "k_latent_pair",
             5: "k_eig_seq4"}
    out = {
        "metric": METRIC, "value": (1 if rows else world) * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if rows else "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": dict(wl["config"], parallelism=("one chain, rows sharded x%d" if rows else "independent chains x%d") % world,
                       l2="inputs (%.2f GB/sweep) exceed the 126 MB L2" % (algo_bytes / 1e9)),
        "gpu_launches": int(launches), "clocks": clocks,
        "e2e": {"value": (1 if rows else world) * e2e_steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h)},
        "roofline": {"bound": "hbm", "kernel": wl["dominant"][1], "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": wl.get("traffic"),
                     "algorithmic_bytes_per_launch": (kbytes / kn) if kn else None,
                     "peak_source": "measured" if peaks else "fallback",
                     "launches_per_sweep": kn / 3.0, "kernel_ms_per_sweep": kms / 3.0,
                     "other_kernels_ms_per_sweep": {names[f]: fam[f][0] / 3.0 for f in fam if fam[f][1] and f != wl["dominant"][0]},
                     "sweep_algorithmic_GBps": algo_bytes / (ms / args.steps * 1e-3) / 1e9},
    }
    if args.workload == "sbm":
        out["n_changed_last_sweep"] = int(getattr(model, "_n_changed", -1))
    if rank == 0 and not args.no_cpu_baseline:
        out["cpu_baseline"] = wl["cpu_baseline"](d, model, args)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle (vectorised T1 restatement of the reference's NumPy path), timed on the
# host cores.  The reference itself cannot run (Python 2, missing modules and packages: SURVEY.md 8c), so kind="port".
# --------------------------------------------------------------------------------------------------
def cpu_baseline_sbm(d, model, args, nodes=None):
    N = d["A"].shape[0]
    nodes = nodes or args.cpu_nodes or min(N, 2048)
    state = dict(c=np.asarray(model.c), m=np.asarray(model.m), p=np.clip(model.p, 1e-12, 1 - 1e-12),
                 mus=model.mus, sigmas=model.sigmas)
    from oracle import sbm
    # block_pair_stats on a row subset needs square indexing: run it on the leading `nodes` rows only
    A, W = d["A"], d["W"]
    t0 = time.perf_counter()
    _stats_rows(A, W, state["c"], len(state["m"]), nodes)
    t_stats = time.perf_counter() - t0
    gc = sbm.gaussian_coefficients(state["mus"], state["sigmas"])
    logm = np.log(state["m"]); c = np.array(state["c"], dtype=np.int64).copy()
    u = np.random.RandomState(1).rand(N)
    t0 = time.perf_counter()
    At = np.ascontiguousarray(A[:, :nodes].T); Wt = np.ascontiguousarray(W[:, :nodes].transpose(1, 0, 2))
    for n1 in range(nodes):
        lp = sbm.node_conditional(n1, A, W, c, logm, state["p"], gc, sbm.SELF_OLD, 1.0, At, Wt)
        c[n1] = sbm.sample_discrete_from_log(lp, u[n1])
    t_sweep = time.perf_counter() - t0
    secs = (t_stats + t_sweep) * (N / float(nodes))
    return {"value": 1.0 / secs, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "oracle (vectorised NumPy restatement) on the first %d of %d nodes: K1 statistics over those rows "
                      "+ their sequential conditionals/draws; %.1f s measured, scaled x%.1f to a full sweep"
                      % (nodes, N, t_stats + t_sweep, N / float(nodes)),
            "measured_s": t_stats + t_sweep}


def cpu_baseline_eigen(d, model, args, nodes=None):
    """Oracle (port) of one logistic-eigenmodel sweep on a bounded sample: the C Devroye sampler (OpenMP, all
    cores) for both Omega entries of the first `nodes` rows, then the per-node conditionals/draws of those nodes
    (vectorised NumPy restatement of eigenmodel.py:876-908); scaled by N/nodes."""
    from oracle import eigenmodel as eg, pg
    A = d["A"]
    N = A.shape[0]
    nodes = nodes or args.cpu_nodes or min(N, 512)
    F = np.asarray(model.F, dtype=np.float64); lm = np.asarray(model.lmbda, dtype=np.float64); mu0 = float(model.mu_0)
    t0 = time.perf_counter()
    O1 = pg.pgdraw_eigen(F, lm, mu0, seed=1, r0=0, r1=nodes)          # omega[n, :]
    O2 = pg.pgdraw_eigen(F, lm, mu0, seed=2, r0=0, r1=nodes)          # omega[:, n]^T (Mu is symmetric)
    t_pg = time.perf_counter() - t0
    t0 = time.perf_counter()
    S = O1 + O2
    Asym = A[:nodes].astype(np.float64) + A[:, :nodes].T
    R = Asym - 1.0 - mu0 * S
    z = np.random.RandomState(1).randn(N, F.shape[1])
    eg.resample_F(S, R, F, lm, model.sigma_F, z, nodes=nodes)
    t_F = time.perf_counter() - t0
    secs = (t_pg + t_F) * (N / float(nodes))
    return {"value": 1.0 / secs, "unit": UNIT, "cores": pg.num_threads(), "kind": "port",
            "sample": "oracle on the first %d of %d nodes: C Polya-gamma sampler (OpenMP x%d) for their 2N Omega entries "
                      "(%.2f s) + sequential conditionals/draws in NumPy (%.2f s, 1 core); scaled x%.1f to a full sweep"
                      % (nodes, N, pg.num_threads(), t_pg, t_F, N / float(nodes)),
            "measured_s": t_pg + t_F}


def cpu_baseline_latent(d, model, args, nodes=None):
    """Oracle (port) of one latent-distance sweep on a bounded sample: the fused value + gradient of
    adjacency.py:244-275 (vectorised NumPy, fp64) over the first `nodes` rows, for the 22 evaluations a sweep needs
    (11 for L, 11 for mu_0), plus the K = 1 NIW statistics of those rows; scaled by N / nodes."""
    from oracle import latent
    A, W = d["A"], d["W"]
    N = A.shape[0]
    nodes = nodes or args.cpu_nodes or min(N, 256)
    adj = model.adjacency
    L = np.asarray(adj.L, dtype=np.float64)
    Ar = A[:nodes].astype(np.float64); Atr = np.ascontiguousarray(A[:, :nodes].T).astype(np.float64)
    t0 = time.perf_counter()
    for _ in range(22):
        latent.rows_value_and_grad(L, adj.mu_0, adj.mu_self, Ar, Atr, 0, adj.sigma)
    e = A[:nodes] != 0
    We = W[:nodes][e].astype(np.float64)
    We.sum(0); We.T.dot(We)
    secs = time.perf_counter() - t0
    full = secs * (N / float(nodes))
    return {"value": 1.0 / full, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "oracle (vectorised NumPy restatement, fp64) on the first %d of %d rows: 22 value+gradient "
                      "evaluations + NIW statistics; %.1f s measured, scaled x%.1f to a full sweep" % (nodes, N, secs, N / float(nodes)),
            "measured_s": secs}


def _stats_rows(A, W, c, K, nodes):
    """K1 of the oracle (bincount over the fused cell index) restricted to the leading rows."""
    from oracle import sbm
    N = A.shape[0]
    B = W.shape[2]
    c = np.asarray(c, dtype=np.int64)
    for r0 in range(0, nodes, 256):
        r1 = min(nodes, r0 + 256)
        Ar = (A[r0:r1] != 0)
        cell = (c[r0:r1, None] * K + c[None, :])
        Ar[np.arange(r1 - r0), np.arange(r0, r1)] = False
        ce = cell[Ar]
        np.bincount(ce, minlength=K * K)
        We = W[r0:r1][Ar].astype(np.float64)
        for b in range(B):
            np.bincount(ce, weights=We[:, b], minlength=K * K)
            for b2 in range(b, B):
                np.bincount(ce, weights=We[:, b] * We[:, b2], minlength=K * K)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N, K, B = args.n, args.k, args.b
    d = make_sbm_data(N, K, B)

    class M:                                           # same initial state as our arm's model (prior draw, seed 5678)
        pass
    np.random.seed(5678)
    rs = np.random
    m = rs.dirichlet(np.ones(K)); c = rs.choice(K, p=m, size=N); p = rs.beta(1.0, 1.0, size=(K, K))
    model = M(); model.c = c; model.m = m; model.p = p
    model.mus = d["mus"]; model.sigmas = d["sigmas"]
    nodes = args.cpu_nodes or min(N, 1024)
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline_sbm(d, model, args, nodes)
        if i >= args.warmup:
            vals.append(r)
        if sum(v["measured_s"] for v in vals) > 150:
            break
    v = float(np.mean([x["value"] for x in vals]))
    out = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
           "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": "gaussian_sbm_gibbs_sweep N=%d K=%d B=%d (configs[1] model at the metric's N)" % (N, K, B),
                      "N": N, "K": K, "B": B},
           "cpu_baseline": dict(vals[-1], value=v),
           "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
