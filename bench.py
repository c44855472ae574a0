#!/usr/bin/env python
"""bench.py -- Gibbs sweeps/s of the B200 engine on BASELINE.json's metric, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload all|sbm|eigen|latent|ais] [--shard chains|rows]

A "step" is one full ``model.resample(network)`` Gibbs sweep on synthetic data (seed 1234, generators of
SURVEY.md 8d).  The metric names two models -- "Gibbs sweeps/sec at N=16k (SBM, logistic eigenmodel)" -- so the
default run (``--workload all``) times the Gaussian SBM (configs[1] model: K=16, B=2) at N=16384 as the top-level
record and the logistic eigenmodel (configs[2]) and the latent-distance model (configs[3]) at N=16384 under
``workloads``.  One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import gc
import json
import os
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "gibbs_sweeps_per_sec"
UNIT = "sweeps/s"
FAMILY = {0: "k_strip_stats", 1: "k_sbm_seq", 2: "k_eig_pg_tile", 3: "k_eig_gfull+k_eig_push", 4: "k_latent_pair",
          5: "k_eig_seq", 6: "k_strip_soft", 7: "k_mf_seq"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="all", choices=["all", "sbm", "eigen", "latent", "ais"])
    ap.add_argument("--n", type=int, default=16384, help="nodes (metric: N=16k)")
    ap.add_argument("--k", type=int, default=16, help="SBM blocks")
    ap.add_argument("--b", type=int, default=2, help="weight dimension")
    ap.add_argument("--d", type=int, default=8, help="eigenmodel latent dimension")
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--shard", default="chains", choices=["chains", "rows"],
                    help="N > 1: independent chains per GPU (weak scaling, default) or ONE chain with rows of A/W sharded "
                         "over the GPUs and NCCL all-reduces of the statistics (strong scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the row-sharded (strong scaling) sub-record")
    ap.add_argument("--no-extras", action="store_true", help="skip the mixing / f64 side records of the SBM workload")
    ap.add_argument("--cpu-nodes", type=int, default=0, help="nodes of the CPU-baseline sample (0 = auto)")
    ap.add_argument("--ref-budget", type=float, default=70.0, help="reference arm: seconds of timed sweeps per workload")
    ap.add_argument("--ais-particles", type=int, default=128, help="co-resident AIS particles per GPU")
    ap.add_argument("--ais-n", type=int, default=65536, help="nodes of the AIS workload (configs[4]: N=65536)")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
# host description (the CPU numbers are quoted with the cores they ran on)
# --------------------------------------------------------------------------------------------------
def host_info():
    info = {"cpu_count": os.cpu_count(), "affinity": len(os.sched_getaffinity(0)),
            "OMP_NUM_THREADS": os.environ.get("OMP_NUM_THREADS")}
    try:
        from threadpoolctl import threadpool_info
        info["threadpools"] = [{"api": t.get("user_api"), "lib": t.get("internal_api"), "threads": t.get("num_threads")}
                               for t in threadpool_info()]
    except Exception:
        info["threadpools"] = None
    return info


def usable_cores():
    return max(1, len(os.sched_getaffinity(0)))


# --------------------------------------------------------------------------------------------------
# synthetic data.  Row chunks are filled by a thread pool, each chunk from its own PCG64 stream
# keyed by (seed, rank, chunk), so the data do not depend on the number of threads.
# --------------------------------------------------------------------------------------------------
def pinned_empty(shape, dtype, pin=True):
    """Host buffer in pinned memory when CUDA is present (the e2e leg copies from it), else plain numpy."""
    if pin:
        try:
            import torch
            if torch.cuda.is_available():
                t = torch.empty(tuple(shape), dtype=getattr(torch, np.dtype(dtype).name), pin_memory=True)
                return t.numpy(), t
        except Exception:
            pass
    return np.empty(shape, dtype), None


def _fill_rows(N, fill, rows=512):
    chunks = [(i, r0, min(N, r0 + rows)) for i, r0 in enumerate(range(0, N, rows))]
    with ThreadPoolExecutor(max_workers=min(32, usable_cores())) as ex:
        list(ex.map(lambda t: fill(*t), chunks))


def _chunk_rng(seed, rank, chunk):
    return np.random.Generator(np.random.PCG64(np.random.SeedSequence([seed, rank, chunk])))


def sbm_truth(N, K, B, seed=1234, rank=0, signal=1.0):
    """m ~ Dir(10), c ~ Cat(m), p ~ Beta(1,1), (mu,Sigma) ~ NIW(0,I,B+2,1).  ``signal`` < 1 shrinks the block
    structure (p towards its mean, mu towards 0, Sigma towards I): the weak-signal dataset of the mixing record."""
    from graphistician_b200.internals.gaussian import invwishart_draw
    rs = np.random.RandomState(seed + 7919 * rank)
    m = rs.dirichlet(10.0 * np.ones(K))
    c = rs.choice(K, p=m, size=N)
    p = rs.beta(1.0, 1.0, size=(K, K))
    mus = np.zeros((K, K, B)); sigmas = np.zeros((K, K, B, B))
    for a in range(K):
        for b in range(K):
            sigmas[a, b] = invwishart_draw(np.eye(B), B + 2, rs)
            mus[a, b] = rs.multivariate_normal(np.zeros(B), sigmas[a, b])
    if signal != 1.0:
        p = p.mean() + signal * (p - p.mean())
        mus = signal * mus
        sigmas = np.eye(B) + signal * (sigmas - np.eye(B))
    return dict(c=c, m=m, p=p, mus=mus, sigmas=sigmas)


def fill_sbm_data(A, W, truth, seed=1234, rank=0):
    N = A.shape[0]
    K, B = truth["p"].shape[0], truth["mus"].shape[2]
    c = truth["c"]
    p32 = truth["p"].astype(np.float32).ravel()
    chol = np.linalg.cholesky(truth["sigmas"]).astype(np.float32).reshape(K * K, B, B)
    mu32 = truth["mus"].astype(np.float32).reshape(K * K, B)

    def fill(i, r0, r1):
        g = _chunk_rng(seed, rank, i)
        cell = (c[r0:r1, None] * K + c[None, :]).astype(np.int32)
        A[r0:r1] = g.random((r1 - r0, N), dtype=np.float32) < p32[cell]
        z = g.standard_normal((r1 - r0, N, B), dtype=np.float32)
        for a in range(B):
            w = mu32[:, a][cell]
            for b in range(a + 1):
                w += chol[:, a, b][cell] * z[:, :, b]
            W[r0:r1, :, a] = w
    _fill_rows(N, fill)


def make_sbm_data(N, K, B, seed=1234, rank=0, pin=True, signal=1.0):
    truth = sbm_truth(N, K, B, seed, rank, signal)
    A, keepA = pinned_empty((N, N), np.uint8, pin)
    W, keepW = pinned_empty((N, N, B), np.float32, pin)
    fill_sbm_data(A, W, truth, seed, rank)
    return dict(truth, A=A, W=W, _keep=(keepA, keepW))


def make_eigen_data(N, D, seed=1234, rank=0, p=0.01, pin=True):
    """F_true ~ N(0, I), lmbda = 1, mu_0 = Phi^-1(p) (eigenmodel.py:56), A ~ Bern(logistic(Mu))."""
    from scipy.stats import norm
    g0 = np.random.Generator(np.random.PCG64(seed + 7919 * rank))
    F = g0.standard_normal((N, D)).astype(np.float32)
    mu0 = np.float32(norm.ppf(p))
    A, keepA = pinned_empty((N, N), np.uint8, pin)

    def fill(i, r0, r1):
        g = _chunk_rng(seed, rank, i)
        eta = F[r0:r1] @ F.T + mu0
        eta[np.arange(r1 - r0), np.arange(r0, r1)] = mu0
        with np.errstate(over="ignore"):           # eta <= 0: exp(-eta) may reach inf, and 1 / (1 + inf) = 0 is the value wanted
            A[r0:r1] = g.random((r1 - r0, N), dtype=np.float32) < 1.0 / (1.0 + np.exp(-eta))
    _fill_rows(N, fill)
    return dict(A=A, W=None, F=F, mu0=float(mu0), _keep=(keepA,))


def make_latent_data(N, dim, B=1, seed=1234, rank=0, sigma=2.0, pin=True):
    """L_true ~ sqrt(sigma) N(0, I) (adjacency.py:202), mu_0 = mu_self = 0,
    A ~ Bern(logistic(-||l_i - l_j||^2 + mu_0)), W ~ N(0, 1)."""
    g0 = np.random.Generator(np.random.PCG64(seed + 7919 * rank))
    L = (np.sqrt(sigma) * g0.standard_normal((N, dim))).astype(np.float32)
    A, keepA = pinned_empty((N, N), np.uint8, pin)
    W, keepW = pinned_empty((N, N, B), np.float32, pin)
    sq = (L * L).sum(1)

    def fill(i, r0, r1):
        g = _chunk_rng(seed, rank, i)
        eta = -(sq[r0:r1, None] + sq[None, :] - 2.0 * (L[r0:r1] @ L.T))
        with np.errstate(over="ignore"):           # eta <= 0: exp(-eta) may reach inf, and 1 / (1 + inf) = 0 is the value wanted
            A[r0:r1] = g.random((r1 - r0, N), dtype=np.float32) < 1.0 / (1.0 + np.exp(-eta))
        W[r0:r1] = g.standard_normal((r1 - r0, N, B), dtype=np.float32)
    _fill_rows(N, fill)
    return dict(A=A, W=W, L=L, _keep=(keepA, keepW))


# --------------------------------------------------------------------------------------------------
# workload descriptions shared by both arms: `config` is what the driver compares between the arms
# --------------------------------------------------------------------------------------------------
def workload_config(name, args):
    N = args.n
    if name == "sbm":
        return {"workload": "gaussian_sbm_gibbs_sweep N=%d K=%d B=%d (configs[1] model at the metric's N)" % (N, args.k, args.b),
                "N": N, "K": args.k, "B": args.b}
    if name == "eigen":
        return {"workload": "logistic_eigenmodel_pg_gibbs_sweep N=%d D=%d lmbda given (configs[2])" % (N, args.d),
                "N": N, "D": args.d}
    if name == "latent":
        return {"workload": "latent_distance_hmc_gibbs_sweep + NIW weights N=%d dim=2 B=1 (configs[3])" % N,
                "N": N, "dim": 2, "B": 1}
    if name == "ais":
        return {"workload": "ais_sbm_particle_sweep N=%d R=3 P=%d per GPU (configs[4])" % (args.ais_n, args.ais_particles),
                "N": args.ais_n, "R": 3, "P": args.ais_particles}
    raise SystemExit("unknown workload " + name)


def algorithmic_bytes(name, args, dtype):
    """bytes one sweep must move (s_A = 1; s_W = s_Omega = 4, or 8 in fp64 parity mode)."""
    N = float(args.n)
    s = 4 if dtype == "f32" else 8
    if name == "sbm":
        return 3.0 * N * N * (1 + args.b * s)              # K1 pass + K2 row pass + column pass
    if name == "eigen":
        return N * N * (2 * s + 1)                         # write S + read S + read A
    if name == "latent":
        return 22.0 * N * N + N * N * (1 + 1 * s)          # 11 + 11 pair passes + NIW statistics
    raise SystemExit(name)


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
    return model, FixedGaussianNetwork(d["A"], d["W"])


def build_eigen_model(d, N, D, dtype):
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    np.random.seed(5678)
    model = LogisticEigenmodel(N, D, p=0.01, sigma_F=1.0, sigma_mu0=1.0, lmbda=np.ones(D), dtype=dtype)
    return model, FixedGaussianNetwork(d["A"], None)


def build_latent_model(d, N, dim, B, dtype):
    from graphistician_b200.networks import FactorizedNetworkDistribution
    from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution
    from graphistician_b200.weights import NIWGaussianWeightDistribution
    from graphistician_b200.old_abstractions import network_for
    np.random.seed(5678)
    model = FactorizedNetworkDistribution(N, LatentDistanceAdjacencyDistribution, {"dim": dim, "sigma": 2.0, "dtype": dtype},
                                          NIWGaussianWeightDistribution, {"B": B, "dtype": dtype})
    return model, network_for(d["A"], d["W"])


def ncu_traffic(kernel):
    """dram read + write bytes per launch of `kernel` from the committed ncu --set full summaries
    (profiles/traffic.json, written by profiles/summarize.py); None when no capture of this kernel is on file."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = t.get(kernel)
        return (ent["dram_bytes_per_launch"], ent["source"]) if ent else (None, None)
    except Exception:
        return None, None


def setup_workload(name, args, rank, dtype):
    N, B = args.n, args.b
    s = 4 if dtype == "f32" else 8
    if name == "sbm":
        K = args.k
        d = make_sbm_data(N, K, B, rank=rank)
        model, network = build_sbm_model(d, N, K, B, dtype)
        return dict(N=N, data=d, model=model, network=network,
                    h2d=N * N * (1 + B * 4) + N * (1 + 8 * B) + N * 12 + 8 * K * K * (2 + B + B * B),
                    d2h=N * 4 + 8 * K * K * (1 + B + B * B) + 8, dominant=0, cpu_baseline=cpu_baseline_sbm)
    if name == "eigen":
        D = args.d
        d = make_eigen_data(N, D, rank=rank)
        model, network = build_eigen_model(d, N, D, dtype)
        return dict(N=N, data=d, model=model, network=network,
                    h2d=N * N + N * D * 16 + 8 * D + 8, d2h=N * D * 8 + 8 * 68 + 8, dominant=3,
                    cpu_baseline=cpu_baseline_eigen)
    if name == "latent":
        dim, Bw = 2, 1
        d = make_latent_data(N, dim, Bw, rank=rank)
        model, network = build_latent_model(d, N, dim, Bw, dtype)
        tup = (d["A"], d["W"])
        return dict(N=N, data=d, model=model, network=network, step=lambda: model.resample(tup),
                    h2d=N * N * (1 + Bw * 4) + N * dim * 8 + N * (1 + 8 * Bw), d2h=N * dim * 8 + 64,
                    dominant=4, cpu_baseline=cpu_baseline_latent)
    raise SystemExit("workload %s: not built" % name)


class Runner:
    """CUDA-event timing of `nsteps` sweeps on the ctx stream: barrier + synchronize on both sides, max over ranks."""

    def __init__(self, args):
        import torch
        from graphistician_b200 import _lib
        self.torch = torch
        self.rank, self.world, self.local = dist_setup(args)
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (this engine has no CPU fallback)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        _lib.set_default_device(self.local)
        self.ctx = _lib.default_context()
        self.stream = torch.cuda.Stream(device=self.dev)
        self.ctx.set_stream(self.stream.cuda_stream)

    def timed(self, step, nsteps, before=None):
        torch = self.torch
        barrier(self.world); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self.stream):
            e0.record(self.stream)
            for _ in range(nsteps):
                if before is not None:
                    before()
                step()
            e1.record(self.stream)
        torch.cuda.synchronize(); barrier(self.world)
        return max_over_ranks(e0.elapsed_time(e1), self.world, self.dev)


def bench_workload(name, args, R, dtype, steps, warmup, rows=False, cpu=True, first_sweep=False, e2e=True):
    """One workload record: value (resident), e2e (host buffers), roofline of the dominant kernel family,
    launches, clocks, and (rank 0, one GPU) the CPU baseline."""
    wl = setup_workload(name, args, 0, dtype)      # every rank: the same data and chain seed, so the per-GPU work is identical
    model, network, d = wl["model"], wl["network"], wl["data"]
    if rows:                                                   # one chain, rows of A / W sharded over the ranks
        from graphistician_b200.sharding import Comm
        R.torch.cuda.set_stream(R.stream)                      # libgk and NCCL share this stream
        network.shard(Comm())
    step = wl.get("step") or (lambda: model.resample(network))
    algo_bytes = algorithmic_bytes(name, args, dtype)
    rec = {}
    if first_sweep:                                            # sweep 1 from the prior draw of the labels (data already resident)
        network.device(dtype)
        rec["first_sweep_ms"] = R.timed(step, 1)
        rec["first_sweep_n_changed"] = int(getattr(model, "_n_changed", -1))
        warmup = max(0, warmup - 1)
    R.timed(step, warmup)
    extra_warm = 0
    if name == "sbm" and not rows and R.world > 1:
        # the ranks' chains run on different data and may take different numbers of sweeps to recover the planted blocks; the
        # timed region is a max over ranks, so every rank is brought to the stationary regime (n_changed = 0) first
        while extra_warm < 12 and max_over_ranks(float(getattr(model, "_n_changed", 0)), R.world, R.dev) > 0:
            R.timed(step, 1)
            extra_warm += 1
    sampler = ClockSampler(R.local); sampler.start()
    l0 = R.ctx.launch_count()
    ms = R.timed(step, steps)
    launches = R.ctx.launch_count() - l0
    clocks = sampler.result()
    # e2e: same call, network re-uploaded from pinned host memory every step, state read back every step
    e2e_steps = max(3, min(steps, 5))
    ms_e2e = float("nan")
    if e2e:
        R.timed(step, 1, before=network.invalidate)
        ms_e2e = R.timed(step, e2e_steps, before=network.invalidate)
    # per-kernel timing pass (CUDA events around every launch of each kernel family, on the ctx stream)
    R.ctx.kernel_timing(True)
    for _ in range(3):
        step()
    fam = {f: R.ctx.kernel_timing_read(f) for f in range(8)}
    R.ctx.kernel_timing(False)
    kms, kn, kbytes = fam[wl["dominant"]]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = kbytes / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
    mult = 1 if rows else R.world
    kname = FAMILY[wl["dominant"]]
    traffic, tsrc = ncu_traffic(kname)
    rec.update({
        "value": mult * steps / (ms * 1e-3), "unit": UNIT, "steps": steps, "ms_per_step": ms / steps, "dtype": dtype,
        "config": workload_config(name, args), "gpu_launches": int(launches), "clocks": clocks,
        "e2e": ({"value": mult * e2e_steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(wl["h2d"]),
                 "d2h_bytes_per_step": int(wl["d2h"])} if e2e else None),
        "roofline": {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": tsrc,
                     "algorithmic_bytes_per_launch": (kbytes / kn) if kn else None,
                     "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                     "launches_per_sweep": kn / 3.0, "kernel_ms_per_sweep": kms / 3.0,
                     "other_kernels_ms_per_sweep": {FAMILY[f]: fam[f][0] / 3.0 for f in fam if fam[f][1] and f != wl["dominant"]},
                     "sweep_algorithmic_bytes": algo_bytes,
                     "sweep_algorithmic_GBps": algo_bytes / (ms / steps * 1e-3) / 1e9,
                     "sweep_frac_of_peak": algo_bytes / (ms / steps * 1e-3) / 1e9 / peak},
        "l2": "inputs (%.2f GB/sweep) exceed the 126 MB L2" % (algo_bytes / 1e9),
    })
    if name == "latent":
        # the pair kernel is bound by the XU (MUFU) pipe, not by HBM: ex2 + rcp per unordered pair, 16 MUFU per clock and SM
        sm_mhz = clocks.get("sm_mhz") or 1965.0
        pairs = 22.0 * args.n * (args.n - 1) / 2.0 / mult
        n_sm = R.torch.cuda.get_device_properties(R.local).multi_processor_count
        bound_ms = pairs * 2.0 / (16.0 * n_sm * sm_mhz * 1e6) * 1e3
        rec["roofline"]["xu_pipe"] = {"mufu_per_pair": 2, "pair_evaluations_per_sweep": pairs, "bound_ms_per_sweep": bound_ms,
                                      "frac": bound_ms / (kms / 3.0) if kms > 0 else None,
                                      "note": "the HBM figure above is SURVEY.md 8(d)'s accounting (22 N^2 bytes); the kernel reads A + A^T "
                                              "once per unordered pair and is bound by the two transcendental operations per pair"}
    if name == "sbm":
        rec["n_changed_last_sweep"] = int(max_over_ranks(float(getattr(model, "_n_changed", -1)), R.world, R.dev))
        rec["extra_warmup_sweeps"] = extra_warm
        rec["regime"] = "stationary (planted blocks recovered during warm-up; see `mixing` for sweeps with movers)"
    if cpu and R.rank == 0 and R.world == 1 and not args.no_cpu_baseline:
        rec["cpu_baseline"] = wl["cpu_baseline"](d, model, args)
    return rec, wl


def sbm_mixing_record(args, R, wl, dtype, signal):
    """Sweeps WITH movers: the planted data are overwritten in place by a draw of the same model whose block structure
    is shrunk to `signal` (0.05: p within +-0.05 of its mean -- the labels settle slowly; 0: no structure at all --
    the labels never settle and almost every node moves in every sweep, the worst case of the serial kernel)."""
    d, network, N = wl["data"], wl["network"], args.n
    truth = sbm_truth(N, args.k, args.b, seed=4321, rank=R.rank, signal=signal)
    fill_sbm_data(d["A"], d["W"], truth, seed=4321, rank=R.rank)
    network.invalidate()
    np.random.seed(8765)
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    model = GaussianStochasticBlockModel(N, args.b, C=args.k, dtype=dtype)
    nch = []

    def step():
        model.resample(network)
        nch.append(int(model._n_changed))
    R.timed(step, 2)
    del nch[:]
    steps = 5
    ms = R.timed(step, steps)
    return {"dataset": "p = mean + s (p - mean), mu x s, Sigma = I + s (Sigma - I) with s = %g, seed 4321" % signal,
            "ms_per_step": ms / steps, "value": R.world * steps / (ms * 1e-3), "unit": UNIT, "steps": steps,
            "n_changed_per_sweep": nch, "us_per_move": 1e3 * (ms / steps) / max(1.0, float(np.mean(nch)))}


def release(wl):
    from graphistician_b200.old_abstractions import drop_cached_networks
    try:
        wl["network"]._drop()
    except Exception:
        pass
    drop_cached_networks()
    wl.clear()
    gc.collect()


def run_ours(args):
    R = Runner(args)
    rows = args.shard == "rows" and R.world > 1
    names = ["sbm", "eigen", "latent"] if args.workload == "all" else [args.workload]
    if names == ["ais"]:
        return run_ais(args, R)
    primary = names[0]
    extras = args.workload in ("all", "sbm") and not args.no_extras and not rows
    rec, wl = bench_workload(primary, args, R, args.dtype, args.steps, args.warmup, rows=rows,
                             first_sweep=(primary == "sbm" and extras))
    out = {"metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": R.world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
           "scaling": "strong" if rows else "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
           "config": rec["config"],
           "parallelism": ("one chain, rows sharded x%d" if rows else "replica chains x%d (same data and seed on every GPU: identical work per GPU, no data-path collective)") % R.world}
    for k in ("l2", "gpu_launches", "clocks", "e2e", "roofline", "n_changed_last_sweep", "extra_warmup_sweeps", "regime", "cpu_baseline"):
        if k in rec:
            out[k] = rec[k]
    if primary == "sbm" and extras:
        mixing = {"first_sweep_from_prior_labels_ms": rec.get("first_sweep_ms"),
                  "first_sweep_n_changed": rec.get("first_sweep_n_changed")}
        mixing["weak_signal"] = sbm_mixing_record(args, R, wl, args.dtype, 0.05)
        mixing["no_signal"] = sbm_mixing_record(args, R, wl, args.dtype, 0.0)
        out["mixing"] = mixing
    release(wl)
    out["workloads"] = {}
    for name in names[1:]:
        r, w = bench_workload(name, args, R, args.dtype, args.steps, args.warmup, rows=rows)
        out["workloads"][name] = r
        release(w)
    if R.world > 1 and not rows and args.workload == "all" and not args.no_strong:
        # ONE chain with the rows of A / W sharded over the ranks (north_star configs 3 / 4) beside the replicas:
        # statistics all-reduced with NCCL in-stream, the sequential draws on the owner of each node block
        out["strong"] = {"scaling": "strong", "parallelism": "one chain, rows sharded x%d" % R.world, "workloads": {}}
        R.torch.cuda.set_stream(R.stream)
        for name in ("latent", "eigen", "sbm"):
            r, w = bench_workload(name, args, R, args.dtype, max(3, args.steps // 2), args.warmup, rows=True, cpu=False, e2e=False)
            out["strong"]["workloads"][name] = {k: r[k] for k in ("value", "unit", "steps", "ms_per_step", "config", "gpu_launches", "roofline")}
            release(w)
    if args.workload == "all" and not rows:                        # config 5 beside the sweeps: batched AIS particle sweeps
        out["workloads"]["ais"] = bench_ais(args, R, max(3, args.steps // 2), args.warmup)
    if primary == "sbm" and extras and args.dtype == "f32":       # the fp64-storage line beside the fp32 one
        r, w = bench_workload("sbm", args, R, "f64", max(3, args.steps // 4), args.warmup, cpu=False)
        out["workloads"]["sbm_f64"] = r
        release(w)
    out["host"] = host_info()
    if R.rank == 0:
        print(json.dumps(out), flush=True)
    if R.world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


AIS_UNIT = "particle-sweeps/s"


def ais_truth(N, R=3, seed=1234):
    """config 5: SBM truth R = 3, alpha0 = 10, b0 = b1 = 1 (examples/demo_ais.py:76-80), seed 1234."""
    rng = np.random.RandomState(seed)
    pi = rng.dirichlet(10 * np.ones(R))
    z = rng.choice(R, size=N, p=pi)
    Bt = rng.beta(1.0, 1.0, (R, R))
    return z, Bt, pi


def _ais_cpu_nodes(N, R, z_true, Bt, z, Bm, pi, beta, lo, hi, seed=3):
    """naive_sample_f (internals/graph_models.py:261-297) for nodes [lo, hi): the count form of oracle.sbm.node_conditional
    (SELF_AIS) on rows / columns of A generated for those nodes only (a 4.3 GB host matrix is never built)."""
    from oracle import sbm
    rng = np.random.RandomState(seed + lo)
    logB, log1mB = np.log(Bm), np.log1p(-Bm)
    z = np.array(z, dtype=np.int64)
    for n in range(lo, hi):
        a_out = rng.rand(N) < Bt[z_true[n], z_true]
        a_in = rng.rand(N) < Bt[z_true, z_true[n]]
        a11 = bool(a_out[n]); a_in[n] = a11
        others = np.ones(N, bool); others[n] = False
        sz = np.bincount(z[others], minlength=R).astype(np.float64)
        co = np.bincount(z[a_out & others], minlength=R).astype(np.float64)
        ci = np.bincount(z[a_in & others], minlength=R).astype(np.float64)
        lp = np.log(pi) + beta * (logB.dot(co) + log1mB.dot(sz - co) + logB.T.dot(ci) + log1mB.T.dot(sz - ci))
        k = np.arange(R)
        lp += 2.0 * beta * (logB[k, k] if a11 else log1mB[k, k])
        z[n] = sbm.log_sum_exp_sample(lp, rng.rand())
    return hi - lo


def cpu_baseline_ais(args, N, R, nodes=None):
    nodes = min(N, nodes or args.cpu_nodes or 4096)
    z_true, Bt, pi = ais_truth(N, R)
    z = np.random.RandomState(1).randint(0, R, N)
    t0 = time.perf_counter()
    _ais_cpu_nodes(N, R, z_true, Bt, z, np.clip(Bt, 1e-6, 1 - 1e-6), pi, 0.5, 0, nodes)
    gen0 = time.perf_counter()
    rng = np.random.RandomState(0)
    for n in range(nodes):                                    # the data generation of the sample, timed alone and subtracted
        rng.rand(N) < Bt[z_true[n], z_true]; rng.rand(N) < Bt[z_true, z_true[n]]
    gen = time.perf_counter() - gen0
    secs = max(1e-9, (gen0 - t0) - gen)
    full = secs * (N / float(nodes))
    return {"value": 1.0 / full, "unit": AIS_UNIT, "cores": 1, "kind": "port",
            "sample": "oracle count form of naive_sample_f (vectorised NumPy) on the first %d of %d nodes of ONE particle; %.2f s "
                      "measured on 1 core (generation of the sample's rows excluded), scaled x%.1f to a full feature sweep"
                      % (nodes, N, secs, N / float(nodes)), "measured_s": secs}


def _ais_device(args, R, P):
    """DeviceAIS holding config 5's matrix: A is generated on the device, 2048 rows at a time; the first chunk also goes
    through the host-buffer path (pinned uint8 rows -> gk_ais_upload_rows) to measure it."""
    import torch
    from graphistician_b200.device import DeviceAIS
    N, Rk = args.ais_n, 3
    z_true, Bt, pi = ais_truth(N, Rk)
    g = torch.Generator(device=R.dev); g.manual_seed(1234 + R.rank)
    dev = DeviceAIS(None, P, Rk, N=N)
    zt = torch.from_numpy(z_true).to(R.dev); Btt = torch.from_numpy(Bt).float().to(R.dev)
    rows_per = 2048
    t_up, chunk_bytes = 0.0, 0
    for r0 in range(0, N, rows_per):
        Pm = Btt[zt[r0:r0 + rows_per]][:, zt]
        rows = (torch.rand(Pm.shape, device=R.dev, generator=g) < Pm).to(torch.uint8).contiguous()
        torch.cuda.synchronize()
        if r0 == 0:
            host_chunk = rows.cpu().pin_memory()
            t0 = time.perf_counter()
            dev.upload_rows(host_chunk.numpy(), 0)
            dev.ctx.synchronize()
            t_up, chunk_bytes = time.perf_counter() - t0, host_chunk.numel()
        dev.upload_rows(None, r0, device_ptr=rows.data_ptr(), nrows=rows.shape[0])
        dev.ctx.synchronize()
        del rows, Pm
    return dev, t_up, chunk_bytes


def bench_ais(args, R, steps, warmup, cpu=True, scan=(8, 32, 512)):
    """config 5: P co-resident AIS particles per GPU over one bit-packed N = 65536 adjacency matrix; a step = one batched
    feature sweep (every particle's exact sequential sweep).  Particles are partitioned over the ranks (no data-path
    collective), so the N-GPU value is the sum of the ranks' particle-sweeps."""
    N, Rk, P = args.ais_n, 3, args.ais_particles
    dev, t_up, chunk_bytes = _ais_device(args, R, P)
    z_true = ais_truth(N, Rk)[0]
    rng = np.random.RandomState(5678 + R.rank)
    dev.set_labels(rng.randint(0, Rk, N))
    model_b0 = model_b1 = 1.0; alpha0 = 10 * np.ones(Rk)
    state = {"beta_i": 0}
    betas = np.linspace(0, 1, 100)[1:]

    def theta():
        """sample_theta of every particle from the counts the device keeps (internals/graph_models.py:337-371)."""
        n_edge, sizes = dev.counts()
        beta = betas[state["beta_i"] % len(betas)]
        lpi = np.log(np.stack([rng.dirichlet(alpha0 + sizes[k]) for k in range(P)]))
        n_pair = sizes[:, :, None] * sizes[:, None, :]
        Bm = rng.beta(model_b1 + beta * n_edge, model_b0 + beta * (n_pair - n_edge)).astype(np.float32).astype(np.float64)
        return lpi, Bm, beta

    lpi, Bm, beta = theta()

    def step_resident():                                       # the device part alone, theta fixed
        dev.sweep(lpi, Bm, beta, seed=7)

    def step_e2e():                                            # one AIS step as the driver runs it: counts -> theta draws -> sweep
        state["beta_i"] += 1
        l, b, be = theta()
        dev.sweep(l, b, be, seed=7)

    R.timed(step_resident, warmup)
    sampler = ClockSampler(R.local); sampler.start()
    l0 = R.ctx.launch_count()
    ms = R.timed(step_resident, steps)
    launches = R.ctx.launch_count() - l0
    clocks = sampler.result()
    e2e_steps = max(3, min(steps, 8))
    R.timed(step_e2e, 1)
    ms_e2e = R.timed(step_e2e, e2e_steps)
    R.ctx.kernel_timing(True)
    for _ in range(3):
        step_resident()
    fam0, fam1 = R.ctx.kernel_timing_read(0), R.ctx.kernel_timing_read(1)
    R.ctx.kernel_timing(False)
    e1, s1 = dev.counts(); e2, s2 = dev.counts(recompute=True)
    assert np.array_equal(e1, e2) and np.array_equal(s1, s2), "AIS cell table drifted"
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    algo = 2.0 * N * N / 8.0                                   # bit-packed row + column pass, shared by the P particles
    kms = fam0[0] / 3.0
    achieved = algo / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
    word_ops = 2.0 * N * (N / 32.0) * P * (Rk - 1)
    rec = {
        "value": R.world * P * steps / (ms * 1e-3), "unit": AIS_UNIT, "steps": steps, "ms_per_step": ms / steps, "dtype": "u32/f64",
        "config": {"workload": "ais_sbm_particle_sweep N=%d R=3 P=%d per GPU (configs[4])" % (N, P), "N": N, "R": Rk, "P": P},
        "gpu_launches": int(launches), "clocks": clocks,
        "particles_total": R.world * P, "device_bytes": dev.device_bytes(),
        "e2e": {"value": R.world * P * e2e_steps / (ms_e2e * 1e-3), "unit": AIS_UNIT,
                "h2d_bytes_per_step": int(P * (1 + Rk + 2 * Rk * Rk) * 8), "d2h_bytes_per_step": int(P * (Rk * Rk + Rk + 1) * 8),
                "what": "counts D2H -> Dirichlet / Beta draws of every particle on the host -> tables H2D -> batched sweep; A stays "
                        "resident (it is the data of the whole AIS run)",
                "upload_rows_host_GBps": (chunk_bytes / t_up / 1e9) if t_up > 0 else None},
        "roofline": {"bound": "hbm", "kernel": "k_ais_rowcnt", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic("k_ais_rowcnt")[0],
                     "algorithmic_bytes_per_launch": algo / ((N + 1023) // 1024),
                     "launches_per_sweep": fam0[1] / 3.0, "kernel_ms_per_sweep": kms,
                     "other_kernels_ms_per_sweep": {"k_ais_seq": fam1[0] / 3.0},
                     "note": "one streamed row of A serves all P particles, so the pass is bound by the integer pipe, not by HBM: "
                             "%.3g AND-popcount word operations per sweep = %.3g /s (Harley-Seal: 5 POPC + 46 LOP3 per 16 words); "
                             "the kernel-timing pass serialises the two particle groups that the timed sweeps overlap"
                             % (word_ops, word_ops / (kms * 1e-3) if kms > 0 else 0.0),
                     "sweep_algorithmic_bytes": algo, "sweep_frac_of_peak": algo / (ms / steps * 1e-3) / 1e9 / peak},
        "l2": "A and A^T bit-packed (%.2f GB) exceed the 126 MB L2" % (2 * N * N / 8 / 1e9),
        "data": "synthetic SBM, R=3, generated on the device in 2048-row chunks (the 4.3 GB uint8 matrix never exists on the host)",
    }
    if cpu and R.rank == 0 and R.world == 1 and not args.no_cpu_baseline:
        rec["cpu_baseline"] = cpu_baseline_ais(args, N, Rk)

    def stationary(dv, Ps):
        """The other end of the ladder: labels at the planted blocks, theta from their tempered posterior; (almost) no node moves."""
        dv.set_labels(z_true)
        n_e, szs = dv.counts()
        lp_s = np.log(np.stack([rng.dirichlet(alpha0 + szs[k]) for k in range(Ps)]))
        n_p = szs[:, :, None] * szs[:, None, :]
        B_s = np.clip(rng.beta(1.0 + 0.5 * n_e, 1.0 + 0.5 * (n_p - n_e)), 1e-6, 1 - 1e-6).astype(np.float32).astype(np.float64)
        moved = []
        st = lambda: moved.append(float(dv.sweep(lp_s, B_s, 0.5, seed=11)[0].mean()))
        R.timed(st, 2)
        ms_s = R.timed(st, 5)
        return {"value": R.world * Ps * 5 / (ms_s * 1e-3), "unit": AIS_UNIT, "ms_per_step": ms_s / 5, "n_changed_mean_last_sweep": moved[-1]}

    rec["regime"] = "labels drawn uniformly and theta from their tempered posterior: about two thirds of the nodes move per sweep " \
                    "(the start of an AIS ladder, the per-node path of k_ais_seq); `stationary` is the other end"
    rec["n_changed_mean_last_sweep"] = float(dev.sweep(lpi, Bm, beta, seed=7)[0].mean())
    rec["stationary"] = stationary(dev, P)
    # the whole of estimate_marginal_likelihood (examples/demo_ais.py:10-62) for these P particles in one C call (gk_ais_run):
    # 100 temperatures x 2 Gibbs steps from uniformly drawn labels, conjugate draws on the host, wall clock
    t0 = time.perf_counter()
    w = dev.run(1, 1.0, 1.0, alpha0, 100, 2, f0=rng.randint(0, Rk, N), seed=1234 + R.rank)
    t_ladder = time.perf_counter() - t0
    from scipy.special import logsumexp
    rec["ladder"] = {"n_temps": 100, "steps_per_temp": 2, "particles": P, "seconds": t_ladder,
                     "particle_sweeps_per_s": P * 198 / t_ladder, "log_Z": float(-np.log(P) + logsumexp(w)),
                     "log_weight_std": float(np.std(w))}
    dev.close()
    rec["particle_scan"] = {str(P): {"value": rec["value"], "ms_per_step": rec["ms_per_step"], "stationary": rec["stationary"]["value"]}}
    for Ps in scan:                                            # P swept per GPU (resident sweeps only)
        if Ps == P or R.world > 1:
            continue
        dv, _, _ = _ais_device(args, R, Ps)
        dv.set_labels(rng.randint(0, Rk, N))
        n_edge, sizes = dv.counts()
        lp_s = np.log(np.stack([rng.dirichlet(alpha0 + sizes[k]) for k in range(Ps)]))
        n_pair = sizes[:, :, None] * sizes[:, None, :]
        B_s = rng.beta(1.0 + beta * n_edge, 1.0 + beta * (n_pair - n_edge)).astype(np.float32).astype(np.float64)
        st = lambda: dv.sweep(lp_s, B_s, beta, seed=7)
        R.timed(st, 2)
        ms_s = R.timed(st, 3)
        rec["particle_scan"][str(Ps)] = {"value": Ps * 3 / (ms_s * 1e-3), "ms_per_step": ms_s / 3, "stationary": stationary(dv, Ps)["value"]}
        dv.close()
    return rec


def run_ais(args, R):
    rec = bench_ais(args, R, args.steps, args.warmup)
    out = {"metric": "AIS particle-sweeps/sec at N=%d (configs[4])" % args.ais_n, "value": rec["value"], "unit": AIS_UNIT,
           "n_gpus": R.world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": rec["ms_per_step"],
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": rec["dtype"], "data": rec["data"],
           "parallelism": "particles partitioned x%d, no data-path collective" % R.world}
    out.update({k: rec[k] for k in rec if k not in out})
    out["host"] = host_info()
    if R.rank == 0:
        print(json.dumps(out), flush=True)
    if R.world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------------
# CPU baseline of our arm: the oracle (vectorised T1 restatement of the reference's NumPy path) on a BOUNDED sample
# of the same workload, one core, sample stated.  The reference itself cannot run (Python 2, missing modules and
# packages: SURVEY.md 8c), so kind = "port".
# --------------------------------------------------------------------------------------------------
class _RowOffset:
    """rows [lo, hi) of a transposed array addressed by their global row index (oracle.sbm.node_conditional's At / Wt)."""

    def __init__(self, a, lo):
        self.a, self.lo = a, lo

    def __getitem__(self, idx):
        return self.a[(idx[0] - self.lo,) + tuple(idx[1:])]


def _stats_rows(A, W, c, K, lo, hi):
    """K1 of the oracle (bincount over the fused cell index) restricted to rows [lo, hi)."""
    B = W.shape[2]
    c = np.asarray(c, dtype=np.int64)
    for r0 in range(lo, hi, 256):
        r1 = min(hi, r0 + 256)
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


def _sbm_nodes(A, W, state, lo, hi):
    """K1 statistics over rows [lo, hi) + the sequential conditionals / draws of those nodes (oracle port)."""
    from oracle import sbm
    K = len(state["m"])
    _stats_rows(A, W, state["c"], K, lo, hi)
    gc_ = sbm.gaussian_coefficients(state["mus"], state["sigmas"])
    logm = np.log(state["m"]); c = np.array(state["c"], dtype=np.int64).copy()
    u = np.random.RandomState(1).rand(A.shape[0])
    At = _RowOffset(np.ascontiguousarray(A[:, lo:hi].T), lo)
    Wt = _RowOffset(np.ascontiguousarray(W[:, lo:hi].transpose(1, 0, 2)), lo)
    for n1 in range(lo, hi):
        lp = sbm.node_conditional(n1, A, W, c, logm, state["p"], gc_, sbm.SELF_OLD, 1.0, At, Wt)
        c[n1] = sbm.sample_discrete_from_log(lp, u[n1])
    return hi - lo


def cpu_baseline_sbm(d, model, args, nodes=None):
    N = d["A"].shape[0]
    nodes = nodes or args.cpu_nodes or min(N, 2048)
    state = dict(c=np.asarray(model.c), m=np.asarray(model.m), p=np.clip(model.p, 1e-12, 1 - 1e-12),
                 mus=model.mus, sigmas=model.sigmas)
    t0 = time.perf_counter()
    _sbm_nodes(d["A"], d["W"], state, 0, nodes)
    secs = time.perf_counter() - t0
    full = secs * (N / float(nodes))
    return {"value": 1.0 / full, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "oracle (vectorised NumPy restatement) on the first %d of %d nodes: K1 statistics over those rows "
                      "+ their sequential conditionals/draws; %.1f s measured on 1 core, scaled x%.1f to a full sweep "
                      "(the fully measured, all-cores number is the --impl reference arm)" % (nodes, N, secs, N / float(nodes)),
            "measured_s": secs}


def _eigen_nodes(A, F, lm, mu0, sigma_F, lo, hi, seed):
    """Both Omega entries of rows [lo, hi) from the C Devroye sampler + the conditionals / draws of those nodes."""
    from oracle import eigenmodel as eg, pg
    N = A.shape[0]
    O1 = pg.pgdraw_eigen(F, lm, mu0, seed=seed, r0=lo, r1=hi)              # omega[n, :]
    O2 = pg.pgdraw_eigen(F, lm, mu0, seed=seed + 1, r0=lo, r1=hi)          # omega[:, n]^T (Mu is symmetric)
    S = np.zeros((hi, N)); S[lo:hi] = O1 + O2                              # rows below lo are never read
    R = np.zeros((hi, N))
    R[lo:hi] = A[lo:hi].astype(np.float64) + A[:, lo:hi].T - 1.0 - mu0 * S[lo:hi]
    z = np.random.RandomState(1).randn(N, F.shape[1])
    eg.resample_F(S, R, F, lm, sigma_F, z, nodes=(lo, hi))
    return hi - lo


def cpu_baseline_eigen(d, model, args, nodes=None):
    from oracle import pg
    A = d["A"]
    N = A.shape[0]
    nodes = nodes or args.cpu_nodes or min(N, 512)
    F = np.asarray(model.F, dtype=np.float64); lm = np.asarray(model.lmbda, dtype=np.float64)
    t0 = time.perf_counter()
    _eigen_nodes(A, F, lm, float(model.mu_0), model.sigma_F, 0, nodes, 1)
    secs = time.perf_counter() - t0
    full = secs * (N / float(nodes))
    return {"value": 1.0 / full, "unit": UNIT, "cores": pg.num_threads(), "kind": "port",
            "sample": "oracle on the first %d of %d nodes: C Polya-gamma sampler (OpenMP x%d) for their 2N Omega entries + "
                      "sequential conditionals/draws in NumPy (1 core); %.2f s measured, scaled x%.1f to a full sweep"
                      % (nodes, N, pg.num_threads(), secs, N / float(nodes)),
            "measured_s": secs}


def _latent_rows(A, W, L, mu0, mu_self, sigma, lo, hi):
    """22 value + gradient evaluations of adjacency.py:244-275 restricted to rows [lo, hi) + their NIW statistics."""
    from oracle import latent
    for r0 in range(lo, hi, 128):
        r1 = min(hi, r0 + 128)
        Ar = A[r0:r1].astype(np.float64); Atr = np.ascontiguousarray(A[:, r0:r1].T).astype(np.float64)
        for _ in range(22):
            latent.rows_value_and_grad(L, mu0, mu_self, Ar, Atr, r0, sigma)
        e = A[r0:r1] != 0
        We = W[r0:r1][e].astype(np.float64)
        We.sum(0); We.T.dot(We)
    return hi - lo


def cpu_baseline_latent(d, model, args, nodes=None):
    A, W = d["A"], d["W"]
    N = A.shape[0]
    nodes = nodes or args.cpu_nodes or min(N, 256)
    adj = model.adjacency
    L = np.asarray(adj.L, dtype=np.float64)
    t0 = time.perf_counter()
    _latent_rows(A, W, L, adj.mu_0, adj.mu_self, adj.sigma, 0, nodes)
    secs = time.perf_counter() - t0
    full = secs * (N / float(nodes))
    return {"value": 1.0 / full, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "oracle (vectorised NumPy restatement, fp64) on the first %d of %d rows: 22 value+gradient "
                      "evaluations + NIW statistics; %.1f s measured, scaled x%.1f to a full sweep" % (nodes, N, secs, N / float(nodes)),
            "measured_s": secs}


# --------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU path (the oracle port -- the reference is Python 2 and cannot run here) on ALL the
# host cores, FULL sweeps measured wall-clock at the config size: the N nodes are spread over one forked worker per
# usable core, every worker runs the per-node arithmetic of its contiguous node range from the same start state.
# (The reference's own sweep is one serial loop; spreading it over the cores is the most favourable reading of "all
# the host threads it can use".)  `steps` in the line is the number of sweeps actually timed inside --ref-budget.
# --------------------------------------------------------------------------------------------------
_SHARED = {}


def _ref_task(task):
    kind, lo, hi = task
    s = _SHARED
    if kind == "sbm":
        return _sbm_nodes(s["A"], s["W"], s["state"], lo, hi)
    if kind == "eigen":
        return _eigen_nodes(s["A"], s["F"], s["lm"], s["mu0"], s["sigma_F"], lo, hi, 1 + 2 * lo)
    if kind == "latent":
        return _latent_rows(s["A"], s["W"], s["L"], 0.0, 0.0, s["sigma"], lo, hi)
    raise ValueError(kind)


def _ref_sweeps(kind, N, procs, args):
    """Wall-clock seconds of full sweeps over all N nodes on `procs` forked workers; (times, warmup_done)."""
    import multiprocessing as mp
    per = max(8, (N + 8 * procs - 1) // (8 * procs))          # ~8 tasks per worker: load balance
    tasks = [(kind, lo, min(N, lo + per)) for lo in range(0, N, per)]
    ctx = mp.get_context("fork")
    times = []
    warm = 0
    budget_t0 = time.perf_counter()
    with ctx.Pool(procs) as pool:
        for i in range(1 + args.steps):
            t0 = time.perf_counter()
            done = sum(pool.map(_ref_task, tasks, chunksize=1))
            dt = time.perf_counter() - t0
            assert done == N
            if i == 0 and args.warmup > 0 and dt < 5.0:       # a cheap first sweep is the warm-up; a long one is kept
                warm = 1
                continue
            times.append(dt)
            if len(times) >= args.steps or time.perf_counter() - budget_t0 > args.ref_budget:
                break
    return times, warm


def _ais_ref_task(task):
    lo, hi, N, R = task
    z_true, Bt, pi = ais_truth(N, R)
    z = np.random.RandomState(1).randint(0, R, N)
    t0 = time.perf_counter()
    _ais_cpu_nodes(N, R, z_true, Bt, z, np.clip(Bt, 1e-6, 1 - 1e-6), pi, 0.5, lo, hi)
    t1 = time.perf_counter()
    rng = np.random.RandomState(0)
    for n in range(lo, hi):
        rng.rand(N) < Bt[z_true[n], z_true]; rng.rand(N) < Bt[z_true, z_true[n]]
    return (t1 - t0) - (time.perf_counter() - t1)


def reference_ais(args, procs):
    """--impl reference --workload ais: the oracle's count form of naive_sample_f on a bounded sample of nodes, one
    forked worker per usable core, scaled to full feature sweeps (a particle's sweep is serial; particles are independent,
    so `procs` cores run `procs` particles side by side)."""
    import multiprocessing as mp
    N, R = args.ais_n, 3
    per = min(1024, N)
    tasks = [((k * per) % (N - per + 1), (k * per) % (N - per + 1) + per, N, R) for k in range(procs)]
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(procs) as pool:
        secs = pool.map(_ais_ref_task, tasks, chunksize=1)
    wall = time.perf_counter() - t0
    per_sweep = float(np.mean(secs)) * (N / float(per))        # one particle's full sweep on one core
    value = procs / per_sweep
    cfg = workload_config("ais", args)
    return {"impl": "reference", "metric": "AIS particle-sweeps/sec at N=%d (configs[4])" % N, "value": value, "unit": AIS_UNIT,
            "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": 1e3 * per_sweep / procs, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": AIS_UNIT, "cores": procs, "kind": "port",
                             "sample": "%d nodes of one particle's feature sweep per core on %d cores (oracle count form, NumPy), "
                                       "%.1f s wall; scaled x%.1f to full sweeps, %d particles side by side" % (per, procs, wall, N / float(per), procs)},
            "e2e": {"value": value, "unit": AIS_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "host": host_info()}


def reference_workload(name, args, procs):
    N = args.n
    os.environ.setdefault("OMP_NUM_THREADS", "1")              # one worker per core; the C sampler stays single-threaded in each
    _SHARED.clear()
    if name == "sbm":
        K, B = args.k, args.b
        d = make_sbm_data(N, K, B, pin=False)
        np.random.seed(5678)                                   # same start state as our arm's model (prior draw, seed 5678)
        m = np.random.dirichlet(np.ones(K)); c = np.random.choice(K, p=m, size=N); p = np.random.beta(1.0, 1.0, size=(K, K))
        _SHARED.update(A=d["A"], W=d["W"], state=dict(c=c, m=m, p=np.clip(p, 1e-12, 1 - 1e-12), mus=d["mus"], sigmas=d["sigmas"]))
        what = "K1 statistics + sequential conditionals/draws of all %d nodes (oracle, vectorised NumPy)" % N
    elif name == "eigen":
        from oracle import pg
        pg.build()
        D = args.d
        d = make_eigen_data(N, D, pin=False)
        np.random.seed(5678)
        F = np.random.randn(N, D)
        _SHARED.update(A=d["A"], F=F, lm=np.ones(D), mu0=float(d["mu0"]), sigma_F=1.0)
        what = "C Polya-gamma sampler for all 2 N^2 Omega entries + sequential conditionals/draws of all %d nodes" % N
    elif name == "latent":
        d = make_latent_data(N, 2, 1, pin=False)
        np.random.seed(5678)
        _SHARED.update(A=d["A"], W=d["W"], L=np.sqrt(2.0) * np.random.randn(N, 2), sigma=2.0)
        what = "22 value+gradient evaluations over all %d rows + NIW statistics (oracle, vectorised NumPy, fp64)" % N
    else:
        raise SystemExit(name)
    times, warm = _ref_sweeps(name, N, procs, args)
    _SHARED.clear()
    secs = float(np.mean(times))
    return {"value": 1.0 / secs, "unit": UNIT, "steps": len(times), "warmup": warm, "ms_per_step": 1e3 * secs,
            "dtype": args.dtype, "config": workload_config(name, args),
            "cpu_baseline": {"value": 1.0 / secs, "unit": UNIT, "cores": procs, "kind": "port",
                             "sample": "FULL sweeps, wall clock: %s, spread over %d forked workers (one per usable core); "
                                       "%d sweep(s) timed: %s s" % (what, procs, len(times), ["%.2f" % t for t in times])},
            "e2e": {"value": 1.0 / secs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    procs = usable_cores()
    if args.workload == "ais":
        print(json.dumps(reference_ais(args, procs)), flush=True)
        return
    names = ["sbm", "eigen", "latent"] if args.workload == "all" else [args.workload]
    recs = {n: reference_workload(n, args, procs) for n in names}
    r = recs[names[0]]
    out = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
           "warmup": r["warmup"], "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": r["config"],
           "arithmetic": "fp64 NumPy on the same fp32 inputs", "cpu_baseline": r["cpu_baseline"], "e2e": r["e2e"],
           "workloads": {n: recs[n] for n in names[1:]}, "host": host_info()}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
