"""Where the SBM sweep's wall time goes on the host: per-sweep wall times from random labels (burn-in), then a cProfile
of 20 steady-state sweeps.   Run on the GPU box:  python profiles/prof_sbm_host.py [N]"""
import cProfile
import pstats
import sys
import time
import types

sys.path.insert(0, ".")
import bench  # noqa: E402
from graphistician_b200 import _lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
args = types.SimpleNamespace(n=N, k=16, b=2, d=8, dtype="f32", workload="sbm")
wl = bench.setup_workload("sbm", args, 0, "f32")
model, network = wl["model"], wl["network"]
ctx = _lib.default_context()
for i in range(6):
    t = time.perf_counter(); model.resample(network); ctx.synchronize()
    print("sweep %d from random labels: %.2f ms" % (i, 1e3 * (time.perf_counter() - t)))
t = time.perf_counter()
for _ in range(20):
    model.resample(network)
ctx.synchronize()
print("steady state: %.3f ms / sweep" % (1e3 * (time.perf_counter() - t) / 20))
pr = cProfile.Profile(); pr.enable()
for _ in range(20):
    model.resample(network)
ctx.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(35)
