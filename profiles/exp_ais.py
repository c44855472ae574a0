import sys, time, numpy as np
sys.path.insert(0, ".")
sys.argv = ["x", "65536", "64"]
import profiles.time_ais as T
dev, z, Bt, pi = T.build(65536, 64, 3)
ctx = dev.ctx
rng = np.random.RandomState(1)
Bm = np.zeros((64, 3, 3)); Bm[:] = Bt
def run(beta, tag, n=4):
    for k in range(n):
        ctx.kernel_timing(True)
        nchg, _ = dev.sweep(np.log(pi), Bm, beta, seed=7)
        ms1 = ctx.kernel_timing_read(1)[0]
        ctx.kernel_timing(False)
        print(tag, beta, k, "seq_ms %.2f" % ms1, "nchg mean %.1f max %d" % (nchg.mean(), nchg.max()), flush=True)
dev.set_labels(z); run(1.0, "truth")
dev.set_labels(z); run(0.02, "truth")
dev.set_labels(z); run(0.001, "truth")
dev.set_labels(rng.randint(0, 3, 65536)); run(0.02, "random")
