"""Times one mean-field SBM update (a18) at N=16384, K=16, B=2 on cuda:0 and prints the kernel-family timings.
Run on the GPU box:  python profiles/time_mf.py [N]"""
import ctypes
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel  # noqa: E402
from graphistician_b200.old_abstractions import FixedGaussianNetwork  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
C, B = 16, 2
rng = np.random.default_rng(0)
np.random.seed(0)
A = (rng.random((N, N), dtype=np.float32) < 0.3).astype(np.uint8)
W = rng.standard_normal((N, N, B), dtype=np.float32)
net = FixedGaussianNetwork(A, W)
model = GaussianStochasticBlockModel(N, B, C=C)
dev = net.device("f32")
ctx = dev.ctx
model.meanfieldupdate(net)
ctx.lib.gk_ctx_kernel_timing(ctx.h, 1)
t = time.time(); model.meanfieldupdate(net); ctx.lib.gk_ctx_synchronize(ctx.h)
print("meanfieldupdate wall s", time.time() - t)
for fam, name in ((6, "k_strip_soft"), (7, "k_mf_seq")):
    ms = ctypes.c_double(); n = ctypes.c_int64(); by = ctypes.c_double()
    ctx.lib.gk_ctx_kernel_timing_read(ctx.h, fam, ctypes.byref(ms), ctypes.byref(n), ctypes.byref(by))
    print(name, "ms", ms.value, "launches", n.value, "GB/s", by.value / 1e9 / (ms.value / 1e3 + 1e-12))
ctx.lib.gk_ctx_kernel_timing(ctx.h, 0)
for nm, fn in (("mf_update_c", lambda: model.mf_update_c(net)), ("mf_update_p", lambda: model.mf_update_p(net))):
    t = time.time(); fn(); print(nm, "wall s", time.time() - t)

# ---- mean-field logistic eigenmodel (a19) at the same N, D=8 ---------------------------------------------------
from graphistician_b200.eigenmodel import LogisticEigenmodel  # noqa: E402

del model, net, dev, W
em = LogisticEigenmodel(N, 8)
enet = FixedGaussianNetwork(A, None)
em.meanfieldupdate(enet)
for nm, fn in (("eig _meanfieldupdate_Omega (1000 draws/entry)", lambda: em._meanfieldupdate_Omega(enet)),
               ("eig _meanfieldupdate_F", lambda: em._meanfieldupdate_F(enet)),
               ("eig _meanfieldupdate_lmbda", lambda: em._meanfieldupdate_lmbda(enet)),
               ("eig meanfieldupdate", lambda: em.meanfieldupdate(enet))):
    t = time.time(); fn(); print(nm, "wall s", time.time() - t)
