"""Timing of the batched AIS sweep (gk_ais_sweep) at config 5's size: N = 65536, SBM truth R = 3, P particles on one GPU.
A is generated on the device in row chunks (torch is plumbing only) and bit-packed by gk_ais_upload_rows."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from graphistician_b200 import _lib  # noqa: E402
from graphistician_b200.device import DeviceAIS  # noqa: E402


def build(N, P, R, seed=1234):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    rng = np.random.RandomState(seed)
    pi = rng.dirichlet(10 * np.ones(R))
    z = rng.choice(R, size=N, p=pi)
    Bt = rng.beta(1, 1, (R, R))
    dev = DeviceAIS(None, P, R, N=N)
    zt = torch.from_numpy(z).cuda()
    Btt = torch.from_numpy(Bt).float().cuda()
    step = 2048
    for r0 in range(0, N, step):
        Pm = Btt[zt[r0:r0 + step]][:, zt]
        rows = (torch.rand(Pm.shape, device="cuda", generator=g) < Pm).to(torch.uint8).contiguous()
        torch.cuda.synchronize()
        dev.upload_rows(None, r0, device_ptr=rows.data_ptr(), nrows=rows.shape[0])
        dev.ctx.synchronize()
    return dev, z, Bt, pi


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    R = 3
    out = []
    for P in [int(a) for a in sys.argv[2:]] or [8, 32, 128]:
        t0 = time.time()
        dev, z, Bt, pi = build(N, P, R)
        rng = np.random.RandomState(1)
        dev.set_labels(rng.randint(0, R, N))
        t_build = time.time() - t0
        ctx = dev.ctx
        for beta, tag in ((0.02, "hot"), (1.0, "cold")):
            Bm = rng.beta(1, 1, (P, R, R))
            Bm[:] = Bt
            for _ in range(2):
                dev.sweep(np.log(pi), Bm, beta, seed=7)
            ctx.kernel_timing(True)
            t0 = time.time(); n = 3
            for _ in range(n):
                nchg, _ = dev.sweep(np.log(pi), Bm, beta, seed=7)
            dt = (time.time() - t0) / n
            ms0, l0, _ = ctx.kernel_timing_read(0); ms1, l1, _ = ctx.kernel_timing_read(1)
            ctx.kernel_timing(False)
            rec = dict(N=N, P=P, regime=tag, beta=beta, ms_per_batched_sweep=dt * 1e3, particle_sweeps_per_s=P / dt,
                       rowcnt_ms=ms0 / n, seq_ms=ms1 / n, n_changed_mean=float(np.mean(nchg)), build_s=t_build,
                       hbm_GBs_vs_2N2_over_8=P * 2 * N * N / 8 / dt / 1e9)
            print(json.dumps(rec), flush=True)
            out.append(rec)
        e1, s1 = dev.counts(); e2, s2 = dev.counts(recompute=True)
        assert np.array_equal(e1, e2) and np.array_equal(s1, s2)
        dev.close()
    return out


if __name__ == "__main__":
    main()
