"""Tensor-core contraction vs the FMA kernel vs the oracle on one feature sweep (fp32 storage, D = 8)."""
import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
N = int(sys.argv[1]); tc = int(sys.argv[2])
from test_gpu_eigen import _problem
from graphistician_b200 import DeviceNetwork
from graphistician_b200.device import DeviceEigenmodel
from oracle import eigenmodel as eg
A, O, F, lm, mu0, z = _problem(N, 8, 5 + N)
net = DeviceNetwork(A, None, dtype="f32"); eig = DeviceEigenmodel(net, 8)
eig.set_option(4, tc)
eig.set_params(F, mu0, lm); eig.set_S(O + O.T)
F_gpu, st, prec, h = eig.resample_F(1.0, z, want_conditionals=True)
S, R = eg.logistic_sym(A.astype(float), O, mu0)
F_ref, prec_ref, h_ref = eg.resample_F(S, R, F, lm, 1.0, z)
print(N, "tc" if tc else "fma", "prec relerr %.2e  h relerr %.2e  F abserr %.2e" % (
    np.abs(prec - prec_ref).max() / np.abs(prec_ref).max(), np.abs(h - h_ref).max() / np.abs(h_ref).max(), np.abs(F_gpu - F_ref).max()))
