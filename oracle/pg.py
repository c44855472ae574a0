"""
oracle/pg.py -- TEST INFRASTRUCTURE (see oracle/__init__.py).

ctypes wrapper over oracle/pg.c (the plain-C PG(1,z) Devroye sampler standing in for the absent
pypolyagamma; eigenmodel.py:859-874) plus the analytic moments used for distributional parity:
E[omega] = tanh(z/2)/(2z)  (the identity the reference itself uses at eigenmodel.py:1062-1063).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")
_SO = os.path.join(_BUILD, "libpgoracle.so")
_lib = None


def build(force=False):
    """gcc -O2 -fopenmp oracle/pg.c -> oracle/_build/libpgoracle.so"""
    src = os.path.join(_HERE, "pg.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(_BUILD, exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-o", _SO, src, "-lm"])
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.pg1_draw_vec.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint64]
        _lib.pg1_draw_vec.restype = None
        _lib.pg1_draw_eigen.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double, ctypes.c_int,
                                        ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.c_uint64]
        _lib.pg1_draw_eigen.restype = None
        _lib.pg_oracle_num_threads.restype = ctypes.c_int
    return _lib


def num_threads():
    return _load().pg_oracle_num_threads()


def pgdraw(z, seed=0):
    """omega[i] ~ PG(1, z[i]) -- same call shape as ppg.pgdrawvpar(ppgs, ones, z, out)."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    out = np.empty_like(z)
    _load().pg1_draw_vec(z.ctypes.data, out.ctypes.data, z.size, seed)
    return out


def pgdraw_eigen(F, lmbda, mu_0, seed=0, r0=0, r1=None):
    """Omega rows [r0, r1) for the eigenmodel, Mu computed on the fly (eigenmodel.py:75-83,859-874)."""
    F = np.ascontiguousarray(F, dtype=np.float64)
    lmbda = np.ascontiguousarray(lmbda, dtype=np.float64)
    N, D = F.shape
    r1 = N if r1 is None else r1
    out = np.empty((r1 - r0, N))
    _load().pg1_draw_eigen(F.ctypes.data, lmbda.ctypes.data, float(mu_0), N, D, r0, r1, out.ctypes.data, seed)
    return out


def pg1_mean(z):
    """E[PG(1,z)] = tanh(z/2) / (2z), -> 1/4 as z -> 0."""
    z = np.abs(np.asarray(z, dtype=np.float64))
    small = z < 1e-6
    zz = np.where(small, 1.0, z)
    return np.where(small, 0.25 - z * z / 48.0, np.tanh(zz / 2) / (2 * zz))


def pg1_var(z):
    """Var[PG(1,z)] = (sinh(z) - z) / (4 z^3 cosh^2(z/2)), -> 1/24 as z -> 0."""
    z = np.abs(np.asarray(z, dtype=np.float64))
    small = z < 1e-3
    zz = np.where(small, 1.0, z)
    v = (np.sinh(zz) - zz) / (4 * zz ** 3 * np.cosh(zz / 2) ** 2)
    return np.where(small, 1.0 / 24.0, v)
