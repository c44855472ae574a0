"""ctypes binding of libgk.so (include/gk.h).  There is no CPU fallback: a missing library or a missing
CUDA device raises, loudly, at first use."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgk.so")

GK_F32, GK_F64 = 0, 1
SELF_SKIP, SELF_ONCE, SELF_TWICE = 0, 1, 2
SAMPLER_CUMSUM, SAMPLER_LSE = 0, 1
CTX_OPT_SBM_OVERLAP = 1
AIS_OPT_INCREMENTAL = 1

c_int, c_i64, c_dbl, c_vp, c_sz = ctypes.c_int, ctypes.c_int64, ctypes.c_double, ctypes.c_void_p, ctypes.c_size_t
PI64 = ctypes.POINTER(ctypes.c_int64)

# name -> (restype, argtypes); every symbol include/gk.h declares (tests/test_abi.py checks the two agree)
SIGNATURES = {
    "gk_ctx_create": (c_int, [c_int, ctypes.POINTER(c_vp)]),
    "gk_ctx_destroy": (None, [c_vp]),
    "gk_last_error": (ctypes.c_char_p, [c_vp]),
    "gk_ctx_set_stream": (c_int, [c_vp, c_vp]),
    "gk_ctx_synchronize": (c_int, [c_vp]),
    "gk_ctx_set_option": (c_int, [c_vp, c_int, c_int]),
    "gk_ctx_launch_count": (c_i64, [c_vp]),
    "gk_ctx_kernel_timing": (c_int, [c_vp, c_int]),
    "gk_ctx_kernel_timing_read": (c_int, [c_vp, c_int, ctypes.POINTER(c_dbl), PI64, ctypes.POINTER(c_dbl)]),
    "gk_network_create": (c_int, [c_vp, c_int, c_int, c_int, c_int, c_int, ctypes.POINTER(c_vp)]),
    "gk_network_upload": (c_int, [c_vp, c_vp, c_vp, c_vp, c_int, c_vp, c_vp]),
    "gk_network_set_diag": (c_int, [c_vp, c_vp, c_vp, c_vp]),
    "gk_network_destroy": (c_int, [c_vp, c_vp]),
    "gk_network_device_bytes": (c_sz, [c_vp]),
    "gk_block_stats": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp]),
    "gk_sbm_sweep": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_dbl, c_vp, c_int,
                             c_vp, PI64]),
    "gk_sbm_sweep_stats": (c_int, [c_vp, c_vp, c_int, c_vp, c_vp, c_vp, ctypes.POINTER(c_int)]),
    "gk_sbm_sweep_begin": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_dbl, c_vp]),
    "gk_sbm_partial_len": (c_sz, [c_vp, c_int]),
    "gk_sbm_block_partials": (c_int, [c_vp, c_vp, c_int, c_int, c_vp]),
    "gk_sbm_block_sample": (c_int, [c_vp, c_vp, c_int, c_int, c_vp]),
    "gk_sbm_get_block": (c_int, [c_vp, c_int, c_int, c_vp]),
    "gk_sbm_set_block": (c_int, [c_vp, c_int, c_int, c_vp]),
    "gk_sbm_export_block_dev": (c_int, [c_vp, c_int, c_int, c_vp]),
    "gk_sbm_import_block_dev": (c_int, [c_vp, c_int, c_int, c_vp]),
    "gk_sbm_sweep_end": (c_int, [c_vp, c_vp, c_vp, c_vp, PI64]),
    "gk_eig_create": (c_int, [c_vp, c_vp, c_int, ctypes.POINTER(c_vp)]),
    "gk_eig_destroy": (c_int, [c_vp, c_vp]),
    "gk_eig_device_bytes": (c_sz, [c_vp]),
    "gk_eig_set_option": (c_int, [c_vp, c_vp, c_int, c_int]),
    "gk_eig_set_params": (c_int, [c_vp, c_vp, c_vp, c_dbl, c_vp]),
    "gk_eig_pg_draw": (c_int, [c_vp, c_vp, ctypes.c_uint64, ctypes.POINTER(c_dbl)]),
    "gk_eig_set_S": (c_int, [c_vp, c_vp, c_vp]),
    "gk_eig_get_S": (c_int, [c_vp, c_vp, c_vp]),
    "gk_eig_resample_F": (c_int, [c_vp, c_vp, c_dbl, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "gk_eig_log_likelihood": (c_int, [c_vp, c_vp, c_int, ctypes.POINTER(c_dbl)]),
    "gk_eig_tn_draw": (c_int, [c_vp, c_vp, ctypes.c_uint64, ctypes.POINTER(c_dbl)]),
    "gk_eig_set_Zsym": (c_int, [c_vp, c_vp, c_vp]),
    "gk_truncnorm_draw": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_sz]),
    "gk_truncnorm_draw_static": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_sz]),
    "gk_truncnorm_static_max": (c_sz, []),
    "gk_eig_sweep_begin": (c_int, [c_vp, c_vp, c_dbl, c_vp]),
    "gk_eig_acc_len": (c_sz, [c_vp]),
    "gk_eig_block_nodes": (c_int, [c_vp]),
    "gk_eig_block_partials": (c_int, [c_vp, c_vp, c_int, c_int, c_vp]),
    "gk_eig_block_sample": (c_int, [c_vp, c_vp, c_int, c_int, c_vp, c_vp]),
    "gk_eig_set_block": (c_int, [c_vp, c_vp, c_int, c_int, c_vp]),
    "gk_eig_sweep_end": (c_int, [c_vp, c_vp, c_vp, c_vp]),
    "gk_latent_create": (c_int, [c_vp, c_vp, c_int, ctypes.POINTER(c_vp)]),
    "gk_latent_destroy": (c_int, [c_vp, c_vp]),
    "gk_latent_set_params": (c_int, [c_vp, c_vp, c_vp, c_dbl, c_dbl, c_dbl]),
    "gk_latent_get_params": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(c_dbl)]),
    "gk_latent_acc_len": (c_sz, [c_vp]),
    "gk_latent_lp_grad": (c_int, [c_vp, c_vp, ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl), c_vp, ctypes.POINTER(c_dbl),
                                  ctypes.POINTER(c_dbl)]),
    "gk_latent_eval_pass": (c_int, [c_vp, c_vp, c_vp]),
    "gk_latent_eval_finish": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl), c_vp,
                                      ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl)]),
    "gk_latent_hmc": (c_int, [c_vp, c_vp, c_int, c_dbl, c_int, c_vp, c_dbl, ctypes.POINTER(c_int), ctypes.POINTER(c_dbl)]),
    "gk_latent_hmc_begin": (c_int, [c_vp, c_vp, c_int, c_dbl, c_int, c_vp]),
    "gk_latent_set_weight_mode": (c_int, [c_vp, c_vp, c_int, c_dbl]),
    "gk_latent_weight_stats": (c_int, [c_vp, c_vp, c_vp]),
    "gk_latent_weight_stats_pass": (c_int, [c_vp, c_vp, c_vp]),
    "gk_latent_hmc_pass": (c_int, [c_vp, c_vp, c_vp]),
    "gk_latent_hmc_advance": (c_int, [c_vp, c_vp, c_vp]),
    "gk_eig_mf_omega": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_dbl, c_dbl, c_int, ctypes.c_uint64,
                                ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl)]),
    "gk_eig_mf_update_F": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_dbl, c_dbl, c_dbl, c_vp, c_vp]),
    "gk_eig_mf_pair_stats": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp]),
    "gk_eig_mf_probit_Z": (c_int, [c_vp, c_vp, c_vp, c_vp, c_dbl, c_int, ctypes.POINTER(c_dbl)]),
    "gk_mf_sbm_soft_stats": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp]),
    "gk_mf_sbm_update_c": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_dbl, c_int, c_vp]),
    "gk_latent_hmc_end": (c_int, [c_vp, c_vp, c_dbl, ctypes.POINTER(c_int), ctypes.POINTER(c_dbl)]),
    "gk_ais_create": (c_int, [c_vp, c_int, c_int, c_int, ctypes.POINTER(c_vp)]),
    "gk_ais_destroy": (c_int, [c_vp, c_vp]),
    "gk_ais_device_bytes": (c_sz, [c_vp]),
    "gk_ais_set_option": (c_int, [c_vp, c_vp, c_int, c_int]),
    "gk_ais_upload_rows": (c_int, [c_vp, c_vp, c_vp, c_int, c_int, c_int]),
    "gk_ais_upload_network": (c_int, [c_vp, c_vp, c_vp]),
    "gk_ais_nnz": (c_int, [c_vp, c_vp, PI64]),
    "gk_ais_set_labels": (c_int, [c_vp, c_vp, c_vp, c_int]),
    "gk_ais_get_labels": (c_int, [c_vp, c_vp, c_vp]),
    "gk_ais_counts": (c_int, [c_vp, c_vp, c_int, c_vp, c_vp]),
    "gk_ais_sweep": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_uint64, c_vp, c_vp]),
    "gk_ais_run": (c_int, [c_vp, c_vp, c_int, c_dbl, c_dbl, c_vp, c_int, c_int, c_vp, ctypes.c_uint64, c_vp]),
    "gk_comm_unique_id": (c_int, [c_vp, c_vp]),
    "gk_comm_init": (c_int, [c_vp, c_vp, c_int, c_int]),
    "gk_comm_destroy": (c_int, [c_vp]),
    "gk_comm_rank": (c_int, [c_vp]),
    "gk_comm_world": (c_int, [c_vp]),
    "gk_comm_allreduce": (c_int, [c_vp, c_vp, c_sz]),
}

_lib = None


class GKError(RuntimeError):
    pass


def load():
    """dlopen libgk.so and type every entry point.  Raises if the library was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GKError("libgk.so not found at %s: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def ptr(a):
    """Host pointer of a C-contiguous ndarray (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


class Context:
    """One gk_ctx: a device, a stream, scratch.  Not thread-safe (as the reference is not)."""

    def __init__(self, device=0):
        self.lib = load()
        h = c_vp()
        rc = self.lib.gk_ctx_create(device, ctypes.byref(h))
        if rc != 0:
            raise GKError("gk_ctx_create(device=%d) failed with %d: no usable CUDA device "
                          "(this engine has no CPU fallback)" % (device, rc))
        self.h = h
        self.device = device

    def check(self, rc):
        if rc != 0:
            raise GKError("libgk error %d: %s" % (rc, self.lib.gk_last_error(self.h).decode()))

    def set_stream(self, cuda_stream):
        self.check(self.lib.gk_ctx_set_stream(self.h, c_vp(cuda_stream)))

    def synchronize(self):
        self.check(self.lib.gk_ctx_synchronize(self.h))

    def set_option(self, option, value):
        self.check(self.lib.gk_ctx_set_option(self.h, int(option), int(value)))

    def launch_count(self):
        return int(self.lib.gk_ctx_launch_count(self.h))

    def kernel_timing(self, enable):
        self.check(self.lib.gk_ctx_kernel_timing(self.h, int(enable)))

    def kernel_timing_read(self, family):
        ms, n, b = c_dbl(), c_i64(), c_dbl()
        self.check(self.lib.gk_ctx_kernel_timing_read(self.h, family, ctypes.byref(ms), ctypes.byref(n), ctypes.byref(b)))
        return ms.value, n.value, b.value

    def init_comm(self, comm):
        """Give this context the engine's own NCCL communicator over the ranks of ``comm`` (sharding.Comm): rank 0 draws
        the id, the host broadcasts it, every rank joins.  From then on libgk issues the collectives of the row-sharded
        sweeps itself, in-stream (include/gk.h 'the engine's own communicator')."""
        if getattr(self, "native_world", 1) > 1 or comm.world < 2:
            return
        uid = np.zeros(128, np.uint8)
        if comm.rank == 0:
            self.check(self.lib.gk_comm_unique_id(self.h, ptr(uid)))
        uid = np.ascontiguousarray(comm.broadcast_np(uid, 0), dtype=np.uint8)
        self.check(self.lib.gk_comm_init(self.h, ptr(uid), int(comm.rank), int(comm.world)))
        self.native_world = comm.world

    def close(self):
        if getattr(self, "h", None):
            self.lib.gk_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx = {}
_default_device = 0


def set_default_device(device):
    """Device used by default_context() when none is named (one process per GPU: call with LOCAL_RANK)."""
    global _default_device
    _default_device = int(device)


def default_context(device=None):
    if device is None:
        device = _default_device
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]
