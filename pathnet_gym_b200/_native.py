"""ctypes binding of libpathnet_b200.so (declarations in include/pathnet_b200.h).

There is no CPU fallback: importing this module needs the built library, and
every compute entry point needs a CUDA device.  PyTorch is used by the callers
only to own device memory and streams.
"""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libpathnet_b200.so")

PN_MAX_LAYERS = 8
PN_MAX_MODULES = 32
PN_ABI_VERSION = 1

PN_FLAG_LSTM = 16
PN_OK, PN_ERR_INVALID, PN_ERR_CUDA, PN_ERR_STATE, PN_ERR_NOMEM = 0, -1, -2, -3, -4


class PnConfig(C.Structure):
    _fields_ = [("L", C.c_int32), ("M", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32),
                ("F", C.c_int32), ("hidden", C.c_int32), ("A", C.c_int32),
                ("kernel", C.c_int32 * PN_MAX_LAYERS), ("stride", C.c_int32 * PN_MAX_LAYERS),
                ("P", C.c_int32), ("B_max", C.c_int32), ("device", C.c_int32), ("flags", C.c_int32)]


class PnTensorDesc(C.Structure):
    _fields_ = [("offset", C.c_int64), ("length", C.c_int64), ("layer", C.c_int32), ("module", C.c_int32),
                ("kind", C.c_int32), ("reserved", C.c_int32)]


class PnProfileEntry(C.Structure):
    _fields_ = [("name", C.c_char * 40), ("total_ms", C.c_float), ("launches", C.c_int32)]


class PathNetError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libpathnet_b200 error %d: %s" % (code, msg))
        self.code = code


# name -> (restype, argtypes); must list every symbol include/pathnet_b200.h declares
_vp, _i, _f = C.c_void_p, C.c_int, C.c_float
SIGNATURES = {
    "pn_abi_version": (_i, []),
    "pn_create": (_i, [C.POINTER(PnConfig), C.POINTER(_vp)]),
    "pn_destroy": (None, [_vp]),
    "pn_last_error": (C.c_char_p, [_vp]),
    "pn_last_global_error": (C.c_char_p, []),
    "pn_slab_floats": (C.c_int64, [_vp]),
    "pn_num_tensors": (_i, [_vp]),
    "pn_tensor_table": (_i, [_vp, C.POINTER(PnTensorDesc), _i]),
    "pn_last_lin": (_i, [_vp]),
    "pn_workspace_bytes": (C.c_int64, [_vp]),
    "pn_debug_read": (_i, [_vp, _vp, _i]),
    "pn_debug_read_relu_masks": (C.c_int64, [_vp, _i, _i, _vp, C.c_int64]),
    "pn_engine_paths": (_i, [_vp, C.c_char_p, _i]),
    "pn_launch_count": (C.c_int64, [_vp]),
    "pn_lin_grads_region": (_i, [_vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "pn_stream_wait_lin_grads": (_i, [_vp, _vp]),
    "pn_peer_buffer_bytes": (C.c_int64, [_vp]),
    "pn_peer_check": (_i, [_vp, _vp]),
    "pn_peer_sum_small": (_i, [_vp, _vp, _i, _i, C.c_uint32, _vp]),
    "pn_peer_sum_part": (_i, [_vp, _vp, _i, _i, C.c_uint32, _i, _vp]),
    "pn_stream_wait_early_grads": (_i, [_vp, _vp, _vp]),
    "pn_peer_allreduce_lin": (_i, [_vp, _vp, _vp, _i, _i, C.c_uint32, _vp]),
    "pn_profile_enable": (_i, [_vp, _i]),
    "pn_profile_read": (_i, [_vp, C.POINTER(PnProfileEntry), _i]),
    "pn_bind_slabs": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "pn_set_genotypes": (_i, [_vp, _vp, _vp]),
    "pn_lstm_bind_state": (_i, [_vp, _vp, _vp]),
    "pn_set_frozen": (_i, [_vp, _vp]),
    "pn_forward": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "pn_train_step": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _i, _f, _vp, _vp, _vp, _vp]),
    "pn_train_step_host": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _i, _f, _vp, _vp]),
    "pn_stage_inputs_host": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _vp]),
    "pn_stage_rollout_host": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    "pn_train_step_staged": (_i, [_vp, _i, _f, _vp, _vp]),
    "pn_rmsprop_apply": (_i, [_vp, _f, _f, _f, _f, _f, _vp]),
    "pn_rmsprop_apply_range": (_i, [_vp, _i, _i, _i, _f, _f, _f, _f, _f, _vp]),
    "pn_rmsprop_apply_noalloc": (_i, [_vp, _vp, _vp, _vp, C.POINTER(PnTensorDesc), _vp, _i, _f, _f, _f, _f, _f, _vp]),
    "pn_reinit_unfrozen": (_i, [_vp, _vp, _vp]),
    "pn_geopath_init": (_i, [_vp, _i, _i, _i, _i, _i, C.c_uint64, C.c_uint32, _vp]),
    "pn_mutate": (_i, [_vp, _i, _i, _i, _i, _i, C.c_uint64, C.c_uint32, _vp]),
    "pn_tournament": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, C.c_uint64, C.c_uint32, _vp]),
    "pn_sample_actions": (_i, [_vp, _i, _i, C.c_uint64, C.c_uint32, C.c_uint32, _vp, _vp]),
    "pn_nstep_returns": (_i, [_vp, _vp, _vp, _vp, _i, _i, C.c_double, _vp, _vp, _vp]),
    "pn_preprocess_frames": (_i, [_vp, _i, _i, _i, _vp, _i, _i, _i, _vp, _vp]),
}

_lib = None


def load():
    """Load the shared library (loading needs no GPU; calling the compute entry points does)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(pathnet_gym_b200 has no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    v = lib.pn_abi_version()
    if v != PN_ABI_VERSION:
        raise ImportError("libpathnet_b200.so ABI %d != binding %d (rebuild)" % (v, PN_ABI_VERSION))
    _lib = lib
    return lib


def check(rc, handle=None):
    if rc < 0:
        lib = load()
        msg = lib.pn_last_error(handle) if handle else lib.pn_last_global_error()
        raise PathNetError(rc, (msg or b"").decode("utf-8", "replace"))
    return rc
