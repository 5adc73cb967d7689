"""ctypes binding of libdlsb200.so (include/dls_b200.h).

This is the Python twin of the JNI / FFM shim shown in INTEGRATION.md: it only marshals plain pointers and sizes.
There is no CPU fallback: importing this module fails loudly if the shared library has not been built
(`python -c "import __graft_entry__ as g; g.build()"` or `make -C deeplearning/scala_b200/csrc`), and creating a
context fails loudly without an sm_100 device.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libdlsb200.so")


class NativeError(RuntimeError):
    """Any non-zero dls_status other than DLS_ERR_INVALID (which maps to ValueError, the mirror of the
    reference's IllegalArgumentException)."""

    def __init__(self, status, message):
        super().__init__("libdlsb200 status %d: %s" % (status, message))
        self.status = status


DLS_OK, DLS_ERR_INVALID, DLS_ERR_NO_DEVICE, DLS_ERR_OOM, DLS_ERR_CUDA, DLS_ERR_UNSUPPORTED, DLS_ERR_COMM = range(7)
DLS_LAUNCH_DEFAULT, DLS_LAUNCH_REDUCE_SUM = 0, 1
DLS_HASH_WANG, DLS_HASH_LCG = 0, 1
DLS_GEMM_FP32, DLS_GEMM_TF32, DLS_GEMM_BF16 = 0, 1, 2
DLS_OPT_MOMENTUM, DLS_OPT_ADAM = 0, 1
DLS_CONV_FORWARD, DLS_CONV_BACKWARD_DATA, DLS_CONV_BACKWARD_WEIGHTS = 0, 1, 2
DLS_SUM_STATIC_MAX_FLOATS = 1 << 28


class CtxOpts(C.Structure):
    _fields_ = [("device", C.c_int32), ("streams_per_device", C.c_int32), ("rng_hash", C.c_int32),
                ("gemm_mode", C.c_int32), ("allocator_cap_bytes", C.c_uint64), ("plan_cache_capacity", C.c_int32),
                ("reserved", C.c_int32)]


class GemmOpts(C.Structure):
    _fields_ = [("mode", C.c_int32), ("bias_mode", C.c_int32), ("activation", C.c_int32), ("flags", C.c_int32)]


class ConvOpts(C.Structure):
    _fields_ = [("mode", C.c_int32), ("activation", C.c_int32), ("in_stride", C.c_int32), ("dz_stride", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("kernel_launches", "plan_cache_hits", "plan_cache_misses", "bytes_in_use",
                                          "bytes_cached", "driver_allocs", "h2d_bytes", "d2h_bytes")]


if not os.path.exists(LIB_PATH):
    raise ImportError("libdlsb200.so is not built (expected at %s). Run `python -c 'import __graft_entry__ as g; "
                      "g.build()'` from the repo root. There is no CPU fallback." % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_P = C.c_void_p
_PP = C.POINTER(C.c_void_p)
_SIG = {
    "dls_abi_version": (C.c_int, []),
    "dls_last_error": (C.c_char_p, []),
    "dls_ctx_create": (C.c_int, [C.POINTER(CtxOpts), _PP]),
    "dls_ctx_destroy": (None, [_P]),
    "dls_ctx_sync": (C.c_int, [_P]),
    "dls_ctx_stats": (C.c_int, [_P, C.POINTER(Stats)]),
    "dls_ctx_trim": (C.c_int, [_P]),
    "dls_ctx_stream": (_P, [_P, C.c_int]),
    "dls_buf_alloc": (C.c_int, [_P, C.c_size_t, _PP]),
    "dls_buf_from_host": (C.c_int, [_P, _P, C.c_size_t, _PP, _PP]),
    "dls_buf_retain": (None, [_P]),
    "dls_buf_release": (None, [_P]),
    "dls_buf_length": (C.c_size_t, [_P]),
    "dls_buf_device_ptr": (_P, [_P]),
    "dls_buf_read": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_size_t, _PP, C.c_int, _PP]),
    "dls_buf_write": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_size_t, _PP]),
    "dls_buf_upload_async": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_size_t, _PP, C.c_int, _PP]),
    "dls_stream_wait_event": (C.c_int, [_P, C.c_int, _P]),
    "dls_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.POINTER(C.c_float))]),
    "dls_host_free": (None, [_P]),
    "dls_event_retain": (None, [_P]),
    "dls_event_release": (None, [_P]),
    "dls_event_wait": (C.c_int, [_P]),
    "dls_event_query": (C.c_int, [_P]),
    "dls_event_on_complete": (C.c_int, [_P, _P, _P]),
    "dls_event_elapsed_ms": (C.c_int, [_P, _P, C.POINTER(C.c_float)]),
    "dls_event_record": (C.c_int, [_P, C.c_int, _PP]),
    "dls_launch_tree": (C.c_int, [_P, C.c_char_p, C.c_size_t, _PP, C.c_int, _PP, C.c_int, C.c_uint32, _PP, _PP]),
    "dls_plan_cache_contains": (C.c_int, [_P, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]),
    "dls_plan_describe": (C.c_int, [_P, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]),
    "dls_sum": (C.c_int, [_P, _P, C.c_size_t, _PP, C.c_int, _PP, _PP]),
    "dls_sum_axis": (C.c_int, [_P, _P, C.c_size_t, C.c_size_t, C.c_size_t, _PP, C.c_int, _PP, _PP]),
    "dls_random": (C.c_int, [_P, C.c_size_t, C.c_uint32, _PP, _PP]),
    "dls_random_normal": (C.c_int, [_P, C.c_size_t, C.c_uint32, _PP, _PP]),
    "dls_sgd_update": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_float, _PP, C.c_int, _PP]),
    "dls_optimizer_step": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, C.c_size_t, _P]),
    "dls_matmul": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int, _PP, C.c_int, _PP, _PP]),
    "dls_matmul_ex": (C.c_int, [_P, _P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int,
                                C.POINTER(GemmOpts), _PP, C.c_int, _PP, _PP]),
    "dls_matmul_gated": (C.c_int, [_P, _P, _P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int,
                                   C.POINTER(GemmOpts), _PP, C.c_int, _PP, _PP]),
    "dls_matmul_sgd": (C.c_int, [_P, _P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_int]),
    "dls_conv3x3_supported": (C.c_int, [_P, C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int]),
    "dls_conv3x3_forward": (C.c_int, [_P, _P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                      C.POINTER(ConvOpts), _PP, C.c_int, _PP, _PP]),
    "dls_conv3x3_backward_data": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                            C.POINTER(ConvOpts), _PP, C.c_int, _PP, _PP]),
    "dls_conv3x3_backward_data_gated": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                                  C.c_size_t, C.POINTER(ConvOpts), _PP, C.c_int, _PP, _PP]),
    "dls_conv3x3_backward_weights": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                               C.POINTER(ConvOpts), _PP, C.c_int, _PP, _PP]),
    "dls_conv3x3_backward_weights_bias": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                                    C.POINTER(ConvOpts), _PP, C.c_int, _PP, _PP, _PP]),
    "dls_comm_unique_id": (C.c_int, [C.c_char_p]),
    "dls_comm_init": (C.c_int, [_P, C.c_char_p, C.c_int, C.c_int]),
    "dls_allreduce_sum": (C.c_int, [_P, _P, C.c_size_t, _PP, C.c_int, _PP]),
    "dls_allreduce_sum_many": (C.c_int, [_P, _PP, C.POINTER(C.c_size_t), C.c_int]),
    "dls_peer_init": (C.c_int, [_P, C.c_size_t]),
    "dls_peer_ready": (C.c_int, [_P]),
    "dls_peer_sgd_step": (C.c_int, [_P, _PP, _PP, C.POINTER(C.c_size_t), C.POINTER(C.c_int), C.c_int, C.c_float]),
    "dls_comm_destroy": (C.c_int, [_P]),
    "dls_graph_begin": (C.c_int, [_P]),
    "dls_graph_end": (C.c_int, [_P, _PP]),
    "dls_graph_abort": (C.c_int, [_P]),
    "dls_graph_launch": (C.c_int, [_P]),
    "dls_graph_info": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "dls_graph_retain": (None, [_P]),
    "dls_graph_release": (None, [_P]),
    "dls_buf_copy": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "dls_sum_static": (C.c_int, [_P, _P, C.c_size_t, _P]),
}
EXPORTS = tuple(_SIG)
for _name, (_res, _args) in _SIG.items():
    _fn = getattr(lib, _name)          # AttributeError here == header and library disagree
    _fn.restype = _res
    _fn.argtypes = _args

if lib.dls_abi_version() != 1:
    raise ImportError("libdlsb200.so ABI version %d, expected 1" % lib.dls_abi_version())


def last_error() -> str:
    return lib.dls_last_error().decode("utf-8", "replace")


def check(status: int) -> None:
    if status == DLS_OK:
        return
    msg = last_error()
    if status == DLS_ERR_INVALID:
        raise ValueError(msg)
    raise NativeError(status, msg)
