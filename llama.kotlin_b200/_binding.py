"""ctypes binding of include/llamakotlin_cuda.h (the C-ABI product library).

There is no CPU fallback: if the shared library is missing this raises, and if
no B200 is present lk_ctx_create fails with LK_ECUDA (IllegalStateException).
"""
from __future__ import annotations

import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
# LK_LIBRARY: another build of the same library (A/B measurements under tools/); symbols it lacks are left unbound
LIB_PATH = os.environ.get("LK_LIBRARY") or os.path.join(_HERE, "libllamakotlin_cuda.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "llamakotlin_cuda.h")

LK_OK, LK_EINVAL_DIM, LK_EINVAL_TYPE, LK_EBOUNDS, LK_ENOTIMPL, LK_ECUDA, LK_ENCCL, LK_EINVAL_ARG, LK_ENOMEM = range(9)
LK_COMM_ID_BYTES = 128
LK_IPC_HANDLE_BYTES = 64


class IllegalArgumentException(ValueError):
    """kotlin.IllegalArgumentException (dim / type mismatch, core/GGMLComputeOps.kt:1440-1449)."""


class IndexOutOfBoundsException(IndexError):
    """kotlin.IndexOutOfBoundsException (buffer bounds, core/GGMLTypes.kt:556-558)."""


class IllegalStateException(RuntimeError):
    """kotlin.IllegalStateException (device / NCCL failures; GGMLStatus.FAILED at backend level)."""


class lk_tensor(C.Structure):
    _fields_ = [
        ("type", C.c_int32), ("reserved", C.c_int32),
        ("ne", C.c_int64 * 4), ("nb", C.c_uint64 * 4),
        ("host_base", C.c_void_p), ("dev_buf", C.c_void_p),
        ("buffer_size", C.c_uint64), ("data_offset", C.c_uint64),
    ]


class lk_matmul_fusion(C.Structure):
    _fields_ = [
        ("norm_x", C.POINTER(lk_tensor)), ("norm_w", C.POINTER(lk_tensor)), ("eps", C.c_float),
        ("residual", C.POINTER(lk_tensor)), ("sum_dst", C.POINTER(lk_tensor)), ("flags", C.c_int32),
        ("n_extra", C.c_int32), ("a_extra", C.POINTER(lk_tensor) * 2), ("dst_extra", C.POINTER(lk_tensor) * 2),
        ("silu_dst", C.POINTER(lk_tensor)), ("mul_dst", C.POINTER(lk_tensor)),
    ]


class lk_matmul_link(C.Structure):
    _fields_ = [("a", C.POINTER(lk_tensor)), ("b", C.POINTER(lk_tensor)), ("dst", C.POINTER(lk_tensor)),
                ("fusion", C.POINTER(lk_matmul_fusion))]


class lk_graph_node(C.Structure):
    _fields_ = [("op", C.c_int32), ("n_src", C.c_int32), ("dst", lk_tensor), ("src", lk_tensor * 4), ("op_params", C.c_int32 * 4)]


class lk_plan_launch(C.Structure):
    _fields_ = [("kind", C.c_int32), ("node", C.c_int32), ("rms", C.c_int32), ("add", C.c_int32), ("n_extra", C.c_int32),
                ("extra", C.c_int32 * 2), ("silu", C.c_int32), ("mul", C.c_int32), ("store_b", C.c_int32), ("chain", C.c_int32),
                ("attn", C.c_int32 * 5)]


class lk_host_io(C.Structure):
    _fields_ = [("direction", C.c_int32), ("reserved", C.c_int32), ("buf", C.c_void_p), ("offset", C.c_uint64),
                ("bytes", C.c_uint64), ("host", C.c_void_p)]


LK_IO_UPLOAD, LK_IO_DOWNLOAD = 0, 1
LK_OP = {"NONE": 0, "ADD": 1, "MUL": 2, "MUL_MAT": 3, "RMS_NORM": 4, "SILU": 5, "SOFT_MAX": 6, "SCALE": 7, "ROPE": 8, "CPY": 9,
         "FLASH_ATTN": 10}
LK_PLAN_NODE, LK_PLAN_FUSED, LK_PLAN_ATTN_STEP = 0, 1, 2
LK_GRAPH_NO_FUSION, LK_GRAPH_CHAIN_ALWAYS, LK_GRAPH_CHAIN_NEVER = 1, 2, 4


class lk_shard_sync(C.Structure):
    _fields_ = [("peer_flags", C.POINTER(C.c_void_p)), ("nranks", C.c_int32), ("rank", C.c_int32),
                ("wait_slot", C.c_int32), ("signal_slot", C.c_int32)]


LK_FLAG_BYTES = 512
LK_MAX_EXCHANGE_SLOTS = 8

_lib = None


def declared_symbols() -> list[str]:
    """Every function the public header declares (used by the CPU-side export test)."""
    with open(HEADER_PATH) as f:
        return sorted(set(re.findall(r"^LK_API[^;(]*?\b(lk_[a-z0-9_]+)\(", f.read(), flags=re.M)))


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise IllegalStateException(
            f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
            "(make -C llama.kotlin_b200/csrc). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, u32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_uint32
    tp = C.POINTER(lk_tensor)
    sig = {
        "lk_abi_version": (i32, []),
        "lk_type_block_bytes": (i32, [i32]),
        "lk_tensor_nbytes": (i64, [tp]),
        "lk_ctx_create": (i32, [i32, C.POINTER(vp)]),
        "lk_ctx_destroy": (i32, [vp]),
        "lk_sync": (i32, [vp]),
        "lk_last_error": (C.c_char_p, [vp]),
        "lk_ctx_stream": (vp, [vp]),
        "lk_ctx_device": (i32, [vp]),
        "lk_ctx_sm_count": (i32, [vp]),
        "lk_ctx_launch_count": (i64, [vp]),
        "lk_timer_start": (i32, [vp]),
        "lk_timer_stop": (i32, [vp, C.POINTER(C.c_float)]),
        "lk_buf_alloc": (i32, [vp, u64, C.POINTER(vp)]),
        "lk_buf_free": (i32, [vp]),
        "lk_buf_size": (u64, [vp]),
        "lk_buf_device_ptr": (vp, [vp]),
        "lk_buf_upload": (i32, [vp, u64, vp, u64]),
        "lk_buf_download": (i32, [vp, u64, vp, u64]),
        "lk_buf_copy": (i32, [vp, u64, vp, u64, u64]),
        "lk_buf_clear": (i32, [vp, C.c_uint8]),
        "lk_host_alloc": (i32, [vp, u64, C.POINTER(vp)]),
        "lk_host_free": (i32, [vp, vp]),
        "lk_compute_matmul": (i32, [vp, tp, tp, tp]),
        "lk_compute_matmul_fused": (i32, [vp, tp, tp, tp, C.POINTER(lk_matmul_fusion)]),
        "lk_compute_matmul_chain": (i32, [vp, C.POINTER(lk_matmul_link), i32]),
        "lk_compute_dequantize": (i32, [vp, tp, tp]),
        "lk_compute_quantize": (i32, [vp, tp, tp]),
        "lk_compute_add": (i32, [vp, tp, tp, tp]),
        "lk_compute_mul": (i32, [vp, tp, tp, tp]),
        "lk_compute_rms_norm": (i32, [vp, tp, tp, C.c_float, tp]),
        "lk_compute_silu": (i32, [vp, tp, tp]),
        "lk_compute_silu_mul": (i32, [vp, tp, tp, tp]),
        "lk_compute_softmax": (i32, [vp, tp, tp]),
        "lk_compute_scale": (i32, [vp, tp, C.c_float, tp]),
        "lk_rope_table": (i32, [i32, C.c_float, i32, vp]),
        "lk_compute_rope": (i32, [vp, tp, tp, tp, i32, tp]),
        "lk_compute_kv_append": (i32, [vp, tp, tp, tp]),
        "lk_compute_attention_decode": (i32, [vp, tp, tp, tp, tp, C.c_float, tp]),
        "lk_compute_attention_decode_noalloc": (i32, [vp, tp, tp, tp, tp, C.c_float, tp]),
        "lk_compute_attention_step": (i32, [vp, tp, tp, tp, tp, tp, i32, tp, tp, C.c_float, tp, tp, tp]),
        "lk_capture_begin": (i32, [vp]),
        "lk_capture_end": (i32, [vp, C.POINTER(vp)]),
        "lk_graph_launch": (i32, [vp, vp]),
        "lk_graph_destroy": (i32, [vp]),
        "lk_graph_plan": (i32, [C.POINTER(lk_graph_node), i32, i32, C.POINTER(lk_plan_launch), i32, C.POINTER(i32)]),
        "lk_graph_compute": (i32, [vp, C.POINTER(lk_graph_node), i32, i32]),
        "lk_graph_compute_io": (i32, [vp, C.POINTER(lk_graph_node), i32, i32, C.POINTER(lk_host_io), i32]),
        "lk_buf_upload_async": (i32, [vp, u64, vp, u64]),
        "lk_buf_download_async": (i32, [vp, u64, vp, u64]),
        "lk_set_option": (i32, [vp, C.c_char_p, i64]),
        "lk_get_option": (i64, [vp, C.c_char_p]),
        "lk_comm_unique_id": (i32, [vp]),
        "lk_comm_init": (i32, [vp, i32, i32, vp]),
        "lk_comm_destroy": (i32, [vp]),
        "lk_allgather_f32": (i32, [vp, vp, u64, i64]),
        "lk_buf_ipc_export": (i32, [vp, vp]),
        "lk_buf_ipc_import": (i32, [vp, vp, u64, C.POINTER(vp)]),
        "lk_peer_barrier": (i32, [vp, vp, C.POINTER(vp), i32, i32, u32]),
        "lk_peer_push_f32": (i32, [vp, vp, u64, i64, C.POINTER(vp), i32]),
        "lk_compute_matmul_sharded": (i32, [vp, tp, tp, tp, i64, i64, C.POINTER(vp), i32]),
        "lk_compute_swiglu_sharded": (i32, [vp, tp, tp, tp, tp, C.c_float, tp, tp, tp, tp, tp, i64, i64, C.POINTER(vp), i32]),
        "lk_compute_matmul_sharded_sync": (i32, [vp, tp, tp, tp, i64, i64, C.POINTER(vp), i32, C.POINTER(lk_shard_sync)]),
        "lk_compute_swiglu_sharded_sync": (i32, [vp, tp, tp, tp, tp, C.c_float, tp, tp, tp, tp, tp, i64, i64, C.POINTER(vp), i32,
                                                 C.POINTER(lk_shard_sync)]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(L, name) and os.environ.get("LK_LIBRARY"):
            continue
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def check(status: int, ctx=None) -> None:
    """Map a status code to the exception class the reference throws (header comment)."""
    if status == LK_OK:
        return
    msg = lib().lk_last_error(ctx)
    msg = msg.decode("utf-8", "replace") if msg else f"status {status}"
    if status in (LK_EINVAL_DIM, LK_EINVAL_TYPE, LK_EINVAL_ARG):
        raise IllegalArgumentException(msg)
    if status == LK_EBOUNDS:
        raise IndexOutOfBoundsException(msg)
    if status == LK_ENOTIMPL:
        raise NotImplementedError(msg)
    raise IllegalStateException(msg)
