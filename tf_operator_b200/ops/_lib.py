"""ctypes binding of libtfop_b200.so (the C ABI in include/tfop_b200.h).

The library is loaded lazily on first use. If it is missing it is built with
nvcc; if that is impossible the import of any op raises — there is no silent
PyTorch fallback on the GPU path.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

_lock = threading.Lock()
_lib = None

F32, BF16 = 0, 1

c_void_p, c_int, c_size_t, c_float = C.c_void_p, C.c_int, C.c_size_t, C.c_float

_SIGS = {
    "tfop_last_error": (C.c_char_p, []),
    "tfop_version": (c_int, []),
    "tfop_comm_handle_bytes": (c_int, []),
    "tfop_comm_create": (c_int, [c_int, c_int, c_size_t, C.POINTER(c_void_p)]),
    "tfop_comm_get_handle": (c_int, [c_void_p, c_void_p]),
    "tfop_comm_open_peers": (c_int, [c_void_p, c_void_p]),
    "tfop_comm_create_emulated": (c_int, [c_int, c_size_t, C.POINTER(c_void_p)]),
    "tfop_comm_buffer": (c_void_p, [c_void_p, c_int]),
    "tfop_comm_buffer_bytes": (c_size_t, [c_void_p]),
    "tfop_comm_rank": (c_int, [c_void_p]),
    "tfop_comm_world": (c_int, [c_void_p]),
    "tfop_comm_check_error": (c_int, [c_void_p, C.POINTER(c_int)]),
    "tfop_comm_destroy": (c_int, [c_void_p]),
    "tfop_comm_quiesce": (c_int, [c_void_p]),
    "tfop_comm_set_timeout_ms": (c_int, [C.c_longlong]),
    "tfop_comm_error_ptr": (c_void_p, [c_void_p]),
    "tfop_allreduce": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_size_t,
                               c_size_t, c_float, c_int, c_int, c_void_p]),
    "tfop_allreduce_emulated": (c_int, [C.POINTER(c_void_p), c_int, C.POINTER(c_void_p), c_int,
                                        C.POINTER(c_void_p), c_int, c_int, c_size_t, c_size_t,
                                        c_float, c_int, c_int, c_void_p]),
    "tfop_allreduce_set_oneshot_max_bytes": (c_int, [c_size_t]),
    "tfop_sgd_momentum": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_size_t,
                                  c_size_t, c_void_p, c_int, c_void_p]),
    "tfop_adamw": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_size_t,
                           c_size_t, c_void_p, c_void_p]),
    "tfop_zero": (c_int, [c_void_p, c_size_t, c_void_p]),
    "tfop_multi_pack_bf16": (c_int, [c_void_p, c_int, c_int, c_void_p, c_float, c_void_p]),
    "tfop_u8_to_bf16_normalize": (c_int, [c_void_p, c_void_p, c_size_t, c_int,
                                          C.POINTER(c_float), C.POINTER(c_float), c_void_p]),
    "tfop_stem_space_to_depth": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int,
                                         C.POINTER(c_float), C.POINTER(c_float), c_void_p]),
    "tfop_bn_workspace_bytes": (c_size_t, []),
    "tfop_bn_fwd_stats": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                  c_void_p, c_float, c_float, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p]),
    "tfop_bn_finalize": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                 c_void_p, c_float, c_float, c_void_p, c_void_p, c_void_p,
                                 c_void_p]),
    "tfop_bn_fwd_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                  c_int, c_int, c_void_p]),
    "tfop_bn_bwd_reduce": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                   c_void_p]),
    "tfop_bn_bwd_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_int,
                                  c_void_p]),
    "tfop_stem_bn_relu_pool_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                           c_int, c_void_p]),
    "tfop_stem_bn_pool_bwd_reduce": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                             c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                             c_void_p, c_void_p]),
    "tfop_stem_bn_pool_bwd_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                            c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                            c_void_p]),
    "tfop_maxpool3x3s2_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                      c_void_p]),
    "tfop_maxpool3x3s2_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                      c_void_p]),
    "tfop_nvls_supported": (c_int, [C.POINTER(c_int)]),
    "tfop_nvls_round_size": (c_int, [c_int, c_size_t, C.POINTER(c_size_t)]),
    "tfop_nvls_create": (c_int, [c_void_p, c_size_t, C.POINTER(c_int)]),
    "tfop_nvls_import": (c_int, [c_void_p, c_size_t, c_int]),
    "tfop_nvls_add_device": (c_int, [c_void_p]),
    "tfop_nvls_bind": (c_int, [c_void_p]),
    "tfop_nvls_local_ptr": (c_void_p, [c_void_p]),
    "tfop_nvls_bytes": (c_size_t, [c_void_p]),
    "tfop_nvls_destroy": (c_int, [c_void_p]),
    "tfop_allreduce_nvls": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_size_t,
                                    c_size_t, c_float, c_int, c_void_p]),
    "tfop_hist_build": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_void_p,
                                c_void_p]),
    "tfop_gemm_allreduce_set_timeout": (c_int, [C.c_longlong]),
    "tfop_gemm_allreduce_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "tfop_gemm_allreduce_bf16_tn": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                            c_int, c_int, c_float, c_size_t, c_size_t, c_int,
                                            c_void_p]),
    "tfop_gemm_allreduce_nvls_bf16_tn": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                                 c_int, c_int, c_int, c_float, c_size_t, c_size_t,
                                                 c_int, c_void_p]),
    "tfop_gemm_allreduce_nvls_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                              c_int, c_int, c_int, c_int, c_int, c_float, c_size_t,
                                              c_size_t, c_int, c_void_p]),
    "tfop_gemm_allreduce_nvls_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "tfop_gemm_allreduce_set_tile_n": (c_int, [c_int]),
    "tfop_gemm_allreduce_set_legacy": (c_int, [c_int]),
    "tfop_gemm_allreduce_emulated": (c_int, [C.POINTER(c_void_p), c_int, C.POINTER(c_void_p),
                                             C.POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int,
                                             c_int, c_float, c_size_t, c_size_t, c_void_p]),
    "tfop_gemm_force_direct_store": (c_int, [c_int]),
    "tfop_gemm_set_cluster_mode": (c_int, [c_int]),
    "tfop_gemm_set_group_m": (c_int, [c_int]),
    "tfop_gemm_set_splits": (c_int, [c_int]),
    "tfop_gemm_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                               c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "tfop_ln_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                            c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "tfop_ln_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                            c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tfop_ln_force_warp_rows": (c_int, [c_int]),
    "tfop_ln_set_bwd_rows": (c_int, [c_int]),
    "tfop_bias_gelu_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "tfop_bias_act_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                  c_int, c_void_p]),
    "tfop_xent_fwd": (c_int, [c_void_p, C.c_long, c_void_p, c_int, c_int, C.c_longlong, c_void_p,
                              c_void_p, c_void_p, c_void_p]),
    "tfop_xent_bwd": (c_int, [c_void_p, C.c_long, c_void_p, c_int, c_int, C.c_longlong, c_void_p,
                              c_void_p, c_void_p, c_void_p, C.c_long, c_void_p]),
    "tfop_conv_stem_s2d_wgrad": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "tfop_conv_stem_set_mode": (c_int, [c_int]),
    "tfop_conv_stem_s2d_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "tfop_conv_fwd": (c_int, [c_void_p] * 4 + [c_int] * 7 + [c_void_p]),
    "tfop_conv_dgrad": (c_int, [c_void_p] * 3 + [c_int] * 8 + [c_void_p]),
    "tfop_conv_wgrad": (c_int, [c_void_p] * 3 + [c_int] * 7 + [c_void_p]),
    "tfop_conv_set_cluster_mode": (c_int, [c_int]),
    "tfop_attn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "tfop_attn_bwd": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_void_p]),
    "tfop_transpose_bf16": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "tfop_ps_gather": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_int, c_int, c_void_p]),
    "tfop_ps_push_adagrad": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p,
                                     c_int, c_float, c_float, c_float, c_int, c_void_p]),
    "tfop_ps_dedup_workspace": (c_int, [c_size_t, c_int, c_void_p]),
    "tfop_ps_set_push_mode": (c_int, [c_int]),
    "tfop_ps_push_adagrad_dedup": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_int, c_float,
                                           c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_int, c_void_p]),
    "tfop_gemm_bf16_tn": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                  c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
}

# entry points added by later kernels register themselves here
EXTRA_SIGS: dict = {}


class PackSeg(C.Structure):
    _fields_ = [("src", c_void_p), ("dst_off", C.c_ulonglong), ("n", C.c_ulonglong)]


class PsTable(C.Structure):
    _fields_ = [("weight", c_void_p * 8), ("accum", c_void_p * 8), ("num_shards", c_int),
                ("dim", c_int), ("rows_per_shard", C.c_longlong)]


class TfopError(RuntimeError):
    pass


def lib_path() -> Path:
    return Path(__file__).resolve().parent.parent / "lib" / "libtfop_b200.so"


def load():
    """Return the loaded library, building it first if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = lib_path()
        if not path.exists():
            from . import _build
            _build.build()
        lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
        for name, (res, args) in {**_SIGS, **EXTRA_SIGS}.items():
            try:
                fn = getattr(lib, name)
            except AttributeError as e:
                raise TfopError(f"{path} lacks the entry point {name}: the library is older "
                                f"than the Python package; rebuild it with "
                                f"`python -m tf_operator_b200.ops._build --force`") from e
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


# number of kernel launches issued through this binding, by entry point
LAUNCHES: dict = {}


def launches_total() -> int:
    return sum(LAUNCHES.values())


def check(rc: int, what: str = "") -> None:
    if what:
        LAUNCHES[what] = LAUNCHES.get(what, 0) + 1
    if rc != 0:
        msg = load().tfop_last_error()
        raise TfopError(f"{what or 'libtfop_b200'} failed (rc={rc}): "
                        f"{msg.decode() if msg else 'unknown error'}")


def ptr(t) -> int:
    """Device pointer of a torch tensor (or None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def stream_handle(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def dtype_tag(dt) -> int:
    import torch
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {dt}")
