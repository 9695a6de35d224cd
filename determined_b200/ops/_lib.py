"""ctypes binding of ``libdetb200.so`` (C ABI declared in ``include/detb200.h``).

The library is built in-tree by :mod:`determined_b200._native.build`.  If CUDA is available and
the library is missing we fail loudly — a silent PyTorch fallback on a GPU box would hide that the
native path is not the one running.
"""
import ctypes
import os
import pathlib
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int64, c_void_p
from typing import Any, Optional

_HERE = pathlib.Path(__file__).resolve().parent
LIB_PATH = _HERE.parent / "_native" / "libdetb200.so"
_lib: Optional[ctypes.CDLL] = None

ABI_VERSION = 1


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return LIB_PATH.exists()


def _sig(lib: ctypes.CDLL, name: str, restype: Any, argtypes: Any) -> None:
    fn = getattr(lib, name, None)
    if fn is None:
        return
    fn.restype = restype
    fn.argtypes = argtypes


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        if os.environ.get("DETB200_AUTOBUILD", "1") == "1":
            from determined_b200._native import build as nb

            nb.build()
        if not LIB_PATH.exists():
            raise NativeError(f"{LIB_PATH} is missing: run `python -m determined_b200._native.build`")
    lib = ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
    lib.detb200_last_error.restype = c_char_p
    lib.detb200_abi_version.restype = c_int
    if lib.detb200_abi_version() != ABI_VERSION:
        raise NativeError("libdetb200.so ABI mismatch; rebuild")
    p, i, i64, f, d = c_void_p, c_int, c_int64, c_float, c_double
    _sig(lib, "detb200_optim_tick", i, [p, p, p])
    _sig(lib, "detb200_adam_flat", i, [p, p, p, p, p, i64, p, p, p, p, f, f, f, f, i, f, i, i, p])
    _sig(lib, "detb200_sgd_flat", i, [p, p, p, i64, p, p, p, p, f, f, f, i, f, i, i, p])
    _sig(lib, "detb200_adadelta_flat", i, [p, p, p, p, i64, p, p, p, f, f, f, f, i, i, p])
    _sig(lib, "detb200_sumsq_scratch_bytes", i64, [])
    _sig(lib, "detb200_sumsq", i, [p, i64, p, p, p, i, p])
    _sig(lib, "detb200_clip_coef", i, [p, i, f, f, p, p, p])
    _sig(lib, "detb200_scale_flat", i, [p, i64, p, f, p])
    _sig(lib, "detb200_ctrl_bytes", i64, [])
    _sig(lib, "detb200_ipc_handle_bytes", i, [])
    _sig(lib, "detb200_symm_alloc", i, [POINTER(c_void_p), i64])
    _sig(lib, "detb200_symm_free", i, [p])
    _sig(lib, "detb200_ipc_export", i, [p, p])
    _sig(lib, "detb200_ipc_open", i, [p, POINTER(c_void_p)])
    _sig(lib, "detb200_ipc_close", i, [p])
    _sig(lib, "detb200_enable_peer_access", i, [i])
    _sig(lib, "detb200_comm_stats_ptr", p, [p])
    _sig(lib, "detb200_comm_read_stats", i, [p, i, POINTER(c_float), p])
    _sig(lib, "detb200_allreduce_fused", i, [POINTER(c_void_p), POINTER(c_void_p), i, i, i64, i64, f, p, p, i, i, d, p])
    _sig(lib, "detb200_rs_adam_ag", i, [POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p), i, i, i, i64, i64,
                                        p, p, p, p, f, p, p, p, p, f, f, f, f, i, i, i, d, p])
    PV = POINTER(c_void_p)
    _sig(lib, "detb200_allreduce_fused_ex", i, [PV, PV, p, i, i, i64, i64, f, p, p, i, i, i, d, p])
    _sig(lib, "detb200_allreduce_fused_emulated", i, [PV, PV, i, i64, i64, f, p, p, i, i, d, p])
    _sig(lib, "detb200_narrow_bf16", i, [p, p, i64, p])
    _sig(lib, "detb200_widen_bf16", i, [p, p, i64, p])
    _sig(lib, "detb200_allreduce_bf16", i, [PV, PV, p, i, i, i64, i64, f, p, p, i, i, i, d, p])
    _sig(lib, "detb200_rs_adam_ag_ex", i, [PV, PV, PV, p, p, i, i, i, i64, i64, i64, p, p, p, p, f, p, p, p, p, f, f, f, f,
                                           i, i, i, i, i, d, p])
    # VMM symmetric memory + multicast objects (csrc/symm.cpp)
    _sig(lib, "detb200_vmm_query", i, [POINTER(c_int)])
    _sig(lib, "detb200_vmm_granularity", i, [i, POINTER(c_int64)])
    _sig(lib, "detb200_vmm_alloc", i, [i64, PV, POINTER(c_int)])
    _sig(lib, "detb200_vmm_import", i, [i, i64, PV])
    _sig(lib, "detb200_vmm_free", i, [p])
    _sig(lib, "detb200_mc_create", i, [i, i64, POINTER(c_int64), POINTER(c_int)])
    _sig(lib, "detb200_mc_import", i, [i, POINTER(c_int64)])
    _sig(lib, "detb200_mc_add_device", i, [i64])
    _sig(lib, "detb200_mc_bind", i, [i64, p, i64])
    _sig(lib, "detb200_mc_map", i, [i64, i64, PV])
    _sig(lib, "detb200_mc_free", i, [i64, p, i64])
    _sig(lib, "detb200_close_fd", i, [i])
    _sig(lib, "detb200_broadcast", i, [POINTER(c_void_p), POINTER(c_void_p), i, i, i, i64, i64, i, d, p])
    _sig(lib, "detb200_norm_fwd", i, [p, p, p, p, p, p, i64, i, f, i, i, p])
    _sig(lib, "detb200_norm_bwd_workspace_elems", i64, [i64, i, i])
    _sig(lib, "detb200_norm_bwd", i, [p, p, p, p, p, p, p, p, p, i64, i, i, i, p])
    _sig(lib, "detb200_xent_fwd", i, [p, p, p, p, i64, i, i64, i, p])
    _sig(lib, "detb200_xent_bwd", i, [p, p, p, p, p, i64, i, i64, i, p])
    _sig(lib, "detb200_bn_workspace_bytes", i64, [i])
    _sig(lib, "detb200_bn_fwd_train", i, [p, p, i64, i, p, p, p, p, p, p, p, p, p, f, f, i, p, i, p])
    _sig(lib, "detb200_bn_apply", i, [p, p, i64, i, p, p, i, i, p])
    _sig(lib, "detb200_bn_fwd_train_fused", i, [p, p, p, i64, i, p, p, p, p, p, p, p, p, p, f, f, i, p, i, p])
    _sig(lib, "detb200_bn_bwd_fused", i, [p, p, p, p, p, i64, i, p, p, p, p, p, p, p, i, p])
    _sig(lib, "detb200_bn_bwd", i, [p, p, p, i64, i, p, p, p, p, p, p, p, i, p])
    # tcgen05 GEMM family (csrc/gemm_sm100.cu)
    _sig(lib, "detb200_gemm_bf16", i, [p, p, p, p, p, i64, i64, i64, i64, i64, i64, i64, i, i, i, i, i, i, p])
    _sig(lib, "detb200_gemm_ctrl_bytes", i64, [])
    _sig(lib, "detb200_gemm_rs_arrivals", i, [i64, i64, i])
    _sig(lib, "detb200_ag_gemm_bf16", i, [POINTER(c_void_p), POINTER(c_void_p), i, i, p, p, p, p, i64, i64, i64, i64, i64, i64, i64,
                                         i, i, i, i, ctypes.c_uint32, ctypes.c_uint32, i, p])
    _sig(lib, "detb200_ag_gemm_bf16_ex", i, [POINTER(c_void_p), POINTER(c_void_p), i, i, p, p, p, p, i64, i64, i64, i64, i64, i64, i64,
                                            i, i, i, i, i, i, ctypes.c_uint32, ctypes.c_uint32, i, p])
    _sig(lib, "detb200_gemm_rs_bf16", i, [p, p, POINTER(c_void_p), POINTER(c_void_p), i, i, p, p, i64, i64, i64, i64, i64, i64,
                                         i, i, i, i, ctypes.c_uint32, ctypes.c_uint32, p])
    # attention (csrc/attention.cu)
    _sig(lib, "detb200_attn_debug_delay_v", i, [i])
    _sig(lib, "detb200_attn_fwd", i, [p, p, p, p, p, i, i, i, i, i, i, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64), POINTER(c_int64), f, i, p])
    P64 = POINTER(c_int64)
    _sig(lib, "detb200_attn_bwd", i, [p, p, p, p, p, p, p, p, p, p, p, i, i, i, i, P64, P64, P64, P64, P64, P64, P64, P64, f, i, p])
    _sig(lib, "detb200_rotary_qkv", i, [p, p, p, i64, i, i, i, i, i64, i64, i64, i64, i, p])
    _sig(lib, "detb200_colsum_bf16", i, [p, p, i64, i, i64, i, p])
    _sig(lib, "detb200_cast_bf16", i, [p, p, i64, p])
    _sig(lib, "detb200_add_bf16", i, [p, p, p, p, i64, p])
    _sig(lib, "detb200_accum_multi", i, [POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int64), ctypes.c_char_p, i, p])
    _lib = lib
    return lib


# kernels launched per native entry point (for the launch accounting bench.py reports); default 1
KERNELS_PER_CALL = {"bn_fwd_train": 2, "bn_bwd": 2, "norm_bwd": 3, "attn_bwd": 3, "optimizer_step": 2, "sumsq": 2}
LAUNCHES = [0]


def count_launches(n: int) -> None:
    LAUNCHES[0] += n


def check(rc: int, what: str = "") -> None:
    LAUNCHES[0] += KERNELS_PER_CALL.get(what, 1)
    if rc != 0:
        lib = load()
        msg = lib.detb200_last_error()
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg.decode() if msg else ''}")


def check0(rc: int, what: str = "") -> None:
    """``check`` for entry points that launch no kernel (set-up calls): nothing is added to the launch count."""
    if rc != 0:
        msg = load().detb200_last_error()
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg.decode() if msg else ''}")


def ptr(t: Any) -> Optional[int]:
    """Device pointer of a tensor (None → NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr(stream: Any = None) -> int:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def dtype_code(dt: Any) -> int:
    import torch

    if dt == torch.float32:
        return 0
    if dt == torch.bfloat16:
        return 1
    raise NativeError(f"unsupported dtype {dt} (float32 and bfloat16 only)")
