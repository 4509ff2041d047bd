"""ctypes binding of ``libb200ddl.so`` (C ABI declared in ``include/b200ddl.h``).

The library is built in-tree by :mod:`.build` and loaded from this directory, never from
site-packages, so the driver can see that the native code is what ran.  On a box with a
GPU a missing library is a hard error (no silent eager fallback); on the CPU-only
authoring box the library still loads (cudart is linked statically) but nothing is
launched.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# DDL_LIB_PATH: an alternative build of the library (A/B timing of kernel variants in one gpurun call; benchmarks only)
LIB_PATH = os.environ.get("DDL_LIB_PATH") or os.path.join(_HERE, "libb200ddl.so")

DDL_F32, DDL_BF16 = 0, 1
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = 0, 1, 2, 3
ACT_NONE, ACT_GELU, ACT_RELU = 0, 1, 2
COMM_MAX_BLOCKS = 320

_lib: Optional[C.CDLL] = None
launch_count = 0  # number of native kernel-launching calls issued from Python (bench "gpu_launches")


class DdlError(RuntimeError):
    pass


class CommT(C.Structure):
    _fields_ = [("signal_pads", C.c_void_p * 8), ("epoch", C.c_void_p), ("rank", C.c_int), ("world", C.c_int)]


_vp, _sz, _i, _f, _ll, _u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_float, C.c_longlong, C.c_ulonglong
_PROTOS = {
    "ddl_version": ([], _i),
    "ddl_device_info": ([_i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_sz)], _i),
    "ddl_symm_alloc": ([_sz, C.POINTER(_vp)], _i),
    "ddl_symm_free": ([_vp], _i),
    "ddl_ipc_export": ([_vp, C.c_char_p], _i),
    "ddl_ipc_open": ([C.c_char_p, C.POINTER(_vp)], _i),
    "ddl_ipc_close": ([_vp], _i),
    "ddl_can_access_peer": ([_i, _i, C.POINTER(_i)], _i),
    "ddl_stream_sync": ([_vp], _i),
    "ddl_comm_signal_pad_bytes": ([], _sz),
    "ddl_comm_epoch_bytes": ([], _sz),
    "ddl_allreduce_oneshot_max_bytes": ([_i, _i], _sz),
    "ddl_allreduce": ([C.POINTER(CommT), C.POINTER(_vp), C.POINTER(_vp), _sz, _i, _i, _f, _i, _vp, _vp, _i, _vp], _i),
    "ddl_allreduce_hybrid": ([C.POINTER(CommT), C.POINTER(_vp), _sz, _f, _vp, _i, _i, _i, _vp], _i),
    "ddl_allgather": ([C.POINTER(CommT), C.POINTER(_vp), _vp, _sz, _i, _vp], _i),
    "ddl_reducescatter": ([C.POINTER(CommT), C.POINTER(_vp), _vp, _sz, _i, _i, _f, _i, _vp], _i),
    "ddl_reducescatter_adamw": ([C.POINTER(CommT), C.POINTER(_vp), _sz, _i, _f, _vp, _vp, _vp, _vp, _vp, _i, _vp], _i),
    "ddl_broadcast": ([C.POINTER(CommT), C.POINTER(_vp), _i, _sz, _i, _vp], _i),
    "ddl_barrier": ([C.POINTER(CommT), _vp], _i),
    "ddl_adamw": ([_vp, _i, _vp, _vp, _vp, _vp, _i, _sz, _vp, _f, _vp], _i),
    "ddl_adamw_tick": ([_vp, _vp], _i),
    "ddl_dropout_effective_p": ([_f], _f),
    "ddl_add_layernorm_fwd": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _f, _f, _u64, _vp, _vp], _i),
    "ddl_layernorm_bwd_scratch_bytes": ([_i, _i], _sz),
    "ddl_layernorm_bwd": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _i, _i, _f, _u64, _vp, _vp], _i),
    "ddl_layernorm_bwd_colsum_scratch_bytes": ([_i, _i], _sz),
    "ddl_layernorm_bwd_colsum": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _i, _i, _f, _u64, _vp, _vp], _i),
    "ddl_embed_layernorm_fwd": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _u64, _vp, _vp], _i),
    "ddl_dropout": ([_vp, _vp, _sz, _f, _u64, _vp, _vp], _i),
    "ddl_cross_entropy": ([_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _f, _ll, _vp], _i),
    "ddl_bn_scratch_bytes": ([_ll, _i], _sz),
    "ddl_bn_apply": ([_vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _vp], _i),
    "ddl_bn_act_fwd_train": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _ll, _i, _f, _f, _i, _vp], _i),
    "ddl_bn_act_fwd_eval": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _f, _i, _vp], _i),
    "ddl_bn_act_bwd": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _ll, _i, _i, _vp], _i),
    "ddl_bn_act_fwd_train_mask": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _ll, _i, _f, _f, _i, _vp], _i),
    "ddl_bn_act_bwd_ex": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _sz, _ll, _i, _i, _vp], _i),
    "ddl_bn_relu_maxpool_supported": ([_i, _i, _i, _i], _i),
    "ddl_bn_relu_maxpool_bwd_scratch_bytes": ([_i, _i, _i, _i], _sz),
    "ddl_bn_relu_maxpool_fwd_train": ([_vp] * 12 + [_i] * 5 + [_f, _f, _vp], _i),
    "ddl_bn_relu_maxpool_bwd": ([_vp] * 12 + [_i, _vp, _sz, _i, _i, _i, _i, _vp], _i),
    "ddl_maxpool3x3s2_bwd_ex": ([_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp], _i),
    "ddl_bn_act_fwd_train_from_partials": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _ll, _i, _f, _f, _i, _vp], _i),
    "ddl_stem_z_bytes": ([_i, _i, _i], _sz),
    "ddl_stem_fwd_grid": ([_i, _i, _i], _i),
    "ddl_stem_wgrad_grid": ([_i, _i, _i], _i),
    "ddl_stem_wgrad_scratch_bytes": ([_i, _i, _i], _sz),
    "ddl_stem_s2d": ([_vp, _i, _vp, _i, _i, _i, _vp], _i),
    "ddl_stem_conv_fwd": ([_vp, _vp, _ll, _ll, _ll, _ll, _vp, _vp, _vp, _i, _i, _i, _vp], _i),
    "ddl_stem_conv_wgrad": ([_vp, _vp, _vp, _ll, _ll, _ll, _ll, _vp, _sz, _i, _i, _i, _vp], _i),
    "ddl_conv1x1_supported": ([_ll, _i, _i], _i),
    "ddl_conv1x1_grid": ([_ll, _i], _i),
    "ddl_conv1x1_fwd": ([_vp, _vp, _vp, _vp, _ll, _i, _i, _vp], _i),
    "ddl_conv1x1_dgrad": ([_vp, _vp, _vp, _ll, _i, _i, _vp], _i),
    "ddl_conv_supported": ([_i] * 9, _i),
    "ddl_conv_fwd_groups": ([_i] * 9, _i),
    "ddl_conv_fwd": ([_vp, _vp, _vp, _vp] + [_i] * 9 + [_vp], _i),
    "ddl_conv_dgrad": ([_vp, _vp, _vp] + [_i] * 9 + [_vp], _i),
    "ddl_conv_wgrad_scratch_bytes": ([_i] * 9, _sz),
    "ddl_conv_wgrad": ([_vp, _vp, _vp, _ll, _ll, _ll, _ll, _vp, _sz] + [_i] * 9 + [_vp], _i),
    "ddl_conv_set_tuning": ([_i, _i], _i),
    "ddl_conv_set_debug": ([_vp, _i], _i),
    "ddl_conv_set_im2col": ([_i], _i),
    "ddl_conv_set_halo": ([_i, _i, _i], _i),
    "ddl_conv_set_stay": ([_i], _i),
    "ddl_probe_umma_shift": ([_vp, _vp, _vp, _i, _i, _i, _vp], _i),
    "ddl_probe_umma_rate": ([_vp, _i, _i, _i, _i, _i, _vp], _i),
    "ddl_probe_tmem_drain": ([_vp, _i, _i, _i, _i, _vp], _i),
    "ddl_embed_bwd_scratch_bytes": ([_i, _i], _sz),
    "ddl_embed_bwd": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i, _i, _i, _i, _i, _i, _vp], _i),
    "ddl_attention_fwd": ([_vp, _vp, _vp, _vp, _i, _i, _i, _f, _u64, _vp, _vp], _i),
    "ddl_attention_bwd_scratch_bytes": ([_i, _i, _i], _sz),
    "ddl_attention_bwd": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i, _i, _i, _f, _u64, _vp, _vp], _i),
    "ddl_maxpool3x3s2_fwd": ([_vp, _vp, _vp, _i, _i, _i, _i, _vp], _i),
    "ddl_maxpool3x3s2_bwd": ([_vp, _vp, _vp, _i, _i, _i, _i, _vp], _i),
    "ddl_avgpool_fwd": ([_vp, _vp, _i, _i, _i, _vp], _i),
    "ddl_avgpool_bwd": ([_vp, _vp, _i, _i, _i, _vp], _i),
    "ddl_nchw_to_nhwc8_bf16": ([_vp, _i, _vp, _i, _i, _ll, _vp], _i),
    "ddl_gemm_bf16": ([_vp, _i, _ll, _vp, _i, _ll, _vp, _ll, _i, _i, _i, _vp, _vp, _vp, _i, _i, _vp, _i, _vp], _i),
    "ddl_gemm_bf16_ex": ([_vp, _i, _ll, _vp, _i, _ll, _vp, _ll, _i, _i, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _i, _vp], _i),
    "ddl_bn_act_fwd_train_from_sums": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _f, _f, _i, _vp], _i),
    "ddl_allgather_gemm_bf16": ([C.POINTER(CommT), C.POINTER(_vp), _vp, _vp, _vp, _ll, _vp, _ll, _i, _i, _i, _vp, _vp, _vp, _i, _vp], _i),
    "ddl_gemm_set_cluster": ([_i], _i),
    "ddl_f32_to_bf16": ([_vp, _vp, _sz, _i, _vp], _i),
    "ddl_dgelu_mul": ([_vp, _vp, _vp, _sz, _vp], _i),
    "ddl_colsum_scratch_bytes": ([_i, _i], _sz),
    "ddl_colsum_bf16": ([_vp, _i, _i, _ll, _vp, _vp, _sz, _vp], _i),
}
# entry points added by later translation units register themselves here
OPTIONAL_PROTOS = {}


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> C.CDLL:
    """Load (once) and return the native library; raise loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DdlError(
            f"{LIB_PATH} is missing: build it with `python -m databricks_distributed_deep_learning_b200.ops.build` "
            "(or __graft_entry__.build()).  There is no eager fallback on the GPU path."
        )
    l = C.CDLL(LIB_PATH)
    l.ddl_last_error.restype = C.c_char_p
    l.ddl_last_error.argtypes = []
    for name, (args, res) in {**_PROTOS, **OPTIONAL_PROTOS}.items():
        try:
            fn = getattr(l, name)
        except AttributeError:
            if name in _PROTOS:
                raise DdlError(f"libb200ddl.so is stale: symbol {name} missing; rebuild") from None
            continue
        fn.argtypes = args
        fn.restype = res
    _lib = l
    return l


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().ddl_last_error().decode(errors="replace")
        raise DdlError(f"{what or 'libb200ddl'} failed (code {rc}): {msg}")


def call(name: str, *args) -> None:
    """Invoke a kernel-launching entry point and count it."""
    global launch_count
    launch_count += 1
    check(getattr(lib(), name)(*args), name)


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return DDL_F32
    if dt == torch.bfloat16:
        return DDL_BF16
    raise DdlError(f"unsupported dtype {dt}; the native kernels take float32 and bfloat16")
