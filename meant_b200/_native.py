"""ctypes binding of libmeant_b200.so (the C ABI declared in include/meant_b200.h).

No fallback: if the library is missing or a call returns non-zero, this raises.  Tensors
stay torch tensors; only raw device pointers, sizes and the current CUDA stream cross the
boundary.
"""
import ctypes as C
import os

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_lib", "libmeant_b200.so")
F32, BF16 = 0, 1
IMPL_AUTO, IMPL_EXACT, IMPL_TCGEN05 = 0, 1, 2
_DT = {torch.float32: F32, torch.bfloat16: BF16}

i64, i32, f32, u64, vp = C.c_int64, C.c_int, C.c_float, C.c_uint64, C.c_void_p


class HeadView(C.Structure):
    _fields_ = [("p", vp), ("sn1", i64), ("sn2", i64), ("ss", i64), ("sh", i64)]


class AttnArgs(C.Structure):
    _fields_ = [("q", HeadView), ("k", HeadView), ("v", HeadView), ("o", HeadView),
                ("dq", HeadView), ("dk", HeadView), ("dv", HeadView), ("out_saved", HeadView),
                ("lse", vp), ("delta", vp), ("qcos", vp), ("qsin", vp), ("kcos", vp), ("ksin", vp),
                ("key_mask", vp),
                ("N", i32), ("n2_count", i32), ("H", i32), ("Sq", i32), ("Sk", i32), ("Dh", i32),
                ("rot_dim", i32), ("pos_is_head", i32), ("causal", i32), ("scale", f32),
                ("q_rot", HeadView), ("k_rot", HeadView),
                ("kx", HeadView), ("vx", HeadView), ("dkx", vp), ("dvx", vp)]


OPT_CHUNK, OPT_MAX_TENSORS, OPT_MAX_BLOCKS = 65536, 48, 320


class TensorTable(C.Structure):
    _fields_ = [("p", vp * OPT_MAX_TENSORS), ("g", vp * OPT_MAX_TENSORS), ("m", vp * OPT_MAX_TENSORS),
                ("v", vp * OPT_MAX_TENSORS), ("shadow", vp * OPT_MAX_TENSORS), ("numel", i64 * OPT_MAX_TENSORS),
                ("block_chunk", C.c_int32 * OPT_MAX_BLOCKS), ("block_tensor", C.c_uint8 * OPT_MAX_BLOCKS),
                ("n_tensors", C.c_int32), ("n_blocks", C.c_int32), ("g_bf16", C.c_int32), ("reserved_", C.c_int32)]


# name -> (restype, argtypes); mirrors include/meant_b200.h one to one
SIGNATURES = {
    "meant_version": (i32, []),
    "meant_launch_count": (i64, []),
    "meant_reset_launch_count": (None, []),
    "meant_rmsnorm_fwd": (i32, [vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, f32, u64, vp, i32, vp]),
    "meant_rmsnorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, f32, u64, vp, vp, i32, vp]),
    "meant_bias_grad": (i32, [vp, vp, i64, i32, i64, i32, vp]),
    "meant_gemm_bf16_tn": (i32, [vp, vp, vp, vp, i64, i32, i32, i64, i64, i64, vp]),
    "meant_gemm_bf16": (i32, [vp, vp, vp, i64, i32, i32, i64, i64, i64, i32, i32, vp, vp, i64, vp]),
    "meant_gemm_bf16_wgrad": (i32, [vp, vp, vp, vp, i64, i32, i32, i64, i64, i64, vp]),
    "meant_embedding_fwd": (i32, [vp, vp, vp, i64, i32, i64, i32, vp]),
    "meant_embedding_bwd": (i32, [vp, vp, vp, i64, i32, i64, i32, vp]),
    "meant_patchify": (i32, [vp, vp, i64, i32, i32, i32, i32, i32, i32, i32, vp]),
    "meant_embed_ln_fwd": (i32, [vp, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, i64, i32, i32, i64, i32, i32, f32, u64, vp, i32, vp]),
    "meant_embed_ln_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, i64, i32, f32, u64, vp,
                                 i32, vp]),
    "meant_layernorm_fwd": (i32, [vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, vp]),
    "meant_layernorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp]),
    "meant_attn_fwd": (i32, [C.POINTER(AttnArgs), i32, i32, vp]),
    "meant_attn_bwd": (i32, [C.POINTER(AttnArgs), i32, i32, vp]),
    "meant_attn_tcgen05_supported": (i32, [C.POINTER(AttnArgs), i32, i32]),
    "meant_attn_prerotate": (i32, [C.POINTER(AttnArgs), i32, vp]),
    "meant_attn_tma_fallbacks": (i64, []),
    "meant_attn_q1_fwd": (i32, [C.POINTER(AttnArgs), vp, i32, vp]),
    "meant_attn_q1_bwd": (i32, [C.POINTER(AttnArgs), vp, i32, i32, vp]),
    "meant_attn_q1_fwd_static": (i32, [C.POINTER(AttnArgs), i32, vp]),
    "meant_meanpool_fwd": (i32, [vp, vp, i64, i32, i32, i32, vp]),
    "meant_meanpool_bwd": (i32, [vp, vp, i64, i32, i32, i32, vp]),
    "meant_geglu_fwd": (i32, [vp, vp, i64, i32, i32, vp]),
    "meant_geglu_bwd": (i32, [vp, vp, vp, i64, i32, i32, vp]),
    "meant_token_shift": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "meant_multi_sqnorm": (i32, [C.POINTER(TensorTable), vp, vp]),
    "meant_adamw_step": (i32, [C.POINTER(TensorTable), vp, vp, f32, f32, f32, f32, f32, f32, vp]),
    "meant_prob_ce": (i32, [vp, vp, vp, vp, i32, i32, f32, vp]),
    "meant_softmax_pool_fwd": (i32, [vp, vp, vp, vp, f32, vp, f32, vp, vp, i64, i32, i32, i32, i32, vp]),
    "meant_softmax_pool_bwd": (i32, [vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, i32, i32, vp]),
    "meant_token_ce_fwd": (i32, [vp, vp, vp, vp, vp, i64, i64, i64, i64, i32, vp]),
    "meant_token_ce_bwd": (i32, [vp, vp, vp, vp, vp, vp, i64, i64, i64, i64, i32, vp]),
    "meant_tc_probe": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    "meant_tc_probe_sw": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
}

_lib = None


ABI_VERSION = 5   # include/meant_b200.h: meant_version()


def lib():
    """Load the shared library (once).  Raises if it has not been built or is stale."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(
                f"{_LIB_PATH} is missing: run `python -m meant_b200.build` (nvcc, sm_100a). "
                "meant_b200 has no CPU or eager fallback.")
        handle = C.CDLL(_LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        if handle.meant_version() != ABI_VERSION:
            raise RuntimeError(f"{_LIB_PATH} has ABI version {handle.meant_version()}, this package expects "
                               f"{ABI_VERSION}: rebuild with `python -m meant_b200.build`")
        _lib = handle
    return _lib


def lib_path():
    return _LIB_PATH


def dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise ValueError(f"meant_b200 kernels take float32 or bfloat16 tensors, got {t.dtype}") from None


def ptr(t):
    return None if t is None else t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def check(code: int, what: str):
    if code != 0:
        if code > 0:
            raise RuntimeError(f"{what}: CUDA error {code} (cudaError_t)")
        raise RuntimeError(f"{what}: {'bad argument' if code == -1 else 'unsupported dtype/shape'} (code {code})")


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("meant_b200 runs on CUDA (sm_100a) tensors only; there is no CPU path")


def launch_count() -> int:
    return int(lib().meant_launch_count())


def reset_launch_count():
    lib().meant_reset_launch_count()


# ---- optional per-kernel CUDA-event timing (bench.py): events are recorded on the launching
# stream around the C call, so the measured interval is that kernel's launch(es) only.
_timers = None


def enable_timers(on=True):
    global _timers
    _timers = {} if on else None


def timers():
    return _timers


def timed(name, fn, *args):
    if _timers is None:
        return fn(*args)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = fn(*args)
    e1.record()
    _timers.setdefault(name, []).append((e0, e1))
    return r
