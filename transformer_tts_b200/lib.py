"""ctypes binding of ``csrc/libttsb200.so`` (C ABI in ``include/ttsb200.h``).

Thin, typed wrappers that take torch CUDA tensors, validate dtype / contiguity / device in Python
and pass raw device pointers plus the current CUDA stream.  There is NO fallback: if the library is
missing or a call fails, a RuntimeError is raised with ``ttsb200_last_error_string()``.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

import torch

from . import _build

EPI_AFFINE_ACT, EPI_RES_LN, EPI_FFN2, EPI_RELU_LN, EPI_RELU_LN_DOT = range(5)
ACT_NONE, ACT_RELU, ACT_TANH, ACT_RELU_GRAD = range(4)

ABI_VERSION = 200  # include/ttsb200.h TTSB200_ABI_VERSION: struct layouts / signatures this binding was written for

LAUNCH_COUNT = 0  # kernels launched through this binding (bench.py reports it as gpu_launches)


class GemmArgs(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("T", C.c_int32), ("N", C.c_int32), ("C", C.c_int32),
        ("taps", C.c_int32), ("pad_left", C.c_int32), ("block_n", C.c_int32),
        ("epi", C.c_int32), ("act", C.c_int32),
        ("x", C.c_void_p), ("ldx", C.c_int64), ("w", C.c_void_p),
        ("bias", C.c_void_p), ("scale", C.c_void_p), ("res_f32", C.c_void_p),
        ("res_bstride", C.c_int64), ("res_tstride", C.c_int64),
        ("res_scale", C.c_void_p), ("res_bf16", C.c_void_p),
        ("ln1_g", C.c_void_p), ("ln1_b", C.c_void_p), ("ln2_g", C.c_void_p), ("ln2_b", C.c_void_p),
        ("eps", C.c_float),
        ("dot_w", C.c_void_p), ("dot_b", C.c_void_p), ("row_mask", C.c_void_p),
        ("out_f32", C.c_void_p), ("out_bf16", C.c_void_p), ("out2_f32", C.c_void_p),
        ("out_row", C.c_void_p),
        ("pre_f32", C.c_void_p), ("gate_bf16", C.c_void_p),
        ("tile_map", C.c_void_p), ("tile_count", C.c_void_p),
    ]


class SmallLinearArgs(C.Structure):
    """ttsb200_small_linear_args of include/ttsb200.h (field order and types must match)."""
    _fields_ = [
        ("x_f32", C.c_void_p), ("x_bf16", C.c_void_p), ("ldx", C.c_int64),
        ("ln_g", C.c_void_p), ("ln_b", C.c_void_p), ("eps", C.c_float),
        ("w_bf16", C.c_void_p), ("ldw", C.c_int64), ("bias", C.c_void_p), ("relu", C.c_int32),
        ("res_f32", C.c_void_p), ("ldres", C.c_int64),
        ("add_row", C.c_void_p), ("add_scale", C.c_void_p),
        ("out_f32", C.c_void_p), ("ldo", C.c_int64), ("out_bf16", C.c_void_p), ("ldob", C.c_int64),
        ("xn_bf16", C.c_void_p), ("pos", C.c_void_p),
        ("x_pos_stride", C.c_int64), ("x_pos_bias", C.c_int32), ("out_pos_stride", C.c_int64),
        ("R", C.c_int32), ("K", C.c_int32), ("N", C.c_int32),
    ]


_LIB: Optional[C.CDLL] = None

_SIGNATURES = {
    "ttsb200_version": ([], C.c_int),
    "ttsb200_last_error_string": ([], C.c_char_p),
    "ttsb200_debug_tmap_cache": ([C.c_void_p, C.c_void_p], None),
    "ttsb200_conv_gemm": ([C.POINTER(GemmArgs), C.c_void_p], C.c_int),
    "ttsb200_small_linear": ([C.POINTER(SmallLinearArgs), C.c_void_p], C.c_int),
    "ttsb200_attn_decode": ([C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                             C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                             C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_advance_pos": ([C.c_void_p, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_bce_logits_loss": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_void_p],
                                C.c_int),
    "ttsb200_relu_gate": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_debug_dcl_trace": ([C.c_void_p, C.c_int], C.c_int),
    "ttsb200_decode_attn_out_q": ([C.c_void_p] * 8 + [C.c_float, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_decode_ffn_block": ([C.c_void_p] * 14 + [C.c_float, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                                    C.c_void_p], C.c_int),
    "ttsb200_ffn_tail_rows": ([C.c_void_p] * 7 + [C.c_float, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_ffn_fused": ([C.c_void_p] * 10 + [C.c_float, C.c_void_p, C.c_void_p] + [C.c_int32] * 4 + [C.c_void_p] * 3, C.c_int),
    "ttsb200_build_tile_map": ([C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                C.c_void_p], C.c_int),
    "ttsb200_attention_fwd": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                               C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_attention_fwd_causal": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_attention_bwd_causal": ([C.c_void_p] * 7 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_attention_fwd_ragged": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_attention_probs": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                 C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_mask_lengths": ([C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_embed_pe_ln": ([C.c_void_p] * 6 + [C.c_float, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                                 C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_layernorm": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p,
                           C.c_int64, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_round_durations": ([C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_lr_scan": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_lr_expand": ([C.c_void_p] * 8 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_length_regulate_static": ([C.c_void_p] * 8 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_bucket_embed_add": ([C.c_void_p] * 11 + [C.c_int64, C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_cast_bf16": ([C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p], C.c_int),
    # ---- training step ----
    "ttsb200_attention_fwd_lse": ([C.c_void_p] * 5 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_conv_wgrad": ([C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                            C.c_int64] + [C.c_int32] * 6 + [C.c_void_p], C.c_int),
    "ttsb200_conv_wgrad_bias": ([C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                 C.c_int64] + [C.c_int32] * 6 + [C.c_void_p], C.c_int),
    "ttsb200_weight_transpose": ([C.c_void_p, C.c_int32, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_attention_bwd": ([C.c_void_p] * 7 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_attn_delta": ([C.c_void_p] * 3 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_layernorm_bwd": ([C.c_void_p] * 9 + [C.c_int32] + [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_float,
                                                                                 C.c_void_p], C.c_int),
    "ttsb200_colsum": ([C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_bn_stats": ([C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_bn_tanh_fwd": ([C.c_void_p] * 5 + [C.c_int64, C.c_int32, C.c_float, C.c_uint64, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_bn_tanh_bwd": ([C.c_void_p] * 7 + [C.c_int64, C.c_int32, C.c_float, C.c_int32, C.c_float, C.c_uint64,
                                                C.c_int64, C.c_void_p], C.c_int),
    # ---- dropout (training) ----
    "ttsb200_layernorm_bwd_drop": ([C.c_void_p] * 9 + [C.c_int32] + [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_float,
                                    C.c_float, C.c_uint64, C.c_int64, C.c_int64, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_dropout": ([C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_uint64, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_dropout_mask": ([C.c_void_p, C.c_int64, C.c_float, C.c_uint64, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_attention_fwd_drop": ([C.c_void_p] * 5 + [C.c_int32] * 4 + [C.c_float, C.c_uint64, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_attention_bwd_drop": ([C.c_void_p] * 7 + [C.c_int32] * 4 + [C.c_float, C.c_uint64, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_attention_fwd_causal_drop": ([C.c_void_p] * 5 + [C.c_int32] * 4 + [C.c_float, C.c_uint64, C.c_int64, C.c_void_p],
                                          C.c_int),
    "ttsb200_attention_bwd_causal_drop": ([C.c_void_p] * 7 + [C.c_int32] * 4 + [C.c_float, C.c_uint64, C.c_int64, C.c_void_p],
                                          C.c_int),
    "ttsb200_lr_bwd": ([C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_embed_bwd": ([C.c_void_p] * 4 + [C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                              C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_l1_loss": ([C.c_void_p] * 5 + [C.c_int64, C.c_float, C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_sumsq": ([C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_adam_step": ([C.c_void_p] * 5 + [C.c_int64] + [C.c_float] * 6 + [C.c_int32, C.c_void_p], C.c_int),
    "ttsb200_radam_step": ([C.c_void_p] * 5 + [C.c_int64] + [C.c_float] * 6 + [C.c_int32, C.c_void_p], C.c_int),
    # ---- fp32-accurate encoder -> duration chain ----
    "ttsb200_gemm_x3": ([C.c_void_p, C.c_void_p, C.c_void_p, C.c_float] + [C.c_int32] * 7 + [C.c_void_p] * 3, C.c_int),
    "ttsb200_attention_x3": ([C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p], C.c_int),
    "ttsb200_xrow": ([C.c_int32] + [C.c_void_p] * 12 + [C.c_float] + [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_float,
                     C.c_uint64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p], C.c_int),
    "ttsb200_split_f16": ([C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_void_p], C.c_int),
    "ttsb200_ffn2_tail": ([C.c_void_p] * 7 + [C.c_float] + [C.c_void_p] * 4 + [C.c_int64, C.c_int32, C.c_void_p],
                          C.c_int),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def lib_path() -> Path:
    return _build.LIB_PATH


def load() -> C.CDLL:
    """Load the shared library (building it if the sources are newer).  Raises if impossible."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _build.LIB_PATH
    if _build.is_stale():
        try:
            _build.build()
        except Exception as e:
            # A snapshot copied to the GPU box may carry a fresh .so with older-looking mtimes and no nvcc: that is
            # fine as long as the library reports the ABI this binding expects (checked below); a MISSING library is
            # always an error.
            if not path.exists():
                raise RuntimeError(f"libttsb200.so is missing and cannot be built: {e}") from e
    lib = C.CDLL(str(path))
    lib.ttsb200_version.restype = C.c_int
    got = lib.ttsb200_version()
    if got != ABI_VERSION:
        raise RuntimeError(f"{path} reports ABI version {got}, this binding needs {ABI_VERSION}: the library is stale "
                           "(rebuild with `python -m transformer_tts_b200._build --force`)")
    for name, (argtypes, restype) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: fail loudly
        fn.argtypes = argtypes
        fn.restype = restype
    _LIB = lib
    return lib


def _check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ttsb200_last_error_string().decode()
        raise RuntimeError(f"ttsb200 {what} failed (rc={rc}): {msg}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor], dtype=None, name: str = "tensor") -> Optional[int]:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"ttsb200: {name} must be a CUDA tensor (there is no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"ttsb200: {name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"ttsb200: {name} must be contiguous")
    return t.data_ptr()


def _count(n: int = 1) -> None:
    global LAUNCH_COUNT
    LAUNCH_COUNT += n


# --------------------------------------------------------------------------------------------------
# GEMM / conv
# --------------------------------------------------------------------------------------------------
def conv_gemm(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, *, taps: int = 1, pad_left: int = 0,
              block_n: int, epi: int = EPI_AFFINE_ACT, act: int = ACT_NONE,
              scale: Optional[torch.Tensor] = None,
              res_f32: Optional[torch.Tensor] = None, res_bstride: int = 0, res_tstride: int = 0,
              res_scale: Optional[torch.Tensor] = None, res_bf16: Optional[torch.Tensor] = None,
              ln1: Optional[tuple] = None, ln2: Optional[tuple] = None, eps: float = 1e-5,
              dot_w: Optional[torch.Tensor] = None, dot_b: Optional[torch.Tensor] = None,
              row_mask: Optional[torch.Tensor] = None,
              out_f32: Optional[torch.Tensor] = None, out_bf16: Optional[torch.Tensor] = None,
              out2_f32: Optional[torch.Tensor] = None, out_row: Optional[torch.Tensor] = None,
              pre_f32: Optional[torch.Tensor] = None, gate_bf16: Optional[torch.Tensor] = None,
              tile_map: Optional[tuple] = None) -> None:
    """x: bf16 (B,T,C); w: bf16 (taps,N,C).  See include/ttsb200.h for the epilogue modes."""
    B, T, Cin = x.shape
    tw, N, Cw = w.shape
    if tw != taps or Cw != Cin:
        raise ValueError(f"conv_gemm: weight {tuple(w.shape)} does not match taps={taps}, C={Cin}")
    a = GemmArgs()
    a.B, a.T, a.N, a.C, a.taps, a.pad_left, a.block_n, a.epi, a.act = B, T, N, Cin, taps, pad_left, block_n, epi, act
    a.x = _ptr(x, torch.bfloat16, "x"); a.ldx = Cin
    a.w = _ptr(w, torch.bfloat16, "w")
    a.bias = _ptr(bias, torch.float32, "bias")
    a.scale = _ptr(scale, torch.float32, "scale")
    a.res_f32 = _ptr(res_f32, torch.float32, "res_f32")
    a.res_bstride, a.res_tstride = res_bstride, res_tstride
    a.res_scale = _ptr(res_scale, torch.float32, "res_scale")
    a.res_bf16 = _ptr(res_bf16, torch.bfloat16, "res_bf16")
    if ln1 is not None:
        a.ln1_g, a.ln1_b = _ptr(ln1[0], torch.float32, "ln1.weight"), _ptr(ln1[1], torch.float32, "ln1.bias")
    if ln2 is not None:
        a.ln2_g, a.ln2_b = _ptr(ln2[0], torch.float32, "ln2.weight"), _ptr(ln2[1], torch.float32, "ln2.bias")
    a.eps = eps
    a.dot_w = _ptr(dot_w, torch.float32, "dot_w")
    a.dot_b = _ptr(dot_b, torch.float32, "dot_b")
    a.row_mask = _ptr(row_mask, torch.uint8, "row_mask")
    a.out_f32 = _ptr(out_f32, torch.float32, "out_f32")
    a.out_bf16 = _ptr(out_bf16, torch.bfloat16, "out_bf16")
    a.out2_f32 = _ptr(out2_f32, torch.float32, "out2_f32")
    a.out_row = _ptr(out_row, torch.float32, "out_row")
    a.pre_f32 = _ptr(pre_f32, torch.float32, "pre_f32")
    a.gate_bf16 = _ptr(gate_bf16, torch.bfloat16, "gate_bf16")
    if tile_map is not None:  # (map, count) from build_tile_map: only the listed 128-row tiles are computed
        a.tile_map, a.tile_count = _ptr(tile_map[0], torch.int32, "tile_map"), _ptr(tile_map[1], torch.int32, "tile_count")
    _check(load().ttsb200_conv_gemm(C.byref(a), _stream()), "conv_gemm")
    _count()


def ffn_fused(x: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor, y: torch.Tensor,
              ln1: tuple, ln2: tuple, z: torch.Tensor, zn: torch.Tensor, eps: float = 1e-5,
              tile_map: Optional[tuple] = None) -> None:
    """z = y + LN(relu(x W1^T + b1) W2^T + b2 + x; ln1); zn = LN(z; ln2) in one kernel (k=1 blocks, D=384)."""
    B, T, D = x.shape
    hidden = w1.shape[-2]
    tm, tc = (None, None) if tile_map is None else tile_map
    _check(load().ttsb200_ffn_fused(_ptr(x, torch.bfloat16, "x"), _ptr(w1, torch.bfloat16, "w1"),
                                    _ptr(b1, torch.float32, "b1"), _ptr(w2, torch.bfloat16, "w2"),
                                    _ptr(b2, torch.float32, "b2"), _ptr(y, torch.float32, "y"),
                                    _ptr(ln1[0], torch.float32, "ln1_g"), _ptr(ln1[1], torch.float32, "ln1_b"),
                                    _ptr(ln2[0], torch.float32, "ln2_g"), _ptr(ln2[1], torch.float32, "ln2_b"), eps,
                                    _ptr(z, torch.float32, "z"), _ptr(zn, torch.bfloat16, "zn"), B, T, D, hidden,
                                    _ptr(tm, torch.int32, "tile_map"), _ptr(tc, torch.int32, "tile_count"), _stream()),
           "ffn_fused")
    _count()


def build_tile_map(lengths: torch.Tensor, T: int, halo: int = 0) -> tuple:
    """Row-tile list of a ragged batch for conv_gemm(tile_map=...).  lengths: int32 (B,) or the (B,2) kv_info of
    mask_lengths (column 0 = 1 + last valid row).  Returns (map int32 (B*m_tiles,), count int32 (1,)); no host sync."""
    if lengths.dtype != torch.int32 or not lengths.is_contiguous():
        raise TypeError("build_tile_map: lengths must be a contiguous int32 tensor")
    B = lengths.shape[0]
    stride = 1 if lengths.dim() == 1 else lengths.shape[1]
    m_tiles = (T + 127) // 128
    tmap = torch.empty(B * m_tiles, dtype=torch.int32, device=lengths.device)
    count = torch.empty(1, dtype=torch.int32, device=lengths.device)
    _check(load().ttsb200_build_tile_map(_ptr(lengths, torch.int32, "lengths"), stride, B, m_tiles, halo,
                                         tmap.data_ptr(), count.data_ptr(), _stream()), "build_tile_map")
    _count()
    return tmap, count


# --------------------------------------------------------------------------------------------------
# attention
# --------------------------------------------------------------------------------------------------
def attention_fwd(qkv: torch.Tensor, mask: torch.Tensor, kv_len: torch.Tensor, ctx: torch.Tensor, H: int,
                  lse2: Optional[torch.Tensor] = None, skip_padded_queries: bool = False,
                  drop: Optional[tuple] = None, causal: bool = False) -> None:
    """lse2 (B,H,T) fp32, when given, receives the log2-domain log-sum-exp of every row (training).
    skip_padded_queries: 128-row query tiles wholly beyond an utterance's last valid row are not computed.
    drop = (p, seed, site): dropout on the probabilities (training, Models/modules.py:18-19).
    causal: key j attends to query i only for j <= i (inference; the target mask of Transformer-TTS)."""
    B, T, W = qkv.shape
    Dh = W // (3 * H)
    if causal:
        if skip_padded_queries:
            raise ValueError("attention_fwd: causal has no ragged query skipping variant")
        if drop is not None and drop[0] > 0:
            _check(load().ttsb200_attention_fwd_causal_drop(
                _ptr(qkv, torch.bfloat16, "qkv"), _ptr(mask, torch.uint8, "mask"), _ptr(kv_len, torch.int32, "kv_len"),
                _ptr(ctx, torch.bfloat16, "ctx"), _ptr(lse2, torch.float32, "lse2"), B, T, H, Dh, float(drop[0]), int(drop[1]),
                int(drop[2]), _stream()), "attention_fwd_causal_drop")
            _count()
            return
        _check(load().ttsb200_attention_fwd_causal(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(mask, torch.uint8, "mask"),
                                                   _ptr(kv_len, torch.int32, "kv_len"), _ptr(ctx, torch.bfloat16, "ctx"),
                                                   _ptr(lse2, torch.float32, "lse2"), B, T, H, Dh, _stream()),
               "attention_fwd_causal")
        _count()
        return
    if drop is not None and drop[0] > 0:
        _check(load().ttsb200_attention_fwd_drop(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(mask, torch.uint8, "mask"),
                                                 _ptr(kv_len, torch.int32, "kv_len"), _ptr(ctx, torch.bfloat16, "ctx"),
                                                 _ptr(lse2, torch.float32, "lse2"), B, T, H, Dh, float(drop[0]),
                                                 int(drop[1]), int(drop[2]), _stream()), "attention_fwd_drop")
        _count()
        return
    if skip_padded_queries:
        if lse2 is not None:
            raise ValueError("attention_fwd: skip_padded_queries is an inference option (no lse2)")
        _check(load().ttsb200_attention_fwd_ragged(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(mask, torch.uint8, "mask"),
                                                   _ptr(kv_len, torch.int32, "kv_len"),
                                                   _ptr(ctx, torch.bfloat16, "ctx"), B, T, H, Dh, _stream()),
               "attention_fwd_ragged")
        _count()
        return
    _check(load().ttsb200_attention_fwd_lse(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(mask, torch.uint8, "mask"),
                                            _ptr(kv_len, torch.int32, "kv_len"), _ptr(ctx, torch.bfloat16, "ctx"),
                                            _ptr(lse2, torch.float32, "lse2"), B, T, H, Dh, _stream()),
           "attention_fwd")
    _count()


def attention_probs(qkv: torch.Tensor, mask: torch.Tensor, H: int) -> torch.Tensor:
    B, T, W = qkv.shape
    Dh = W // (3 * H)
    probs = torch.empty((B, H, T, T), device=qkv.device, dtype=torch.float32)
    _check(load().ttsb200_attention_probs(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(mask, torch.uint8, "mask"),
                                          _ptr(probs), B, T, H, Dh, _stream()), "attention_probs")
    _count()
    return probs


# --------------------------------------------------------------------------------------------------
# row kernels
# --------------------------------------------------------------------------------------------------
def mask_lengths(mask: torch.Tensor) -> torch.Tensor:
    B, T = mask.shape
    out = torch.empty((B, 2), device=mask.device, dtype=torch.int32)  # (1 + last valid, clean-prefix length)
    _check(load().ttsb200_mask_lengths(_ptr(mask, torch.uint8, "mask"), _ptr(out), B, T, _stream()), "mask_lengths")
    _count()
    return out


def embed_pe_ln(ids, table, pe, alpha, g, b, eps, x0, xn) -> None:
    B, T = ids.shape
    V, D = table.shape
    if pe.shape[0] < T or pe.shape[1] != D:
        raise ValueError(f"embed_pe_ln: positional table {tuple(pe.shape)} too small for T={T}, D={D}")
    _check(load().ttsb200_embed_pe_ln(_ptr(ids, torch.int64, "ids"), _ptr(table, torch.float32, "table"),
                                      _ptr(pe, torch.float32, "pe"), _ptr(alpha, torch.float32, "alpha"),
                                      _ptr(g, torch.float32, "g"), _ptr(b, torch.float32, "b"), eps,
                                      _ptr(x0, torch.float32, "x0"), _ptr(xn, torch.bfloat16, "xn"),
                                      B, T, D, V, _stream()), "embed_pe_ln")
    _count()


def layernorm(x, g, b, eps, y_bf16=None, y_f32=None) -> None:
    D = x.shape[-1]
    rows = x.numel() // D
    _check(load().ttsb200_layernorm(_ptr(x, torch.float32, "x"), _ptr(g, torch.float32, "g"),
                                    _ptr(b, torch.float32, "b"), eps, _ptr(y_bf16, torch.bfloat16, "y_bf16"),
                                    _ptr(y_f32, torch.float32, "y_f32"), rows, D, _stream()), "layernorm")
    _count()


def ffn2_tail(u, yn, y, ln1, ln2, eps, z, zn, zn32=None, pre_f32=None) -> None:
    """z = y + LN(u + yn; ln1); zn = LN(z; ln2) — the FFN2 epilogue as a row kernel (short batches)."""
    D = u.shape[-1]
    rows = u.numel() // D
    for name, t in (("yn", yn), ("y", y), ("z", z), ("zn", zn), ("zn32", zn32), ("pre_f32", pre_f32)):
        if t is not None and t.numel() != u.numel():
            raise ValueError(f"ffn2_tail: {name} has {t.numel()} elements, expected {u.numel()}")
    _check(load().ttsb200_ffn2_tail(_ptr(u, torch.float32, "u"), _ptr(yn, torch.bfloat16, "yn"),
                                    _ptr(y, torch.float32, "y"), _ptr(ln1[0], torch.float32, "ln1_g"),
                                    _ptr(ln1[1], torch.float32, "ln1_b"), _ptr(ln2[0], torch.float32, "ln2_g"),
                                    _ptr(ln2[1], torch.float32, "ln2_b"), eps, _ptr(z, torch.float32, "z"),
                                    _ptr(zn, torch.bfloat16, "zn"), _ptr(zn32, torch.float32, "zn32"),
                                    _ptr(pre_f32, torch.float32, "pre_f32"), rows, D, _stream()), "ffn2_tail")
    _count()


def round_durations(log_d: torch.Tensor) -> torch.Tensor:
    out = torch.empty(log_d.shape, device=log_d.device, dtype=torch.int32)
    _check(load().ttsb200_round_durations(_ptr(log_d, torch.float32, "log_d"), _ptr(out), log_d.numel(), _stream()),
           "round_durations")
    _count()
    return out


def lr_scan(dur: torch.Tensor):
    B, L = dur.shape
    cum = torch.empty((B, L), device=dur.device, dtype=torch.int32)
    mel_len = torch.empty((B,), device=dur.device, dtype=torch.int32)
    _check(load().ttsb200_lr_scan(_ptr(dur, torch.int32, "dur"), _ptr(cum), _ptr(mel_len), B, L, _stream()), "lr_scan")
    _count()
    return cum, mel_len


def lr_expand(x: torch.Tensor, cum: torch.Tensor, mel_len: torch.Tensor, Tmax: int, want_aux: bool = True,
              out_f32: Optional[torch.Tensor] = None):
    """out_f32 (B,Tmax,D) fp32, when given, receives the expanded rows widened to fp32 as well."""
    B, L, D = x.shape
    dev = x.device
    out = torch.empty((B, Tmax, D), device=dev, dtype=torch.bfloat16)
    frame_idx = torch.empty((B, Tmax), device=dev, dtype=torch.int32) if want_aux else None
    mel_pos = torch.empty((B, Tmax), device=dev, dtype=torch.int64) if want_aux else None
    mel_mask = torch.empty((B, Tmax), device=dev, dtype=torch.uint8)
    _check(load().ttsb200_lr_expand(_ptr(x, torch.bfloat16, "x"), _ptr(cum, torch.int32, "cum"),
                                    _ptr(mel_len, torch.int32, "mel_len"), _ptr(out), _ptr(frame_idx),
                                    _ptr(mel_pos), _ptr(mel_mask), _ptr(out_f32, torch.float32, "out_f32"), B, L, Tmax,
                                    D, _stream()), "lr_expand")
    _count()
    return out, frame_idx, mel_pos, mel_mask


def bucket_embed_add(x, p, e, pbins, ebins, ptab, etab, want_idx: bool = False, out_f32=None):
    D = x.shape[-1]
    rows = x.numel() // D
    out = torch.empty_like(x)
    p_idx = torch.empty(p.shape, device=x.device, dtype=torch.int32) if want_idx else None
    e_idx = torch.empty(e.shape, device=x.device, dtype=torch.int32) if want_idx else None
    _check(load().ttsb200_bucket_embed_add(_ptr(x, torch.bfloat16, "x"), _ptr(p, torch.float32, "p"),
                                           _ptr(e, torch.float32, "e"), _ptr(pbins, torch.float32, "pitch_bins"),
                                           _ptr(ebins, torch.float32, "energy_bins"),
                                           _ptr(ptab, torch.float32, "pitch_table"),
                                           _ptr(etab, torch.float32, "energy_table"), _ptr(out), _ptr(p_idx),
                                           _ptr(e_idx), _ptr(out_f32, torch.float32, "out_f32"), rows, D, pbins.numel(),
                                           _stream()), "bucket_embed_add")
    _count()
    return out, p_idx, e_idx


def cast_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    y = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16) if out is None else out
    if y.numel() != x.numel():
        raise ValueError("cast_bf16: out must have as many elements as x")
    _check(load().ttsb200_cast_bf16(_ptr(x, torch.float32, "x"), _ptr(y, torch.bfloat16, "out"), x.numel(), _stream()), "cast_bf16")
    _count()
    return y


# --------------------------------------------------------------------------------------------------
# training step (include/ttsb200.h, "TRAINING STEP")
# --------------------------------------------------------------------------------------------------
def conv_wgrad(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, *, taps: int = 1, pad_left: int = 0,
               n_off: int = 0, N: Optional[int] = None, conv_layout: bool = False, dbias: Optional[torch.Tensor] = None) -> None:
    """dW += dy^T (*) x.  dy bf16 (B,T,ld) of which columns [n_off, n_off+N) are used; x bf16 (B,T,C).
    dw fp32: (N,C) for a linear, (taps,N,C) or — conv_layout — the reference's Conv1d (N,C,taps).
    dbias fp32 (N), optional: += the column sums of the same dy columns (the bias gradient), in the same launch."""
    B, T, ldy = dy.shape
    Cin = x.shape[2]
    N = ldy - n_off if N is None else N
    _ptr(dy, torch.bfloat16, "dy"); _ptr(x, torch.bfloat16, "x"); _ptr(dw, torch.float32, "dw")
    if dw.numel() != taps * N * Cin:
        raise ValueError(f"conv_wgrad: dw has {dw.numel()} elements, expected {taps}*{N}*{Cin}")
    target = dw
    if conv_layout and taps > 1:
        # accumulate tap-major (contiguous x-channels -> 16-byte vector reductions), then one strided add into the
        # reference's Conv1d (out, in, k) layout
        target = torch.zeros((taps, N, Cin), device=dw.device, dtype=torch.float32)
    sj, sn, sc = N * Cin, Cin, 1
    if dbias is not None and dbias.numel() != N:
        raise ValueError(f"conv_wgrad: dbias has {dbias.numel()} elements, expected {N}")
    _check(load().ttsb200_conv_wgrad_bias(dy.data_ptr() + 2 * n_off, ldy, x.data_ptr(), x.shape[2], target.data_ptr(),
                                          _ptr(dbias, torch.float32, "dbias"), sj, sn, sc, B, T, N, Cin, taps, pad_left,
                                          _stream()), "conv_wgrad")
    _count()
    if target is not dw:
        dw.view(N, Cin, taps).add_(target.permute(1, 2, 0))


def attn_delta(dctx: torch.Tensor, ctx: torch.Tensor, H: int) -> torch.Tensor:
    B, T, D = ctx.shape
    delta = torch.empty((B, H, T), device=ctx.device, dtype=torch.float32)
    _check(load().ttsb200_attn_delta(_ptr(dctx, torch.bfloat16, "dctx"), _ptr(ctx, torch.bfloat16, "ctx"),
                                     _ptr(delta), B, T, H, D // H, _stream()), "attn_delta")
    _count()
    return delta


def attention_bwd(qkv, dctx, mask, kv_len, lse2, delta, H: int, drop: Optional[tuple] = None,
                  causal: bool = False) -> torch.Tensor:
    B, T, W = qkv.shape
    dqkv = torch.empty_like(qkv)
    p, seed, site = drop if drop is not None else (0.0, 0, -1)
    if causal and p > 0:
        _check(load().ttsb200_attention_bwd_causal_drop(
            _ptr(qkv, torch.bfloat16, "qkv"), _ptr(dctx, torch.bfloat16, "dctx"), _ptr(mask, torch.uint8, "mask"),
            _ptr(kv_len, torch.int32, "kv_len"), _ptr(lse2, torch.float32, "lse2"), _ptr(delta, torch.float32, "delta"),
            _ptr(dqkv), B, T, H, W // (3 * H), float(p), int(seed), int(site), _stream()), "attention_bwd_causal_drop")
        _count(2)
        return dqkv
    if causal:
        _check(load().ttsb200_attention_bwd_causal(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(dctx, torch.bfloat16, "dctx"),
                                                   _ptr(mask, torch.uint8, "mask"), _ptr(kv_len, torch.int32, "kv_len"),
                                                   _ptr(lse2, torch.float32, "lse2"), _ptr(delta, torch.float32, "delta"),
                                                   _ptr(dqkv), B, T, H, W // (3 * H), _stream()), "attention_bwd_causal")
        _count(2)
        return dqkv
    _check(load().ttsb200_attention_bwd_drop(_ptr(qkv, torch.bfloat16, "qkv"), _ptr(dctx, torch.bfloat16, "dctx"),
                                             _ptr(mask, torch.uint8, "mask"), _ptr(kv_len, torch.int32, "kv_len"),
                                             _ptr(lse2, torch.float32, "lse2"), _ptr(delta, torch.float32, "delta"),
                                             _ptr(dqkv), B, T, H, W // (3 * H), float(p), int(seed), int(site),
                                             _stream()), "attention_bwd")
    _count(2)
    return dqkv


def layernorm_bwd(x, gamma, *, beta=None, g_f32=None, g_bf16=None, g_row=None, g_vec=None, row_mask=None,
                  add_f32=None, relu_gate=False, dx_f32=None, dx_bf16=None, dgamma=None, dbeta=None, dvec=None,
                  dscal=None, eps: float = 1e-5, drop_p: float = 0.0, drop_seed: int = 0, site_g: int = -1,
                  site_dx: int = -1, dx_mode: int = 0) -> None:
    """site_g / site_dx / dx_mode: the forward pass's dropout masks, see ttsb200_layernorm_bwd_drop."""
    D = x.shape[-1]
    rows = x.numel() // D
    _check(load().ttsb200_layernorm_bwd_drop(
        _ptr(x, torch.float32, "x"), _ptr(gamma, torch.float32, "gamma"), _ptr(beta, torch.float32, "beta"),
        _ptr(g_f32, torch.float32, "g_f32"), _ptr(g_bf16, torch.bfloat16, "g_bf16"),
        _ptr(g_row, torch.float32, "g_row"), _ptr(g_vec, torch.float32, "g_vec"),
        _ptr(row_mask, torch.uint8, "row_mask"), _ptr(add_f32, torch.float32, "add_f32"), int(relu_gate),
        _ptr(dx_f32, torch.float32, "dx_f32"), _ptr(dx_bf16, torch.bfloat16, "dx_bf16"),
        _ptr(dgamma, torch.float32, "dgamma"), _ptr(dbeta, torch.float32, "dbeta"),
        _ptr(dvec, torch.float32, "dvec"), _ptr(dscal, torch.float32, "dscal"), rows, D, eps, float(drop_p),
        int(drop_seed), int(site_g), int(site_dx), int(dx_mode), _stream()), "layernorm_bwd")
    _count()


def colsum(x: torch.Tensor, out: torch.Tensor, cols: Optional[int] = None, col_off: int = 0) -> None:
    """out[c] += sum over rows of x[..., col_off + c]; x bf16 or fp32 (.., ld)."""
    ld = x.shape[-1]
    rows = x.numel() // ld
    cols = ld - col_off if cols is None else cols
    if x.dtype not in (torch.bfloat16, torch.float32):
        raise TypeError("colsum: bf16 or fp32")
    _ptr(x, None, "x"); _ptr(out, torch.float32, "out")
    _check(load().ttsb200_colsum(x.data_ptr() + col_off * x.element_size(), int(x.dtype == torch.bfloat16), ld,
                                 out.data_ptr(), rows, cols, _stream()), "colsum")
    _count()


def bn_stats(x: torch.Tensor, stats: torch.Tensor) -> None:
    Cc = x.shape[-1]
    _check(load().ttsb200_bn_stats(_ptr(x, torch.float32, "x"), _ptr(stats, torch.float32, "stats"),
                                   x.numel() // Cc, Cc, _stream()), "bn_stats")
    _count()


def bn_tanh_fwd(x, mean_rstd, gamma, beta, h, drop: tuple = (0.0, 0, -1)) -> None:
    Cc = x.shape[-1]
    _check(load().ttsb200_bn_tanh_fwd(_ptr(x, torch.float32, "x"), _ptr(mean_rstd, torch.float32, "mean_rstd"),
                                      _ptr(gamma, torch.float32, "gamma"), _ptr(beta, torch.float32, "beta"),
                                      _ptr(h, torch.bfloat16, "h"), x.numel() // Cc, Cc, float(drop[0]), int(drop[1]),
                                      int(drop[2]), _stream()), "bn_tanh_fwd")
    _count()


def bn_tanh_bwd(dh, h, x, mean_rstd, gamma, red, dx, inv_count: float, apply: bool, drop: tuple = (0.0, 0, -1)) -> None:
    Cc = x.shape[-1]
    _check(load().ttsb200_bn_tanh_bwd(_ptr(dh, torch.bfloat16, "dh"), _ptr(h, torch.bfloat16, "h"),
                                      _ptr(x, torch.float32, "x"), _ptr(mean_rstd, torch.float32, "mean_rstd"),
                                      _ptr(gamma, torch.float32, "gamma"), _ptr(red, torch.float32, "red"),
                                      _ptr(dx, torch.bfloat16, "dx"), x.numel() // Cc, Cc, inv_count, int(apply),
                                      float(drop[0]), int(drop[1]), int(drop[2]), _stream()), "bn_tanh_bwd")
    _count()


def lr_bwd(g: torch.Tensor, cum: torch.Tensor, add_f32: Optional[torch.Tensor] = None) -> torch.Tensor:
    B, Tmax, D = g.shape
    L = cum.shape[1]
    dx = torch.empty((B, L, D), device=g.device, dtype=torch.float32)
    _check(load().ttsb200_lr_bwd(_ptr(g, torch.float32, "g"), _ptr(cum, torch.int32, "cum"), _ptr(dx),
                                 _ptr(add_f32, torch.float32, "add_f32"), B, L, Tmax, D, _stream()), "lr_bwd")
    _count()
    return dx


def embed_bwd(g: torch.Tensor, T: int, *, idx: Optional[torch.Tensor] = None, table: Optional[torch.Tensor] = None,
              skip_idx: int = -1, pe: Optional[torch.Tensor] = None, dalpha: Optional[torch.Tensor] = None) -> None:
    D = g.shape[-1]
    rows = g.numel() // D
    i64 = idx if idx is not None and idx.dtype == torch.int64 else None
    i32 = idx if idx is not None and idx.dtype == torch.int32 else None
    if idx is not None and i64 is None and i32 is None:
        raise TypeError("embed_bwd: idx must be int64 or int32")
    V = table.shape[0] if table is not None else 0
    _check(load().ttsb200_embed_bwd(_ptr(g, torch.float32, "g"), _ptr(i64), _ptr(i32),
                                    _ptr(table, torch.float32, "table"), skip_idx, _ptr(pe, torch.float32, "pe"),
                                    _ptr(dalpha, torch.float32, "dalpha"), rows, T, D, V, _stream()), "embed_bwd")
    _count()


def l1_loss(pred, target, loss, *, grad=None, grad_bf16=None, log_target: bool = False) -> None:
    n = pred.numel()
    if target.numel() != n:
        raise ValueError("l1_loss: shape mismatch")
    _check(load().ttsb200_l1_loss(_ptr(pred, torch.float32, "pred"), _ptr(target, torch.float32, "target"),
                                  _ptr(grad, torch.float32, "grad"), _ptr(grad_bf16, torch.bfloat16, "grad_bf16"),
                                  _ptr(loss, torch.float32, "loss"), n, 1.0 / n, int(log_target), _stream()),
           "l1_loss")
    _count()


def sumsq(g: torch.Tensor, out: torch.Tensor) -> None:
    _check(load().ttsb200_sumsq(_ptr(g, torch.float32, "g"), _ptr(out, torch.float32, "out"), g.numel(), _stream()),
           "sumsq")
    _count()


def adam_step(p, g, m, v, sumsq_t, *, gscale: float, clip: float, lr: float, beta1: float, beta2: float, eps: float,
              step: int) -> None:
    _check(load().ttsb200_adam_step(_ptr(p, torch.float32, "p"), _ptr(g, torch.float32, "g"),
                                    _ptr(m, torch.float32, "m"), _ptr(v, torch.float32, "v"),
                                    _ptr(sumsq_t, torch.float32, "sumsq"), p.numel(), gscale, clip, lr, beta1, beta2,
                                    eps, step, _stream()), "adam_step")
    _count()


def radam_step(p, g, m, v, sumsq_t, *, gscale: float, clip: float, lr: float, beta1: float, beta2: float, eps: float,
               step: int) -> None:
    """radam.RAdam.step (radam.py:30-92) over the flat buffers, same clip/scale prologue as adam_step."""
    _check(load().ttsb200_radam_step(_ptr(p, torch.float32, "p"), _ptr(g, torch.float32, "g"),
                                     _ptr(m, torch.float32, "m"), _ptr(v, torch.float32, "v"),
                                     _ptr(sumsq_t, torch.float32, "sumsq"), p.numel(), gscale, clip, lr, beta1, beta2,
                                     eps, step, _stream()), "radam_step")
    _count()


# --------------------------------------------------------------------------------------------------
# fp32-accurate encoder -> duration chain (include/ttsb200.h, "fp32-ACCURATE ENCODER -> DURATION CHAIN")
# "split" tensors are fp16 (.., 2C) = [hi | lo] with value = hi + lo
# --------------------------------------------------------------------------------------------------
XROW_ADD_LN, XROW_FFN_TAIL, XROW_LN_DOT = range(3)


def split_f16(x: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    D = x.shape[-1]
    out = torch.empty(x.shape[:-1] + (2 * D,), device=x.device, dtype=torch.float16)
    _check(load().ttsb200_split_f16(_ptr(x, torch.float32, "x"), _ptr(out), x.numel() // D, D, scale, _stream()),
           "split_f16")
    _count()
    return out


def gemm_x3(x_split: torch.Tensor, w_split: torch.Tensor, bias: torch.Tensor, inv_scale: float, *, taps: int = 1,
            pad_left: int = 0, relu: bool = False, split_out: bool = False) -> torch.Tensor:
    """x_split fp16 (B,T,2C); w_split fp16 (taps,N,2C).  Returns fp32 (B,T,N) or the fp16 split (B,T,2N)."""
    B, T, C2 = x_split.shape
    tw, N, Cw2 = w_split.shape
    if tw != taps or Cw2 != C2:
        raise ValueError(f"gemm_x3: weight {tuple(w_split.shape)} does not match taps={taps}, 2C={C2}")
    dev = x_split.device
    out32 = None if split_out else torch.empty((B, T, N), device=dev, dtype=torch.float32)
    outs = torch.empty((B, T, 2 * N), device=dev, dtype=torch.float16) if split_out else None
    _check(load().ttsb200_gemm_x3(_ptr(x_split, torch.float16, "x_split"), _ptr(w_split, torch.float16, "w_split"),
                                  _ptr(bias, torch.float32, "bias"), inv_scale, B, T, N, C2 // 2, taps, pad_left,
                                  int(relu), _ptr(out32), _ptr(outs), _stream()), "gemm_x3")
    _count()
    return outs if split_out else out32


def attention_x3(qkv_split: torch.Tensor, mask: torch.Tensor, kv_len: torch.Tensor, H: int) -> torch.Tensor:
    """qkv_split fp16 (B,T,2*3D) = [q k v hi | q k v lo]; returns the split context fp16 (B,T,2D)."""
    B, T, W = qkv_split.shape
    D = W // 6
    ctx = torch.empty((B, T, 2 * D), device=qkv_split.device, dtype=torch.float16)
    _check(load().ttsb200_attention_x3(_ptr(qkv_split, torch.float16, "qkv_split"), _ptr(mask, torch.uint8, "mask"),
                                       _ptr(kv_len, torch.int32, "kv_len"), _ptr(ctx), B, T, H, D // H, _stream()),
           "attention_x3")
    _count()
    return ctx


def xrow(mode: int, a: torch.Tensor, ln1: tuple, *, b0=None, b0_bf16=None, b0_split=None, c=None, ln2=None, dot_w=None,
         dot_b=None, row_mask=None, eps: float = 1e-5, s_out=None, z_out=None, out_split=None, out_bf16=None,
         out_f32=None, out_row=None, drop_p: float = 0.0, drop_seed: int = 0, site_b: int = -1, site_s: int = -1,
         site_t: int = -1) -> None:
    D = a.shape[-1]
    rows = a.numel() // D
    g2, b2 = (ln2 if ln2 is not None else (None, None))
    _check(load().ttsb200_xrow(mode, _ptr(a, torch.float32, "a"), _ptr(b0, torch.float32, "b0"),
                               _ptr(b0_bf16, torch.bfloat16, "b0_bf16"),
                               _ptr(b0_split, torch.float16, "b0_split"), _ptr(c, torch.float32, "c"),
                               _ptr(ln1[0], torch.float32, "g1"), _ptr(ln1[1], torch.float32, "b1"),
                               _ptr(g2, torch.float32, "g2"), _ptr(b2, torch.float32, "b2"),
                               _ptr(dot_w, torch.float32, "dot_w"), _ptr(dot_b, torch.float32, "dot_b"),
                               _ptr(row_mask, torch.uint8, "row_mask"), eps, _ptr(s_out, torch.float32, "s_out"),
                               _ptr(z_out, torch.float32, "z_out"), _ptr(out_split, torch.float16, "out_split"),
                               _ptr(out_bf16, torch.bfloat16, "out_bf16"), _ptr(out_f32, torch.float32, "out_f32"),
                               _ptr(out_row, torch.float32, "out_row"), rows, D, float(drop_p), int(drop_seed),
                               int(site_b), int(site_s), int(site_t), _stream()), "xrow")
    _count()


# --------------------------------------------------------------------------------------------------
# dropout (training): counter-based masks, regenerated from (p, seed, site) wherever they are needed
# --------------------------------------------------------------------------------------------------
def dropout_(x: torch.Tensor, p: float, seed: int, site: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x * m (m = 0 or 1/(1-p)), in place unless `out` is given.  fp32 or bf16."""
    if x.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError("dropout_: fp32 or bf16")
    dst = x if out is None else out
    if dst.dtype != x.dtype or dst.numel() != x.numel():
        raise ValueError("dropout_: out must match x")
    _check(load().ttsb200_dropout(_ptr(x, None, "x"), _ptr(dst, None, "out"), int(x.dtype == torch.bfloat16), x.numel(),
                                  float(p), int(seed), int(site), _stream()), "dropout")
    _count()
    return dst


def dropout_mask(n: int, p: float, seed: int, site: int, device="cuda") -> torch.Tensor:
    out = torch.empty(n, device=device, dtype=torch.float32)
    _check(load().ttsb200_dropout_mask(_ptr(out), n, float(p), int(seed), int(site), _stream()), "dropout_mask")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
# autoregressive decode step (Transformer-TTS)
# ------------------------------------------------------------------------------------------------
def pad_weight_rows(w: torch.Tensor) -> torch.Tensor:
    """bf16 [N][roundup(K,32)] copy of a Linear weight (N,K), rows zero-padded: the layout small_linear streams."""
    N, K = w.shape
    Kp = (K + 31) // 32 * 32
    out = torch.zeros((N, Kp), device=w.device, dtype=torch.bfloat16)
    out[:, :K] = w.to(torch.bfloat16)
    return out


def small_linear(w: torch.Tensor, bias: Optional[torch.Tensor], K: int, *, x_f32: Optional[torch.Tensor] = None,
                 x_bf16: Optional[torch.Tensor] = None, ldx: Optional[int] = None, R: Optional[int] = None, ln=None,
                 eps: float = 1e-5, relu: bool = False, res_f32: Optional[torch.Tensor] = None,
                 add_row: Optional[torch.Tensor] = None, add_scale: Optional[torch.Tensor] = None,
                 out_f32: Optional[torch.Tensor] = None, ldo: Optional[int] = None,
                 out_bf16: Optional[torch.Tensor] = None, xn_bf16: Optional[torch.Tensor] = None,
                 pos: Optional[torch.Tensor] = None, x_pos_stride: int = 0, x_pos_bias: int = 0,
                 out_pos_stride: int = 0) -> None:
    """y = act(LN?(x) W^T + b) (+ res) (+ add_scale * add_row[pos]) for a few rows; see ttsb200_small_linear."""
    x = x_f32 if x_f32 is not None else x_bf16
    N = w.shape[0]
    a = SmallLinearArgs()
    a.x_f32 = _ptr(x_f32, torch.float32, "x_f32")
    a.x_bf16 = _ptr(x_bf16, torch.bfloat16, "x_bf16")
    a.ldx = int(ldx if ldx is not None else x.shape[-1])
    a.R = int(R if R is not None else x.shape[0])
    a.ln_g = _ptr(ln[0], torch.float32, "ln_g") if ln is not None else None
    a.ln_b = _ptr(ln[1], torch.float32, "ln_b") if ln is not None else None
    a.eps = eps
    a.w_bf16 = _ptr(w, torch.bfloat16, "w")
    a.ldw = int(w.shape[1])
    a.bias = _ptr(bias, torch.float32, "bias")
    a.relu = int(relu)
    a.res_f32 = _ptr(res_f32, torch.float32, "res_f32")
    a.ldres = int(res_f32.shape[-1]) if res_f32 is not None else 0
    a.add_row = _ptr(add_row, torch.float32, "add_row")
    a.add_scale = _ptr(add_scale, torch.float32, "add_scale")
    a.out_f32 = _ptr(out_f32, torch.float32, "out_f32")
    a.ldo = int(ldo if ldo is not None else (out_f32.shape[-1] if out_f32 is not None else 0))
    a.out_bf16 = _ptr(out_bf16, torch.bfloat16, "out_bf16")
    a.ldob = int(out_bf16.shape[-1]) if out_bf16 is not None else 0
    a.xn_bf16 = _ptr(xn_bf16, torch.bfloat16, "xn_bf16")
    a.pos = _ptr(pos, torch.int32, "pos")
    a.x_pos_stride = int(x_pos_stride)
    a.x_pos_bias = int(x_pos_bias)
    a.out_pos_stride = int(out_pos_stride)
    a.K = int(K)
    a.N = int(N)
    _check(load().ttsb200_small_linear(C.byref(a), _stream()), "small_linear")
    _count()


def attn_decode(q: torch.Tensor, q_off: int, ldq: int, k: torch.Tensor, v_off: int, strides, cap: int, H: int, Dh: int,
                ctx: torch.Tensor, *, kv_new: Optional[torch.Tensor] = None, knew_off: int = 0, vnew_off: int = 0,
                pos: Optional[torch.Tensor] = None, kv_len: Optional[torch.Tensor] = None) -> None:
    """One query row per (utterance, head).  q / kv_new: column ranges (element offsets) of a (R, ldq) bf16 buffer such as
    the step's q|k|v projection.  k: the bf16 buffer holding keys AND values: values start v_off elements after the keys;
    strides = (utterance, head, key) element strides; cap = key slots."""
    R = ctx.shape[0]
    es = 2
    qp = _ptr(q, torch.bfloat16, "q") + q_off * es
    kp = vp = None
    if kv_new is not None:
        base = _ptr(kv_new, torch.bfloat16, "kv_new")
        kp, vp = base + knew_off * es, base + vnew_off * es
    kb = _ptr(k, torch.bfloat16, "k")
    _check(load().ttsb200_attn_decode(qp, ldq, kp, vp, int(kv_new.shape[-1]) if kv_new is not None else 0, kb,
                                      kb + v_off * es, int(strides[0]), int(strides[1]), int(strides[2]),
                                      _ptr(pos, torch.int32, "pos"), _ptr(kv_len, torch.int32, "kv_len"),
                                      _ptr(ctx, torch.bfloat16, "ctx"), int(ctx.shape[-1]), R, H, Dh, cap, _stream()),
           "attn_decode")
    _count()


def advance_pos(pos: torch.Tensor, delta: int = 1) -> None:
    _check(load().ttsb200_advance_pos(_ptr(pos, torch.int32, "pos"), delta, _stream()), "advance_pos")
    _count()


def ffn_tail_rows(u, yn, y, ln1, ln2, eps, z, zn) -> None:
    """z = y + LN(u + yn; ln1); zn = bf16(LN(z; ln2)) for the rows of a decode step (see ttsb200_ffn_tail_rows)."""
    R, D = u.shape
    _check(load().ttsb200_ffn_tail_rows(_ptr(u, torch.float32, "u"), _ptr(yn, torch.bfloat16, "yn"), _ptr(y, torch.float32, "y"),
                                        _ptr(ln1[0], torch.float32, "ln1_g"), _ptr(ln1[1], torch.float32, "ln1_b"),
                                        _ptr(ln2[0], torch.float32, "ln2_g"), _ptr(ln2[1], torch.float32, "ln2_b"), eps,
                                        _ptr(z, torch.float32, "z"), _ptr(zn, torch.bfloat16, "zn"), R, D, _stream()),
           "ffn_tail_rows")
    _count()


def decode_attn_out_q(ctx, x_in, wo, bo, ln, wq, bq, eps, x_out, q_out) -> None:
    """x_out = x_in + ctx Wo^T + bo; q_out = bf16(LN(x_out) Wq^T + bq): one 16-CTA cluster per 16 rows."""
    R, D = x_in.shape
    f32, bf = torch.float32, torch.bfloat16
    _check(load().ttsb200_decode_attn_out_q(_ptr(ctx, bf, "ctx"), _ptr(x_in, f32, "x_in"), _ptr(wo, bf, "wo"), _ptr(bo, f32, "bo"),
                                            _ptr(ln[0], f32, "ln_g"), _ptr(ln[1], f32, "ln_b"), _ptr(wq, bf, "wq"),
                                            _ptr(bq, f32, "bq"), eps, _ptr(x_out, f32, "x_out"), _ptr(q_out, bf, "q_out"), R, D,
                                            _stream()), "decode_attn_out_q")
    _count()


def decode_ffn_block(ctx, x_in, wo, bo, ln3, w1, b1, w2, b2, ln_ff, ln_next, eps, z_out, zn_out) -> None:
    """Second attention's out-projection, FeedForward and the next LayerNorm of a decoder block in one cluster kernel."""
    R, D = x_in.shape
    f32, bf = torch.float32, torch.bfloat16
    _check(load().ttsb200_decode_ffn_block(
        _ptr(ctx, bf, "ctx"), _ptr(x_in, f32, "x_in"), _ptr(wo, bf, "wo"), _ptr(bo, f32, "bo"), _ptr(ln3[0], f32, "ln3_g"),
        _ptr(ln3[1], f32, "ln3_b"), _ptr(w1, bf, "w1"), _ptr(b1, f32, "b1"), _ptr(w2, bf, "w2"), _ptr(b2, f32, "b2"),
        _ptr(ln_ff[0], f32, "lnff_g"), _ptr(ln_ff[1], f32, "lnff_b"), _ptr(ln_next[0], f32, "lnn_g"),
        _ptr(ln_next[1], f32, "lnn_b"), eps, _ptr(z_out, f32, "z_out"), _ptr(zn_out, bf, "zn_out"),
        R, D, int(w1.shape[0]), _stream()), "decode_ffn_block")
    _count()


def bce_logits_loss(x, y, loss, *, grad=None, pos_weight: float = 1.0) -> None:
    """loss += mean BCE-with-logits(x, y; pos_weight) (train.py:217); grad (optional) receives d loss / d x."""
    n = x.numel()
    if y.numel() != n:
        raise ValueError("bce_logits_loss: shape mismatch")
    _check(load().ttsb200_bce_logits_loss(_ptr(x, torch.float32, "x"), _ptr(y, torch.float32, "y"),
                                          _ptr(grad, torch.float32, "grad"), _ptr(loss, torch.float32, "loss"), n, 1.0 / n,
                                          float(pos_weight), _stream()), "bce_logits_loss")
    _count()


def relu_gate(g: torch.Tensor, gate: torch.Tensor) -> torch.Tensor:
    """bf16(g where gate > 0 else 0): backward of a ReLU whose output is `gate`."""
    out = torch.empty(g.shape, device=g.device, dtype=torch.bfloat16)
    _check(load().ttsb200_relu_gate(_ptr(g, torch.float32, "g"), _ptr(gate, torch.bfloat16, "gate"), _ptr(out), g.numel(),
                                    _stream()), "relu_gate")
    _count()
    return out


_WT_DTYPE = None
_WT_KEEP: list = []


def weight_transpose(srcs, dsts) -> None:
    """dsts[i][taps-1-j, c, n] = srcs[i][j, n, c] for every pair, ONE launch (the dgrad operands of a backward pass).
    srcs: bf16 (taps, N, C) contiguous; dsts: bf16 (taps, C, N) contiguous."""
    import numpy as np

    global _WT_DTYPE
    if _WT_DTYPE is None:
        _WT_DTYPE = np.dtype([("src", "<u8"), ("dst", "<u8"), ("taps", "<i4"), ("N", "<i4"), ("C", "<i4"), ("tile0", "<i4")])
    if not srcs or len(srcs) != len(dsts):
        raise ValueError("weight_transpose: need as many destinations as sources")
    tab = np.zeros(len(srcs), dtype=_WT_DTYPE)
    tiles = 0
    for i, (a, b) in enumerate(zip(srcs, dsts)):
        taps, N, Cin = a.shape
        if tuple(b.shape) != (taps, Cin, N):
            raise ValueError(f"weight_transpose: destination {tuple(b.shape)} for source {tuple(a.shape)}")
        tab[i] = (_ptr(a, torch.bfloat16, "src"), _ptr(b, torch.bfloat16, "dst"), taps, N, Cin, tiles)
        tiles += taps * ((N + 63) // 64) * ((Cin + 63) // 64)
    host = torch.from_numpy(tab.view(np.uint8))
    if torch.cuda.is_current_stream_capturing():
        # a captured copy reads its source again at every replay: page-locked and kept alive with the process
        host = host.pin_memory()
        _WT_KEEP.append(host)
        tab_dev = host.to(srcs[0].device, non_blocking=True)
        _WT_KEEP.append(tab_dev)
    else:
        tab_dev = host.to(srcs[0].device)
    _check(load().ttsb200_weight_transpose(tab_dev.data_ptr(), len(srcs), tiles, _stream()), "weight_transpose")
    _count()
