"""Kernel-sequencing helpers shared by the drop-in modules.

Holds the derived (bf16, re-laid-out, fused) weight caches and the functions that chain the
sm_100a kernels of ``csrc/`` for one FFT block, one variance predictor and the PostNet.  The
``state_dict`` stays in the reference layout (fp32, Conv1d weights (out,in,k), separate q/k/v);
caches are rebuilt whenever a source parameter's storage or version counter changes
(``load_state_dict`` and optimizer steps both bump ``Tensor._version``).

The inference sequences live here; the training step (saved activations, explicit backward) is
``transformer_tts_b200/training.py`` and reuses these caches.
"""
from __future__ import annotations

from functools import lru_cache
from typing import Optional, Sequence

import numpy as np
import torch

from .. import lib

LN_EPS = 1e-5
BN_EPS = 1e-5


# --------------------------------------------------------------------------------------------------
# positional table (Models/modules.py:97-105): exponents 2*i/d for sin (i already even) and
# 2*(i+1)/d for cos, float64 math stored as fp32.  Vectorised here; tests check it bit-for-bit
# against the oracle's scalar restatement of the reference loop.
# --------------------------------------------------------------------------------------------------
@lru_cache(maxsize=4)
def positional_table_cpu(d_model: int, max_seq_len: int = 5000) -> torch.Tensor:
    pos = np.arange(max_seq_len, dtype=np.float64)[:, None]
    i = np.arange(0, d_model, 2, dtype=np.float64)[None, :]
    pe = np.zeros((max_seq_len, d_model), dtype=np.float64)
    pe[:, 0::2] = np.sin(pos / np.power(10000.0, (2 * i) / d_model))
    pe[:, 1::2] = np.cos(pos / np.power(10000.0, (2 * (i + 1)) / d_model))
    return torch.from_numpy(pe).to(torch.float32)


_PE_DEVICE = {}


def positional_table(d_model: int, device: torch.device, max_seq_len: int = 5000) -> torch.Tensor:
    key = (d_model, max_seq_len, str(device))
    t = _PE_DEVICE.get(key)
    if t is None:
        t = positional_table_cpu(d_model, max_seq_len).to(device).contiguous()
        _PE_DEVICE[key] = t
    return t


# --------------------------------------------------------------------------------------------------
# guards
# --------------------------------------------------------------------------------------------------
def require_inference(module: torch.nn.Module, what: str) -> None:
    if module.training:
        raise NotImplementedError(
            f"{what}: standalone sub-modules run the inference kernels only (call .eval()); the training step is "
            "built for the whole model — FastSpeech2.forward in train() mode / transformer_tts_b200.training.Trainer")


def require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{what}: tensors must live on a CUDA device — this package has no CPU path")


def mask_to_u8(mask: Optional[torch.Tensor], B: int, T: int, device) -> torch.Tensor:
    """(B,1,T) / (B,T) bool-like -> contiguous uint8 (B,T); None -> all valid."""
    if mask is None:
        return torch.ones((B, T), dtype=torch.uint8, device=device)
    if mask.dim() == 3:
        if mask.shape[1] != 1:
            raise NotImplementedError("only key-padding masks (B,1,T) are supported (fix_mask is out of scope)")
        mask = mask[:, 0]
    if tuple(mask.shape) != (B, T):
        raise ValueError(f"mask shape {tuple(mask.shape)} does not match (B={B}, T={T})")
    return (mask != 0).to(torch.uint8).contiguous()


def _sig(params: Sequence[torch.Tensor]):
    return tuple((p.data_ptr(), p._version) for p in params)


def _mirror(w: torch.Tensor) -> Optional[torch.Tensor]:
    """The bf16 mirror of a parameter that lives in a training.Trainer's flat buffer (ONE cast kernel refreshes the mirror of
    the whole model after an optimizer step), if it is still current (`_version` guards writes that bypass the trainer)."""
    m = getattr(w, "_ttsb_bf16", None)
    if m is not None and m[1] == w._version and m[0].shape == w.shape:
        return m[0]
    return None


def _bf16(w: torch.Tensor) -> torch.Tensor:
    m = _mirror(w)
    return m if m is not None else w.detach().to(torch.bfloat16).contiguous()


def _f32(w: torch.Tensor) -> torch.Tensor:
    return w.detach().to(torch.float32).contiguous()


def _adjacent(ts) -> bool:
    return all(a.is_contiguous() and a.data_ptr() + a.numel() * a.element_size() == b.data_ptr() for a, b in zip(ts, ts[1:])) \
        and ts[-1].is_contiguous()


def _cat_rows(ts) -> torch.Tensor:
    """torch.cat(ts, 0) — or, when the tensors already sit next to each other in memory (a trainer's flat layout places
    q | k | v so), a view of that memory: no kernel."""
    if len(ts) > 1 and _adjacent(ts) and all(t.shape[1:] == ts[0].shape[1:] and t.dtype == ts[0].dtype for t in ts):
        rows = sum(t.shape[0] for t in ts)
        shape = (rows,) + tuple(ts[0].shape[1:])
        stride, n = [], 1
        for d in reversed(shape):
            stride.insert(0, n)
            n *= d
        return torch.as_strided(ts[0], shape, tuple(stride))
    return torch.cat(list(ts), 0)


def _bf16_cat(ws) -> torch.Tensor:
    """bf16 of the row-wise concatenation of parameters (q | k | v operands)."""
    ms = [_mirror(w) for w in ws]
    if all(m is not None for m in ms):
        return _cat_rows(ms)
    return torch.cat([w.detach() for w in ws], 0).to(torch.bfloat16).contiguous()


def _f32_cat(ws) -> torch.Tensor:
    ds = [w.detach() for w in ws]
    if all(d.dtype == torch.float32 for d in ds):
        return _cat_rows(ds)
    return torch.cat(ds, 0).to(torch.float32).contiguous()


def conv_w_to_taps(w: torch.Tensor) -> torch.Tensor:
    """Conv1d weight (out, in, k) -> bf16 (k, out, in): one K-major GEMM operand per tap."""
    m = _mirror(w)
    if m is not None:
        return m.permute(2, 0, 1).contiguous()      # k = 1: the same memory, no kernel
    return w.detach().permute(2, 0, 1).to(torch.bfloat16).contiguous()


class Cache:
    """Per-module derived tensors with a staleness signature."""

    def __init__(self):
        self.sig = None
        self.t = {}

    def stale(self, params) -> bool:
        return self.sig != _sig(params)

    def commit(self, params) -> None:
        self.sig = _sig(params)


# --------------------------------------------------------------------------------------------------
# FFT block (Models/layers.py:29-41) = 4 GEMM launches + 1 attention launch
# --------------------------------------------------------------------------------------------------
def layer_cache(layer) -> dict:
    c: Cache = layer._ttsb_cache
    a, ff = layer.attn, layer.ff
    params = [a.q_linear.weight, a.q_linear.bias, a.k_linear.weight, a.k_linear.bias, a.v_linear.weight,
              a.v_linear.bias, a.out.weight, a.out.bias, ff.f_1.weight, ff.f_1.bias, ff.f_2.weight, ff.f_2.bias,
              ff.layer_norm.weight, ff.layer_norm.bias, layer.norm_1.weight, layer.norm_1.bias,
              layer.norm_2.weight, layer.norm_2.bias]
    if c.stale(params):
        if a.concat_after:
            raise NotImplementedError("concat_after=True is not supported (hparams default is False)")
        t = c.t
        t["wqkv"] = _bf16_cat([a.q_linear.weight, a.k_linear.weight, a.v_linear.weight]).unsqueeze(0).contiguous()
        t["bqkv"] = _f32_cat([a.q_linear.bias, a.k_linear.bias, a.v_linear.bias])
        t["wo"] = _bf16(a.out.weight).unsqueeze(0).contiguous()
        t["bo"] = _f32(a.out.bias)
        t["w1"] = conv_w_to_taps(ff.f_1.weight)
        t["b1"] = _f32(ff.f_1.bias)
        t["w2"] = conv_w_to_taps(ff.f_2.weight)
        t["b2"] = _f32(ff.f_2.bias)
        t["ln_ff"] = (_f32(ff.layer_norm.weight), _f32(ff.layer_norm.bias))
        t["ln1"] = (_f32(layer.norm_1.weight), _f32(layer.norm_1.bias))
        t["ln2"] = (_f32(layer.norm_2.weight), _f32(layer.norm_2.bias))
        t["k"] = ff.f_1.weight.shape[-1]
        c.commit(params)
    return c.t


# Ragged batches: the frame-level stack (k=1 convs, masked keys) and the causal PostNet never read a row beyond an
# utterance's own length from a valid row, so 128-row tiles that are pure padding are dropped from the kernels' tile
# lists (lib.build_tile_map).  At cfg2 the mean utterance is 995 frames against Tmax 1102: ~10 % of the tiles.
RAGGED_TILES = True


def run_fft_block(t: dict, heads: int, x: torch.Tensor, xn: torch.Tensor, mask_u8: torch.Tensor,
                  kv_len: torch.Tensor, next_ln, want_norm_f32: bool = False, tile_map=None, zero_dropped=False):
    """x: fp32 residual stream (B,T,D); xn = norm_1(x) in bf16.  Returns (x', LN_next(x') bf16[, fp32]).
    tile_map (k=1 blocks only): rows of unlisted 128-row tiles are neither read nor written."""
    B, T, D = x.shape
    dev = x.device
    k = t["k"]
    qkv = torch.empty((B, T, 3 * D), device=dev, dtype=torch.bfloat16)
    if tile_map is not None and k != 1:
        raise ValueError("run_fft_block: a ragged tile list needs k=1 convs (a k>1 conv reads its padding rows)")
    lib.conv_gemm(xn, t["wqkv"], t["bqkv"], block_n=192, out_bf16=qkv, tile_map=tile_map)  # modules.py:49-51
    ctx = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
    lib.attention_fwd(qkv, mask_u8, kv_len, ctx, heads, skip_padded_queries=tile_map is not None)  # modules.py:7-21,58-65
    y = torch.empty((B, T, D), device=dev, dtype=torch.float32)
    yn = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
    lib.conv_gemm(ctx, t["wo"], t["bo"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x, res_bstride=T * D,
                  res_tstride=D, ln1=t["ln2"], out_f32=y, out_bf16=yn, tile_map=tile_map)  # modules.py:68 + layers.py:33,35
    zn32 = torch.empty((B, T, D), device=dev, dtype=torch.float32) if want_norm_f32 else None
    if FUSED_FFN and k == 1 and D == 384 and t["w1"].shape[1] == 4 * D and not want_norm_f32 and \
            B * ((T + 127) // 128) > SHORT_BATCH_TILES:
        # modules.py:81-88 + layers.py:40 + the next norm_1 as ONE kernel: the 1536-wide hidden activation stays in
        # tensor memory instead of making a 434 MB round trip through HBM (bit-identical to the two-kernel path)
        z = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        zn = (torch.zeros if (tile_map is not None and zero_dropped) else torch.empty)((B, T, D), device=dev,
                                                                                       dtype=torch.bfloat16)
        lib.ffn_fused(yn, t["w1"], t["b1"], t["w2"], t["b2"], y, t["ln_ff"], next_ln, z, zn, eps=LN_EPS, tile_map=tile_map)
        return z, zn, zn32
    cb = ffn_chunk_utterances(B, T, D)
    if cb < B:
        # The 1536-wide hidden activation of the whole batch (217 MB at cfg2) is the largest HBM item of the step and
        # conv1 is bound by writing it.  Run conv1 -> conv2 on groups of utterances whose hidden activation fits the
        # 126 MB L2, through ONE reused buffer: conv2 reads it back from L2 and the next group overwrites the lines
        # before they are ever written to DRAM.
        z = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        zn = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
        h = torch.empty((cb, T, 4 * D), device=dev, dtype=torch.bfloat16)
        for b0 in range(0, B, cb):
            b1 = min(B, b0 + cb)
            hb = h[:b1 - b0]
            lib.conv_gemm(yn[b0:b1], t["w1"], t["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU,
                          out_bf16=hb)  # modules.py:83
            run_ffn2(t, hb, yn[b0:b1], y[b0:b1], next_ln, zn32=None if zn32 is None else zn32[b0:b1],
                     out=(z[b0:b1], zn[b0:b1]))
        return z, zn, zn32
    h = torch.empty((B, T, 4 * D), device=dev, dtype=torch.bfloat16)
    lib.conv_gemm(yn, t["w1"], t["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU, out_bf16=h,
                  tile_map=tile_map)  # modules.py:83
    z, zn = run_ffn2(t, h, yn, y, next_ln, zn32=zn32, tile_map=tile_map,
                     zero_dropped=zero_dropped)  # modules.py:84-87 + layers.py:40 + next norm
    return z, zn, zn32


# k=1 blocks (the frame-level stack): the FeedForward as one kernel (csrc/ffn_fused.cuh).  False = conv1 and conv2 as two
# conv_gemm launches (the path the training step and the short batches take).
FUSED_FFN = True

# conv1 -> conv2 grouping (see run_fft_block): a group is FFN_GROUP_WAVES full waves of 128-row tiles over the 148 SMs
# (a group of 17 utterances x 9 tiles = 153 tiles would run a second, almost empty wave); 0 disables the grouping.
FFN_GROUP_WAVES = 0
NUM_SMS = 148


def ffn_chunk_utterances(B: int, T: int, D: int) -> int:
    """Utterances per conv1 -> conv2 group: all of them when grouping is off or the hidden activation is small."""
    m_tiles = (T + 127) // 128
    if FFN_GROUP_WAVES <= 0 or B * T * 4 * D * 2 <= (96 << 20):
        return B
    return max(1, min(B, (NUM_SMS // m_tiles) * FFN_GROUP_WAVES))


# Whole-row LayerNorm epilogues run one CTA per 128-row tile.  With fewer tiles than this, the k-tap conv2 of
# the text encoder (K = 1536*k) would occupy a fraction of the 148 SMs for its whole (tensor-bound) duration, so
# the GEMM runs column-split over 2-3x the CTAs and a row kernel applies the two LayerNorms.
SHORT_BATCH_TILES = 74


def run_ffn2(t: dict, h: torch.Tensor, yn: torch.Tensor, y: torch.Tensor, next_ln, zn32=None, pre_f32=None,
             out=None, tile_map=None, zero_dropped=False):
    """z = y + LN_ff(conv2(h) + b2 + yn); zn = LN_next(z).  Returns (z fp32, zn bf16) (written into `out` if given)."""
    B, T, D = y.shape
    dev = y.device
    k = t["k"]
    if out is not None:
        z, zn = out
    else:
        z = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        # rows of dropped (padding-only) tiles are never written: the stack's final output is defined (zero) there
        zn = (torch.zeros if (tile_map is not None and zero_dropped) else torch.empty)((B, T, D), device=dev,
                                                                                       dtype=torch.bfloat16)
    tiles = B * ((T + 127) // 128)
    if tiles <= SHORT_BATCH_TILES and D == 384 and tile_map is None:
        u = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        lib.conv_gemm(h, t["w2"], t["b2"], taps=k, pad_left=k // 2, block_n=128 if tiles * 3 <= 148 else 192,
                      out_f32=u)
        lib.ffn2_tail(u, yn, y, t["ln_ff"], next_ln, LN_EPS, z, zn, zn32=zn32, pre_f32=pre_f32)
    else:
        lib.conv_gemm(h, t["w2"], t["b2"], taps=k, pad_left=k // 2, block_n=D, epi=lib.EPI_FFN2, res_bf16=yn,
                      res_f32=y, res_bstride=T * D, res_tstride=D, ln1=t["ln_ff"], ln2=next_ln, out_f32=z,
                      out_bf16=zn, out2_f32=zn32, pre_f32=pre_f32, tile_map=tile_map)
    return z, zn


# --------------------------------------------------------------------------------------------------
# VariancePredictor (Models/varianceadaptor.py:216-231) = 2 GEMM launches
# --------------------------------------------------------------------------------------------------
def predictor_cache(pred) -> dict:
    c: Cache = pred._ttsb_cache
    params = [pred.conv1.weight, pred.conv1.bias, pred.layer_norm1.weight, pred.layer_norm1.bias,
              pred.conv2.weight, pred.conv2.bias, pred.layer_norm2.weight, pred.layer_norm2.bias,
              pred.linear_layer.weight, pred.linear_layer.bias]
    if c.stale(params):
        t = c.t
        t["w1"], t["b1"] = conv_w_to_taps(pred.conv1.weight), _f32(pred.conv1.bias)
        t["w2"], t["b2"] = conv_w_to_taps(pred.conv2.weight), _f32(pred.conv2.bias)
        t["ln1"] = (_f32(pred.layer_norm1.weight), _f32(pred.layer_norm1.bias))
        t["ln2"] = (_f32(pred.layer_norm2.weight), _f32(pred.layer_norm2.bias))
        t["dw"] = _f32(pred.linear_layer.weight.reshape(-1))
        t["db"] = _f32(pred.linear_layer.bias.reshape(-1))
        c.commit(params)
    return c.t


def run_predictor(t: dict, x_bf16: torch.Tensor, mask_u8: Optional[torch.Tensor]) -> torch.Tensor:
    B, T, _ = x_bf16.shape
    fs = t["b1"].numel()
    dev = x_bf16.device
    h = torch.empty((B, T, fs), device=dev, dtype=torch.bfloat16)
    lib.conv_gemm(x_bf16, t["w1"], t["b1"], taps=3, pad_left=1, block_n=fs, epi=lib.EPI_RELU_LN, ln1=t["ln1"],
                  out_bf16=h)  # varianceadaptor.py:217-219
    out = torch.empty((B, T), device=dev, dtype=torch.float32)
    lib.conv_gemm(h, t["w2"], t["b2"], taps=3, pad_left=1, block_n=fs, epi=lib.EPI_RELU_LN_DOT, ln1=t["ln2"],
                  dot_w=t["dw"], dot_b=t["db"], row_mask=mask_u8, out_row=out)  # :220-229
    return out


# --------------------------------------------------------------------------------------------------
# PostConvNet (Models/postnets.py:64-79), eval-mode BatchNorm folded into the conv epilogue
# --------------------------------------------------------------------------------------------------
def postnet_cache(pn) -> dict:
    c: Cache = pn._ttsb_cache
    bns = [pn.pre_batchnorm] + list(pn.batch_norm_list)
    convs = [pn.conv1] + list(pn.conv_list)
    params = [pn.out.weight, pn.out.bias, pn.conv2.weight, pn.conv2.bias]
    for cv, bn in zip(convs, bns):
        params += [cv.weight, cv.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var]
    if c.stale(params):
        t = c.t
        t["wout"], t["bout"] = _bf16(pn.out.weight).unsqueeze(0).contiguous(), _f32(pn.out.bias)
        t["convs"] = []
        for cv, bn in zip(convs, bns):
            scale = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
            shift = (cv.bias.detach().float() - bn.running_mean.detach().float()) * scale + bn.bias.detach().float()
            t["convs"].append((conv_w_to_taps(cv.weight), scale.contiguous(), shift.contiguous()))
        t["w2"], t["b2"] = conv_w_to_taps(pn.conv2.weight), _f32(pn.conv2.bias)
        c.commit(params)
    return c.t


def run_postnet(t: dict, d_bf16: torch.Tensor, tile_map=None):
    """tile_map: the convs are causal (pad 4 + drop the last 4), so a listed tile only reads listed rows; the mel
    outputs of unlisted (padding) tiles are zero."""
    B, T, D = d_bf16.shape
    dev = d_bf16.device
    md = t["bout"].numel()
    new = torch.empty if tile_map is None else torch.zeros
    mel_pre = new((B, T, md), device=dev, dtype=torch.float32)
    mel_pre_bf = torch.empty((B, T, md), device=dev, dtype=torch.bfloat16)
    lib.conv_gemm(d_bf16, t["wout"], t["bout"], block_n=md, out_f32=mel_pre, out_bf16=mel_pre_bf,
                  tile_map=tile_map)  # postnets.py:67
    h = mel_pre_bf
    for w, scale, shift in t["convs"]:  # postnets.py:71-73 (pad 4 + drop last 4 == causal, pad_left 4)
        o = torch.empty((B, T, w.shape[1]), device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(h, w, shift, taps=5, pad_left=4, block_n=192, act=lib.ACT_TANH, scale=scale, out_bf16=o,
                      tile_map=tile_map)
        h = o
    mel_post = new((B, T, md), device=dev, dtype=torch.float32)
    lib.conv_gemm(h, t["w2"], t["b2"], taps=5, pad_left=4, block_n=md, res_f32=mel_pre, res_bstride=T * md,
                  res_tstride=md, out_f32=mel_post, tile_map=tile_map)  # postnets.py:74-75
    return mel_pre, mel_post


# --------------------------------------------------------------------------------------------------
# fp32-accurate encoder -> duration chain (csrc/precise.cu): Models/encoder.py:83-112 + the duration predictor
# Models/varianceadaptor.py:216-229 evaluated to fp32 accuracy on the tensor cores (fp16 hi/lo operand splitting,
# three products per term, fp32 round-to-nearest accumulation outside the tensor core).  The reference rounds an
# fp32 log_d (varianceadaptor.py:75): with this chain free-running durations match it bit for bit wherever the
# reference's own exp(log_d) - 1 is not within ~1e-4 of a rounding boundary.
# --------------------------------------------------------------------------------------------------
def split_weight(w: torch.Tensor):
    """fp32 (taps, N, C) -> (fp16 (taps, N, 2C) = [hi | lo] of w * 2^e, 2^-e).  The power of two puts max|w| in
    [4096, 8192) so the lo halves (2^-11 of the value) stay normal fp16 numbers."""
    w = w.detach().to(torch.float32).contiguous()
    mx = float(w.abs().max())
    e = 0 if mx == 0.0 or not np.isfinite(mx) else int(np.floor(np.log2(8192.0 / mx)))
    e = max(-24, min(24, e))
    return lib.split_f16(w, float(2.0 ** e)), float(2.0 ** -e)


def layer_cache_x3(layer) -> dict:
    c = getattr(layer, "_ttsb_cache_x3", None)
    if c is None:
        c = Cache()
        object.__setattr__(layer, "_ttsb_cache_x3", c)
    a, ff = layer.attn, layer.ff
    params = [a.q_linear.weight, a.k_linear.weight, a.v_linear.weight, a.out.weight, ff.f_1.weight, ff.f_2.weight]
    if c.stale(params):
        t = c.t
        t["wqkv"], t["sqkv"] = split_weight(torch.cat([a.q_linear.weight, a.k_linear.weight, a.v_linear.weight], 0).unsqueeze(0))
        t["wo"], t["so"] = split_weight(a.out.weight.unsqueeze(0))
        t["w1"], t["s1"] = split_weight(ff.f_1.weight.detach().permute(2, 0, 1))
        t["w2"], t["s2"] = split_weight(ff.f_2.weight.detach().permute(2, 0, 1))
        c.commit(params)
    return c.t


def predictor_cache_x3(pred) -> dict:
    c = getattr(pred, "_ttsb_cache_x3", None)
    if c is None:
        c = Cache()
        object.__setattr__(pred, "_ttsb_cache_x3", c)
    params = [pred.conv1.weight, pred.conv2.weight]
    if c.stale(params):
        c.t["w1"], c.t["s1"] = split_weight(pred.conv1.weight.detach().permute(2, 0, 1))
        c.t["w2"], c.t["s2"] = split_weight(pred.conv2.weight.detach().permute(2, 0, 1))
        c.commit(params)
    return c.t


def run_encoder_precise(enc, src: torch.Tensor, mask_u8: torch.Tensor, kv_len: torch.Tensor):
    """The text encoder (embedding=True) to fp32 accuracy.  Returns (LN_final(x) bf16 (B,T,D) for the frame-level
    stack, the same as an fp16 hi|lo split (B,T,2D) for the duration predictor, and fp32)."""
    if not enc.embedding:
        raise NotImplementedError("run_encoder_precise: the text encoder only (embedding=True)")
    t = enc._cache()
    B, T = src.shape
    D = enc.norm.weight.numel()
    dev = src.device
    pe = enc.pe.table(dev)
    if T > pe.shape[0]:
        raise ValueError(f"sequence length {T} exceeds the positional table ({pe.shape[0]}, modules.py:92)")
    caches = [layer_cache(l) for l in enc.layers]
    xcs = [layer_cache_x3(l) for l in enc.layers]
    ln_first = caches[0]["ln1"] if enc.N > 0 else t["norm"]
    f32 = lambda *s: torch.empty(s, device=dev, dtype=torch.float32)  # noqa: E731
    f16 = lambda *s: torch.empty(s, device=dev, dtype=torch.float16)  # noqa: E731
    x = f32(B, T, D)
    out_bf = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
    lib.embed_pe_ln(src.contiguous(), t["table"], pe, t["alpha"], ln_first[0], ln_first[1], LN_EPS, x, out_bf)
    xs = f16(B, T, 2 * D)
    out32 = f32(B, T, D)
    lib.xrow(lib.XROW_ADD_LN, x, ln_first, out_split=xs, out_f32=out32 if enc.N == 0 else None, eps=LN_EPS)
    H = enc.heads
    for i, (lc, xc) in enumerate(zip(caches, xcs)):
        last = i == enc.N - 1
        k = lc["k"]
        next_ln = t["norm"] if last else caches[i + 1]["ln1"]
        qkv = lib.gemm_x3(xs, xc["wqkv"], lc["bqkv"], xc["sqkv"], split_out=True)       # modules.py:49-51
        ctx = lib.attention_x3(qkv, mask_u8, kv_len, H)                                  # modules.py:7-21,58-65
        o = lib.gemm_x3(ctx, xc["wo"], lc["bo"], xc["so"])                               # modules.py:68
        y, yns = f32(B, T, D), f16(B, T, 2 * D)
        lib.xrow(lib.XROW_ADD_LN, x, lc["ln2"], b0=o, s_out=y, out_split=yns, eps=LN_EPS)  # layers.py:33,35
        hs = lib.gemm_x3(yns, xc["w1"], lc["b1"], xc["s1"], taps=k, pad_left=k // 2, relu=True, split_out=True)
        u = lib.gemm_x3(hs, xc["w2"], lc["b2"], xc["s2"], taps=k, pad_left=k // 2)      # modules.py:83-84
        z, zns = f32(B, T, D), f16(B, T, 2 * D)
        lib.xrow(lib.XROW_FFN_TAIL, u, lc["ln_ff"], b0_split=yns, c=y, ln2=next_ln, z_out=z, out_split=zns,
                 out_bf16=out_bf if last else None, out_f32=out32 if last else None, eps=LN_EPS)  # modules.py:85-87, layers.py:40
        x, xs = z, zns
    return out_bf, xs, out32


def run_predictor_precise(pred, x_split: torch.Tensor, mask_u8: Optional[torch.Tensor]) -> torch.Tensor:
    """VariancePredictor (varianceadaptor.py:216-229) to fp32 accuracy on a split input.  Returns fp32 (B,T)."""
    t = predictor_cache(pred)
    xc = predictor_cache_x3(pred)
    B, T, _ = x_split.shape
    fs = t["b1"].numel()
    dev = x_split.device
    c1 = lib.gemm_x3(x_split, xc["w1"], t["b1"], xc["s1"], taps=3, pad_left=1, relu=True)   # :217-218
    h1 = torch.empty((B, T, 2 * fs), device=dev, dtype=torch.float16)
    lib.xrow(lib.XROW_ADD_LN, c1, t["ln1"], out_split=h1, eps=LN_EPS)                        # :219
    c2 = lib.gemm_x3(h1, xc["w2"], t["b2"], xc["s2"], taps=3, pad_left=1, relu=True)        # :220-221
    out = torch.empty((B, T), device=dev, dtype=torch.float32)
    lib.xrow(lib.XROW_LN_DOT, c2, t["ln2"], dot_w=t["dw"], dot_b=t["db"], row_mask=mask_u8, out_row=out,
             eps=LN_EPS)                                                                      # :222-229
    return out
