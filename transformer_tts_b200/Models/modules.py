"""Drop-in for the hot-path classes of the reference's ``Models/modules.py``:
``attention`` (:7-21), ``MultiHeadAttention`` (:23-70), ``FeedForward`` (:72-88),
``PositionalEncoder`` (:90-111).  Same constructor signatures, parameter names and return values;
``forward`` runs the sm_100a kernels of ``csrc/`` (no PyTorch compute ops, no CPU path).

Dropout: the reference's attention dropout fires even in eval(), so parity is only
defined at dropout 0 — which is what its inference script passes.  Non-zero dropout is accepted by
the constructors (the training script passes 0.1) but the inference kernels ignore it.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .. import lib
from . import _engine as E


def attention(q, k, v, d_k, mask, dropout, debug=False):
    """(B,H,T,d_k) q/k/v, mask (B,1,T) -> (output (B,H,T,d_k), probabilities (B,H,T,T)).

    Standalone form of modules.py:7-21.  The model itself calls the fused tcgen05 kernel on the
    packed q|k|v projection; this wrapper packs its arguments the same way and additionally
    materialises the probabilities with the debug kernel, as the reference returns them."""
    E.require_cuda(q, "attention")
    B, H, T, Dh = q.shape
    if dropout:
        raise NotImplementedError("attention dropout > 0 is not supported by the inference kernels (SURVEY F4)")
    qkv = torch.cat([t.transpose(1, 2).reshape(B, T, H * Dh) for t in (q, k, v)], dim=-1).to(torch.bfloat16).contiguous()
    m = E.mask_to_u8(mask, B, T, q.device)
    kv = lib.mask_lengths(m)
    ctx = torch.empty((B, T, H * Dh), device=q.device, dtype=torch.bfloat16)
    lib.attention_fwd(qkv, m, kv, ctx, H)
    probs = lib.attention_probs(qkv, m, H)
    return ctx.view(B, T, H, Dh).transpose(1, 2).float(), probs


class MultiHeadAttention(nn.Module):
    def __init__(self, heads, q_dim, k_dim, v_dim, d_model, dropout=0.1, concat_after=False):
        super().__init__()
        self.d_model = d_model
        self.d_k = d_model // heads
        self.h = heads
        self.q_linear = nn.Linear(q_dim, d_model)
        self.v_linear = nn.Linear(k_dim, d_model)  # the reference swaps k_dim/v_dim (modules.py:33-34)
        self.k_linear = nn.Linear(v_dim, d_model)
        self.dropout = dropout
        self.concat_after = concat_after
        if self.concat_after:
            self.out = nn.Linear(2 * d_model, d_model)
        else:
            self.out = nn.Linear(d_model, d_model)
        self._ttsb_cache = E.Cache()

    def _cache(self):
        c = self._ttsb_cache
        params = [self.q_linear.weight, self.q_linear.bias, self.k_linear.weight, self.k_linear.bias,
                  self.v_linear.weight, self.v_linear.bias, self.out.weight, self.out.bias]
        if c.stale(params):
            if self.concat_after:
                raise NotImplementedError("concat_after=True is not supported")
            c.t["wqkv"] = E._bf16(torch.cat([self.q_linear.weight, self.k_linear.weight, self.v_linear.weight], 0)).unsqueeze(0).contiguous()
            c.t["bqkv"] = E._f32(torch.cat([self.q_linear.bias, self.k_linear.bias, self.v_linear.bias], 0))
            c.t["wo"] = E._bf16(self.out.weight).unsqueeze(0).contiguous()
            c.t["bo"] = E._f32(self.out.bias)
            c.commit(params)
        return c.t

    def forward(self, q, k, v, mask=None, debug=False, return_attn=True):
        """Self-attention only (q is k is v), as every FastSpeech2 call site uses it (layers.py:32)."""
        E.require_inference(self, "MultiHeadAttention")
        E.require_cuda(q, "MultiHeadAttention")
        if not (q is k and k is v):
            raise NotImplementedError("cross-attention is outside the FastSpeech2 path (SURVEY §8f f2)")
        t = self._cache()
        B, T, D = q.shape
        x = q.to(torch.bfloat16).contiguous()
        m = E.mask_to_u8(mask, B, T, q.device)
        kv = lib.mask_lengths(m)
        qkv = torch.empty((B, T, 3 * D), device=q.device, dtype=torch.bfloat16)
        lib.conv_gemm(x, t["wqkv"], t["bqkv"], block_n=128, out_bf16=qkv)
        ctx = torch.empty((B, T, D), device=q.device, dtype=torch.bfloat16)
        lib.attention_fwd(qkv, m, kv, ctx, self.h)
        out = torch.empty((B, T, D), device=q.device, dtype=torch.float32)
        lib.conv_gemm(ctx, t["wo"], t["bo"], block_n=128, out_f32=out)
        attn = lib.attention_probs(qkv, m, self.h) if return_attn else None
        return out, attn


class FeedForward(nn.Module):
    def __init__(self, d_model, ff_conv_kernel_size, dropout=0.1):
        super().__init__()
        self.f_1 = nn.Conv1d(d_model, d_model * 4, kernel_size=ff_conv_kernel_size, padding=int(ff_conv_kernel_size / 2))
        self.dropout = nn.Dropout(dropout)
        self.f_2 = nn.Conv1d(d_model * 4, d_model, kernel_size=ff_conv_kernel_size, padding=int(ff_conv_kernel_size / 2))
        self.layer_norm = nn.LayerNorm(d_model)
        self._ttsb_cache = E.Cache()

    def _cache(self):
        c = self._ttsb_cache
        params = [self.f_1.weight, self.f_1.bias, self.f_2.weight, self.f_2.bias, self.layer_norm.weight,
                  self.layer_norm.bias]
        if c.stale(params):
            c.t["w1"], c.t["b1"] = E.conv_w_to_taps(self.f_1.weight), E._f32(self.f_1.bias)
            c.t["w2"], c.t["b2"] = E.conv_w_to_taps(self.f_2.weight), E._f32(self.f_2.bias)
            c.t["ln"] = (E._f32(self.layer_norm.weight), E._f32(self.layer_norm.bias))
            c.commit(params)
        return c.t

    def forward(self, x):
        """LN(conv2(relu(conv1(x))) + x) (modules.py:81-88).  Standalone: the fused tail also wants a
        second LayerNorm, so the identity affine is passed for it and the first LN's output is taken
        from the fp32 stream: z = 0 + LN(v) -> out_f32."""
        E.require_inference(self, "FeedForward")
        E.require_cuda(x, "FeedForward")
        t = self._cache()
        B, T, D = x.shape
        k = self.f_1.kernel_size[0]
        dev = x.device
        xb = x.to(torch.bfloat16).contiguous()
        h = torch.empty((B, T, 4 * D), device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(xb, t["w1"], t["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU, out_bf16=h)
        zeros = torch.zeros((T, D), device=dev, dtype=torch.float32)
        ident = (torch.ones(D, device=dev), torch.zeros(D, device=dev))
        out = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        scratch = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(h, t["w2"], t["b2"], taps=k, pad_left=k // 2, block_n=D, epi=lib.EPI_FFN2, res_bf16=xb,
                      res_f32=zeros, res_bstride=0, res_tstride=D, ln1=t["ln"], ln2=ident, out_f32=out,
                      out_bf16=scratch)
        return out


class PositionalEncoder(nn.Module):
    def __init__(self, d_model, max_seq_len=5000, dropout=0.1):
        super().__init__()
        self.d_model = d_model
        self.max_seq_len = max_seq_len
        self.dropout = nn.Dropout(dropout)
        self.alpha = nn.Parameter(torch.ones(1))
        # The reference builds a (1, max_seq_len, d_model) table with a Python loop per instance
        # (modules.py:97-105); it is not a buffer and not in the state_dict.  Here it is built once,
        # vectorised, and shared per (d_model, device).
        self.pe = E.positional_table_cpu(d_model, max_seq_len).unsqueeze(0)

    def table(self, device) -> torch.Tensor:
        return E.positional_table(self.d_model, device, self.max_seq_len)

    def forward(self, x):
        """x + alpha * pe[:T] (modules.py:107-111); elementwise glue kept in torch for the standalone
        module — inside Encoder it is fused into the embedding / input-linear kernel."""
        E.require_cuda(x, "PositionalEncoder")
        seq_len = x.shape[1]
        return x + self.alpha * self.table(x.device)[:seq_len].unsqueeze(0)
