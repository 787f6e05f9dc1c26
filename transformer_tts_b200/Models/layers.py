"""Drop-in for ``EncoderLayer`` of the reference's ``Models/layers.py:8-41``."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import lib
from . import _engine as E
from .modules import FeedForward, MultiHeadAttention


class EncoderLayer(nn.Module):
    def __init__(self, d_model, heads, ff_conv_kernel_size, dropout=0.1, concat_after=False, multi_speaker=False,
                 spk_emb_dim=None):
        super().__init__()
        if multi_speaker:
            raise NotImplementedError("multi_speaker FFT blocks are out of scope (SURVEY §2a: default hparams only)")
        self.norm_1 = nn.LayerNorm(d_model)
        self.norm_2 = nn.LayerNorm(d_model)
        self.attn = MultiHeadAttention(heads, d_model, d_model, d_model, d_model, dropout=dropout,
                                       concat_after=concat_after)
        self.ff = FeedForward(d_model, ff_conv_kernel_size, dropout=dropout)
        self.dropout_1 = nn.Dropout(dropout)
        self.dropout_2 = nn.Dropout(dropout)
        self.multi_speaker = multi_speaker
        self._ttsb_cache = E.Cache()

    def forward(self, x, mask, spkr_emb=None, return_attn=True):
        """x: (B,T,D) fp32 -> (x', attn) as layers.py:29-41.  Standalone entry: norm_1 runs as its own
        kernel and the block's tail LayerNorm slot is fed the identity; inside Encoder the norms are
        fused into the neighbouring GEMM epilogues."""
        E.require_inference(self, "EncoderLayer")
        E.require_cuda(x, "EncoderLayer")
        if spkr_emb is not None:
            raise NotImplementedError("speaker embeddings are out of scope")
        t = E.layer_cache(self)
        B, T, D = x.shape
        x = x.float().contiguous()
        m = E.mask_to_u8(mask, B, T, x.device)
        kv = lib.mask_lengths(m)
        xn = torch.empty((B, T, D), device=x.device, dtype=torch.bfloat16)
        lib.layernorm(x, t["ln1"][0], t["ln1"][1], E.LN_EPS, y_bf16=xn)
        ident = (torch.ones(D, device=x.device), torch.zeros(D, device=x.device))
        z, _, _ = E.run_fft_block(t, self.attn.h, x, xn, m, kv, ident)
        attn = None
        if return_attn:
            qkv = torch.empty((B, T, 3 * D), device=x.device, dtype=torch.bfloat16)
            lib.conv_gemm(xn, t["wqkv"], t["bqkv"], block_n=128, out_bf16=qkv)
            attn = lib.attention_probs(qkv, m, self.attn.h)
        return z, attn


class DecoderLayer(nn.Module):
    """Parameter holder for ``DecoderLayer`` of the reference's ``Models/layers.py:84-124`` (causal self-attention,
    encoder-decoder attention, FeedForward, pre-LayerNorm residual blocks).  The kernels run from ``Decoder``
    (Models/decoder.py), one decoder position at a time against cached keys / values."""

    def __init__(self, d_model, heads, ff_conv_kernel_size, dropout=0.1, concat_after=False, multi_speaker=False,
                 spk_emb_dim=None):
        super().__init__()
        if multi_speaker:
            raise NotImplementedError("multi_speaker decoder blocks are out of scope (SURVEY §2a: default hparams only)")
        if concat_after:
            raise NotImplementedError("concat_after=True is not supported (hparams default is False)")
        self.norm_1 = nn.LayerNorm(d_model)
        self.norm_2 = nn.LayerNorm(d_model)
        self.norm_3 = nn.LayerNorm(d_model)
        self.dropout_1 = nn.Dropout(dropout)
        self.dropout_2 = nn.Dropout(dropout)
        self.dropout_3 = nn.Dropout(dropout)
        self.attn_1 = MultiHeadAttention(heads, d_model, d_model, d_model, d_model, dropout=dropout, concat_after=concat_after)
        self.attn_2 = MultiHeadAttention(heads, d_model, d_model, d_model, d_model, dropout=dropout, concat_after=concat_after)
        self.ff = FeedForward(d_model, ff_conv_kernel_size, dropout=dropout)
        self.multi_speaker = multi_speaker

    def forward(self, x, e_outputs, src_mask, trg_mask, spkr_emb=None):
        raise NotImplementedError("DecoderLayer runs inside Decoder (Models/decoder.py): its self-attention needs the "
                                  "key / value cache of the earlier positions")
