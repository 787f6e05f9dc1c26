"""Drop-in for ``Encoder`` of the reference's ``Models/encoder.py:31-112`` (the FFT stack used both as
the phoneme encoder, ``embedding=True``, and as the FastSpeech2 "decoder", ``embedding=False``)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import lib
from . import _engine as E
from .layers import EncoderLayer
from .modules import PositionalEncoder


class MultiSequential(torch.nn.Sequential):
    """Multi-input multi-output Sequential (same container class as the reference, encoder.py:9-16)."""

    def forward(self, *args):
        for m in self:
            args = m(*args)
        return args


def repeat(N, fn):
    return MultiSequential(*[fn() for _ in range(N)])


class Encoder(nn.Module):
    def __init__(self, vocab_size, d_model, N, heads, ff_conv_kernel_size, concat_after_encoder, dropout,
                 multi_speaker=False, spk_emb_dim=None, embedding=True, accent_emb=False, spk_emb_layer=None,
                 gender_emb=False, intermediate_layers_out=None):
        super().__init__()
        if multi_speaker or accent_emb or gender_emb or spk_emb_layer is not None or intermediate_layers_out:
            raise NotImplementedError("speaker/accent/gender embeddings and intermediate outputs are out of scope "
                                      "(SURVEY §2a: default hparams only)")
        self.N = N
        self.heads = heads
        self.accent_emb_flag = accent_emb
        self.gender_emb_flag = gender_emb
        self.intermediate_layers_out = intermediate_layers_out
        self.embedding = embedding
        if embedding:
            self.embed = nn.Embedding(vocab_size, d_model, padding_idx=0)
        else:
            self.embed = nn.Linear(vocab_size, d_model)
        self.pe = PositionalEncoder(d_model, dropout=dropout)
        self.layers = repeat(N, lambda: EncoderLayer(d_model, heads, ff_conv_kernel_size, dropout=dropout,
                                                     concat_after=concat_after_encoder, multi_speaker=multi_speaker,
                                                     spk_emb_dim=spk_emb_dim))
        self.norm = nn.LayerNorm(d_model)
        self._ttsb_cache = E.Cache()
        self.return_attn = False  # F10: the (b,N,H,t,t) maps are only materialised on request

    def _cache(self):
        c = self._ttsb_cache
        params = [self.embed.weight, self.norm.weight, self.norm.bias, self.pe.alpha]
        if not self.embedding:
            params.append(self.embed.bias)
        if c.stale(params):
            t = c.t
            t["norm"] = (E._f32(self.norm.weight), E._f32(self.norm.bias))
            t["alpha"] = E._f32(self.pe.alpha)
            if self.embedding:
                t["table"] = E._f32(self.embed.weight)
            else:
                t["wemb"] = E._bf16(self.embed.weight).unsqueeze(0).contiguous()
                t["bemb"] = E._f32(self.embed.bias)
            c.commit(params)
        return c.t

    def run(self, src, mask_u8, kv_len, want_f32=False, return_attn=False, tile_map=None):
        """Kernel sequence for the whole stack.  src: int64 ids (B,T) or bf16 features (B,T,C).
        Returns (LN_final(x) bf16, optional fp32 copy, attn maps or None)."""
        t = self._cache()
        B, T = src.shape[0], src.shape[1]
        D = self.norm.weight.numel()
        dev = src.device
        pe = self.pe.table(dev)
        if T > pe.shape[0]:
            raise ValueError(f"sequence length {T} exceeds the positional table ({pe.shape[0]}, modules.py:92)")
        caches = [E.layer_cache(l) for l in self.layers]
        x = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        xn = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
        ln_first = caches[0]["ln1"] if self.N > 0 else t["norm"]
        if self.embedding:
            lib.embed_pe_ln(src.contiguous(), t["table"], pe, t["alpha"], ln_first[0], ln_first[1], E.LN_EPS, x, xn)
        else:
            lib.conv_gemm(src, t["wemb"], t["bemb"], block_n=D, epi=lib.EPI_RES_LN, res_f32=pe, res_bstride=0,
                          res_tstride=D, res_scale=t["alpha"], ln1=ln_first, out_f32=x, out_bf16=xn,
                          tile_map=tile_map)
        attns = [] if return_attn else None
        xn32 = None
        for i, lc in enumerate(caches):
            last = i == self.N - 1
            if return_attn:
                qkv = torch.empty((B, T, 3 * D), device=dev, dtype=torch.bfloat16)
                lib.conv_gemm(xn, lc["wqkv"], lc["bqkv"], block_n=128, out_bf16=qkv)
                attns.append(lib.attention_probs(qkv, mask_u8, self.heads))
            next_ln = t["norm"] if last else caches[i + 1]["ln1"]
            x, xn, xn32 = E.run_fft_block(lc, self.heads, x, xn, mask_u8, kv_len, next_ln,
                                          want_norm_f32=(last and want_f32), tile_map=tile_map, zero_dropped=last)
        if self.N == 0 and want_f32:
            xn32 = xn.float()
        return xn, xn32, (torch.stack(attns, 1) if return_attn else None)

    def forward(self, src, mask, spkr_emb=None, accent=None, gender_id=None, attn_detach=True):
        """(norm(x), attns) as encoder.py:83-112.  attns is None unless ``self.return_attn`` is set
        (no live caller reads the 6 GB of maps)."""
        E.require_inference(self, "Encoder")
        E.require_cuda(src, "Encoder")
        if spkr_emb is not None or accent is not None or gender_id is not None:
            raise NotImplementedError("speaker/accent/gender inputs are out of scope")
        B, T = src.shape[0], src.shape[1]
        m = E.mask_to_u8(mask, B, T, src.device)
        kv = lib.mask_lengths(m)
        if not self.embedding:
            src = src.to(torch.bfloat16).contiguous()
        _, out32, attns = self.run(src, m, kv, want_f32=True, return_attn=self.return_attn)
        return out32, attns
