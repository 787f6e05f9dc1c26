"""Drop-in for ``Transformer`` of the reference's ``Models/transformer.py:14-118`` (Transformer-TTS; SURVEY §8f row f3:
autoregressive inference with an incremental key / value cache and the stop token, BASELINE.json configs[4]).

Same constructor signature, sub-module names and state_dict layout (321 entries canonical, checked against the reference
module in tests/test_oracle_tts.py).  `forward` is the reference's call — a teacher-forced prefix, the call test.py:141,156
makes at every step — and returns the same 6-tuple (attention maps None).  `generate` is the loop of
test.py:128-191 itself, which the reference has only as script code: one decoder position per step against cached keys
and values instead of a forward over the whole prefix per step.

In train() mode `forward` is the teacher-forced training forward of train.py:197 with autograd
connected to every parameter through one Function (tts_training.autograd_forward: dropout with counter-based masks, backward
on the same kernels); the loop-owning form with flat buffers and the fused clip + Adam is tts_training.TTSTrainer.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .. import lib
from . import _engine as E
from .decoder import Decoder
from .encoder import Encoder
from .postnets import PostConvNet


class Transformer(nn.Module):
    def __init__(self, hp, src_vocab, trg_vocab, d_model_encoder, N_e, n_head_encoder, ff_conv_kernel_size_encoder,
                 concat_after_encoder, d_model_decoder, N_d, n_head_decoder, ff_conv_kernel_size_decoder,
                 concat_after_decoder, reduction_rate, dropout, dropout_prenet=0.5, dropout_postnet=0.5,
                 multi_speaker=False, spk_emb_dim=None, spk_emb_architecture=None, output_type=None,
                 decoder_type='transformer'):
        super().__init__()
        if getattr(hp, "gst", False):
            raise NotImplementedError("GST style embeddings are out of scope (SURVEY §2)")
        if multi_speaker or (spk_emb_architecture and ("encoder" in spk_emb_architecture or "decoder" in spk_emb_architecture)):
            raise NotImplementedError("speaker embeddings are out of scope (SURVEY §2a: default hparams only)")
        if decoder_type.lower() != 'transformer' or getattr(hp, "encoder_type", "transformer").lower() == 'conformer':
            raise NotImplementedError("only the transformer encoder / decoder are built (SURVEY §2)")
        self.gst = False
        self.spk_emb_vers = 1
        self.decoder_type = decoder_type
        self.hp = hp
        self.reduction_rate = reduction_rate
        self.mel_dim = trg_vocab
        self.encoder = Encoder(src_vocab, d_model_encoder, N_e, n_head_encoder, ff_conv_kernel_size_encoder,
                               concat_after_encoder, dropout=dropout, multi_speaker=False, spk_emb_dim=spk_emb_dim)
        self.linear = nn.Linear(d_model_encoder, d_model_decoder) if d_model_encoder != d_model_decoder else None
        self.decoder = Decoder(trg_vocab, d_model_decoder, N_d, n_head_decoder, ff_conv_kernel_size_decoder,
                               concat_after_decoder, dropout=dropout, dropout_prenet=dropout_prenet, multi_speaker=False,
                               spk_emb_dim=spk_emb_dim, output_type=output_type)
        self.out = nn.Linear(d_model_decoder, trg_vocab * reduction_rate)
        self.stop_token = nn.Linear(d_model_decoder, reduction_rate)
        self.postnet = PostConvNet(hp, d_model_decoder, trg_vocab, reduction_rate, dropout_postnet, prev_version=False)
        self._ttsb_cache = E.Cache()
        self.use_cuda_graphs = True
        self.prefix_mode = "batched"   # forward(): "batched" = whole prefix on the tcgen05 kernels; "steps" = decode-step machine

    # ------------------------------------------------------------------------------------------------------------------
    def _head(self) -> dict:
        """mel and stop projections (transformer.py:109,111) as ONE operand: rows [out | stop_token]."""
        c = self._ttsb_cache
        params = [self.out.weight, self.out.bias, self.stop_token.weight, self.stop_token.bias]
        if self.linear is not None:
            params += [self.linear.weight, self.linear.bias]
        if c.stale(params):
            c.t["w"] = lib.pad_weight_rows(torch.cat([self.out.weight.detach(), self.stop_token.weight.detach()], 0))
            c.t["b"] = E._f32(torch.cat([self.out.bias.detach(), self.stop_token.bias.detach()], 0))
            # the same rows zero-padded to a multiple of the 192-column GEMM tile, as a (1, N, C) conv_gemm operand
            n = c.t["w"].shape[0]
            npad = (n + 191) // 192 * 192
            wg = torch.zeros((1, npad, c.t["w"].shape[1]), device=c.t["w"].device, dtype=torch.bfloat16)
            wg[0, :n] = c.t["w"]
            bg = torch.zeros(npad, device=c.t["b"].device, dtype=torch.float32)
            bg[:n] = c.t["b"]
            c.t["w_g"], c.t["b_g"] = wg, bg
            if self.linear is not None:
                c.t["wlin"] = E._bf16(self.linear.weight).unsqueeze(0).contiguous()
                c.t["blin"] = E._f32(self.linear.bias)
            c.commit(params)
        return c.t

    def encode(self, src, src_mask) -> torch.Tensor:
        """transformer.py:92-94: encoder outputs (B, L, d_model_decoder), bf16 (the operand precision of the k|v
        projections that consume them)."""
        B, L = src.shape
        m = E.mask_to_u8(src_mask, B, L, src.device)
        kv = lib.mask_lengths(m)
        e_bf16, _, _ = self.encoder.run(src, m, kv)
        if self.linear is not None:
            t = self._head()
            out = torch.empty((B, L, self.linear.out_features), device=src.device, dtype=torch.bfloat16)
            lib.conv_gemm(e_bf16, t["wlin"], t["blin"], block_n=128, out_bf16=out)
            e_bf16 = out
        return e_bf16

    def _forward_train(self, src, trg, src_mask, trg_mask):
        """train.py:197: the same 6-tuple with autograd attached.  The three attention maps, which train.py:198-200 moves to
        the host and nothing reads (the image logging is commented out), are EMPTY tensors (not materialised)."""
        from ..tts_training import autograd_forward

        if trg_mask is None:
            raise NotImplementedError("the training forward needs the target mask (train.py:195)")
        tm = trg_mask if trg_mask.dim() == 3 else trg_mask.unsqueeze(1)
        key_valid = tm[:, -1, :].to(torch.uint8).contiguous()           # last row of causal & padding = the padding mask
        kvt = lib.mask_lengths(key_valid)
        if not torch.equal(kvt[:, 0], kvt[:, 1]):
            raise NotImplementedError("the target mask must be causal AND a padding mask (train.py:26-58)")
        pre, stop = autograd_forward(self, src, trg, src_mask, key_valid)
        if self.reduction_rate == 1:
            stop = stop.squeeze(2)
        none = pre.new_empty(0)
        return pre, pre, stop, none, none, none

    def _split(self, out_all: torch.Tensor, n: int):
        """(R, cap, mel*r + r) -> outputs (R, n, mel*r), stop (R, n, r) -> the reference's shapes (transformer.py:109-111)."""
        md = self.mel_dim * self.reduction_rate
        pre = out_all[:, :n, :md]
        stop = out_all[:, :n, md:md + self.reduction_rate]   # (the row pitch is padded to a multiple of 4 floats)
        if self.reduction_rate == 1:
            stop = stop.squeeze(2)
        return pre, stop

    def forward(self, src, trg, src_mask, trg_mask, spkr_emb=None, training=True, ref_mel=None):
        """The reference's call (transformer.py:90-118) on a teacher-forced prefix: trg (B, T, mel_dim) decoder inputs,
        trg_mask (B, T, T) = causal AND key padding (train.py:26-58).  Returns (outputs_prenet, outputs_postnet,
        stop_token, None, None, None): `PostConvNet(prev_version=False)` returns its input (postnets.py:64-79), so the
        two mel outputs are the same tensor values; the attention maps are not materialised.

        prefix_mode "batched" (default): the whole prefix at once on the tensor-core kernels (Decoder.run_prefix: tcgen05
        GEMMs, causal tcgen05 attention, encoder-decoder attention on a packed sequence, fused FeedForward);
        "steps": T decode steps of the generate() machine (each position sees the cached keys / values of the earlier ones)."""
        E.require_cuda(src, "Transformer")
        if spkr_emb is not None or ref_mel is not None:
            raise NotImplementedError("speaker / reference-mel inputs are out of scope")
        if self.training:
            return self._forward_train(src, trg, src_mask, trg_mask)
        B, T, _ = trg.shape
        trg_len = key_valid = None
        if trg_mask is not None:
            tm = trg_mask if trg_mask.dim() == 3 else trg_mask.unsqueeze(1)
            key_valid = tm[:, -1, :].to(torch.uint8).contiguous()          # last row of causal & padding = the padding mask
            kvt = lib.mask_lengths(key_valid)
            if not torch.equal(kvt[:, 0], kvt[:, 1]):
                raise NotImplementedError("the target mask must be causal AND a padding mask (train.py:26-58)")
            trg_len = kvt[:, 0].contiguous()
        e = self.encode(src, src_mask)
        if self.prefix_mode == "batched" and self.decoder.norm.weight.numel() == 384:
            # the whole prefix at once on the tensor-core kernels (Decoder.run_prefix)
            key_valid = (torch.ones((B, T), device=trg.device, dtype=torch.uint8) if trg_mask is None else key_valid)
            out_all = self.decoder.run_prefix(self._head(), trg.float(), e, src_mask, key_valid)
        else:  # position by position through the decode-step machine (what generate() uses)
            st = self.decoder.begin(self._head(), e, src_mask, cap=T, forced_in=trg.float().contiguous(), trg_len=trg_len,
                                    use_cuda_graph=self.use_cuda_graphs and T > 2)
            st.run(T)
            out_all = st.out_all
        pre, stop = self._split(out_all, T)
        pre = pre.contiguous()
        return pre, self.postnet(pre), stop.contiguous(), None, None, None

    @torch.no_grad()
    def generate(self, text, src_mask, max_steps: int, stop_threshold: float = 0.5, check_every: int = 0):
        """The autoregressive loop of test.py:128-191, batched: position 0 is fed zeros (:136), position p > 0 the first
        of the r frames predicted at p - 1 (:165 `outputs_postnet[:, ::r, :]`); a position's stop probability is the mean
        over its r frames (:168-171), tested from position 1 on (:149 starts at i = 2).

        Returns dict(mel (B, max_steps * r, mel_dim), stop_logits (B, max_steps, r), stop_step (B,) = first position whose
        stop probability exceeds the threshold (max_steps if none), steps = positions actually decoded).
        check_every > 0: every that many positions the host reads the stop flags and ends the loop once EVERY utterance
        has stopped (the script, batch 1, breaks at once); 0 = run max_steps positions without a host sync."""
        if self.training:
            raise NotImplementedError("generate() runs the inference kernels: call .eval()")
        E.require_cuda(text, "Transformer.generate")
        e = self.encode(text, src_mask)
        st = self.decoder.begin(self._head(), e, src_mask, cap=max_steps, use_cuda_graph=self.use_cuda_graphs)
        done = 0
        chunk = check_every if check_every > 0 else max_steps
        while done < max_steps:
            n = min(chunk, max_steps - done)
            st.run(n)
            done += n
            if check_every > 0 and bool((self._stop_steps(st.out_all, done, stop_threshold) < done).all()):
                break
        r, md = self.reduction_rate, self.mel_dim * self.reduction_rate
        mel = st.out_all[:, :done, :md].reshape(text.shape[0], done * r, self.mel_dim)
        stop = st.out_all[:, :done, md:md + r].contiguous()
        self.last_state = st
        return dict(mel=mel, stop_logits=stop, stop_step=self._stop_steps(st.out_all, done, stop_threshold), steps=done)

    def _stop_steps(self, out_all: torch.Tensor, n: int, thr: float) -> torch.Tensor:
        md = self.mel_dim * self.reduction_rate
        p = torch.sigmoid(out_all[:, :n, md:md + self.reduction_rate]).mean(dim=-1) > thr
        p[:, 0] = False
        idx = torch.arange(n, device=out_all.device).expand_as(p)
        return torch.where(p, idx, torch.full_like(idx, n)).min(dim=1).values
