"""Drop-in for ``FastSpeech2`` of the reference's ``Models/fastspeech2.py:19-241``.

Same 32-argument constructor, same ``forward`` signature, same 14-tuple, same ``state_dict`` keys
(289 entries with the canonical hparams).  ``forward`` chains the sm_100a kernels:

    embed+PE+LN -> 6 x [QKV GEMM, attention, out-proj GEMM(+res+LN), FFN1 GEMM(+ReLU), FFN2 GEMM(+LN+res+LN)]
    -> duration predictor (2 GEMMs) -> round -> LengthRegulator (scan + expand)
    -> pitch / energy predictors (4 GEMMs) -> bucketize+embedding add
    -> decoder input GEMM(+PE+LN) -> 6 x FFT block -> PostNet (7 GEMMs)

Research branches of the reference (SQ-VAE, hop embedding, debug post-model, conformer, speaker
embeddings, fix_mask, perturbation) raise NotImplementedError instead of silently differing.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import lib
from . import _engine as E
from .encoder import Encoder
from .postnets import PostConvNet
from .varianceadaptor import VarianceAdaptor


class FastSpeech2(nn.Module):
    def __init__(self, hp, src_vocab, trg_vocab, d_model_encoder, N_e, n_head_encoder, ff_conv_kernel_size_encoder,
                 concat_after_encoder, d_model_decoder, N_d, n_head_decoder, ff_conv_kernel_size_decoder,
                 concat_after_decoder, reduction_rate, dropout, dropout_postnet, dropout_variance_adaptor,
                 n_bins, f0_min, f0_max, energy_min, energy_max, pitch_pred=True, energy_pred=True, accent_emb=False,
                 output_type=None, num_group=None, log_offset=1., multi_speaker=False, spk_emb_dim=None,
                 spk_emb_architecture=None, debug=False):
        super().__init__()
        self.hp = hp
        self.spk_emb_architecture = spk_emb_architecture if spk_emb_architecture is not None else ''

        def _off(name, default=False):
            return getattr(hp, name, default)

        unsupported = []
        if str(_off('encoder_type', 'transformer')).lower() != 'transformer':
            unsupported.append("encoder_type != 'transformer'")
        if str(_off('decoder_type', 'transformer')).lower() != 'transformer':
            unsupported.append("decoder_type != 'transformer'")
        if _off('use_sq_vae'):
            unsupported.append('use_sq_vae')
        if _off('use_hop'):
            unsupported.append('use_hop')
        if debug:
            unsupported.append('debug post-model')
        if multi_speaker or self.spk_emb_architecture:
            unsupported.append('speaker embeddings')
        if accent_emb:
            unsupported.append('accent_emb')
        if not _off('postnet_pred', True):
            unsupported.append('postnet_pred=False')
        if unsupported:
            raise NotImplementedError("FastSpeech2 (B200): out-of-scope options requested: " + ", ".join(unsupported))

        self.encoder = Encoder(src_vocab, d_model_encoder, N_e, n_head_encoder, ff_conv_kernel_size_encoder,
                               concat_after_encoder, dropout, multi_speaker=False, spk_emb_dim=None,
                               accent_emb=accent_emb)
        self.use_sq_vae = False
        self.variance_adaptor = VarianceAdaptor(d_model_encoder, n_bins, f0_min, f0_max, energy_min, energy_max,
                                                log_offset, pitch_pred, energy_pred, dropout=dropout_variance_adaptor,
                                                use_rnn_length=_off('use_rnn_length'), use_pos=_off('use_pos'))
        self.decoder = Encoder(d_model_encoder, d_model_decoder, N_d, n_head_decoder, ff_conv_kernel_size_decoder,
                               concat_after_decoder, dropout, multi_speaker=False, spk_emb_dim=None, embedding=False,
                               accent_emb=False)
        self.postnet = PostConvNet(hp=hp, num_hidden=d_model_decoder, mel_dim=trg_vocab, reduction_rate=reduction_rate,
                                   dropout=dropout_postnet)
        self.debug = debug
        # Free-running inference decides the durations by ROUNDING log_d (varianceadaptor.py:75).  "fp32": the encoder
        # and the duration predictor run the fp32-accurate tensor-core chain (csrc/precise.cu) so the durations, frame
        # indices and mel lengths equal the fp32 reference's; "bf16": the plain bf16 chain (faster, ~5 % of the
        # durations land on the other side of a .5 boundary).  Teacher-forced calls always take the bf16
        # chain: there log_d is only an output.
        self.duration_precision = "fp32"
        # "fp32": the pitch / energy predictors run the fp32-accurate chain as well (needs duration_precision "fp32"), so
        # the pitch / energy buckets — the other discrete decision of a free-running forward, worth ~2 % of mel rel-L2
        # per flipped bucket — equal the fp32 reference's too.  Costs ~0.2 ms at cfg2; off by default.
        self.variance_precision = "bf16"
        self.aux_outputs_fp32 = True   # tuple slots 5 / 6 in the reference's dtype (see _aux_f32)
        self.use_cuda_graphs = True    # free-running inference: the pre-LengthRegulator stage replays a CUDA graph
        self._graphs = {}
        self.return_attn = False   # materialise the (B,N,H,T,T) maps of slots 7/8 (parity tests only)
        self.last_aux = None       # integer by-products of the last forward (durations, frame indices, ...)

    def forward(self, src, src_mask, mel_mask=None, d_target=None, p_target=None, e_target=None, accent=None,
                spkr_emb=None, fix_mask=None, spkr_emb_post=None, temperature=None, pitch_perturbation=False,
                duration_perturbation=False, hop_size=None):
        E.require_cuda(src, "FastSpeech2")
        # Arguments the reference itself ignores in this configuration are ignored here too (they are passed by
        # train_fastspeech2.py:167 on every call): `hop_size` is read only under hp.use_hop (fastspeech2.py:168-172;
        # collate_fn hands over a list of None otherwise), `temperature` only under hp.use_sq_vae (:142-166), `accent`
        # only by an Encoder built with accent_emb (encoder.py:106-108), `spkr_emb` only by multi-speaker layers
        # (layers.py:36-39) and the 'middle' speaker projection (:138-140), `spkr_emb_post` only by the debug
        # post-model (:201-236) — all of which the constructor rejects.
        if fix_mask is not None:
            raise NotImplementedError("fix_mask (band-limited attention, fastspeech2.py:124-135) is out of scope")
        if pitch_perturbation or duration_perturbation:
            assert not self.training, "fastspeech2.py:120"
            raise NotImplementedError("pitch / duration perturbation (varianceadaptor.py:76-81,103-107) is out of scope")
        if self.training:
            from ..training import autograd_forward

            mel_pre, mel_post, log_d, p, e, va_x, text_dur = autograd_forward(self, src, src_mask, mel_mask, d_target,
                                                                              p_target, e_target)
            return (mel_pre, mel_post, log_d, p, e, self._aux_f32(va_x), self._aux_f32(text_dur), None, None, None,
                    None, None, None, None)
        B, L = src.shape
        dev = src.device
        sm = E.mask_to_u8(src_mask, B, L, dev)
        kv_src_fn = lambda: lib.mask_lengths(sm)  # noqa: E731
        if self.duration_precision not in ("fp32", "bf16"):
            raise ValueError(f"duration_precision must be 'fp32' or 'bf16', got {self.duration_precision!r}")
        mm = None if mel_mask is None else E.mask_to_u8(mel_mask, B, mel_mask.shape[-1], dev)
        attn_enc = None
        if d_target is None and not self.return_attn and self.encoder.N > 0:
            # free-running: encoder -> duration predictor -> rounding -> prefix sum, up to the one host sync (Tmax)
            if self.variance_precision not in ("fp32", "bf16"):
                raise ValueError(f"variance_precision must be 'fp32' or 'bf16', got {self.variance_precision!r}")
            if self.variance_precision == "fp32" and self.duration_precision != "fp32":
                raise ValueError("variance_precision 'fp32' needs duration_precision 'fp32' (the fp32-accurate encoder)")
            e_bf, e_split, pre = self._pre_lr(src.contiguous(), sm)
            va = self.variance_adaptor.run(e_bf, sm, mm, None, p_target, e_target, want_aux=True, pre=pre,
                                           x_split=e_split if self.variance_precision == "fp32" else None,
                                           aux_f32=self.aux_outputs_fp32)  # :176-181
        else:
            e_bf, _, attn_enc = self.encoder.run(src, sm, kv_src_fn(), want_f32=False, return_attn=self.return_attn)  # :137
            va = self.variance_adaptor.run(e_bf, sm, mm, d_target, p_target, e_target, want_aux=True,
                                           aux_f32=self.aux_outputs_fp32)  # :176-181
        mel_u8 = va["mel_mask"]
        kv_mel = lib.mask_lengths(mel_u8)
        tile_map = None
        Tm = mel_u8.shape[1]
        if E.RAGGED_TILES and not self.return_attn and self.decoder.N > 0 and \
                B * ((Tm + 127) // 128) > E.SHORT_BATCH_TILES and \
                all(l.ff.f_1.kernel_size[0] == 1 for l in self.decoder.layers):  # short batches keep their own path
            tile_map = lib.build_tile_map(kv_mel, mel_u8.shape[1])  # padding-only row tiles are not computed
        d_bf, _, attn_dec = self.decoder.run(va["x"], mel_u8, kv_mel, want_f32=False, return_attn=self.return_attn,
                                             tile_map=tile_map)  # :195
        mel_pre, mel_post = self.postnet.run(d_bf, tile_map=tile_map)  # :197
        self.last_aux = dict(dur=va["dur"], mel_len=va["mel_len"], frame_idx=va["frame_idx"], mel_pos=va["mel_pos"],
                             mel_mask=mel_u8, p_idx=va["p_idx"], e_idx=va["e_idx"], encoder_output=e_bf,
                             decoder_output=d_bf, va_x=va["x"], text_dur=va["text_dur"])
        va_x = va["x_f32"] if va["x_f32"] is not None else va["x"]
        text_dur = va["text_dur_f32"] if va["text_dur_f32"] is not None else va["text_dur"]
        return (mel_pre, mel_post, va["log_d"], va["p"], va["e"], va_x, text_dur, attn_enc, attn_dec, None, None, None,
                None, None)  # fastspeech2.py:241

    # ------------------------------------------------------------------------------------------
    # Everything before the LengthRegulator's host sync has shapes that depend on (B, L) only, so it is captured ONCE
    # per (B, L, duration_precision) as a CUDA graph and replayed: ~50 kernel launches (fp32-accurate chain) become
    # one, without launch gaps on the device.  Replays are bit-identical to the eager sequence (same kernels, same
    # operands; the TMA descriptors are kernel parameters and live in the graph).
    # ------------------------------------------------------------------------------------------
    def _pre_lr_eager(self, src, sm):
        kv_src = lib.mask_lengths(sm)
        log_d = e_split = None
        if self.duration_precision == "fp32":
            e_bf, e_split, _ = E.run_encoder_precise(self.encoder, src, sm, kv_src)  # :137, to fp32 accuracy
            log_d = E.run_predictor_precise(self.variance_adaptor.duration_predictor, e_split, sm)  # varianceadaptor.py:69
        else:
            e_bf, _, _ = self.encoder.run(src, sm, kv_src, want_f32=False)  # :137
        return e_bf, e_split, self.variance_adaptor.run_pre(e_bf, sm, log_d)

    def _pre_lr_params(self):
        ps = list(self.encoder.parameters()) + list(self.variance_adaptor.duration_predictor.parameters())
        return ps

    def _pre_lr(self, src, sm):
        if not self.use_cuda_graphs or torch.cuda.is_current_stream_capturing():
            return self._pre_lr_eager(src, sm)
        B, L = src.shape
        key = (B, L, self.duration_precision, src.device.index)
        sig = E._sig(self._pre_lr_params())
        ent = self._graphs.get(key)
        if ent is None or ent["sig"] != sig:
            static_src, static_sm = src.clone(), sm.clone()
            self._pre_lr_eager(static_src, static_sm)  # builds the derived-weight caches, first-call kernel attributes
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = lib.LAUNCH_COUNT
            with torch.cuda.graph(g):
                outs = self._pre_lr_eager(static_src, static_sm)
            n_kernels = lib.LAUNCH_COUNT - n0  # kernels of ours inside one replay (bench.py's gpu_launches)
            if len(self._graphs) >= 8:
                self._graphs.pop(next(iter(self._graphs)))
            ent = dict(sig=sig, graph=g, src=static_src, sm=static_sm, outs=outs, n=n_kernels)
            self._graphs[key] = ent
        ent["src"].copy_(src)
        ent["sm"].copy_(sm)
        ent["graph"].replay()
        lib._count(ent["n"])
        e_bf, e_split, (log_d, dur, cum, mel_len, tmax) = ent["outs"]
        # the graph's output buffers are overwritten by the next replay: what leaves this call is copied out (e_split and
        # cum are consumed by the LengthRegulator inside this forward)
        return e_bf.clone(), e_split, (log_d.clone(), dur.clone(), cum, mel_len.clone(), tmax)

    def _aux_f32(self, t):
        """Slots 5 / 6 of the tuple (variance_adaptor_output, text_dur_predicted) are fp32 in the reference; the
        kernels compute on bf16.  Inference: the producing kernels (bucket_embed_add, lr_expand) write the fp32 copies
        themselves; training: a cast.  No live caller reads them (train_fastspeech2.py:167, test_fastspeech2.py:169), so
        `aux_outputs_fp32 = False` skips them (216 MB of writes at cfg2)."""
        return t.float() if self.aux_outputs_fp32 else t

    def invalidate_caches(self) -> None:
        """Drop every derived (bf16 / split / fused) weight copy.  The caches notice `load_state_dict`, optimizer steps
        and any other in-place update that bumps `Tensor._version`; writes through `param.data` (EMA / averaging code)
        do not bump it — call this after such an update."""
        for m in self.modules():
            for name in ("_ttsb_cache", "_ttsb_cache_x3"):
                c = getattr(m, name, None)
                if c is not None:
                    c.sig = None
        self._graphs.clear()
