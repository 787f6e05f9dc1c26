"""Drop-in for the reference's ``Models/varianceadaptor.py``: ``VarianceAdaptor`` (:34-129),
``LengthRegulator`` (:131-184), ``VariancePredictor`` (:186-231), ``pad`` (:233-249),
``get_mask_from_lengths`` (:251-259)."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import lib
from . import _engine as E


class VariancePredictor(nn.Module):
    """Duration, Pitch and Energy Predictor: conv(k3)-ReLU-LN-conv(k3)-ReLU-LN-Linear(->1)-mask."""

    def __init__(self, encoder_hidden_size, variance_predictor_filter_size=256, variance_predictor_kernel_size=3,
                 variance_predictor_dropout=0.5):
        super().__init__()
        if variance_predictor_kernel_size != 3 or variance_predictor_filter_size != 256:
            raise NotImplementedError("this build supports the reference's fixed predictor shape (k=3, 256 filters)")
        self.input_size = encoder_hidden_size
        self.filter_size = variance_predictor_filter_size
        self.kernel = variance_predictor_kernel_size
        self.conv_output_size = variance_predictor_filter_size
        self.dropout = variance_predictor_dropout
        self.conv1 = nn.Conv1d(self.input_size, self.filter_size, kernel_size=self.kernel, padding=1)
        self.relu1 = nn.ReLU()
        self.layer_norm1 = nn.LayerNorm(self.filter_size)
        self.dropout1 = nn.Dropout(self.dropout)
        self.conv2 = nn.Conv1d(self.filter_size, self.filter_size, kernel_size=self.kernel, padding=1)
        self.relu2 = nn.ReLU()
        self.layer_norm2 = nn.LayerNorm(self.filter_size)
        self.dropout2 = nn.Dropout(self.dropout)
        self.linear_layer = nn.Linear(self.conv_output_size, 1)
        self._ttsb_cache = E.Cache()

    def run(self, x_bf16, mask_u8):
        return E.run_predictor(E.predictor_cache(self), x_bf16, mask_u8)

    def forward(self, encoder_output, mask):
        E.require_inference(self, "VariancePredictor")
        E.require_cuda(encoder_output, "VariancePredictor")
        B, T, _ = encoder_output.shape
        m = None if mask is None else E.mask_to_u8(mask, B, T, encoder_output.device)
        return self.run(encoder_output.to(torch.bfloat16).contiguous(), m)


def pad(input_ele, mel_max_length=None):
    """Host-side helper kept for API parity (varianceadaptor.py:233-249); the kernels never call it."""
    if mel_max_length:
        max_len = mel_max_length
    else:
        max_len = max(e.size(0) for e in input_ele)
    out_list = []
    for batch in input_ele:
        if batch.dim() == 1:
            out_list.append(F.pad(batch, (0, max_len - batch.size(0)), "constant", 0.0))
        elif batch.dim() == 2:
            out_list.append(F.pad(batch, (0, 0, 0, max_len - batch.size(0)), "constant", 0.0))
    return torch.stack(out_list)


def get_mask_from_lengths(lengths, max_len=None):
    """`arange(max) <= lengths` (varianceadaptor.py:251-259).  The reference calls this with padded
    POSITIONS; the LengthRegulator kernel emits the same mask directly."""
    batch_size = lengths.shape[0]
    if max_len is None:
        max_len = int(torch.max(lengths).item())
    ids = torch.arange(0, max_len, device=lengths.device).unsqueeze(0).expand(batch_size, -1)
    return ids <= lengths


class LengthRegulator(nn.Module):
    """Length Regulator: warp-scan prefix sum of durations + vectorised frame gather."""

    def __init__(self, use_lstm=False, hidden_dim=None):
        super().__init__()
        if use_lstm:
            raise NotImplementedError("use_lstm is dead code in the reference (varianceadaptor.py:136-139)")
        self.use_lstm = use_lstm

    @staticmethod
    def run(x_bf16, dur_i32, max_length=None, want_aux=True, scan=None, f32_out=None):
        """Returns (expanded bf16 (B,Tmax,D), frame_idx, mel_pos, mel_mask u8, mel_len i32, cum).
        scan: (cum, mel_len, max(mel_len) as a device scalar) when the caller already ran the prefix sum.
        f32_out: a list that receives the expanded rows as an fp32 tensor too (written by the same kernel)."""
        cum, mel_len, tmax_dev = scan if scan is not None else (*lib.lr_scan(dur_i32), None)
        if max_length:
            Tmax = int(max_length)
        else:
            if tmax_dev is None:
                tmax_dev = mel_len.max()
            Tmax = int(tmax_dev.item())  # the one host sync of free-running inference (varianceadaptor.py:254)
            Tmax = max(Tmax, 1)
        o32 = None
        if f32_out is not None:
            o32 = torch.empty((x_bf16.shape[0], Tmax, x_bf16.shape[2]), device=x_bf16.device, dtype=torch.float32)
            f32_out.append(o32)
        out, fidx, mel_pos, mel_mask = lib.lr_expand(x_bf16, cum, mel_len, Tmax, want_aux=want_aux, out_f32=o32)
        return out, fidx, mel_pos, mel_mask, mel_len, cum

    def LR(self, x, duration, max_length=None):
        E.require_cuda(x, "LengthRegulator")
        dur = torch.clamp(duration, min=0).to(torch.int32).contiguous()
        xb = x.to(torch.bfloat16).contiguous()
        out, _, mel_pos, _, _, _ = self.run(xb, dur, max_length)
        return out.to(x.dtype), mel_pos

    def forward(self, x, duration, max_length=None):
        return self.LR(x, duration, max_length)


class VarianceAdaptor(nn.Module):
    def __init__(self, d_model_encoder, n_bins=256, f0_min=71.0, f0_max=795.8, energy_min=0.0, energy_max=315.0,
                 log_offset=1., pitch_pred=True, energy_pred=True, dropout=0.5, use_rnn_length=False, use_pos=False):
        super().__init__()
        if use_rnn_length or use_pos:
            raise NotImplementedError("use_rnn_length / use_pos are out of scope (off by default, SURVEY §2a)")
        if not (pitch_pred and energy_pred):
            raise NotImplementedError("pitch_pred and energy_pred must both be on (hparams default)")
        self.pitch_pred = pitch_pred
        self.energy_pred = energy_pred
        self.use_rnn_length = use_rnn_length
        self.use_pos = use_pos
        self.duration_predictor = VariancePredictor(d_model_encoder, variance_predictor_dropout=dropout)
        self.length_regulator = LengthRegulator()
        self.pitch_predictor = VariancePredictor(d_model_encoder, variance_predictor_dropout=dropout)
        # bins are plain tensors (not buffers) computed on the host exactly as the reference does (:56,:61)
        self.pitch_bins = torch.exp(torch.linspace(np.log(f0_min), np.log(f0_max), n_bins - 1))
        self.pitch_embedding = nn.Embedding(n_bins, d_model_encoder)
        self.energy_predictor = VariancePredictor(d_model_encoder, variance_predictor_dropout=dropout)
        self.energy_bins = torch.linspace(energy_min, energy_max, n_bins - 1)
        self.energy_embedding = nn.Embedding(n_bins, d_model_encoder)
        self.log_offset = 1.  # hard-coded in the reference regardless of the argument (:64)
        self._ttsb_cache = E.Cache()
        self._bins_dev = {}

    def _bins(self, device):
        k = str(device)
        if k not in self._bins_dev:
            self._bins_dev[k] = (self.pitch_bins.to(device).float().contiguous(),
                                 self.energy_bins.to(device).float().contiguous())
        return self._bins_dev[k]

    def _tables(self):
        c = self._ttsb_cache
        params = [self.pitch_embedding.weight, self.energy_embedding.weight]
        if c.stale(params):
            c.t["ptab"], c.t["etab"] = E._f32(self.pitch_embedding.weight), E._f32(self.energy_embedding.weight)
            c.commit(params)
        return c.t

    def run(self, x_bf16, src_mask_u8, mel_mask_u8=None, duration_target=None, pitch_target=None,
            energy_target=None, max_len=None, want_aux=True, log_d=None, pre=None, x_split=None, aux_f32=False):
        """Kernel sequence of varianceadaptor.py:66-129 on bf16 activations.  Returns a dict.
        log_d: the duration predictor's output when the caller already evaluated it (fp32-accurate chain).
        pre: (log_d, dur, cum, mel_len, max mel_len) from `run_pre` (free-running: everything before the one host sync).
        x_split: the encoder output as an fp16 hi|lo split (B,L,2D): the pitch / energy predictors then run the
        fp32-accurate chain too, so the pitch / energy BUCKETS equal the fp32 reference's (variance_precision "fp32")."""
        dev = x_bf16.device
        scan = None
        if pre is not None:
            log_d, dur, cum, mel_len, tmax_dev = pre
            scan = (cum, mel_len, tmax_dev)
            max_len = None
        else:
            if log_d is None:
                log_d = self.duration_predictor.run(x_bf16, src_mask_u8)  # :69
            if duration_target is not None:
                dur = torch.clamp(duration_target, min=0).to(torch.int32).contiguous()
                if mel_mask_u8 is not None:
                    max_len = mel_mask_u8.shape[1]  # :71-72
            else:
                dur = lib.round_durations(log_d)  # :75
                max_len = None
        td32 = [] if aux_f32 else None  # aux_f32: tuple slots 5 / 6 in fp32, written by the kernels that produce them
        xe, fidx, mel_pos, lr_mask, mel_len, _ = self.length_regulator.run(x_bf16, dur, max_len, want_aux, scan=scan,
                                                                           f32_out=td32)  # :73/:82
        if duration_target is None or mel_mask_u8 is None:
            mel_mask_u8 = lr_mask  # :84 (the kernel applies the `arange <= mel_pos` rule)
        if x_split is not None:
            # the split is expanded as 16-bit words by the same gather kernel (a frame copies its phoneme's hi and lo halves)
            xs16 = x_split.view(torch.bfloat16)
            xe_s, _, _, _ = lib.lr_expand(xs16, pre[2] if pre is not None else lib.lr_scan(dur)[0], mel_len, xe.shape[1],
                                          want_aux=False)
            xe_s = xe_s.view(torch.float16)
            p = E.run_predictor_precise(self.pitch_predictor, xe_s, mel_mask_u8)   # :96
            e = E.run_predictor_precise(self.energy_predictor, xe_s, mel_mask_u8)  # :115
        else:
            p = self.pitch_predictor.run(xe, mel_mask_u8)  # :96
            e = self.energy_predictor.run(xe, mel_mask_u8)  # :115
        p_src = p if pitch_target is None else pitch_target.to(torch.float32).contiguous()  # :97-109
        e_src = e if energy_target is None else energy_target.to(torch.float32).contiguous()  # :116-118
        pb, eb = self._bins(dev)
        tabs = self._tables()
        x32 = torch.empty(xe.shape, device=dev, dtype=torch.float32) if aux_f32 else None
        out, p_idx, e_idx = lib.bucket_embed_add(xe, p_src, e_src, pb, eb, tabs["ptab"], tabs["etab"],
                                                 want_idx=want_aux, out_f32=x32)  # :100-126
        return dict(x=out, log_d=log_d, p=p, e=e, mel_pos=mel_pos, mel_mask=mel_mask_u8, text_dur=xe,
                    frame_idx=fidx, mel_len=mel_len, dur=dur, p_idx=p_idx, e_idx=e_idx, x_f32=x32,
                    text_dur_f32=td32[0] if td32 else None)

    def run_pre(self, x_bf16, src_mask_u8, log_d=None):
        """Free-running: everything between the encoder and the host read of Tmax — duration predictor (:69) unless
        `log_d` is given, rounding (:75), the LengthRegulator's prefix sum and the batch maximum.  No host sync."""
        if log_d is None:
            log_d = self.duration_predictor.run(x_bf16, src_mask_u8)
        dur = lib.round_durations(log_d)
        cum, mel_len = lib.lr_scan(dur)
        return log_d, dur, cum, mel_len, mel_len.max()

    def forward(self, x, src_mask, mel_mask=None, duration_target=None, pitch_target=None, energy_target=None,
                max_len=None, p_scheduled_sampling=0.0, pitch_perturbation=False, duration_perturbation=False):
        E.require_inference(self, "VarianceAdaptor")
        E.require_cuda(x, "VarianceAdaptor")
        if pitch_perturbation or duration_perturbation or p_scheduled_sampling:
            raise NotImplementedError("perturbation / scheduled sampling are out of scope (off by default)")
        B, L, _ = x.shape
        sm = E.mask_to_u8(src_mask, B, L, x.device)
        mm = None if mel_mask is None else E.mask_to_u8(mel_mask, B, mel_mask.shape[-1], x.device)
        r = self.run(x.to(torch.bfloat16).contiguous(), sm, mm, duration_target, pitch_target, energy_target, max_len)
        out_mask = mel_mask if (duration_target is not None and mel_mask is not None) else r["mel_mask"].bool()
        return (r["x"].float(), r["log_d"], r["p"], r["e"], r["mel_pos"], out_mask, r["text_dur"].float())
