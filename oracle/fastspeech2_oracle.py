"""ORACLE — test infrastructure, NOT the product.

A plain-PyTorch fp32 restatement of the reference's FastSpeech2 mel-generation path
(syoamakase/Transformer_TTS, `Models/fastspeech2.py` and the modules it calls), written as pure
functions over a reference-layout ``state_dict``.  Every function cites the reference file:line it
follows.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / reference arm
may import this file; the product (``transformer_tts_b200``) never does.

Parity status: PINNED against the reference itself.  The reference ships no tests or golden vectors
(SURVEY.md §4), so ``oracle/make_golden.py`` imports the unmodified reference from /root/reference in
the build container, runs it on seeded inputs and commits the outputs under ``tests/golden/``;
``tests/test_oracle.py`` checks this restatement against those files (and against the live reference
when /root/reference is present).

Deliberate scope: default hparams branch only — transformer encoder/decoder, no
speaker/accent embeddings, no SQ-VAE/hop/debug/fix_mask.  Dropout: 0 unless the caller supplies the masks
(`drop` argument: every dropout call site of the reference is named, see `_drop`); attention dropout fires
even in eval() in the reference, so inference parity is only defined at 0.
"""
from __future__ import annotations

import math
from functools import lru_cache
from typing import Dict, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


def _drop(drop, name: str, x: Tensor, channels_first: bool = False) -> Tensor:
    """Training-mode dropout call site `name` of the reference (nn.Dropout / F.dropout).  The reference draws its masks
    from torch's generator, which no other implementation can reproduce; for parity tests the caller supplies them:
    `drop(name, shape) -> mask` (0 or 1/(1-p)) in the time-major (B,T,C) / (B,H,T,T) layout.  None = dropout 0."""
    if drop is None:
        return x
    if channels_first:  # (B,C,T) activations, mask given as (B,T,C)
        return x * drop(name, tuple(x.transpose(1, 2).shape)).transpose(1, 2)
    return x * drop(name, tuple(x.shape))


# --------------------------------------------------------------------------------------------------
# Models/modules.py:90-111  PositionalEncoder
# --------------------------------------------------------------------------------------------------
@lru_cache(maxsize=4)
def positional_table(d_model: int, max_seq_len: int = 5000) -> Tensor:
    """modules.py:97-105 — note the non-standard exponents 2*i/d (i already even) and 2*(i+1)/d,
    evaluated in Python float64 and stored as fp32."""
    pe = torch.zeros(max_seq_len, d_model)
    rows = []
    for pos in range(max_seq_len):
        row = [0.0] * d_model
        for i in range(0, d_model, 2):
            row[i] = math.sin(pos / (10000 ** ((2 * i) / d_model)))
            row[i + 1] = math.cos(pos / (10000 ** ((2 * (i + 1)) / d_model)))
        rows.append(row)
    pe = torch.tensor(rows, dtype=torch.float64).to(torch.float32)
    return pe  # (max_seq_len, d_model); the reference keeps it as (1, max_seq_len, d_model)


def positional_encoder(x: Tensor, alpha: Tensor, drop=None, tag: str = "enc") -> Tensor:
    """modules.py:107-111."""
    pe = positional_table(x.shape[-1])[: x.shape[1]].to(x.device)
    return _drop(drop, f"{tag}.pe", x + alpha * pe.unsqueeze(0))


# --------------------------------------------------------------------------------------------------
# Models/modules.py:7-21  attention ; :43-70 MultiHeadAttention.forward
# --------------------------------------------------------------------------------------------------
def attention(q: Tensor, k: Tensor, v: Tensor, d_k: int, mask: Optional[Tensor], drop=None, site: str = ""):
    scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(d_k)  # :8-9
    if mask is not None:
        mask = mask.unsqueeze(1)  # :12
        scores = scores.masked_fill(mask == 0, -1e4)  # :14
    scores = torch.softmax(scores, dim=-1)  # :15/:17
    scores = _drop(drop, site, scores)  # :18-19
    return torch.matmul(scores, v), scores  # :20-21


def multi_head_attention(sd: Dict[str, Tensor], pfx: str, x: Tensor, mask: Optional[Tensor], heads: int, drop=None,
                         site: str = ""):
    bs, _, d_model = x.shape
    d_k = d_model // heads
    k = F.linear(x, sd[pfx + "k_linear.weight"], sd[pfx + "k_linear.bias"]).view(bs, -1, heads, d_k)  # :49
    q = F.linear(x, sd[pfx + "q_linear.weight"], sd[pfx + "q_linear.bias"]).view(bs, -1, heads, d_k)  # :50
    v = F.linear(x, sd[pfx + "v_linear.weight"], sd[pfx + "v_linear.bias"]).view(bs, -1, heads, d_k)  # :51
    k, q, v = k.transpose(1, 2), q.transpose(1, 2), v.transpose(1, 2)  # :58-60
    ctx, attn = attention(q, k, v, d_k, mask, drop, site)  # :63
    concat = ctx.transpose(1, 2).contiguous().view(bs, -1, d_model)  # :65
    return F.linear(concat, sd[pfx + "out.weight"], sd[pfx + "out.bias"]), attn  # :68


# --------------------------------------------------------------------------------------------------
# Models/modules.py:72-88  FeedForward
# --------------------------------------------------------------------------------------------------
def feed_forward(sd: Dict[str, Tensor], pfx: str, x: Tensor, drop=None, site: str = "") -> Tensor:
    k = sd[pfx + "f_1.weight"].shape[-1]
    res = x
    h = F.relu(F.conv1d(x.transpose(1, 2), sd[pfx + "f_1.weight"], sd[pfx + "f_1.bias"], padding=int(k / 2)))  # :83
    h = F.conv1d(h, sd[pfx + "f_2.weight"], sd[pfx + "f_2.bias"], padding=int(k / 2)).transpose(1, 2)  # :84
    h = _drop(drop, site, h + res)  # :85-86
    return F.layer_norm(h, (h.shape[-1],), sd[pfx + "layer_norm.weight"], sd[pfx + "layer_norm.bias"])  # :87


# --------------------------------------------------------------------------------------------------
# Models/layers.py:29-41  EncoderLayer.forward
# --------------------------------------------------------------------------------------------------
def encoder_layer(sd, pfx: str, x: Tensor, mask: Optional[Tensor], heads: int, drop=None, site: str = ""):
    """`site` = "<enc|dec>.<layer index>": prefix of this block's four dropout call sites."""
    d = x.shape[-1]
    res = x
    xn = F.layer_norm(x, (d,), sd[pfx + "norm_1.weight"], sd[pfx + "norm_1.bias"])  # :31
    out, attn = multi_head_attention(sd, pfx + "attn.", xn, mask, heads, drop, site + ".attn")  # :32
    x = res + _drop(drop, site + ".drop1", out)  # :33
    res = x
    xn = F.layer_norm(x, (d,), sd[pfx + "norm_2.weight"], sd[pfx + "norm_2.bias"])  # :35
    x = res + _drop(drop, site + ".drop2", feed_forward(sd, pfx + "ff.", xn, drop, site + ".ff"))  # :40
    return x, attn


# --------------------------------------------------------------------------------------------------
# Models/encoder.py:83-112  Encoder.forward (embedding=True for the text encoder, Linear for the decoder)
# --------------------------------------------------------------------------------------------------
def encoder(sd, pfx: str, src: Tensor, mask: Optional[Tensor], n_layers: int, heads: int, return_attn=False, drop=None):
    w = sd[pfx + "embed.weight"]
    if src.dtype in (torch.int64, torch.int32):
        x = F.embedding(src, w, padding_idx=0)  # :55,:84 (padding_idx only affects gradients)
    else:
        x = F.linear(src, w, sd[pfx + "embed.bias"])  # :57,:84
    tag = "enc" if pfx.startswith("encoder") else "dec"
    x = positional_encoder(x, sd[pfx + "pe.alpha"], drop, tag)  # :94
    attns = []
    for i in range(n_layers):
        x, attn = encoder_layer(sd, f"{pfx}layers.{i}.", x, mask, heads, drop, f"{tag}.{i}")  # :99
        if return_attn:
            attns.append(attn)
    x = F.layer_norm(x, (x.shape[-1],), sd[pfx + "norm.weight"], sd[pfx + "norm.bias"])  # :112
    return x, (torch.stack(attns, 1) if return_attn else None)  # attns: (b, N, heads, t, t) :97,:105


# --------------------------------------------------------------------------------------------------
# Models/varianceadaptor.py:186-231  VariancePredictor
# --------------------------------------------------------------------------------------------------
def variance_predictor(sd, pfx: str, x: Tensor, mask: Optional[Tensor], drop=None) -> Tensor:
    fs = sd[pfx + "conv1.weight"].shape[0]
    name = pfx.rstrip(".").split(".")[-1].replace("_predictor", "")  # duration | pitch | energy
    h = F.conv1d(x.transpose(1, 2), sd[pfx + "conv1.weight"], sd[pfx + "conv1.bias"], padding=1)  # :217
    h = F.relu(h).transpose(1, 2)  # :218
    h = F.layer_norm(h, (fs,), sd[pfx + "layer_norm1.weight"], sd[pfx + "layer_norm1.bias"])  # :219
    h = _drop(drop, name + ".drop1", h).transpose(1, 2)
    h = F.conv1d(h, sd[pfx + "conv2.weight"], sd[pfx + "conv2.bias"], padding=1)  # :220
    h = F.relu(h).transpose(1, 2)  # :221
    h = F.layer_norm(h, (fs,), sd[pfx + "layer_norm2.weight"], sd[pfx + "layer_norm2.bias"])  # :222
    h = _drop(drop, name + ".drop2", h)
    out = F.linear(h, sd[pfx + "linear_layer.weight"], sd[pfx + "linear_layer.bias"]).squeeze(-1)  # :223-224
    if mask is not None:
        m = mask.squeeze(1) if mask.dim() == 3 else mask
        out = out.masked_fill(m == 0, 0.0)  # :228-229
    return out


# --------------------------------------------------------------------------------------------------
# Models/varianceadaptor.py:75  duration rounding ; :141-177 LengthRegulator ; :233-259 pad / mask
# --------------------------------------------------------------------------------------------------
def round_durations(log_d: Tensor) -> Tensor:
    return torch.clamp(torch.round(torch.exp(log_d) - 1.0), min=0)  # :75 (log_offset hard-coded 1, :64)


def length_regulator(x: Tensor, duration: Tensor, max_length: Optional[int] = None):
    """:141-156 LR + :167-177 expand + :233-249 pad.  Returns (output, mel_pos, frame_idx).

    output[b] = cat_i x[b,i].expand(int(duration[b,i]), -1), zero padded (or cropped: F.pad with a
    negative amount crops) to max_length / the batch maximum.  mel_pos = arange(1, len+1) padded with
    0 in the free-running call; with max_length given the reference pads `output` twice and returns
    the padded output in its place (:150-151) — callers ignore it, we return positions there too.
    frame_idx (source phoneme of every frame, -1 on padding) is an extra the CUDA kernel exposes."""
    B, L, D = x.shape
    dur = duration.to(torch.int64).clamp(min=0)
    lens = dur.sum(1)
    max_len = int(max_length) if max_length else int(lens.max().item())
    out = x.new_zeros(B, max_len, D)
    mel_pos = torch.zeros(B, max_len, dtype=torch.int64, device=x.device)
    frame_idx = torch.full((B, max_len), -1, dtype=torch.int64, device=x.device)
    for b in range(B):
        idx = torch.repeat_interleave(torch.arange(L, device=x.device), dur[b])  # expand + cat (:170-175)
        n = min(int(idx.numel()), max_len)
        out[b, :n] = x[b, idx[:n]]
        frame_idx[b, :n] = idx[:n]
        mel_pos[b, :n] = torch.arange(1, n + 1, device=x.device)
    return out, mel_pos, frame_idx


def get_mask_from_lengths(mel_pos: Tensor) -> Tensor:
    """:251-259 — called with padded POSITIONS, not lengths: ids <= mel_pos."""
    max_len = int(torch.max(mel_pos).item())
    ids = torch.arange(0, max_len, device=mel_pos.device).unsqueeze(0).expand(mel_pos.shape[0], -1)
    return ids <= mel_pos


def variance_bins(n_bins: int, f0_min: float, f0_max: float, energy_min: float, energy_max: float):
    """:56,:61 — computed on the host exactly as the reference does."""
    import numpy as np

    pitch_bins = torch.exp(torch.linspace(np.log(f0_min), np.log(f0_max), n_bins - 1))
    energy_bins = torch.linspace(energy_min, energy_max, n_bins - 1)
    return pitch_bins, energy_bins


# --------------------------------------------------------------------------------------------------
# Models/varianceadaptor.py:66-129  VarianceAdaptor.forward
# --------------------------------------------------------------------------------------------------
def variance_adaptor(sd, pfx: str, x: Tensor, src_mask, mel_mask, d_target, p_target, e_target, bins, drop=None):
    pitch_bins, energy_bins = (b.to(x.device) for b in bins)
    log_d = variance_predictor(sd, pfx + "duration_predictor.", x, src_mask, drop)  # :69
    if d_target is not None:
        max_len = mel_mask.shape[2] if mel_mask is not None else None  # :71-72
        x, mel_pos, frame_idx = length_regulator(x, d_target, max_len)  # :73
    else:
        dur = round_durations(log_d)  # :75
        x, mel_pos, frame_idx = length_regulator(x, dur, None)  # :82
        mel_mask = get_mask_from_lengths(mel_pos)  # :84 -> 2-D (B,T)
    p = variance_predictor(sd, pfx + "pitch_predictor.", x, mel_mask, drop)  # :96
    p_src = p_target if p_target is not None else p  # :97-109 (scheduled sampling p=0 returns the target)
    p_idx = torch.bucketize(p_src, pitch_bins)
    e = variance_predictor(sd, pfx + "energy_predictor.", x, mel_mask, drop)  # :115
    e_src = e_target if e_target is not None else e  # :116-118
    e_idx = torch.bucketize(e_src, energy_bins)
    text_dur = x  # :122
    x = x + F.embedding(p_idx, sd[pfx + "pitch_embedding.weight"])  # :123-124
    x = x + F.embedding(e_idx, sd[pfx + "energy_embedding.weight"])  # :125-126
    return dict(x=x, log_d=log_d, p=p, e=e, mel_pos=mel_pos, mel_mask=mel_mask, text_dur=text_dur,
                frame_idx=frame_idx, p_idx=p_idx, e_idx=e_idx)


# --------------------------------------------------------------------------------------------------
# Models/postnets.py:64-79  PostConvNet.forward (prev_version=True, eval-mode BatchNorm, dropout 0)
# --------------------------------------------------------------------------------------------------
def _bn_eval(sd, pfx: str, x: Tensor, training: bool = False) -> Tensor:
    if training:  # nn.BatchNorm1d in train(): batch statistics over all B*T positions (running stats not updated here)
        return F.batch_norm(x, None, None, sd[pfx + "weight"], sd[pfx + "bias"], training=True, eps=1e-5)
    return F.batch_norm(x, sd[pfx + "running_mean"], sd[pfx + "running_var"], sd[pfx + "weight"],
                        sd[pfx + "bias"], training=False, eps=1e-5)


def postnet(sd, pfx: str, x: Tensor, training: bool = False, drop=None):
    mel_pred = F.linear(x, sd[pfx + "out.weight"], sd[pfx + "out.bias"]).transpose(1, 2)  # :67
    h = F.conv1d(mel_pred, sd[pfx + "conv1.weight"], sd[pfx + "conv1.bias"], padding=4)[:, :, :-4]  # :71
    h = _drop(drop, "postnet.0", torch.tanh(_bn_eval(sd, pfx + "pre_batchnorm.", h, training)), channels_first=True)
    for i in range(3):  # :72-73
        h = F.conv1d(h, sd[f"{pfx}conv_list.{i}.weight"], sd[f"{pfx}conv_list.{i}.bias"], padding=4)[:, :, :-4]
        h = _drop(drop, f"postnet.{i + 1}", torch.tanh(_bn_eval(sd, f"{pfx}batch_norm_list.{i}.", h, training)),
                  channels_first=True)
    h = F.conv1d(h, sd[pfx + "conv2.weight"], sd[pfx + "conv2.bias"], padding=4)[:, :, :-4]  # :74
    post = mel_pred + h  # :75
    return mel_pred.transpose(1, 2), post.transpose(1, 2)  # :77


# --------------------------------------------------------------------------------------------------
# Models/fastspeech2.py:118-241  FastSpeech2.forward (default branch)
# --------------------------------------------------------------------------------------------------
def fastspeech2_forward(sd: Dict[str, Tensor], hp: dict, src: Tensor, src_mask: Tensor,
                        mel_mask: Optional[Tensor] = None, d_target: Optional[Tensor] = None,
                        p_target: Optional[Tensor] = None, e_target: Optional[Tensor] = None,
                        return_attn: bool = False, training: bool = False, drop=None) -> dict:
    """hp keys: n_layer_encoder, n_head_encoder, n_layer_decoder, n_head_decoder, nbins, f0_min,
    f0_max, energy_min, energy_max.  Returns a dict with the reference's tuple slots by name."""
    bins = variance_bins(hp["nbins"], hp["f0_min"], hp["f0_max"], hp["energy_min"], hp["energy_max"])
    e_out, attn_enc = encoder(sd, "encoder.", src, src_mask, hp["n_layer_encoder"], hp["n_head_encoder"],
                              return_attn, drop)  # :137
    va = variance_adaptor(sd, "variance_adaptor.", e_out, src_mask, mel_mask, d_target, p_target, e_target,
                          bins, drop)  # :176-181
    # Free-running: the reference hands the 2-D (B,T) mask to attention(), whose unsqueeze(1) only
    # broadcasts as a key mask at B == 1 (B=2 raises, B=H mis-assigns).  The oracle uses the
    # intended per-utterance key mask (B,1,T), which is bit-identical to the reference at B == 1.
    dec_mask = mel_mask if d_target is not None else va["mel_mask"].unsqueeze(1)
    d_out, attn_dec = encoder(sd, "decoder.", va["x"], dec_mask, hp["n_layer_decoder"], hp["n_head_decoder"],
                              return_attn, drop)  # :195
    mel_pre, mel_post = postnet(sd, "postnet.", d_out, training, drop)  # :197
    return dict(outputs_prenet=mel_pre, outputs_postnet=mel_post, log_d_prediction=va["log_d"],
                p_prediction=va["p"], e_prediction=va["e"], variance_adaptor_output=va["x"],
                text_dur_predicted=va["text_dur"], attn_enc=attn_enc, attn_dec=attn_dec,
                mel_pos=va["mel_pos"], mel_mask=va["mel_mask"], frame_idx=va["frame_idx"],
                p_idx=va["p_idx"], e_idx=va["e_idx"], encoder_output=e_out, decoder_output=d_out)


# --------------------------------------------------------------------------------------------------
# train_fastspeech2.py:212-217,245-259  the five unmasked L1 terms of the training loss
# --------------------------------------------------------------------------------------------------
def training_losses(out: dict, mel: Tensor, alignment: Tensor, f0: Tensor, energy: Tensor) -> dict:
    mel_loss = F.l1_loss(out["outputs_prenet"], mel)  # :212
    post_mel_loss = F.l1_loss(out["outputs_postnet"], mel)  # :214
    duration_loss = F.l1_loss(out["log_d_prediction"], torch.log(alignment.float() + 1))  # :245
    f0_loss = F.l1_loss(out["p_prediction"], f0)  # :249
    energy_loss = F.l1_loss(out["e_prediction"], energy)  # :254
    loss = mel_loss + post_mel_loss + f0_loss + energy_loss + duration_loss  # :215,250,255,259
    return dict(loss=loss, mel_loss=mel_loss, post_mel_loss=post_mel_loss, duration_loss=duration_loss,
                f0_loss=f0_loss, energy_loss=energy_loss)


def training_step_grads(sd: Dict[str, Tensor], hp: dict, text, src_mask, mel_mask, alignment, f0, energy, mel, drop=None):
    """Forward in train() semantics (BatchNorm batch statistics; dropout masks supplied by `drop`) + losses + autograd backward,
    as train_fastspeech2.py:167,212-259,304 does.  Returns (losses, {name: grad}) for the float parameters."""
    names = [k for k, v in sd.items() if v.is_floating_point() and "running_" not in k]
    p = {k: (sd[k].detach().clone().requires_grad_(True) if k in names else sd[k]) for k in sd}
    out = fastspeech2_forward(p, hp, text, src_mask, mel_mask=mel_mask, d_target=alignment, p_target=f0,
                              e_target=energy, training=True, drop=drop)
    losses = training_losses(out, mel, alignment, f0, energy)
    grads = torch.autograd.grad(losses["loss"], [p[k] for k in names], allow_unused=True)
    return ({k: v.detach() for k, v in losses.items()},
            {k: (g if g is not None else torch.zeros_like(sd[k])) for k, g in zip(names, grads)}, out)


def radam_step(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, lr: float = 1e-3, beta1: float = 0.9,
               beta2: float = 0.98, eps: float = 1e-9):
    """One radam.RAdam.step() for one tensor (radam.py:56-92; weight_decay 0, degenerated_to_sgd=True — the
    optimizer train_fastspeech2.py:412-414 builds).  `step` is the 1-based count AFTER the increment at :61.
    Returns (p, m, v) as new tensors."""
    v = v * beta2 + (1 - beta2) * g * g                      # :58
    m = m * beta1 + (1 - beta1) * g                          # :59
    beta2_t = beta2 ** step                                  # :67
    n_sma_max = 2 / (1 - beta2) - 1                          # :68
    n_sma = n_sma_max - 2 * step * beta2_t / (1 - beta2_t)   # :69
    if n_sma >= 5:                                           # :73-74, :83-88
        step_size = math.sqrt((1 - beta2_t) * (n_sma - 4) / (n_sma_max - 4) * (n_sma - 2) / n_sma * n_sma_max /
                              (n_sma_max - 2)) / (1 - beta1 ** step)
        p = p - step_size * lr * m / (v.sqrt() + eps)
    else:                                                    # :75-76, :89-93
        step_size = 1.0 / (1 - beta1 ** step)
        p = p - step_size * lr * m
    return p, m, v
