"""Synthetic hparams, weights and batches for tests and benchmarks.

There is no network and the reference's corpora are private, so every test and benchmark runs on
random-init weights of the canonical architecture and synthetic phoneme batches.  The weight
generator does not depend on module construction order: each ``state_dict`` entry is drawn from its
own seeded CPU generator, so the reference model (build container), the oracle and the CUDA model
(GPU box) all load bit-identical tensors.  Distributions follow what the reference's
``train_fastspeech2.py:399`` leaves in place after ``model.apply(init_weight)`` (utils/utils.py:153-177):
Conv1d kaiming-normal with zero bias, Linear/Embedding at PyTorch defaults — with the norm layers
perturbed away from identity so that parity tests exercise them.
"""
from __future__ import annotations

import hashlib
import math
from collections import OrderedDict
from types import SimpleNamespace
from typing import Dict

import torch

CANONICAL_HP: Dict[str, object] = dict(
    vocab_size=152, mel_dim=80,
    d_model_encoder=384, n_layer_encoder=6, n_head_encoder=4, ff_conv_kernel_size_encoder=5,
    concat_after_encoder=False,
    d_model_decoder=384, n_layer_decoder=6, n_head_decoder=4, ff_conv_kernel_size_decoder=1,
    concat_after_decoder=False,
    reduction_rate=1, nbins=256, f0_min=71.0, f0_max=799.8, energy_min=0.0, energy_max=403.8,
    postnet_pred=True,
    # fill_variables defaults (utils/utils.py:184-201) the model constructor reads
    encoder_type='transformer', decoder_type='transformer', use_sq_vae=False, use_hop=False,
    use_rnn_length=False, use_pos=False, p_scheduled_sampling=0.0, spk_emb_architecture='',
    pitch_pred=True, energy_pred=True, accent_emb=False, is_multi_speaker=False, spk_emb_dim=None,
    output_type=None, num_group=None,
)


def make_hp(**overrides) -> SimpleNamespace:
    d = dict(CANONICAL_HP)
    d.update(overrides)
    return SimpleNamespace(**d)


def _attn_spec(spec, pfx: str, d: int) -> None:
    for nm in ("q_linear", "v_linear", "k_linear", "out"):
        spec[pfx + nm + ".weight"] = ((d, d), "linear_w")
        spec[pfx + nm + ".bias"] = ((d,), ("linear_b", d))


def _ff_spec(spec, pfx: str, d: int, k: int) -> None:
    spec[pfx + "f_1.weight"] = ((4 * d, d, k), "conv_w")
    spec[pfx + "f_1.bias"] = ((4 * d,), "conv_b")
    spec[pfx + "f_2.weight"] = ((d, 4 * d, k), "conv_w")
    spec[pfx + "f_2.bias"] = ((d,), "conv_b")
    spec[pfx + "layer_norm.weight"] = ((d,), "norm_w")
    spec[pfx + "layer_norm.bias"] = ((d,), "norm_b")


def _postnet_spec(spec, dd: int, md: int, with_out: bool) -> None:
    spec["postnet.conv1.weight"] = ((dd, md, 5), "conv_w")
    spec["postnet.conv1.bias"] = ((dd,), "conv_b")
    for i in range(3):
        spec[f"postnet.conv_list.{i}.weight"] = ((dd, dd, 5), "conv_w")
        spec[f"postnet.conv_list.{i}.bias"] = ((dd,), "conv_b")
    spec["postnet.conv2.weight"] = ((md, dd, 5), "conv_w")
    spec["postnet.conv2.bias"] = ((md,), "conv_b")
    if with_out:
        spec["postnet.out.weight"] = ((md, dd), "linear_w")
        spec["postnet.out.bias"] = ((md,), ("linear_b", dd))
    for bn in ("batch_norm_list.0", "batch_norm_list.1", "batch_norm_list.2", "pre_batchnorm"):
        spec[f"postnet.{bn}.weight"] = ((dd,), "norm_w")
        spec[f"postnet.{bn}.bias"] = ((dd,), "norm_b")
        spec[f"postnet.{bn}.running_mean"] = ((dd,), "bn_mean")
        spec[f"postnet.{bn}.running_var"] = ((dd,), "bn_var")
        spec[f"postnet.{bn}.num_batches_tracked"] = ((), "bn_count")


TTS_HP: Dict[str, object] = dict(reduction_rate=2, gst=False, dropout_prenet=0.0)


def make_tts_hp(**overrides) -> SimpleNamespace:
    """hparams of the Transformer-TTS configs (BASELINE.json configs[3], [4]): the canonical dimensions with the yaml's
    reduction_rate = 2 (config/hparams_template.py) and the fields Models/transformer.py reads."""
    d = dict(CANONICAL_HP)
    d.update(TTS_HP)
    d.update(overrides)
    return SimpleNamespace(**d)


def tts_state_dict_spec(hp) -> "OrderedDict[str, tuple]":
    """Key -> (shape, kind) of `Models/transformer.py` Transformer (decoder_type 'transformer', no GST / speakers) in the
    reference's state_dict order; tests/test_oracle_tts.py checks it against the reference module itself."""
    spec: "OrderedDict[str, tuple]" = OrderedDict()
    fs2 = state_dict_spec(hp)
    for k, v in fs2.items():  # the text encoder is the FastSpeech2 one (Models/encoder.py)
        if k.startswith("encoder."):
            spec[k] = v
    de, dd = hp.d_model_encoder, hp.d_model_decoder
    if de != dd:
        spec["linear.weight"] = ((dd, de), "linear_w")
        spec["linear.bias"] = ((dd,), ("linear_b", de))
    ph = 256  # DecoderPreNet hidden_size (Models/prenets.py:12)
    spec["decoder.decoder_prenet.layer.fc1.weight"] = ((ph, hp.mel_dim), "linear_w")
    spec["decoder.decoder_prenet.layer.fc1.bias"] = ((ph,), ("linear_b", hp.mel_dim))
    spec["decoder.decoder_prenet.layer.fc2.weight"] = ((dd, ph), "linear_w")
    spec["decoder.decoder_prenet.layer.fc2.bias"] = ((dd,), ("linear_b", ph))
    spec["decoder.pe.alpha"] = ((1,), "alpha")
    for i in range(hp.n_layer_decoder):
        lp = f"decoder.layers.{i}."
        for nm in ("norm_1", "norm_2", "norm_3"):
            spec[lp + nm + ".weight"] = ((dd,), "norm_w")
            spec[lp + nm + ".bias"] = ((dd,), "norm_b")
        _attn_spec(spec, lp + "attn_1.", dd)
        _attn_spec(spec, lp + "attn_2.", dd)
        _ff_spec(spec, lp + "ff.", dd, hp.ff_conv_kernel_size_decoder)
    spec["decoder.norm.weight"] = ((dd,), "norm_w")
    spec["decoder.norm.bias"] = ((dd,), "norm_b")
    md = hp.mel_dim * hp.reduction_rate
    spec["out.weight"] = ((md, dd), "linear_w")
    spec["out.bias"] = ((md,), ("linear_b", dd))
    spec["stop_token.weight"] = ((hp.reduction_rate, dd), "linear_w")
    spec["stop_token.bias"] = ((hp.reduction_rate,), ("linear_b", dd))
    _postnet_spec(spec, dd, md, with_out=False)  # PostConvNet(prev_version=False) has no `out` (Models/postnets.py:40-41)
    return spec


def synth_tts_state_dict(hp=None, seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Deterministic fp32 Transformer-TTS state_dict in the reference layout (same generators as synth_state_dict)."""
    hp = hp or make_tts_hp()
    return _draw(tts_state_dict_spec(hp), seed)


def state_dict_spec(hp) -> "OrderedDict[str, tuple]":
    """Key -> (shape, kind) in the reference's state_dict order (SURVEY §8b, 289 entries canonical)."""
    spec: "OrderedDict[str, tuple]" = OrderedDict()

    def enc(pfx, vocab, d, n_layers, k, embedding):
        if embedding:
            spec[pfx + "embed.weight"] = ((vocab, d), "embedding")
        else:
            spec[pfx + "embed.weight"] = ((d, vocab), "linear_w")
            spec[pfx + "embed.bias"] = ((d,), ("linear_b", vocab))
        spec[pfx + "pe.alpha"] = ((1,), "alpha")
        for i in range(n_layers):
            lp = f"{pfx}layers.{i}."
            for nm in ("norm_1", "norm_2"):
                spec[lp + nm + ".weight"] = ((d,), "norm_w")
                spec[lp + nm + ".bias"] = ((d,), "norm_b")
            for nm in ("q_linear", "v_linear", "k_linear", "out"):
                spec[lp + "attn." + nm + ".weight"] = ((d, d), "linear_w")
                spec[lp + "attn." + nm + ".bias"] = ((d,), ("linear_b", d))
            spec[lp + "ff.f_1.weight"] = ((4 * d, d, k), "conv_w")
            spec[lp + "ff.f_1.bias"] = ((4 * d,), "conv_b")
            spec[lp + "ff.f_2.weight"] = ((d, 4 * d, k), "conv_w")
            spec[lp + "ff.f_2.bias"] = ((d,), "conv_b")
            spec[lp + "ff.layer_norm.weight"] = ((d,), "norm_w")
            spec[lp + "ff.layer_norm.bias"] = ((d,), "norm_b")
        spec[pfx + "norm.weight"] = ((d,), "norm_w")
        spec[pfx + "norm.bias"] = ((d,), "norm_b")

    de, dd = hp.d_model_encoder, hp.d_model_decoder
    enc("encoder.", hp.vocab_size, de, hp.n_layer_encoder, hp.ff_conv_kernel_size_encoder, True)
    fs = 256
    for nm in ("duration", "pitch", "energy"):
        pp = f"variance_adaptor.{nm}_predictor."
        spec[pp + "conv1.weight"] = ((fs, de, 3), "conv_w")
        spec[pp + "conv1.bias"] = ((fs,), "conv_b")
        spec[pp + "layer_norm1.weight"] = ((fs,), "norm_w")
        spec[pp + "layer_norm1.bias"] = ((fs,), "norm_b")
        spec[pp + "conv2.weight"] = ((fs, fs, 3), "conv_w")
        spec[pp + "conv2.bias"] = ((fs,), "conv_b")
        spec[pp + "layer_norm2.weight"] = ((fs,), "norm_w")
        spec[pp + "layer_norm2.bias"] = ((fs,), "norm_b")
        spec[pp + "linear_layer.weight"] = ((1, fs), "linear_w")
        spec[pp + "linear_layer.bias"] = ((1,), ("linear_b", fs))
        if nm != "duration":
            spec[f"variance_adaptor.{nm}_embedding.weight"] = ((hp.nbins, de), "embedding_nopad")
    enc("decoder.", de, dd, hp.n_layer_decoder, hp.ff_conv_kernel_size_decoder, False)
    md = hp.mel_dim * hp.reduction_rate
    spec["postnet.conv1.weight"] = ((dd, md, 5), "conv_w")
    spec["postnet.conv1.bias"] = ((dd,), "conv_b")
    for i in range(3):
        spec[f"postnet.conv_list.{i}.weight"] = ((dd, dd, 5), "conv_w")
        spec[f"postnet.conv_list.{i}.bias"] = ((dd,), "conv_b")
    spec["postnet.conv2.weight"] = ((md, dd, 5), "conv_w")
    spec["postnet.conv2.bias"] = ((md,), "conv_b")
    spec["postnet.out.weight"] = ((md, dd), "linear_w")
    spec["postnet.out.bias"] = ((md,), ("linear_b", dd))
    for bn in ("batch_norm_list.0", "batch_norm_list.1", "batch_norm_list.2", "pre_batchnorm"):
        spec[f"postnet.{bn}.weight"] = ((dd,), "norm_w")
        spec[f"postnet.{bn}.bias"] = ((dd,), "norm_b")
        spec[f"postnet.{bn}.running_mean"] = ((dd,), "bn_mean")
        spec[f"postnet.{bn}.running_var"] = ((dd,), "bn_var")
        spec[f"postnet.{bn}.num_batches_tracked"] = ((), "bn_count")
    return spec


def _gen(key: str, seed: int) -> torch.Generator:
    h = hashlib.sha256(f"{seed}:{key}".encode()).digest()
    g = torch.Generator(device="cpu")
    g.manual_seed(int.from_bytes(h[:7], "little"))
    return g


def synth_state_dict(hp=None, seed: int = 0, frames_per_phoneme: float = 9.6) -> "OrderedDict[str, torch.Tensor]":
    """Deterministic fp32 state_dict in the reference layout.

    ``frames_per_phoneme`` sets the duration predictor's output bias (log(x+1)); with the default
    the canonical model expands 128 phonemes to ~1000 frames (measured: mean 7.8 frames per phoneme; SURVEY §8d used log(11) with its own init)."""
    hp = hp or make_hp()
    sd = _draw(state_dict_spec(hp), seed)
    sd["variance_adaptor.duration_predictor.linear_layer.bias"].fill_(math.log(frames_per_phoneme + 1.0))
    return sd


def _draw(spec, seed: int) -> "OrderedDict[str, torch.Tensor]":
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for key, (shape, kind) in spec.items():
        g = _gen(key, seed)
        if isinstance(kind, tuple):  # ("linear_b", fan_in)
            bound = 1.0 / math.sqrt(kind[1])
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif kind == "embedding":
            t = torch.randn(shape, generator=g)
            t[0].zero_()  # padding_idx=0 (Models/encoder.py:55)
        elif kind == "embedding_nopad":
            t = torch.randn(shape, generator=g)
        elif kind == "linear_w":
            bound = 1.0 / math.sqrt(shape[1])  # kaiming_uniform(a=sqrt(5)) == U(-1/sqrt(fan_in), +)
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif kind == "conv_w":
            fan_in = shape[1] * shape[2]
            t = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_in)  # kaiming_normal_ (utils.py:168)
        elif kind == "conv_b":
            t = torch.randn(shape, generator=g) * 0.02  # reference: 0 at init; non-zero exercises the bias path
        elif kind == "norm_w":
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif kind == "norm_b":
            t = 0.05 * torch.randn(shape, generator=g)
        elif kind == "bn_mean":
            t = 0.1 * torch.randn(shape, generator=g)
        elif kind == "bn_var":
            t = 0.5 + torch.rand(shape, generator=g)
        elif kind == "bn_count":
            t = torch.tensor(100, dtype=torch.int64)
        elif kind == "alpha":
            t = torch.ones(shape)
        else:
            raise KeyError(kind)
        sd[key] = t.contiguous()
    return sd


def state_dict_digest(sd) -> str:
    h = hashlib.sha256()
    for k, v in sd.items():
        h.update(k.encode())
        h.update(v.detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def synth_text_batch(B: int, L: int, vocab: int = 152, seed: int = 1234, ragged: bool = False, min_len=None):
    """text ids (B,L) int64 in [1,vocab), pos_text (B,L) = 1..len then 0, lengths (B,) — the inputs
    `test_fastspeech2.py:146-165` feeds the model.  Ragged: lengths uniform in [min_len, L]."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    text = torch.randint(1, vocab, (B, L), generator=g)
    if ragged:
        lo = min_len if min_len is not None else max(1, L // 2)
        lens = torch.randint(lo, L + 1, (B,), generator=g)
        lens[0] = L
    else:
        lens = torch.full((B,), L, dtype=torch.int64)
    ar = torch.arange(1, L + 1).unsqueeze(0).expand(B, L)
    pos = torch.where(ar <= lens.unsqueeze(1), ar, torch.zeros_like(ar))
    text = torch.where(pos != 0, text, torch.zeros_like(text))
    return text, pos, lens


def synth_targets(pos_text: torch.Tensor, seed: int = 1234, max_dur: int = 15):
    """Teacher-forcing targets shaped like the training batch (train_fastspeech2.py:124-142):
    alignment (B,L) int64 (a few zeros on valid phonemes, 0 on padding), pos_mel (B,T), f0, energy (B,T)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed + 17)
    B, L = pos_text.shape
    dur = torch.randint(0, max_dur + 1, (B, L), generator=g)
    dur = torch.where(pos_text != 0, dur, torch.zeros_like(dur))
    mel_len = dur.sum(1)
    T = int(mel_len.max().item())
    ar = torch.arange(1, T + 1).unsqueeze(0).expand(B, T)
    pos_mel = torch.where(ar <= mel_len.unsqueeze(1), ar, torch.zeros_like(ar))
    f0 = torch.rand((B, T), generator=g) * (800.0 - 71.0) + 71.0
    energy = torch.rand((B, T), generator=g) * 400.0
    valid = pos_mel != 0
    f0 = torch.where(valid, f0, torch.zeros_like(f0))
    energy = torch.where(valid, energy, torch.zeros_like(energy))
    return dur, pos_mel, f0, energy


def synth_mel(pos_mel: torch.Tensor, mel_dim: int = 80, seed: int = 1234) -> torch.Tensor:
    """Training target mel (B,T,mel_dim): N(0,1) on valid frames, -0.5 on padding (the collate pad value,
    datasets/datasets_fastspeech2.py:728-739)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed + 29)
    B, T = pos_mel.shape
    mel = torch.randn((B, T, mel_dim), generator=g)
    return torch.where((pos_mel != 0).unsqueeze(-1), mel, torch.full_like(mel, -0.5))
