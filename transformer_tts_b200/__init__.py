"""B200-native (sm_100a) FastSpeech2 mel-generation path behind the reference's module API.

``transformer_tts_b200.Models`` mirrors the reference's ``Models`` package for the hot path;
``transformer_tts_b200.lib`` is the ctypes binding of the CUDA library (C ABI: include/ttsb200.h).
"""
from __future__ import annotations

__all__ = ["build_fastspeech2", "build_transformer_tts"]


def build_fastspeech2(hp, dropout: float = 0.0, dropout_postnet: float = 0.0, dropout_variance_adaptor: float = 0.0):
    """Construct FastSpeech2 from an hparams object the way test_fastspeech2.py:115-122 does."""
    from .Models.fastspeech2 import FastSpeech2

    g = lambda k, d=None: getattr(hp, k, d)  # noqa: E731
    return FastSpeech2(
        hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder,
        N_e=hp.n_layer_encoder, n_head_encoder=hp.n_head_encoder,
        ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder, concat_after_encoder=hp.concat_after_encoder,
        d_model_decoder=hp.d_model_decoder, N_d=hp.n_layer_decoder, n_head_decoder=hp.n_head_decoder,
        ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder, concat_after_decoder=hp.concat_after_decoder,
        reduction_rate=hp.reduction_rate, dropout=dropout, dropout_postnet=dropout_postnet,
        dropout_variance_adaptor=dropout_variance_adaptor, n_bins=hp.nbins, f0_min=hp.f0_min, f0_max=hp.f0_max,
        energy_min=hp.energy_min, energy_max=hp.energy_max, pitch_pred=g("pitch_pred", True),
        energy_pred=g("energy_pred", True), accent_emb=g("accent_emb", False), output_type=g("output_type"),
        num_group=g("num_group"), multi_speaker=g("is_multi_speaker", False), spk_emb_dim=g("spk_emb_dim"),
        spk_emb_architecture=g("spk_emb_architecture", ""))


def build_transformer_tts(hp):
    """Construct the Transformer-TTS model from an hparams object the way test.py:102-109 does (all dropouts 0)."""
    from .Models.transformer import Transformer

    g = lambda k, d=None: getattr(hp, k, d)  # noqa: E731
    return Transformer(
        hp=hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder, N_e=hp.n_layer_encoder,
        n_head_encoder=hp.n_head_encoder, ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder,
        concat_after_encoder=hp.concat_after_encoder, d_model_decoder=hp.d_model_decoder, N_d=hp.n_layer_decoder,
        n_head_decoder=hp.n_head_decoder, ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder,
        concat_after_decoder=hp.concat_after_decoder, reduction_rate=hp.reduction_rate, dropout=0.0, dropout_prenet=0.0,
        dropout_postnet=0.0, multi_speaker=g("is_multi_speaker", False), spk_emb_dim=g("spk_emb_dim"),
        spk_emb_architecture=g("spk_emb_architecture", ""))
