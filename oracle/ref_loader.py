"""ORACLE support — test infrastructure, NOT the product.

Imports the UNMODIFIED reference (`/root/reference/Models/*`) in the build container.  The reference
does not import as shipped: two stray IDE auto-imports (`itsdangerous.NoneAlgorithm`,
`turtle.distance`) are stubbed in ``sys.modules``; neither name is used.  `/root/reference` does not
exist on the GPU box, so only ``oracle/make_golden.py`` and the build-container-only tests use this.
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("TTSB_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "Models"))


def import_reference():
    """Returns the reference's `Models.fastspeech2` module (and leaves `Models`, `utils` importable)."""
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    if "itsdangerous" not in sys.modules:
        m = types.ModuleType("itsdangerous")
        m.NoneAlgorithm = object
        sys.modules["itsdangerous"] = m
    if "turtle" not in sys.modules:
        m = types.ModuleType("turtle")
        m.distance = lambda *a, **k: None
        sys.modules["turtle"] = m
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import Models.fastspeech2 as fs2  # noqa: E402

    return fs2


def build_reference_model(hp, sd=None):
    """The reference FastSpeech2 built the way test_fastspeech2.py:115-122 builds it (all dropouts 0)."""
    fs2 = import_reference()
    model = fs2.FastSpeech2(
        hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder,
        N_e=hp.n_layer_encoder, n_head_encoder=hp.n_head_encoder,
        ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder, concat_after_encoder=hp.concat_after_encoder,
        d_model_decoder=hp.d_model_decoder, N_d=hp.n_layer_decoder, n_head_decoder=hp.n_head_decoder,
        ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder, concat_after_decoder=hp.concat_after_decoder,
        reduction_rate=hp.reduction_rate, dropout=0.0, dropout_postnet=0.0, dropout_variance_adaptor=0.0,
        n_bins=hp.nbins, f0_min=hp.f0_min, f0_max=hp.f0_max, energy_min=hp.energy_min, energy_max=hp.energy_max,
        pitch_pred=hp.pitch_pred, energy_pred=hp.energy_pred, accent_emb=hp.accent_emb,
        output_type=hp.output_type, num_group=hp.num_group, multi_speaker=hp.is_multi_speaker,
        spk_emb_dim=hp.spk_emb_dim, spk_emb_architecture=hp.spk_emb_architecture)
    model.eval()
    if sd is not None:
        model.load_state_dict(sd)
    return model
