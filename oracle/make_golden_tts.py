"""ORACLE support — test infrastructure, NOT the product.

Generates tests/golden/tts_*.npz by running the UNMODIFIED reference Transformer-TTS (`Models/transformer.py`, imported in
place from /root/reference through oracle/ref_loader.py) on the synthetic weights of transformer_tts_b200/synth.py:
one teacher-forced forward (the call of train.py:201 / test.py:141) and the autoregressive loop of test.py:128-174
(restated here around the reference MODEL, because test.py is a script that runs at import).  Build container only:

    python oracle/make_golden_tts.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from oracle import ref_loader  # noqa: E402
from transformer_tts_b200 import synth  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"


def build_reference_tts(hp, sd=None):
    """The reference Transformer built the way test.py:102-109 builds it (all dropouts 0)."""
    ref_loader.import_reference()
    import Models.transformer as tr

    model = tr.Transformer(
        hp=hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder, N_e=hp.n_layer_encoder,
        n_head_encoder=hp.n_head_encoder, ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder,
        concat_after_encoder=hp.concat_after_encoder, d_model_decoder=hp.d_model_decoder, N_d=hp.n_layer_decoder,
        n_head_decoder=hp.n_head_decoder, ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder,
        concat_after_decoder=hp.concat_after_decoder, reduction_rate=hp.reduction_rate, dropout=0.0, dropout_prenet=0.0,
        dropout_postnet=0.0, multi_speaker=hp.is_multi_speaker, spk_emb_dim=hp.spk_emb_dim,
        spk_emb_architecture=hp.spk_emb_architecture)
    model.eval()
    if sd is not None:
        model.load_state_dict(sd)
    return model


def reference_create_masks(src_pos, trg_pos):
    """train.py:26-58 (the function test.py imports), with DEVICE = cpu."""
    src_mask = (src_pos != 0).unsqueeze(-2)
    trg_mask = (trg_pos != 0).unsqueeze(-2)
    size = trg_pos.size(1)
    np_mask = torch.from_numpy(np.triu(np.ones((1, size, size)), k=1).astype("uint8")) == 0
    return src_mask, trg_mask & np_mask


def reference_loop(model, hp, text, pos_text, n_iter):
    """test.py:136-166 verbatim in structure (batch 1, spk_emb None, no stop-token break): returns the last forward."""
    r = hp.reduction_rate
    dummy_input = torch.zeros((1, 1, hp.mel_dim), dtype=torch.float)
    pos_mel = torch.arange(start=1, end=2, dtype=torch.long).unsqueeze(0)
    src_mask, trg_mask = reference_create_masks(pos_text, pos_mel)
    mel_input = dummy_input
    outputs_prenet, outputs_postnet, outputs_stop_token, _, _, _ = model(text, mel_input, src_mask, trg_mask, None, training=False)
    b, t, c = outputs_prenet.shape
    outputs_postnet = outputs_postnet.reshape(b, t * r, c // r)
    mel_input = outputs_postnet[:, -1, :].unsqueeze(1)
    mel_input = torch.cat((dummy_input, mel_input), dim=1)
    for i in range(2, n_iter):
        pos_mel = torch.arange(start=1, end=i + 1, dtype=torch.long).unsqueeze(0)
        src_mask, trg_mask = reference_create_masks(pos_text, pos_mel)
        outputs_prenet, outputs_postnet, outputs_stop_token, _, _, _ = model(text, mel_input, src_mask, trg_mask, None, training=False)
        b, t, c = outputs_prenet.shape
        outputs_postnet = outputs_postnet.reshape(b, t * r, c // r)
        mel_input = outputs_postnet[:, ::r, :]
        mel_input = torch.cat((dummy_input, mel_input), dim=1)
    return outputs_postnet, outputs_stop_token


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    hp = synth.make_tts_hp()
    sd = synth.synth_tts_state_dict(hp)
    digest = synth.state_dict_digest(sd)
    model = build_reference_tts(hp, sd)
    GOLDEN.mkdir(parents=True, exist_ok=True)

    # ---- case A: teacher-forced forward, ragged batch (B=2, 14 phonemes, 11 decoder positions) ----
    text, pos, _ = synth.synth_text_batch(2, 14, ragged=True, seed=5)
    g = torch.Generator().manual_seed(6)
    T = 11
    trg = torch.randn(2, T, hp.mel_dim, generator=g)
    pos_mel = torch.arange(1, T + 1).unsqueeze(0).repeat(2, 1)
    pos_mel[1, 8:] = 0
    src_mask, trg_mask = reference_create_masks(pos, pos_mel)
    with torch.no_grad():
        pre, post, stop, attn_enc, a11, a12 = model(text, trg, src_mask, trg_mask, None, training=False)
    assert torch.equal(pre, post)  # PostConvNet(prev_version=False) returns its input (Models/postnets.py:64-79)
    np.savez_compressed(GOLDEN / "tts_teacher_forced.npz", digest=digest, text=text.numpy(), pos=pos.numpy(),
                        trg=trg.numpy(), pos_mel=pos_mel.numpy(), outputs_prenet=pre.numpy(), stop_token=stop.numpy(),
                        attn_dec_dec_l0=a11[:, 0].numpy(), attn_dec_enc_l5=a12[:, 5].numpy())

    # ---- case B: the script's autoregressive loop, batch 1, 20 phonemes, 24 decoder positions ----
    text, pos, _ = synth.synth_text_batch(1, 20, seed=8)
    with torch.no_grad():
        mel, stop = reference_loop(model, hp, text, pos, n_iter=25)  # the last forward sees a 24-position prefix
    np.savez_compressed(GOLDEN / "tts_ar_loop.npz", digest=digest, text=text.numpy(), pos=pos.numpy(), mel=mel.numpy(),
                        stop_token=stop.numpy())
    for f in sorted(GOLDEN.glob("tts_*.npz")):
        print(f.name, f.stat().st_size // 1024, "KiB")
    print("state_dict digest", digest, "keys", len(sd))


if __name__ == "__main__":
    main()
