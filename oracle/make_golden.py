"""ORACLE support — test infrastructure, NOT the product.

Generates the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(/root/reference, imported in place through oracle/ref_loader.py) on the synthetic weights and
batches of transformer_tts_b200/synth.py.  Run in the build container only:

    python oracle/make_golden.py

The reference ships no tests or fixtures of its own, so these files are what pins the
oracle (tests/test_oracle.py) — and through it the CUDA path — to the reference's behaviour.
Weights are not stored: they are regenerated from the seed and checked by digest.
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
SLOTS = ["outputs_prenet", "outputs_postnet", "log_d_prediction", "p_prediction", "e_prediction",
         "variance_adaptor_output", "text_dur_predicted", "attn_enc", "attn_dec"]


def to_np(t):  # fp32 as computed (large tensors are subsampled by the callers, never re-quantised)
    return t.detach().cpu().numpy()


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    digest = synth.state_dict_digest(sd)
    model = ref_loader.build_reference_model(hp, sd)
    GOLDEN.mkdir(parents=True, exist_ok=True)

    # ---- case A: cfg1, the reference's own inference call, B=1, 64 phonemes, free-running ----
    text, pos, _ = synth.synth_text_batch(1, 64, seed=1234)
    src_mask = (pos != 0).unsqueeze(-2)
    with torch.no_grad():
        out = model(text, src_mask)
    np.savez_compressed(GOLDEN / "fs2_cfg1_free_running.npz", digest=digest, text=text.numpy(), pos=pos.numpy(),
                        mel_pre=to_np(out[0]), mel_post=to_np(out[1]), log_d=to_np(out[2]), p=to_np(out[3]), e=to_np(out[4]),
                        va_out_sub=to_np(out[5][:, ::8, ::16]), attn_enc_l0h0=to_np(out[7][0, 0, 0]),
                        attn_dec_l5h3_sub=to_np(out[8][0, 5, 3, ::16, ::16]))

    # ---- case B: ragged batch, teacher-forced durations + f0/energy targets (training-style call) ----
    text, pos, lens = synth.synth_text_batch(3, 40, ragged=True, seed=99)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=99, max_dur=9)
    src_mask = (pos != 0).unsqueeze(-2)
    mel_mask = (pos_mel != 0).unsqueeze(-2)
    with torch.no_grad():
        out = model(text, src_mask, mel_mask=mel_mask, d_target=dur, p_target=f0, e_target=en)
        out2 = model(text, src_mask, mel_mask=mel_mask, d_target=dur)  # predicted pitch/energy drive the buckets
    np.savez_compressed(GOLDEN / "fs2_ragged_teacher_forced.npz", digest=digest, text=text.numpy(), pos=pos.numpy(),
                        dur=dur.numpy(), pos_mel=pos_mel.numpy(), f0=f0.numpy(), energy=en.numpy(),
                        mel_pre=to_np(out[0]), mel_post=to_np(out[1]), log_d=to_np(out[2]), p=to_np(out[3]), e=to_np(out[4]),
                        va_out_sub=to_np(out[5][:, ::4, ::16]), text_dur_sub=to_np(out[6][:, ::4, ::16]),
                        mel_post_pred_buckets=to_np(out2[1]))

    # ---- case C: op-level goldens from the reference's own submodules ----
    g = torch.Generator().manual_seed(7)
    x = torch.randn(2, 21, 384, generator=g)
    m = torch.ones(2, 1, 21, dtype=torch.bool)
    m[1, 0, 13:] = False
    with torch.no_grad():
        lay = model.encoder.layers[2]
        y_layer, attn = lay(x, m)
        y_ff = model.decoder.layers[1].ff(x)
        y_vp = model.variance_adaptor.pitch_predictor(x, m)
        y_pre, y_post = model.postnet(x)
        y_pe = model.encoder.pe(x)
        d = torch.tensor([[3, 0, 2, 5] + [1] * 17, [0, 0, 4, 1] + [2] * 9 + [0] * 8])
        lr_x, lr_pos = model.variance_adaptor.length_regulator(x, d)
        lr_mask = sys.modules["Models.varianceadaptor"].get_mask_from_lengths(lr_pos)
        lr_x_max, _ = model.variance_adaptor.length_regulator(x, d, 30)
    np.savez_compressed(GOLDEN / "fs2_ops.npz", digest=digest, x=x.numpy(), mask=m.numpy(),
                        enc_layer2=to_np(y_layer), enc_layer2_attn=to_np(attn), dec_ff1=to_np(y_ff), pitch_pred=to_np(y_vp),
                        post_pre=to_np(y_pre), post_post=to_np(y_post), pe=to_np(y_pe), lr_dur=d.numpy(), lr_x=to_np(lr_x),
                        lr_pos=lr_pos.numpy(), lr_mask=lr_mask.numpy(), lr_x_max30=to_np(lr_x_max),
                        pitch_bins=model.variance_adaptor.pitch_bins.cpu().numpy(),
                        energy_bins=model.variance_adaptor.energy_bins.cpu().numpy(),
                        pe_table_sub=model.encoder.pe.pe[0, ::97].numpy())
    for f in sorted(GOLDEN.glob("*.npz")):
        print(f.name, f.stat().st_size // 1024, "KiB")
    print("state_dict digest", digest)


if __name__ == "__main__":
    main()
