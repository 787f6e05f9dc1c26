"""ORACLE support — test infrastructure, NOT the product.

Golden fixture for the TRAINING step: runs the UNMODIFIED reference model (imported in place through
oracle/ref_loader.py) in train() mode with all dropouts 0 on a seeded synthetic batch, forms the loss exactly as
train_fastspeech2.py:212-259 does (five unmasked nn.L1Loss terms), calls loss.backward() and stores the losses,
the L2 norm of every parameter gradient, and a few gradients in full / subsampled.  Build container only:

    python oracle/make_golden_train.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch
import torch.nn as nn

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from oracle import ref_loader  # noqa: E402
from transformer_tts_b200 import synth  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
FULL = ["encoder.pe.alpha", "decoder.pe.alpha", "encoder.norm.weight", "decoder.norm.bias",
        "encoder.layers.0.norm_1.weight", "decoder.layers.5.ff.layer_norm.bias", "decoder.layers.2.attn.out.bias",
        "variance_adaptor.duration_predictor.linear_layer.weight", "variance_adaptor.pitch_predictor.linear_layer.bias",
        "variance_adaptor.energy_predictor.layer_norm1.weight", "postnet.pre_batchnorm.weight",
        "postnet.batch_norm_list.2.bias", "postnet.out.bias", "postnet.conv2.bias"]
SUB = ["encoder.embed.weight", "encoder.layers.3.attn.q_linear.weight", "encoder.layers.1.ff.f_1.weight",
       "decoder.layers.0.attn.v_linear.weight", "decoder.layers.4.ff.f_2.weight", "decoder.embed.weight",
       "variance_adaptor.pitch_embedding.weight", "variance_adaptor.energy_predictor.conv1.weight",
       "postnet.conv1.weight", "postnet.conv_list.1.weight", "postnet.conv2.weight", "postnet.out.weight"]


def train_batch(B=3, L=24, seed=321, max_dur=8):
    text, pos, _ = synth.synth_text_batch(B, L, ragged=True, seed=seed)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=seed, max_dur=max_dur)
    mel = synth.synth_mel(pos_mel, 80, seed=seed)
    return text, pos, dur, pos_mel, f0, en, mel


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    digest = synth.state_dict_digest(sd)
    model = ref_loader.build_reference_model(hp, sd)
    model.train()  # BatchNorm batch statistics; every dropout is 0
    text, pos, dur, pos_mel, f0, en, mel = train_batch()
    src_mask, mel_mask = (pos != 0).unsqueeze(-2), (pos_mel != 0).unsqueeze(-2)
    out = model(text, src_mask, mel_mask, dur, f0, en)  # train_fastspeech2.py:167 (positional, accent=None)
    mel_loss = nn.L1Loss()(out[0], mel)
    post_mel_loss = nn.L1Loss()(out[1], mel)
    duration_loss = nn.L1Loss()(out[2], torch.log(dur.float() + 1))
    f0_loss = nn.L1Loss()(out[3], f0)
    energy_loss = nn.L1Loss()(out[4], en)
    loss = mel_loss + post_mel_loss
    loss = loss + f0_loss
    loss = loss + energy_loss
    loss = loss + duration_loss
    loss.backward()
    names, norms = [], []
    save = dict(digest=digest, losses=np.array([x.item() for x in (mel_loss, post_mel_loss, duration_loss, f0_loss,
                                                                  energy_loss, loss)], dtype=np.float64),
                mel_pre=out[0].detach().numpy(), mel_post=out[1].detach().numpy())
    total = 0.0
    for n, p in model.named_parameters():
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        names.append(n)
        norms.append(float(g.double().norm()))
        total += float((g.double() ** 2).sum())
    save["names"] = np.array(names)
    save["grad_norms"] = np.array(norms)
    save["global_grad_norm"] = np.array(total ** 0.5)
    pd = dict(model.named_parameters())
    for n in FULL:
        save["full:" + n] = pd[n].grad.numpy()
    for n in SUB:
        g = pd[n].grad
        save["sub:" + n] = g.reshape(g.shape[0], -1)[::7, ::13].numpy()
    # BatchNorm running statistics after this one training forward
    save["bn_running_mean"] = model.postnet.pre_batchnorm.running_mean.numpy()
    save["bn_running_var"] = model.postnet.pre_batchnorm.running_var.numpy()
    np.savez_compressed(GOLDEN / "fs2_train_step.npz", **save)
    f = GOLDEN / "fs2_train_step.npz"
    print(f.name, f.stat().st_size // 1024, "KiB; loss", loss.item(), "global grad norm", total ** 0.5)


if __name__ == "__main__":
    main()
