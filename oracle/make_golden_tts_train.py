"""ORACLE support — test infrastructure, NOT the product.

Golden fixture for the Transformer-TTS TRAINING step: runs the UNMODIFIED reference module
(Models/transformer.py, imported in place through oracle/ref_loader.py) in train() mode with all dropouts 0 on a seeded
synthetic batch, forms the inputs and the loss exactly as train.py:184-219 does (mel[:, :-r:r] decoder inputs, create_masks,
the reshapes, nn.L1Loss twice, binary_cross_entropy_with_logits with pos_weight), calls loss.backward() and stores the losses,
the L2 norm of every parameter gradient, and a few gradients in full / subsampled.  Build container only:

    python oracle/make_golden_tts_train.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from transformer_tts_b200 import synth  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
POS_WEIGHT = 5.0  # utils/default.yaml:76
FULL = ["encoder.pe.alpha", "decoder.pe.alpha", "decoder.norm.weight", "decoder.layers.0.norm_1.bias",
        "decoder.layers.3.norm_2.weight", "decoder.layers.5.norm_3.bias", "decoder.layers.2.attn_1.out.bias",
        "decoder.layers.4.attn_2.q_linear.bias", "decoder.layers.1.attn_2.v_linear.bias", "decoder.layers.5.ff.layer_norm.weight",
        "decoder.decoder_prenet.layer.fc1.bias", "decoder.decoder_prenet.layer.fc2.bias", "out.bias", "stop_token.weight",
        "stop_token.bias", "encoder.norm.weight"]
SUB = ["encoder.embed.weight", "encoder.layers.2.attn.k_linear.weight", "encoder.layers.5.ff.f_2.weight",
       "decoder.decoder_prenet.layer.fc1.weight", "decoder.decoder_prenet.layer.fc2.weight",
       "decoder.layers.0.attn_1.q_linear.weight", "decoder.layers.3.attn_1.v_linear.weight",
       "decoder.layers.1.attn_2.k_linear.weight", "decoder.layers.4.attn_2.v_linear.weight", "decoder.layers.5.attn_2.out.weight",
       "decoder.layers.2.ff.f_1.weight", "decoder.layers.0.ff.f_2.weight", "out.weight"]


def tts_train_batch(hp, B=3, L=24, frames=96, seed=321):
    """What train.py:168-195 hands the model: mel (B, frames, mel_dim), frames a multiple of r, ragged lengths (padding 0),
    stop-token targets 1 from the last frame on; decoder inputs mel[:, :-r:r] and their positions (:186-187)."""
    r = hp.reduction_rate
    g = torch.Generator().manual_seed(seed)
    text, pos, _ = synth.synth_text_batch(B, L, ragged=True, seed=seed)
    lens = torch.randint(frames // 2 // r, frames // r + 1, (B,), generator=g) * r
    lens[0] = frames
    pos_mel = torch.arange(1, frames + 1).unsqueeze(0).repeat(B, 1)
    pos_mel[pos_mel > lens.unsqueeze(1)] = 0
    mel = torch.randn(B, frames, hp.mel_dim, generator=g) * (pos_mel != 0).unsqueeze(-1)
    stop = (pos_mel == 0).float()
    stop[torch.arange(B), lens - 1] = 1.0
    return dict(text=text, pos=pos, mel=mel, stop=stop, mel_input=mel[:, :-r:r, :].contiguous(), pos_in=pos_mel[:, :-r:r].contiguous())


def main():
    from oracle.make_golden_tts import build_reference_tts, reference_create_masks

    torch.manual_seed(0)
    torch.set_num_threads(8)
    hp = synth.make_tts_hp()
    r = hp.reduction_rate
    sd = synth.synth_tts_state_dict(hp)
    model = build_reference_tts(hp, sd)
    model.train()  # every dropout is 0; PostConvNet's BatchNorm runs on batch statistics, its output is discarded (prev_version=False)
    b = tts_train_batch(hp)
    mel, stop_token = b["mel"], b["stop"]
    mel_input = mel[:, :-r:r, :]                                                               # train.py:186
    src_mask, trg_mask = reference_create_masks(b["pos"], b["pos_in"])                         # :195
    outputs_prenet, outputs_postnet, outputs_stop_token, _, _, _ = model(b["text"], mel_input, src_mask, trg_mask, None)  # :197
    bb, t, c = outputs_prenet.shape                                                            # :204-208
    outputs_prenet = outputs_prenet.reshape(bb, t * r, c // r)
    outputs_postnet = outputs_postnet.reshape(bb, t * r, c // r)
    outputs_stop_token = outputs_stop_token.reshape(bb, t * r)
    mel_loss = nn.L1Loss()(outputs_prenet, mel[:, r:, :])                                      # :211-217
    post_mel_loss = nn.L1Loss()(outputs_postnet, mel[:, r:, :])
    loss_token = F.binary_cross_entropy_with_logits(outputs_stop_token, stop_token[:, r:], reduction="mean",
                                                    pos_weight=torch.tensor(POS_WEIGHT))
    loss = mel_loss + post_mel_loss
    loss += loss_token
    loss.backward()                                                                            # :252
    names, norms, total = [], [], 0.0
    for n, p in model.named_parameters():
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        names.append(n)
        norms.append(float(g.double().norm()))
        total += float((g.double() ** 2).sum())
    save = dict(digest=synth.state_dict_digest(sd),
                losses=np.array([x.item() for x in (mel_loss, post_mel_loss, loss_token, loss)], dtype=np.float64),
                outputs_prenet=outputs_prenet.detach().numpy(), stop_logits=outputs_stop_token.detach().numpy(),
                names=np.array(names), grad_norms=np.array(norms), global_grad_norm=np.array(total ** 0.5))
    pd = dict(model.named_parameters())
    for n in FULL:
        save["full:" + n] = pd[n].grad.numpy()
    for n in SUB:
        g = pd[n].grad
        save["sub:" + n] = g.reshape(g.shape[0], -1)[::7, ::13].numpy()
    f = GOLDEN / "tts_train_step.npz"
    np.savez_compressed(f, **save)
    print(f.name, f.stat().st_size // 1024, "KiB; loss", loss.item(), "global grad norm", total ** 0.5,
          "postnet grads all None/zero:", all(p.grad is None or float(p.grad.abs().max()) == 0.0 for n, p in pd.items() if n.startswith("postnet.")))


if __name__ == "__main__":
    main()
