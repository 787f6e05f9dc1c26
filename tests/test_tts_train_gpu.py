"""Transformer-TTS teacher-forced training step on the GPU: transformer_tts_b200.tts_training.TTSTrainer
against torch autograd through the fp32 oracle (oracle/transformer_tts_oracle.py, pinned to the unmodified reference by
tests/golden/tts_teacher_forced.npz) with the losses of train.py:214-219, dropout 0.

Tolerances: as for the FastSpeech2 training step (tests/test_train_gpu.py): bf16 operands with fp32 accumulation on both
passes; losses 2e-3 relative, per-tensor gradient relative L2 <= 1.2e-1 against max(reference norm, 1e-3 of the global norm),
global gradient vector <= 3e-2.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

POS_WEIGHT = 5.0  # utils/default.yaml:76


def make_batch(hp, B, L, frames, seed):
    """What train.py:168-195 hands the model: mel (B, frames, mel_dim) with `frames` a multiple of r, ragged lengths."""
    from transformer_tts_b200 import synth

    r = hp.reduction_rate
    g = torch.Generator().manual_seed(seed)
    text, pos, _ = synth.synth_text_batch(B, L, ragged=True, seed=seed)
    lens = torch.randint(frames // 2 // r, frames // r + 1, (B,), generator=g) * r
    lens[0] = frames
    pos_mel = torch.arange(1, frames + 1).unsqueeze(0).repeat(B, 1)
    pos_mel[pos_mel > lens.unsqueeze(1)] = 0
    mel = torch.randn(B, frames, hp.mel_dim, generator=g) * (pos_mel != 0).unsqueeze(-1)
    stop = (pos_mel == 0).float()                       # 1 on the padding frames ...
    stop[torch.arange(B), lens - 1] = 1.0               # ... and on the last frame (utils: the stop token target)
    mel_input = mel[:, :-r:r, :].contiguous()           # train.py:186
    pos_in = pos_mel[:, :-r:r].contiguous()             # train.py:187
    return dict(text=text, pos=pos, mel=mel, stop=stop, mel_input=mel_input, pos_in=pos_in)


def oracle_grads(sd, hp, b):
    from oracle import transformer_tts_oracle as T

    r = hp.reduction_rate
    skip = ("postnet.", "num_batches_tracked")
    leaf = {k: (v.clone().float().requires_grad_(True) if v.is_floating_point() and not k.startswith(skip) and "running_" not in k
                else v) for k, v in sd.items()}
    src_mask, trg_mask = T.create_masks(b["pos"], b["pos_in"])
    out = T.transformer_forward(leaf, vars(hp), b["text"], b["mel_input"], src_mask, trg_mask)
    pre, stop = out["outputs_prenet"], out["stop_token"]
    bb, t, c = pre.shape
    pre = pre.reshape(bb, t * r, c // r)
    stop = stop.reshape(bb, t * r)
    mel_loss = F.l1_loss(pre, b["mel"][:, r:, :])
    tok = F.binary_cross_entropy_with_logits(stop, b["stop"][:, r:], reduction="mean", pos_weight=torch.tensor(POS_WEIGHT))
    loss = 2 * mel_loss + tok                            # mel_loss + post_mel_loss (the same values) + loss_token
    loss.backward()
    grads = {k: v.grad for k, v in leaf.items() if isinstance(v, torch.Tensor) and v.requires_grad}
    return dict(loss=loss.item(), mel_loss=mel_loss.item(), loss_token=tok.item()), grads, src_mask


def build():
    from transformer_tts_b200 import build_transformer_tts, synth

    hp = synth.make_tts_hp()
    sd = synth.synth_tts_state_dict(hp)
    model = build_transformer_tts(hp)
    model.load_state_dict(sd)
    return hp, sd, model.cuda().train()


def gpu_args(b, src_mask, r):
    c = lambda t: t.cuda()  # noqa: E731
    return (c(b["text"]), c(b["mel_input"]), c(src_mask), c((b["pos_in"] != 0).to(torch.uint8)), c(b["mel"][:, r:, :].contiguous()),
            c(b["stop"][:, r:].contiguous()))


@pytest.mark.parametrize("B,L,frames,seed", [(3, 24, 96, 11), (4, 40, 300, 12)])
def test_tts_train_step_vs_oracle_autograd(B, L, frames, seed):
    from transformer_tts_b200.tts_training import TTSTrainer

    hp, sd, model = build()
    b = make_batch(hp, B, L, frames, seed)
    torch.set_num_threads(8)
    losses, grads, src_mask = oracle_grads(sd, hp, b)
    tr = TTSTrainer(model, positive_weight=POS_WEIGHT)
    tr.zero_grad()
    out = tr.forward_backward(*gpu_args(b, src_mask, hp.reduction_rate))
    torch.cuda.synchronize()
    for k in ("mel_loss", "loss_token", "loss"):
        assert abs(float(out[k]) - losses[k]) <= 2e-3 * max(1.0, abs(losses[k])), (k, float(out[k]), losses[k])
    floor = 1e-3 * float(sum(float((v.double() ** 2).sum()) for v in grads.values()) ** 0.5)
    errs, num, den = [], 0.0, 0.0
    for n, p in model.named_parameters():
        if n.startswith("postnet."):
            assert float(tr.G(p).abs().max()) == 0.0    # PostConvNet(prev_version=False) is outside the loss graph
            continue
        a, w = tr.G(p).double().cpu(), grads[n].double()
        errs.append((float((a - w).norm() / max(float(w.norm()), floor)), n))
        num += float(((a - w) ** 2).sum())
        den += float((w ** 2).sum())
    errs.sort(reverse=True)
    print("worst per-tensor gradient rel-L2:", errs[:8], "global", (num / den) ** 0.5)
    assert errs[0][0] < 1.2e-1, errs[:8]
    assert (num / den) ** 0.5 < 3e-2


def test_tts_optimizer_steps_reduce_loss():
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200.tts_training import TTSTrainer

    hp, sd, model = build()
    b = make_batch(hp, 4, 32, 160, 21)
    src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
    args = gpu_args(b, src_mask, hp.reduction_rate)
    tr = TTSTrainer(model, positive_weight=POS_WEIGHT, warmup_step=4000)
    w0 = model.decoder.layers[0].attn_2.k_linear.weight.detach().clone()
    losses = [float(tr.train_step(*args)["loss"]) for _ in range(8)]
    print("losses:", losses)
    assert losses[-1] < losses[0]
    assert not torch.equal(w0, model.decoder.layers[0].attn_2.k_linear.weight)
    # after the update the eval-mode forward (decode path) sees the new weights
    model.eval()
    with torch.no_grad():
        _, trg_mask = T.create_masks(b["pos"], b["pos_in"])
        pre, *_ = model(args[0], args[1], args[2], trg_mask.cuda(), None, training=False)
    assert torch.isfinite(pre).all()


def test_tts_trainer_rejects_dropout():
    from transformer_tts_b200 import synth
    from transformer_tts_b200.Models.transformer import Transformer
    from transformer_tts_b200.tts_training import TTSTrainer

    hp = synth.make_tts_hp()
    m = Transformer(hp, hp.vocab_size, hp.mel_dim, 384, 1, 4, 5, False, 384, 1, 4, 1, False, 2, dropout=0.1).cuda()
    with pytest.raises(NotImplementedError, match="dropout"):
        TTSTrainer(m)
