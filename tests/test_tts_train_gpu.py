"""Transformer-TTS teacher-forced training step on the GPU: transformer_tts_b200.tts_training.TTSTrainer
against torch autograd through the fp32 oracle (oracle/transformer_tts_oracle.py, pinned to the unmodified reference by
tests/golden/tts_teacher_forced.npz) with the losses of train.py:214-219 — at dropout 0, and at the reference's training configuration (dropout 0.1, prenet 0.5) with OUR counter-based masks handed to the oracle's named dropout call sites (the reference's own masks come from torch's
generator and cannot be reproduced; same method as tests/test_dropout_gpu.py).

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
    """What train.py:168-195 hands the model (the builder the golden script uses)."""
    from oracle.make_golden_tts_train import tts_train_batch

    return tts_train_batch(hp, B, L, frames, seed)


def oracle_grads(sd, hp, b, drop=None, dev="cpu"):
    from oracle import transformer_tts_oracle as T

    r = hp.reduction_rate
    skip = ("postnet.", "num_batches_tracked")
    leaf = {k: (v.to(dev).clone().float().requires_grad_(True)
                if v.is_floating_point() and not k.startswith(skip) and "running_" not in k else v.to(dev)) for k, v in sd.items()}
    src_mask, trg_mask = T.create_masks(b["pos"], b["pos_in"])
    c = lambda t: t.to(dev)  # noqa: E731
    out = T.transformer_forward(leaf, vars(hp), c(b["text"]), c(b["mel_input"]), c(src_mask), c(trg_mask), drop=drop)
    pre, stop = out["outputs_prenet"], out["stop_token"]
    bb, t, ch = pre.shape
    pre = pre.reshape(bb, t * r, ch // r)
    stop = stop.reshape(bb, t * r)
    mel_loss = F.l1_loss(pre, c(b["mel"][:, r:, :]))
    tok = F.binary_cross_entropy_with_logits(stop, c(b["stop"][:, r:]), reduction="mean",
                                             pos_weight=torch.tensor(POS_WEIGHT, device=dev))
    loss = 2 * mel_loss + tok                            # mel_loss + post_mel_loss (the same values) + loss_token
    loss.backward()
    grads = {k: v.grad for k, v in leaf.items() if isinstance(v, torch.Tensor) and v.requires_grad}
    return dict(loss=loss.item(), mel_loss=mel_loss.item(), loss_token=tok.item()), grads, src_mask


def build(dropout=0.0, dropout_prenet=0.0):
    from transformer_tts_b200 import synth
    from transformer_tts_b200.Models.transformer import Transformer

    hp = synth.make_tts_hp()
    sd = synth.synth_tts_state_dict(hp)
    # the constructor call of train.py:117-126
    model = Transformer(hp=hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder,
                        N_e=hp.n_layer_encoder, n_head_encoder=hp.n_head_encoder,
                        ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder, concat_after_encoder=hp.concat_after_encoder,
                        d_model_decoder=hp.d_model_decoder, N_d=hp.n_layer_decoder, n_head_decoder=hp.n_head_decoder,
                        ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder, concat_after_decoder=hp.concat_after_decoder,
                        reduction_rate=hp.reduction_rate, dropout=dropout, dropout_prenet=dropout_prenet, dropout_postnet=0.5,
                        multi_speaker=False, spk_emb_dim=None, spk_emb_architecture="")
    model.load_state_dict(sd)
    return hp, sd, model.cuda().train()


def gpu_args(b, src_mask, r):
    c = lambda t: t.cuda()  # noqa: E731
    return (c(b["text"]), c(b["mel_input"]), c(src_mask), c((b["pos_in"] != 0).to(torch.uint8)), c(b["mel"][:, r:, :].contiguous()),
            c(b["stop"][:, r:].contiguous()))


# (1, 3, 4): ONE decoder position and three phonemes; (2, 130, 520): 259 positions / 130 phonemes, both past a 128-row tile
@pytest.mark.parametrize("B,L,frames,seed", [(3, 24, 96, 11), (4, 40, 300, 12), (1, 3, 4, 13), (2, 130, 520, 14)])
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
    # a single decoder position has no averaging over rows: every gradient is ONE outer product of bf16-rounded vectors
    # carried through 12 blocks (pe.alpha, a cancelling scalar sum, is the worst); the bars of the batch cases do not apply
    bar_worst, bar_global = (2.5e-1, 8e-2) if B * ((frames - hp.reduction_rate) // hp.reduction_rate) < 16 else (1.2e-1, 3e-2)
    assert errs[0][0] < bar_worst, errs[:8]
    assert (num / den) ** 0.5 < bar_global


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


def check_grads(tr, model, grads, bar_worst, bar_global):
    floor = 1e-3 * float(sum(float((v.double() ** 2).sum()) for v in grads.values()) ** 0.5)
    errs, num, den = [], 0.0, 0.0
    for n, p in model.named_parameters():
        if n.startswith("postnet."):
            continue
        a, w = tr.G(p).double().cpu(), grads[n].double().cpu()
        errs.append((float((a - w).norm() / max(float(w.norm()), floor)), n))
        num += float(((a - w) ** 2).sum())
        den += float((w ** 2).sum())
    errs.sort(reverse=True)
    print("worst per-tensor gradient rel-L2:", errs[:6], "global", (num / den) ** 0.5)
    assert errs[0][0] < bar_worst, errs[:8]
    assert (num / den) ** 0.5 < bar_global


@pytest.mark.parametrize("dropout,dropout_prenet", [(0.1, 0.5), (0.0, 0.5), (0.1, 0.0)])
def test_tts_train_step_with_reference_dropout_matches_oracle_with_the_same_masks(dropout, dropout_prenet):
    """train.py:117-126 builds the model with dropout = hp.dropout (0.1) and the prenet's default 0.5.  Bars as in
    tests/test_dropout_gpu.py: losses 4e-3, global gradient 4e-2, worst tensor 2e-1."""
    from transformer_tts_b200.tts_training import TTSTrainer

    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    hp, sd, model = build(dropout, dropout_prenet)
    tr = TTSTrainer(model, positive_weight=POS_WEIGHT, dropout_seed=42)
    assert tr.drop.active and tr.drop.p_stack == {"enc": dropout, "dec": dropout} and tr.drop.p_prenet == dropout_prenet
    b = make_batch(hp, 3, 24, 96, 31)
    from oracle import transformer_tts_oracle as T

    src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
    args = gpu_args(b, src_mask, hp.reduction_rate)
    tr.zero_grad()
    o = tr.forward_backward(*args)
    seed = tr.last_dropout_seed
    torch.cuda.synchronize()

    def drop(name, shape):
        if name.split(".")[0] in ("enc", "dec") and dropout == 0.0:
            return torch.ones(shape, device="cuda")
        if name.startswith("prenet") and dropout_prenet == 0.0:
            return torch.ones(shape, device="cuda")
        if name.endswith(".attn2"):   # the packed-sequence launch: element (b, h, L + t, l) of a (B, H, L+T, L+T) mask
            bq, hq, tq, lq = shape
            return tr.drop.mask(name, (bq, hq, lq + tq, lq + tq), seed)[:, :, lq:, :lq].float()
        return tr.drop.mask(name, shape, seed).float()

    losses, grads, _ = oracle_grads(sd, hp, b, drop=drop, dev="cuda")
    for k in ("mel_loss", "loss_token", "loss"):
        assert abs(float(o[k]) - losses[k]) <= 4e-3 * max(1.0, abs(losses[k])), (k, float(o[k]), losses[k])
    check_grads(tr, model, grads, 2e-1, 4e-2)
    # a second forward draws new masks; the same seed reproduces the step exactly
    tr.zero_grad()
    o2 = tr.forward_backward(*args)
    assert tr.last_dropout_seed != seed and float(o2["loss"]) != float(o["loss"])
    _, _, model_b = build(dropout, dropout_prenet)
    tr_b = TTSTrainer(model_b, positive_weight=POS_WEIGHT, dropout_seed=42)
    tr_b.zero_grad()
    # (the loss sums are atomic accumulations: equal up to the order of the additions)
    assert abs(float(tr_b.forward_backward(*args)["loss"]) - float(o["loss"])) < 2e-6 * abs(float(o["loss"]))


def test_tts_reference_training_configuration_runs():
    """dropout 0.1 / prenet 0.5, Noam rate, clip 1.0: eight optimizer steps, finite losses, parameters move."""
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200.tts_training import TTSTrainer

    hp, sd, model = build(0.1, 0.5)
    b = make_batch(hp, 4, 32, 160, 22)
    src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
    tr = TTSTrainer(model, positive_weight=POS_WEIGHT)
    w0 = model.decoder.decoder_prenet.layer.fc1.weight.detach().clone()
    losses = [float(tr.train_step(*gpu_args(b, src_mask, hp.reduction_rate))["loss"]) for _ in range(8)]
    print("losses with dropout:", losses)
    assert all(l == l and abs(l) < 1e3 for l in losses)
    assert not torch.equal(w0, model.decoder.decoder_prenet.layer.fc1.weight)


def test_reference_training_call_is_a_drop_in():
    """The statements of train.py:186-259 on the drop-in module: model(...) in train() mode, the reshapes, the three loss
    terms with torch, loss.backward(), clip_grad_norm_, torch.optim.Adam.  Gradients against the oracle's autograd (dropout 0),
    then two optimizer steps with the reference's construction (dropout 0.1 / 0.5)."""
    import torch.nn as nn

    from oracle import transformer_tts_oracle as T

    hp, sd, model = build()
    r = hp.reduction_rate
    b = make_batch(hp, 3, 24, 96, 41)
    c = lambda t: t.cuda()  # noqa: E731

    def step(model):
        text, mel, stop_token = c(b["text"]), c(b["mel"]), c(b["stop"])
        mel_input = mel[:, :-r:r, :]                                             # train.py:186
        src_mask, trg_mask = T.create_masks(c(b["pos"]), c(b["pos_in"]))        # :195
        outputs_prenet, outputs_postnet, outputs_stop_token, attn_enc, attn_dec_dec, attn_dec_enc = model(
            text, mel_input, src_mask, trg_mask, None)                          # :197
        attn_dec_dec, attn_dec_enc, attn_enc = attn_dec_dec.cpu(), attn_dec_enc.cpu(), attn_enc.cpu()   # :198-200
        bb, t, ch = outputs_prenet.shape                                        # :204-208
        outputs_prenet = outputs_prenet.reshape(bb, t * r, ch // r)
        outputs_postnet = outputs_postnet.reshape(bb, t * r, ch // r)
        outputs_stop_token = outputs_stop_token.reshape(bb, t * r)
        mel_loss = nn.L1Loss()(outputs_prenet, mel[:, r:, :])                   # :211-217
        post_mel_loss = nn.L1Loss()(outputs_postnet, mel[:, r:, :])
        loss_token = F.binary_cross_entropy_with_logits(outputs_stop_token, stop_token[:, r:], reduction="mean",
                                                        pos_weight=torch.tensor(POS_WEIGHT, device="cuda"))
        loss = mel_loss + post_mel_loss + loss_token
        loss.backward()                                                         # :252
        return loss

    loss = step(model)
    losses, grads, _ = oracle_grads(sd, hp, b)
    assert abs(float(loss.detach()) - losses["loss"]) < 2e-3 * max(1.0, abs(losses["loss"]))

    class _G:   # check_grads reads gradients through .G(p)
        @staticmethod
        def G(p):
            return p.grad

    check_grads(_G, model, grads, 1.2e-1, 3e-2)
    hp, sd, model = build(0.1, 0.5)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    w0 = model.out.weight.detach().clone()
    for _ in range(2):
        opt.zero_grad()
        loss = step(model)
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)                 # :255
        opt.step()
        assert torch.isfinite(loss).item()
    assert not torch.equal(w0, model.out.weight)
    # eval() after training steps: the decode path sees the updated parameters
    model.eval()
    with torch.no_grad():
        src_mask, trg_mask = T.create_masks(c(b["pos"]), c(b["pos_in"]))
        pre, *_ = model(c(b["text"]), c(b["mel_input"]), src_mask, trg_mask, None, training=False)
    assert torch.isfinite(pre).all()


def test_tts_cuda_graphed_train_step_matches_eager():
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200.tts_training import TTSTrainer

    runs = {}
    for graphed in (False, True):
        hp, sd, model = build()
        b = make_batch(hp, 4, 32, 160, 23)
        src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
        args = gpu_args(b, src_mask, hp.reduction_rate)
        tr = TTSTrainer(model, positive_weight=POS_WEIGHT, cuda_graph=graphed)
        losses = [float(tr.train_step(*args)["loss"]) for _ in range(6)]
        assert (len(tr._graphs) == 1) == graphed
        runs[graphed] = (losses, tr.flat_p.clone())
    print("eager  ", runs[False][0])
    print("graphed", runs[True][0])
    for a, b in zip(runs[False][0], runs[True][0]):
        assert abs(a - b) <= 1e-3 * max(1.0, abs(a))
    assert float((runs[True][1] - runs[False][1]).norm() / runs[False][1].norm()) < 1e-3
    hp, sd, model = build(0.1, 0.5)
    assert not TTSTrainer(model, cuda_graph=True)._graph_ok()      # dropout: a replay would repeat the masks


def test_tts_train_step_vs_reference_golden():
    """tests/golden/tts_train_step.npz: losses and gradients of the UNMODIFIED reference module's loss.backward() on the step
    statements of train.py:186-252 (oracle/make_golden_tts_train.py).  Bars as for the FastSpeech2 golden (test_train_gpu.py)."""
    from pathlib import Path

    import numpy as np

    from oracle import transformer_tts_oracle as T
    from oracle.make_golden_tts_train import FULL, SUB, tts_train_batch
    from transformer_tts_b200 import synth
    from transformer_tts_b200.tts_training import TTSTrainer

    g = np.load(Path(__file__).resolve().parent / "golden" / "tts_train_step.npz")
    hp, sd, model = build()
    assert str(g["digest"]) == synth.state_dict_digest(sd)
    b = tts_train_batch(hp)
    src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
    tr = TTSTrainer(model, positive_weight=POS_WEIGHT)
    tr.zero_grad()
    out = tr.forward_backward(*gpu_args(b, src_mask, hp.reduction_rate))
    torch.cuda.synchronize()
    ref = g["losses"]
    for i, k in enumerate(["mel_loss", "post_mel_loss", "loss_token", "loss"]):
        assert abs(float(out[k]) - ref[i]) <= 2e-3 * max(1.0, abs(ref[i])), (k, float(out[k]), ref[i])
    bb, t, c = out["outputs_prenet"].shape
    r = hp.reduction_rate
    e_out = float((out["outputs_prenet"].reshape(bb, t * r, c // r).double().cpu() - torch.from_numpy(g["outputs_prenet"]).double()).norm()
                  / torch.from_numpy(g["outputs_prenet"]).double().norm())
    assert e_out < 1e-2, e_out
    norms = dict(zip([str(n) for n in g["names"]], g["grad_norms"]))
    floor = 1e-3 * float(g["global_grad_norm"])
    worst = []
    for n, p in model.named_parameters():
        gn = float(tr.G(p).double().norm())
        worst.append((abs(gn - norms[n]) / max(norms[n], floor), n, gn, norms[n]))
    worst.sort(reverse=True)
    print("output rel-L2", e_out, "worst gradient-norm deviations:", worst[:4])
    assert worst[0][0] < 5e-2, worst[:5]
    pd = dict(model.named_parameters())
    bad = []
    for n in FULL:
        ref_t = torch.from_numpy(g["full:" + n]).double()
        e = float((tr.G(pd[n]).double().cpu() - ref_t).norm() / max(float(ref_t.norm()), floor))
        if e >= 8e-2:
            bad.append((n, e))
    for n in SUB:
        gr = tr.G(pd[n])
        ref_t = torch.from_numpy(g["sub:" + n]).double()
        e = float((gr.reshape(gr.shape[0], -1)[::7, ::13].double().cpu() - ref_t).norm() / (float(ref_t.norm()) + 1e-30))
        if e >= 8e-2:
            bad.append((n, e))
    assert not bad, bad
    gg = float(torch.sqrt((tr.flat_g.double() ** 2).sum()))
    assert abs(gg - float(g["global_grad_norm"])) / float(g["global_grad_norm"]) < 3e-2


def test_tts_trainer_state_dict_roundtrip():
    """train.py saves model and optimizer state dicts and resumes `step` from the optimizer's: a TTSTrainer restored from
    model.state_dict() + TTSTrainer.state_dict() takes the same next step as the one that kept running (torch.optim.Adam
    layout, inherited from training.Trainer)."""
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200.tts_training import TTSTrainer

    hp, sd, model = build()
    b = make_batch(hp, 3, 24, 96, 51)
    src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
    args = gpu_args(b, src_mask, hp.reduction_rate)
    tr = TTSTrainer(model, positive_weight=POS_WEIGHT)
    for _ in range(2):
        tr.train_step(*args)
    msd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    osd = tr.state_dict()
    assert len(osd["state"]) == len(tr.params) and int(osd["state"][0]["step"]) == 2
    tr.train_step(*args)
    want = {k: v.detach().clone() for k, v in model.state_dict().items()}
    _, _, model2 = build()
    model2.load_state_dict(msd)
    tr2 = TTSTrainer(model2, positive_weight=POS_WEIGHT)
    tr2.load_state_dict(osd)
    assert tr2.step_count == 2
    tr2.train_step(*args)
    torch.cuda.synchronize()
    for k, v in model2.state_dict().items():
        if v.is_floating_point() and "running_" not in k:
            assert torch.allclose(v, want[k], rtol=0, atol=2e-6), k
    opt = torch.optim.Adam(build()[2].parameters(), lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    opt.load_state_dict({"state": osd["state"], "param_groups": [dict(opt.state_dict()["param_groups"][0])]})
    assert int(opt.state_dict()["state"][0]["step"]) == 2
