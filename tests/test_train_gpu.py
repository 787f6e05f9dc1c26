"""Training-step parity on the GPU: transformer_tts_b200.training.Trainer against
(1) the golden produced by the UNMODIFIED reference's loss.backward() (oracle/make_golden_train.py) and
(2) the fp32 oracle's autograd on a second seeded batch.

Tolerances.  Forward/backward run in bf16 on the tensor cores with fp32 accumulation; the reference's own
bf16-autocast floor on forward outputs is 7e-3 rel-L2.  Gradients pass through 12 attention + 24
conv GEMM backward stages each rounding to bf16, and the L1 loss gradient is sign(pred - target) (a few signs flip
when pred is perturbed).  The reference's OWN bf16-autocast floor on the batch of the second test, measured by
running the oracle under torch.autocast(bfloat16) against its fp32 self, is 1.88e-2 on the global gradient vector
and 8.6e-2 on the worst tensor (duration_predictor.conv1.bias, only B*L = 160 rows); this path measures 1.72e-2 /
8.6e-2 on the same batch.  Bars: losses 2e-3 relative, per-tensor gradient rel-L2 <= 1.2e-1 (1.4x that floor),
global gradient vector rel-L2 <= 3e-2, gradient norms within 5 %.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from pathlib import Path

GOLDEN = Path(__file__).resolve().parent / "golden"


def build(seed_sd=0):
    from transformer_tts_b200 import build_fastspeech2, lib, synth

    lib.load()
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp, seed=seed_sd)
    model = build_fastspeech2(hp)
    model.load_state_dict(sd)
    model = model.to("cuda:0").train()
    return hp, sd, model


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def run_step(trainer, batch):
    text, pos, dur, pos_mel, f0, en, mel = batch
    c = lambda t: t.cuda()  # noqa: E731
    trainer.zero_grad()
    return trainer.forward_backward(c(text), c((pos != 0).unsqueeze(-2)), c((pos_mel != 0).unsqueeze(-2)), c(dur),
                                    c(f0), c(en), c(mel))


def test_train_step_vs_reference_golden():
    from oracle.make_golden_train import FULL, SUB, train_batch
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    g = np.load(GOLDEN / "fs2_train_step.npz")
    assert str(g["digest"]) == synth.state_dict_digest(sd)
    tr = Trainer(model)
    out = run_step(tr, train_batch())
    torch.cuda.synchronize()
    ref = g["losses"]
    for i, k in enumerate(["mel_loss", "post_mel_loss", "duration_loss", "f0_loss", "energy_loss", "loss"]):
        assert abs(float(out[k]) - ref[i]) <= 2e-3 * max(1.0, abs(ref[i])), (k, float(out[k]), ref[i])
    assert rel(out["outputs_postnet"], torch.from_numpy(g["mel_post"])) < 1e-2
    norms = dict(zip([str(n) for n in g["names"]], g["grad_norms"]))
    worst = []
    # k_linear.bias gradients are identically zero in exact arithmetic (softmax is invariant to a per-row constant):
    # deviations are measured against max(reference norm, 1e-3 of the global gradient norm)
    floor = 1e-3 * float(g["global_grad_norm"])
    for n, p in model.named_parameters():
        gn = float(tr.G(p).double().norm())
        worst.append((abs(gn - norms[n]) / max(norms[n], floor), n, gn, norms[n]))
    worst.sort(reverse=True)
    print("worst gradient-norm deviations:", worst[:5])
    assert worst[0][0] < 5e-2, worst[:5]
    pd = dict(model.named_parameters())
    bad = []
    for n in FULL:
        ref_t = torch.from_numpy(g["full:" + n])
        # pe.alpha gradients are scalars formed by heavily cancelling sums: measured against the same floor
        e = float((tr.G(pd[n]).double().cpu() - ref_t.double()).norm() / max(float(ref_t.double().norm()), floor))
        if e >= 8e-2:
            bad.append((n, e))
    for n in SUB:
        gr = tr.G(pd[n])
        e = rel(gr.reshape(gr.shape[0], -1)[::7, ::13], torch.from_numpy(g["sub:" + n]))
        if e >= 8e-2:
            bad.append((n, e))
    assert not bad, bad
    gg = float(torch.sqrt((tr.flat_g.double() ** 2).sum()))
    assert abs(gg - float(g["global_grad_norm"])) / float(g["global_grad_norm"]) < 3e-2


def test_train_step_vs_oracle_autograd():
    from oracle import fastspeech2_oracle as O
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    text, pos, _ = synth.synth_text_batch(4, 40, ragged=True, seed=77)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=77, max_dur=12)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=77)
    torch.set_num_threads(8)
    losses, grads, _ = O.training_step_grads(sd, vars(hp), text, (pos != 0).unsqueeze(-2),
                                             (pos_mel != 0).unsqueeze(-2), dur, f0, en, mel)
    tr = Trainer(model)
    out = run_step(tr, (text, pos, dur, pos_mel, f0, en, mel))
    torch.cuda.synchronize()
    for k in ["mel_loss", "post_mel_loss", "duration_loss", "f0_loss", "energy_loss", "loss"]:
        assert abs(float(out[k]) - float(losses[k])) <= 2e-3 * max(1.0, abs(float(losses[k]))), k
    errs = []
    num = den = 0.0
    floor = 1e-3 * float(sum(float((v.double() ** 2).sum()) for v in grads.values()) ** 0.5)
    for n, p in model.named_parameters():
        a, b = tr.G(p).double().cpu(), grads[n].double()
        errs.append((float((a - b).norm() / max(float(b.norm()), floor)), n))
        num += float(((a - b) ** 2).sum())
        den += float((b ** 2).sum())
    errs.sort(reverse=True)
    print("worst per-tensor gradient rel-L2:", errs[:8], "global", (num / den) ** 0.5)
    assert errs[0][0] < 1.2e-1, errs[:8]
    assert (num / den) ** 0.5 < 3e-2


def test_optimizer_steps_reduce_loss_and_update_running_stats():
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer, noam_lr

    hp, sd, model = build()
    text, pos, _ = synth.synth_text_batch(4, 32, ragged=True, seed=5)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=5, max_dur=10)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=5)
    c = lambda t: t.cuda()  # noqa: E731
    args = (c(text), c((pos != 0).unsqueeze(-2)), c((pos_mel != 0).unsqueeze(-2)), c(dur), c(f0), c(en), c(mel))
    tr = Trainer(model, warmup_step=4000)
    rm0 = model.postnet.pre_batchnorm.running_mean.clone()
    w0 = model.decoder.layers[0].attn.q_linear.weight.detach().clone()
    losses = []
    for i in range(6):
        o = tr.train_step(*args)
        losses.append(float(o["mel_loss"]) + float(o["post_mel_loss"]) + float(o["duration_loss"]))
        assert abs(o["lr"] - noam_lr(i + 1, 384, 1.0, 4000)) < 1e-12
    assert all(np.isfinite(losses))
    assert losses[-1] < losses[0], losses
    assert not torch.equal(w0, model.decoder.layers[0].attn.q_linear.weight.detach())
    assert not torch.equal(rm0, model.postnet.pre_batchnorm.running_mean)
    assert int(model.postnet.pre_batchnorm.num_batches_tracked) == 100 + 6
    # the state_dict keeps the reference layout and reloads into a fresh model
    sd2 = model.state_dict()
    assert len(sd2) == 289 and sd2["decoder.layers.0.attn.q_linear.weight"].shape == (384, 384)


def test_train_step_matches_torch_adam_on_own_gradients():
    """clip_grad_norm_(1.0) + Adam(betas=(0.9,0.98), eps=1e-9) at the Noam rate (train_fastspeech2.py:117-120,304-315)
    applied by the fused kernel equals torch's implementation applied to the same gradients."""
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    text, pos, _ = synth.synth_text_batch(2, 16, ragged=True, seed=9)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=9, max_dur=6)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=9)
    tr = Trainer(model)
    run_step(tr, (text, pos, dur, pos_mel, f0, en, mel))
    p_ref = tr.flat_p.clone().requires_grad_(True)
    p_ref.grad = tr.flat_g.clone()
    opt = torch.optim.Adam([p_ref], lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    torch.nn.utils.clip_grad_norm_([p_ref], 1.0)
    lr = tr.optimizer_step()
    for gp in opt.param_groups:
        gp["lr"] = lr
    opt.step()
    torch.cuda.synchronize()
    assert (tr.flat_p - p_ref.detach()).abs().max().item() < 1e-6


def test_train_step_radam_matches_oracle_on_own_gradients():
    """hp.optimizer == 'radam' (train_fastspeech2.py:412-414, radam.py:56-92): fixed rate 1e-3 (no Noam schedule,
    :117-120), clip_grad_norm_(1.0) first.  The fused kernel equals the oracle's restatement on the same gradients."""
    from oracle import fastspeech2_oracle as O
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    text, pos, _ = synth.synth_text_batch(2, 16, ragged=True, seed=9)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=9, max_dur=6)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=9)
    tr = Trainer(model, optimizer="radam")
    p, m, v = tr.flat_p.cpu().clone(), torch.zeros(tr.total), torch.zeros(tr.total)
    for step in (1, 2):
        tr.zero_grad()
        run_step(tr, (text, pos, dur, pos_mel, f0, en, mel))
        g = tr.flat_g.cpu().clone()
        g = g * min(1.0, 1.0 / (g.norm().item() + 1e-6))
        lr = tr.optimizer_step()
        assert lr == 1e-3
        p, m, v = O.radam_step(p, g, m, v, step, lr=1e-3)
        torch.cuda.synchronize()
        assert (tr.flat_p.cpu() - p).abs().max().item() < 2e-6, step


def test_dropin_autograd_mode_matches_trainer():
    """The reference's own loop shape (train_fastspeech2.py:167-315): model.train(); model(...) ; nn.L1Loss terms;
    loss.backward(); clip_grad_norm_; torch.optim.Adam.step() — with this package's modules in place of Models/."""
    import torch.nn as nn

    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    text, pos, _ = synth.synth_text_batch(3, 24, ragged=True, seed=21)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=21, max_dur=8)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=21)
    c = lambda t: t.cuda()  # noqa: E731
    text, src_mask, mel_mask, dur, f0, en, mel = (c(t) for t in (text, (pos != 0).unsqueeze(-2),
                                                                 (pos_mel != 0).unsqueeze(-2), dur, f0, en, mel))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    out = model(text, src_mask, mel_mask, dur, f0, en)
    assert len(out) == 14 and out[0].requires_grad and out[7] is None
    loss = nn.L1Loss()(out[0], mel) + nn.L1Loss()(out[1], mel)
    loss = loss + nn.L1Loss()(out[3], f0) + nn.L1Loss()(out[4], en)
    loss = loss + nn.L1Loss()(out[2], torch.log(dur.float() + 1))
    loss.backward()
    grads = {n: p.grad.detach().clone() for n, p in model.named_parameters()}
    assert all(g is not None for g in grads.values()) and len(grads) == 277
    w0 = model.postnet.conv2.weight.detach().clone()
    torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
    opt.step()
    assert not torch.equal(w0, model.postnet.conv2.weight.detach())
    out2 = model(text, src_mask, mel_mask, dur, f0, en)  # derived bf16 operands were refreshed (version counters)
    assert not torch.equal(out2[1], out[1])
    # same kernels through the Trainer on a fresh copy of the weights
    _, _, model2 = build()
    tr = Trainer(model2)
    tr.zero_grad()
    o = tr.forward_backward(text, src_mask, mel_mask, dur, f0, en, mel)
    torch.cuda.synchronize()
    assert abs(float(o["loss"]) - float(loss.detach())) < 1e-3 * abs(float(loss.detach()))
    num = den = 0.0
    for n, p in model2.named_parameters():
        num += float(((tr.G(p) - grads[n]).double() ** 2).sum())
        den += float((grads[n].double() ** 2).sum())
    assert (num / den) ** 0.5 < 2e-3  # identical kernels; only fp32 atomic order and BN running stats differ
    model.eval()
    with torch.no_grad():
        ev = model(text[:1], src_mask[:1])
    assert ev[1].shape[-1] == hp.mel_dim


def test_trainer_state_dict_roundtrip_resumes_bit_identically():
    """train_fastspeech2.py:323,444 saves optimizer.state_dict() and resumes `step` from it: a Trainer restored from
    model.state_dict() + Trainer.state_dict() must take the SAME next step as the one that kept running, and the
    exported state must load into a real torch.optim.Adam over the same parameters."""
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    text, pos, _ = synth.synth_text_batch(3, 24, ragged=True, seed=5)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=5, max_dur=8)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=5)
    c = lambda t: t.cuda()  # noqa: E731
    batch = dict(text=c(text), src_mask=c((pos != 0).unsqueeze(-2)), mel_mask=c((pos_mel != 0).unsqueeze(-2)),
                 alignment=c(dur), f0=c(f0), energy=c(en), mel=c(mel))
    tr = Trainer(model, optimizer="adam")
    for _ in range(2):
        tr.train_step(**batch)
    msd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    osd = tr.state_dict()
    assert set(osd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"} and len(osd["state"]) == len(tr.params)
    assert int(osd["state"][0]["step"]) == 2 and osd["param_groups"][0]["params"] == list(range(len(tr.params)))
    tr.train_step(**batch)  # the run that kept going
    want = {k: v.detach().clone() for k, v in model.state_dict().items()}

    _, _, model2 = build(seed_sd=1)
    model2.load_state_dict(msd)
    tr2 = Trainer(model2, optimizer="adam")
    tr2.load_state_dict(osd)
    assert tr2.step_count == 2
    tr2.train_step(**batch)
    torch.cuda.synchronize()
    for k, v in model2.state_dict().items():
        if v.is_floating_point() and "running_" not in k:
            # identical kernels and inputs; only the order of fp32 atomics in the weight gradients differs run to run
            assert torch.allclose(v, want[k], rtol=0, atol=2e-6), k
    # the exported layout is torch.optim.Adam's
    _, _, model3 = build()
    opt = torch.optim.Adam(model3.parameters(), lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    opt.load_state_dict({"state": osd["state"], "param_groups": [dict(opt.state_dict()["param_groups"][0])]})
    assert int(opt.state_dict()["state"][0]["step"]) == 2


def test_reference_training_call_signature_runs_unchanged():
    """The exact call of train_fastspeech2.py:167 on the exact batch tuple `collate_fn` returns
    (datasets_fastspeech2.py:521-616): `accent`, `spkr_emb`, `fix_mask=hp.fix_mask` (None), `temperature` (None) and
    `hop_size` — a Python list of None when hp.use_hop is off — are all passed on every step.  The model is built with
    the kwargs of train_fastspeech2.py:381-389.  One optimizer step must run and the loss must be finite."""
    import torch.nn as nn

    from transformer_tts_b200 import data, synth
    from transformer_tts_b200.Models.fastspeech2 import FastSpeech2

    hp = synth.make_hp(dropout=0.1, dropout_variance_adaptor=0.5, fix_mask=None, use_hop=False)  # utils/default.yaml:73, fill_variables
    model = FastSpeech2(hp=hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder,
                        N_e=hp.n_layer_encoder, n_head_encoder=hp.n_head_encoder,
                        ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder,
                        concat_after_encoder=hp.concat_after_encoder, d_model_decoder=hp.d_model_decoder,
                        N_d=hp.n_layer_decoder, n_head_decoder=hp.n_head_decoder,
                        ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder,
                        concat_after_decoder=hp.concat_after_decoder, dropout_variance_adaptor=hp.dropout_variance_adaptor,
                        reduction_rate=hp.reduction_rate, dropout=hp.dropout, dropout_postnet=0.5, n_bins=hp.nbins,
                        f0_min=hp.f0_min, f0_max=hp.f0_max, energy_min=hp.energy_min, energy_max=hp.energy_max,
                        pitch_pred=hp.pitch_pred, energy_pred=hp.energy_pred, accent_emb=hp.accent_emb,
                        output_type=hp.output_type, num_group=hp.num_group, multi_speaker=hp.is_multi_speaker,
                        spk_emb_dim=hp.spk_emb_dim, spk_emb_architecture=hp.spk_emb_architecture)
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda().train()
    # item dicts shaped like TrainDatasets.__getitem__'s through the package's collate_fn (bit-exact against the
    # reference's own, tests/test_data.py)
    from oracle import make_golden_data as MG

    batch = data.collate_fn(MG.synth_items(n=3, seed=11))
    (text, mel, pos_text, pos_mel, text_lengths, mel_lengths, stop_token, spk_emb, f0, energy, alignment, accent, gender,
     spk_emb_postprocess, mel_name, hop_size) = batch
    assert isinstance(hop_size, list) and all(h is None for h in hop_size)
    text, mel, pos_text, pos_mel, f0, energy, alignment = (t.cuda() for t in (text, mel, pos_text, pos_mel, f0, energy,
                                                                               alignment))
    src_mask, trg_mask = data.create_masks(pos_text, pos_mel)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    opt.zero_grad()
    out = model(text, src_mask, trg_mask, alignment, f0, energy, accent, spkr_emb=spk_emb, fix_mask=hp.fix_mask,
                temperature=None, hop_size=hop_size)  # train_fastspeech2.py:167, verbatim
    (outputs_prenet, outputs_postnet, log_d_prediction, p_prediction, e_prediction, variance_adaptor_output,
     text_dur_predicted, attn_enc, attn_dec_dec, _, _, _, sq_vae_loss, sq_vae_perplexity) = out
    assert variance_adaptor_output.dtype == torch.float32 and text_dur_predicted.dtype == torch.float32
    loss = nn.L1Loss()(outputs_prenet, mel) + nn.L1Loss()(outputs_postnet, mel)
    loss = loss + nn.L1Loss()(p_prediction, f0) + nn.L1Loss()(e_prediction, energy)
    loss = loss + nn.L1Loss()(log_d_prediction, torch.log(alignment.float() + 1))
    loss.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
    opt.step()
    assert torch.isfinite(loss).item()
    model.eval()
    with torch.no_grad():
        ev = model(text[:1], src_mask[:1], hop_size=[None], accent=accent, spkr_emb=spk_emb)  # test_fastspeech2.py:169-171
    assert ev[5].dtype == torch.float32 and torch.isfinite(ev[1]).all()


def test_cuda_graphed_train_step_matches_eager():
    """Trainer(cuda_graph=True): zero_grad + forward + losses + backward replayed from one captured graph per input shape, the
    optimizer step eager.  Same loss sequence and parameters as the eager trainer (atomic accumulation order is the only
    difference), a second shape gets its own graph, and dropout switches the graph off (a replay would repeat the masks)."""
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    def batch(seed, B, L):
        text, pos, _ = synth.synth_text_batch(B, L, ragged=True, seed=seed)
        dur, pos_mel, f0, en = synth.synth_targets(pos, seed=seed, max_dur=10)
        mel = synth.synth_mel(pos_mel, 80, seed=seed)
        c = lambda t: t.cuda()  # noqa: E731
        return (c(text), c((pos != 0).unsqueeze(-2)), c((pos_mel != 0).unsqueeze(-2)), c(dur), c(f0), c(en), c(mel))

    b1, b2 = batch(5, 4, 32), batch(6, 3, 24)
    runs = {}
    for graphed in (False, True):
        hp, sd, model = build()
        tr = Trainer(model, cuda_graph=graphed)
        losses = []
        for i in range(6):
            o = tr.train_step(*(b1 if i != 3 else b2))
            losses.append(float(o["loss"]))
        assert (len(tr._graphs) == 2) == graphed
        runs[graphed] = (losses, tr.flat_p.clone(), model.postnet.pre_batchnorm.running_mean.clone())
    (le, pe, re_), (lg, pg, rg) = runs[False], runs[True]
    print("eager  ", le)
    print("graphed", lg)
    for a, b in zip(le, lg):
        assert abs(a - b) <= 1e-3 * max(1.0, abs(a))
    assert rel(pg, pe) < 1e-3 and rel(rg, re_) < 1e-3   # (bf16 roundings of activations amplify the atomic-order noise)
    hp = synth.make_hp(dropout=0.1)
    from transformer_tts_b200 import build_fastspeech2

    md = build_fastspeech2(hp)
    if any(m.p > 0 for m in md.modules() if isinstance(m, torch.nn.Dropout)):
        tr = Trainer(md.cuda().train(), cuda_graph=True)
        assert not tr._graph_ok()


def test_parameters_written_behind_the_trainers_back_are_seen():
    """The GEMM operands are views of the trainer's bf16 mirror of the flat parameter buffer, refreshed after every optimizer
    step.  A write that bypasses the trainer (model.load_state_dict after the Trainer was built) bumps the parameters'
    versions: the stale mirror must not be used."""
    from transformer_tts_b200 import synth
    from transformer_tts_b200.training import Trainer

    hp, sd, model = build()
    tr = Trainer(model)
    batch = train_batch_small()
    tr.zero_grad()
    l0 = float(tr.forward_backward(*batch)["loss"])
    sd2 = synth.synth_state_dict(hp, seed=3)
    model.load_state_dict(sd2)                       # in place, through the views of the flat buffer
    tr.zero_grad()
    l1 = float(tr.forward_backward(*batch)["loss"])
    _, _, fresh = build(seed_sd=3)
    tf = Trainer(fresh)
    tf.zero_grad()
    l2 = float(tf.forward_backward(*batch)["loss"])
    assert abs(l1 - l2) <= 1e-5 * max(1.0, abs(l2)) and abs(l1 - l0) > 1e-3, (l0, l1, l2)
    # ... and after an optimizer step the mirror is current again (operands are views of it: no per-tensor casts)
    tr.optimizer_step()
    from transformer_tts_b200.Models import _engine as E

    w = model.decoder.layers[0].attn.out.weight
    assert E._mirror(w) is not None and torch.equal(E._mirror(w), w.detach().to(torch.bfloat16))
    lc = E.layer_cache(model.decoder.layers[0])
    assert lc["wqkv"].data_ptr() == E._mirror(model.decoder.layers[0].attn.q_linear.weight).data_ptr()   # q|k|v: a view


def train_batch_small():
    from transformer_tts_b200 import synth

    text, pos, _ = synth.synth_text_batch(3, 24, ragged=True, seed=5)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=5, max_dur=8)
    mel = synth.synth_mel(pos_mel, 80, seed=5)
    c = lambda t: t.cuda()  # noqa: E731
    return (c(text), c((pos != 0).unsqueeze(-2)), c((pos_mel != 0).unsqueeze(-2)), c(dur), c(f0), c(en), c(mel))
