"""CPU tests: the oracle (fp32 restatement) against the golden vectors produced by the unmodified
reference (oracle/make_golden.py), against the live reference when /root/reference is present, plus
the host-side pieces of the package (state_dict layout, positional table, hparams file, C-ABI symbols).
"""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import fastspeech2_oracle as O
from oracle import ref_loader
from transformer_tts_b200 import synth

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"
TOL = 2e-5  # same ops, same library, fp32: differences are summation-order noise only


@pytest.fixture(scope="module")
def world():
    torch.set_num_threads(8)
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    return hp, sd, synth.state_dict_digest(sd)


def close(a, b, tol=TOL):
    a = torch.as_tensor(a).double()
    b = torch.as_tensor(b).double()
    assert a.shape == b.shape, (a.shape, b.shape)
    err = (a - b).abs().max().item()
    assert err <= tol * max(1.0, b.abs().max().item()), err


def test_state_dict_layout(world):
    hp, sd, _ = world
    assert len(sd) == 289
    assert sd["encoder.embed.weight"].shape == (152, 384)
    assert sd["encoder.layers.0.ff.f_1.weight"].shape == (1536, 384, 5)
    assert sd["decoder.layers.5.ff.f_2.weight"].shape == (384, 1536, 1)
    assert sd["postnet.conv2.weight"].shape == (80, 384, 5)
    assert sd["postnet.pre_batchnorm.num_batches_tracked"].dtype == torch.int64
    assert sum(v.numel() for k, v in sd.items() if v.is_floating_point() and "running" not in k) > 54_000_000


def test_golden_cfg1_free_running(world):
    hp, sd, digest = world
    g = np.load(GOLDEN / "fs2_cfg1_free_running.npz")
    assert str(g["digest"]) == digest, "synthetic weights differ from the ones the golden run used"
    text, pos = torch.from_numpy(g["text"]), torch.from_numpy(g["pos"])
    with torch.no_grad():
        o = O.fastspeech2_forward(sd, vars(hp), text, (pos != 0).unsqueeze(-2), return_attn=True)
    close(o["outputs_prenet"], g["mel_pre"])
    close(o["outputs_postnet"], g["mel_post"])
    close(o["log_d_prediction"], g["log_d"])
    close(o["p_prediction"], g["p"])
    close(o["e_prediction"], g["e"])
    close(o["variance_adaptor_output"][:, ::8, ::16], g["va_out_sub"])
    close(o["attn_enc"][0, 0, 0], g["attn_enc_l0h0"])
    close(o["attn_dec"][0, 5, 3, ::16, ::16], g["attn_dec_l5h3_sub"])
    # integer decisions: exact
    d_ref = torch.clamp(torch.round(torch.exp(torch.from_numpy(g["log_d"])) - 1), min=0)
    assert torch.equal(O.round_durations(o["log_d_prediction"]), d_ref)
    assert o["outputs_prenet"].shape[1] == int(d_ref.sum())


def test_golden_ragged_teacher_forced(world):
    hp, sd, digest = world
    g = np.load(GOLDEN / "fs2_ragged_teacher_forced.npz")
    assert str(g["digest"]) == digest
    t = lambda k: torch.from_numpy(g[k])  # noqa: E731
    src_mask = (t("pos") != 0).unsqueeze(-2)
    mel_mask = (t("pos_mel") != 0).unsqueeze(-2)
    with torch.no_grad():
        o = O.fastspeech2_forward(sd, vars(hp), t("text"), src_mask, mel_mask=mel_mask, d_target=t("dur"),
                                  p_target=t("f0"), e_target=t("energy"))
        o2 = O.fastspeech2_forward(sd, vars(hp), t("text"), src_mask, mel_mask=mel_mask, d_target=t("dur"))
    close(o["outputs_prenet"], g["mel_pre"])
    close(o["outputs_postnet"], g["mel_post"])
    close(o["log_d_prediction"], g["log_d"])
    close(o["p_prediction"], g["p"])
    close(o["e_prediction"], g["e"])
    close(o["variance_adaptor_output"][:, ::4, ::16], g["va_out_sub"])
    close(o["text_dur_predicted"][:, ::4, ::16], g["text_dur_sub"])
    close(o2["outputs_postnet"], g["mel_post_pred_buckets"])


def test_golden_ops(world):
    hp, sd, digest = world
    g = np.load(GOLDEN / "fs2_ops.npz")
    assert str(g["digest"]) == digest
    x, m = torch.from_numpy(g["x"]), torch.from_numpy(g["mask"])
    with torch.no_grad():
        y, attn = O.encoder_layer(sd, "encoder.layers.2.", x, m, 4)
        close(y, g["enc_layer2"])
        close(attn, g["enc_layer2_attn"])
        close(O.feed_forward(sd, "decoder.layers.1.ff.", x), g["dec_ff1"])
        close(O.variance_predictor(sd, "variance_adaptor.pitch_predictor.", x, m), g["pitch_pred"])
        pre, post = O.postnet(sd, "postnet.", x)
        close(pre, g["post_pre"])
        close(post, g["post_post"])
        assert torch.equal(O.positional_encoder(x, sd["encoder.pe.alpha"]), torch.from_numpy(g["pe"]))
        d = torch.from_numpy(g["lr_dur"])
        lx, lpos, lidx = O.length_regulator(x, d)
        assert torch.equal(lx, torch.from_numpy(g["lr_x"]))
        assert torch.equal(lpos, torch.from_numpy(g["lr_pos"]))
        assert torch.equal(O.get_mask_from_lengths(lpos), torch.from_numpy(g["lr_mask"]))
        assert torch.equal(O.length_regulator(x, d, 30)[0], torch.from_numpy(g["lr_x_max30"]))
        # frame_idx is consistent with the expansion
        for b in range(2):
            n = int((lidx[b] >= 0).sum())
            assert torch.equal(lx[b, :n], x[b, lidx[b, :n]])
    pb, eb = O.variance_bins(hp.nbins, hp.f0_min, hp.f0_max, hp.energy_min, hp.energy_max)
    assert torch.equal(pb, torch.from_numpy(g["pitch_bins"])) and torch.equal(eb, torch.from_numpy(g["energy_bins"]))
    assert torch.equal(O.positional_table(384)[::97], torch.from_numpy(g["pe_table_sub"]))


def test_package_positional_table_matches_reference_loop():
    from transformer_tts_b200.Models import _engine as E

    assert torch.equal(E.positional_table_cpu(384), O.positional_table(384))


def test_package_state_dict_is_reference_layout(world):
    from transformer_tts_b200 import build_fastspeech2

    hp, sd, _ = world
    model = build_fastspeech2(hp)
    msd = model.state_dict()
    assert list(msd.keys()) == list(sd.keys())
    for k in sd:
        assert msd[k].shape == sd[k].shape and msd[k].dtype == sd[k].dtype, k
    model.load_state_dict(sd)
    # bins are host-computed exactly like the reference's
    pb, eb = O.variance_bins(hp.nbins, hp.f0_min, hp.f0_max, hp.energy_min, hp.energy_max)
    assert torch.equal(model.variance_adaptor.pitch_bins, pb) and torch.equal(model.variance_adaptor.energy_bins, eb)


def test_package_has_no_cpu_path(world):
    from transformer_tts_b200 import build_fastspeech2

    hp, sd, _ = world
    model = build_fastspeech2(hp).eval()
    with pytest.raises(RuntimeError, match="no CPU path"):
        model(torch.ones(1, 4, dtype=torch.long), torch.ones(1, 1, 4, dtype=torch.bool))
    with pytest.raises(NotImplementedError):
        build_fastspeech2(synth.make_hp(use_sq_vae=True))
    with pytest.raises(NotImplementedError):
        build_fastspeech2(synth.make_hp(encoder_type="conformer"))


def test_product_does_not_import_oracle():
    for p in (ROOT / "transformer_tts_b200").rglob("*.py"):
        src = p.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f"{p} imports the oracle"


def test_hparams_template_loads():
    ns = {}
    exec((ROOT / "config" / "hparams_template.py").read_text(), ns)
    hp = synth.make_hp()
    for k in ("vocab_size", "mel_dim", "d_model_encoder", "n_layer_encoder", "n_head_encoder",
              "ff_conv_kernel_size_encoder", "d_model_decoder", "n_layer_decoder", "n_head_decoder",
              "ff_conv_kernel_size_decoder", "nbins", "f0_min", "f0_max", "energy_min", "energy_max", "reduction_rate"):
        assert ns[k] == getattr(hp, k), k
    assert "_configured" not in ns  # utils/__init__.py:40-42 rejects that name


def test_c_abi_exports_every_declared_symbol():
    from transformer_tts_b200 import _build, lib

    path = _build.build()
    dll = ctypes.CDLL(str(path))
    header = (ROOT / "include" / "ttsb200.h").read_text()
    declared = set(re.findall(r"\b(ttsb200_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    assert declared == set(lib.EXPORTED_SYMBOLS), declared ^ set(lib.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(dll, name), name
    dll.ttsb200_version.restype = ctypes.c_int
    # the binding refuses a library built from other struct layouts / signatures (lib.load checks the same)
    assert dll.ttsb200_version() == lib.ABI_VERSION
    assert f"#define TTSB200_ABI_VERSION {lib.ABI_VERSION}" in header


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference checkout not present (GPU box)")
def test_oracle_matches_live_reference(world):
    hp, sd, _ = world
    model = ref_loader.build_reference_model(hp, sd)
    text, pos, _ = synth.synth_text_batch(1, 20, seed=11)
    m = (pos != 0).unsqueeze(-2)
    with torch.no_grad():
        out = model(text, m)
        o = O.fastspeech2_forward(sd, vars(hp), text, m)
    close(o["outputs_postnet"], out[1])
    close(o["log_d_prediction"], out[2])


def test_golden_train_step(world):
    """Oracle (autograd through the restatement, train() semantics) vs the unmodified reference's
    loss.backward() on the same seeded batch (oracle/make_golden_train.py)."""
    from oracle.make_golden_train import FULL, SUB, train_batch

    hp, sd, digest = world
    g = np.load(GOLDEN / "fs2_train_step.npz")
    assert str(g["digest"]) == digest
    text, pos, dur, pos_mel, f0, en, mel = train_batch()
    losses, grads, out = O.training_step_grads(sd, vars(hp), text, (pos != 0).unsqueeze(-2),
                                               (pos_mel != 0).unsqueeze(-2), dur, f0, en, mel)
    ref = g["losses"]
    for i, k in enumerate(["mel_loss", "post_mel_loss", "duration_loss", "f0_loss", "energy_loss", "loss"]):
        assert abs(float(losses[k]) - ref[i]) <= 1e-5 * max(1.0, abs(ref[i])), k
    close(out["outputs_postnet"].detach(), g["mel_post"], 1e-4)
    norms = dict(zip([str(n) for n in g["names"]], g["grad_norms"]))
    assert set(norms) == set(grads)
    for n, gr in grads.items():
        assert abs(float(gr.double().norm()) - norms[n]) <= 2e-3 * max(norms[n], 1e-6) + 1e-7, n
    for n in FULL:
        close(grads[n], g["full:" + n], 1e-3)
    for n in SUB:
        close(grads[n].reshape(grads[n].shape[0], -1)[::7, ::13], g["sub:" + n], 1e-3)


def test_radam_oracle_matches_reference_golden():
    """oracle.radam_step (restatement of radam.py:56-92) against the parameters the UNMODIFIED reference optimizer
    produced (oracle/make_golden_radam.py), and against the live reference when it is present."""
    gold = np.load(GOLDEN / "radam_steps.npz")
    p, grads, want = (torch.from_numpy(gold[k]) for k in ("p0", "grads", "p_steps"))
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    live = None
    if ref_loader.reference_available():
        import sys
        import warnings

        if ref_loader.REFERENCE_ROOT not in sys.path:
            sys.path.insert(0, ref_loader.REFERENCE_ROOT)
        import radam

        lp = torch.nn.Parameter(p.clone())
        live = (lp, radam.RAdam([lp], lr=1e-3, betas=(0.9, 0.98), eps=1e-9), warnings)
    for t in range(grads.shape[0]):
        p, m, v = O.radam_step(p, grads[t], m, v, t + 1)
        assert (p - want[t]).abs().max().item() < 1e-6, t
        if live is not None:
            lp, opt, warnings = live
            lp.grad = grads[t].clone()
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                opt.step()
            assert (p - lp.detach()).abs().max().item() < 1e-6, t
    close(m, gold["exp_avg"], 1e-6)
    close(v, gold["exp_avg_sq"], 1e-6)
