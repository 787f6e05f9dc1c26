"""CPU tests of the Transformer-TTS oracle (oracle/transformer_tts_oracle.py) against the golden vectors produced by the
unmodified reference (oracle/make_golden_tts.py), of its incremental restatement against the re-run-the-prefix loop, and of
the synthetic state_dict layout against the reference module itself (build container only)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import ref_loader
from oracle import transformer_tts_oracle as T
from transformer_tts_b200 import synth

GOLDEN = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def world():
    torch.set_num_threads(8)
    hp = synth.make_tts_hp()
    sd = synth.synth_tts_state_dict(hp)
    return hp, sd, synth.state_dict_digest(sd)


def _close(a, b, tol):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    err = float((a - b).abs().max() / (b.abs().max() + 1e-30))
    assert err < tol, err
    return err


def test_teacher_forced_forward_matches_reference_golden(world):
    hp, sd, digest = world
    g = np.load(GOLDEN / "tts_teacher_forced.npz")
    assert str(g["digest"]) == digest
    text, pos = torch.from_numpy(g["text"]), torch.from_numpy(g["pos"])
    trg, pos_mel = torch.from_numpy(g["trg"]), torch.from_numpy(g["pos_mel"])
    src_mask, trg_mask = T.create_masks(pos, pos_mel)
    with torch.no_grad():
        out = T.transformer_forward(sd, vars(hp), text, trg, src_mask, trg_mask, return_attn=True)
    _close(out["outputs_prenet"], g["outputs_prenet"], 2e-5)
    _close(out["outputs_postnet"], g["outputs_prenet"], 2e-5)
    _close(out["stop_token"], g["stop_token"], 2e-5)
    _close(out["attn_dec_dec"][:, 0], g["attn_dec_dec_l0"], 2e-5)
    _close(out["attn_dec_enc"][:, 5], g["attn_dec_enc_l5"], 2e-5)
    # causal: no probability above the diagonal
    assert float(out["attn_dec_dec"].triu(1).abs().max()) == 0.0


def test_autoregressive_loop_matches_reference_golden(world):
    hp, sd, digest = world
    g = np.load(GOLDEN / "tts_ar_loop.npz")
    assert str(g["digest"]) == digest
    text, pos = torch.from_numpy(g["text"]), torch.from_numpy(g["pos"])
    n = g["stop_token"].shape[1]
    with torch.no_grad():
        a = T.greedy_decode(sd, vars(hp), text, pos, n)
        b = T.greedy_decode_cached(sd, vars(hp), text, pos, n)
    _close(a["mel"], g["mel"], 5e-5)
    _close(a["stop_logits"], g["stop_token"], 5e-5)
    # the incremental restatement (what the CUDA decode is compared with at long lengths) equals the re-run loop
    _close(b["mel"], a["mel"], 5e-5)
    _close(b["stop_logits"], a["stop_logits"], 5e-5)
    assert torch.equal(a["stop_step"], b["stop_step"])


def test_cached_decode_equals_rerun_on_a_ragged_batch(world):
    hp, sd, _ = world
    text, pos, _ = synth.synth_text_batch(3, 17, ragged=True, seed=21)
    with torch.no_grad():
        a = T.greedy_decode(sd, vars(hp), text, pos, 9)
        b = T.greedy_decode_cached(sd, vars(hp), text, pos, 9)
    _close(b["mel"], a["mel"], 5e-5)
    assert a["mel"].shape == (3, 9 * hp.reduction_rate, hp.mel_dim)
    assert a["stop_logits"].shape == (3, 9, hp.reduction_rate)


def test_stop_steps_rule():
    logits = torch.full((2, 6, 2), -5.0)
    logits[0, 0] = 9.0        # position 0 is never tested (the loop starts at i = 2)
    logits[0, 3, 0] = 9.0     # mean(sigmoid) = 0.5 + tiny  -> stops
    logits[0, 3, 1] = 0.1
    assert T.stop_steps(logits).tolist() == [3, 6]


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference checkout not present (GPU box)")
def test_state_dict_layout_and_forward_match_live_reference(world):
    from oracle.make_golden_tts import build_reference_tts, reference_create_masks

    hp, sd, _ = world
    model = build_reference_tts(hp, None)
    ref_sd = model.state_dict()
    assert list(ref_sd.keys()) == list(sd.keys())
    for k in sd:
        assert tuple(ref_sd[k].shape) == tuple(sd[k].shape), k
    model.load_state_dict(sd)
    text, pos, _ = synth.synth_text_batch(2, 9, ragged=True, seed=3)
    trg = torch.randn(2, 5, hp.mel_dim, generator=torch.Generator().manual_seed(4))
    pos_mel = torch.arange(1, 6).unsqueeze(0).repeat(2, 1)
    sm, tm = reference_create_masks(pos, pos_mel)
    sm2, tm2 = T.create_masks(pos, pos_mel)
    assert torch.equal(sm, sm2) and torch.equal(tm, tm2)
    with torch.no_grad():
        pre, post, stop, *_ = model(text, trg, sm, tm, None, training=False)
        out = T.transformer_forward(sd, vars(hp), text, trg, sm, tm)
    _close(out["outputs_prenet"], pre, 2e-5)
    _close(out["stop_token"], stop, 2e-5)


def test_oracle_autograd_matches_reference_training_golden():
    """tests/golden/tts_train_step.npz (the unmodified reference's loss.backward() on the statements of train.py:186-252,
    made by oracle/make_golden_tts_train.py): torch autograd through the oracle reproduces the losses and the gradients —
    this is what makes the oracle the yardstick of the GPU training-step tests (tests/test_tts_train_gpu.py)."""
    import torch.nn.functional as F

    from oracle import transformer_tts_oracle as T
    from oracle.make_golden_tts_train import FULL, POS_WEIGHT, SUB, tts_train_batch
    from transformer_tts_b200 import synth

    g = np.load(GOLDEN / "tts_train_step.npz")
    hp = synth.make_tts_hp()
    r = hp.reduction_rate
    sd = synth.synth_tts_state_dict(hp)
    assert str(g["digest"]) == synth.state_dict_digest(sd)
    b = tts_train_batch(hp)
    leaf = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and not k.startswith("postnet.") else v) for k, v in sd.items()}
    src_mask, trg_mask = T.create_masks(b["pos"], b["pos_in"])
    torch.set_num_threads(8)
    o = T.transformer_forward(leaf, vars(hp), b["text"], b["mel_input"], src_mask, trg_mask)
    bb, t, c = o["outputs_prenet"].shape
    pre = o["outputs_prenet"].reshape(bb, t * r, c // r)
    mel_loss = F.l1_loss(pre, b["mel"][:, r:, :])
    tok = F.binary_cross_entropy_with_logits(o["stop_token"].reshape(bb, t * r), b["stop"][:, r:], reduction="mean",
                                             pos_weight=torch.tensor(POS_WEIGHT))
    loss = 2 * mel_loss + tok
    loss.backward()
    np.testing.assert_allclose([mel_loss.item(), mel_loss.item(), tok.item(), loss.item()], g["losses"], rtol=1e-5)
    np.testing.assert_allclose(pre.detach().numpy(), g["outputs_prenet"], rtol=0, atol=2e-4)
    norms = dict(zip([str(n) for n in g["names"]], g["grad_norms"]))
    gl = float(g["global_grad_norm"])
    for n, v in leaf.items():
        if isinstance(v, torch.Tensor) and v.requires_grad and n in norms:
            assert abs(float(v.grad.double().norm()) - norms[n]) <= 1e-4 * max(norms[n], 1e-3 * gl), n
    for n in FULL:
        np.testing.assert_allclose(leaf[n].grad.numpy(), g["full:" + n], rtol=0, atol=1e-4 * gl)
    for n in SUB:
        gr = leaf[n].grad
        np.testing.assert_allclose(gr.reshape(gr.shape[0], -1)[::7, ::13].numpy(), g["sub:" + n], rtol=0, atol=1e-5 * gl)
    assert all(float(g["grad_norms"][i]) == 0.0 for i, n in enumerate(g["names"]) if str(n).startswith("postnet."))
