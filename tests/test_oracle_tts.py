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
