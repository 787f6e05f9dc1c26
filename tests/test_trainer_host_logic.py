"""Host-side logic of the trainers that needs no GPU: the flat-buffer layout (q | k | v placed next to each other, 8-element
alignment), the view-instead-of-cat helpers of Models/_engine and the dropout call-site plan of the Transformer-TTS step."""
import torch


def test_flat_order_groups_qkv_and_keeps_everything_once():
    from transformer_tts_b200 import build_fastspeech2, build_transformer_tts, synth
    from transformer_tts_b200.training import Trainer

    for model in (build_fastspeech2(synth.make_hp()), build_transformer_tts(synth.make_tts_hp())):
        names = [n for n, _ in model.named_parameters()]
        order = Trainer._flat_order(names)
        assert sorted(order) == sorted(names) and len(order) == len(set(order))
        pos = {n: i for i, n in enumerate(order)}
        groups = [n[: -len("q_linear.weight")] for n in names if n.endswith("q_linear.weight")]
        assert groups
        for pre in groups:
            want = [pre + f"{x}_linear.{y}" for y in ("weight", "bias") for x in ("q", "k", "v")]
            idx = [pos[w] for w in want]
            assert idx == list(range(idx[0], idx[0] + 6)), pre       # q.w k.w v.w q.b k.b v.b, consecutive
        # parameters outside attention modules keep their relative order
        rest = [n for n in names if "_linear." not in n or not any(t in n for t in ("q_linear", "k_linear", "v_linear"))]
        assert [n for n in order if n in set(rest)] == rest
    # a module that lacks one of the six tensors is left alone
    odd = ["a.q_linear.weight", "a.q_linear.bias", "a.k_linear.weight", "b.weight"]
    assert Trainer._flat_order(odd) == odd


def test_cat_rows_is_a_view_when_the_rows_are_adjacent():
    from transformer_tts_b200.Models import _engine as E

    flat = torch.arange(3 * 4 * 5 + 8, dtype=torch.float32)
    parts = [flat[i * 20:(i + 1) * 20].view(4, 5) for i in range(3)]
    assert E._adjacent(parts)
    v = E._cat_rows(parts)
    assert v.shape == (12, 5) and v.data_ptr() == flat.data_ptr() and torch.equal(v, torch.cat(parts, 0))
    v[5, 2] = -1.0                                   # a view: writes land in the flat buffer
    assert flat[5 * 5 + 2] == -1.0
    gap = [flat[0:20].view(4, 5), flat[28:48].view(4, 5)]
    assert not E._adjacent(gap)
    c = E._cat_rows(gap)
    assert c.data_ptr() != flat.data_ptr() and torch.equal(c, torch.cat(gap, 0))
    b = E._cat_rows([flat[0:4], flat[4:8], flat[8:12]])          # 1-D (biases)
    assert b.shape == (12,) and b.data_ptr() == flat.data_ptr()
    assert E._cat_rows([flat[0:4]]).shape == (4,)


def test_mirror_is_ignored_once_the_parameter_changed():
    from transformer_tts_b200.Models import _engine as E

    p = torch.nn.Parameter(torch.randn(6, 4))
    assert E._mirror(p) is None
    m = p.detach().to(torch.bfloat16)
    p._ttsb_bf16 = (m, p._version)
    assert E._mirror(p) is m and E._bf16(p) is m
    with torch.no_grad():
        p.mul_(2.0)                                  # an in-place write bumps the version: the mirror is stale
    assert E._mirror(p) is None
    assert torch.equal(E._bf16(p), p.detach().to(torch.bfloat16))
    w = torch.nn.Parameter(torch.randn(8, 4, 1))     # Conv1d weight with k = 1: the operand is the mirror's memory
    mw = w.detach().to(torch.bfloat16)
    w._ttsb_bf16 = (mw, w._version)
    assert E.conv_w_to_taps(w).data_ptr() == mw.data_ptr() and E.conv_w_to_taps(w).shape == (1, 8, 4)


def test_tts_dropout_plan_sites_and_probabilities():
    from transformer_tts_b200 import synth
    from transformer_tts_b200.Models.transformer import Transformer
    from transformer_tts_b200.tts_training import TTSDropoutPlan

    hp = synth.make_tts_hp()
    args = (hp, hp.vocab_size, hp.mel_dim, 384, 2, 4, 5, False, 384, 2, 4, 1, False, 2)
    plan = TTSDropoutPlan(Transformer(*args, dropout=0.1, dropout_prenet=0.5), base_seed=3)
    assert plan.active and plan.p_stack == {"enc": 0.1, "dec": 0.1} and plan.p_prenet == 0.5
    assert plan.prob("prenet.2") == 0.5 and plan.prob("dec.1.attn2") == 0.1 and plan.prob("enc.0.ff") == 0.1
    assert plan.site("dec.0.attn1") == plan.site("dec.0.attn1") != plan.site("dec.0.attn2")
    s1, s2 = plan.next_seed(), plan.next_seed()
    assert s1 != s2 and TTSDropoutPlan(Transformer(*args, dropout=0.1, dropout_prenet=0.5), base_seed=3).next_seed() == s1
    assert not TTSDropoutPlan(Transformer(*args, dropout=0.0, dropout_prenet=0.0)).active
