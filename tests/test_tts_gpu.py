"""GPU parity tests of the Transformer-TTS decode path: the decode-step kernels through the C-ABI
binding against fp64 torch, and `Transformer.forward` / `Transformer.generate` against the golden vectors of the unmodified
reference (tests/golden/tts_*.npz) and the oracle (oracle/transformer_tts_oracle.py).

Tolerances: the kernels use bf16 operands with fp32 accumulation and an fp32 residual stream, like the FastSpeech2 path:
kernel-level 2e-5 against an fp64 evaluation of the SAME bf16-rounded operands; model-level 1e-2 relative L2 (north_star's
bf16 bar) with the decoder inputs forced, and the free-running drift reported separately.
"""
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

GOLDEN = Path(__file__).resolve().parent / "golden"
DEV = "cuda"


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-300))


@pytest.fixture(scope="module")
def world():
    from transformer_tts_b200 import build_transformer_tts, synth

    hp = synth.make_tts_hp()
    sd = synth.synth_tts_state_dict(hp)
    model = build_transformer_tts(hp).eval()
    model.load_state_dict(sd)
    return hp, sd, model.cuda()


# ---------------------------------------------------------------------------------------------------------------------
# kernels
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("R,K,N", [(16, 384, 1152), (16, 80, 256), (5, 256, 384), (16, 1536, 384), (16, 384, 162), (33, 384, 1536)])
def test_small_linear_matches_fp64(R, K, N):
    from transformer_tts_b200 import lib

    g = torch.Generator().manual_seed(R * 1000 + K + N)
    x = torch.randn(R, K, generator=g).to(DEV)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(DEV)
    b = torch.randn(N, generator=g).to(DEV)
    res = torch.randn(R, N, generator=g).to(DEV)
    lng, lnb = (1 + 0.1 * torch.randn(K, generator=g)).to(DEV), (0.1 * torch.randn(K, generator=g)).to(DEV)
    wp = lib.pad_weight_rows(w)
    wd = w.to(torch.bfloat16).double()
    # plain: bf16 input rows
    out = torch.empty(R, N, device=DEV)
    lib.small_linear(wp, b, K, x_bf16=x.to(torch.bfloat16).contiguous(), out_f32=out)
    want = x.to(torch.bfloat16).double() @ wd.T + b.double()
    assert rel(out, want) < 2e-5
    if K > 1024:   # fp32 rows are staged in shared memory next to the operand tiles: the model's widest fp32 rows are 384
        with pytest.raises(RuntimeError, match="too large"):
            lib.small_linear(wp, b, K, x_f32=x, ln=(lng, lnb), out_f32=out)
        return
    # LayerNorm + ReLU + residual, both outputs, normalised rows exported
    out_bf = torch.empty(R, N, device=DEV, dtype=torch.bfloat16)
    xn = torch.empty(R, K, device=DEV, dtype=torch.bfloat16)
    lib.small_linear(wp, b, K, x_f32=x, ln=(lng, lnb), relu=True, res_f32=res, out_f32=out, out_bf16=out_bf, xn_bf16=xn)
    xnd = F.layer_norm(x.double(), (K,), lng.double(), lnb.double(), 1e-5)
    assert rel(xn, xnd) < 3e-3                                   # bf16 rounding of the normalised rows
    want = torch.relu(xn.double() @ wd.T + b.double()) + res.double()   # the kernel's own rounded rows as the operand
    assert rel(out, want) < 2e-5
    assert rel(out_bf, want) < 3e-3
    torch.cuda.synchronize()


def test_small_linear_position_addressing():
    from transformer_tts_b200 import lib

    g = torch.Generator().manual_seed(3)
    R, K, N, cap = 4, 80, 96, 7
    hist = torch.randn(R, cap, 162, generator=g).to(DEV)          # [mel | stop] rows of earlier positions
    w = torch.randn(N, K, generator=g).to(DEV) / 9
    pe = torch.randn(cap, N, generator=g).to(DEV)
    alpha = torch.tensor([0.7], device=DEV)
    wp = lib.pad_weight_rows(w)
    out = torch.zeros(R, cap, N, device=DEV)
    for p in range(cap):
        pos = torch.tensor([p], device=DEV, dtype=torch.int32)
        lib.small_linear(wp, None, K, x_f32=hist, ldx=cap * 162, R=R, x_pos_stride=162, x_pos_bias=-1, add_row=pe,
                         add_scale=alpha, out_f32=out, ldo=cap * N, out_pos_stride=N, pos=pos)
    torch.cuda.synchronize()
    xin = torch.cat([torch.zeros(R, 1, K, device=DEV), hist[:, :-1, :K]], 1)      # position p reads position p - 1; zeros at 0
    want = xin.to(torch.bfloat16).double() @ w.to(torch.bfloat16).double().T + 0.7 * pe.double()
    assert rel(out, want) < 2e-5


def test_attn_decode_self_and_cross_match_fp64():
    from transformer_tts_b200 import lib

    g = torch.Generator().manual_seed(11)
    R, H, Dh, cap = 5, 4, 96, 300
    D = H * Dh
    cache = torch.randn(2, R, H, cap, Dh, generator=g).to(torch.bfloat16).to(DEV)
    ref_cache = cache.clone()
    ctx = torch.empty(R, D, device=DEV, dtype=torch.bfloat16)
    sb, sh, sj = H * cap * Dh, cap * Dh, Dh
    for p, with_len in ((0, False), (1, False), (130, False), (299, False), (200, True)):
        qkv = torch.randn(R, 3 * D, generator=g).to(torch.bfloat16).to(DEV)
        pos = torch.tensor([p], device=DEV, dtype=torch.int32)
        lens = torch.tensor([1, 50, 201, 300, 77], device=DEV, dtype=torch.int32) if with_len else None
        lib.attn_decode(qkv, 0, 3 * D, cache, R * sb, (sb, sh, sj), cap, H, Dh, ctx, kv_new=qkv, knew_off=D, vnew_off=2 * D,
                        pos=pos, kv_len=lens)
        torch.cuda.synchronize()
        ref_cache[0, :, :, p] = qkv[:, D:2 * D].view(R, H, Dh)
        ref_cache[1, :, :, p] = qkv[:, 2 * D:].view(R, H, Dh)
        assert torch.equal(cache, ref_cache)                      # the append wrote slot p and nothing else
        q = qkv[:, :D].view(R, H, 1, Dh).double()
        for r in range(R):
            n = p + 1 if lens is None else min(p + 1, int(lens[r]))
            k, v = ref_cache[0, r, :, :n].double(), ref_cache[1, r, :, :n].double()
            pr = torch.softmax(q[r] @ k.transpose(-1, -2) / Dh ** 0.5, -1)
            want = (pr @ v).reshape(D)
            assert rel(ctx[r], want) < 4e-3, (p, r)                # bf16 output rounding
    # cross-attention over the (R, L, 2D) layout of a k|v projection
    L = 37
    kv = torch.randn(R, L, 2 * D, generator=g).to(torch.bfloat16).to(DEV)
    q = torch.randn(R, D, generator=g).to(torch.bfloat16).to(DEV)
    lens = torch.tensor([37, 1, 20, 36, 5], device=DEV, dtype=torch.int32)
    lib.attn_decode(q, 0, D, kv, D, (L * 2 * D, Dh, 2 * D), L, H, Dh, ctx, kv_len=lens)
    torch.cuda.synchronize()
    for r in range(R):
        n = int(lens[r])
        k = kv[r, :n, :D].view(n, H, Dh).transpose(0, 1).double()
        v = kv[r, :n, D:].view(n, H, Dh).transpose(0, 1).double()
        pr = torch.softmax(q[r].view(H, 1, Dh).double() @ k.transpose(-1, -2) / Dh ** 0.5, -1)
        assert rel(ctx[r], (pr @ v).reshape(D)) < 4e-3, r


@pytest.mark.parametrize("R", [16, 5, 20])
def test_cluster_kernels_match_the_unfused_step_kernels(R):
    """decode_attn_out_q / decode_ffn_block (one 16-CTA cluster per 16 rows, slices exchanged through distributed shared
    memory) against the chain of small_linear / ffn_tail_rows launches they replace: same operands, same arithmetic up to
    the order of the k-slice partial sums."""
    from transformer_tts_b200 import lib

    g = torch.Generator().manual_seed(100 + R)
    D, Hd = 384, 1536
    rnd = lambda *s, sc=1.0: (torch.randn(*s, generator=g) * sc).to(DEV)   # noqa: E731
    ctx = rnd(R, D).to(torch.bfloat16)
    x = rnd(R, D)
    wo, wq, w1, w2 = (lib.pad_weight_rows(rnd(D, D, sc=D ** -0.5)), lib.pad_weight_rows(rnd(D, D, sc=D ** -0.5)),
                      lib.pad_weight_rows(rnd(Hd, D, sc=D ** -0.5)), lib.pad_weight_rows(rnd(D, Hd, sc=Hd ** -0.5)))
    bo, bq, b1, b2 = rnd(D, sc=0.1), rnd(D, sc=0.1), rnd(Hd, sc=0.1), rnd(D, sc=0.1)
    ln = lambda: (1 + 0.1 * rnd(D), 0.1 * rnd(D))   # noqa: E731
    ln2, ln3, lnff, lnn = ln(), ln(), ln(), ln()
    f32 = lambda: torch.empty(R, D, device=DEV)   # noqa: E731
    bf = lambda n=D: torch.empty(R, n, device=DEV, dtype=torch.bfloat16)   # noqa: E731
    # attention output projection + residual, LayerNorm, query projection
    x2_ref, q_ref = f32(), bf()
    lib.small_linear(wo, bo, D, x_bf16=ctx, res_f32=x, out_f32=x2_ref)
    lib.small_linear(wq, bq, D, x_f32=x2_ref, ln=ln2, out_bf16=q_ref)
    x2, q = f32(), bf()
    lib.decode_attn_out_q(ctx, x, wo, bo, ln2, wq, bq, 1e-5, x2, q)
    torch.cuda.synchronize()
    assert rel(x2, x2_ref) < 1e-6 and rel(q, q_ref) < 2e-3          # q: bf16 roundings of values that differ in the last fp32 bits
    # second attention's output projection, FeedForward, next LayerNorm
    x3_ref, hid, xn, u, z_ref, zn_ref = f32(), bf(Hd), bf(), f32(), f32(), bf()
    lib.small_linear(wo, bo, D, x_bf16=ctx, res_f32=x, out_f32=x3_ref)
    lib.small_linear(w1, b1, D, x_f32=x3_ref, ln=ln3, relu=True, out_bf16=hid, xn_bf16=xn)
    lib.small_linear(w2, b2, Hd, x_bf16=hid, out_f32=u)
    lib.ffn_tail_rows(u, xn, x3_ref, lnff, lnn, 1e-5, z_ref, zn_ref)
    z, zn = f32(), bf()
    lib.decode_ffn_block(ctx, x, wo, bo, ln3, w1, b1, w2, b2, lnff, lnn, 1e-5, z, zn)
    torch.cuda.synchronize()
    assert rel(z, z_ref) < 2e-3 and rel(zn, zn_ref) < 4e-3          # a hidden value on a bf16 rounding boundary may flip
    # and against fp64 on the same bf16 operands
    xd = x.double() + ctx.double() @ wo.double().T + bo.double()
    xnd = F.layer_norm(xd, (D,), ln3[0].double(), ln3[1].double(), 1e-5).to(torch.bfloat16).double()
    hd = torch.relu(xnd @ w1.double().T + b1.double()).to(torch.bfloat16).double()
    zd = xd + F.layer_norm(hd @ w2.double().T + b2.double() + xnd, (D,), lnff[0].double(), lnff[1].double(), 1e-5)
    assert rel(z, zd) < 2e-3


# ---------------------------------------------------------------------------------------------------------------------
# model
# ---------------------------------------------------------------------------------------------------------------------
def test_teacher_forced_forward_matches_reference_golden(world):
    from oracle import transformer_tts_oracle as T

    hp, sd, model = world
    g = np.load(GOLDEN / "tts_teacher_forced.npz")
    text, pos = torch.from_numpy(g["text"]), torch.from_numpy(g["pos"])
    trg, pos_mel = torch.from_numpy(g["trg"]), torch.from_numpy(g["pos_mel"])
    src_mask, trg_mask = T.create_masks(pos, pos_mel)
    with torch.no_grad():
        pre, post, stop, a0, a1, a2 = model(text.cuda(), trg.cuda(), src_mask.cuda(), trg_mask.cuda(), None, training=False)
    assert a0 is None and a1 is None and a2 is None
    assert tuple(pre.shape) == g["outputs_prenet"].shape and tuple(stop.shape) == g["stop_token"].shape
    assert torch.equal(pre, post)                                 # PostConvNet(prev_version=False) returns its input
    e1, e2 = rel(pre, torch.from_numpy(g["outputs_prenet"])), rel(stop, torch.from_numpy(g["stop_token"]))
    print(f"teacher-forced vs reference golden: mel rel-L2 {e1:.2e}, stop logits {e2:.2e} (padded target rows included)")
    assert e1 < 1e-2 and e2 < 1e-2


def test_generate_matches_reference_loop_golden(world):
    hp, sd, model = world
    g = np.load(GOLDEN / "tts_ar_loop.npz")
    text, pos = torch.from_numpy(g["text"]).cuda(), torch.from_numpy(g["pos"]).cuda()
    n = g["stop_token"].shape[1]
    out = model.generate(text, (pos != 0).unsqueeze(-2), max_steps=n)
    assert out["mel"].shape == g["mel"].shape
    e1, e2 = rel(out["mel"], torch.from_numpy(g["mel"])), rel(out["stop_logits"], torch.from_numpy(g["stop_token"]).view(1, n, -1))
    print(f"free-running {n} positions vs the reference's loop: mel rel-L2 {e1:.2e}, stop logits {e2:.2e}")
    assert e1 < 1e-2 and e2 < 1e-2


def test_generate_ragged_batch_vs_oracle_and_graph_vs_eager(world):
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200 import synth

    hp, sd, model = world
    text, pos, _ = synth.synth_text_batch(3, 23, ragged=True, seed=31)
    n = 40
    src_mask = (pos != 0).unsqueeze(-2).cuda()
    out = model.generate(text.cuda(), src_mask, max_steps=n)
    model.use_cuda_graphs = False
    try:
        eager = model.generate(text.cuda(), src_mask, max_steps=n)
    finally:
        model.use_cuda_graphs = True
    assert torch.equal(out["mel"], eager["mel"]) and torch.equal(out["stop_logits"], eager["stop_logits"])
    with torch.no_grad():
        o = T.greedy_decode_cached(sd, vars(hp), text, pos, n)
        # decoder inputs forced to OUR generated frames: per-position parity without the feedback loop
        mel_in = torch.cat([torch.zeros(3, 1, hp.mel_dim), out["mel"].cpu()[:, ::hp.reduction_rate][:, :-1]], 1)
        f = T.greedy_decode_cached(sd, vars(hp), text, pos, n, forced_mel_in=mel_in)
    e_free, e_forced = rel(out["mel"], o["mel"]), rel(out["mel"], f["mel"])
    print(f"ragged batch, {n} positions: free-running mel rel-L2 {e_free:.2e}; inputs forced {e_forced:.2e}; "
          f"stop logits forced {rel(out['stop_logits'], f['stop_logits']):.2e}")
    assert e_forced < 1e-2 and e_free < 2e-2
    assert rel(out["stop_logits"], f["stop_logits"]) < 1e-2


def test_cfg5_1000_positions_batch16_vs_oracle(world):
    """BASELINE.json configs[4] at full size: batch 16, 128 phonemes, 1000 decoder positions (2000 frames).  The oracle is
    fed OUR decoder inputs (teacher-forced on the generated frames), so every position is compared without chaotic drift;
    stop positions must agree wherever the oracle's stop probability is not within 1e-3 of the threshold."""
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200 import synth

    hp, sd, model = world
    B, L, n = 16, 128, 1000
    text, pos, _ = synth.synth_text_batch(B, L, ragged=True, seed=77, min_len=64)
    out = model.generate(text.cuda(), (pos != 0).unsqueeze(-2).cuda(), max_steps=n)
    assert out["steps"] == n and out["mel"].shape == (B, n * hp.reduction_rate, hp.mel_dim)
    assert bool(torch.isfinite(out["mel"]).all())
    torch.set_num_threads(16)
    with torch.no_grad():
        mel_in = torch.cat([torch.zeros(B, 1, hp.mel_dim), out["mel"].cpu()[:, ::hp.reduction_rate][:, :-1]], 1)
        f = T.greedy_decode_cached(sd, vars(hp), text, pos, n, forced_mel_in=mel_in)
    e_mel, e_stop = rel(out["mel"], f["mel"]), rel(out["stop_logits"], f["stop_logits"])
    worst = max(rel(out["mel"][:, s:s + 100], f["mel"][:, s:s + 100]) for s in range(0, 2 * n, 100))
    print(f"cfg5 B=16 x 1000 positions: mel rel-L2 {e_mel:.2e} (worst 100-frame window {worst:.2e}), stop logits {e_stop:.2e}")
    assert e_mel < 1e-2 and worst < 2e-2 and e_stop < 1e-2
    p_or = torch.sigmoid(f["stop_logits"]).mean(-1)
    safe = ((p_or - 0.5).abs() > 1e-3).all(dim=1)
    assert torch.equal(out["stop_step"].cpu()[safe], f["stop_step"][safe])


def test_cluster_and_unfused_decode_steps_agree(world):
    from transformer_tts_b200 import synth

    hp, sd, model = world
    text, pos, _ = synth.synth_text_batch(16, 40, ragged=True, seed=55)
    src_mask = (pos != 0).unsqueeze(-2).cuda()
    assert model.decoder.use_cluster_kernels
    a = model.generate(text.cuda(), src_mask, max_steps=30)
    launches_cluster = model.last_state.launches_per_step
    model.decoder.use_cluster_kernels = False
    try:
        b = model.generate(text.cuda(), src_mask, max_steps=30)
        launches_unfused = model.last_state.launches_per_step
    finally:
        model.decoder.use_cluster_kernels = True
    e = rel(a["mel"], b["mel"])
    print(f"cluster kernels ({launches_cluster} launches per position) vs unfused ({launches_unfused}): mel rel-L2 {e:.2e}")
    assert launches_cluster < launches_unfused and e < 1e-2      # free-running: both are within 5e-3 of the oracle


def test_generate_early_exit_on_stop(world):
    hp, sd, model = world
    from transformer_tts_b200 import synth

    text, pos, _ = synth.synth_text_batch(2, 16, seed=41)
    src_mask = (pos != 0).unsqueeze(-2).cuda()
    full = model.generate(text.cuda(), src_mask, max_steps=64)
    # force a stop: a large positive stop bias makes every position stop, so the loop must end at the first check
    with torch.no_grad():
        old = model.stop_token.bias.clone()
        model.stop_token.bias.fill_(30.0)
    try:
        out = model.generate(text.cuda(), src_mask, max_steps=64, check_every=8)
    finally:
        with torch.no_grad():
            model.stop_token.bias.copy_(old)
    assert out["steps"] == 8 and out["stop_step"].tolist() == [1, 1]
    assert torch.equal(out["mel"], full["mel"][:, :8 * hp.reduction_rate])      # the mel head is untouched by the stop bias
    assert full["steps"] == 64


def test_other_shapes_batch_above_one_cluster_and_reduction_rate_1():
    """Batch 19 (two row tiles / clusters) and a model with reduction_rate = 1 (stop logits (B, T) as the reference squeezes
    them, transformer.py:111) and 2 decoder blocks: generate and the teacher-forced forward against the oracle."""
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200 import build_transformer_tts, synth

    hp = synth.make_tts_hp(reduction_rate=1, n_layer_decoder=2, n_layer_encoder=2)
    sd = synth.synth_tts_state_dict(hp, seed=3)
    model = build_transformer_tts(hp).eval()
    model.load_state_dict(sd)
    model = model.cuda()
    text, pos, _ = synth.synth_text_batch(19, 21, ragged=True, seed=9)
    n = 7
    out = model.generate(text.cuda(), (pos != 0).unsqueeze(-2).cuda(), max_steps=n)
    with torch.no_grad():
        o = T.greedy_decode_cached(sd, vars(hp), text, pos, n)
    assert out["mel"].shape == (19, n, hp.mel_dim) and out["stop_logits"].shape == (19, n, 1)
    assert rel(out["mel"], o["mel"]) < 1e-2 and rel(out["stop_logits"], o["stop_logits"]) < 1e-2
    trg = torch.randn(19, 5, hp.mel_dim, generator=torch.Generator().manual_seed(1))
    pos_mel = torch.arange(1, 6).unsqueeze(0).repeat(19, 1)
    pos_mel[3, 2:] = 0
    src_mask, trg_mask = T.create_masks(pos, pos_mel)
    with torch.no_grad():
        pre, post, stop, *_ = model(text.cuda(), trg.cuda(), src_mask.cuda(), trg_mask.cuda(), None, training=False)
        ref = T.transformer_forward(sd, vars(hp), text, trg, src_mask, trg_mask)
    assert tuple(stop.shape) == (19, 5) == tuple(ref["stop_token"].shape)       # squeezed, as the reference returns it
    assert rel(pre, ref["outputs_prenet"]) < 1e-2 and rel(stop, ref["stop_token"]) < 1e-2


def test_batched_prefix_forward_matches_step_machine_and_oracle(world):
    """Transformer.forward on a 300-position ragged prefix: the tensor-core path (causal tcgen05 attention, encoder-decoder
    attention on the packed sequence, fused FeedForward) against the decode-step machine and against the oracle."""
    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200 import synth

    hp, sd, model = world
    B, L, Tn = 3, 50, 300
    text, pos, _ = synth.synth_text_batch(B, L, ragged=True, seed=61)
    trg = torch.randn(B, Tn, hp.mel_dim, generator=torch.Generator().manual_seed(62))
    pos_mel = torch.arange(1, Tn + 1).unsqueeze(0).repeat(B, 1)
    pos_mel[1, 170:] = 0
    pos_mel[2, 1:] = 0
    src_mask, trg_mask = T.create_masks(pos, pos_mel)
    args = (text.cuda(), trg.cuda(), src_mask.cuda(), trg_mask.cuda(), None)
    assert model.prefix_mode == "batched"
    with torch.no_grad():
        pre_b, _, stop_b, *_ = model(*args, training=False)
        model.prefix_mode = "steps"
        try:
            pre_s, _, stop_s, *_ = model(*args, training=False)
        finally:
            model.prefix_mode = "batched"
        ref = T.transformer_forward(sd, vars(hp), text, trg, src_mask, trg_mask)
    e_bs, e_bo, e_so = rel(pre_b, pre_s), rel(pre_b, ref["outputs_prenet"]), rel(pre_s, ref["outputs_prenet"])
    print(f"prefix of {Tn} positions: batched vs steps {e_bs:.2e}; batched vs oracle {e_bo:.2e}; steps vs oracle {e_so:.2e}; "
          f"stop logits batched vs oracle {rel(stop_b, ref['stop_token']):.2e}")
    assert e_bo < 1e-2 and e_so < 1e-2 and e_bs < 1e-2
    assert rel(stop_b, ref["stop_token"]) < 1e-2
