"""Kernel-level parity on the GPU: every entry point of include/ttsb200.h against a plain torch
restatement of the reference op it replaces (float64 math on the same bf16-rounded operands).

Tolerances: fp32 outputs 2e-4 relative to the output scale (fp32 accumulation-order noise),
bf16 outputs 1e-2 (one bf16 rounding = 2^-9 relative), integer outputs exact.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    from transformer_tts_b200 import lib

    lib.load()
    return lib


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a = a.double()
    b = b.double()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def max_err(a, b):
    return (a.double() - b.double()).abs().max().item()


def conv_ref(x, w, pad_left):
    """x (B,T,C) bf16, w (taps,N,C) bf16 -> float64 (B,T,N): sum_j x[t+j-pad] @ w[j]^T, zero padded."""
    B, T, C = x.shape
    taps, N, _ = w.shape
    xd, wd = x.double(), w.double()
    out = torch.zeros(B, T, N, dtype=torch.float64, device=x.device)
    for j in range(taps):
        sh = j - pad_left
        lo, hi = max(0, -sh), min(T, T - sh)
        if hi > lo:
            out[:, lo:hi] += xd[:, lo + sh:hi + sh] @ wd[j].T
    return out


def ln_ref(v, g, b, eps=1e-5):
    m = v.mean(-1, keepdim=True)
    var = ((v - m) ** 2).mean(-1, keepdim=True)
    return (v - m) / torch.sqrt(var + eps) * g.double() + b.double()


def rnd(*shape, scale=1.0, dtype=torch.float32, seed=None):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed if seed is not None else (hash(shape) & 0xFFFF))
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(dev())


@pytest.mark.parametrize("B,T,C,N,taps,pad,bn,act", [
    (2, 200, 384, 1152, 1, 0, 128, 0),   # fused q/k/v linears
    (3, 130, 384, 1536, 5, 2, 256, 1),   # encoder FFN conv1 k=5 + ReLU
    (2, 257, 384, 1536, 1, 0, 256, 1),   # decoder FFN conv1 k=1 + ReLU
    (2, 140, 384, 80, 1, 0, 80, 0),      # postnet out linear
    (2, 140, 80, 384, 5, 4, 128, 2),     # postnet conv1 (causal, C=80 -> partial K box) + tanh
    (2, 140, 384, 384, 5, 4, 192, 2),    # postnet conv_list
    (1, 64, 384, 384, 1, 0, 128, 0),     # single short utterance
])
def test_gemm_affine_act(L, B, T, C, N, taps, pad, bn, act):
    x = rnd(B, T, C, dtype=torch.bfloat16, seed=1)
    w = rnd(taps, N, C, scale=1.0 / math.sqrt(C * taps), dtype=torch.bfloat16, seed=2)
    bias = rnd(N, seed=3)
    scale = (1.0 + 0.1 * rnd(N, seed=4))
    res = rnd(B, T, N, seed=5)
    of = torch.empty(B, T, N, device=dev())
    ob = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
    L.conv_gemm(x, w, bias, taps=taps, pad_left=pad, block_n=bn, epi=L.EPI_AFFINE_ACT, act=act, scale=scale,
                res_f32=res, res_bstride=T * N, res_tstride=N, out_f32=of, out_bf16=ob)
    torch.cuda.synchronize()
    v = conv_ref(x, w, pad) * scale.double() + bias.double()
    v = {0: v, 1: torch.relu(v), 2: torch.tanh(v)}[act] + res.double()
    assert rel_err(of, v) < 2e-4, rel_err(of, v)
    assert rel_err(ob, v) < 1e-2


@pytest.mark.parametrize("B,T,C,N,taps,pad,bn,act", [
    (3, 333, 384, 1152, 1, 0, 192, 0),    # q/k/v: A-stationary sweep over 6 weight tiles, TMA stores
    (2, 1103, 384, 1536, 1, 0, 256, 1),   # decoder FFN conv1: A-stationary, ragged last tile clipped by TMA
    (70, 300, 384, 1536, 1, 0, 256, 1),   # more m-tiles than SMs: persistent loop, several units per CTA
    (3, 130, 384, 1536, 5, 2, 256, 1),    # encoder FFN conv1 k=5: streaming A, double-buffered accumulators
    (2, 140, 80, 384, 5, 4, 192, 2),      # postnet conv1, scale+shift+tanh
    (40, 200, 384, 384, 5, 4, 192, 2),    # postnet conv_list, many tiles per CTA
    (2, 200, 384, 384, 1, 0, 128, 0),
    (2, 200, 256, 384, 1, 0, 192, 0),     # C = 256: four K blocks, A-stationary
])
def test_gemm_bf16_tma_store_paths(L, B, T, C, N, taps, pad, bn, act):
    """bf16-only output: accumulator -> swizzled smem staging -> TMA store (the production path)."""
    x = rnd(B, T, C, dtype=torch.bfloat16, seed=101)
    w = rnd(taps, N, C, scale=1.0 / math.sqrt(C * taps), dtype=torch.bfloat16, seed=102)
    bias = rnd(N, seed=103)
    scale = (1.0 + 0.1 * rnd(N, seed=104)) if act == 2 else None
    ob = torch.full((B * T + 128, N), 7.0, device=dev(), dtype=torch.bfloat16)
    out = ob[: B * T].view(B, T, N)
    L.conv_gemm(x, w, bias, taps=taps, pad_left=pad, block_n=bn, epi=L.EPI_AFFINE_ACT, act=act, scale=scale,
                out_bf16=out)
    torch.cuda.synchronize()
    v = conv_ref(x, w, pad)
    if scale is not None:
        v = v * scale.double()
    v = v + bias.double()
    v = {0: v, 1: torch.relu(v), 2: torch.tanh(v)}[act]
    assert rel_err(out, v) < 1e-2, rel_err(out, v)
    assert (out.double() - v).abs().max().item() < 0.05 * max(1.0, v.abs().max().item())
    assert (ob[B * T:] == 7.0).all(), "TMA store wrote past the tensor"


@pytest.mark.parametrize("B,T", [(2, 300), (40, 500)])  # second case: more tiles than SMs (persistent loop)
def test_gemm_res_ln(L, B, T):
    C, N = 384, 384
    x = rnd(B, T, C, dtype=torch.bfloat16, seed=11)
    w = rnd(1, N, C, scale=1 / math.sqrt(C), dtype=torch.bfloat16, seed=12)
    bias, g, b = rnd(N, seed=13), 1 + 0.1 * rnd(N, seed=14), 0.1 * rnd(N, seed=15)
    res = rnd(B, T, N, seed=16)
    of = torch.empty(B, T, N, device=dev())
    ob = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
    L.conv_gemm(x, w, bias, block_n=384, epi=L.EPI_RES_LN, res_f32=res, res_bstride=T * N, res_tstride=N,
                ln1=(g, b), out_f32=of, out_bf16=ob)
    torch.cuda.synchronize()
    v = conv_ref(x, w, 0) + bias.double() + res.double()
    assert rel_err(of, v) < 2e-4
    assert rel_err(ob, ln_ref(v, g, b)) < 1e-2
    # positional-encoding flavour: residual broadcast over the batch and scaled by a device scalar
    pe = rnd(T, N, seed=17)
    alpha = torch.tensor([0.7], device=dev())
    L.conv_gemm(x, w, bias, block_n=384, epi=L.EPI_RES_LN, res_f32=pe, res_bstride=0, res_tstride=N,
                res_scale=alpha, ln1=(g, b), out_f32=of, out_bf16=ob)
    torch.cuda.synchronize()
    v = conv_ref(x, w, 0) + bias.double() + 0.7 * pe.double()
    assert rel_err(of, v) < 2e-4
    assert rel_err(ob, ln_ref(v, g, b)) < 1e-2


@pytest.mark.parametrize("taps,pad,B,T", [(1, 0, 2, 150), (5, 2, 2, 150), (1, 0, 40, 500)])
def test_gemm_ffn2(L, taps, pad, B, T):
    C, N = 1536, 384
    h = torch.relu(rnd(B, T, C, seed=21)).to(torch.bfloat16)
    w = rnd(taps, N, C, scale=1 / math.sqrt(C * taps), dtype=torch.bfloat16, seed=22)
    bias = rnd(N, seed=23)
    g1, b1, g2, b2 = 1 + 0.1 * rnd(N, seed=24), 0.1 * rnd(N, seed=25), 1 + 0.1 * rnd(N, seed=26), 0.1 * rnd(N, seed=27)
    yn = rnd(B, T, N, dtype=torch.bfloat16, seed=28)
    y = rnd(B, T, N, seed=29)
    z = torch.empty(B, T, N, device=dev())
    zn = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
    zn32 = torch.empty(B, T, N, device=dev())
    L.conv_gemm(h, w, bias, taps=taps, pad_left=pad, block_n=384, epi=L.EPI_FFN2, res_bf16=yn, res_f32=y,
                res_bstride=T * N, res_tstride=N, ln1=(g1, b1), ln2=(g2, b2), out_f32=z, out_bf16=zn, out2_f32=zn32)
    torch.cuda.synchronize()
    v = conv_ref(h, w, pad) + bias.double() + yn.double()
    zr = y.double() + ln_ref(v, g1, b1)
    assert rel_err(z, zr) < 2e-4
    assert rel_err(zn32, ln_ref(zr, g2, b2)) < 2e-4
    assert rel_err(zn, ln_ref(zr, g2, b2)) < 1e-2


@pytest.mark.parametrize("taps,pad,B,T,bn", [(5, 2, 2, 150, 128), (5, 2, 30, 128, 192), (1, 0, 3, 70, 128)])
def test_ffn2_short_batch_path(L, taps, pad, B, T, bn):
    """Short batches: conv2 as a column-split AFFINE GEMM + the ffn2_tail row kernel; same arithmetic as the
    EPI_FFN2 epilogue (modules.py:84-87 + layers.py:40,31), compared with it and with the float64 restatement."""
    C, N = 1536, 384
    h = torch.relu(rnd(B, T, C, seed=21)).to(torch.bfloat16)
    w = rnd(taps, N, C, scale=1 / math.sqrt(C * taps), dtype=torch.bfloat16, seed=22)
    bias = rnd(N, seed=23)
    g1, b1, g2, b2 = 1 + 0.1 * rnd(N, seed=24), 0.1 * rnd(N, seed=25), 1 + 0.1 * rnd(N, seed=26), 0.1 * rnd(N, seed=27)
    yn = rnd(B, T, N, dtype=torch.bfloat16, seed=28)
    y = rnd(B, T, N, seed=29)
    u = torch.empty(B, T, N, device=dev())
    z, zn32, pre = (torch.empty(B, T, N, device=dev()) for _ in range(3))
    zn = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
    L.conv_gemm(h, w, bias, taps=taps, pad_left=pad, block_n=bn, out_f32=u)
    L.ffn2_tail(u, yn, y, (g1, b1), (g2, b2), 1e-5, z, zn, zn32=zn32, pre_f32=pre)
    z2 = torch.empty(B, T, N, device=dev())
    zn2 = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
    pre2 = torch.empty(B, T, N, device=dev())
    L.conv_gemm(h, w, bias, taps=taps, pad_left=pad, block_n=384, epi=L.EPI_FFN2, res_bf16=yn, res_f32=y,
                res_bstride=T * N, res_tstride=N, ln1=(g1, b1), ln2=(g2, b2), out_f32=z2, out_bf16=zn2, pre_f32=pre2)
    torch.cuda.synchronize()
    v = conv_ref(h, w, pad) + bias.double() + yn.double()
    zr = y.double() + ln_ref(v, g1, b1)
    assert rel_err(pre, v) < 2e-4
    assert rel_err(z, zr) < 2e-4
    assert rel_err(zn32, ln_ref(zr, g2, b2)) < 2e-4
    assert rel_err(zn, ln_ref(zr, g2, b2)) < 1e-2
    assert rel_err(z, z2) < 1e-5 and rel_err(pre, pre2) < 1e-5 and rel_err(zn, zn2) < 5e-3


@pytest.mark.parametrize("B,T", [(3, 170), (40, 500)])
def test_gemm_predictor(L, B, T):
    C, N = 384, 256
    x = rnd(B, T, C, dtype=torch.bfloat16, seed=31)
    w1 = rnd(3, N, C, scale=1 / math.sqrt(3 * C), dtype=torch.bfloat16, seed=32)
    b1, g1, be1 = rnd(N, seed=33), 1 + 0.1 * rnd(N, seed=34), 0.1 * rnd(N, seed=35)
    h = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
    L.conv_gemm(x, w1, b1, taps=3, pad_left=1, block_n=256, epi=L.EPI_RELU_LN, ln1=(g1, be1), out_bf16=h)
    torch.cuda.synchronize()
    ref1 = ln_ref(torch.relu(conv_ref(x, w1, 1) + b1.double()), g1, be1)
    assert rel_err(h, ref1) < 1e-2
    w2 = rnd(3, N, N, scale=1 / math.sqrt(3 * N), dtype=torch.bfloat16, seed=36)
    dw, db = rnd(N, scale=0.1, seed=37), torch.tensor([2.4], device=dev())
    mask = torch.ones(B, T, dtype=torch.uint8, device=dev())
    mask[1, 100:] = 0
    mask[2, 33:] = 0
    out = torch.full((B, T), 7.0, device=dev())
    L.conv_gemm(h, w2, b1, taps=3, pad_left=1, block_n=256, epi=L.EPI_RELU_LN_DOT, ln1=(g1, be1), dot_w=dw,
                dot_b=db, row_mask=mask, out_row=out)
    torch.cuda.synchronize()
    u = ln_ref(torch.relu(conv_ref(h, w2, 1) + b1.double()), g1, be1)
    ref = (u * dw.double()).sum(-1) + 2.4
    ref = ref.masked_fill(mask == 0, 0.0)
    assert max_err(out, ref) < 2e-4 * max(1.0, ref.abs().max().item())
    assert (out[1, 100:] == 0).all() and (out[2, 33:] == 0).all()


def attention_ref(qkv, mask, H):
    B, T, W = qkv.shape
    D = W // 3
    Dh = D // H
    q, k, v = [t.double().view(B, T, H, Dh).transpose(1, 2) for t in qkv.split(D, dim=-1)]
    s = q @ k.transpose(-1, -2) / math.sqrt(Dh)
    s = s.masked_fill(mask.view(B, 1, 1, T) == 0, -1e4)
    p = torch.softmax(s, dim=-1)
    return (p @ v).transpose(1, 2).reshape(B, T, D), p


@pytest.mark.parametrize("B,T,lens", [
    (2, 128, [128, 77]),
    (3, 333, [333, 129, 1]),
    (2, 1000, [1000, 640]),
    (2, 200, [200, 0]),      # an utterance with no valid key: uniform over all T keys (the -1e4 fill)
])
def test_attention(L, B, T, lens):
    H, Dh = 4, 96
    qkv = rnd(B, T, 3 * H * Dh, dtype=torch.bfloat16, seed=41)
    mask = torch.zeros(B, T, dtype=torch.uint8, device=dev())
    for i, n in enumerate(lens):
        mask[i, :n] = 1
    kv = L.mask_lengths(mask)
    assert kv[:, 0].tolist() == lens and kv[:, 1].tolist() == lens  # prefix masks: last valid + 1 == clean prefix
    ctx = torch.empty(B, T, H * Dh, device=dev(), dtype=torch.bfloat16)
    L.attention_fwd(qkv, mask, kv, ctx, H)
    torch.cuda.synchronize()
    ref, p_ref = attention_ref(qkv, mask, H)
    # P is rounded to bf16 before P.V (as autocast does): 2^-9 relative per term, averaged
    assert rel_err(ctx, ref) < 1e-2, rel_err(ctx, ref)
    if T <= 333:
        probs = L.attention_probs(qkv, mask, H)
        torch.cuda.synchronize()
        assert max_err(probs, p_ref) < 1e-5


@pytest.mark.parametrize("B,T,lens", [(2, 128, [128, 77]), (3, 333, [333, 129, 1]), (2, 1000, [1000, 640]), (1, 65, [65])])
def test_attention_causal(L, B, T, lens):
    """The causal AND key-padding mask of Transformer-TTS (train.py:26-58): hidden keys get the -1e4 fill; rows beyond an
    utterance's length (padding queries) attend to its valid keys up to their own index, as in the reference."""
    H, Dh = 4, 96
    qkv = rnd(B, T, 3 * H * Dh, dtype=torch.bfloat16, seed=45)
    mask = torch.zeros(B, T, dtype=torch.uint8, device=dev())
    for i, n in enumerate(lens):
        mask[i, :n] = 1
    kv = L.mask_lengths(mask)
    ctx = torch.empty(B, T, H * Dh, device=dev(), dtype=torch.bfloat16)
    L.attention_fwd(qkv, mask, kv, ctx, H, causal=True)
    torch.cuda.synchronize()
    D = H * Dh
    q, k, v = [t.double().view(B, T, H, Dh).transpose(1, 2) for t in qkv.split(D, dim=-1)]
    s = q @ k.transpose(-1, -2) / math.sqrt(Dh)
    full = (mask.view(B, 1, 1, T) != 0) & torch.tril(torch.ones(T, T, dtype=torch.bool, device=dev())).view(1, 1, T, T)
    p = torch.softmax(s.masked_fill(~full, -1e4), dim=-1)
    ref = (p @ v).transpose(1, 2).reshape(B, T, D)
    assert rel_err(ctx, ref) < 1e-2, rel_err(ctx, ref)
    # row 0 sees only key 0: its context is v[0] exactly (bf16)
    assert torch.equal(ctx[:, 0], qkv[:, 0, 2 * D:])


def test_attention_sparse_mask(L):
    """Non-prefix masks are honoured (the reference accepts any boolean key mask)."""
    B, T, H, Dh = 1, 260, 4, 96
    qkv = rnd(B, T, 3 * H * Dh, dtype=torch.bfloat16, seed=43)
    mask = (torch.arange(T, device=dev()) % 3 != 1).to(torch.uint8).view(1, T).contiguous()
    kv = L.mask_lengths(mask)
    ctx = torch.empty(B, T, H * Dh, device=dev(), dtype=torch.bfloat16)
    L.attention_fwd(qkv, mask, kv, ctx, H)
    torch.cuda.synchronize()
    ref, _ = attention_ref(qkv, mask, H)
    assert rel_err(ctx, ref) < 1e-2


def test_embed_pe_ln(L):
    B, T, D, V = 3, 70, 384, 152
    ids = torch.randint(0, V, (B, T), device=dev())
    table, pe = rnd(V, D, seed=51), rnd(100, D, seed=52)
    alpha = torch.tensor([1.3], device=dev())
    g, b = 1 + 0.1 * rnd(D, seed=53), 0.1 * rnd(D, seed=54)
    x0 = torch.empty(B, T, D, device=dev())
    xn = torch.empty(B, T, D, device=dev(), dtype=torch.bfloat16)
    L.embed_pe_ln(ids, table, pe, alpha, g, b, 1e-5, x0, xn)
    torch.cuda.synchronize()
    ref = table[ids] + alpha * pe[:T]
    assert torch.equal(x0, ref)
    assert rel_err(xn, ln_ref(ref.double(), g, b)) < 1e-2


def test_layernorm(L):
    x = rnd(1000, 384, seed=61) * 3 + 0.5
    g, b = 1 + 0.1 * rnd(384, seed=62), 0.1 * rnd(384, seed=63)
    yb = torch.empty(1000, 384, device=dev(), dtype=torch.bfloat16)
    yf = torch.empty(1000, 384, device=dev())
    L.layernorm(x, g, b, 1e-5, yb, yf)
    torch.cuda.synchronize()
    ref = ln_ref(x.double(), g, b)
    assert max_err(yf, ref) < 1e-5
    assert rel_err(yb, ref) < 1e-2
    assert max_err(yf, torch.nn.functional.layer_norm(x, (384,), g, b, 1e-5)) < 1e-5


def test_round_durations(L):
    log_d = torch.cat([torch.linspace(-3, 4.5, 4001), torch.log(torch.arange(1, 40).float() + 0.5),
                       torch.log(torch.arange(1, 40).float() + 1.5)]).to(dev())
    got = L.round_durations(log_d)
    torch.cuda.synchronize()
    ref = torch.clamp(torch.round(torch.exp(log_d.cpu()) - 1.0), min=0).to(torch.int32)  # varianceadaptor.py:75
    diff = (got.cpu() != ref).nonzero().flatten()
    # exp() may differ by 1 ulp between libraries; a flip needs exp(x)-1 within 1 ulp of k+0.5
    for i in diff.tolist():
        e = math.exp(float(log_d[i])) - 1.0
        assert abs((e % 1.0) - 0.5) < 1e-5, (i, e)
    assert len(diff) <= 4


@pytest.mark.parametrize("B,L_,D", [(4, 37, 384), (64, 128, 384)])
def test_length_regulator(L, B, L_, D):
    g = torch.Generator().manual_seed(7)
    dur = torch.randint(0, 16, (B, L_), generator=g).to(torch.int32)
    dur[0, :5] = 0
    if B > 2:
        dur[2] = 0  # an all-zero-duration utterance
    x = rnd(B, L_, D, dtype=torch.bfloat16, seed=71)
    durd = dur.to(dev())
    cum, mel_len = L.lr_scan(durd)
    torch.cuda.synchronize()
    assert torch.equal(cum.cpu(), torch.cumsum(dur, 1).to(torch.int32))
    assert torch.equal(mel_len.cpu(), dur.sum(1).to(torch.int32))
    Tmax = int(mel_len.max().item())
    out, fidx, mel_pos, mel_mask = L.lr_expand(x, cum, mel_len, Tmax)
    torch.cuda.synchronize()
    # reference algorithm: Models/varianceadaptor.py:141-177 (expand/cat/pad)
    xc = x.cpu()
    for b in range(B):
        rows = [xc[b, i:i + 1].expand(int(dur[b, i]), -1) for i in range(L_)]
        exp = torch.cat(rows, 0)
        n = exp.shape[0]
        assert torch.equal(out[b, :n].cpu(), exp)
        assert (out[b, n:] == 0).all()
        idx = torch.repeat_interleave(torch.arange(L_), dur[b].long())
        assert torch.equal(fidx[b, :n].cpu().long(), idx)
        assert (fidx[b, n:] == -1).all()
        pos = torch.cat([torch.arange(1, n + 1), torch.zeros(Tmax - n, dtype=torch.long)])
        assert torch.equal(mel_pos[b].cpu(), pos)
        assert torch.equal(mel_mask[b].cpu().bool(), torch.arange(Tmax) <= pos)  # varianceadaptor.py:257
    # teacher-forced max_len shorter / longer than the expansion (pad crops or zero-fills)
    for Tm in (Tmax - 3, Tmax + 5):
        if Tm <= 0:
            continue
        out2, _, _, _ = L.lr_expand(x, cum, mel_len, Tm)
        torch.cuda.synchronize()
        k = min(Tm, Tmax)
        assert torch.equal(out2[:, :k], out[:, :k])
        assert (out2[:, k:] == 0).all()


def test_bucket_embed_add(L):
    rows, D, nb = 777, 384, 255
    x = rnd(rows, D, dtype=torch.bfloat16, seed=81)
    pbins = torch.exp(torch.linspace(math.log(71.0), math.log(799.8), nb)).to(dev())
    ebins = torch.linspace(0.0, 403.8, nb).to(dev())
    p = (torch.rand(rows) * 900 - 20).to(dev())
    e = (torch.rand(rows) * 450 - 20).to(dev())
    p[:nb] = pbins  # exact bin edges exercise right=False
    e[:nb] = ebins
    ptab, etab = rnd(nb + 1, D, seed=82), rnd(nb + 1, D, seed=83)
    out, pi, ei = L.bucket_embed_add(x, p, e, pbins, ebins, ptab, etab, want_idx=True)
    torch.cuda.synchronize()
    pr, er = torch.bucketize(p, pbins), torch.bucketize(e, ebins)  # varianceadaptor.py:109,118
    assert torch.equal(pi.long(), pr) and torch.equal(ei.long(), er)
    ref = ((x.float() + ptab[pr]) + etab[er]).to(torch.bfloat16)
    assert torch.equal(out, ref)


def test_cast_bf16(L):
    x = rnd(1001, seed=91)
    x4 = rnd(4096, seed=92)
    assert torch.equal(L.cast_bf16(x4), x4.to(torch.bfloat16))
    assert torch.equal(L.cast_bf16(x[:1000].contiguous()), x[:1000].to(torch.bfloat16))


def test_errors_are_loud(L):
    x = torch.zeros(1, 8, 384, dtype=torch.bfloat16)  # CPU tensor: no CPU path exists
    with pytest.raises(RuntimeError):
        L.conv_gemm(x, x, torch.zeros(384), block_n=128)
    xd = torch.zeros(1, 8, 384, dtype=torch.bfloat16, device=dev())
    wd = torch.zeros(1, 384, 384, dtype=torch.bfloat16, device=dev())
    with pytest.raises(RuntimeError, match="unsupported|block_n"):
        L.conv_gemm(xd, wd, torch.zeros(384, device=dev()), block_n=96, out_bf16=torch.empty_like(xd))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_second_device_in_one_process(L):
    """The >48 KB shared-memory opt-in is per device: a process that already ran on cuda:0 must still be able to launch
    every tensor-core kernel on cuda:1 (it is set once per device, not once per process)."""
    outs = []
    for dev in (0, 1):
        with torch.cuda.device(dev):
            g = torch.Generator().manual_seed(3)
            B, T, H, Dh = 2, 200, 4, 96
            qkv = (torch.randn(B, T, 3 * H * Dh, generator=g) * 0.5).to(torch.bfloat16).cuda()
            mask = torch.ones(B, T, dtype=torch.uint8).cuda()
            ctx = torch.empty(B, T, H * Dh, dtype=torch.bfloat16, device="cuda")
            L.attention_fwd(qkv, mask, L.mask_lengths(mask), ctx, H)
            x = (torch.randn(B, T, 384, generator=g) * 0.5).to(torch.bfloat16).cuda()
            w = (torch.randn(1, 1152, 384, generator=g) * 0.05).to(torch.bfloat16).cuda()
            o = torch.empty(B, T, 1152, dtype=torch.bfloat16, device="cuda")
            L.conv_gemm(x, w, torch.zeros(1152, device="cuda"), block_n=192, out_bf16=o)
            torch.cuda.synchronize()
            outs.append((ctx.cpu(), o.cpu()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("B,T,ragged", [(2, 128, False), (3, 300, True), (5, 1000, True)])
def test_ffn_fused_is_bit_identical_to_the_two_kernel_path(B, T, ragged):
    """csrc/ffn_fused.cuh: conv1 + ReLU + conv2 + residual + LayerNorm + residual + LayerNorm of a k=1 FFT block in one
    kernel, the hidden activation kept in tensor memory.  Same MMAs in the same order as conv_gemm x 2, so z (fp32) and
    zn (bf16) must be BIT-IDENTICAL on every computed row; both are also checked against an fp64 reference (whose
    hidden activation is rounded to bf16 like the kernels')."""
    import torch.nn.functional as F

    from transformer_tts_b200 import lib

    torch.manual_seed(B * 100 + T)
    dev, D, Hd = "cuda", 384, 1536
    x = (torch.randn(B, T, D, device=dev) * 0.7).bfloat16()
    y = torch.randn(B, T, D, device=dev)
    w1 = (torch.randn(1, Hd, D, device=dev) * 0.05).bfloat16()
    w2 = (torch.randn(1, D, Hd, device=dev) * 0.03).bfloat16()
    b1, b2 = torch.randn(Hd, device=dev) * 0.1, torch.randn(D, device=dev) * 0.1
    ln1 = (torch.randn(D, device=dev), torch.randn(D, device=dev))
    ln2 = (torch.randn(D, device=dev), torch.randn(D, device=dev))
    lens = torch.randint(T // 3, T + 1, (B,)) if ragged else torch.full((B,), T)
    lens[0] = T
    mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).to(dev)
    kv = lib.mask_lengths(mask)
    for tmap in (None, lib.build_tile_map(kv, T)):
        z = torch.full((B, T, D), 7.0, device=dev)
        zn = torch.full((B, T, D), 7.0, device=dev, dtype=torch.bfloat16)
        lib.ffn_fused(x, w1, b1, w2, b2, y, ln1, ln2, z, zn, tile_map=tmap)
        h = torch.empty(B, T, Hd, device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(x, w1, b1, block_n=256, act=lib.ACT_RELU, out_bf16=h)
        z2, zn2 = torch.empty_like(z), torch.empty_like(zn)
        lib.conv_gemm(h, w2, b2, block_n=D, epi=lib.EPI_FFN2, res_bf16=x, res_f32=y, res_bstride=T * D, res_tstride=D,
                      ln1=ln1, ln2=ln2, out_f32=z2, out_bf16=zn2)
        m_tiles = (T + 127) // 128
        tiles = ((lens + 127) // 128).clamp(max=m_tiles) if tmap is not None else torch.full((B,), m_tiles)
        rows = torch.arange(T)[None] < (tiles * 128)[:, None]
        rows = rows.to(dev)
        assert torch.equal(z[rows], z2[rows]) and torch.equal(zn[rows], zn2[rows])
        if tmap is not None:  # rows of dropped tiles are neither read nor written
            assert bool((z[~rows] == 7.0).all()) and bool((zn[~rows].float() == 7.0).all())
        xd = x.double()
        hd = torch.relu(xd @ w1[0].double().T + b1.double()).bfloat16().double()
        u = hd @ w2[0].double().T + b2.double() + xd
        zr = y.double() + F.layer_norm(u, (D,), ln1[0].double(), ln1[1].double(), 1e-5)
        znr = F.layer_norm(zr, (D,), ln2[0].double(), ln2[1].double(), 1e-5)
        e_z = float((z[rows].double() - zr[rows]).norm() / zr[rows].norm())
        e_zn = float((zn[rows].double() - znr[rows]).norm() / znr[rows].norm())
        assert e_z < 1e-4 and e_zn < 4e-3, (e_z, e_zn)


def test_fp32_second_outputs_of_lr_expand_and_bucket_embed_add():
    """Slots 5 / 6 of the reference's return tuple are fp32: the producing kernels write the fp32 copy themselves; it must
    be exactly the bf16 output widened."""
    from transformer_tts_b200 import lib as L

    torch.manual_seed(4)
    B, Lp, D = 3, 20, 384
    x = torch.randn(B, Lp, D, device="cuda").bfloat16()
    dur = torch.randint(0, 7, (B, Lp), device="cuda", dtype=torch.int32)
    cum, mel_len = L.lr_scan(dur)
    Tmax = int(mel_len.max())
    o32 = torch.full((B, Tmax, D), 9.0, device="cuda")
    out, _, _, _ = L.lr_expand(x, cum, mel_len, Tmax, out_f32=o32)
    assert torch.equal(o32, out.float())
    p = torch.rand(B, Tmax, device="cuda") * 700 + 80
    e = torch.rand(B, Tmax, device="cuda") * 400
    pb = torch.exp(torch.linspace(4.26, 6.68, 255)).cuda()
    eb = torch.linspace(0.0, 403.8, 255).cuda()
    pt, et = torch.randn(256, D, device="cuda"), torch.randn(256, D, device="cuda")
    x32 = torch.full((B, Tmax, D), 9.0, device="cuda")
    y, _, _ = L.bucket_embed_add(out, p, e, pb, eb, pt, et, out_f32=x32)
    assert torch.equal(x32, y.float())
