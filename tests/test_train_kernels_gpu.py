"""Kernel-level parity of the TRAINING-step kernels (include/ttsb200.h, "TRAINING STEP") against torch
autograd on the same operands: what the reference obtains from loss.backward() (train_fastspeech2.py:304).

Tolerances: fp32 reductions 1e-3 relative (fp32 atomics, order noise), bf16 outputs 1e-2 (2^-9 per rounding;
the attention backward rounds P and dS to bf16 before the second GEMM like every flash-attention backward).
"""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    from transformer_tts_b200 import lib

    lib.load()
    return lib


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def rnd(*shape, scale=1.0, dtype=torch.float32, seed=0):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(dev())


def conv_torch(x, w_oik, pad_left, taps):
    """x (B,T,C) fp64, w (N,C,k): the reference conv1d with `pad_left` zeros on the left, output length T."""
    xp = F.pad(x.transpose(1, 2), (pad_left, taps - 1 - pad_left))
    return F.conv1d(xp, w_oik).transpose(1, 2)


# ------------------------------------------------------------------------------------------------
# wgrad
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,T,N,C,taps,pad", [
    (2, 200, 384, 384, 1, 0),
    (3, 130, 1536, 384, 5, 2),    # encoder FFN conv1
    (2, 257, 384, 1536, 1, 0),    # decoder FFN conv2
    (2, 140, 80, 384, 5, 4),      # postnet conv2 (N = 80 < tile)
    (2, 140, 384, 80, 5, 4),      # postnet conv1 (C = 80 < tile)
    (2, 90, 256, 256, 3, 1),      # predictor conv2
    (16, 300, 384, 384, 1, 0),    # several split-K ranges
])
def test_conv_wgrad(L, B, T, N, C, taps, pad):
    x = rnd(B, T, C, dtype=torch.bfloat16, seed=1)
    dy = rnd(B, T, N, dtype=torch.bfloat16, seed=2)
    w = torch.zeros(N, C, taps, dtype=torch.float64, device=dev(), requires_grad=True)
    y = conv_torch(x.double(), w, pad, taps)
    (y * dy.double()).sum().backward()
    dw = torch.zeros(N, C, taps, device=dev())
    L.conv_wgrad(dy, x, dw, taps=taps, pad_left=pad, conv_layout=True)
    assert rel_err(dw, w.grad) < 1e-3
    dw2 = torch.zeros(taps, N, C, device=dev())
    L.conv_wgrad(dy, x, dw2, taps=taps, pad_left=pad)
    assert rel_err(dw2, w.grad.permute(2, 0, 1)) < 1e-3
    # the bias gradient in the same launch (the ones operand): accumulated onto what is there, dW unchanged by it
    db = torch.full((N,), 0.5, device=dev())
    dw3 = torch.zeros(taps, N, C, device=dev())
    L.conv_wgrad(dy, x, dw3, taps=taps, pad_left=pad, dbias=db)
    assert rel_err(db, 0.5 + dy.double().sum((0, 1))) < 1e-5
    assert rel_err(dw3, w.grad.permute(2, 0, 1)) < 1e-3


def test_conv_wgrad_column_slice(L):
    B, T, D = 2, 150, 384
    x = rnd(B, T, D, dtype=torch.bfloat16, seed=1)
    dqkv = rnd(B, T, 3 * D, dtype=torch.bfloat16, seed=2)
    for i in range(3):
        dw = torch.zeros(D, D, device=dev())
        L.conv_wgrad(dqkv, x, dw, n_off=i * D, N=D)
        ref = dqkv[..., i * D:(i + 1) * D].double().reshape(-1, D).T @ x.double().reshape(-1, D)
        assert rel_err(dw, ref) < 1e-3
        db = torch.zeros(D, device=dev())
        L.conv_wgrad(dqkv, x, torch.zeros(D, D, device=dev()), n_off=i * D, N=D, dbias=db)
        assert rel_err(db, dqkv[..., i * D:(i + 1) * D].double().sum((0, 1))) < 1e-5


# ------------------------------------------------------------------------------------------------
# dgrad through conv_gemm with transposed, tap-flipped weights; ReLU gate; pre_f32
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,T,N,C,taps,pad,bn", [
    (2, 200, 1536, 384, 5, 2, 192),
    (2, 140, 384, 384, 5, 4, 192),
    (2, 257, 1536, 384, 1, 0, 128),
    (2, 140, 80, 384, 5, 4, 128),
])
def test_conv_dgrad(L, B, T, N, C, taps, pad, bn):
    """forward conv: C -> N; dgrad: dy (B,T,N) -> dx (B,T,C)."""
    w = rnd(N, C, taps, scale=1 / math.sqrt(C * taps), dtype=torch.bfloat16, seed=3)
    dy = rnd(B, T, N, dtype=torch.bfloat16, seed=4)
    x = torch.zeros(B, T, C, dtype=torch.float64, device=dev(), requires_grad=True)
    (conv_torch(x, w.double(), pad, taps) * dy.double()).sum().backward()
    wt = w.permute(2, 1, 0).flip(0).contiguous()  # (taps, C, N): w'[j'][c][n] = w[n][c][taps-1-j']
    zero = torch.zeros(C, device=dev())
    dx = torch.empty(B, T, C, device=dev())
    L.conv_gemm(dy, wt, zero, taps=taps, pad_left=taps - 1 - pad, block_n=bn, out_f32=dx)
    assert rel_err(dx, x.grad) < 2e-3


def test_relu_gate(L):
    B, T, N, C = 2, 200, 1536, 384
    w = rnd(1, N, C, scale=1 / math.sqrt(C), dtype=torch.bfloat16, seed=3)
    x = rnd(B, T, C, dtype=torch.bfloat16, seed=4)
    gate = rnd(B, T, N, dtype=torch.bfloat16, seed=5)
    zero = torch.zeros(N, device=dev())
    for use_f32 in (False, True):
        ob = torch.empty(B, T, N, device=dev(), dtype=torch.bfloat16)
        of = torch.empty(B, T, N, device=dev()) if use_f32 else None
        L.conv_gemm(x, w, zero, block_n=256, act=L.ACT_RELU_GRAD, gate_bf16=gate, out_bf16=ob, out_f32=of)
        ref = (x.double() @ w[0].double().T) * (gate.double() > 0)
        assert rel_err(ob, ref) < 1e-2
        if use_f32:
            assert rel_err(of, ref) < 2e-3


def test_bce_logits_loss_and_relu_gate_rows(L):
    """train.py:217 binary_cross_entropy_with_logits(reduction='mean', pos_weight) and its gradient; the elementwise ReLU
    backward used by the decoder prenet (prenets.py:33-38)."""
    g = torch.Generator().manual_seed(9)
    for n, pw in ((4 * 301 * 2, 5.0), (7, 1.0), (100003, 0.25)):
        x = (4 * torch.randn(n, generator=g)).to(dev())
        x[: min(n, 4)] = torch.tensor([60.0, -60.0, 0.0, 1e-4])[: min(n, 4)].to(dev())   # both tails of log1p(exp(.))
        y = (torch.rand(n, generator=g) < 0.3).float().to(dev())
        loss = torch.zeros(1, device=dev())
        grad = torch.empty_like(x)
        L.bce_logits_loss(x, y, loss, grad=grad, pos_weight=pw)
        xr = x.double().requires_grad_(True)
        ref = F.binary_cross_entropy_with_logits(xr, y.double(), reduction="mean", pos_weight=torch.tensor(pw, device=dev()))
        ref.backward()
        assert abs(loss.item() - ref.item()) < 2e-6 * max(1.0, abs(ref.item())), (n, loss.item(), ref.item())
        assert rel_err(grad, xr.grad) < 1e-6
        l2 = torch.zeros(1, device=dev())
        L.bce_logits_loss(x, y, l2, pos_weight=pw)       # loss only
        assert abs(l2.item() - loss.item()) <= 2e-6 * max(1.0, abs(loss.item()))   # (block sums meet in one atomic: order varies)
    gr = rnd(3, 77, 384, seed=5)
    gate = torch.relu(rnd(3, 77, 384, dtype=torch.bfloat16, seed=6))
    out = L.relu_gate(gr, gate)
    assert out.dtype == torch.bfloat16
    assert torch.equal(out, (gr * (gate > 0)).to(torch.bfloat16))


def test_pre_f32(L):
    B, T, D = 2, 200, 384
    x = rnd(B, T, 4 * D, dtype=torch.bfloat16, seed=1)
    w = rnd(1, D, 4 * D, scale=1 / math.sqrt(4 * D), dtype=torch.bfloat16, seed=2)
    bias = rnd(D, seed=3)
    yn = rnd(B, T, D, dtype=torch.bfloat16, seed=4)
    y = rnd(B, T, D, seed=5)
    g1, b1, g2, b2 = (1 + 0.1 * rnd(D, seed=6)), rnd(D, seed=7), (1 + 0.1 * rnd(D, seed=8)), rnd(D, seed=9)
    z = torch.empty(B, T, D, device=dev())
    zn = torch.empty(B, T, D, device=dev(), dtype=torch.bfloat16)
    pre = torch.empty(B, T, D, device=dev())
    L.conv_gemm(x, w, bias, block_n=D, epi=L.EPI_FFN2, res_bf16=yn, res_f32=y, res_bstride=T * D, res_tstride=D,
                ln1=(g1, b1), ln2=(g2, b2), out_f32=z, out_bf16=zn, pre_f32=pre)
    ref = x.double() @ w[0].double().T + bias.double() + yn.double()
    assert rel_err(pre, ref) < 2e-4
    u = F.layer_norm(ref, (D,), g1.double(), b1.double())
    assert rel_err(z, y.double() + u) < 2e-4


# ------------------------------------------------------------------------------------------------
# attention forward lse + backward
# ------------------------------------------------------------------------------------------------
def attn_ref(qkv, mask, H):
    B, T, W = qkv.shape
    D = W // 3
    dk = D // H
    q, k, v = (qkv[..., i * D:(i + 1) * D].reshape(B, T, H, dk).transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) / math.sqrt(dk)
    s = s.masked_fill(mask[:, None, None, :] == 0, -1e4)
    p = torch.softmax(s, -1)
    lse = torch.logsumexp(s, -1)
    return (p @ v).transpose(1, 2).reshape(B, T, D), lse


@pytest.mark.parametrize("B,T,lens", [
    (2, 128, [128, 100]),
    (3, 300, [300, 257, 64]),
    (2, 130, [130, 1]),
])
def test_attention_bwd(L, B, T, lens):
    H, D = 4, 384
    qkv = rnd(B, T, 3 * D, scale=1.0, dtype=torch.bfloat16, seed=1)
    dctx = rnd(B, T, D, scale=1.0, dtype=torch.bfloat16, seed=2)
    mask = (torch.arange(T)[None, :] < torch.tensor(lens)[:, None]).to(torch.uint8).to(dev())
    kv = L.mask_lengths(mask)
    ctx = torch.empty(B, T, D, device=dev(), dtype=torch.bfloat16)
    lse2 = torch.empty(B, H, T, device=dev())
    L.attention_fwd(qkv, mask, kv, ctx, H, lse2=lse2)
    x = qkv.double().requires_grad_(True)
    ref_ctx, ref_lse = attn_ref(x, mask, H)
    assert rel_err(ctx, ref_ctx) < 1e-2
    assert (lse2.double() * math.log(2.0) - ref_lse).abs().max().item() < 2e-3
    (ref_ctx * dctx.double()).sum().backward()
    delta = L.attn_delta(dctx, ctx, H)
    ref_delta = (dctx.double() * ref_ctx.detach()).reshape(B, T, H, D // H).sum(-1).transpose(1, 2)
    assert rel_err(delta, ref_delta) < 1e-2
    dqkv = L.attention_bwd(qkv, dctx, mask, kv, lse2, delta, H)
    torch.cuda.synchronize()
    for i, name in enumerate("qkv"):
        e = rel_err(dqkv[..., i * D:(i + 1) * D], x.grad[..., i * D:(i + 1) * D])
        assert e < 2e-2, f"d{name}: {e}"


@pytest.mark.parametrize("B,T,lens", [
    (2, 128, [128, 100]),
    (3, 300, [300, 257, 64]),
    (2, 400, [400, 129]),
    (1, 65, [65]),
])
def test_attention_bwd_causal(L, B, T, lens):
    """Causal AND key-padding mask (the target mask of Transformer-TTS, train.py:26-58): forward with lse, backward."""
    H, D = 4, 384
    dk = D // H
    qkv = rnd(B, T, 3 * D, scale=1.0, dtype=torch.bfloat16, seed=11)
    dctx = rnd(B, T, D, scale=1.0, dtype=torch.bfloat16, seed=12)
    mask = (torch.arange(T)[None, :] < torch.tensor(lens)[:, None]).to(torch.uint8).to(dev())
    kv = L.mask_lengths(mask)
    ctx = torch.empty(B, T, D, device=dev(), dtype=torch.bfloat16)
    lse2 = torch.empty(B, H, T, device=dev())
    L.attention_fwd(qkv, mask, kv, ctx, H, lse2=lse2, causal=True)
    x = qkv.double().requires_grad_(True)
    q, k, v = (x[..., i * D:(i + 1) * D].reshape(B, T, H, dk).transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) / math.sqrt(dk)
    full = (mask[:, None, None, :] != 0) & torch.tril(torch.ones(T, T, dtype=torch.bool, device=dev()))[None, None]
    s = s.masked_fill(~full, -1e4)
    ref_ctx = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, T, D)
    ref_lse = torch.logsumexp(s, -1)
    assert rel_err(ctx, ref_ctx) < 1e-2
    assert (lse2.double() * math.log(2.0) - ref_lse).abs().max().item() < 2e-3
    (ref_ctx * dctx.double()).sum().backward()
    delta = L.attn_delta(dctx, ctx, H)
    dqkv = L.attention_bwd(qkv, dctx, mask, kv, lse2, delta, H, causal=True)
    torch.cuda.synchronize()
    for i, name in enumerate("qkv"):
        e = rel_err(dqkv[..., i * D:(i + 1) * D], x.grad[..., i * D:(i + 1) * D])
        assert e < 2e-2, f"d{name}: {e}"


# ------------------------------------------------------------------------------------------------
# LayerNorm backward
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("D,rows", [(384, 1000), (256, 777)])
def test_layernorm_bwd(L, D, rows):
    x = rnd(rows, D, seed=1)
    g = rnd(rows, D, seed=2)
    add = rnd(rows, D, seed=3)
    gamma, beta = 1 + 0.1 * rnd(D, seed=4), rnd(D, seed=5)
    xr = x.double().requires_grad_(True)
    gr, br = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    (F.layer_norm(xr, (D,), gr, br) * g.double()).sum().backward()
    dx = torch.empty(rows, D, device=dev())
    dxb = torch.empty(rows, D, device=dev(), dtype=torch.bfloat16)
    dg, db = torch.zeros(D, device=dev()), torch.zeros(D, device=dev())
    L.layernorm_bwd(x, gamma, g_f32=g, add_f32=add, dx_f32=dx, dx_bf16=dxb, dgamma=dg, dbeta=db)
    assert rel_err(dx, xr.grad + add.double()) < 1e-4
    assert rel_err(dxb, xr.grad + add.double()) < 1e-2
    assert rel_err(dg, gr.grad) < 1e-3 and rel_err(db, br.grad) < 1e-3
    # bf16 gradient source
    dx2 = torch.empty(rows, D, device=dev())
    L.layernorm_bwd(x, gamma, g_bf16=g.to(torch.bfloat16), dx_f32=dx2)
    xr2 = x.double().requires_grad_(True)
    (F.layer_norm(xr2, (D,), gamma.double(), beta.double()) * g.to(torch.bfloat16).double()).sum().backward()
    assert rel_err(dx2, xr2.grad) < 1e-4


def test_layernorm_bwd_predictor_head(L):
    """relu -> LN -> linear(256->1) -> masked_fill (varianceadaptor.py:218-229), gradient w.r.t. the conv output."""
    D, rows = 256, 900
    c = rnd(rows, D, seed=1)
    gout = rnd(rows, seed=2)
    mask = (torch.rand(rows, generator=torch.Generator().manual_seed(3)) > 0.2).to(torch.uint8).to(dev())
    gamma, beta, w = 1 + 0.1 * rnd(D, seed=4), rnd(D, seed=5), rnd(D, seed=6)
    cr = c.double().requires_grad_(True)
    gr, br, wr = (t.double().requires_grad_(True) for t in (gamma, beta, w))
    bias = torch.zeros(1, dtype=torch.float64, device=dev(), requires_grad=True)
    out = (F.layer_norm(F.relu(cr), (D,), gr, br) @ wr + bias).masked_fill(mask == 0, 0.0)
    (out * gout.double()).sum().backward()
    r = F.relu(c)
    dxb = torch.empty(rows, D, device=dev(), dtype=torch.bfloat16)
    dx = torch.empty(rows, D, device=dev())
    dg, db, dw, dbias = (torch.zeros(n, device=dev()) for n in (D, D, D, 1))
    L.layernorm_bwd(r, gamma, beta=beta, g_row=gout, g_vec=w, row_mask=mask, relu_gate=True, dx_f32=dx, dx_bf16=dxb,
                    dgamma=dg, dbeta=db, dvec=dw, dscal=dbias)
    assert rel_err(dx, cr.grad) < 1e-4
    assert rel_err(dg, gr.grad) < 1e-3 and rel_err(db, br.grad) < 1e-3
    assert rel_err(dw, wr.grad) < 1e-3 and rel_err(dbias, bias.grad) < 1e-3


# ------------------------------------------------------------------------------------------------
# column sums, BatchNorm + tanh, LengthRegulator, embeddings, loss, optimizer
# ------------------------------------------------------------------------------------------------
def test_colsum(L):
    x = rnd(3, 333, 1152, dtype=torch.bfloat16, seed=1)
    out = torch.zeros(384, device=dev())
    L.colsum(x, out, cols=384, col_off=384)
    assert rel_err(out, x[..., 384:768].double().sum((0, 1))) < 1e-3
    xf = rnd(1000, 80, seed=2)
    o2 = torch.zeros(80, device=dev())
    L.colsum(xf, o2)
    assert rel_err(o2, xf.double().sum(0)) < 1e-4


def test_bn_tanh(L):
    rows, C = 3000, 384
    x = rnd(rows, C, scale=2.0, seed=1) + 0.3
    gamma, beta = 1 + 0.1 * rnd(C, seed=2), 0.1 * rnd(C, seed=3)
    dh = rnd(rows, C, dtype=torch.bfloat16, seed=4)
    xr = x.double().requires_grad_(True)
    gr, br = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    h_ref = torch.tanh(F.batch_norm(xr, None, None, gr, br, training=True, eps=1e-5))
    stats = torch.zeros(2, C, device=dev())
    L.bn_stats(x, stats)
    mean = stats[0] / rows
    var = stats[1] / rows - mean * mean
    assert rel_err(mean, x.double().mean(0)) < 1e-4
    assert rel_err(var, x.double().var(0, unbiased=False)) < 1e-3
    ms = torch.stack([mean, torch.rsqrt(var + 1e-5)]).contiguous()
    h = torch.empty(rows, C, device=dev(), dtype=torch.bfloat16)
    L.bn_tanh_fwd(x, ms, gamma, beta, h)
    assert rel_err(h, h_ref) < 1e-2
    # backward with the same bf16 h the kernel saw
    hb = h.double()
    dt = dh.double() * (1 - hb * hb)
    xh = (x.double() - mean.double()) * ms[1].double()
    red = torch.zeros(2, C, device=dev())
    L.bn_tanh_bwd(dh, h, x, ms, gamma, red, None, 1.0 / rows, apply=False)
    assert rel_err(red[0], dt.sum(0)) < 1e-3 and rel_err(red[1], (dt * xh).sum(0)) < 1e-3
    dx = torch.empty(rows, C, device=dev(), dtype=torch.bfloat16)
    L.bn_tanh_bwd(dh, h, x, ms, gamma, red, dx, 1.0 / rows, apply=True)
    (h_ref * dh.double()).sum().backward()
    assert rel_err(dx, xr.grad) < 2e-2
    assert rel_err(red[1], gr.grad) < 2e-2 and rel_err(red[0], br.grad) < 2e-2


def test_lr_bwd(L):
    B, Lp, D = 3, 40, 384
    gen = torch.Generator().manual_seed(1)
    dur = torch.randint(0, 9, (B, Lp), generator=gen).to(torch.int32).to(dev())
    cum, mel_len = L.lr_scan(dur)
    Tmax = int(mel_len.max().item()) - 5  # crop a little, like max_len from a shorter mel mask
    g = rnd(B, Tmax, D, seed=2)
    add = rnd(B, Lp, D, seed=3)
    dx = L.lr_bwd(g, cum, add)
    ref = add.double().clone()
    for b in range(B):
        idx = torch.repeat_interleave(torch.arange(Lp, device=dev()), dur[b].long())[:Tmax]
        ref[b].index_add_(0, idx, g[b, :idx.numel()].double())
    assert rel_err(dx, ref) < 1e-5


def test_embed_bwd(L):
    B, T, D, V = 3, 50, 384, 152
    ids = torch.randint(0, V, (B, T), generator=torch.Generator().manual_seed(1)).to(dev())
    g = rnd(B, T, D, seed=2)
    pe = rnd(T, D, seed=3)
    table = torch.zeros(V, D, device=dev())
    dalpha = torch.zeros(1, device=dev())
    L.embed_bwd(g, T, idx=ids, table=table, skip_idx=0, pe=pe, dalpha=dalpha)
    ref = torch.zeros(V, D, dtype=torch.float64, device=dev())
    ref.index_add_(0, ids.reshape(-1), g.double().reshape(-1, D))
    ref[0] = 0
    assert rel_err(table, ref) < 1e-5
    assert abs(dalpha.item() - (g.double() * pe.double()[None]).sum().item()) < 1e-2
    t2 = torch.zeros(256, D, device=dev())
    idx32 = torch.randint(0, 256, (B, T), generator=torch.Generator().manual_seed(4)).to(torch.int32).to(dev())
    L.embed_bwd(g, T, idx=idx32, table=t2)
    ref2 = torch.zeros(256, D, dtype=torch.float64, device=dev())
    ref2.index_add_(0, idx32.long().reshape(-1), g.double().reshape(-1, D))
    assert rel_err(t2, ref2) < 1e-5


def test_l1_loss(L):
    pred, target = rnd(7, 333, 80, seed=1), rnd(7, 333, 80, seed=2)
    loss = torch.zeros(1, device=dev())
    grad = torch.empty_like(pred)
    L.l1_loss(pred, target, loss, grad=grad)
    pr = pred.double().requires_grad_(True)
    ref = F.l1_loss(pr, target.double())
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-5
    assert rel_err(grad, pr.grad) < 1e-6
    d = torch.randint(0, 12, (4, 50), generator=torch.Generator().manual_seed(3)).float().to(dev())
    p = rnd(4, 50, seed=4)
    l2 = torch.zeros(1, device=dev())
    L.l1_loss(p, d, l2, log_target=True)
    assert abs(l2.item() - F.l1_loss(p.double(), torch.log(d.double() + 1)).item()) < 1e-5


def test_adam_clip(L):
    n = 100003
    p0, g = rnd(n, seed=1), rnd(n, scale=0.01, seed=2)
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pr], lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    p, m, v = p0.clone(), torch.zeros(n, device=dev()), torch.zeros(n, device=dev())
    for step in range(1, 4):
        gs = g * step
        pr.grad = gs.clone()
        torch.nn.utils.clip_grad_norm_([pr], 1.0)
        opt.step()
        ss = torch.zeros(1, device=dev())
        L.sumsq(gs, ss)
        assert abs(ss.item() - (gs.double() ** 2).sum().item()) / ss.item() < 1e-4
        L.adam_step(p, gs, m, v, ss, gscale=1.0, clip=1.0, lr=1e-3, beta1=0.9, beta2=0.98, eps=1e-9, step=step)
    assert (p - pr.detach()).abs().max().item() < 1e-5


def test_radam_matches_reference_golden_and_oracle(L):
    """ttsb200_radam_step against tests/golden/radam_steps.npz (the UNMODIFIED radam.RAdam of the reference,
    oracle/make_golden_radam.py) and against the oracle's restatement of radam.py:56-92; steps 1-5 take the
    degenerated-to-SGD branch, 6-8 the rectified one."""
    import numpy as np
    from pathlib import Path

    from oracle import fastspeech2_oracle as O

    gold = np.load(Path(__file__).resolve().parent / "golden" / "radam_steps.npz")
    p0, grads, want = (torch.from_numpy(gold[k]) for k in ("p0", "grads", "p_steps"))
    n = p0.numel()
    buf = torch.zeros(4, (n + 3) // 4 * 4, device=dev())  # 16-byte aligned rows: p, g, m, v
    p, g, m, v = (buf[i, :n] for i in range(4))
    p.copy_(p0)
    po, mo, vo = p0.clone(), torch.zeros(n), torch.zeros(n)
    for t in range(grads.shape[0]):
        g.copy_(grads[t])
        L.radam_step(p, g, m, v, None, gscale=1.0, clip=0.0, lr=1e-3, beta1=0.9, beta2=0.98, eps=1e-9, step=t + 1)
        po, mo, vo = O.radam_step(po, grads[t], mo, vo, t + 1)
        assert (p.cpu() - want[t]).abs().max().item() < 2e-6, t
        assert (p.cpu() - po).abs().max().item() < 2e-6, t
    assert (m.cpu() - torch.from_numpy(gold["exp_avg"])).abs().max().item() < 1e-6
    assert rel_err(v.cpu(), torch.from_numpy(gold["exp_avg_sq"])) < 1e-5
    # clip + gradient-mean scale prologue shared with adam_step
    ss = torch.zeros(1, device=dev())
    g.copy_(grads[0] * 50)
    L.sumsq(g, ss)
    p1 = p.clone()
    m1, v1 = m.clone(), v.clone()
    L.radam_step(p, g, m, v, ss, gscale=0.5, clip=1.0, lr=1e-3, beta1=0.9, beta2=0.98, eps=1e-9, step=9)
    gs = grads[0] * 25
    gs = gs * min(1.0, 1.0 / (gs.norm().item() + 1e-6))
    pw, _, _ = O.radam_step(p1.cpu(), gs, m1.cpu(), v1.cpu(), 9)
    assert (p.cpu() - pw).abs().max().item() < 2e-6


def test_weight_transpose_batched(L):
    """ttsb200_weight_transpose: every dgrad operand of a backward pass in one launch == permute + flip of each operand."""
    shapes = [(1, 384, 384), (9, 1536, 384), (1, 160, 384), (5, 80, 384), (3, 256, 384), (1, 1152, 384), (1, 256, 80), (5, 384, 80)]
    srcs = [rnd(*sh, dtype=torch.bfloat16, seed=10 + i) for i, sh in enumerate(shapes)]
    dsts = [torch.full((t, c, n), 7.0, device=dev(), dtype=torch.bfloat16) for (t, n, c) in shapes]
    L.weight_transpose(srcs, dsts)
    for a, b in zip(srcs, dsts):
        assert torch.equal(b, a.permute(0, 2, 1).flip(0).contiguous())
    with pytest.raises(ValueError):
        L.weight_transpose(srcs[:1], [dsts[1]])
