"""The fp32-accurate encoder -> duration chain (csrc/precise.cu) against fp64 references, through the C ABI.

Bars: the split-fp16 tensor-core GEMM / attention must be at least as accurate as plain fp32 arithmetic (torch fp32
with TF32 off is evaluated beside it against the same fp64 reference); row kernels 1e-6; the whole chain's log_d within
1e-5 of the fp64 oracle, which is what makes free-running durations bit-exact (tests/test_model_gpu.py)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-300))


@pytest.fixture(scope="module", autouse=True)
def _no_tf32():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False


def unsplit(t):
    n = t.shape[-1] // 2
    return t[..., :n].double() + t[..., n:].double()


def test_split_f16_roundtrip():
    from transformer_tts_b200 import lib

    x = torch.randn(37, 384, device="cuda") * torch.logspace(-6, 3, 384, device="cuda")
    s = lib.split_f16(x)
    assert s.dtype == torch.float16 and s.shape == (37, 768)
    err = (unsplit(s) - x.double()).abs()
    # 22 mantissa bits where the lo half is a normal fp16 number, an absolute 2^-25 below that
    assert bool((err <= x.double().abs() * 2.0 ** -21 + 2.0 ** -24).all())
    big = torch.tensor([[1e6, -1e6, 7e4, 3.0]], device="cuda")
    assert torch.isfinite(lib.split_f16(big)).all()  # saturates instead of overflowing to inf


@pytest.mark.parametrize("B,T,C,N,taps", [(2, 200, 384, 256, 3), (3, 128, 1536, 384, 5), (2, 130, 384, 1152, 1),
                                          (1, 64, 384, 1536, 5), (5, 77, 256, 256, 3)])
def test_gemm_x3_vs_fp64(B, T, C, N, taps):
    from transformer_tts_b200 import lib
    from transformer_tts_b200.Models import _engine as E

    torch.manual_seed(B * 1000 + T)
    x = torch.randn(B, T, C, device="cuda")
    w = torch.randn(taps, N, C, device="cuda") * 0.05
    bias = torch.randn(N, device="cuda")
    xs = lib.split_f16(x)
    ws, inv = E.split_weight(w)
    pad = taps // 2
    ref = F.conv1d(x.double().transpose(1, 2), w.double().permute(1, 2, 0), bias.double(), padding=pad).transpose(1, 2)
    r32 = F.conv1d(x.transpose(1, 2), w.permute(1, 2, 0), bias, padding=pad).transpose(1, 2)
    out = lib.gemm_x3(xs, ws, bias, inv, taps=taps, pad_left=pad)
    e_ours, e_32 = rel(out, ref), rel(r32, ref)
    outs = lib.gemm_x3(xs, ws, bias, inv, taps=taps, pad_left=pad, relu=True, split_out=True)
    e_split = rel(unsplit(outs), ref.clamp(min=0))
    print(f"\n[gemm_x3 B{B} T{T} C{C} N{N} k{taps}] rel ours {e_ours:.2e} split+relu {e_split:.2e} torch-fp32 {e_32:.2e}")
    assert e_ours < 4e-7 and e_split < 4e-7
    assert e_ours < 1.5 * e_32 + 1e-7


def test_attention_x3_vs_fp64():
    from transformer_tts_b200 import lib

    torch.manual_seed(5)
    B, T, H, Dh = 4, 200, 4, 96
    D = H * Dh
    qkv = torch.randn(B, T, 3 * D, device="cuda")
    lens = torch.tensor([200, 131, 64, 0])
    mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).cuda()
    mask[1, 5] = 0  # a hole inside the valid prefix
    kv = lib.mask_lengths(mask)
    ctx = lib.attention_x3(lib.split_f16(qkv), mask, kv, H)
    q, k, v = (t.double().view(B, T, H, Dh).transpose(1, 2) for t in qkv.split(D, -1))
    sc = ((q @ k.transpose(-1, -2)) / Dh ** 0.5).masked_fill(mask[:, None, None, :] == 0, -1e4)
    ref = (torch.softmax(sc, -1) @ v).transpose(1, 2).reshape(B, T, D)
    q32, k32, v32 = (t.view(B, T, H, Dh).transpose(1, 2) for t in qkv.split(D, -1))
    sc32 = ((q32 @ k32.transpose(-1, -2)) / Dh ** 0.5).masked_fill(mask[:, None, None, :] == 0, -1e4)
    r32 = (torch.softmax(sc32, -1) @ v32).transpose(1, 2).reshape(B, T, D)
    e_ours, e_32 = rel(unsplit(ctx), ref), rel(r32, ref)
    print(f"\n[attention_x3] rel ours {e_ours:.2e} torch-fp32 {e_32:.2e}")
    assert e_ours < 1e-6 and e_ours < 1.5 * e_32 + 1e-7
    # the utterance without any valid key degenerates to a uniform average over all T keys, like the reference
    assert rel(unsplit(ctx)[3], v.transpose(1, 2).reshape(B, T, D)[3].mean(0, keepdim=True).expand(T, D)) < 1e-6


def test_xrow_modes_vs_fp64():
    from transformer_tts_b200 import lib

    torch.manual_seed(9)
    rows, D = 300, 384
    dev = "cuda"
    a, b0, c = (torch.randn(rows, D, device=dev) for _ in range(3))
    g1, b1, g2, b2 = (torch.randn(D, device=dev) for _ in range(4))
    ln = lambda x, g, b: F.layer_norm(x.double(), (x.shape[-1],), g.double(), b.double(), 1e-5)  # noqa: E731
    # ADD_LN
    s_out = torch.empty_like(a)
    o_s = torch.empty(rows, 2 * D, device=dev, dtype=torch.float16)
    o_b = torch.empty(rows, D, device=dev, dtype=torch.bfloat16)
    o_f = torch.empty_like(a)
    lib.xrow(lib.XROW_ADD_LN, a, (g1, b1), b0=b0, s_out=s_out, out_split=o_s, out_bf16=o_b, out_f32=o_f)
    ref = ln(a.double() + b0.double(), g1, b1)
    assert torch.equal(s_out, a + b0)
    assert rel(o_f, ref) < 1e-6 and rel(unsplit(o_s), ref) < 1e-6 and rel(o_b, ref) < 4e-3
    # FFN_TAIL with the split inner residual
    yn_s = lib.split_f16(b0)
    z = torch.empty_like(a)
    lib.xrow(lib.XROW_FFN_TAIL, a, (g1, b1), b0_split=yn_s, c=c, ln2=(g2, b2), z_out=z, out_split=o_s, out_f32=o_f)
    zr = c.double() + ln(a.double() + b0.double(), g1, b1)
    assert rel(z, zr) < 1e-6 and rel(o_f, ln(zr, g2, b2)) < 1e-6 and rel(unsplit(o_s), ln(zr, g2, b2)) < 1e-6
    # LN_DOT (the predictor head), D = 256, with masked rows
    D2 = 256
    a2 = torch.relu(torch.randn(rows, D2, device=dev))
    g, bb, w = (torch.randn(D2, device=dev) for _ in range(3))
    wb = torch.randn(1, device=dev)
    m = (torch.rand(rows, device=dev) > 0.2).to(torch.uint8)
    out = torch.empty(rows, device=dev)
    lib.xrow(lib.XROW_LN_DOT, a2, (g, bb), dot_w=w, dot_b=wb, row_mask=m, out_row=out)
    r = (ln(a2, g, bb) @ w.double() + wb.double()) * m.double()
    assert float((out.double() - r).abs().max()) < 2e-5 * float(r.abs().max()) and bool((out[m == 0] == 0).all())


@pytest.mark.parametrize("B,L,ragged,seed", [(1, 64, False, 1234), (4, 48, True, 77), (16, 128, True, 3)])
def test_chain_log_d_vs_fp64_oracle(B, L, ragged, seed):
    """Encoder + duration predictor: our log_d against the oracle evaluated in fp64; the oracle's own fp32 run is the
    yardstick.  Also the bf16 copy of the encoder output that feeds the frame-level stack."""
    from oracle import fastspeech2_oracle as O
    from transformer_tts_b200 import build_fastspeech2, lib, synth
    from transformer_tts_b200.Models import _engine as E

    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    model = build_fastspeech2(hp).eval()
    model.load_state_dict(sd)
    model = model.cuda()
    sd32 = {k: v.cuda() for k, v in sd.items()}
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd32.items()}
    text, pos, _ = synth.synth_text_batch(B, L, ragged=ragged, seed=seed)
    text, pos = text.cuda(), pos.cuda()
    sm = (pos != 0).unsqueeze(-2)
    m8 = E.mask_to_u8(sm, B, L, text.device)
    kv = lib.mask_lengths(m8)
    with torch.no_grad():
        e_bf, e_s, e32 = E.run_encoder_precise(model.encoder, text, m8, kv)
        log_d = E.run_predictor_precise(model.variance_adaptor.duration_predictor, e_s, m8)
        pe = O.positional_table(hp.d_model_encoder)[:L].cuda()

        def run(sdx, dt):
            x = F.embedding(text, sdx["encoder.embed.weight"]) + sdx["encoder.pe.alpha"] * pe.to(dt).unsqueeze(0)
            for i in range(hp.n_layer_encoder):
                x, _ = O.encoder_layer(sdx, f"encoder.layers.{i}.", x, sm, hp.n_head_encoder)
            x = F.layer_norm(x, (x.shape[-1],), sdx["encoder.norm.weight"], sdx["encoder.norm.bias"])
            return x, O.variance_predictor(sdx, "variance_adaptor.duration_predictor.", x, sm)

        e64, ld64 = run(sd64, torch.float64)
        _, ld32 = run(sd32, torch.float32)
    valid = pos != 0
    err = float((log_d.double() - ld64)[valid].abs().max())
    err32 = float((ld32.double() - ld64)[valid].abs().max())
    print(f"\n[chain B{B} L{L}] log_d max abs err: ours {err:.2e}, oracle fp32 {err32:.2e}; encoder rel {rel(e32[valid], e64[valid]):.2e}")
    assert err < 1e-5 and err < 2 * err32 + 1e-6
    assert rel(e32[valid], e64[valid]) < 1e-6 and rel(unsplit(e_s)[valid], e64[valid]) < 1e-6
    assert rel(e_bf[valid], e64[valid]) < 4e-3  # one bf16 rounding of the fp32-accurate value
    assert bool((log_d[~valid] == 0).all())
