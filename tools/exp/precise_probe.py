"""GPU check of the fp32-accurate chain (csrc/precise.cu): gemm_x3 / attention_x3 against fp64, then log_d of the
whole encoder + duration predictor against the oracle evaluated in fp64 and fp32."""
import sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib, synth, build_fastspeech2
from transformer_tts_b200.Models import _engine as E
from oracle import fastspeech2_oracle as O

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
torch.manual_seed(0)
dev = "cuda"

def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())

# ---- gemm_x3 ----
for (B, T, C, N, taps) in ((2, 200, 384, 256, 3), (4, 128, 1536, 384, 5), (2, 128, 384, 1152, 1)):
    x = torch.randn(B, T, C, device=dev)
    w = torch.randn(taps, N, C, device=dev) * 0.05
    bias = torch.randn(N, device=dev)
    xs = lib.split_f16(x)
    ws, inv = E.split_weight(w)
    out = lib.gemm_x3(xs, ws, bias, inv, taps=taps, pad_left=taps // 2)
    ref = torch.nn.functional.conv1d(x.double().transpose(1, 2), w.double().permute(1, 2, 0), bias.double(), padding=taps // 2).transpose(1, 2)
    r32 = torch.nn.functional.conv1d(x.transpose(1, 2), w.permute(1, 2, 0), bias, padding=taps // 2).transpose(1, 2)
    outs = lib.gemm_x3(xs, ws, bias, inv, taps=taps, pad_left=taps // 2, relu=True, split_out=True)
    rec = outs[..., :N].double() + outs[..., N:].double()
    print(f"gemm_x3 B{B} T{T} C{C} N{N} k{taps}: rel {rel(out, ref):.2e} (torch fp32 {rel(r32, ref):.2e}) split+relu {rel(rec, ref.clamp(min=0)):.2e}")

# ---- attention_x3 ----
B, T, H, Dh = 3, 200, 4, 96
qkv = torch.randn(B, T, 3 * H * Dh, device=dev)
lens = torch.tensor([200, 131, 64])
mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).to(dev)
mask[1, 5] = 0
kv = lib.mask_lengths(mask)
ctx = lib.attention_x3(lib.split_f16(qkv), mask, kv, H)
D = H * Dh
got = ctx[..., :D].double() + ctx[..., D:].double()
q, k, v = (t.double().view(B, T, H, Dh).transpose(1, 2) for t in qkv.split(D, -1))
sc = (q @ k.transpose(-1, -2)) / Dh ** 0.5
sc = sc.masked_fill(mask[:, None, None, :] == 0, -1e4)
ref = (torch.softmax(sc, -1) @ v).transpose(1, 2).reshape(B, T, D)
q32, k32, v32 = (t.view(B, T, H, Dh).transpose(1, 2) for t in qkv.split(D, -1))
sc32 = ((q32 @ k32.transpose(-1, -2)) / Dh ** 0.5).masked_fill(mask[:, None, None, :] == 0, -1e4)
r32 = (torch.softmax(sc32, -1) @ v32).transpose(1, 2).reshape(B, T, D)
print(f"attention_x3: rel {rel(got, ref):.2e} (torch fp32 {rel(r32, ref):.2e}) max abs {float((got - ref).abs().max()):.2e}")

# ---- the chain ----
hp = synth.make_hp()
sd = synth.synth_state_dict(hp)
model = build_fastspeech2(hp).eval()
model.load_state_dict(sd)
model = model.cuda()
sd32 = {k: v.cuda() for k, v in sd.items()}
sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd32.items()}
import oracle.fastspeech2_oracle as OO
for (Bb, L, seed) in ((1, 64, 1234), (4, 48, 77), (64, 128, 1234)):
    text, pos, lens = synth.synth_text_batch(Bb, L, ragged=Bb == 4, seed=seed)
    text, pos = text.cuda(), pos.cuda()
    sm = (pos != 0).unsqueeze(-2)
    m8 = E.mask_to_u8(sm, Bb, L, text.device)
    kvl = lib.mask_lengths(m8)
    with torch.no_grad():
        torch.cuda.synchronize(); t0 = time.time()
        e_bf, e_s, e32 = E.run_encoder_precise(model.encoder, text, m8, kvl)
        log_d = E.run_predictor_precise(model.variance_adaptor.duration_predictor, e_s, m8)
        torch.cuda.synchronize(); t1 = time.time()
        e_bf, e_s, e32 = E.run_encoder_precise(model.encoder, text, m8, kvl)
        log_d = E.run_predictor_precise(model.variance_adaptor.duration_predictor, e_s, m8)
        torch.cuda.synchronize(); t2 = time.time()
        # oracle in fp64 and fp32 (PE table is fp32 in both)
        pe = OO.positional_table(384)
        def run(sdx, dt):
            x = torch.nn.functional.embedding(text, sdx["encoder.embed.weight"])
            x = x + sdx["encoder.pe.alpha"] * pe[:L].to(x.device).to(dt).unsqueeze(0)
            for i in range(hp.n_layer_encoder):
                x, _ = OO.encoder_layer(sdx, f"encoder.layers.{i}.", x, sm, hp.n_head_encoder)
            x = torch.nn.functional.layer_norm(x, (384,), sdx["encoder.norm.weight"], sdx["encoder.norm.bias"])
            return x, OO.variance_predictor(sdx, "variance_adaptor.duration_predictor.", x, sm)
        e64, ld64 = run(sd64, torch.float64)
        e32o, ld32 = run(sd32, torch.float32)
        model.duration_precision = "bf16"
    valid = pos != 0
    d64 = torch.exp(ld64) - 1
    dist = ((d64 % 1.0) - 0.5).abs()[valid]
    dur_ours = O.round_durations(log_d).long()
    dur_32 = O.round_durations(ld32).long()
    dur_64 = torch.clamp(torch.round(d64), min=0).long()
    print(f"chain B{Bb} L{L}: {1e3 * (t2 - t1):.2f} ms (first {1e3 * (t1 - t0):.1f}); enc rel ours {rel(e32[valid], e64[valid]):.2e} oracle32 {rel(e32o[valid], e64[valid]):.2e}; "
          f"log_d max abs ours {float((log_d - ld64)[valid].abs().max()):.2e} oracle32 {float((ld32 - ld64)[valid].abs().max()):.2e}; "
          f"flips ours-vs-32 {int((dur_ours != dur_32)[valid].sum())} ours-vs-64 {int((dur_ours != dur_64)[valid].sum())} 32-vs-64 {int((dur_32 != dur_64)[valid].sum())}; "
          f"min boundary dist {float(dist.min()):.2e}, #<1e-4: {int((dist < 1e-4).sum())}")
