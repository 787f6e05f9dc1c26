"""ffn_fused against the two-kernel path and an fp64 reference; timing at the decoder shape of cfg2."""
import sys
from pathlib import Path
import torch
import torch.nn.functional as F
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib, synth, build_fastspeech2
from transformer_tts_b200.Models import _engine as E

torch.manual_seed(0)
dev = "cuda"
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
lc = E.layer_cache(model.decoder.layers[0])
ln_next = E.layer_cache(model.decoder.layers[1])["ln1"]
D = 384

def rel(a, b): return float((a.double() - b.double()).norm() / b.double().norm())

for (B, T) in ((3, 300), (5, 128), (64, 1103)):
    lens = torch.randint(int(0.7 * T), T + 1, (B,)); lens[0] = T
    mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).to(dev)
    kv = lib.mask_lengths(mask)
    x = (torch.randn(B, T, D, device=dev) * 0.7).bfloat16()
    y = torch.randn(B, T, D, device=dev)
    for tmap in (None, lib.build_tile_map(kv, T)):
        z, zn = torch.zeros(B, T, D, device=dev), torch.zeros(B, T, D, device=dev, dtype=torch.bfloat16)
        lib.ffn_fused(x, lc["w1"], lc["b1"], lc["w2"], lc["b2"], y, lc["ln_ff"], ln_next, z, zn, tile_map=tmap)
        torch.cuda.synchronize()
        # two-kernel path
        h = torch.empty(B, T, 4 * D, device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(x, lc["w1"], lc["b1"], block_n=256, act=lib.ACT_RELU, out_bf16=h)
        z2, zn2 = torch.empty_like(z), torch.empty_like(zn)
        lib.conv_gemm(h, lc["w2"], lc["b2"], block_n=D, epi=lib.EPI_FFN2, res_bf16=x, res_f32=y, res_bstride=T * D, res_tstride=D,
                      ln1=lc["ln_ff"], ln2=ln_next, out_f32=z2, out_bf16=zn2)
        # fp64 reference (h rounded to bf16 like both kernels do)
        xd = x.double()
        hd = torch.relu(xd @ lc["w1"][0].double().T + lc["b1"].double()).bfloat16().double()
        u = hd @ lc["w2"][0].double().T + lc["b2"].double() + xd
        zr = y.double() + F.layer_norm(u, (D,), lc["ln_ff"][0].double(), lc["ln_ff"][1].double(), 1e-5)
        znr = F.layer_norm(zr, (D,), ln_next[0].double(), ln_next[1].double(), 1e-5)
        v = mask.bool() if tmap is not None else torch.ones_like(mask).bool()
        print(f"B{B} T{T} tile_map={tmap is not None}: z vs fp64 {rel(z[v], zr[v]):.2e} (two-kernel {rel(z2[v], zr[v]):.2e}); "
              f"zn vs fp64 {rel(zn[v], znr[v]):.2e} (two-kernel {rel(zn2[v], znr[v]):.2e}); fused==two-kernel z maxabs {float((z[v] - z2[v]).abs().max()):.2e}")

B, T = 64, 1103
lens = torch.randint(int(0.8 * T), T + 1, (B,)); lens[0] = T
mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).to(dev)
kv = lib.mask_lengths(mask); tmap = lib.build_tile_map(kv, T)
NS = 3
xs = [(torch.randn(B, T, D, device=dev) * 0.7).bfloat16() for _ in range(NS)]
ys = [torch.randn(B, T, D, device=dev) for _ in range(NS)]
zs = [torch.empty(B, T, D, device=dev) for _ in range(NS)]
zns = [torch.empty(B, T, D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
hs = [torch.empty(B, T, 4 * D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
def fused(i): lib.ffn_fused(xs[i], lc["w1"], lc["b1"], lc["w2"], lc["b2"], ys[i], lc["ln_ff"], ln_next, zs[i], zns[i], tile_map=tmap)
def two(i):
    lib.conv_gemm(xs[i], lc["w1"], lc["b1"], block_n=256, act=lib.ACT_RELU, out_bf16=hs[i], tile_map=tmap)
    lib.conv_gemm(hs[i], lc["w2"], lc["b2"], block_n=D, epi=lib.EPI_FFN2, res_bf16=xs[i], res_f32=ys[i], res_bstride=T * D, res_tstride=D,
                  ln1=lc["ln_ff"], ln2=ln_next, out_f32=zs[i], out_bf16=zns[i], tile_map=tmap)
for name, fn in (("fused", fused), ("two-kernel", two)):
    for i in range(NS): fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(4):
            for i in range(NS): fn(i)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(5):
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) / 12 * 1e3)
    rows = float(lens.sum())
    print(f"{name}: {min(ts):.1f} us per layer  ({4 * rows * 384 * 1536 / (min(ts) * 1e-6) / 1e12:.0f} TFLOP/s on valid rows)")

import ctypes
L_ = lib.load()
buf = (ctypes.c_longlong * (148 * 16))()
L_.ttsb200_debug_ffn_trace(buf, 0, 1)
fused(0); torch.cuda.synchronize()
L_.ttsb200_debug_ffn_trace(buf, 0, 1)
fused(0); torch.cuda.synchronize()
L_.ttsb200_debug_ffn_trace(buf, 148 * 16, 2)
L_.ttsb200_debug_ffn_trace(buf, 0, 0)
for pair in (0, 20, 73):
    o = buf[pair * 16: pair * 16 + 16]
    n = max(o[6], 1)
    print(f"pair {pair}: issuer total {o[0]} clk over {o[6]} units ({o[0] / n:.0f}/unit): wait X {o[1] / n:.0f} W1 {o[2] / n:.0f} W2 {o[3] / n:.0f} H {o[4] / n:.0f} A2E {o[5] / n:.0f} | LN epilogue {o[8] / n:.0f} clk/unit = pass1 {o[9] / n:.0f} + pass2 {o[10] / n:.0f} (ring wait {o[12] / n:.0f}) + pass3 {o[11] / n:.0f}")
