"""q|k|v projection (N = 1152, K = 384) at the decoder shape of cfg2: tile width variants of conv_gemm."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib

dev = "cuda"
B, T, D = 64, 1103, 384
torch.manual_seed(0)
lens = torch.randint(int(0.8 * T), T + 1, (B,)); lens[0] = T
mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).to(dev)
kv = lib.mask_lengths(mask); tmap = lib.build_tile_map(kv, T)
NS = 3
xs = [(torch.randn(B, T, D, device=dev) * 0.7).bfloat16() for _ in range(NS)]
w = (torch.randn(1, 3 * D, D, device=dev) * D ** -0.5).bfloat16()
b = torch.randn(3 * D, device=dev)
outs = [torch.empty(B, T, 3 * D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
ref = None
for bn in (128, 192, 256, 384):
    try:
        def fn(i): lib.conv_gemm(xs[i], w, b, block_n=bn, out_bf16=outs[i], tile_map=tmap)
        for i in range(NS): fn(i)
        torch.cuda.synchronize()
        if ref is None: ref = outs[0].clone()
        same = torch.equal(ref[mask.bool()], outs[0][mask.bool()])
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
        print(f"block_n {bn}: {min(ts):.1f} us  ({2 * rows * 1152 * 384 / (min(ts) * 1e-6) / 1e12:.0f} TFLOP/s on valid rows) identical={same}")
    except Exception as e:
        print(f"block_n {bn}: {str(e)[:150]}")
