"""Decoder-shape FFN conv1 (+ReLU) and q|k|v GEMMs, back to back in a CUDA graph over rotating buffers."""
import sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
L.load()
dev = "cuda"
B, T, D, NS = 64, 1103, 384, 3
xs = [(torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16) for _ in range(NS)]
w1 = (torch.randn(1, 4 * D, D, device=dev) * 0.05).to(torch.bfloat16)
wq = (torch.randn(1, 3 * D, D, device=dev) * 0.05).to(torch.bfloat16)
b1, bq = torch.zeros(4 * D, device=dev), torch.zeros(3 * D, device=dev)
hs = [torch.empty(B, T, 4 * D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
qs = [torch.empty(B, T, 3 * D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
def t(fn, n=12, reps=5):
    g_ = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3): fn(i)
        torch.cuda.synchronize()
        with torch.cuda.graph(g_, stream=s):
            for i in range(n): fn(i % NS)
    torch.cuda.synchronize(); g_.replay(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g_.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / n * 1e3)
    return best
f1 = t(lambda i: L.conv_gemm(xs[i], w1, b1, block_n=256, act=L.ACT_RELU, out_bf16=hs[i]))
fq = t(lambda i: L.conv_gemm(xs[i], wq, bq, block_n=192, out_bf16=qs[i]))
fl = 2.0 * B * T * D * 4 * D
print(f"ffn1 {f1:.1f} us ({fl / f1 / 1e6:.0f} TFLOP/s)   qkv {fq:.1f} us ({fl * 0.75 / fq / 1e6:.0f} TFLOP/s)")
