"""K=64 GEMM with the FFN1 output shape: the MMA work is 1/6, so the time is the epilogue + store path alone."""
import ctypes as C, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
lib = L.load()
lib.ttsb200_debug_gemm_pair.argtypes = [C.c_int]; lib.ttsb200_debug_gemm_pair.restype = None
dev = "cuda"
B, T = 64, 1103
def t(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20 * 1e3
for K in (64, 128, 384):
    x = (torch.randn(B, T, K, device=dev) * 0.5).to(torch.bfloat16)
    for N, bn in ((1536, 256), (1152, 192), (384, 192)):
        w = (torch.randn(1, N, K, device=dev) * 0.05).to(torch.bfloat16)
        bias = torch.zeros(N, device=dev)
        o = torch.empty(B, T, N, device=dev, dtype=torch.bfloat16)
        r = []
        for pair in (0, 1):
            lib.ttsb200_debug_gemm_pair(pair)
            r.append(t(lambda: L.conv_gemm(x, w, bias, block_n=bn, out_bf16=o)))
        mb = (x.numel() + o.numel()) * 2 / 1e6
        print(f"K={K:3d} N={N:4d}: single {r[0]:6.1f} us pair {r[1]:6.1f} us | {mb:6.1f} MB -> {mb/r[1]*1e-3:5.2f} TB/s (pair)", flush=True)
# plain device copy of the same output size for scale
o = torch.empty(B, T, 1536, device=dev, dtype=torch.bfloat16); o2 = torch.empty_like(o)
print(f"copy_ 217 MB -> 217 MB: {t(lambda: o2.copy_(o)):6.1f} us;  fill_ 217 MB: {t(lambda: o.fill_(1.0)):6.1f} us")
