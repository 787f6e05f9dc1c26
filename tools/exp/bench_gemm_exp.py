"""Timing experiments on the AFFINE GEMM (results are wrong by construction): which smem consumer bounds it."""
import ctypes as C, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L, _build
tag = sys.argv[1]
if tag != "base":
    _build.LIB_PATH = type(_build.LIB_PATH)(f"tools/exp/libgemm_{tag}.so")
    _build.is_stale = lambda: False
lib = L.load()
lib.ttsb200_debug_gemm_pair.argtypes = [C.c_int]; lib.ttsb200_debug_gemm_pair.restype = None
dev = "cuda"
B, T, D = 64, 1103, 384
x = (torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16)
w1 = (torch.randn(1, 4 * D, D, device=dev) * 0.05).to(torch.bfloat16)
wq = (torch.randn(1, 3 * D, D, device=dev) * 0.05).to(torch.bfloat16)
b1, bq = torch.zeros(4 * D, device=dev), torch.zeros(3 * D, device=dev)
h = torch.empty(B, T, 4 * D, device=dev, dtype=torch.bfloat16)
qkv = torch.empty(B, T, 3 * D, device=dev, dtype=torch.bfloat16)
def t(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20 * 1e3
for pair in (0, 1):
    lib.ttsb200_debug_gemm_pair(pair)
    a = t(lambda: L.conv_gemm(x, wq, bq, block_n=192, out_bf16=qkv))
    b = t(lambda: L.conv_gemm(x, w1, b1, block_n=256, act=L.ACT_RELU, out_bf16=h))
    print(f"{tag:5s} pair={pair}: qkv {a:6.1f} us  ffn1 {b:6.1f} us", flush=True)
