"""Back-to-back launch times (CUDA graph, GPU-bound) of the encoder-shaped GEMMs: B=64 utterances, T phonemes."""
import ctypes as C, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
lib = L.load()
lib.ttsb200_debug_gemm_pair.argtypes = [C.c_int]; lib.ttsb200_debug_gemm_pair.restype = None
dev = "cuda"
B, T, D = 64, int(sys.argv[1]) if len(sys.argv) > 1 else 100, 384
xs = (torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16)
hs = torch.relu(torch.randn(B, T, 4 * D, device=dev)).to(torch.bfloat16)
x32 = torch.randn(B, T, D, device=dev)
w1 = (torch.randn(9, 4 * D, D, device=dev) * 0.05).to(torch.bfloat16)
wq = (torch.randn(1, 3 * D, D, device=dev) * 0.05).to(torch.bfloat16)
w2 = (torch.randn(1, D, 4 * D, device=dev) * 0.05).to(torch.bfloat16)
wo = (torch.randn(1, D, D, device=dev) * 0.05).to(torch.bfloat16)
b1, bq, b2 = torch.zeros(4 * D, device=dev), torch.zeros(3 * D, device=dev), torch.zeros(D, device=dev)
g = (torch.ones(D, device=dev), torch.zeros(D, device=dev))
h = torch.empty(B, T, 4 * D, device=dev, dtype=torch.bfloat16)
qkv = torch.empty(B, T, 3 * D, device=dev, dtype=torch.bfloat16)
ob = torch.empty(B, T, D, device=dev, dtype=torch.bfloat16)
o32 = torch.empty(B, T, D, device=dev)
def t(fn, n=50):
    g_ = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(g_, stream=s):
            for _ in range(n): fn()
    torch.cuda.synchronize()
    g_.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g_.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
lib.ttsb200_debug_gemm_astat_min.argtypes = [C.c_int]; lib.ttsb200_debug_gemm_astat_min.restype = None
for pair, amin in ((1, 0), (1, 148), (2, 148)):
    lib.ttsb200_debug_gemm_pair(pair); lib.ttsb200_debug_gemm_astat_min(amin)
    print(f"astat_min={amin}", end=" ")
    print(f"T={T} pair={pair}: qkv {t(lambda: L.conv_gemm(xs, wq, bq, block_n=192, out_bf16=qkv)):6.1f}  "
          f"out-proj {t(lambda: L.conv_gemm(xs, wo, b2, block_n=D, epi=L.EPI_RES_LN, res_f32=x32, res_bstride=T * D, res_tstride=D, ln1=g, out_f32=o32, out_bf16=ob)):6.1f}  "
          f"ffn1 k9 {t(lambda: L.conv_gemm(xs, w1, b1, taps=9, pad_left=4, block_n=256, act=L.ACT_RELU, out_bf16=h)):6.1f}  "
          f"ffn2 {t(lambda: L.conv_gemm(hs, w2, b2, block_n=D, epi=L.EPI_FFN2, res_bf16=xs, res_f32=x32, res_bstride=T * D, res_tstride=D, ln1=g, ln2=g, out_f32=o32, out_bf16=ob)):6.1f} us", flush=True)
# out-proj + residual + LayerNorm: whole-row epilogue on B*ceil(T/128) CTAs vs column-split AFFINE GEMM + row kernel
lib.ttsb200_debug_gemm_pair(1)
def split():
    L.conv_gemm(xs, wo, b2, block_n=192, res_f32=x32, res_bstride=T * D, res_tstride=D, out_f32=o32)
    L.layernorm(o32, g[0], g[1], 1e-5, y_bf16=ob)
def split128():
    L.conv_gemm(xs, wo, b2, block_n=128, res_f32=x32, res_bstride=T * D, res_tstride=D, out_f32=o32)
    L.layernorm(o32, g[0], g[1], 1e-5, y_bf16=ob)
fused = lambda: L.conv_gemm(xs, wo, b2, block_n=D, epi=L.EPI_RES_LN, res_f32=x32, res_bstride=T * D, res_tstride=D, ln1=g, out_f32=o32, out_bf16=ob)
print(f"out-proj fused {t(fused):.1f} us, split(192)+LN {t(split):.1f} us, split(128)+LN {t(split128):.1f} us")
