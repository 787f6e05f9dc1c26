"""A/B of the CTA-pair (cta_group::2) AFFINE GEMM kernels against the single-CTA ones: same outputs, time per launch."""
import ctypes as C, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
lib = L.load()
lib.ttsb200_debug_gemm_pair.argtypes = [C.c_int]
lib.ttsb200_debug_gemm_pair.restype = None
dev = "cuda"
torch.manual_seed(0)
def run(name, B, T, C_, N, taps, bn, act, pad=0):
    x = (torch.randn(B, T, C_, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn(taps, N, C_, device=dev) * 0.05).to(torch.bfloat16)
    b = torch.randn(N, device=dev) * 0.1
    outs, times = [], []
    for pair in (0, 2):
        lib.ttsb200_debug_gemm_pair(pair)
        o = torch.full((B, T, N), 7.0, device=dev, dtype=torch.bfloat16)
        fn = lambda: L.conv_gemm(x, w, b, taps=taps, pad_left=pad, block_n=bn, act=act, out_bf16=o)
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): fn()
        e1.record(); torch.cuda.synchronize()
        outs.append(o.float().clone()); times.append(e0.elapsed_time(e1) / 20 * 1e3)
    ref = torch.zeros(B, T, N, device=dev)
    xp = torch.nn.functional.pad(x.float(), (0, 0, pad, taps - 1 - pad))
    for j in range(taps):
        ref += xp[:, j:j + T] @ w[j].float().t()
    ref += b
    if act == L.ACT_RELU: ref = torch.relu(ref)
    if act == L.ACT_TANH: ref = torch.tanh(ref)
    err = lambda o: ((o - ref).norm() / ref.norm()).item()
    print(f"{name:34s} single {times[0]:7.1f} us  pair {times[1]:7.1f} us  | max|pair-single| {(outs[0]-outs[1]).abs().max().item():.3g}  "
          f"rel-L2 vs fp32: single {err(outs[0]):.2e} pair {err(outs[1]):.2e}", flush=True)
run("tiny odd tiles B=3 T=200 qkv", 3, 200, 384, 1152, 1, 192, L.ACT_NONE)
run("tiny B=1 T=129 ffn1", 1, 129, 384, 1536, 1, 256, L.ACT_RELU)
run("tiny conv k5 B=2 T=300", 2, 300, 512, 512, 5, 256, L.ACT_TANH, pad=2)
run("qkv  B=64 T=1103 <192,A-stat>", 64, 1103, 384, 1152, 1, 192, L.ACT_NONE)
run("ffn1 B=64 T=1103 <256,A-stat>", 64, 1103, 384, 1536, 1, 256, L.ACT_RELU)
run("postnet k5 512->512 <256>", 64, 1103, 512, 512, 5, 256, L.ACT_TANH, pad=2)
run("postnet k5 80->512 <256>", 64, 1103, 80, 512, 5, 256, L.ACT_TANH, pad=2)
run("enc qkv B=64 T=146 <192,A-stat>", 64, 146, 384, 1152, 1, 192, L.ACT_NONE)
run("ffn1 k9? conv 384->1536 k9 <256>", 64, 1103, 384, 1536, 9, 256, L.ACT_RELU, pad=4)
