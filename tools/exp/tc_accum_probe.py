"""How does tcgen05 accumulate?  Inputs exactly representable in bf16 -> products exact; any error of the fp32 TMEM
result against the fp64 sum is the tensor core's own accumulation (rounding mode, guard bits).
Prints mean signed / rms relative error for K = 64 .. 7680, signed and all-positive operands."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib

torch.manual_seed(0)
dev = "cuda"
for positive in (False, True):
    for taps, C in ((1, 64), (1, 384), (1, 1536), (5, 1536)):
        B, T, N = 8, 128, 384
        x = torch.rand(B, T + taps - 1, C, device=dev) if positive else torch.randn(B, T + taps - 1, C, device=dev)
        w = torch.rand(taps, N, C, device=dev) if positive else torch.randn(taps, N, C, device=dev)
        xb, wb = x.bfloat16().contiguous(), w.bfloat16().contiguous()
        bias = torch.zeros(N, device=dev)
        out = torch.empty(B, T + taps - 1, N, device=dev)
        lib.conv_gemm(xb, wb, bias, taps=taps, pad_left=0, block_n=128, out_f32=out)
        xd, wd = xb.double(), wb.double()
        ref = sum(xd[:, j:j + T] @ wd[j].T for j in range(taps))
        got = out[:, :T].double()
        f32 = sum(xb.float()[:, j:j + T] @ wb.float()[j].T for j in range(taps)).double()  # cuBLAS fp32 (no tf32)
        scale = ref.abs().mean()
        e = (got - ref) / scale
        e32 = (f32 - ref) / scale
        sgn = ((got - ref) * ref.sign()).mean() / scale
        print(f"positive={positive} K={taps * C:5d}: tcgen05 rms {e.pow(2).mean().sqrt():.3e} max {e.abs().max():.3e} "
              f"signed-toward-growth {sgn:+.3e} | cuBLAS fp32 rms {e32.pow(2).mean().sqrt():.3e}")
