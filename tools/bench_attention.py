"""Times the fused attention kernel alone at the decoder shape of the bench workload (B=64, T=1103, H=4, Dh=96)."""
import sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
B, T, H, Dh = 64, 1103, 4, 96
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
torch.manual_seed(0)
g = torch.Generator().manual_seed(1)
lens = torch.randint(892, 1104, (B,), generator=g); lens[0] = T
qkv = (torch.randn(B, T, 3 * H * Dh) * 0.5).to(torch.bfloat16).cuda()
mask = (torch.arange(T)[None, :] < lens[:, None]).to(torch.uint8).cuda()
kv = L.mask_lengths(mask)
ctx = torch.empty(B, T, H * Dh, dtype=torch.bfloat16, device='cuda')
for _ in range(3): L.attention_fwd(qkv, mask, kv, ctx, H)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
torch.cuda.synchronize(); ev[0].record()
for i in range(n):
    L.attention_fwd(qkv, mask, kv, ctx, H); ev[i + 1].record()
torch.cuda.synchronize()
ms = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(n))[n // 2]
fl = float((4 * lens.double() ** 2 * H * Dh).sum())
print(f"attention: {ms*1e3:.1f} us, valid {fl/1e9:.1f} GFLOP -> {fl/ms/1e9:.0f} TFLOP/s")
