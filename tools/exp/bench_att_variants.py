import ctypes as C, sys, torch
sys.path.insert(0, '.')
B, T, H, Dh = 64, 1103, 4, 96
g = torch.Generator().manual_seed(1)
lens = torch.randint(892, 1104, (B,), generator=g); lens[0] = T
qkv = (torch.randn(B, T, 3 * H * Dh) * 0.5).to(torch.bfloat16).cuda()
mask = (torch.arange(T)[None, :] < lens[:, None]).to(torch.uint8).cuda()
from transformer_tts_b200 import lib as L
kv = L.mask_lengths(mask)
ctx = torch.empty(B, T, H * Dh, dtype=torch.bfloat16, device='cuda')
fl = float((4 * lens.double() ** 2 * H * Dh).sum())
for v in sys.argv[1:]:
    lib = C.CDLL(f"tools/exp/libatt_{v}.so")
    fn = lib.ttsb200_attention_fwd
    fn.argtypes = [C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p]
    st = torch.cuda.current_stream().cuda_stream
    run = lambda: fn(qkv.data_ptr(), mask.data_ptr(), kv.data_ptr(), ctx.data_ptr(), B, T, H, Dh, st)
    for _ in range(3): assert run() == 0
    n = 20
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    torch.cuda.synchronize(); ev[0].record()
    for i in range(n):
        run(); ev[i + 1].record()
    torch.cuda.synchronize()
    ms = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(n))[n // 2]
    c = ctx.float().clone()
    if 'ref' not in globals(): ref = c
    err = float((c - ref).norm() / ref.norm())
    print(f"{v}: {ms*1e3:.1f} us  {fl/ms/1e9:.0f} TFLOP/s  rel-L2 vs first variant {err:.2e}")
