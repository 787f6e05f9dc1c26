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
lib = C.CDLL("tools/exp/libatt_tr" + (sys.argv[2] if len(sys.argv) > 2 else "1") + ".so")
fn = lib.ttsb200_attention_fwd
fn.argtypes = [C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p]
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    assert fn(qkv.data_ptr(), mask.data_ptr(), kv.data_ptr(), ctx.data_ptr(), B, T, H, Dh, st) == 0
torch.cuda.synchronize()
buf = (C.c_longlong * 400)()
lib.ttsb200_debug_att_trace.argtypes = [C.c_void_p, C.c_int]
assert lib.ttsb200_debug_att_trace(buf, 400) == 0
ev = [(v >> 56, v & ((1 << 56) - 1)) for v in buf if v]
t0 = ev[0][1]
names = {9: "entry", 20: "synced", 21: "Q_landed", 10: "loop_top", 11: "PF_ok", 12: "VF_ok", 13: "PV_issued", 14: "KF_ok", 15: "SE_ok", 16: "S_issued", 0: "start", 1: "tile_begin", 2: "S_ready", 3: "S_loaded", 4: "max_done", 5: "OF_ok", 6: "exp_done", 7: "PF_arrived", 8: "end"}
prev = t0
for slot, t in ev[:int(sys.argv[1]) if len(sys.argv) > 1 else 10000]:
    print(f"{names[slot]:>11s} t={t - t0:7d} d={t - prev:6d}")
    prev = t
    if slot == 8: break
