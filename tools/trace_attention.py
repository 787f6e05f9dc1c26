import ctypes, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
lib = L.load()
B, T, H, Dh = 64, 1103, 4, 96
g = torch.Generator().manual_seed(1)
lens = torch.randint(892, 1104, (B,), generator=g); lens[0] = T
qkv = (torch.randn(B, T, 3 * H * Dh) * 0.5).to(torch.bfloat16).cuda()
mask = (torch.arange(T)[None, :] < lens[:, None]).to(torch.uint8).cuda()
kv = L.mask_lengths(mask)
ctx = torch.empty(B, T, H * Dh, dtype=torch.bfloat16, device='cuda')
for _ in range(3): L.attention_fwd(qkv, mask, kv, ctx, H)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 512)()
lib.ttsb200_attention_trace.argtypes = [ctypes.c_void_p]; lib.ttsb200_attention_trace.restype = None
lib.ttsb200_attention_trace(buf)   # allocate
L.attention_fwd(qkv, mask, kv, ctx, H)
lib.ttsb200_attention_trace(buf)   # read back
t = list(buf)
t0 = t[0]
print('prologue sync', t[1]-t0, 'final OF', t[2]-t0, 'end', t[3]-t0)
names = {0:'mma KF', 1:'mma SE', 2:'mma PF', 3:'mma VF', 8:'sm top', 9:'sm SF', 10:'sm SE-arr', 11:'sm preOF', 12:'sm OF', 13:'sm PF-arr'}
for j in range(10):
    row = []
    for k in (8, 0, 1, 9, 10, 11, 12, 13, 2, 3):
        v = t[16 + j*16 + k]
        row.append(f"{names[k]}={v-t0 if v else -1}")
    print(j, ' '.join(row))
