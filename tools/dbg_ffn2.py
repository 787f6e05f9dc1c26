import math, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L
torch.manual_seed(0)
dev = 'cuda'
B, T, C, N = 40, 500, 1536, 384
taps, pad = 1, 0
h = torch.relu(torch.randn(B, T, C)).to(torch.bfloat16).to(dev)
w = (torch.randn(taps, N, C) / math.sqrt(C)).to(torch.bfloat16).to(dev)
bias = torch.randn(N, device=dev)
g1, b1, g2, b2 = [torch.randn(N, device=dev) * 0.1 + (1 if i % 2 == 0 else 0) for i in range(4)]
yn = torch.randn(B, T, N).to(torch.bfloat16).to(dev)
y = torch.randn(B, T, N, device=dev)
z = torch.empty(B, T, N, device=dev)
zn = torch.empty(B, T, N, device=dev, dtype=torch.bfloat16)
L.conv_gemm(h, w, bias, taps=taps, pad_left=pad, block_n=384, epi=L.EPI_FFN2, res_bf16=yn, res_f32=y,
            res_bstride=T * N, res_tstride=N, ln1=(g1, b1), ln2=(g2, b2), out_f32=z, out_bf16=zn)
torch.cuda.synchronize()
v = (h.double() @ w[0].double().T) + bias.double() + yn.double()
def ln(v, g, b):
    m = v.mean(-1, keepdim=True); var = ((v - m) ** 2).mean(-1, keepdim=True)
    return (v - m) / torch.sqrt(var + 1e-5) * g.double() + b.double()
zr = y.double() + ln(v, g1, b1)
znr = ln(zr, g2, b2)
err = (z.double() - zr).abs()
errn = (zn.double() - znr).abs()
bad = []
for b in range(B):
    for mt in range(4):
        for c in range(12):
            e = err[b, mt*128:(mt+1)*128, c*32:(c+1)*32].max().item()
            en = errn[b, mt*128:(mt+1)*128, c*32:(c+1)*32].max().item()
            if e > 1e-3 or en > 0.1:
                bad.append((b, mt, b*4+mt, c, round(e, 3), round(en, 3)))
print(len(bad), 'bad blocks of', B*4*12)
for x in bad[:80]: print(x)
