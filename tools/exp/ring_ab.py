"""A/B of the LayerNorm-epilogue GEMMs (out-proj RES_LN, FFN2) at the cfg2 decoder shape; run once per build variant."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib, synth, build_fastspeech2
from transformer_tts_b200.Models import _engine as E
dev = "cuda"
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
lc = E.layer_cache(model.decoder.layers[0]); D = 384
B, T = 64, 1103
g = torch.Generator().manual_seed(7)
lens = torch.randint(int(0.8 * T), T + 1, (B,), generator=g); lens[0] = T
mask = (torch.arange(T)[None] < lens[:, None]).to(torch.uint8).to(dev)
kv = lib.mask_lengths(mask); tmap = lib.build_tile_map(kv, T)
NS = 3
xs = [(torch.randn(B, T, D, device=dev) * 0.5).bfloat16() for _ in range(NS)]
x32 = [torch.randn(B, T, D, device=dev) for _ in range(NS)]
hs = [torch.relu(torch.randn(B, T, 4 * D, device=dev)).bfloat16() for _ in range(NS)]
ob = [torch.empty(B, T, D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
o32 = [torch.empty(B, T, D, device=dev) for _ in range(NS)]
def t_out(i): lib.conv_gemm(xs[i], lc["wo"], lc["bo"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x32[i], res_bstride=T * D, res_tstride=D, ln1=lc["ln2"], out_f32=o32[i], out_bf16=ob[i], tile_map=tmap)
def t_ffn2(i): lib.conv_gemm(hs[i], lc["w2"], lc["b2"], block_n=D, epi=lib.EPI_FFN2, res_bf16=xs[i], res_f32=x32[i], res_bstride=T * D, res_tstride=D, ln1=lc["ln_ff"], ln2=lc["ln1"], out_f32=o32[i], out_bf16=ob[i], tile_map=tmap)
for name, fn in (("out_proj", t_out), ("ffn2", t_ffn2)):
    for i in range(NS): fn(i)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(4):
            for i in range(NS): fn(i)
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(5):
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) / 12 * 1e3)
    print(f"{name}: {min(ts):.1f} us")
