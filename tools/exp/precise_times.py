"""Per-kernel GPU times of the fp32-accurate chain at cfg2 (B=64, L=128), CUDA events around graph replays."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib, synth, build_fastspeech2
from transformer_tts_b200.Models import _engine as E

dev = "cuda"
hp = synth.make_hp()
model = build_fastspeech2(hp).eval()
model.load_state_dict(synth.synth_state_dict(hp))
model = model.cuda()
B, L, D, H = 64, 128, 384, 4
text, pos, _ = synth.synth_text_batch(B, L, seed=1234)
text, pos = text.cuda(), pos.cuda()
m8 = E.mask_to_u8((pos != 0).unsqueeze(-2), B, L, text.device)
kv = lib.mask_lengths(m8)
lc = E.layer_cache(model.encoder.layers[0]); xc = E.layer_cache_x3(model.encoder.layers[0])

def timeit(name, fn, reps=10):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(5):
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / reps * 1e3)
    print(f"{name:40s} {min(ts):8.1f} us")
    return min(ts)

x = torch.randn(B, L, D, device=dev)
xs = lib.split_f16(x)
qkv = lib.gemm_x3(xs, xc["wqkv"], lc["bqkv"], xc["sqkv"], split_out=True)
ctx = lib.attention_x3(qkv, m8, kv, H)
hs = lib.gemm_x3(xs, xc["w1"], lc["b1"], xc["s1"], taps=5, pad_left=2, relu=True, split_out=True)
u = lib.gemm_x3(hs, xc["w2"], lc["b2"], xc["s2"], taps=5, pad_left=2)
y = torch.empty_like(x); yns = torch.empty_like(xs); z = torch.empty_like(x)
tot = 0
tot += timeit("gemm_x3 qkv (K=384,N=1152)", lambda: lib.gemm_x3(xs, xc["wqkv"], lc["bqkv"], xc["sqkv"], split_out=True))
tot += timeit("attention_x3", lambda: lib.attention_x3(qkv, m8, kv, H))
tot += timeit("gemm_x3 out (K=384,N=384)", lambda: lib.gemm_x3(ctx, xc["wo"], lc["bo"], xc["so"]))
tot += timeit("xrow ADD_LN", lambda: lib.xrow(lib.XROW_ADD_LN, x, lc["ln2"], b0=u, s_out=y, out_split=yns))
tot += timeit("gemm_x3 ffn1 k5 (K=1920,N=1536) split", lambda: lib.gemm_x3(xs, xc["w1"], lc["b1"], xc["s1"], taps=5, pad_left=2, relu=True, split_out=True))
tot += timeit("gemm_x3 ffn2 k5 (K=7680,N=384)", lambda: lib.gemm_x3(hs, xc["w2"], lc["b2"], xc["s2"], taps=5, pad_left=2))
tot += timeit("xrow FFN_TAIL", lambda: lib.xrow(lib.XROW_FFN_TAIL, u, lc["ln_ff"], b0_split=yns, c=y, ln2=lc["ln1"], z_out=z, out_split=yns))
print(f"one layer: {tot:.1f} us, x6 = {6 * tot / 1e3:.2f} ms")
def chain():
    e_bf, e_s, _ = E.run_encoder_precise(model.encoder, text, m8, kv)
    return E.run_predictor_precise(model.variance_adaptor.duration_predictor, e_s, m8)
with torch.no_grad():
    timeit("whole precise chain (graph)", chain, reps=3)
    def bf():
        e_bf, _, _ = model.encoder.run(text, m8, kv)
        return model.variance_adaptor.duration_predictor.run(e_bf, m8)
    timeit("whole bf16 chain (graph)", bf, reps=3)

import ctypes
L_ = lib.load()
for flags, name in ((2, "no MMAs"), (1, "no TMEM read-back"), (3, "no MMA, no read-back")):
    L_.ttsb200_debug_x3(flags)
    print("---", name)
    timeit("gemm_x3 ffn1 k5", lambda: lib.gemm_x3(xs, xc["w1"], lc["b1"], xc["s1"], taps=5, pad_left=2, relu=True, split_out=True))
    timeit("gemm_x3 ffn2 k5", lambda: lib.gemm_x3(hs, xc["w2"], lc["b2"], xc["s2"], taps=5, pad_left=2))
L_.ttsb200_debug_x3(0)
