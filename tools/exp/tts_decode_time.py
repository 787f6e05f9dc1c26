"""Time Transformer.generate at BASELINE.json configs[4] (batch 16, 128 phonemes, 1000 positions) and per-kernel shares."""
import sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import build_transformer_tts, synth, lib

hp = synth.make_tts_hp()
model = build_transformer_tts(hp).eval(); model.load_state_dict(synth.synth_tts_state_dict(hp)); model = model.cuda()
for B in (16, 1, 64):
    text, pos, _ = synth.synth_text_batch(B, 128, ragged=True, seed=77, min_len=64)
    text, mask = text.cuda(), (pos != 0).unsqueeze(-2).cuda()
    n = 1000
    model.generate(text, mask, max_steps=50)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(3):
        e0.record(); out = model.generate(text, mask, max_steps=n); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts)
    print(f"B={B}: {ms:.1f} ms for {n} positions = {ms / n * 1e3:.1f} us per position; {B * n * hp.reduction_rate / ms * 1e3:.0f} frames/s; "
          f"launches per position {model.last_state.launches_per_step}")

# ---- per-kernel times at B=16 (events around 200 launches each) ----
B = 16
text, pos, _ = synth.synth_text_batch(B, 128, ragged=True, seed=77, min_len=64)
out = model.generate(text.cuda(), (pos != 0).unsqueeze(-2).cuda(), max_steps=1000)
st = model.last_state
t = st.t; lc = t["layers"][0]; D, H, Dh, cap, R = st.D, st.H, st.Dh, st.cap, st.R
sb, sh, sj = H * cap * Dh, cap * Dh, Dh
def timeit(name, fn, n=200):
    """GPU time per launch inside a CUDA graph of n back-to-back launches (warm caches, no host in the loop)."""
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f"  {name}: {e0.elapsed_time(e1) / n * 1e3:.2f} us")
for p in (500,):
    st.pos.fill_(p)
    print(f"pos {p}")
    timeit("LN1 + qkv (384 -> 1152)", lambda: lib.small_linear(lc["wqkv1"], lc["bqkv1"], D, x_f32=st.x, ln=lc["ln1"], out_bf16=st.qkv))
    timeit("qkv from fp32 rows, no LN", lambda: lib.small_linear(lc["wqkv1"], lc["bqkv1"], D, x_f32=st.x, out_bf16=st.qkv))
    timeit("qkv from bf16 rows", lambda: lib.small_linear(lc["wqkv1"], lc["bqkv1"], D, x_bf16=st.xn, out_bf16=st.qkv))
    timeit("LN + 384 -> 384", lambda: lib.small_linear(lc["wq2"], lc["bq2"], D, x_f32=st.x, ln=lc["ln2"], out_bf16=st.q2))
    timeit("self attention", lambda: lib.attn_decode(st.qkv, 0, 3 * D, st.kv_self[0], R * sb, (sb, sh, sj), cap, H, Dh, st.ctx, kv_new=st.qkv, knew_off=D, vnew_off=2 * D, pos=st.pos))
    timeit("out-proj + res (384 -> 384)", lambda: lib.small_linear(lc["wo1"], lc["bo1"], D, x_bf16=st.ctx, res_f32=st.x, out_f32=st.x2))
    timeit("cross attention (128 keys)", lambda: lib.attn_decode(st.q2, 0, D, st.kv_cross[0], D, (st.L * 2 * D, Dh, 2 * D), st.L, H, Dh, st.ctx, kv_len=st.src_len))
    timeit("LN3 + FFN1 (384 -> 1536)", lambda: lib.small_linear(lc["w1"], lc["b1"], D, x_f32=st.x, ln=lc["ln3"], relu=True, out_bf16=st.hid, xn_bf16=st.xn))
    timeit("FFN2 (1536 -> 384)", lambda: lib.small_linear(lc["w2"], lc["b2"], 4 * D, x_bf16=st.hid, out_f32=st.u))
    timeit("ffn2_tail", lambda: lib.ffn2_tail(st.u, st.xn, st.x, lc["ln_ff"], t["ident"], 1e-5, st.x2, st.zn))
    nl = t["layers"][1]["ln1"]
    timeit("cluster: out-proj + LN2 + q2", lambda: lib.decode_attn_out_q(st.ctx, st.x, lc["wo1"], lc["bo1"], lc["ln2"], lc["wq2"], lc["bq2"], 1e-5, st.x2, st.q2))
    timeit("cluster: out-proj + LN3 + FFN + tail", lambda: lib.decode_ffn_block(st.ctx, st.x2, lc["wo2"], lc["bo2"], lc["ln3"], lc["w1"], lc["b1"], lc["w2"], lc["b2"], lc["ln_ff"], nl, 1e-5, st.x, st.zn))
    timeit("advance_pos(0)", lambda: lib.advance_pos(st.pos, 0))

import ctypes
L_ = lib.load()
L_.ttsb200_debug_dcl_trace(None, 1)
for _ in range(3):
    lib.decode_ffn_block(st.ctx, st.x2, lc["wo2"], lc["bo2"], lc["ln3"], lc["w1"], lc["b1"], lc["w2"], lc["b2"], lc["ln_ff"], nl, 1e-5, st.x, st.zn)
    torch.cuda.synchronize()
buf = (ctypes.c_longlong * 16)()
L_.ttsb200_debug_dcl_trace(buf, 2); L_.ttsb200_debug_dcl_trace(None, 0)
names = ["start", "prefetch issued", "dependency wait", "ctx landed", "out-proj + push", "barrier 1", "LayerNorm 3", "FFN1 mma", "hidden slice", "FFN2 first half + scatter", "W2 second half landed", "FFN2 second half + scatter", "barrier 2", "tail"]
print("ffn_block timeline (clk, isolated launch): " + ", ".join(f"{n} {buf[i] - (buf[i-1] if i else 0)}" for i, n in enumerate(names)))
