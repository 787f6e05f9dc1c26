"""clock64 accounting of the gemm_x3 issuer / epilogue waits (debug flag 8)."""
import sys, ctypes
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import lib, synth, build_fastspeech2
from transformer_tts_b200.Models import _engine as E
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
B, L, D = 64, 128, 384
lc = E.layer_cache(model.encoder.layers[0]); xc = E.layer_cache_x3(model.encoder.layers[0])
x = torch.randn(B, L, D, device="cuda"); xs = lib.split_f16(x)
hs = lib.gemm_x3(xs, xc["w1"], lc["b1"], xc["s1"], taps=5, pad_left=2, relu=True, split_out=True)
L_ = lib.load()
buf = (ctypes.c_longlong * (148 * 16))()
for name, fn in (("ffn1", lambda: lib.gemm_x3(xs, xc["w1"], lc["b1"], xc["s1"], taps=5, pad_left=2, relu=True, split_out=True)),
                 ("ffn2", lambda: lib.gemm_x3(hs, xc["w2"], lc["b2"], xc["s2"], taps=5, pad_left=2)),
                 ("qkv", lambda: lib.gemm_x3(xs, xc["wqkv"], lc["bqkv"], xc["sqkv"]))):
    for flags in (8, 8 | 2, 8 | 1 | 2):
        L_.ttsb200_debug_x3(flags)
        fn(); torch.cuda.synchronize()
        L_.ttsb200_debug_x3_trace(buf, 0, 1)
        fn(); torch.cuda.synchronize()
        L_.ttsb200_debug_x3_trace(buf, 148 * 16, 0)
        for pair in (0, 37):
            o = buf[pair * 16: pair * 16 + 16]
            n = max(o[5], 1)
            print(f"{name} flags={flags} pair {pair}: total {o[0]} clk, k-iters {o[5]} ({o[0] / n:.0f} clk/iter): wait A {o[1] / n:.0f} acce {o[2] / n:.0f} wfull {o[3] / n:.0f} issue {o[4] / n:.0f} | epi wait accf {o[8] / n:.0f}")
L_.ttsb200_debug_x3(0)
