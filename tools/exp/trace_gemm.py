"""MMA-warp wait accounting of conv_gemm (TTSB_GEMM_TRACE build): per CTA total clk, clk waiting for the resident A
tile, for a free accumulator (epilogue), for a filled smem stage (TMA)."""
import ctypes as C, sys, torch
sys.path.insert(0, '.')
from transformer_tts_b200 import lib as L, _build
from transformer_tts_b200.Models import _engine as E
_build.LIB_PATH = type(_build.LIB_PATH)("tools/exp/libgemm_tr.so")
_build.is_stale = lambda: False
L.load()
import os
B, T, D = 64, int(os.environ.get('TT', '1103')), 384
dev = "cuda"
xs = (torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16)
hs = torch.relu(torch.randn(B, T, 4 * D, device=dev)).to(torch.bfloat16)
x32 = torch.randn(B, T, D, device=dev)
w1 = (torch.randn(1, 4 * D, D, device=dev) * 0.05).to(torch.bfloat16)
wq = (torch.randn(1, 3 * D, D, device=dev) * 0.05).to(torch.bfloat16)
w2 = (torch.randn(1, D, 4 * D, device=dev) * 0.05).to(torch.bfloat16)
wo = (torch.randn(1, D, D, device=dev) * 0.05).to(torch.bfloat16)
w5 = (torch.randn(5, D, D, device=dev) * 0.05).to(torch.bfloat16)
b1, bq, b2 = torch.zeros(4 * D, device=dev), torch.zeros(3 * D, device=dev), torch.zeros(D, device=dev)
g = (torch.ones(D, device=dev), torch.zeros(D, device=dev))
h = torch.empty(B, T, 4 * D, device=dev, dtype=torch.bfloat16)
qkv = torch.empty(B, T, 3 * D, device=dev, dtype=torch.bfloat16)
ob = torch.empty(B, T, D, device=dev, dtype=torch.bfloat16)
o32 = torch.empty(B, T, D, device=dev)
lib = L.load()
lib.ttsb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_int]
def report(name, fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    buf = (C.c_longlong * (148 * 32))()
    assert lib.ttsb200_debug_gemm_trace(buf, 148 * 32) == 0
    rows = [buf[i * 32:(i + 1) * 32] for i in range(0, 148, 2)]  # leader CTAs (pair mode); even CTAs otherwise
    n = len(rows)
    tot = sum(r[0] for r in rows) / n
    tl = max(rows[0][5] // 2, 1)
    nc = max(rows[0][23], 1)
    print(f"{'':28s} per 32-col chunk (wg0 thread 0): tmem_ld {rows[0][16]/nc:5.0f} math {rows[0][17]/nc:5.0f} free-wait {rows[0][18]/nc:5.0f} "
          f"sts {rows[0][19]/nc:5.0f} fence {rows[0][20]/nc:5.0f} arrive {rows[0][21]/nc:5.0f}  chunks {nc}")
    if "LN" in name or "FFN2" in name:
        nt_ = max(rows[0][23], 1)
        print(f"{'':28s} LN epilogue per tile (wg0 t0): wait-acc {rows[0][16]/nt_:6.0f} pass1 {rows[0][17]/nt_:6.0f} xchg {rows[0][18]/nt_:6.0f} "
              f"pass2 {rows[0][19]/nt_:6.0f} last {rows[0][20]/nt_:6.0f} (ring waits {rows[0][21]/nt_:6.0f}) tiles {nt_}")
    print(f"{'':28s} epilogue wg0 per tile: pre {rows[0][8]/tl:6.0f} wait-acc {rows[0][9]/tl:6.0f} work {rows[0][10]/tl:6.0f} clk | "
          f"wg1: pre {rows[0][11]/tl:6.0f} wait-acc {rows[0][12]/tl:6.0f} work {rows[0][13]/tl:6.0f}; MMA total/tile {rows[0][0]/max(rows[0][5],1):6.0f}")
    print(f"{name:28s} {e0.elapsed_time(e1)*1e3:7.1f} us | MMA warp clk/CTA: total {tot:9.0f}  wait A-tile {sum(r[1] for r in rows)/n/tot:5.1%}  "
          f"wait epilogue {sum(r[2] for r in rows)/n/tot:5.1%}  wait TMA stage {sum(r[3] for r in rows)/n/tot:5.1%}  k-iters {rows[0][4]} tiles {rows[0][5]}")
report("qkv  <192,AFFINE,A-stat>", lambda: L.conv_gemm(xs, wq, bq, block_n=192, out_bf16=qkv))
report("ffn1 <256,AFFINE,A-stat>", lambda: L.conv_gemm(xs, w1, b1, block_n=256, act=L.ACT_RELU, out_bf16=h))
report("ffn2 <384,FFN2>", lambda: L.conv_gemm(hs, w2, b2, block_n=D, epi=L.EPI_FFN2, res_bf16=xs, res_f32=x32, res_bstride=T * D, res_tstride=D, ln1=g, ln2=g, out_f32=o32, out_bf16=ob))
report("out  <384,RES_LN>", lambda: L.conv_gemm(xs, wo, b2, block_n=D, epi=L.EPI_RES_LN, res_f32=x32, res_bstride=T * D, res_tstride=D, ln1=g, out_f32=o32, out_bf16=ob))
report("post <192,AFFINE> k5", lambda: L.conv_gemm(xs, w5, b2, taps=5, pad_left=4, block_n=192, act=L.ACT_TANH, out_bf16=ob))
