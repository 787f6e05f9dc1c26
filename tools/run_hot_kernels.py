#!/usr/bin/env python
"""Launch every hot kernel of the inference step (decoder shape of cfg2: B=64, T=1103) and of the training step
(B=32, T=1073) twice, for `ncu --set full` captures:

  ncu --set full --clock-control none --import-source on \
      -k regex:'conv_gemm|ffn_fused|attention_fwd|gemm_x3|attention_x3|xrow|wgrad|attn_bwd|ln_bwd|lr_expand|bucket_embed|layernorm_kernel' \
      -o gpurun_out/r02_hot python tools/run_hot_kernels.py
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from transformer_tts_b200 import build_fastspeech2, lib, synth  # noqa: E402
from transformer_tts_b200.Models import _engine as E  # noqa: E402


def main():
    lib.load()
    hp = synth.make_hp()
    model = build_fastspeech2(hp).eval()
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda()
    dev = "cuda"
    H, D = 4, 384
    for B, T, train in ((64, 1103, False), (32, 1073, True)):
        lc = E.layer_cache(model.decoder.layers[0])
        g = torch.Generator().manual_seed(7)
        lens = torch.randint(int(0.8 * T), T + 1, (B,), generator=g)
        lens[0] = T
        mask = (torch.arange(T)[None, :] < lens[:, None]).to(torch.uint8).to(dev)
        kv = lib.mask_lengths(mask)
        xs = (torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16)
        x32 = torch.randn(B, T, D, device=dev)
        qkv = (torch.randn(B, T, 3 * D, device=dev) * 0.5).to(torch.bfloat16)
        hs = torch.relu(torch.randn(B, T, 4 * D, device=dev)).to(torch.bfloat16)
        ob = torch.empty(B, T, D, device=dev, dtype=torch.bfloat16)
        o32 = torch.empty(B, T, D, device=dev)
        lse = torch.empty(B, H, T, device=dev)
        for rep in range(2):
            if not train:
                lib.conv_gemm(xs, lc["wqkv"], lc["bqkv"], block_n=192, out_bf16=qkv)
                lib.attention_fwd(qkv, mask, kv, ob, H)
                lib.conv_gemm(xs, lc["wo"], lc["bo"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x32, res_bstride=T * D,
                              res_tstride=D, ln1=lc["ln2"], out_f32=o32, out_bf16=ob)
                lib.ffn_fused(xs, lc["w1"], lc["b1"], lc["w2"], lc["b2"], x32, lc["ln_ff"], lc["ln1"], o32, ob)
                lib.conv_gemm(xs, lc["w1"], lc["b1"], block_n=256, act=lib.ACT_RELU, out_bf16=hs)
                lib.conv_gemm(hs, lc["w2"], lc["b2"], block_n=D, epi=lib.EPI_FFN2, res_bf16=xs, res_f32=x32,
                              res_bstride=T * D, res_tstride=D, ln1=lc["ln_ff"], ln2=lc["ln1"], out_f32=o32, out_bf16=ob)
                lib.layernorm(x32, lc["ln1"][0], lc["ln1"][1], E.LN_EPS, y_bf16=ob)
                xe = torch.randn(B, 128, D, device=dev).to(torch.bfloat16)
                dur = torch.full((B, 128), T // 128, dtype=torch.int32, device=dev)
                cum, mel_len = lib.lr_scan(dur)
                lib.lr_expand(xe, cum, mel_len, int(mel_len.max().item()), want_aux=True)
                va = model.variance_adaptor
                pb, eb = va._bins(torch.device(dev))
                tabs = va._tables()
                lib.bucket_embed_add(xs, torch.rand(B, T, device=dev) * 700 + 80, torch.rand(B, T, device=dev) * 400, pb,
                                     eb, tabs["ptab"], tabs["etab"])
            else:
                lib.attention_fwd(qkv, mask, kv, ob, H, lse2=lse)
                dctx = (torch.randn(B, T, D, device=dev) * 0.1).to(torch.bfloat16)
                delta = lib.attn_delta(dctx, ob, H)
                lib.attention_bwd(qkv, dctx, mask, kv, lse, delta, H)
                dw = torch.zeros(D, D, device=dev)
                lib.conv_wgrad(xs, xs, dw)                                  # 384 x 384 linear
                dw2 = torch.zeros(4 * D, D, device=dev)
                lib.conv_wgrad(hs, xs, dw2)                                 # FFN conv1 (k=1)
                dg, db = torch.zeros(D, device=dev), torch.zeros(D, device=dev)
                lib.layernorm_bwd(x32, lc["ln1"][0], g_f32=o32.normal_(), add_f32=x32, dx_f32=o32, dx_bf16=ob, dgamma=dg,
                                  dbeta=db)
                lib.colsum(hs, torch.zeros(4 * D, device=dev))
        torch.cuda.synchronize()
    # the fp32-accurate encoder -> duration chain at the encoder shape of cfg2 (B=64, L=128): csrc/precise.cu
    B, L = 64, 128
    lc = E.layer_cache(model.encoder.layers[0])
    xc = E.layer_cache_x3(model.encoder.layers[0])
    m8 = torch.ones(B, L, dtype=torch.uint8, device=dev)
    kvl = lib.mask_lengths(m8)
    x = torch.randn(B, L, D, device=dev)
    for rep in range(2):
        xs = lib.split_f16(x)
        qkv = lib.gemm_x3(xs, xc["wqkv"], lc["bqkv"], xc["sqkv"], split_out=True)
        ctx = lib.attention_x3(qkv, m8, kvl, H)
        o = lib.gemm_x3(ctx, xc["wo"], lc["bo"], xc["so"])
        y, yns = torch.empty_like(x), torch.empty_like(xs)
        lib.xrow(lib.XROW_ADD_LN, x, lc["ln2"], b0=o, s_out=y, out_split=yns)
        hs_ = lib.gemm_x3(yns, xc["w1"], lc["b1"], xc["s1"], taps=5, pad_left=2, relu=True, split_out=True)
        u = lib.gemm_x3(hs_, xc["w2"], lc["b2"], xc["s2"], taps=5, pad_left=2)
        z, zns = torch.empty_like(x), torch.empty_like(xs)
        lib.xrow(lib.XROW_FFN_TAIL, u, lc["ln_ff"], b0_split=yns, c=y, ln2=lc["ln1"], z_out=z, out_split=zns)
    torch.cuda.synchronize()
    print("ok", lib.LAUNCH_COUNT)


if __name__ == "__main__":
    main()
