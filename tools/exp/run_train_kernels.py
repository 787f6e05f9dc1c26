"""The training-only kernels changed this round, a few launches each, for an `ncu --set full` capture:
wgrad with the bias gradient in the same launch (attention-projection and FeedForward shapes of the FastSpeech2 step),
the causal attention forward / backward of the Transformer-TTS step, and the batched dgrad-operand transpose."""
import sys

import torch

sys.path.insert(0, ".")


def main():
    from transformer_tts_b200 import lib

    lib.load()
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s: (torch.randn(*s, device=dev, generator=g) * 0.1).to(torch.bfloat16)  # noqa: E731
    B, T, D = 32, 1073, 384
    for rep in range(2):
        dy, x = rnd(B, T, D), rnd(B, T, D)
        dw, db = torch.zeros(D, D, device=dev), torch.zeros(D, device=dev)
        lib.conv_wgrad(dy, x, dw, dbias=db)                                   # out-projection shape
        dh, xn = rnd(B, T, 4 * D), rnd(B, T, D)
        dw1, db1 = torch.zeros(9, 4 * D, D, device=dev), torch.zeros(4 * D, device=dev)
        lib.conv_wgrad(dh, xn, dw1, taps=9, pad_left=4, dbias=db1)            # FeedForward conv1, k = 9
        # Transformer-TTS decoder self-attention, B = 32 x 399 positions, H = 4
        Tt, H = 399, 4
        qkv = rnd(B, Tt, 3 * D)
        mask = torch.ones(B, Tt, device=dev, dtype=torch.uint8)
        kv = lib.mask_lengths(mask)
        ctx = torch.empty(B, Tt, D, device=dev, dtype=torch.bfloat16)
        lse = torch.empty(B, H, Tt, device=dev)
        lib.attention_fwd(qkv, mask, kv, ctx, H, lse2=lse, causal=True)
        dctx = rnd(B, Tt, D)
        delta = lib.attn_delta(dctx, ctx, H)
        lib.attention_bwd(qkv, dctx, mask, kv, lse, delta, H, causal=True)
        srcs = [rnd(9, 4 * D, D), rnd(9, D, 4 * D), rnd(1, 3 * D, D), rnd(1, D, D)]
        dsts = [torch.empty(a.shape[0], a.shape[2], a.shape[1], device=dev, dtype=torch.bfloat16) for a in srcs]
        lib.weight_transpose(srcs, dsts)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
