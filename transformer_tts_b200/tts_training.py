"""Transformer-TTS teacher-forced training step on the sm_100a kernels (BASELINE.json configs[3]).

What the reference does in ``train.py:160-263`` with autograd — encoder, teacher-forced decoder (causal self-attention,
encoder-decoder attention, FeedForward), mel / stop heads, two L1 losses and the stop-token BCE, backward, clip + Adam — as an
explicit kernel sequence, in the style of ``training.Trainer`` (whose flat parameter / gradient / optimizer buffers, text
encoder forward / backward, NCCL gradient exchange and optimizer step this class inherits):

    encoder        Trainer._stack_fwd / _stack_bwd (the FastSpeech2 FFT stack, Models/encoder.py:83-112)
    decoder        prenet (Models/prenets.py:33-38) + positional encoding; per block (Models/layers.py:109-124):
                   q|k|v GEMM, CAUSAL tcgen05 attention (+ lse), out-projection + residual + LayerNorm epilogue,
                   query GEMM, encoder-decoder attention on the packed sequence [encoder rows | decoder rows] (the same
                   attention kernels: the key mask admits only the encoder rows), out-projection + residual + LayerNorm,
                   FeedForward conv1 + ReLU, conv2 + LayerNorm + residual + next LayerNorm epilogue
    heads / loss   mel head (Models/transformer.py:109), stop head (:111), 2 x nn.L1Loss on the same tensor
                   (PostConvNet(prev_version=False) returns its input, Models/postnets.py:64-79) and
                   binary_cross_entropy_with_logits with pos_weight (train.py:214-219)
    backward       dgrad = conv_gemm with transposed weights, wgrad = tcgen05 MN-major GEMM, causal attention backward,
                   the packed-sequence trick for attn_2 (zero dctx on the encoder rows: the kernel's dK / dV of the encoder
                   rows and dQ of the decoder rows are exactly the encoder-decoder attention's gradients), LayerNorm backward

Scope of this first version: dropout 0 only (the reference trains with dropout 0.1 and prenet dropout 0.5: the counter-based
masks of training.DropoutPlan are not wired into this sequence yet, a non-zero probability raises), d_model_encoder ==
d_model_decoder, decoder FeedForward kernel 1.  PostConvNet is not evaluated (its result is discarded by the reference and
it receives no gradient; its BatchNorm running statistics, which the reference updates as a side effect, stay as loaded).
"""
from __future__ import annotations

from typing import Dict

import torch

from . import lib
from .Models import _engine as E
from .training import Trainer, _t


class _NoDropout:
    """Dropout plan of a model whose every dropout probability is 0 (what this training step supports)."""

    def __init__(self, model):
        ps = {m.p for m in model.modules() if isinstance(m, torch.nn.Dropout)}
        ps |= {float(m.dropout or 0.0) for m in model.modules() if hasattr(m, "q_linear")}
        if any(p > 0 for p in ps):
            raise NotImplementedError("TTSTrainer: dropout > 0 is not built for the Transformer-TTS training step "
                                      f"(probabilities found: {sorted(ps)}); construct the model with dropout 0")
        self.p_stack = {"enc": 0.0, "dec": 0.0}
        self.calls = 0

    active = False

    def site(self, name: str) -> int:
        return -1

    def next_seed(self) -> int:
        self.calls += 1
        return 0


class TTSTrainer(Trainer):
    """Flat buffers + training steps of a ``Models.transformer.Transformer`` (see the module docstring)."""

    def __init__(self, model, *, positive_weight: float = 1.0, **kw):
        if model.linear is not None:
            raise NotImplementedError("TTSTrainer: d_model_encoder != d_model_decoder is not built")
        self.positive_weight = float(positive_weight)
        super().__init__(model, sync_bn=False, **kw)

    def _dropout_plan(self, model, seed: int):
        return _NoDropout(model)

    # ------------------------------------------------------------------------------------------------------------------
    def _head_cache(self) -> dict:
        m = self.model
        c = m._ttsb_cache
        params = [m.out.weight, m.out.bias, m.stop_token.weight, m.stop_token.bias]
        if c.stale(params) or "out_g" not in c.t:
            m._head()
            dev = m.out.weight.device
            c.t["out_g"] = E._bf16(m.out.weight).unsqueeze(0).contiguous()       # (1, mel*r, D)
            c.t["out_b"] = E._f32(m.out.bias)
            r = m.stop_token.weight.shape[0]
            sg = torch.zeros((1, 80, m.stop_token.weight.shape[1]), device=dev, dtype=torch.bfloat16)
            sg[0, :r] = m.stop_token.weight.detach().to(torch.bfloat16)         # stop head padded to one 80-column tile
            sb = torch.zeros(80, device=dev, dtype=torch.float32)
            sb[:r] = m.stop_token.bias.detach()
            c.t["stop_g"], c.t["stop_b"] = sg, sb
        return c.t

    def forward_train(self, text, trg, src_mask, trg_key_mask):
        """text (B,L) ids; trg (B,T,mel_dim) decoder inputs; src_mask (B,1,L) / (B,L); trg_key_mask (B,T) uint8 (the padding
        half of the target mask: the causal half is the kernel's).  Returns (mel (B,T,mel*r) fp32, stop (B,T,r) fp32, tape)."""
        model = self.model
        dec = model.decoder
        B, L = text.shape
        T = trg.shape[1]
        dev = text.device
        D, H = dec.norm.weight.numel(), dec.heads
        bf16, f32 = torch.bfloat16, torch.float32
        bf = lambda *s: torch.empty(s, device=dev, dtype=bf16)   # noqa: E731
        fl = lambda *s: torch.empty(s, device=dev, dtype=f32)    # noqa: E731
        self.drop.next_seed()
        m_src = E.mask_to_u8(src_mask, B, L, dev)
        kv_src = lib.mask_lengths(m_src)
        e_bf16, enc_tape = self._stack_fwd(model.encoder, text, m_src, kv_src, tag="enc", seed=0)
        t = dec._cache()
        hd = self._head_cache()
        if D != 384:
            raise NotImplementedError("TTSTrainer: d_model 384 only (the row-complete LayerNorm epilogue's tile)")
        if T > t["pe"].shape[0]:
            raise ValueError(f"{T} decoder positions exceed the positional table ({t['pe'].shape[0]})")
        kv_trg = lib.mask_lengths(trg_key_mask)
        maskp = torch.zeros((B, L + T), device=dev, dtype=torch.uint8)
        maskp[:, :L] = m_src
        kvp = lib.mask_lengths(maskp)
        # prenet + positional encoding
        trg_bf = trg.to(bf16).contiguous()
        h1 = bf(B, T, t["fc1_g"].shape[1])
        lib.conv_gemm(trg_bf, t["fc1_g"], t["fc1_b"], block_n=128, act=lib.ACT_RELU, out_bf16=h1)
        pe = t["pe"][:T]
        pe_s = (t["alpha"] * pe).contiguous()
        x = fl(B, T, D)                                           # x0 = relu(fc2) + alpha * pe (modules.py:107-111), fp32
        lib.conv_gemm(h1, t["fc2_g"], t["fc2_b"], block_n=128, act=lib.ACT_RELU, res_f32=pe_s, res_bstride=0, res_tstride=D,
                      out_f32=x)
        a2 = (x - pe_s).to(bf16)                                  # the ReLU's output again (exactly 0 where it was 0): the gate
        l0 = t["layers"][0]
        xn = bf(B, T, D)
        lib.layernorm(x, l0["ln1"][0], l0["ln1"][1], E.LN_EPS, y_bf16=xn)
        tape = dict(enc=enc_tape, e=e_bf16, trg=trg_bf, h1=h1, a2=a2, pe=pe, m_trg=trg_key_mask, kv_trg=kv_trg, maskp=maskp,
                    kvp=kvp, blocks=[], L=L, T=T)
        n_layers = len(t["layers"])
        for i, lc in enumerate(t["layers"]):
            qkv = bf(B, T, 3 * D)
            lib.conv_gemm(xn, lc["wqkv1_g"], lc["bqkv1"], block_n=192, out_bf16=qkv)
            ctx, lse = bf(B, T, D), fl(B, H, T)
            lib.attention_fwd(qkv, trg_key_mask, kv_trg, ctx, H, lse2=lse, causal=True)
            x2, x2n = fl(B, T, D), bf(B, T, D)
            lib.conv_gemm(ctx, lc["wo1_g"], lc["bo1"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x, res_bstride=T * D,
                          res_tstride=D, ln1=lc["ln2"], out_f32=x2, out_bf16=x2n)
            q2 = bf(B, T, D)
            lib.conv_gemm(x2n, lc["wq2_g"], lc["bq2"], block_n=128, out_bf16=q2)
            kv = bf(B, L, 2 * D)
            lib.conv_gemm(e_bf16, lc["wkv2"], lc["bkv2"], block_n=128, out_bf16=kv)
            packed = torch.zeros((B, L + T, 3 * D), device=dev, dtype=bf16)
            packed[:, :L, D:] = kv
            packed[:, L:, :D] = q2
            ctxp, lsep = bf(B, L + T, D), fl(B, H, L + T)
            lib.attention_fwd(packed, maskp, kvp, ctxp, H, lse2=lsep)
            ctx2 = ctxp[:, L:, :].contiguous()
            x3, x3n = fl(B, T, D), bf(B, T, D)
            lib.conv_gemm(ctx2, lc["wo2_g"], lc["bo2"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x2, res_bstride=T * D,
                          res_tstride=D, ln1=lc["ln3"], out_f32=x3, out_bf16=x3n)
            h = bf(B, T, 4 * D)
            lib.conv_gemm(x3n, lc["w1_g"], lc["b1"], block_n=256, act=lib.ACT_RELU, out_bf16=h)
            next_ln = t["layers"][i + 1]["ln1"] if i + 1 < n_layers else t["norm"]
            upre = fl(B, T, D)
            z, zn = E.run_ffn2(dict(k=1, w2=lc["w2_g"], b2=lc["b2"], ln_ff=lc["ln_ff"]), h, x3n, x3, next_ln, pre_f32=upre)
            tape["blocks"].append(dict(x=x, xn=xn, qkv=qkv, ctx=ctx, lse=lse, x2=x2, x2n=x2n, packed=packed, ctxp=ctxp,
                                       lsep=lsep, ctx2=ctx2, x3=x3, x3n=x3n, h=h, upre=upre))
            x, xn = z, zn
        tape["z_last"], tape["zn_last"] = x, xn
        mel = fl(B, T, hd["out_g"].shape[1])
        lib.conv_gemm(xn, hd["out_g"], hd["out_b"], block_n=80, out_f32=mel)
        stop80 = fl(B, T, 80)
        lib.conv_gemm(xn, hd["stop_g"], hd["stop_b"], block_n=80, out_f32=stop80)
        r = model.stop_token.weight.shape[0]
        return mel, stop80[:, :, :r].contiguous(), tape

    # ------------------------------------------------------------------------------------------------------------------
    def backward_train(self, tape, g_mel: torch.Tensor, g_stop: torch.Tensor) -> None:
        """g_mel (B,T,mel*r) fp32, g_stop (B,T,r) fp32: gradients of the loss w.r.t. the two head outputs."""
        model = self.model
        dec = model.decoder
        G = self.G
        t = dec._cache()
        hd = self._head_cache()
        B, T, L = g_mel.shape[0], tape["T"], tape["L"]
        dev = g_mel.device
        D, H = dec.norm.weight.numel(), dec.heads
        bf16, f32 = torch.bfloat16, torch.float32
        bf = lambda *s: torch.empty(s, device=dev, dtype=bf16)   # noqa: E731
        fl = lambda *s: torch.empty(s, device=dev, dtype=f32)    # noqa: E731
        # ---- heads: out (transformer.py:109) and stop_token (:111) on zn = LN_norm(z_last) ----
        zn = tape["zn_last"]
        g_mel_bf = g_mel.to(bf16)
        lib.colsum(g_mel_bf, G(model.out.bias))
        lib.conv_wgrad(g_mel_bf, zn, G(model.out.weight))
        r = g_stop.shape[2]
        g_stop80 = torch.zeros((B, T, 80), device=dev, dtype=bf16)
        g_stop80[:, :, :r] = g_stop.to(bf16)
        dws = torch.zeros((80, D), device=dev, dtype=f32)
        dbs = torch.zeros(80, device=dev, dtype=f32)
        lib.colsum(g_stop80, dbs)
        lib.conv_wgrad(g_stop80, zn, dws)
        G(model.stop_token.weight).add_(dws[:r])
        G(model.stop_token.bias).add_(dbs[:r])
        dzn1 = fl(B, T, D)
        lib.conv_gemm(g_mel_bf, _t(hd["out_g"]), self.zeros(D), block_n=192, out_f32=dzn1)
        dzn = fl(B, T, D)
        lib.conv_gemm(g_stop80, _t(hd["stop_g"]), self.zeros(D), block_n=192, res_f32=dzn1, res_bstride=T * D,
                      res_tstride=D, out_f32=dzn)
        dz = fl(B, T, D)
        lib.layernorm_bwd(tape["z_last"], t["norm"][0], g_f32=dzn, dx_f32=dz, dgamma=G(dec.norm.weight), dbeta=G(dec.norm.bias))
        de = None      # fp32 gradient w.r.t. the encoder output, accumulated over the blocks
        e_bf16 = tape["e"]
        for i in range(len(t["layers"]) - 1, -1, -1):
            layer, lc, s = dec.layers[i], t["layers"][i], tape["blocks"][i]
            a1, a2, ff = layer.attn_1, layer.attn_2, layer.ff
            # z = x3 + LN_ff(upre);  upre = conv2(h) + b2 + x3n
            dupre, dupre_bf = fl(B, T, D), bf(B, T, D)
            lib.layernorm_bwd(s["upre"], lc["ln_ff"][0], g_f32=dz, dx_f32=dupre, dx_bf16=dupre_bf,
                              dgamma=G(ff.layer_norm.weight), dbeta=G(ff.layer_norm.bias))
            lib.colsum(dupre_bf, G(ff.f_2.bias))
            lib.conv_wgrad(dupre_bf, s["h"], G(ff.f_2.weight), conv_layout=True)
            dh = bf(B, T, 4 * D)
            lib.conv_gemm(dupre_bf, _t(lc["w2_g"]), self.zeros(4 * D), block_n=256, act=lib.ACT_RELU_GRAD, gate_bf16=s["h"],
                          out_bf16=dh)
            lib.colsum(dh, G(ff.f_1.bias))
            lib.conv_wgrad(dh, s["x3n"], G(ff.f_1.weight), conv_layout=True)
            dx3n = fl(B, T, D)
            lib.conv_gemm(dh, _t(lc["w1_g"]), self.zeros(D), block_n=192, res_f32=dupre, res_bstride=T * D, res_tstride=D,
                          out_f32=dx3n)
            # x3n = norm_3(x3);  z = x3 + ff(x3n)
            dx3, dx3_bf = fl(B, T, D), bf(B, T, D)
            lib.layernorm_bwd(s["x3"], lc["ln3"][0], g_f32=dx3n, add_f32=dz, dx_f32=dx3, dx_bf16=dx3_bf,
                              dgamma=G(layer.norm_3.weight), dbeta=G(layer.norm_3.bias))
            # x3 = x2 + ctx2 Wo2^T + bo2
            lib.colsum(dx3_bf, G(a2.out.bias))
            lib.conv_wgrad(dx3_bf, s["ctx2"], G(a2.out.weight))
            dctxp = torch.zeros((B, L + T, D), device=dev, dtype=bf16)      # zero on the encoder rows: they are not outputs
            dctx2 = bf(B, T, D)
            lib.conv_gemm(dx3_bf, _t(lc["wo2_g"]), self.zeros(D), block_n=192, out_bf16=dctx2)
            dctxp[:, L:, :] = dctx2
            delta = lib.attn_delta(dctxp, s["ctxp"], H)
            dqkvp = lib.attention_bwd(s["packed"], dctxp, tape["maskp"], tape["kvp"], s["lsep"], delta, H)
            dq2 = dqkvp[:, L:, :D].contiguous()
            dkv = dqkvp[:, :L, D:].contiguous()                            # (B, L, 2D): d k | d v of the encoder rows
            # k / v projections of the encoder output (modules.py:49,51)
            lib.colsum(dkv, G(a2.k_linear.bias), cols=D, col_off=0)
            lib.colsum(dkv, G(a2.v_linear.bias), cols=D, col_off=D)
            lib.conv_wgrad(dkv, e_bf16, G(a2.k_linear.weight), n_off=0, N=D)
            lib.conv_wgrad(dkv, e_bf16, G(a2.v_linear.weight), n_off=D, N=D)
            de_new = fl(B, L, D)
            if de is None:
                lib.conv_gemm(dkv, _t(lc["wkv2"]), self.zeros(D), block_n=192, out_f32=de_new)
            else:
                lib.conv_gemm(dkv, _t(lc["wkv2"]), self.zeros(D), block_n=192, res_f32=de, res_bstride=L * D, res_tstride=D,
                              out_f32=de_new)
            de = de_new
            # q projection of x2n = norm_2(x2)
            lib.colsum(dq2, G(a2.q_linear.bias))
            lib.conv_wgrad(dq2, s["x2n"], G(a2.q_linear.weight))
            dx2n = fl(B, T, D)
            lib.conv_gemm(dq2, _t(lc["wq2_g"]), self.zeros(D), block_n=192, out_f32=dx2n)
            dx2, dx2_bf = fl(B, T, D), bf(B, T, D)
            lib.layernorm_bwd(s["x2"], lc["ln2"][0], g_f32=dx2n, add_f32=dx3, dx_f32=dx2, dx_bf16=dx2_bf,
                              dgamma=G(layer.norm_2.weight), dbeta=G(layer.norm_2.bias))
            # x2 = x + ctx Wo1^T + bo1; causal self-attention
            lib.colsum(dx2_bf, G(a1.out.bias))
            lib.conv_wgrad(dx2_bf, s["ctx"], G(a1.out.weight))
            dctx = bf(B, T, D)
            lib.conv_gemm(dx2_bf, _t(lc["wo1_g"]), self.zeros(D), block_n=192, out_bf16=dctx)
            delta = lib.attn_delta(dctx, s["ctx"], H)
            dqkv = lib.attention_bwd(s["qkv"], dctx, tape["m_trg"], tape["kv_trg"], s["lse"], delta, H, causal=True)
            for j, lin in enumerate((a1.q_linear, a1.k_linear, a1.v_linear)):
                lib.colsum(dqkv, G(lin.bias), cols=D, col_off=j * D)
                lib.conv_wgrad(dqkv, s["xn"], G(lin.weight), n_off=j * D, N=D)
            dxn = fl(B, T, D)
            lib.conv_gemm(dqkv, _t(lc["wqkv1_g"]), self.zeros(D), block_n=192, out_f32=dxn)
            dx = fl(B, T, D)
            lib.layernorm_bwd(s["x"], lc["ln1"][0], g_f32=dxn, add_f32=dx2, dx_f32=dx, dgamma=G(layer.norm_1.weight),
                              dbeta=G(layer.norm_1.bias))
            dz = dx
        # ---- x0 = relu(fc2(h1)) + alpha * pe;  h1 = relu(fc1(trg)) ----
        pn = dec.decoder_prenet.layer
        lib.embed_bwd(dz, T, pe=tape["pe"].contiguous(), dalpha=G(dec.pe.alpha))
        dz2 = lib.relu_gate(dz, tape["a2"])
        lib.colsum(dz2, G(pn.fc2.bias))
        lib.conv_wgrad(dz2, tape["h1"], G(pn.fc2.weight))
        dh1 = bf(B, T, tape["h1"].shape[2])
        lib.conv_gemm(dz2, _t(t["fc2_g"]), self.zeros(tape["h1"].shape[2]), block_n=128, act=lib.ACT_RELU_GRAD,
                      gate_bf16=tape["h1"], out_bf16=dh1)
        lib.colsum(dh1, G(pn.fc1.bias))
        lib.conv_wgrad(dh1, tape["trg"], G(pn.fc1.weight))
        # ---- text encoder ----
        self._stack_bwd(model.encoder, tape["enc"], de)

    # ------------------------------------------------------------------------------------------------------------------
    def forward_backward(self, text, trg, src_mask, trg_key_mask, mel_target, stop_target) -> Dict[str, torch.Tensor]:
        """mel_target (B, T*r, mel_dim) = mel[:, r:, :]; stop_target (B, T*r) = stop_token[:, r:] (train.py:214-217)."""
        mel, stop, tape = self.forward_train(text, trg, src_mask, trg_key_mask)
        B, T = mel.shape[0], mel.shape[1]
        self.loss_buf.zero_()
        g_mel = torch.empty_like(mel)
        # mel_loss + post_mel_loss: two nn.L1Loss terms on the same values (PostConvNet(prev_version=False) returns its input)
        lib.l1_loss(mel.view(-1), mel_target.reshape(-1).float().contiguous(), self.loss_buf[0:1], grad=g_mel.view(-1))
        g_mel.mul_(2.0)
        g_stop = torch.empty_like(stop)
        lib.bce_logits_loss(stop.view(-1), stop_target.reshape(-1).float().contiguous(), self.loss_buf[1:2], grad=g_stop.view(-1),
                            pos_weight=self.positive_weight)
        self.backward_train(tape, g_mel, g_stop)
        mel_loss, tok = self.loss_buf[0], self.loss_buf[1]
        return dict(loss=2 * mel_loss + tok, mel_loss=mel_loss, post_mel_loss=mel_loss, loss_token=tok, outputs_prenet=mel,
                    stop_token=stop)

    def train_step(self, text, trg, src_mask, trg_key_mask, mel_target, stop_target) -> Dict[str, torch.Tensor]:
        self.zero_grad()
        out = self.forward_backward(text, trg, src_mask, trg_key_mask, mel_target, stop_target)
        out["lr"] = self.optimizer_step()
        return out
