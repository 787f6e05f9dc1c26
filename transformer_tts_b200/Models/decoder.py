"""Drop-in for ``Decoder`` of the reference's ``Models/decoder.py:29-56`` (Transformer-TTS).

The reference evaluates the decoder on a whole prefix at once and, at inference, re-runs it on the growing prefix at
every step (test.py:128-191).  Here the decoder is a STEP machine: `DecodeState.step()` processes one new position of
every utterance against cached keys / values (csrc/decode.cu: `small_linear`, `attn_decode`), one CUDA graph replay per
position.  `Decoder.forward` (teacher-forced prefix) feeds the given frames through the same steps; the autoregressive
loop is `Transformer.generate` (Models/transformer.py).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .. import lib
from . import _engine as E
from .layers import DecoderLayer
from .modules import PositionalEncoder
from .prenets import DecoderPreNet


class MultiSequential(torch.nn.Sequential):
    """Same container class as the reference (decoder.py:10-17): keys `layers.<i>.` in the state_dict."""

    def forward(self, *args):
        for m in self:
            args = m(*args)
        return args


def repeat(N, fn):
    return MultiSequential(*[fn() for _ in range(N)])


class DecodeState:
    """Buffers and the captured graph of one decode session: R utterances, up to `cap` positions.

    Per position (launch list, 9 per layer): prenet fc1 / fc2 (+ positional encoding); per layer LN1 + q|k|v, self-attention
    (appends this position's key / value), out-proj + residual, LN2 + q, encoder-decoder attention, out-proj + residual,
    LN3 + FeedForward conv1 + ReLU, conv2, FeedForward LayerNorm + residual; final LN + mel / stop heads; position += 1.
    """

    def __init__(self, dec: "Decoder", head: dict, e_bf16: torch.Tensor, src_len: torch.Tensor, cap: int,
                 forced_in: Optional[torch.Tensor] = None, trg_len: Optional[torch.Tensor] = None,
                 use_cuda_graph: bool = True):
        t = dec._cache()
        self.dec, self.t, self.head = dec, t, head
        R, L, D = e_bf16.shape
        H = dec.heads
        dev = e_bf16.device
        self.R, self.L, self.D, self.H, self.Dh, self.cap = R, L, D, H, D // H, cap
        self.n_head = head["w"].shape[0]            # mel_dim * r + r
        self.n_out = (self.n_head + 3) // 4 * 4     # row pitch of out_all: 16-byte aligned rows for the fed-back mel
        self.mel_dim = dec.decoder_prenet.input_size
        self.src_len = src_len
        self.trg_len = trg_len
        self.forced_in = forced_in
        if cap > t["pe"].shape[0]:
            raise ValueError(f"{cap} decoder positions exceed the positional table ({t['pe'].shape[0]}, modules.py:92)")
        f32, bf16 = torch.float32, torch.bfloat16
        self.pos = torch.zeros(1, device=dev, dtype=torch.int32)
        self.done = 0                               # host copy of *pos (no device read in the loop)
        self.out_all = torch.zeros((R, cap, self.n_out), device=dev, dtype=f32)   # [mel (r frames) | stop (r)] per position
        self.x = torch.empty((R, D), device=dev, dtype=f32)        # residual stream of the current position
        self.x2 = torch.empty((R, D), device=dev, dtype=f32)
        self.h1 = torch.empty((R, t["fc1_w"].shape[0]), device=dev, dtype=bf16)
        self.qkv = torch.empty((R, 3 * D), device=dev, dtype=bf16)
        self.q2 = torch.empty((R, D), device=dev, dtype=bf16)
        self.ctx = torch.empty((R, D), device=dev, dtype=bf16)
        self.xn = torch.empty((R, D), device=dev, dtype=bf16)
        self.hid = torch.empty((R, 4 * D), device=dev, dtype=bf16)
        self.u = torch.empty((R, D), device=dev, dtype=f32)
        self.zn = torch.empty((R, D), device=dev, dtype=bf16)
        # self-attention cache: [layer][k|v][R][H][cap][Dh]
        self.kv_self = torch.zeros((dec.N, 2, R, H, cap, self.Dh), device=dev, dtype=bf16)
        # encoder-decoder keys / values: one k|v projection of the encoder output per layer, (R, L, 2D), computed once
        self.kv_cross = []
        for lc in t["layers"]:
            kv = torch.empty((R, L, 2 * D), device=dev, dtype=bf16)
            lib.conv_gemm(e_bf16, lc["wkv2"], lc["bkv2"], block_n=128, out_bf16=kv)
            self.kv_cross.append(kv)
        self.graph = None
        self.use_cuda_graph = use_cuda_graph

    # -- one position -------------------------------------------------------------------------------------------------
    def _launch_step(self) -> None:
        t, R, D, H, Dh, cap = self.t, self.R, self.D, self.H, self.Dh, self.cap
        pos = self.pos
        if self.forced_in is not None:   # teacher-forced: decoder input of position p = trg[:, p]
            T = self.forced_in.shape[1]
            src = dict(x_f32=self.forced_in, ldx=T * self.mel_dim, x_pos_stride=self.mel_dim, x_pos_bias=0)
        else:                            # autoregressive: frame 0 of the previous position's r frames, zeros at position 0
            src = dict(x_f32=self.out_all, ldx=cap * self.n_out, x_pos_stride=self.n_out, x_pos_bias=-1)
        lib.small_linear(t["fc1_w"], t["fc1_b"], self.mel_dim, R=R, relu=True, out_bf16=self.h1, pos=pos, **src)
        lib.small_linear(t["fc2_w"], t["fc2_b"], t["fc1_w"].shape[0], x_bf16=self.h1, relu=True, add_row=t["pe"],
                         add_scale=t["alpha"], out_f32=self.x, pos=pos)
        x, x2 = self.x, self.x2
        sb, sh, sj = H * cap * Dh, cap * Dh, Dh
        n_layers = len(t["layers"])
        for i, lc in enumerate(t["layers"]):
            if i == 0:
                lib.small_linear(lc["wqkv1"], lc["bqkv1"], D, x_f32=x, ln=lc["ln1"], out_bf16=self.qkv)
            else:      # norm_1 of this block was applied by the previous block's FeedForward tail
                lib.small_linear(lc["wqkv1"], lc["bqkv1"], D, x_bf16=self.zn, out_bf16=self.qkv)
            lib.attn_decode(self.qkv, 0, 3 * D, self.kv_self[i], R * sb, (sb, sh, sj), cap, H, Dh, self.ctx,
                            kv_new=self.qkv, knew_off=D, vnew_off=2 * D, pos=pos, kv_len=self.trg_len)
            # zn = the NEXT LayerNorm of the block's output: norm_1 of the following block (layers.py:110) or the decoder's
            # final norm (decoder.py:56)
            next_ln = t["layers"][i + 1]["ln1"] if i + 1 < n_layers else t["norm"]
            if self.dec.use_cluster_kernels:
                # 5 launches per block: the stretches between the attentions run as ONE 16-CTA cluster kernel each
                # (csrc/decode_cluster.cu): result slices are exchanged through distributed shared memory
                lib.decode_attn_out_q(self.ctx, x, lc["wo1"], lc["bo1"], lc["ln2"], lc["wq2"], lc["bq2"], E.LN_EPS, x2, self.q2)
                lib.attn_decode(self.q2, 0, D, self.kv_cross[i], D, (self.L * 2 * D, Dh, 2 * D), self.L, H, Dh, self.ctx,
                                kv_len=self.src_len)
                lib.decode_ffn_block(self.ctx, x2, lc["wo2"], lc["bo2"], lc["ln3"], lc["w1"], lc["b1"], lc["w2"], lc["b2"],
                                     lc["ln_ff"], next_ln, E.LN_EPS, x, self.zn)
                continue
            lib.small_linear(lc["wo1"], lc["bo1"], D, x_bf16=self.ctx, res_f32=x, out_f32=x2)
            lib.small_linear(lc["wq2"], lc["bq2"], D, x_f32=x2, ln=lc["ln2"], out_bf16=self.q2)
            lib.attn_decode(self.q2, 0, D, self.kv_cross[i], D, (self.L * 2 * D, Dh, 2 * D), self.L, H, Dh, self.ctx,
                            kv_len=self.src_len)
            lib.small_linear(lc["wo2"], lc["bo2"], D, x_bf16=self.ctx, res_f32=x2, out_f32=x)
            lib.small_linear(lc["w1"], lc["b1"], D, x_f32=x, ln=lc["ln3"], relu=True, out_bf16=self.hid, xn_bf16=self.xn)
            lib.small_linear(lc["w2"], lc["b2"], 4 * D, x_bf16=self.hid, out_f32=self.u)
            # x2 = x + LN_ff(u + LN3(x))  (modules.py:85-87, layers.py:122)
            lib.ffn_tail_rows(self.u, self.xn, x, lc["ln_ff"], next_ln, E.LN_EPS, x2, self.zn)
            x, x2 = x2, x
        lib.small_linear(self.head["w"], self.head["b"], D, x_bf16=self.zn, out_f32=self.out_all,
                         ldo=cap * self.n_out, pos=pos, out_pos_stride=self.n_out)
        lib.advance_pos(pos, 1)

    def step_eager(self) -> None:
        self._launch_step()
        self.done += 1

    def capture(self) -> None:
        """Capture one decode step as a CUDA graph (on the current stream); processes no position."""
        # an odd number of residual swaps would make the captured step end in the other buffer: 9 launches per layer
        # leave x in self.x after an even number of layers, in self.x2 after an odd one — the head reads whichever
        # it is inside the capture, and the next replay starts from the prenet again, so no state crosses steps
        g = torch.cuda.CUDAGraph()
        launches = lib.LAUNCH_COUNT
        with torch.cuda.graph(g):
            self._launch_step()
        self.launches_per_step = lib.LAUNCH_COUNT - launches
        lib.LAUNCH_COUNT = launches   # capture launched nothing
        self.graph = g

    def replay(self) -> None:
        """One position from the captured graph, on the current stream."""
        self.graph.replay()
        self.done += 1
        lib.LAUNCH_COUNT += self.launches_per_step

    def run(self, n_steps: int) -> None:
        """Process the next n_steps positions (all utterances)."""
        if self.done + n_steps > self.cap:
            raise ValueError(f"{self.done} + {n_steps} positions exceed the session's capacity {self.cap}")
        if n_steps <= 0:
            return
        if not self.use_cuda_graph:
            for _ in range(n_steps):
                self.step_eager()
            return
        if self.graph is None:
            self.step_eager()            # warm-up (TMA-free kernels: only the smem opt-ins and lazy module load)
            n_steps -= 1
            if n_steps == 0:
                return
            self.capture()
        for _ in range(n_steps):
            self.replay()


class Decoder(nn.Module):
    def __init__(self, vocab_size, d_model, N, heads, ff_conv_kernel_size, concat_after_decoder, dropout,
                 dropout_prenet=0.5, multi_speaker=False, spk_emb_dim=None, output_type=None):
        super().__init__()
        if multi_speaker or output_type:
            raise NotImplementedError("speaker embeddings / discrete targets are out of scope (SURVEY §2a)")
        if ff_conv_kernel_size != 1:
            raise NotImplementedError("a decoder FeedForward kernel wider than 1 looks at FUTURE positions (the conv is "
                                      "centred, modules.py:76-78): it has no incremental form; hparams default is 1")
        self.N = N
        self.heads = heads
        self.output_type = output_type
        self.decoder_prenet = DecoderPreNet(vocab_size, d_model, p=dropout_prenet, output_type=output_type)
        self.pe = PositionalEncoder(d_model, dropout=dropout)
        self.layers = repeat(N, lambda: DecoderLayer(d_model, heads, ff_conv_kernel_size, dropout, concat_after_decoder,
                                                     multi_speaker, spk_emb_dim))
        self.norm = nn.LayerNorm(d_model)
        self._ttsb_cache = E.Cache()
        # the stretches between the attentions of a block as cluster kernels (5 launches per block instead of 9)
        self.use_cluster_kernels = d_model == 384

    def _cache(self) -> dict:
        c = self._ttsb_cache
        params = list(self.parameters())
        if c.stale(params):
            t = c.t
            dev = self.norm.weight.device
            D = self.norm.weight.numel()
            pn = self.decoder_prenet.layer

            def opnd(*ws):
                """bf16 [N][K] operand of the rows of `ws` (K a multiple of 32: the layout small_linear streams is the plain
                one — a view of a trainer's bf16 mirror when there is one); otherwise the zero-padded copy."""
                if ws[0].shape[1] % 32 == 0:
                    return E._bf16_cat(list(ws)) if len(ws) > 1 else E._bf16(ws[0])
                return lib.pad_weight_rows(torch.cat([w.detach() for w in ws], 0))

            t["fc1_w"], t["fc1_b"] = opnd(pn.fc1.weight), E._f32(pn.fc1.bias)
            t["fc2_w"], t["fc2_b"] = opnd(pn.fc2.weight), E._f32(pn.fc2.bias)
            t["alpha"] = E._f32(self.pe.alpha)
            t["pe"] = self.pe.table(dev)
            t["norm"] = (E._f32(self.norm.weight), E._f32(self.norm.bias))
            t["ident"] = (torch.ones(D, device=dev), torch.zeros(D, device=dev))
            t["layers"] = []
            for l in self.layers:
                a1, a2, ff = l.attn_1, l.attn_2, l.ff
                catb = lambda *ws: E._f32_cat(list(ws))  # noqa: E731
                lc = dict(
                    ln1=(E._f32(l.norm_1.weight), E._f32(l.norm_1.bias)),
                    ln2=(E._f32(l.norm_2.weight), E._f32(l.norm_2.bias)),
                    ln3=(E._f32(l.norm_3.weight), E._f32(l.norm_3.bias)),
                    ln_ff=(E._f32(ff.layer_norm.weight), E._f32(ff.layer_norm.bias)),
                    wqkv1=opnd(a1.q_linear.weight, a1.k_linear.weight, a1.v_linear.weight),
                    bqkv1=catb(a1.q_linear.bias, a1.k_linear.bias, a1.v_linear.bias),
                    wo1=opnd(a1.out.weight), bo1=E._f32(a1.out.bias),
                    wq2=opnd(a2.q_linear.weight), bq2=E._f32(a2.q_linear.bias),
                    # k|v projection of the encoder output: a (1, 2D, D) conv_gemm operand, run once per utterance batch
                    wkv2=E._bf16_cat([a2.k_linear.weight, a2.v_linear.weight]).unsqueeze(0).contiguous(),
                    bkv2=catb(a2.k_linear.bias, a2.v_linear.bias),
                    wo2=opnd(a2.out.weight), bo2=E._f32(a2.out.bias),
                    w1=E.conv_w_to_taps(ff.f_1.weight)[0], b1=E._f32(ff.f_1.bias),
                    w2=E.conv_w_to_taps(ff.f_2.weight)[0], b2=E._f32(ff.f_2.bias),
                )
                    # the same weights as (taps=1, N, C) conv_gemm operands, for the batched teacher-forced prefix (run_prefix)
                for name in ("wqkv1", "wo1", "wq2", "wo2", "w1", "w2"):
                    lc[name + "_g"] = lc[name].unsqueeze(0).contiguous()
                t["layers"].append(lc)
            fc1 = pn.fc1.weight.detach()
            t["fc1_g"] = E._bf16(fc1).unsqueeze(0).contiguous()          # (1, 256, mel_dim): C = 80, partial K box
            t["fc2_g"] = t["fc2_w"].unsqueeze(0).contiguous()
            c.commit(params)
        return c.t

    def run_prefix(self, head: dict, trg: torch.Tensor, e_outputs: torch.Tensor, src_mask, trg_key_mask_u8: torch.Tensor):
        """The decoder on a whole teacher-forced prefix at once (Models/decoder.py:43-56 as the reference evaluates it):
        trg (B, T, mel_dim) decoder inputs -> (B, T, head rows) fp32 = [mel (r frames) | stop (r)] per position.

        Tensor-core path: the tcgen05 GEMMs and the fused FeedForward of the FastSpeech2 blocks, the tcgen05 attention
        with the CAUSAL flag for attn_1, and for attn_2 (queries = decoder rows, keys = encoder rows) the same attention
        kernel on a packed sequence [encoder rows | decoder rows] whose key mask admits only the encoder rows — valid keys
        form a prefix, so every query tile loads ceil(L / 64) key tiles; the encoder rows' own outputs are discarded."""
        E.require_inference(self, "Decoder")
        t = self._cache()
        B, T, _ = trg.shape
        L, D, H = e_outputs.shape[1], self.norm.weight.numel(), self.heads
        dev = trg.device
        bf16, f32 = torch.bfloat16, torch.float32
        if T > t["pe"].shape[0]:
            raise ValueError(f"{T} decoder positions exceed the positional table ({t['pe'].shape[0]}, modules.py:92)")
        e_bf16 = e_outputs.to(bf16).contiguous()
        m_src = E.mask_to_u8(src_mask, B, L, dev)
        kv_trg = lib.mask_lengths(trg_key_mask_u8)
        maskp = torch.zeros((B, L + T), device=dev, dtype=torch.uint8)
        maskp[:, :L] = m_src
        kvp = lib.mask_lengths(maskp)
        # prenet (prenets.py:33-38) + positional encoding (modules.py:107-111)
        h1 = torch.empty((B, T, t["fc1_g"].shape[1]), device=dev, dtype=bf16)
        lib.conv_gemm(trg.to(bf16).contiguous(), t["fc1_g"], t["fc1_b"], block_n=128, act=lib.ACT_RELU, out_bf16=h1)
        pe = (t["alpha"] * t["pe"][:T]).contiguous()
        x = torch.empty((B, T, D), device=dev, dtype=f32)
        lib.conv_gemm(h1, t["fc2_g"], t["fc2_b"], block_n=128, act=lib.ACT_RELU, res_f32=pe, res_bstride=0, res_tstride=D,
                      out_f32=x)
        xn = torch.empty((B, T, D), device=dev, dtype=bf16)
        l0 = t["layers"][0]
        lib.layernorm(x, l0["ln1"][0], l0["ln1"][1], E.LN_EPS, y_bf16=xn)
        packed = torch.zeros((B, L + T, 3 * D), device=dev, dtype=bf16)
        n_layers = len(t["layers"])
        for i, lc in enumerate(t["layers"]):
            qkv = torch.empty((B, T, 3 * D), device=dev, dtype=bf16)
            lib.conv_gemm(xn, lc["wqkv1_g"], lc["bqkv1"], block_n=192, out_bf16=qkv)
            ctx = torch.empty((B, T, D), device=dev, dtype=bf16)
            lib.attention_fwd(qkv, trg_key_mask_u8, kv_trg, ctx, H, causal=True)                       # layers.py:111
            x2 = torch.empty((B, T, D), device=dev, dtype=f32)
            x2n = torch.empty((B, T, D), device=dev, dtype=bf16)
            lib.conv_gemm(ctx, lc["wo1_g"], lc["bo1"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x, res_bstride=T * D,
                          res_tstride=D, ln1=lc["ln2"], out_f32=x2, out_bf16=x2n)                      # :112, :114
            q2 = torch.empty((B, T, D), device=dev, dtype=bf16)
            lib.conv_gemm(x2n, lc["wq2_g"], lc["bq2"], block_n=128, out_bf16=q2)
            kv = torch.empty((B, L, 2 * D), device=dev, dtype=bf16)
            lib.conv_gemm(e_bf16, lc["wkv2"], lc["bkv2"], block_n=128, out_bf16=kv)
            packed[:, :L, D:] = kv
            packed[:, L:, :D] = q2
            ctxp = torch.empty((B, L + T, D), device=dev, dtype=bf16)
            lib.attention_fwd(packed, maskp, kvp, ctxp, H)                                             # :115
            ctx2 = ctxp[:, L:, :].contiguous()
            x3 = torch.empty((B, T, D), device=dev, dtype=f32)
            x3n = torch.empty((B, T, D), device=dev, dtype=bf16)
            lib.conv_gemm(ctx2, lc["wo2_g"], lc["bo2"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x2, res_bstride=T * D,
                          res_tstride=D, ln1=lc["ln3"], out_f32=x3, out_bf16=x3n)                      # :116, :118
            next_ln = t["layers"][i + 1]["ln1"] if i + 1 < n_layers else t["norm"]
            z = torch.empty((B, T, D), device=dev, dtype=f32)
            zn = torch.empty((B, T, D), device=dev, dtype=bf16)
            lib.ffn_fused(x3n, lc["w1_g"], lc["b1"], lc["w2_g"], lc["b2"], x3, lc["ln_ff"], next_ln, z, zn)   # :122, next :110
            x, xn = z, zn
        out = torch.empty((B, T, head["w_g"].shape[1]), device=dev, dtype=f32)
        lib.conv_gemm(xn, head["w_g"], head["b_g"], block_n=192, out_f32=out)                          # decoder.py:56 + heads
        return out

    def begin(self, head: dict, e_outputs: torch.Tensor, src_mask, cap: int, forced_in=None, trg_len=None,
              use_cuda_graph: bool = True) -> DecodeState:
        """Open a decode session over encoder outputs (R, L, D); head = the mel / stop projection (Transformer._head)."""
        E.require_inference(self, "Decoder")
        E.require_cuda(e_outputs, "Decoder")
        R, L, _ = e_outputs.shape
        m = E.mask_to_u8(src_mask, R, L, e_outputs.device)
        kv = lib.mask_lengths(m)
        if not torch.equal(kv[:, 0], kv[:, 1]):
            raise NotImplementedError("the source mask must be a padding mask (valid prefix) — train.py:50")
        if int(kv[:, 0].min()) < 1:
            raise ValueError("an utterance without a single valid phoneme")
        return DecodeState(self, head, e_outputs.to(torch.bfloat16).contiguous(), kv[:, 0].contiguous(), cap,
                           forced_in=forced_in, trg_len=trg_len, use_cuda_graph=use_cuda_graph)

    def forward(self, trg, e_outputs, src_mask, trg_mask, spk_emb=None, attn_detach=True):
        raise NotImplementedError("Decoder.forward needs the output heads to run its step machine: call Transformer.forward "
                                  "(teacher-forced prefix) or Transformer.generate (autoregressive)")
