"""ORACLE — test infrastructure, NOT the product.

A plain-PyTorch fp32 restatement of the reference's Transformer-TTS forward (syoamakase/Transformer_TTS,
`Models/transformer.py` and the modules it calls) and of the autoregressive loop of `test.py:128-191`, written as pure
functions over a reference-layout ``state_dict``.  It shares the encoder / attention / FeedForward restatements with
``oracle/fastspeech2_oracle.py``.  Every function cites the reference file:line it follows.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / reference arm may import this file.

Parity status: PINNED against the reference itself — ``oracle/make_golden_tts.py`` imports the unmodified reference
``Models.transformer.Transformer`` in the build container, runs its forward and the test.py loop on seeded inputs and
commits the outputs under ``tests/golden/tts_*.npz``; ``tests/test_oracle_tts.py`` checks this restatement against those
files (and against the live reference when /root/reference is present).

Scope: decoder_type 'transformer', transformer encoder, no GST, no speaker embeddings,
output_type None, all dropouts 0 (test.py:107 builds the model that way).  `PostConvNet(prev_version=False)` returns its
INPUT (Models/postnets.py:64-79 computes the convolutions and discards them), so outputs_postnet == outputs_prenet.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from . import fastspeech2_oracle as O

Tensor = torch.Tensor


# --------------------------------------------------------------------------------------------------
# train.py:26-58  nopeak_mask / create_masks  (test.py uses the same function)
# --------------------------------------------------------------------------------------------------
def create_masks(src_pos: Tensor, trg_pos: Optional[Tensor], src_pad: int = 0, trg_pad: int = 0):
    src_mask = (src_pos != src_pad).unsqueeze(-2)  # :50  (B,1,L)
    if trg_pos is None:
        return src_mask, None
    trg_mask = (trg_pos != trg_pad).unsqueeze(-2)  # :52  (B,1,T)
    size = trg_pos.size(1)
    np_mask = torch.tril(torch.ones((1, size, size), dtype=torch.bool, device=trg_pos.device))  # :34-35 (triu(k=1) == 0)
    return src_mask, trg_mask & np_mask  # :55  (B,T,T)


# --------------------------------------------------------------------------------------------------
# Models/modules.py:43-70  MultiHeadAttention.forward with separate query and key/value inputs
# --------------------------------------------------------------------------------------------------
def multi_head_attention(sd: Dict[str, Tensor], pfx: str, q_in: Tensor, kv_in: Tensor, mask: Optional[Tensor], heads: int,
                         drop=None, site: str = ""):
    bs, _, d_model = q_in.shape
    d_k = d_model // heads
    k = F.linear(kv_in, sd[pfx + "k_linear.weight"], sd[pfx + "k_linear.bias"]).view(bs, -1, heads, d_k)  # :49
    q = F.linear(q_in, sd[pfx + "q_linear.weight"], sd[pfx + "q_linear.bias"]).view(bs, -1, heads, d_k)  # :50
    v = F.linear(kv_in, sd[pfx + "v_linear.weight"], sd[pfx + "v_linear.bias"]).view(bs, -1, heads, d_k)  # :51
    k, q, v = k.transpose(1, 2), q.transpose(1, 2), v.transpose(1, 2)  # :58-60
    ctx, attn = O.attention(q, k, v, d_k, mask, drop, site)  # :63 (dropout on the probabilities, :18-19)
    concat = ctx.transpose(1, 2).contiguous().view(bs, -1, d_model)  # :65
    return F.linear(concat, sd[pfx + "out.weight"], sd[pfx + "out.bias"]), attn  # :68


# --------------------------------------------------------------------------------------------------
# Models/prenets.py:8-44  DecoderPreNet (output_type None; dropout 0)
# --------------------------------------------------------------------------------------------------
def decoder_prenet(sd, pfx: str, x: Tensor, drop=None) -> Tensor:
    h = O._drop(drop, "prenet.1", F.relu(F.linear(x, sd[pfx + "layer.fc1.weight"], sd[pfx + "layer.fc1.bias"])))  # :33-35
    return O._drop(drop, "prenet.2", F.relu(F.linear(h, sd[pfx + "layer.fc2.weight"], sd[pfx + "layer.fc2.bias"])))  # :36-38


# --------------------------------------------------------------------------------------------------
# Models/layers.py:109-124  DecoderLayer.forward
# --------------------------------------------------------------------------------------------------
def decoder_layer(sd, pfx: str, x: Tensor, e_outputs: Tensor, src_mask, trg_mask, heads: int, drop=None, site: str = ""):
    """`drop` / `site` = "dec.<layer index>": the training-mode dropout call sites, named as in fastspeech2_oracle._drop."""
    d = x.shape[-1]
    res = x
    xn = F.layer_norm(x, (d,), sd[pfx + "norm_1.weight"], sd[pfx + "norm_1.bias"])  # :110
    out, attn_1 = multi_head_attention(sd, pfx + "attn_1.", xn, xn, trg_mask, heads, drop, site + ".attn1")  # :111
    x = res + O._drop(drop, site + ".drop1", out)  # :112
    res = x
    xn = F.layer_norm(x, (d,), sd[pfx + "norm_2.weight"], sd[pfx + "norm_2.bias"])  # :114
    out, attn_2 = multi_head_attention(sd, pfx + "attn_2.", xn, e_outputs, src_mask, heads, drop, site + ".attn2")  # :115
    x = res + O._drop(drop, site + ".drop2", out)  # :116
    res = x
    xn = F.layer_norm(x, (d,), sd[pfx + "norm_3.weight"], sd[pfx + "norm_3.bias"])  # :118
    x = res + O._drop(drop, site + ".drop3", O.feed_forward(sd, pfx + "ff.", xn, drop, site + ".ff"))  # :122
    return x, attn_1, attn_2


# --------------------------------------------------------------------------------------------------
# Models/decoder.py:43-56  Decoder.forward
# --------------------------------------------------------------------------------------------------
def decoder(sd, pfx: str, trg: Tensor, e_outputs: Tensor, src_mask, trg_mask, n_layers: int, heads: int,
            return_attn: bool = False, drop=None):
    x = decoder_prenet(sd, pfx + "decoder_prenet.", trg, drop)  # :44
    x = O.positional_encoder(x, sd[pfx + "pe.alpha"], drop, "dec")  # :47
    a1, a2 = [], []
    for i in range(n_layers):
        x, attn_1, attn_2 = decoder_layer(sd, f"{pfx}layers.{i}.", x, e_outputs, src_mask, trg_mask, heads, drop, f"dec.{i}")  # :53
        if return_attn:
            a1.append(attn_1)
            a2.append(attn_2)
    x = F.layer_norm(x, (x.shape[-1],), sd[pfx + "norm.weight"], sd[pfx + "norm.bias"])  # :56
    if return_attn:
        return x, torch.stack(a1, 1), torch.stack(a2, 1)
    return x, None, None


# --------------------------------------------------------------------------------------------------
# Models/transformer.py:90-118  Transformer.forward
# --------------------------------------------------------------------------------------------------
def encode(sd, hp: dict, src: Tensor, src_mask: Tensor, drop=None) -> Tensor:
    e_outputs, _ = O.encoder(sd, "encoder.", src, src_mask, hp["n_layer_encoder"], hp["n_head_encoder"], drop=drop)  # :92
    if "linear.weight" in sd:  # :93-94 (d_model_encoder != d_model_decoder)
        e_outputs = F.linear(e_outputs, sd["linear.weight"], sd["linear.bias"])
    return e_outputs


def transformer_forward(sd: Dict[str, Tensor], hp: dict, src: Tensor, trg: Tensor, src_mask: Tensor, trg_mask: Tensor,
                        return_attn: bool = False, e_outputs: Optional[Tensor] = None, drop=None) -> dict:
    """drop: None (every dropout probability 0) or `drop(name, shape) -> mask` for the named training-mode dropout call
    sites (fastspeech2_oracle._drop): "enc.*" of the text encoder, "prenet.1|2", "dec.pe", "dec.<i>.attn1|drop1|attn2|drop2|ff|drop3"."""
    if e_outputs is None:
        e_outputs = encode(sd, hp, src, src_mask, drop)
    d_output, a1, a2 = decoder(sd, "decoder.", trg, e_outputs, src_mask, trg_mask, hp["n_layer_decoder"],
                               hp["n_head_decoder"], return_attn, drop)  # :107
    outputs_prenet = F.linear(d_output, sd["out.weight"], sd["out.bias"])  # :109
    outputs_postnet = outputs_prenet  # :110 with postnets.py:64-79, prev_version=False: the input comes back
    stop_token = F.linear(d_output, sd["stop_token.weight"], sd["stop_token.bias"])  # :111 (.squeeze(2) is a no-op for r > 1)
    if stop_token.shape[2] == 1:
        stop_token = stop_token.squeeze(2)
    return dict(outputs_prenet=outputs_prenet, outputs_postnet=outputs_postnet, stop_token=stop_token, e_outputs=e_outputs,
                attn_dec_dec=a1, attn_dec_enc=a2)


# --------------------------------------------------------------------------------------------------
# test.py:128-191  the autoregressive loop, batched (the script runs batch 1), without the break on the stop token so
# that a fixed number of steps can be compared; `stop_step` reports where the script would have stopped.
# --------------------------------------------------------------------------------------------------
def greedy_decode(sd: Dict[str, Tensor], hp: dict, text: Tensor, pos_text: Tensor, n_steps: int) -> dict:
    """Runs decoder positions 0 .. n_steps-1 the reference's way: EVERY step re-runs the model on the whole prefix
    (test.py:139-141 for position 0, :151-156 inside the loop).  Returns the last forward's outputs:
    mel (B, n_steps * r, mel_dim) and stop logits (B, n_steps, r) — the prefix outputs of a causal model do not change
    from step to step, which is also what the script relies on when it rebuilds `mel_input` from `outputs_postnet`."""
    r = hp["reduction_rate"]
    mel_dim = hp["mel_dim"]
    B = text.shape[0]
    dev = text.device
    dummy = torch.zeros((B, 1, mel_dim), dtype=torch.float32, device=dev)  # :136
    mel_input = dummy
    src_mask, _ = create_masks(pos_text, None)
    e_outputs = encode(sd, hp, text, src_mask)  # (the script recomputes the encoder every step; it is deterministic)
    out = None
    for i in range(1, n_steps + 1):
        pos_mel = torch.arange(1, i + 1, dtype=torch.long, device=dev).unsqueeze(0).expand(B, -1)  # :138 / :150
        _, trg_mask = create_masks(pos_text, pos_mel)
        out = transformer_forward(sd, hp, text, mel_input, src_mask, trg_mask, e_outputs=e_outputs)
        post = out["outputs_postnet"]
        b, t, c = post.shape
        post = post.reshape(b, t * r, c // r)  # :143-146 / :161-164
        mel_input = torch.cat((dummy, post[:, ::r, :]), dim=1)  # :165-166
    stop = out["stop_token"].reshape(B, n_steps, -1)
    return dict(mel=post, stop_logits=stop, stop_step=stop_steps(stop), e_outputs=e_outputs)


def stop_steps(stop_logits: Tensor) -> Tensor:
    """test.py:168-174: the script stops after the first step i >= 2 (decoder position i - 1) whose stop probability,
    averaged over the r frames of the position, exceeds 0.5.  Returns that position per utterance (n_steps if none)."""
    B, T, _ = stop_logits.shape
    hit = torch.sigmoid(stop_logits).mean(dim=-1) > 0.5
    hit[:, 0] = False  # the loop starts at i = 2: position 0 is never tested
    idx = torch.arange(T, device=stop_logits.device).expand(B, T)
    return torch.where(hit, idx, torch.full_like(idx, T)).min(dim=1).values


def greedy_decode_cached(sd: Dict[str, Tensor], hp: dict, text: Tensor, pos_text: Tensor, n_steps: int,
                         forced_mel_in: Optional[Tensor] = None) -> dict:
    """The same decode evaluated incrementally (keys / values of earlier positions kept), still fp32 torch on the CPU:
    lets the tests compare long decodes in seconds.  tests/test_oracle_tts.py checks it against `greedy_decode`.
    forced_mel_in (B, n_steps, mel_dim): feed these decoder inputs instead of the model's own outputs."""
    r, mel_dim = hp["reduction_rate"], hp["mel_dim"]
    H, N = hp["n_head_decoder"], hp["n_layer_decoder"]
    B, dev = text.shape[0], text.device
    src_mask, _ = create_masks(pos_text, None)
    e = encode(sd, hp, text, src_mask)
    d = e.shape[-1]
    dk = d // H
    pe = O.positional_table(d).to(dev)
    alpha = sd["decoder.pe.alpha"]

    def heads(x):  # (B,t,d) -> (B,H,t,dk)
        return x.view(B, -1, H, dk).transpose(1, 2)

    cross = []
    for i in range(N):
        p = f"decoder.layers.{i}.attn_2."
        cross.append((heads(F.linear(e, sd[p + "k_linear.weight"], sd[p + "k_linear.bias"])),
                      heads(F.linear(e, sd[p + "v_linear.weight"], sd[p + "v_linear.bias"]))))
    kc = [None] * N
    vc = [None] * N
    mel_in = torch.zeros((B, 1, mel_dim), device=dev)
    mels, stops = [], []
    for t in range(n_steps):
        if forced_mel_in is not None:
            mel_in = forced_mel_in[:, t:t + 1]
        x = decoder_prenet(sd, "decoder.decoder_prenet.", mel_in) + alpha * pe[t]
        for i in range(N):
            p = f"decoder.layers.{i}."
            xn = F.layer_norm(x, (d,), sd[p + "norm_1.weight"], sd[p + "norm_1.bias"])
            q = heads(F.linear(xn, sd[p + "attn_1.q_linear.weight"], sd[p + "attn_1.q_linear.bias"]))
            k = heads(F.linear(xn, sd[p + "attn_1.k_linear.weight"], sd[p + "attn_1.k_linear.bias"]))
            v = heads(F.linear(xn, sd[p + "attn_1.v_linear.weight"], sd[p + "attn_1.v_linear.bias"]))
            kc[i] = k if kc[i] is None else torch.cat((kc[i], k), dim=2)
            vc[i] = v if vc[i] is None else torch.cat((vc[i], v), dim=2)
            ctx, _ = O.attention(q, kc[i], vc[i], dk, None)
            ctx = ctx.transpose(1, 2).reshape(B, 1, d)
            x = x + F.linear(ctx, sd[p + "attn_1.out.weight"], sd[p + "attn_1.out.bias"])
            xn = F.layer_norm(x, (d,), sd[p + "norm_2.weight"], sd[p + "norm_2.bias"])
            q = heads(F.linear(xn, sd[p + "attn_2.q_linear.weight"], sd[p + "attn_2.q_linear.bias"]))
            ctx, _ = O.attention(q, cross[i][0], cross[i][1], dk, src_mask)
            ctx = ctx.transpose(1, 2).reshape(B, 1, d)
            x = x + F.linear(ctx, sd[p + "attn_2.out.weight"], sd[p + "attn_2.out.bias"])
            xn = F.layer_norm(x, (d,), sd[p + "norm_3.weight"], sd[p + "norm_3.bias"])
            x = x + O.feed_forward(sd, p + "ff.", xn)
        dn = F.layer_norm(x, (d,), sd["decoder.norm.weight"], sd["decoder.norm.bias"])
        o = F.linear(dn, sd["out.weight"], sd["out.bias"])  # (B,1,mel_dim*r)
        mels.append(o.reshape(B, r, mel_dim))
        stops.append(F.linear(dn, sd["stop_token.weight"], sd["stop_token.bias"]).reshape(B, 1, r))
        mel_in = o[:, :, :mel_dim]  # frame 0 of the position's r frames (test.py:165: outputs_postnet[:, ::r, :])
    mel = torch.cat(mels, dim=1)
    stop = torch.cat(stops, dim=1)
    return dict(mel=mel, stop_logits=stop, stop_step=stop_steps(stop), e_outputs=e)
