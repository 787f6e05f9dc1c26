"""FastSpeech2 training step on the sm_100a kernels.

What the reference does in ``train_fastspeech2.py:117-315`` with autocast + autograd + DDP/SyncBatchNorm +
``torch.optim.Adam`` + Noam LR + ``clip_grad_norm_`` is restated here as an explicit kernel sequence:

    forward  (training mode: saves pre-LayerNorm values, attention log-sum-exp, BatchNorm batch statistics)
    losses   (five unmasked nn.L1Loss terms, train_fastspeech2.py:212-259) and their gradients in one kernel each
    backward (dgrad GEMMs = conv_gemm with transposed tap-flipped weights; wgrad = tcgen05 MN-major split-K GEMM;
              attention backward = two tcgen05 kernels; LayerNorm / BatchNorm / LengthRegulator / embedding
              backward row kernels), every parameter gradient lands in ONE flat fp32 buffer
    all-reduce of that buffer over NCCL (data parallel, one process per GPU), SyncBatchNorm statistics
    optimizer (global-norm clip + Adam over the flat buffer, Noam learning rate)

PyTorch is plumbing only (buffers, streams, torch.distributed).  There is no autograd graph and no fallback.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import lib
from .Models import _engine as E


def noam_lr(step: int, d_model: int, warmup_factor: float, warmup_step: int) -> float:
    """utils/utils.py:204-215 (get_learning_rate)."""
    return warmup_factor * min(step ** -0.5, step * warmup_step ** -1.5) * (d_model ** -0.5)


def _t(w_taps: torch.Tensor) -> torch.Tensor:
    """Forward operand (taps, N, C) -> dgrad operand (taps, C, N), taps flipped: w'[j'][c][n] = w[taps-1-j'][n][c]."""
    wt = w_taps.permute(0, 2, 1)
    if w_taps.shape[0] > 1:  # a linear (one tap) needs no flip: one copy kernel instead of two
        wt = wt.flip(0)
    return wt.contiguous()


def _src(d, k):
    """d[k], or d[k0][k1]... for a tuple path."""
    if isinstance(k, tuple):
        for kk in k:
            d = d[kk]
        return d
    return d[k]


class DropoutPlan:
    """Dropout call sites of the reference's training forward and their probabilities.

    The reference constructs the model with dropout = hp.dropout (0.1), dropout_postnet = 0.5 and
    dropout_variance_adaptor (train_fastspeech2.py:381-389).  Every call site of one module instance gets a `site`
    number; a mask is a pure function of (seed, site, element index) (csrc/common.cuh), so the forward kernels apply
    it and the backward kernels regenerate it.  The seed advances with every training forward.  Sites per FFT block:
    attention probabilities (Models/modules.py:18-19), dropout_1 on the attention output (layers.py:33), the
    FeedForward's own dropout before its LayerNorm (modules.py:86), dropout_2 on the FeedForward output (layers.py:40);
    per stack the PositionalEncoder's (modules.py:111); per VariancePredictor dropout1 / dropout2
    (varianceadaptor.py:219,222); per PostNet conv the dropout after tanh (postnets.py:71-73)."""

    def __init__(self, model, base_seed: int = 0):
        self.base_seed = int(base_seed)
        self.calls = 0
        self._sites: Dict[str, int] = {}
        enc, dec, va, pn = model.encoder, model.decoder, model.variance_adaptor, model.postnet
        self.p_stack = {"enc": float(enc.pe.dropout.p), "dec": float(dec.pe.dropout.p)}
        for tag, st in (("enc", enc), ("dec", dec)):
            for l in st.layers:
                ps = {float(l.dropout_1.p), float(l.dropout_2.p), float(l.ff.dropout.p), float(l.attn.dropout or 0.0),
                      self.p_stack[tag]}
                if len(ps) != 1:
                    raise NotImplementedError(f"{tag}: one dropout probability per FFT stack (the reference passes a "
                                              f"single `dropout` to all of its call sites), got {sorted(ps)}")
        self.p_pred = {n: float(getattr(va, n + "_predictor").dropout1.p) for n in ("duration", "pitch", "energy")}
        for n in self.p_pred:
            if float(getattr(va, n + "_predictor").dropout2.p) != self.p_pred[n]:
                raise NotImplementedError("VariancePredictor: dropout1 and dropout2 share one probability in the reference")
        self.p_post = float(pn.dropout1.p)
        if any(float(d.p) != self.p_post for d in pn.dropout_list):
            raise NotImplementedError("PostConvNet: one dropout probability")
        for name, p in list(self.p_stack.items()) + list(self.p_pred.items()) + [("postnet", self.p_post)]:
            if not 0.0 <= p < 1.0:
                raise ValueError(f"dropout probability of {name} must be in [0, 1), got {p}")

    @property
    def active(self) -> bool:
        return any(p > 0 for p in self.p_stack.values()) or any(p > 0 for p in self.p_pred.values()) or self.p_post > 0

    def site(self, name: str) -> int:
        return self._sites.setdefault(name, len(self._sites))

    def prob(self, name: str) -> float:
        """Probability of the call site `name` ("enc.3.attn", "dec.pe", "pitch.drop1", "postnet.2", ...)."""
        head = name.split(".")[0]
        if head in self.p_stack:
            return self.p_stack[head]
        if head in self.p_pred:
            return self.p_pred[head]
        if head == "postnet":
            return self.p_post
        raise KeyError(name)

    def mask(self, name: str, shape, seed: int, device="cuda") -> torch.Tensor:
        """The mask (0 or 1/(1-p)) the kernels apply at `name` for `seed`, materialised (tests / debugging only)."""
        n = 1
        for d in shape:
            n *= int(d)
        return lib.dropout_mask(n, self.prob(name), seed, self.site(name), device=device).view(*shape)

    def next_seed(self) -> int:
        self.calls += 1
        return (self.base_seed * 0x9E3779B97F4A7C15 + self.calls * 0xD1B54A32D192ED03) & 0xFFFFFFFFFFFFFFFF


def allreduce_mean_scale(flat_g: torch.Tensor, world: int, group=None) -> float:
    """Data-parallel gradient exchange: ONE all-reduce (SUM) over the flat gradient buffer; the mean is
    not applied here but returned as the scale the optimizer kernel folds into its gradient read."""
    if world > 1:
        dist.all_reduce(flat_g, group=group)
    return 1.0 / world


def allreduce_tail_begin(flat_g: torch.Tensor, tail_start: int, group=None):
    """First half of the overlapped exchange: asynchronous all-reduce (SUM) of flat_g[tail_start:], the part of the
    gradient buffer that is final once the decoder stack has run backward.  Returns the work handle."""
    return dist.all_reduce(flat_g[tail_start:], group=group, async_op=True)


def allreduce_tail_finish(flat_g: torch.Tensor, tail_start: int, work, world: int, group=None) -> float:
    """Second half: all-reduce the head flat_g[:tail_start], join the tail; returns the mean scale like
    `allreduce_mean_scale`."""
    dist.all_reduce(flat_g[:tail_start], group=group)
    work.wait()
    return 1.0 / world


def sync_bn_moments(stats: torch.Tensor, rows_local: int, world: int, sync: bool, eps: float, group=None):
    """[sum x | sum x^2] (2,C) of the local rows -> ([mean | rstd] (2,C), biased var, global row count).  With
    `sync` the sums are all-reduced first, which is what SyncBatchNorm's forward does."""
    count = float(rows_local)
    if sync and world > 1:
        dist.all_reduce(stats, group=group)
        count *= world
    mean = stats[0] / count
    var = (stats[1] / count - mean * mean).clamp_(min=0.0)
    return torch.stack([mean, torch.rsqrt(var + eps)]).contiguous(), var, count


class Trainer:
    """Owns the flat parameter / gradient / Adam-state buffers of a ``FastSpeech2`` and runs training steps."""

    def __init__(self, model, *, lr_max: float = 1e-3, betas=(0.9, 0.98), eps: float = 1e-9, clip: float = 1.0,
                 warmup_step: int = 4000, warmup_factor: float = 1.0, fixed_lr: Optional[float] = None,
                 process_group=None, sync_bn: bool = True, bn_momentum: float = 0.1, flatten: bool = True,
                 optimizer: str = "adam", overlap_allreduce: bool = False, dropout_seed: int = 0,
                 cuda_graph: bool = False):
        """flatten=True (training loop owner): parameters are re-pointed into one flat fp32 buffer, `p.grad` are
        views of the flat gradient buffer and optimizer_step() runs the fused clip+Adam kernel.
        flatten=False (drop-in autograd mode, see `autograd_forward`): parameters are left alone, only the flat
        gradient buffer exists and the caller's torch optimizer owns the update.
        cuda_graph=True: train_step() captures zero_grad + forward + losses + backward (weight-operand refresh included) as ONE
        CUDA graph per input shape and replays it; the optimizer step (learning rate and bias corrections change per step)
        stays three eager launches.  Used only where a replay is the same computation as an eager step: dropout inactive (the
        mask seed is a kernel argument, a replay would repeat the masks) and a single process — capturing the SyncBatchNorm
        moment all-reduces that sit inside the forward and backward was tried on 2 GPUs and hung (NCCL 2.28 under
        torch.cuda.graph), so data-parallel steps stay eager.  If a capture fails the trainer falls back to eager steps."""
        lib.load()
        self.cuda_graph = bool(cuda_graph)
        self._graphs: Dict[tuple, tuple] = {}
        self.drop = self._dropout_plan(model, dropout_seed)
        # hp.optimizer (train_fastspeech2.py:412-416): 'radam' -> radam.RAdam at the fixed rate lr_max (the Noam
        # schedule is skipped for it, :117-120); anything else -> torch.optim.Adam at the Noam rate
        self.optimizer = optimizer.lower()
        if self.optimizer not in ("adam", "noam", "radam"):
            raise ValueError(f"Trainer: optimizer {optimizer!r} is not built (adam / noam / radam)")
        self.flatten = flatten
        self.overlap_allreduce = overlap_allreduce
        self.model = model
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if (dist.is_available() and dist.is_initialized()) else 1
        self.sync_bn = sync_bn and self.world > 1
        self.betas, self.eps, self.clip, self.lr_max = betas, eps, clip, lr_max
        self.warmup_step, self.warmup_factor, self.fixed_lr = warmup_step, warmup_factor, fixed_lr
        self.bn_momentum = bn_momentum
        self.step_count = 0
        params = [(n, p) for n, p in model.named_parameters()]
        if not params or not params[0][1].is_cuda:
            raise RuntimeError("Trainer: move the model to a CUDA device first (there is no CPU path)")
        dev = params[0][1].device
        self.dev = dev
        # every tensor starts on an 8-element boundary inside the flat buffers: 16 bytes in the bf16 mirror (TMA bases), 32 in
        # the fp32 buffers (vector loads).  The q | k | v weights (and biases) of an attention module sit next to each other,
        # so their concatenated GEMM operand is a VIEW of the mirror, not a torch.cat per step.
        offs, total = {}, 0
        sizes = dict((n, p.numel()) for n, p in params)
        for n in self._flat_order([n for n, _ in params]):
            offs[n] = total
            total += (sizes[n] + 7) // 8 * 8
        self.params = [p for _, p in params]
        self.offsets = [(offs[n], p.numel()) for n, p in params]
        self.total = total
        self.grads: Dict[int, torch.Tensor] = {}
        self.names: Dict[int, str] = {id(p): n for n, p in params}
        self.flat_g = torch.zeros(total, device=dev, dtype=torch.float32)
        # Gradient exchange overlapped with the backward: the backward runs PostNet -> decoder -> variance adaptor ->
        # encoder and named_parameters() lists encoder, variance_adaptor, decoder, postnet, so once the decoder stack
        # is done the TAIL of the flat buffer (decoder + postnet, ~45 % of the elements) is final and its all-reduce
        # can run on NCCL's stream under the rest of the backward.  tail_start = None disables the overlap.
        # Off by default: at 2 GPUs it measured 0.7 % SLOWER (16.54 vs 16.43 ms/step) — the NCCL kernel takes SMs
        # from persistent backward kernels sized for all 148, and the whole exchange is only 0.6 ms of the step.
        names = [n for n, _ in params]
        first = next((i for i, n in enumerate(names) if n.startswith(("decoder.", "postnet."))), None)
        self.tail_start = None
        if first is not None and first > 0 and all(n.startswith(("decoder.", "postnet.")) for n in names[first:]):
            self.tail_start = offs[names[first]]
        self._overlap = False      # armed by train_step() only: an accumulating caller must not reduce early
        self._pending = None
        self._bind_grads()
        self.flat_p16 = None
        self._mirrors = []
        if flatten:
            self.flat_p = torch.zeros(total, device=dev, dtype=torch.float32)
            # bf16 mirror of the parameters: refreshed by ONE cast kernel whenever the flat buffer changes (optimizer step,
            # broadcast); Models/_engine builds the GEMM operands as views of it instead of per-tensor casts / cats
            self.flat_p16 = torch.zeros(total, device=dev, dtype=torch.bfloat16)
            self.flat_m = torch.zeros(total, device=dev, dtype=torch.float32)
            self.flat_v = torch.zeros(total, device=dev, dtype=torch.float32)
            with torch.no_grad():
                for n, p in params:
                    o, k = offs[n], p.numel()
                    self.flat_p[o:o + k].copy_(p.detach().reshape(-1).float())
                    p.data = self.flat_p[o:o + k].view(p.shape)  # parameters become views of the flat buffer
                    p.grad = self.grads[id(p)]
                    self._mirrors.append((p, self.flat_p16[o:o + k].view(p.shape)))
        self.sumsq = torch.zeros(1, device=dev, dtype=torch.float32)
        self.loss_buf = torch.zeros(8, device=dev, dtype=torch.float32)
        self._zeros = {}
        # dgrad operands (transposed, tap-flipped weights): the (cache dict, key) pairs a backward pass asks for are
        # recorded on the first pass; from then on ONE batched kernel builds them all at the start of the backward
        self._wt_pairs: Dict[tuple, tuple] = {}
        self._wt_cache: Dict[tuple, tuple] = {}
        self.invalidate_caches()

    @staticmethod
    def _flat_order(names):
        """named_parameters() order, except that the six tensors {q,k,v}_linear.{weight,bias} of one attention module are
        emitted together as q.w k.w v.w q.b k.b v.b at the position of the first of them."""
        have, done, out = set(names), set(), []
        for n in names:
            if n in done:
                continue
            for tail in ("q_linear.weight", "k_linear.weight", "v_linear.weight", "q_linear.bias", "k_linear.bias",
                         "v_linear.bias"):
                if n.endswith("." + tail) or n == tail:
                    pre = n[: len(n) - len(tail)]
                    grp = [pre + f"{x}_linear.{y}" for y in ("weight", "bias") for x in ("q", "k", "v")]
                    if all(g in have and g not in done for g in grp):
                        out += grp
                        done.update(grp)
                    break
            if n not in done:
                out.append(n)
                done.add(n)
        return out

    def _refresh_mirror(self) -> None:
        if self.flat_p16 is None:
            return
        lib.cast_bf16(self.flat_p, out=self.flat_p16)
        for p, m in self._mirrors:
            p._ttsb_bf16 = (m, p._version)

    def _dropout_plan(self, model, seed: int):
        """The dropout call sites of the model (FastSpeech2 here; tts_training.TTSTrainer overrides it)."""
        return DropoutPlan(model, base_seed=seed)

    # ------------------------------------------------------------------------------------------
    def _bind_grads(self) -> None:
        for p, (o, k) in zip(self.params, self.offsets):
            self.grads[id(p)] = self.flat_g[o:o + k].view(p.shape)

    def fresh_grads(self) -> None:
        """A new zeroed gradient buffer (autograd mode hands the previous one to the caller's optimizer)."""
        self.flat_g = torch.zeros(self.total, device=self.dev, dtype=torch.float32)
        self._bind_grads()

    def G(self, p: torch.Tensor) -> torch.Tensor:
        return self.grads[id(p)]

    def zeros(self, n: int) -> torch.Tensor:
        z = self._zeros.get(n)
        if z is None:
            z = torch.zeros(n, device=self.dev, dtype=torch.float32)
            self._zeros[n] = z
        return z

    def _wt(self, d, k) -> torch.Tensor:
        """The dgrad operand of the forward operand d[k] (bf16 (taps, N, C)): (taps, C, N), taps flipped.  `d` must be a
        persistent cache dict (Models/_engine caches): the pair is remembered for `_prepare_wt`."""
        src = _src(d, k)
        key = (id(d), k)
        ent = self._wt_cache.get(key)
        if ent is not None and ent[0] is src:
            return ent[1]
        if key not in self._wt_pairs and len(self._wt_pairs) < 512:
            self._wt_pairs[key] = (d, k)
        return _t(src)

    def _wgrad_rows(self, dy: torch.Tensor, x: torch.Tensor, linears, n_off: int = 0) -> None:
        """Weight and bias gradients of Linear layers whose outputs are consecutive column blocks of `dy` (q | k | v).  The
        flat layout places their parameters — hence their gradient views — next to each other, so the blocks are the rows of
        ONE (sum N, C) matrix: one wgrad launch that reads x once, instead of one per layer."""
        gw = [self.G(l.weight) for l in linears]
        gb = [self.G(l.bias) for l in linears]
        if len(linears) > 1 and E._adjacent(gw) and E._adjacent(gb):
            lib.conv_wgrad(dy, x, E._cat_rows(gw), n_off=n_off, N=sum(g.shape[0] for g in gw), dbias=E._cat_rows(gb))
            return
        for l, w, b in zip(linears, gw, gb):
            lib.conv_wgrad(dy, x, w, n_off=n_off, N=w.shape[0], dbias=b)
            n_off += w.shape[0]

    def _prepare_wt(self) -> None:
        """All dgrad operands of this backward pass in one launch (ttsb200_weight_transpose) instead of one or two torch
        copy kernels per weight.  Call after the forward (the forward operands are fresh then)."""
        if not self._wt_pairs:
            return
        keys, srcs, dsts = [], [], []
        for key, (d, k) in self._wt_pairs.items():
            try:
                src = _src(d, k)
            except (KeyError, IndexError):
                continue
            ent = self._wt_cache.get(key)
            if ent is not None and ent[0] is src:
                continue
            taps, N, Cin = src.shape
            dst = ent[1] if ent is not None and tuple(ent[1].shape) == (taps, Cin, N) else \
                torch.empty((taps, Cin, N), device=src.device, dtype=torch.bfloat16)
            keys.append(key); srcs.append(src); dsts.append(dst)
        if srcs:
            lib.weight_transpose(srcs, dsts)
            for key, src, dst in zip(keys, srcs, dsts):
                self._wt_cache[key] = (src, dst)

    def invalidate_caches(self) -> None:
        """The optimizer kernel updates parameters through raw pointers: derived bf16 operands are stale."""
        for m in self.model.modules():
            c = getattr(m, "_ttsb_cache", None)
            if c is not None:
                c.sig = None
        self._refresh_mirror()

    def lr(self) -> float:
        if self.fixed_lr is not None:
            return self.fixed_lr
        if self.optimizer == "radam":
            return self.lr_max
        d_model = self.model.decoder.norm.weight.numel()
        return noam_lr(max(self.step_count, 1), d_model, self.warmup_factor, self.warmup_step)

    # ------------------------------------------------------------------------------------------
    # FFT stack (Models/encoder.py:83-112, layers.py:29-41)
    # ------------------------------------------------------------------------------------------
    def _stack_fwd(self, enc, src, mask_u8, kv, tag: str = "enc", seed: int = 0):
        t = enc._cache()
        B, T = src.shape[0], src.shape[1]
        D = enc.norm.weight.numel()
        dev = src.device
        pe = enc.pe.table(dev)
        if T > pe.shape[0]:
            raise ValueError(f"sequence length {T} exceeds the positional table ({pe.shape[0]})")
        caches = [E.layer_cache(l) for l in enc.layers]
        x = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        xn = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
        ln_first = caches[0]["ln1"]
        pd = self.drop.p_stack[tag]          # one probability for every dropout call site of this stack
        S = lambda name: self.drop.site(f"{tag}.{name}")  # noqa: E731
        if enc.embedding:
            lib.embed_pe_ln(src.contiguous(), t["table"], pe, t["alpha"], ln_first[0], ln_first[1], E.LN_EPS, x, xn)
        else:
            lib.conv_gemm(src, t["wemb"], t["bemb"], block_n=D, epi=lib.EPI_RES_LN, res_f32=pe, res_bstride=0,
                          res_tstride=D, res_scale=t["alpha"], ln1=ln_first, out_f32=x, out_bf16=xn)
        if pd > 0:  # PositionalEncoder's dropout (modules.py:111): x0 := drop(x0), then the first norm_1 again
            lib.dropout_(x, pd, seed, S("pe"))
            lib.layernorm(x, ln_first[0], ln_first[1], E.LN_EPS, y_bf16=xn)
        tape = dict(src=src, mask=mask_u8, kv=kv, pe=pe, T=T, blocks=[], caches=caches, top=t, tag=tag, seed=seed, pd=pd)
        H = enc.heads
        for i, lc in enumerate(caches):
            k = lc["k"]
            next_ln = t["norm"] if i == enc.N - 1 else caches[i + 1]["ln1"]
            qkv = torch.empty((B, T, 3 * D), device=dev, dtype=torch.bfloat16)
            lib.conv_gemm(xn, lc["wqkv"], lc["bqkv"], block_n=192, out_bf16=qkv)
            ctx = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
            lse = torch.empty((B, H, T), device=dev, dtype=torch.float32)
            lib.attention_fwd(qkv, mask_u8, kv, ctx, H, lse2=lse, drop=(pd, seed, S(f"{i}.attn")))
            y = torch.empty((B, T, D), device=dev, dtype=torch.float32)
            yn = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
            h = torch.empty((B, T, 4 * D), device=dev, dtype=torch.bfloat16)
            upre = torch.empty((B, T, D), device=dev, dtype=torch.float32)
            if pd > 0:
                # With dropout between the GEMM and its residual / LayerNorm the fused epilogues do not apply: the GEMMs
                # leave acc + bias in fp32 and one fp32 row kernel per half block does dropout + residual + LayerNorm.
                o = torch.empty((B, T, D), device=dev, dtype=torch.float32)
                lib.conv_gemm(ctx, lc["wo"], lc["bo"], block_n=128, out_f32=o)
                lib.xrow(lib.XROW_ADD_LN, x, lc["ln2"], b0=o, s_out=y, out_bf16=yn, eps=E.LN_EPS, drop_p=pd,
                         drop_seed=seed, site_b=S(f"{i}.drop1"))                                     # layers.py:33,35
                lib.conv_gemm(yn, lc["w1"], lc["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU, out_bf16=h)
                u = torch.empty((B, T, D), device=dev, dtype=torch.float32)
                lib.conv_gemm(h, lc["w2"], lc["b2"], taps=k, pad_left=k // 2, block_n=128, out_f32=u)
                z = torch.empty((B, T, D), device=dev, dtype=torch.float32)
                zn = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
                lib.xrow(lib.XROW_FFN_TAIL, u, lc["ln_ff"], b0_bf16=yn, c=y, ln2=next_ln, s_out=upre, z_out=z, out_bf16=zn,
                         eps=E.LN_EPS, drop_p=pd, drop_seed=seed, site_s=S(f"{i}.ff"),
                         site_t=S(f"{i}.drop2"))                                    # modules.py:85-87, layers.py:40
            else:
                lib.conv_gemm(ctx, lc["wo"], lc["bo"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x, res_bstride=T * D,
                              res_tstride=D, ln1=lc["ln2"], out_f32=y, out_bf16=yn)
                lib.conv_gemm(yn, lc["w1"], lc["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU, out_bf16=h)
                z, zn = E.run_ffn2(lc, h, yn, y, next_ln, pre_f32=upre)
            tape["blocks"].append(dict(x=x, xn=xn, qkv=qkv, ctx=ctx, lse=lse, y=y, yn=yn, h=h, upre=upre))
            x, xn = z, zn
        tape["z_last"] = x
        return xn, tape

    def _stack_bwd(self, enc, tape, g_out: torch.Tensor) -> torch.Tensor:
        """g_out: fp32 gradient w.r.t. the stack output LN_norm(z_last).  Returns the fp32 gradient w.r.t. the
        stack input features (decoder) or None (text encoder: the embedding table gradient is accumulated)."""
        G = self.G
        B, T, D = g_out.shape
        dev = g_out.device
        H = enc.heads
        mask, kv = tape["mask"], tape["kv"]
        pd, seed, tag = tape["pd"], tape["seed"], tape["tag"]
        S = lambda name: self.drop.site(f"{tag}.{name}") if pd > 0 else -1  # noqa: E731
        dk = dict(drop_p=pd, drop_seed=seed)
        f32 = lambda *s: torch.empty(s, device=dev, dtype=torch.float32)  # noqa: E731
        bf = lambda *s: torch.empty(s, device=dev, dtype=torch.bfloat16)  # noqa: E731
        dz = f32(B, T, D)
        lib.layernorm_bwd(tape["z_last"], tape["top"]["norm"][0], g_f32=g_out, dx_f32=dz, dgamma=G(enc.norm.weight),
                          dbeta=G(enc.norm.bias))
        dz_bf = None
        for i in range(enc.N - 1, -1, -1):
            layer, lc, s = enc.layers[i], tape["caches"][i], tape["blocks"][i]
            a, ff = layer.attn, layer.ff
            k = lc["k"]
            # z = y + LN_ff(upre)
            dupre, dupre_bf = f32(B, T, D), bf(B, T, D)
            # dropout: LN_ff's output was dropped (dropout_2), its input is the FeedForward dropout's output
            lib.layernorm_bwd(s["upre"], lc["ln_ff"][0], g_f32=dz, dx_f32=dupre, dx_bf16=dupre_bf,
                              dgamma=G(ff.layer_norm.weight), dbeta=G(ff.layer_norm.bias), site_g=S(f"{i}.drop2"),
                              site_dx=S(f"{i}.ff"), dx_mode=1 if pd > 0 else 0, **dk)
            # upre = conv2(h) + b2 + yn
            lib.conv_wgrad(dupre_bf, s["h"], G(ff.f_2.weight), taps=k, pad_left=k // 2, conv_layout=True,
                           dbias=G(ff.f_2.bias))
            dh = bf(B, T, 4 * D)
            lib.conv_gemm(dupre_bf, self._wt(lc, "w2"), self.zeros(4 * D), taps=k, pad_left=k - 1 - k // 2, block_n=256,
                          act=lib.ACT_RELU_GRAD, gate_bf16=s["h"], out_bf16=dh)
            # h = relu(conv1(yn) + b1)
            lib.conv_wgrad(dh, s["yn"], G(ff.f_1.weight), taps=k, pad_left=k // 2, conv_layout=True,
                           dbias=G(ff.f_1.bias))
            dyn = f32(B, T, D)
            lib.conv_gemm(dh, self._wt(lc, "w1"), self.zeros(D), taps=k, pad_left=k - 1 - k // 2, block_n=192,
                          res_f32=dupre, res_bstride=T * D, res_tstride=D, out_f32=dyn)
            # yn = norm_2(y); z = y + ff(yn)
            dy, dy_bf = f32(B, T, D), bf(B, T, D)
            # dropout: dy_bf feeds the gradients of the attention branch, whose output dropout_1 masked (dx_mode 2)
            lib.layernorm_bwd(s["y"], lc["ln2"][0], g_f32=dyn, add_f32=dz, dx_f32=dy, dx_bf16=dy_bf,
                              dgamma=G(layer.norm_2.weight), dbeta=G(layer.norm_2.bias), site_dx=S(f"{i}.drop1"),
                              dx_mode=2 if pd > 0 else 0, **dk)
            # y = x + ctx Wo^T + bo
            lib.conv_wgrad(dy_bf, s["ctx"], G(a.out.weight), dbias=G(a.out.bias))
            dctx = bf(B, T, D)
            lib.conv_gemm(dy_bf, self._wt(lc, "wo"), self.zeros(D), block_n=192, out_bf16=dctx)
            delta = lib.attn_delta(dctx, s["ctx"], H)
            dqkv = lib.attention_bwd(s["qkv"], dctx, mask, kv, s["lse"], delta, H, drop=(pd, seed, S(f"{i}.attn")))
            self._wgrad_rows(dqkv, s["xn"], (a.q_linear, a.k_linear, a.v_linear))
            dxn = f32(B, T, D)
            lib.conv_gemm(dqkv, self._wt(lc, "wqkv"), self.zeros(D), block_n=192, out_f32=dxn)
            # xn = norm_1(x); y = x + attn(xn)
            dx = f32(B, T, D)
            dz_bf = bf(B, T, D) if i == 0 else None
            # dropout: the first block's input is the PositionalEncoder dropout's output
            first = i == 0 and pd > 0
            lib.layernorm_bwd(s["x"], lc["ln1"][0], g_f32=dxn, add_f32=dy, dx_f32=dx, dx_bf16=dz_bf,
                              dgamma=G(layer.norm_1.weight), dbeta=G(layer.norm_1.bias),
                              site_dx=S("pe") if first else -1, dx_mode=1 if first else 0, **dk)
            dz = dx
        # x0 = embed(src) + alpha * pe
        t = tape["top"]
        if enc.embedding:
            lib.embed_bwd(dz, T, idx=tape["src"], table=G(enc.embed.weight), skip_idx=0, pe=tape["pe"],
                          dalpha=G(enc.pe.alpha))
            return None
        lib.embed_bwd(dz, T, pe=tape["pe"], dalpha=G(enc.pe.alpha))
        lib.colsum(dz, G(enc.embed.bias))
        lib.conv_wgrad(dz_bf, tape["src"], G(enc.embed.weight))
        dsrc = f32(B, T, tape["src"].shape[2])
        lib.conv_gemm(dz_bf, self._wt(t, "wemb"), self.zeros(tape["src"].shape[2]), block_n=192, out_f32=dsrc)
        return dsrc

    # ------------------------------------------------------------------------------------------
    # VariancePredictor (Models/varianceadaptor.py:216-231)
    # ------------------------------------------------------------------------------------------
    def _pred_fwd(self, pred, x_bf16, mask_u8, name: str = "duration", seed: int = 0):
        t = E.predictor_cache(pred)
        B, T, _ = x_bf16.shape
        fs = t["b1"].numel()
        dev = x_bf16.device
        pd = self.drop.p_pred[name]
        h = torch.empty((B, T, fs), device=dev, dtype=torch.bfloat16)
        r1 = torch.empty((B, T, fs), device=dev, dtype=torch.float32)
        out = torch.empty((B, T), device=dev, dtype=torch.float32)
        r2 = torch.empty((B, T, fs), device=dev, dtype=torch.float32)
        if pd > 0:  # dropout after each LayerNorm (varianceadaptor.py:219,222): GEMM + ReLU in fp32, then a row kernel
            s1, s2 = self.drop.site(f"{name}.drop1"), self.drop.site(f"{name}.drop2")
            lib.conv_gemm(x_bf16, t["w1"], t["b1"], taps=3, pad_left=1, block_n=128, act=lib.ACT_RELU, out_f32=r1)
            lib.xrow(lib.XROW_ADD_LN, r1, t["ln1"], out_bf16=h, eps=E.LN_EPS, drop_p=pd, drop_seed=seed, site_t=s1)
            lib.conv_gemm(h, t["w2"], t["b2"], taps=3, pad_left=1, block_n=128, act=lib.ACT_RELU, out_f32=r2)
            lib.xrow(lib.XROW_LN_DOT, r2, t["ln2"], dot_w=t["dw"], dot_b=t["db"], row_mask=mask_u8, out_row=out,
                     eps=E.LN_EPS, drop_p=pd, drop_seed=seed, site_t=s2)
        else:
            s1 = s2 = -1
            lib.conv_gemm(x_bf16, t["w1"], t["b1"], taps=3, pad_left=1, block_n=fs, epi=lib.EPI_RELU_LN, ln1=t["ln1"],
                          out_bf16=h, pre_f32=r1)
            lib.conv_gemm(h, t["w2"], t["b2"], taps=3, pad_left=1, block_n=fs, epi=lib.EPI_RELU_LN_DOT, ln1=t["ln2"],
                          dot_w=t["dw"], dot_b=t["db"], row_mask=mask_u8, out_row=out, pre_f32=r2)
        return out, dict(x=x_bf16, h=h, r1=r1, r2=r2, mask=mask_u8, t=t, dk=dict(drop_p=pd, drop_seed=seed), s1=s1, s2=s2)

    def _pred_bwd(self, pred, tape, g_out: torch.Tensor, add_f32: Optional[torch.Tensor]) -> torch.Tensor:
        """g_out (B,T) fp32 -> fp32 gradient w.r.t. the predictor input (+ add_f32)."""
        G = self.G
        t = tape["t"]
        B, T, Din = tape["x"].shape
        fs = t["b1"].numel()
        dev = g_out.device
        dc2 = torch.empty((B, T, fs), device=dev, dtype=torch.bfloat16)
        lib.layernorm_bwd(tape["r2"], t["ln2"][0], beta=t["ln2"][1], g_row=g_out.contiguous(), g_vec=t["dw"],
                          row_mask=tape["mask"], relu_gate=True, dx_bf16=dc2, dgamma=G(pred.layer_norm2.weight),
                          dbeta=G(pred.layer_norm2.bias), dvec=G(pred.linear_layer.weight),
                          dscal=G(pred.linear_layer.bias), site_g=tape["s2"], **tape["dk"])
        lib.conv_wgrad(dc2, tape["h"], G(pred.conv2.weight), taps=3, pad_left=1, conv_layout=True,
                       dbias=G(pred.conv2.bias))
        dh = torch.empty((B, T, fs), device=dev, dtype=torch.float32)
        lib.conv_gemm(dc2, self._wt(t, "w2"), self.zeros(fs), taps=3, pad_left=1, block_n=128, out_f32=dh)
        dc1 = torch.empty((B, T, fs), device=dev, dtype=torch.bfloat16)
        lib.layernorm_bwd(tape["r1"], t["ln1"][0], g_f32=dh, relu_gate=True, dx_bf16=dc1,
                          dgamma=G(pred.layer_norm1.weight), dbeta=G(pred.layer_norm1.bias), site_g=tape["s1"],
                          **tape["dk"])
        lib.conv_wgrad(dc1, tape["x"], G(pred.conv1.weight), taps=3, pad_left=1, conv_layout=True,
                       dbias=G(pred.conv1.bias))
        dx = torch.empty((B, T, Din), device=dev, dtype=torch.float32)
        kw = {} if add_f32 is None else dict(res_f32=add_f32, res_bstride=T * Din, res_tstride=Din)
        lib.conv_gemm(dc1, self._wt(t, "w1"), self.zeros(Din), taps=3, pad_left=1, block_n=192, out_f32=dx, **kw)
        return dx

    # ------------------------------------------------------------------------------------------
    # PostConvNet with BatchNorm batch statistics (Models/postnets.py:64-79, nn.BatchNorm1d in train())
    # ------------------------------------------------------------------------------------------
    def _bn_finish_stats(self, stats: torch.Tensor, rows_local: int, bn) -> tuple:
        """[sum | sum^2] -> [mean | rstd] (+ running-stat update as nn.BatchNorm1d / SyncBatchNorm do)."""
        ms, var, count = sync_bn_moments(stats, rows_local, self.world, self.sync_bn, bn.eps, self.pg)
        mean = ms[0]
        with torch.no_grad():
            m = self.bn_momentum
            bn.running_mean.mul_(1 - m).add_(mean, alpha=m)
            bn.running_var.mul_(1 - m).add_(var * (count / max(count - 1.0, 1.0)), alpha=m)
            bn.num_batches_tracked += 1
        return ms, count

    def _postnet_fwd(self, pn, d_bf16, seed: int = 0):
        t = E.postnet_cache(pn)
        B, T, D = d_bf16.shape
        dev = d_bf16.device
        md = t["bout"].numel()
        mel_pre = torch.empty((B, T, md), device=dev, dtype=torch.float32)
        mel_pre_bf = torch.empty((B, T, md), device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(d_bf16, t["wout"], t["bout"], block_n=md, out_f32=mel_pre, out_bf16=mel_pre_bf)
        convs = [pn.conv1] + list(pn.conv_list)
        bns = [pn.pre_batchnorm] + list(pn.batch_norm_list)
        h = mel_pre_bf
        layers = []
        pp = self.drop.p_post
        for li, ((w, _, _), cv, bn) in enumerate(zip(t["convs"], convs, bns)):
            dr = (pp, seed, self.drop.site(f"postnet.{li}") if pp > 0 else -1)
            c = torch.empty((B, T, D), device=dev, dtype=torch.float32)
            lib.conv_gemm(h, w, cv.bias, taps=5, pad_left=4, block_n=192, out_f32=c)
            stats = torch.zeros((2, D), device=dev, dtype=torch.float32)
            lib.bn_stats(c, stats)
            ms, count = self._bn_finish_stats(stats, B * T, bn)
            o = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
            lib.bn_tanh_fwd(c, ms, bn.weight, bn.bias, o, drop=dr)  # postnets.py:71-73 incl. the dropout after tanh
            layers.append(dict(x=h, c=c, ms=ms, h=o, count=count, w=w, drop=dr))
            h = o
        mel_post = torch.empty((B, T, md), device=dev, dtype=torch.float32)
        lib.conv_gemm(h, t["w2"], t["b2"], taps=5, pad_left=4, block_n=md, res_f32=mel_pre, res_bstride=T * md,
                      res_tstride=md, out_f32=mel_post)
        return mel_pre, mel_post, dict(d=d_bf16, layers=layers, t=t, h_last=h)

    def _postnet_bwd(self, pn, tape, g_pre: torch.Tensor, g_post: torch.Tensor, g_post_bf: torch.Tensor):
        """fp32 gradient w.r.t. the decoder output from dL/dmel_pre and dL/dmel_post."""
        G = self.G
        t = tape["t"]
        B, T, D = tape["d"].shape
        md = g_pre.shape[-1]
        dev = g_pre.device
        convs = [pn.conv1] + list(pn.conv_list)
        bns = [pn.pre_batchnorm] + list(pn.batch_norm_list)
        # mel_post = mel_pre + conv2(h4)
        lib.colsum(g_post, G(pn.conv2.bias))
        lib.conv_wgrad(g_post_bf, tape["h_last"], G(pn.conv2.weight), taps=5, pad_left=4, conv_layout=True)
        dh = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
        lib.conv_gemm(g_post_bf, self._wt(t, "w2"), self.zeros(D), taps=5, pad_left=0, block_n=192, out_bf16=dh)
        g_mel = g_pre + g_post  # both paths reach mel_pre (postnets.py:75)
        dmel = torch.empty((B, T, md), device=dev, dtype=torch.float32)
        dmel_bf = torch.empty((B, T, md), device=dev, dtype=torch.bfloat16)
        for i in range(len(convs) - 1, -1, -1):
            s, cv, bn = tape["layers"][i], convs[i], bns[i]
            red = torch.zeros((2, D), device=dev, dtype=torch.float32)
            lib.bn_tanh_bwd(dh, s["h"], s["c"], s["ms"], bn.weight, red, None, 1.0 / s["count"], apply=False, drop=s["drop"])
            G(bn.bias).add_(red[0])
            G(bn.weight).add_(red[1])
            if self.sync_bn:  # SyncBatchNorm backward: the two reductions are summed over all ranks
                dist.all_reduce(red, group=self.pg)
            dc = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
            lib.bn_tanh_bwd(dh, s["h"], s["c"], s["ms"], bn.weight, red, dc, 1.0 / s["count"], apply=True, drop=s["drop"])
            lib.conv_wgrad(dc, s["x"], G(cv.weight), taps=5, pad_left=4, conv_layout=True, dbias=G(cv.bias))
            if i > 0:
                dh = torch.empty((B, T, D), device=dev, dtype=torch.bfloat16)
                lib.conv_gemm(dc, self._wt(t, ("convs", i, 0)), self.zeros(D), taps=5, pad_left=0, block_n=192, out_bf16=dh)
            else:
                lib.conv_gemm(dc, self._wt(t, ("convs", i, 0)), self.zeros(md), taps=5, pad_left=0, block_n=md, res_f32=g_mel,
                              res_bstride=T * md, res_tstride=md, out_f32=dmel, out_bf16=dmel_bf)
        # mel_pre = out(d)
        lib.colsum(dmel, G(pn.out.bias))
        lib.conv_wgrad(dmel_bf, tape["d"], G(pn.out.weight))
        dd = torch.empty((B, T, D), device=dev, dtype=torch.float32)
        lib.conv_gemm(dmel_bf, self._wt(t, "wout"), self.zeros(D), block_n=192, out_f32=dd)
        return dd

    # ------------------------------------------------------------------------------------------
    # whole model (Models/fastspeech2.py:118-241, teacher-forced as train_fastspeech2.py:167 calls it)
    # ------------------------------------------------------------------------------------------
    def forward_train(self, text, src_mask, mel_mask, alignment, f0, energy):
        """Teacher-forced forward in train() semantics.  Returns (outputs dict, tape for backward_train)."""
        model = self.model
        enc, va, dec, pn = model.encoder, model.variance_adaptor, model.decoder, model.postnet
        B, L = text.shape
        dev = text.device
        sm = E.mask_to_u8(src_mask, B, L, dev)
        kv_src = lib.mask_lengths(sm)
        seed = self.drop.next_seed() if self.drop.active else 0   # a new set of dropout masks per training forward
        self.last_dropout_seed = seed
        e_bf, tape_enc = self._stack_fwd(enc, text, sm, kv_src, "enc", seed)
        mm = E.mask_to_u8(mel_mask, B, mel_mask.shape[-1], dev)
        Tm = mm.shape[1]
        # ---- variance adaptor (varianceadaptor.py:66-129, teacher-forced) ----
        log_d, tape_d = self._pred_fwd(va.duration_predictor, e_bf, sm, "duration", seed)
        dur = torch.clamp(alignment, min=0).to(torch.int32).contiguous()
        cum, mel_len = lib.lr_scan(dur)
        xe, _, _, _ = lib.lr_expand(e_bf, cum, mel_len, Tm, want_aux=False)
        p, tape_p = self._pred_fwd(va.pitch_predictor, xe, mm, "pitch", seed)
        e, tape_e = self._pred_fwd(va.energy_predictor, xe, mm, "energy", seed)
        pb, eb = va._bins(dev)
        tabs = va._tables()
        f0 = f0.to(torch.float32).contiguous()
        energy = energy.to(torch.float32).contiguous()
        va_x, p_idx, e_idx = lib.bucket_embed_add(xe, f0, energy, pb, eb, tabs["ptab"], tabs["etab"], want_idx=True)
        kv_mel = lib.mask_lengths(mm)
        d_bf, tape_dec = self._stack_fwd(dec, va_x, mm, kv_mel, "dec", seed)
        mel_pre, mel_post, tape_pn = self._postnet_fwd(pn, d_bf, seed)
        tape = dict(enc=tape_enc, dec=tape_dec, pn=tape_pn, d=tape_d, p=tape_p, e=tape_e, cum=cum, p_idx=p_idx,
                    e_idx=e_idx, Tm=Tm, seed=seed, shapes=[mel_pre.shape, mel_post.shape, log_d.shape, p.shape, e.shape])
        return dict(mel_pre=mel_pre, mel_post=mel_post, log_d=log_d, p=p, e=e, va_x=va_x, text_dur=xe,
                    f0=f0, energy=energy), tape

    def backward_train(self, tape, g_pre, g_post, g_d, g_p, g_e, g_post_bf=None) -> None:
        """Gradients of every parameter from the gradients of the five outputs (accumulated into the flat buffer)."""
        self._prepare_wt()
        model = self.model
        enc, va, dec, pn = model.encoder, model.variance_adaptor, model.decoder, model.postnet
        G = self.G
        Tm = tape["Tm"]
        if g_post_bf is None:
            g_post_bf = lib.cast_bf16(g_post)
        dd = self._postnet_bwd(pn, tape["pn"], g_pre, g_post, g_post_bf)
        dva = self._stack_bwd(dec, tape["dec"], dd)                    # fp32 (B,Tm,D) w.r.t. va_x
        if self._overlap and self.world > 1 and self.flatten and self.tail_start is not None:
            # decoder + postnet gradients are final: reduce them while the adaptor and the encoder run backward
            self._pending = allreduce_tail_begin(self.flat_g, self.tail_start, self.pg)
        lib.embed_bwd(dva, Tm, idx=tape["p_idx"], table=G(va.pitch_embedding.weight))
        lib.embed_bwd(dva, Tm, idx=tape["e_idx"], table=G(va.energy_embedding.weight))
        dxe = self._pred_bwd(va.pitch_predictor, tape["p"], g_p, dva)
        dxe = self._pred_bwd(va.energy_predictor, tape["e"], g_e, dxe)
        de = lib.lr_bwd(dxe, tape["cum"])                              # fp32 (B,L,D)
        de = self._pred_bwd(va.duration_predictor, tape["d"], g_d, de)
        self._stack_bwd(enc, tape["enc"], de)

    def forward_backward(self, text, src_mask, mel_mask, alignment, f0, energy, mel) -> Dict[str, torch.Tensor]:
        """One forward + losses + backward.  Gradients ACCUMULATE into the flat buffer (zero_grad() first).
        Returns the five loss terms and their sum as device scalars plus the model outputs."""
        o, tape = self.forward_train(text, src_mask, mel_mask, alignment, f0, energy)
        mel_pre, mel_post, log_d, p, e, f0, energy = (o[k] for k in ("mel_pre", "mel_post", "log_d", "p", "e", "f0",
                                                                     "energy"))
        dev = text.device

        # ---- losses (train_fastspeech2.py:212-217,245-259): unmasked L1 means, summed ----
        mel = mel.to(torch.float32).contiguous()
        lb = self.loss_buf
        lb.zero_()
        g_pre = torch.empty_like(mel_pre)
        g_post = torch.empty_like(mel_post)
        g_post_bf = torch.empty(mel_post.shape, device=dev, dtype=torch.bfloat16)
        g_d, g_p, g_e = torch.empty_like(log_d), torch.empty_like(p), torch.empty_like(e)
        lib.l1_loss(mel_pre, mel, lb[0:1], grad=g_pre)
        lib.l1_loss(mel_post, mel, lb[1:2], grad=g_post, grad_bf16=g_post_bf)
        lib.l1_loss(log_d, alignment.to(torch.float32).contiguous(), lb[2:3], grad=g_d, log_target=True)
        lib.l1_loss(p, f0, lb[3:4], grad=g_p)
        lib.l1_loss(e, energy, lb[4:5], grad=g_e)

        self.backward_train(tape, g_pre, g_post, g_d, g_p, g_e, g_post_bf)
        total = lb[0:5].sum()
        return dict(loss=total, mel_loss=lb[0], post_mel_loss=lb[1], duration_loss=lb[2], f0_loss=lb[3],
                    energy_loss=lb[4], outputs_prenet=mel_pre, outputs_postnet=mel_post, log_d_prediction=log_d,
                    p_prediction=p, e_prediction=e)

    def zero_grad(self) -> None:
        self.flat_g.zero_()

    def optimizer_step(self) -> float:
        """Gradient all-reduce (mean over ranks) + clip_grad_norm_ + Adam with the Noam learning rate
        (train_fastspeech2.py:117-120,304-315).  Returns the learning rate used."""
        if not self.flatten:
            raise RuntimeError("optimizer_step() needs flatten=True; in autograd mode the caller's optimizer steps")
        self.step_count += 1
        if self._pending is not None:  # the tail went out during the backward: reduce the head, join both
            gscale = allreduce_tail_finish(self.flat_g, self.tail_start, self._pending, self.world, self.pg)
            self._pending = None
        else:
            gscale = allreduce_mean_scale(self.flat_g, self.world, self.pg)  # NCCL over NVLink
        self.sumsq.zero_()
        lib.sumsq(self.flat_g, self.sumsq)
        lr = self.lr()
        step_fn = lib.radam_step if self.optimizer == "radam" else lib.adam_step
        step_fn(self.flat_p, self.flat_g, self.flat_m, self.flat_v, self.sumsq, gscale=gscale,
                clip=float(self.clip or 0.0), lr=lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                step=self.step_count)
        self.invalidate_caches()
        return lr

    def grad_norm(self) -> torch.Tensor:
        return torch.sqrt(self.sumsq) / self.world

    # ------------------------------------------------------------------------------------------
    # CUDA-graphed forward + backward (see __init__)
    # ------------------------------------------------------------------------------------------
    def _graph_ok(self) -> bool:
        return self.cuda_graph and self.flatten and self.world == 1 and not self.drop.active

    def _graphed_forward_backward(self, args: tuple) -> Dict[str, torch.Tensor]:
        """zero_grad() + forward_backward(*args) through a captured graph.  The returned tensors are the graph's own
        buffers: the next replay overwrites them."""
        key = tuple((tuple(a.shape), a.dtype) if torch.is_tensor(a) else a for a in args)
        ent = self._graphs.get(key)
        if ent is None:
            static = [a.clone() if torch.is_tensor(a) else a for a in args]
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):          # one eager pass first: lazy initialisation, allocator warm-up
                saved = [b.clone() for b in self.model.buffers()]     # (its BatchNorm running-stat updates are undone)
                self.invalidate_caches()
                self.zero_grad()
                self.forward_backward(*static)
                with torch.no_grad():
                    for b, sv in zip(self.model.buffers(), saved):
                        b.copy_(sv)
            torch.cuda.current_stream().wait_stream(side)
            self.invalidate_caches()               # so that the capture contains the refresh of the weight operands
            graph = torch.cuda.CUDAGraph()
            l0 = lib.LAUNCH_COUNT
            try:
                with torch.cuda.graph(graph):
                    self.zero_grad()
                    out = self.forward_backward(*static)
            except Exception as ex:  # noqa: BLE001 — e.g. a collective that cannot be captured on this stack
                import warnings

                warnings.warn(f"Trainer: CUDA graph capture failed ({str(ex)[:200]}); continuing with eager steps")
                self.cuda_graph = False
                torch.cuda.synchronize()
                self.invalidate_caches()
                self.zero_grad()
                return self.forward_backward(*args)
            ent = self._graphs[key] = (graph, static, out, lib.LAUNCH_COUNT - l0)
        else:
            for st, a in zip(ent[1], args):
                if torch.is_tensor(a):
                    st.copy_(a)
        ent[0].replay()
        self.graph_launches = ent[3]
        return dict(ent[2])

    def train_step(self, text, src_mask, mel_mask, alignment, f0, energy, mel) -> Dict[str, torch.Tensor]:
        if self._graph_ok():
            out = self._graphed_forward_backward((text, src_mask, mel_mask, alignment, f0, energy, mel))
            out["lr"] = self.optimizer_step()
            return out
        self.zero_grad()
        self._overlap = self.overlap_allreduce
        try:
            out = self.forward_backward(text, src_mask, mel_mask, alignment, f0, energy, mel)
        finally:
            self._overlap = False
        out["lr"] = self.optimizer_step()
        return out

    # ------------------------------------------------------------------------------------------
    # checkpointing: the reference saves optimizer.state_dict() and resumes `step` from it
    # (train_fastspeech2.py:323,444).  Same layout as torch.optim.Adam / radam.RAdam: per-parameter
    # {'step', 'exp_avg', 'exp_avg_sq'} in model.parameters() order, views of the flat buffers cloned out.
    # ------------------------------------------------------------------------------------------
    def state_dict(self) -> dict:
        if not self.flatten:
            raise RuntimeError("state_dict() needs flatten=True; in autograd mode the caller's optimizer owns the state")
        state = {}
        for i, (p, (o, k)) in enumerate(zip(self.params, self.offsets)):
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.flat_m[o:o + k].view(p.shape).clone(),
                        "exp_avg_sq": self.flat_v[o:o + k].view(p.shape).clone()}
        group = {"lr": self.lr() if self.step_count else (self.fixed_lr or self.lr_max), "betas": tuple(self.betas),
                 "eps": self.eps, "weight_decay": 0, "amsgrad": False, "params": list(range(len(self.params)))}
        return {"state": state if self.step_count else {}, "param_groups": [group],
                "ttsb200": {"optimizer": self.optimizer, "step_count": self.step_count}}

    def load_state_dict(self, sd: dict) -> None:
        """Accepts what `state_dict()` wrote, or a torch.optim.Adam / radam.RAdam state_dict of the same model."""
        if not self.flatten:
            raise RuntimeError("load_state_dict() needs flatten=True")
        state = sd.get("state", {})
        if state and len(state) != len(self.params):
            raise ValueError(f"optimizer state holds {len(state)} tensors, the model has {len(self.params)} parameters")
        steps = set()
        self.flat_m.zero_()
        self.flat_v.zero_()
        for i, (p, (o, k)) in enumerate(zip(self.params, self.offsets)):
            st = state.get(i, state.get(str(i)))
            if st is None:
                continue
            for name, flat in (("exp_avg", self.flat_m), ("exp_avg_sq", self.flat_v)):
                t = st[name]
                if tuple(t.shape) != tuple(p.shape):
                    raise ValueError(f"optimizer state {i}.{name} has shape {tuple(t.shape)}, parameter {tuple(p.shape)}")
                flat[o:o + k].copy_(t.reshape(-1).to(device=self.dev, dtype=torch.float32))
            steps.add(int(float(st["step"])))
        extra = sd.get("ttsb200", {})
        if len(steps) > 1:
            raise ValueError(f"per-parameter step counts differ ({sorted(steps)}): not a state of this training loop")
        self.step_count = steps.pop() if steps else int(extra.get("step_count", 0))
        g = sd.get("param_groups", [{}])[0]
        if "betas" in g:
            self.betas = tuple(g["betas"])
        if "eps" in g:
            self.eps = g["eps"]

    def broadcast_parameters(self, src: int = 0) -> None:
        """DDP's initial parameter broadcast, one collective over the flat buffer."""
        if self.world > 1:
            dist.broadcast(self.flat_p, src=src, group=self.pg)
            for bn in [self.model.postnet.pre_batchnorm] + list(self.model.postnet.batch_norm_list):
                dist.broadcast(bn.running_mean, src=src, group=self.pg)
                dist.broadcast(bn.running_var, src=src, group=self.pg)
            self.invalidate_caches()


# --------------------------------------------------------------------------------------------------
# Drop-in autograd mode: model.train(); out = model(text, src_mask, mel_mask, alignment, f0, energy);
# loss.backward(); torch optimizer.step()  — the reference's own loop (train_fastspeech2.py:167-315).
# --------------------------------------------------------------------------------------------------
class _FastSpeech2TrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, text, src_mask, mel_mask, alignment, f0, energy, *params):
        o, tape = engine.forward_train(text, src_mask, mel_mask, alignment, f0, energy)
        ctx.engine, ctx.tape = engine, tape
        ctx.mark_non_differentiable(o["va_x"], o["text_dur"])
        return o["mel_pre"], o["mel_post"], o["log_d"], o["p"], o["e"], o["va_x"], o["text_dur"]

    @staticmethod
    def backward(ctx, g_pre, g_post, g_d, g_p, g_e, _g_va, _g_td):
        engine, tape = ctx.engine, ctx.tape
        o_shapes = tape["shapes"]

        def z(g, shape):
            return torch.zeros(shape, device=engine.dev, dtype=torch.float32) if g is None else g.float().contiguous()

        engine.fresh_grads()
        engine.backward_train(tape, z(g_pre, o_shapes[0]), z(g_post, o_shapes[1]), z(g_d, o_shapes[2]),
                              z(g_p, o_shapes[3]), z(g_e, o_shapes[4]))
        return (None,) * 7 + tuple(engine.G(p) for p in engine.params)


def autograd_forward(model, text, src_mask, mel_mask, d_target, p_target, e_target):
    """FastSpeech2.forward in train() mode: the 14-tuple of Models/fastspeech2.py:241 with autograd connected
    to every parameter through one Function whose backward is `Trainer.backward_train`."""
    if mel_mask is None or d_target is None or p_target is None or e_target is None:
        raise NotImplementedError("training forward is teacher-forced: mel_mask, d_target, p_target, e_target are "
                                  "required (train_fastspeech2.py:167)")
    eng = getattr(model, "_ttsb_train_engine", None)
    if eng is None:
        eng = Trainer(model, flatten=False)
        object.__setattr__(model, "_ttsb_train_engine", eng)
    fn_out = _FastSpeech2TrainFn.apply(eng, text, src_mask, mel_mask, d_target, p_target, e_target, *eng.params)
    return fn_out
