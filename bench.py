#!/usr/bin/env python
"""bench.py — FastSpeech2 mel frames/sec on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (config.workload): BASELINE.json configs[1] — batched free-running inference, bf16 tensor-core
compute, batch 64 per GPU, 128 synthetic phonemes per utterance expanded to ~1000 mel frames, random-init
weights of the canonical architecture (config/hparams_template.py), synthetic data.
A "step" is one full forward (text ids -> mel_post) over one batch.  `value` = valid mel frames
(sum of mel_len over the batch and all ranks) per second with inputs resident in HBM, timed with
CUDA events; `e2e` = the same through the public module API with pinned HOST inputs copied H2D and
mel_post copied D2H inside the timed region.  Multi-GPU: utterances are sharded across ranks with no
data-path collective (weak scaling); the only collective is the MAX over ranks of the elapsed time.

The timed mode is `config.duration_precision` = "fp32": the encoder -> duration chain runs to fp32 accuracy so the
durations, frame indices and mel lengths equal the fp32 reference's (tests/test_model_gpu.py::
test_cfg2_full_size_vs_oracle checks exactly this batch).  `bf16_duration_chain` reports the faster mode whose
durations may flip.  Sub-records: `torch_eager_b200` (the oracle port of the reference under stock
PyTorch eager on the same GPU, bf16 autocast, decisions teacher-forced — BASELINE.md §3), `train` (the data-parallel
training step of configs[2] with the NCCL all-reduce, at every N), `latency_b1` (the reference's own batch-1 call).

`--impl reference` times the reference algorithm's CPU implementation (the oracle port, fp32, all host
threads) on the same 64-utterance batch — the unmodified reference is Python that cannot travel to the GPU
box (it lives outside the repo), the oracle is its line-by-line restatement pinned by tests/golden.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "fastspeech2_mel_frames_per_sec"
UNIT = "frames/s"
B_PER_GPU = 64
L_PHON = 128
WORKLOAD = "fastspeech2_inference_bf16_B64_L128_T~1000_free_running"


def read_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return dict(hbm=d.get("hbm_gbs", 6650.0), tc_burst=d.get("bf16_tflops", 1590.0),
                    tc_sust=d.get("bf16_tflops_sustained", 1400.0), src="measured")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sust=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """Samples SM clocks / throttle reasons of this rank's GPU while the timed regions run.  NVML is queried in-process
    (pynvml: microseconds per sample, no fork); `nvidia-smi` subprocesses are the fallback.  Round 1 forked a
    10 Hz nvidia-smi on every rank inside the timed region, which the 8-GPU run may have paid for."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, period_s: float = 0.05):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.period = period_s
        self.stop_flag = threading.Event()
        self.samples = []   # (sm_mhz, max_mhz, set of reason names)
        self.nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[gpu_index]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else gpu_index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
        get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        r = get(self.handle)
        names = set()
        for nm, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                        ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                names.add(nm)
        self.samples.append((float(sm), float(mx), names))

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                              str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
            f = [t.strip() for t in out.split(",")]
            names = {n for n, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[2:6])
                     if v.lower().startswith("active")}
            self.samples.append((float(f[0]), float(f[1]), names))

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self.stop_flag.wait(self.period if self.nvml is not None else 0.5)

    def summary(self):
        sm = [x[0] for x in self.samples]
        reasons = set()
        for x in self.samples:
            reasons |= x[2]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_min_mhz": min(sm) if sm else None,
                "sm_max_mhz": max((x[1] for x in self.samples), default=None), "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def make_batch(rank: int, world: int = 1):
    """One pool of world x 64 utterances, dealt to ranks by transformer_tts_b200.sharding (longest first,
    snake order); rank 0 of a 1-GPU run gets exactly the batch of tests/test_model_gpu.py::test_cfg2_full_size_vs_oracle."""
    import torch

    from transformer_tts_b200 import sharding, synth

    text, pos, lens = synth.synth_text_batch(B_PER_GPU * world, L_PHON, seed=1234)
    mine = sorted(sharding.shard_by_length([int(n) for n in lens], world)[rank])
    idx = torch.tensor(mine)
    return text[idx], pos[idx], lens[idx]


def cpu_oracle_run(steps: int, warmup: int, sample_b: int):
    """Times the oracle (reference algorithm, fp32, torch CPU) on `sample_b` utterances of the workload."""
    import torch

    from oracle import fastspeech2_oracle as O
    from transformer_tts_b200 import synth

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    text, pos, _ = make_batch(0, 1)
    text, pos = text[:sample_b], pos[:sample_b]
    src_mask = (pos != 0).unsqueeze(-2)
    times, frames = [], 0
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            o = O.fastspeech2_forward(sd, vars(hp), text, src_mask)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
                frames = int((o["mel_pos"] != 0).sum())
    tot = sum(times)
    return dict(value=frames * len(times) / tot, ms_per_step=1e3 * tot / len(times), cores=cores,
                sample=f"{sample_b} utterances x {L_PHON} phonemes ({frames} frames) per step, {len(times)} steps, "
                       f"fp32 torch CPU oracle port of Models/fastspeech2.py, free-running")


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_oracle_run(max(1, min(args.steps, 10)), 1, sample_b=B_PER_GPU)  # the whole batch of the GPU arm
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": B_PER_GPU, "phonemes": L_PHON,
                       "note": "CPU fp32 oracle port, the same 64-utterance batch as the GPU arm's rank 0"},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def tmap_stats():
    import ctypes

    from transformer_tts_b200 import lib

    h, m = ctypes.c_ulonglong(0), ctypes.c_ulonglong(0)
    lib.load().ttsb200_debug_tmap_cache(ctypes.byref(h), ctypes.byref(m))
    return {"hits": h.value, "misses": m.value}


def load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the hot kernels, from ONE `ncu --set full` capture of
    tools/run_hot_kernels.py on this tree (profiles/r02_hot_kernels_traffic.json, written by tools/ncu_traffic.py with
    the tree it was captured on).  Absent file -> traffic is reported as null, never as a stale literal."""
    p = ROOT / "profiles" / "r02_hot_kernels_traffic.json"
    if not p.exists():
        return {}, None
    d = json.loads(p.read_text())
    return d.get("traffic_bytes", {}), d.get("tree")


def kernel_rooflines(model, peaks, mel_len):
    """Per-kernel rooflines at the decoder shape of the workload: the REAL ragged batch of the timed step (mel_len of
    its 64 utterances, Tmax = their maximum, the same ragged tile list the step uses).  Every hot kernel is timed alone
    with CUDA events around a captured graph of 12 launches, rotating over three buffer sets so that consecutive
    launches never find their operands in the 126 MB L2.  Algorithmic work counts VALID rows only (SURVEY §8d:
    padding work does not count) and is reported against both roofs."""
    import torch

    from transformer_tts_b200 import lib
    from transformer_tts_b200.Models import _engine as E

    lens = mel_len.to("cpu", torch.int64)
    B, T, D, H = int(lens.numel()), int(lens.max()), 384, 4
    dev = "cuda"
    lc = E.layer_cache(model.decoder.layers[0])
    k = lc["k"]
    NS = 3
    mask = (torch.arange(T)[None, :] < lens[:, None]).to(torch.uint8).to(dev)
    kv = lib.mask_lengths(mask)
    tmap = lib.build_tile_map(kv, T)
    M = float(lens.sum())                      # valid rows
    rows_padded = float(B * T)
    xs = [(torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16) for _ in range(NS)]
    x32 = [torch.randn(B, T, D, device=dev) for _ in range(NS)]
    qkvs = [(torch.randn(B, T, 3 * D, device=dev) * 0.5).to(torch.bfloat16) for _ in range(NS)]
    hs = [torch.relu(torch.randn(B, T, 4 * D, device=dev)).to(torch.bfloat16) for _ in range(NS)]
    ob = [torch.empty(B, T, D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
    o32 = [torch.empty(B, T, D, device=dev) for _ in range(NS)]
    L = L_PHON
    xe = [(torch.randn(B, L, D, device=dev)).to(torch.bfloat16) for _ in range(NS)]
    # durations that expand L phonemes to exactly this batch's frame counts
    dur = (lens[:, None] // L).expand(B, L).clone()
    dur[:, 0] += lens - dur.sum(1)
    dur = dur.to(torch.int32).to(dev)
    cum, ml = lib.lr_scan(dur)
    Tlr = T
    va = model.variance_adaptor
    pb, eb = va._bins(torch.device(dev))
    tabs = va._tables()
    pv = [torch.rand(B, T, device=dev) * 700 + 80 for _ in range(NS)]
    ev_ = [torch.rand(B, T, device=dev) * 400 for _ in range(NS)]
    ln = lc["ln1"]

    def t_qkv(i):
        lib.conv_gemm(xs[i], lc["wqkv"], lc["bqkv"], block_n=192, out_bf16=qkvs[i], tile_map=tmap)

    def t_att(i):
        lib.attention_fwd(qkvs[i], mask, kv, ob[i], H, skip_padded_queries=True)

    def t_out(i):
        lib.conv_gemm(xs[i], lc["wo"], lc["bo"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x32[i], res_bstride=T * D,
                      res_tstride=D, ln1=lc["ln2"], out_f32=o32[i], out_bf16=ob[i], tile_map=tmap)

    def t_ffn1(i):
        lib.conv_gemm(xs[i], lc["w1"], lc["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU, out_bf16=hs[i],
                      tile_map=tmap)

    def t_ffn2(i):
        lib.conv_gemm(hs[i], lc["w2"], lc["b2"], taps=k, pad_left=k // 2, block_n=D, epi=lib.EPI_FFN2, res_bf16=xs[i],
                      res_f32=x32[i], res_bstride=T * D, res_tstride=D, ln1=lc["ln_ff"], ln2=ln, out_f32=o32[i],
                      out_bf16=ob[i], tile_map=tmap)

    z32 = [torch.empty(B, T, D, device=dev) for _ in range(NS)]

    def t_ffn(i):  # the kernel the inference step runs for k=1 blocks (csrc/ffn_fused.cuh)
        lib.ffn_fused(xs[i], lc["w1"], lc["b1"], lc["w2"], lc["b2"], x32[i], lc["ln_ff"], ln, z32[i], ob[i], tile_map=tmap)

    def t_lr(i):
        lib.lr_expand(xe[i], cum, ml, Tlr, want_aux=True)

    def t_bucket(i):
        lib.bucket_embed_add(xs[i], pv[i], ev_[i], pb, eb, tabs["ptab"], tabs["etab"])

    def t_ln(i):
        lib.layernorm(x32[i], ln[0], ln[1], E.LN_EPS, y_bf16=ob[i])

    att_flops = float((4.0 * lens.double() ** 2 * D).sum())
    s2 = 2  # bf16 bytes
    # (key, name, launcher, algorithmic FLOPs, algorithmic HBM bytes).  Bytes are what the op must move when NOTHING is
    # re-read: operands once in, results once out.  The FFN pair's bytes are those of the fused op the reference writes
    # as one function (Models/modules.py:81-88): x in (bf16), residual in + out (fp32), next-LN out (bf16) — the
    # 1536-wide hidden activation is not algorithmic traffic.
    specs0 = [
        ("qkv", "conv_gemm<192,AFFINE,A-stationary> decoder q|k|v linears", t_qkv, 2 * M * D * 3 * D, M * D * s2 * 4),
        ("attention", "attention_fwd decoder masked self-attention", t_att, att_flops, M * D * s2 * 4),
        ("out_proj", "conv_gemm<384,RES_LN> decoder out-proj + residual + LayerNorm", t_out, 2 * M * D * D,
         M * D * (2 * s2 + 8)),
        ("ffn_fused", "ffn_fused decoder FeedForward, the reference's ONE op (conv1 + ReLU + conv2 + res + LN + res + LN), "
                      "hidden activation kept in TMEM", t_ffn, 4 * M * D * 4 * D * k, M * D * (s2 + 4 + 4 + s2)),
        ("ffn1", "conv_gemm<256,AFFINE,A-stationary> FFN conv1 + ReLU (two-kernel path: training, short batches)", t_ffn1,
         2 * M * D * 4 * D * k, M * D * s2 * 5),
        ("ffn2", "conv_gemm<384,FFN2> FFN conv2 + LN + residual + LN (two-kernel path)", t_ffn2, 2 * M * D * 4 * D * k,
         M * D * (s2 * 6 + 8)),
        ("lr_expand", "lr_expand LengthRegulator frame gather", t_lr, 0,
         B * L * D * s2 + B * L * 4 + M * D * s2 + rows_padded * (4 + 8 + 1) + B * 4),
        ("bucket_embed_add", "bucket_embed_add pitch/energy bucketize + embedding add", t_bucket, 0,
         rows_padded * D * s2 * 2 + rows_padded * 8),
        ("layernorm", "layernorm standalone fp32 -> bf16", t_ln, 0, rows_padded * D * (4 + s2)),
    ]
    traffic, traffic_tree = load_traffic()
    out = []
    reps = 4
    for key, name, fn, flops, nbytes in specs0:
        for i in range(NS):
            fn(i)  # warm-up: first-call attribute setup, allocator pools
        torch.cuda.synchronize()
        # reps x NS launches captured in one CUDA graph: back-to-back on the device, no host launch gaps
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(reps):
                for i in range(NS):
                    fn(i)
        graph.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        times = []
        for _ in range(3):
            e0.record()
            graph.replay()
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1) / (reps * NS))
        ms = statistics.median(times)
        tf = flops / (ms * 1e-3) / 1e12
        gb = nbytes / (ms * 1e-3) / 1e9
        t_tc = flops / (peaks["tc_burst"] * 1e12)
        t_hbm = nbytes / (peaks["hbm"] * 1e9)
        bound = "tensor" if t_tc >= t_hbm else "hbm"
        ach, peak, unit = (tf, peaks["tc_burst"], "TFLOP/s") if bound == "tensor" else (gb, peaks["hbm"], "GB/s")
        tr = traffic.get(key)
        out.append({"kernel": name, "key": key, "bound": bound, "achieved": ach, "peak": peak, "unit": unit,
                    "frac": ach / peak, "traffic": tr, "ms_per_launch": ms,
                    "algorithmic_flops_per_launch": flops, "algorithmic_bytes_per_launch": nbytes,
                    "frac_of_tensor_roof": tf / peaks["tc_burst"], "frac_of_hbm_roof": gb / peaks["hbm"],
                    "rows_valid": M, "rows_padded": rows_padded, "traffic_tree": traffic_tree,
                    "peak_source": peaks["src"] + (" burst bf16 (kernel timed alone)" if bound == "tensor" else " copy bandwidth")})
    return out


def torch_eager_b200(hp, text_d, mask_d, aux, p_pred, e_pred, steps: int = 3):
    """BASELINE.md §3 "reference on the same box": the oracle port of the reference's modules under STOCK PyTorch eager
    on this GPU, bf16 autocast, durations / pitch / energy teacher-forced to the decisions of our own run (oracle mode
    3, so both arms produce the same frames), CUDA events.  A reported baseline; never on the product path."""
    import torch

    from oracle import fastspeech2_oracle as O
    from transformer_tts_b200 import synth

    sd = {k: v.cuda() for k, v in synth.synth_state_dict(hp).items()}
    dur = aux["dur"].long()
    mel_len = dur.sum(1)
    B, T = int(dur.shape[0]), int(mel_len.max())
    ar = torch.arange(1, T + 1, device="cuda").unsqueeze(0)
    mel_mask = (ar <= mel_len.unsqueeze(1)).unsqueeze(-2)
    frames = int(mel_len.sum())

    def step():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            return O.fastspeech2_forward(sd, vars(hp), text_d, mask_d, mel_mask=mel_mask, d_target=dur, p_target=p_pred,
                                         e_target=e_pred)

    step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"value": frames / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": steps, "frames_per_step": frames,
            "what": "oracle port of Models/fastspeech2.py, stock PyTorch eager on cuda:0, torch.autocast(bfloat16), "
                    "durations / pitch / energy teacher-forced (oracle mode 3), attention maps not kept"}


def latency_b1(model, hp, iters: int = 20):
    """BASELINE.json configs[0] shape on the GPU: the reference's own serving call (test_fastspeech2.py:165-171),
    batch 1, 64 phonemes, free-running; wall-clock per call with a synchronize after each (a serving loop's latency)."""
    import torch

    from transformer_tts_b200 import lib, synth

    text, pos, _ = synth.synth_text_batch(1, 64, seed=1234)
    text, mask = text.cuda(), (pos != 0).unsqueeze(-2).cuda()
    for _ in range(3):
        model(text, mask)
    torch.cuda.synchronize()
    ts = []
    l0 = lib.LAUNCH_COUNT
    for _ in range(iters):
        t0 = time.perf_counter()
        o = model(text, mask)
        torch.cuda.synchronize()
        ts.append(1e3 * (time.perf_counter() - t0))
    return {"ms_median": statistics.median(ts), "ms_min": min(ts), "frames": int(o[0].shape[1]),
            "launches_per_call": (lib.LAUNCH_COUNT - l0) // iters, "workload": "B=1, 64 phonemes, free-running, eager"}


def ar_decode_record(cpu: bool, positions: int = 1000, batch: int = 16, phonemes: int = 128):
    """BASELINE.json configs[4]: Transformer-TTS autoregressive inference with the incremental KV cache and the stop
    token, batch 16, 1000 decoder positions (2000 frames at reduction rate 2).  Device time of `Transformer.generate`
    (encoder + cross k|v projections + 1000 graph replays), inputs resident; `e2e` adds the H2D of the ids and the D2H of
    the mel.  cpu_reference: the oracle evaluating the loop the REFERENCE's way (test.py:128-191: the whole prefix re-run
    per step) on a bounded sample, the first `cpu_positions` positions of the same batch."""
    import torch

    from transformer_tts_b200 import build_transformer_tts, lib, synth

    hp = synth.make_tts_hp()
    sd = synth.synth_tts_state_dict(hp)
    model = build_transformer_tts(hp).eval()
    model.load_state_dict(sd)
    model = model.cuda()
    text, pos, _ = synth.synth_text_batch(batch, phonemes, ragged=True, seed=77, min_len=phonemes // 2)
    text_h, mask_h = text.pin_memory(), (pos != 0).unsqueeze(-2).pin_memory()
    text_d, mask_d = text_h.cuda(), mask_h.cuda()
    model.generate(text_d, mask_d, max_steps=32)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts, te = [], []
    for _ in range(3):
        l0 = lib.LAUNCH_COUNT
        e0.record()
        out = model.generate(text_d, mask_d, max_steps=positions)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
        launches = lib.LAUNCH_COUNT - l0
        t0 = time.perf_counter()
        o2 = model.generate(text_h.cuda(non_blocking=True), mask_h.cuda(non_blocking=True), max_steps=positions)
        mel_host = o2["mel"].cpu()
        te.append(1e3 * (time.perf_counter() - t0))
    ms, ems = min(ts), min(te)
    frames = batch * positions * hp.reduction_rate
    rec = {"metric": "transformer_tts_ar_mel_frames_per_sec", "value": frames / ms * 1e3, "unit": "frames/s",
           "ms_total": ms, "us_per_position": ms / positions * 1e3, "batch": batch, "phonemes": phonemes,
           "positions": positions, "reduction_rate": hp.reduction_rate, "frames": frames,
           "launches_per_position": model.last_state.launches_per_step, "gpu_launches": launches,
           "e2e": {"value": frames / ems * 1e3, "unit": "frames/s", "ms_total": ems,
                   "h2d_bytes": int(text_h.numel() * 8 + mask_h.numel()), "d2h_bytes": int(mel_host.numel() * 4)},
           "stop_positions_rank0": [int(v) for v in out["stop_step"].cpu()],
           "what": "Transformer.generate: KV-cache decode, one CUDA graph replay (small_linear / attn_decode kernels, programmatic "
                   "dependent launch) per position; random-init weights, so the stop token fires at random: all positions are decoded"}
    # teacher-forced prefix forward at BASELINE.json configs[3]'s shape (mel length 800 = 400 decoder positions at r = 2,
    # 128 phonemes), batch 32: the tensor-core path of Transformer.forward in eval mode (the training step at this shape is
    # the next sub-record) next to the same modules under stock PyTorch eager fp32 on this GPU (the oracle port)
    try:
        from oracle import transformer_tts_oracle as T

        Bp, Tp = 32, 400
        textp, posp, _ = synth.synth_text_batch(Bp, phonemes, ragged=True, seed=78, min_len=phonemes // 2)
        trg = torch.randn(Bp, Tp, hp.mel_dim, generator=torch.Generator().manual_seed(79))
        pos_mel = torch.arange(1, Tp + 1).unsqueeze(0).repeat(Bp, 1)
        sm, tm = T.create_masks(posp, pos_mel)
        a = (textp.cuda(), trg.cuda(), sm.cuda(), tm.cuda(), None)
        for _ in range(2):
            model(*a, training=False)
        torch.cuda.synchronize()
        tp = []
        for _ in range(5):
            e0.record()
            model(*a, training=False)
            e1.record()
            torch.cuda.synchronize()
            tp.append(e0.elapsed_time(e1))
        sd_gpu = {k: v.cuda() for k, v in sd.items()}
        with torch.no_grad():
            T.transformer_forward(sd_gpu, vars(hp), *a[:4])
            torch.cuda.synchronize()
            e0.record()
            for _ in range(3):
                T.transformer_forward(sd_gpu, vars(hp), *a[:4])
            e1.record()
            torch.cuda.synchronize()
        eager_ms = e0.elapsed_time(e1) / 3
        rec["teacher_forced_forward"] = {"ms": min(tp), "batch": Bp, "positions": Tp, "phonemes": phonemes,
                                         "frames_per_s": Bp * Tp * hp.reduction_rate / min(tp) * 1e3,
                                         "torch_eager_fp32_same_gpu_ms": eager_ms,
                                         "what": "Transformer.forward(prefix_mode='batched'): tcgen05 GEMMs, causal tcgen05 attention, "
                                                 "encoder-decoder attention on a packed sequence, fused FeedForward (eval mode)"}
    except Exception as ex:  # noqa: BLE001 — a sub-record must not take the headline line down
        rec["teacher_forced_forward"] = {"error": str(ex)[:200]}
    try:
        rec["teacher_forced_train_step"] = tts_train_record(cpu)
    except Exception as ex:  # noqa: BLE001
        rec["teacher_forced_train_step"] = {"error": str(ex)[:300]}
    if cpu:
        from oracle import transformer_tts_oracle as T

        n_cpu = 32
        torch.set_num_threads(os.cpu_count() or 1)
        with torch.no_grad():
            T.greedy_decode(sd, vars(hp), text[:, :], pos, 2)
            t0 = time.perf_counter()
            T.greedy_decode(sd, vars(hp), text, pos, n_cpu)
            dt = time.perf_counter() - t0
        rec["cpu_reference"] = {"value": batch * n_cpu * hp.reduction_rate / dt, "unit": "frames/s", "cores": os.cpu_count(),
                                "kind": "port", "seconds": dt,
                                "sample": f"the reference's loop (whole prefix re-run per step) for the FIRST {n_cpu} positions of the "
                                          f"same batch; its cost per step grows with the prefix, so 1000 positions are slower than this rate"}
    return rec


def tts_train_record(cpu: bool, batch: int = 32, phonemes: int = 128, mel_len: int = 800):
    """BASELINE.json configs[3]: Transformer-TTS teacher-forced training step (train.py:160-263: encoder, decoder with the
    encoder-decoder attention, mel / stop heads, 2 x L1 + BCE, backward, clip + Adam), mel length 800, batch 32, dropout 0
    (the scope of tts_training.TTSTrainer).  Device time per step, inputs resident; next to it the same step under stock
    PyTorch eager fp32 autograd on this GPU (the oracle port, forward + backward only) and, with `cpu`, on the host cores."""
    import torch
    import torch.nn.functional as F

    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200 import build_transformer_tts, lib, synth
    from transformer_tts_b200.tts_training import TTSTrainer

    hp = synth.make_tts_hp()
    r = hp.reduction_rate
    sd = synth.synth_tts_state_dict(hp)
    model = build_transformer_tts(hp)
    model.load_state_dict(sd)
    model = model.cuda().train()
    text, pos, _ = synth.synth_text_batch(batch, phonemes, ragged=True, seed=81, min_len=phonemes // 2)
    g = torch.Generator().manual_seed(82)
    lens = torch.randint(mel_len // 2 // r, mel_len // r + 1, (batch,), generator=g) * r
    lens[0] = mel_len
    pos_mel = torch.arange(1, mel_len + 1).unsqueeze(0).repeat(batch, 1)
    pos_mel[pos_mel > lens.unsqueeze(1)] = 0
    mel = torch.randn(batch, mel_len, hp.mel_dim, generator=g) * (pos_mel != 0).unsqueeze(-1)
    stop = (pos_mel == 0).float()
    stop[torch.arange(batch), lens - 1] = 1.0
    mel_in, pos_in = mel[:, :-r:r, :].contiguous(), pos_mel[:, :-r:r].contiguous()
    src_mask, trg_mask = T.create_masks(pos, pos_in)
    a = (text.cuda(), mel_in.cuda(), src_mask.cuda(), (pos_in != 0).to(torch.uint8).cuda(), mel[:, r:, :].contiguous().cuda(),
         stop[:, r:].contiguous().cuda())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 5

    def timed(tr):
        tr.world = 1          # rank 0 runs this record alone, also under torchrun: no gradient exchange with absent peers
        for _ in range(5):
            out = tr.train_step(*a)
        torch.cuda.synchronize()
        best = None
        for _ in range(2):    # two rounds, the faster one: the first trainer of a process still grows the allocator's pools
            l0 = lib.LAUNCH_COUNT
            e0.record()
            for _ in range(steps):
                out = tr.train_step(*a)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            best = ms if best is None else min(best, ms)
        return best, (lib.LAUNCH_COUNT - l0) // steps, out

    # the reference's training construction (train.py:117-126: dropout 0.1, prenet 0.5), counter-based masks
    from transformer_tts_b200.Models.transformer import Transformer
    md = Transformer(hp=hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder,
                     N_e=hp.n_layer_encoder, n_head_encoder=hp.n_head_encoder,
                     ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder, concat_after_encoder=hp.concat_after_encoder,
                     d_model_decoder=hp.d_model_decoder, N_d=hp.n_layer_decoder, n_head_decoder=hp.n_head_decoder,
                     ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder, concat_after_decoder=hp.concat_after_decoder,
                     reduction_rate=r, dropout=0.1, dropout_prenet=0.5, dropout_postnet=0.5)
    md.load_state_dict(sd)
    ms_drop, launches_drop, _ = timed(TTSTrainer(md.cuda().train(), positive_weight=5.0))
    del md
    ms_eager, launches, out = timed(TTSTrainer(model, positive_weight=5.0))
    model = build_transformer_tts(hp)
    model.load_state_dict(sd)
    trg = TTSTrainer(model.cuda().train(), positive_weight=5.0, cuda_graph=True)
    ms, _, out = timed(trg)
    launches = getattr(trg, "graph_launches", launches) + 3
    frames = float((pos_mel[:, r:] != 0).sum())
    rec = {"metric": "transformer_tts_train_mel_frames_per_sec", "value": frames / ms * 1e3, "unit": "frames/s", "ms_per_step": ms,
           "batch": batch, "phonemes": phonemes, "mel_len": mel_len, "decoder_positions": int(mel_in.shape[1]),
           "valid_target_frames": frames, "gpu_launches_per_step": launches, "loss": float(out["loss"]),
           "dropout": 0.0, "optimizer": "adam (Noam rate), clip 1.0", "cuda_graph": True, "ms_per_step_eager": ms_eager,
           "with_reference_dropout": {"ms_per_step": ms_drop, "frames_per_s": frames / ms_drop * 1e3, "dropout": 0.1,
                                      "dropout_prenet": 0.5, "gpu_launches_per_step": launches_drop},
           "what": "tts_training.TTSTrainer.train_step: tcgen05 GEMMs / wgrad, causal tcgen05 attention forward + backward, "
                   "encoder-decoder attention on a packed sequence, fused clip + Adam over the flat buffers"}

    def eager_step(sdx, dev):
        leaf = {k: (v.to(dev).clone().requires_grad_(True) if v.is_floating_point() and not k.startswith("postnet.") else v.to(dev))
                for k, v in sdx.items()}
        c = lambda t: t.to(dev)  # noqa: E731

        def step():
            o = T.transformer_forward(leaf, vars(hp), c(text), c(mel_in), c(src_mask), c(trg_mask))
            pre, st = o["outputs_prenet"], o["stop_token"]
            b, t, ch = pre.shape
            loss = 2 * F.l1_loss(pre.reshape(b, t * r, ch // r), c(mel[:, r:, :])) + F.binary_cross_entropy_with_logits(
                st.reshape(b, t * r), c(stop[:, r:]), reduction="mean", pos_weight=torch.tensor(5.0, device=dev))
            loss.backward()
            return loss
        return step

    try:
        step = eager_step(sd, "cuda")
        step()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(2):
            step()
        e1.record()
        torch.cuda.synchronize()
        rec["torch_eager_fp32_same_gpu_ms"] = e0.elapsed_time(e1) / 2
    except Exception as ex:  # noqa: BLE001
        rec["torch_eager_fp32_same_gpu_ms"] = None
        rec["eager_error"] = str(ex)[:200]
    if cpu:
        torch.set_num_threads(os.cpu_count() or 1)
        step = eager_step(sd, "cpu")
        step()
        n_cpu = 3   # bounded sample: 3 steps (forward + backward, no optimizer) of the same batch after one warm-up step
        t0 = time.perf_counter()
        for _ in range(n_cpu):
            step()
        dt = (time.perf_counter() - t0) / n_cpu
        rec["cpu_reference"] = {"value": frames / dt, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
                                "seconds_per_step": dt,
                                "sample": f"forward + backward (no optimizer) of the same batch, mean of {n_cpu} steps after one warm-up"}
    return rec


def train_record(args, rank: int, world: int, steps: int = 5):
    """BASELINE.json configs[2] inside the default line: the data-parallel training step (fwd + bwd + NCCL all-reduce
    of the flat gradient buffer + SyncBatchNorm + fused RAdam), 32 utterances per GPU, so the driver's 1/2/4/8 scaling
    run records the training curve next to the inference one."""
    import torch
    import torch.distributed as dist

    from transformer_tts_b200 import build_fastspeech2, synth
    from transformer_tts_b200.training import Trainer

    hp = synth.make_hp()
    model = build_fastspeech2(hp)
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda().train()
    tr = Trainer(model, optimizer="radam")
    tr.broadcast_parameters()
    host, frames = make_train_batch(rank, 32)
    dev = {k: v.cuda() for k, v in host.items()}
    for _ in range(3):
        out = tr.train_step(**dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = tr.train_step(**dev)
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    my_ms = e0.elapsed_time(e1) / steps
    ar_ms = 0.0
    if world > 1:
        e0.record()
        for _ in range(3):
            dist.all_reduce(tr.flat_g)
        e1.record()
        torch.cuda.synchronize()
        ar_ms = e0.elapsed_time(e1) / 3
    t = torch.tensor([my_ms, ar_ms, float(frames)], device="cuda", dtype=torch.float64)
    tot = t.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms, ar = float(t[0]), float(t[1])
    total_frames = float(tot[2])
    rec = {"metric": "fastspeech2_train_mel_frames_per_sec", "value": total_frames / (ms * 1e-3), "unit": UNIT,
           "ms_per_step": ms, "allreduce_ms": ar, "steps": steps, "batch_per_gpu": 32, "frames_per_step": total_frames,
           "optimizer": "radam", "loss": float(out["loss"]), "cuda_graph": False,
           "parallelism": f"dp{world}: one NCCL all-reduce of the flat fp32 gradient buffer "
                          f"({tr.flat_g.numel() * 4 / 1e6:.0f} MB) + SyncBatchNorm moments" if world > 1 else "dp1"}
    if world == 1 and not getattr(args, "no_train_graph", False):
        # single process: zero_grad + forward + losses + backward replayed from ONE captured CUDA graph, optimizer eager
        # (Trainer(cuda_graph=True)).  Not the headline of this record: data-parallel steps cannot use it (SyncBatchNorm
        # collectives inside the step), and `ms_per_step` above stays comparable across the 1/2/4/8 runs.
        try:
            del tr, model
            model = build_fastspeech2(hp)
            model.load_state_dict(synth.synth_state_dict(hp))
            tg = Trainer(model.cuda().train(), optimizer="radam", cuda_graph=True)
            for _ in range(3):
                tg.train_step(**dev)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(steps):
                tg.train_step(**dev)
            e1.record()
            torch.cuda.synchronize()
            gms = e0.elapsed_time(e1) / steps
            rec["single_process_cuda_graph"] = {"ms_per_step": gms, "value": total_frames / (gms * 1e-3), "unit": UNIT,
                                                "graph_kernel_nodes": getattr(tg, "graph_launches", None),
                                                "captured": bool(tg._graphs)}
            del tg
        except Exception as ex:  # noqa: BLE001
            rec["single_process_cuda_graph"] = {"error": str(ex)[:200]}
        try:
            # the reference's training construction (train_fastspeech2.py:381-389: dropout 0.1, PostNet 0.5, variance adaptor
            # 0.5): counter-based masks in the kernels, eager steps (a graph replay would repeat the masks)
            md = build_fastspeech2(hp, dropout=0.1, dropout_postnet=0.5, dropout_variance_adaptor=0.5)
            md.load_state_dict(synth.synth_state_dict(hp))
            td = Trainer(md.cuda().train(), optimizer="radam")
            for _ in range(5):
                td.train_step(**dev)
            torch.cuda.synchronize()
            dms = None
            for _ in range(2):
                e0.record()
                for _ in range(steps):
                    td.train_step(**dev)
                e1.record()
                torch.cuda.synchronize()
                dms = e0.elapsed_time(e1) / steps if dms is None else min(dms, e0.elapsed_time(e1) / steps)
            rec["with_reference_dropout"] = {"ms_per_step": dms, "value": total_frames / (dms * 1e-3), "unit": UNIT,
                                             "dropout": 0.1, "dropout_postnet": 0.5, "dropout_variance_adaptor": 0.5,
                                             "active": bool(td.drop.active)}
            del td, md
        except Exception as ex:  # noqa: BLE001
            rec["with_reference_dropout"] = {"error": str(ex)[:200]}
    tr = model = dev = out = None     # drop the references (a failed sub-measurement above may have deleted some already)
    torch.cuda.empty_cache()
    return rec


def ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from transformer_tts_b200 import build_fastspeech2, lib, synth

    lib.load()
    hp = synth.make_hp()
    model = build_fastspeech2(hp).eval()
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda()
    model.duration_precision = args.duration_precision
    model.variance_precision = args.variance_precision if args.duration_precision == "fp32" else "bf16"
    text_h, pos_h, _ = make_batch(rank, world)
    text_h, mask_h = text_h.pin_memory(), (pos_h != 0).unsqueeze(-2).pin_memory()
    text_d, mask_d = text_h.cuda(), mask_h.cuda()
    peaks = read_peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x: float, op) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=op)
        return float(t.item())

    def gather(x: float):
        if world == 1:
            return [round(x, 4)]
        g = [torch.zeros(1, device="cuda", dtype=torch.float64) for _ in range(world)]
        dist.all_gather(g, torch.tensor([x], device="cuda", dtype=torch.float64))
        return [round(float(t.item()), 4) for t in g]

    MAX, SUM = (dist.ReduceOp.MAX, dist.ReduceOp.SUM) if world > 1 else (None, None)

    def timed_steps(n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            model(text_d, mask_d)
        e1.record()
        barrier()
        return e0.elapsed_time(e1)

    with torch.no_grad():
        for _ in range(max(3, args.warmup)):
            out = model(text_d, mask_d)
        aux = dict(model.last_aux)
        p_pred, e_pred = out[3].clone(), out[4].clone()
        frames = int(aux["mel_len"].sum().item())
        tmax = int(aux["mel_len"].max().item())
        # ---- device-resident throughput (the parity mode) ----
        sampler = ClockSampler(local_rank)
        sampler.start()
        # the first collective of a process group sets the NCCL communicator up (seconds of host work, some of it after the
        # call returns): do it, and two more untimed steps, before the timed region starts
        barrier()
        for _ in range(2):
            model(text_d, mask_d)
        barrier()
        l0 = lib.LAUNCH_COUNT
        my_ms = timed_steps(args.steps)
        launches = lib.LAUNCH_COUNT - l0
        dev_ms = reduce(my_ms, MAX)
        rank_ms = gather(my_ms / args.steps)
        # ---- end to end: pinned host inputs -> H2D -> forward -> mel_post D2H ----
        # Every step copies its inputs H2D and its mel D2H inside the timed region.  The D2H of step i runs on a
        # copy stream (two pinned buffers) while step i+1 computes, as a serving loop would; the timed region ends
        # only after the last copy has landed.
        out_h = [torch.empty((B_PER_GPU, tmax, hp.mel_dim), dtype=torch.float32).pin_memory() for _ in range(2)]
        main_s, copy_s = torch.cuda.current_stream(), torch.cuda.Stream()
        done = [torch.cuda.Event(), torch.cuda.Event()]

        def e2e_step(i: int) -> None:
            o = model(text_h.cuda(non_blocking=True), mask_h.cuda(non_blocking=True))
            copy_s.wait_stream(main_s)
            with torch.cuda.stream(copy_s):
                done[i & 1].synchronize()  # the host buffer's previous copy (two steps ago) has landed
                out_h[i & 1].copy_(o[1], non_blocking=True)
                done[i & 1].record(copy_s)
            o[1].record_stream(copy_s)

        for i in range(2):
            e2e_step(i)
        copy_s.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            e2e_step(i)
        main_s.wait_stream(copy_s)
        e1.record()
        barrier()
        e2e_ms = reduce(e0.elapsed_time(e1), MAX)
        sampler.stop_flag.set()
        sampler.join(timeout=2)
        clocks = sampler.summary()
        clocks_by_rank = gather(float(clocks["sm_mhz"] or 0.0))
        # ---- the other precision modes, for comparison ----
        timed_mode = (model.duration_precision, model.variance_precision)
        modes = {}
        for dp, vp in (("fp32", "fp32"), ("fp32", "bf16"), ("bf16", "bf16")):
            if (dp, vp) == timed_mode or args.step_only:
                continue
            model.duration_precision, model.variance_precision = dp, vp
            for _ in range(3):
                model(text_d, mask_d)
            frames_o = reduce(float(model.last_aux["mel_len"].sum().item()), SUM)
            n_o = max(5, min(args.steps, 20))
            other_ms = reduce(timed_steps(n_o), MAX)
            modes[f"durations_{dp}_pitch_energy_{vp}"] = {"value": frames_o * n_o / (other_ms * 1e-3), "unit": UNIT,
                                                          "ms_per_step": other_ms / n_o, "steps": n_o}
        model.duration_precision, model.variance_precision = timed_mode
        lat = latency_b1(model, hp) if rank == 0 and not args.step_only else None
        roofs = kernel_rooflines(model, peaks, aux["mel_len"]) if rank == 0 and not args.step_only else None
        eager = torch_eager_b200(hp, text_d, mask_d, aux, p_pred, e_pred) \
            if rank == 0 and not (args.no_eager_baseline or args.step_only) else None
    total_frames = reduce(float(frames), SUM)
    value = total_frames * args.steps / (dev_ms * 1e-3)
    e2e_value = total_frames * args.steps / (e2e_ms * 1e-3)
    del out_h
    torch.cuda.empty_cache()
    trn = None if (args.no_train_record or args.step_only) else train_record(args, rank, world)
    ar = None
    if rank == 0 and not (args.step_only or args.no_ar_record):   # configs[4]; one GPU's batch (utterances are independent)
        ar = ar_decode_record(cpu=world == 1 and not args.no_cpu_baseline)
    cpu = None
    if rank == 0 and world == 1 and not (args.no_cpu_baseline or args.step_only):
        r = cpu_oracle_run(3, 1, sample_b=B_PER_GPU)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}
    if rank == 0:
        by_key = {r["key"]: r for r in (roofs or [])}
        # dominant kernel = the largest share of the timed step: the fused decoder FeedForward (6 launches x ~165 us of a
        # ~6 ms step; the launch list under profiles/ has the shares), then the decoder attention (6 x ~127 us)
        dominant = by_key.get("ffn_fused")
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "ms_per_step_by_rank": rank_ms, "sm_mhz_by_rank": clocks_by_rank,
                "config": {"workload": WORKLOAD, "batch_per_gpu": B_PER_GPU, "phonemes": L_PHON,
                           "duration_precision": model.duration_precision,
                           "variance_precision": model.variance_precision,
                           "precision_note": "fp32 = that chain (encoder + duration predictor; pitch / energy predictors) "
                                             "runs to fp32 accuracy on split-fp16 x3 tensor-core GEMMs, so durations, frame "
                                             "indices, mel lengths and pitch / energy buckets equal the fp32 reference's; "
                                             "decoder and PostNet are bf16 in every mode",
                           "frames_per_step_rank0": frames, "t_max_rank0": tmax,
                           "sharding": f"one pool of {B_PER_GPU * world} utterances dealt by sharding.shard_by_length "
                                       f"over {world} rank(s), no data-path collective",
                           "l2": "working set per step (>1 GB of activations) exceeds the 126 MB L2; no explicit flush"},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(text_h.numel() * 8 + mask_h.numel()),
                        "d2h_bytes_per_step": int(B_PER_GPU * tmax * hp.mel_dim * 4), "ms_per_step": e2e_ms / args.steps},
                "gpu_launches": launches, "launches_per_step": launches / args.steps, "clocks": clocks,
                "tma_descriptor_cache": tmap_stats(),
                "roofline": dominant, "kernel_rooflines": roofs,
                # the same step in the other precision modes: fp32 = that predictor chain runs to fp32 accuracy so its
                # discrete decisions (durations / pitch-energy buckets) equal the fp32 reference's; bf16 = plain bf16
                "other_precision_modes": modes,
                "torch_eager_b200": eager, "latency_b1": lat, "train": trn, "tts_ar_decode": ar, "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def make_train_batch(rank: int, B: int):
    """cfg3: ragged text, alignment in [2,15) on valid phonemes (T ~ 1000), N(0,1) mel, -0.5 padding."""
    import torch

    from transformer_tts_b200 import synth

    text, pos, _ = synth.synth_text_batch(B, L_PHON, ragged=True, seed=1234 + rank, min_len=L_PHON // 2)
    g = torch.Generator().manual_seed(4321 + rank)
    dur = torch.randint(2, 15, (B, L_PHON), generator=g)
    dur = torch.where(pos != 0, dur, torch.zeros_like(dur))
    mel_len = dur.sum(1)
    T = int(mel_len.max())
    ar = torch.arange(1, T + 1).unsqueeze(0).expand(B, T)
    pos_mel = torch.where(ar <= mel_len.unsqueeze(1), ar, torch.zeros_like(ar))
    valid = pos_mel != 0
    f0 = torch.where(valid, torch.rand((B, T), generator=g) * 729.0 + 71.0, torch.zeros(B, T))
    en = torch.where(valid, torch.rand((B, T), generator=g) * 400.0, torch.zeros(B, T))
    mel = synth.synth_mel(pos_mel, 80, seed=1234 + rank)
    return dict(text=text, src_mask=(pos != 0).unsqueeze(-2), mel_mask=valid.unsqueeze(-2), alignment=dur, f0=f0,
                energy=en, mel=mel), int(mel_len.sum())


def train(args):
    """BASELINE.json configs[2]: FastSpeech2 training step (fwd + bwd + clip + Adam/Noam), batch 32 per GPU,
    data parallel over N GPUs with the gradient all-reduce through NCCL.  Not the headline metric: run with
    `--workload train`.  value = valid mel frames trained per second over all ranks."""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from transformer_tts_b200 import build_fastspeech2, lib, synth
    from transformer_tts_b200.training import Trainer

    lib.load()
    hp = synth.make_hp()
    model = build_fastspeech2(hp)
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda().train()
    tr = Trainer(model, optimizer=args.optimizer, overlap_allreduce=args.overlap)
    tr.broadcast_parameters()
    Bt = args.train_batch
    host, frames = make_train_batch(rank, Bt)
    host = {k: v.pin_memory() for k, v in host.items()}
    dev = {k: v.cuda() for k, v in host.items()}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allr(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=op)
        return float(t.item())

    for _ in range(max(3, args.warmup)):
        out = tr.train_step(**dev)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    l0 = lib.LAUNCH_COUNT
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = tr.train_step(**dev)
    e1.record()
    barrier()
    launches = lib.LAUNCH_COUNT - l0
    dev_ms = allr(e0.elapsed_time(e1), dist.ReduceOp.MAX)
    loss_h = torch.empty(1).pin_memory()
    barrier()
    e0.record()
    for _ in range(args.steps):
        d = {k: v.cuda(non_blocking=True) for k, v in host.items()}
        out = tr.train_step(**d)
        loss_h.copy_(out["loss"].reshape(1), non_blocking=True)
    e1.record()
    barrier()
    e2e_ms = allr(e0.elapsed_time(e1), dist.ReduceOp.MAX)
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    total_frames = allr(float(frames), dist.ReduceOp.SUM)
    if rank == 0:
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        line = {"metric": "fastspeech2_train_mel_frames_per_sec", "value": total_frames * args.steps / (dev_ms * 1e-3),
                "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"fastspeech2_train_step_bf16_B{Bt}_per_gpu_L128_T~1000_"
                                       + ("radam_lr1e-3_clip" if args.optimizer == "radam" else "adam_noam_clip"),
                           "batch_per_gpu": Bt, "frames_per_step_per_gpu": frames, "t_max": int(dev["mel"].shape[1]),
                           "parallelism": f"dp{world}, NCCL all-reduce of one flat fp32 gradient buffer "
                                          f"({tr.flat_g.numel() * 4 / 1e6:.0f} MB)" + (", SyncBatchNorm" if world > 1 else ""),
                           "l2": "working set per step (several GB of saved activations) exceeds the 126 MB L2"},
                "e2e": {"value": total_frames * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms / args.steps},
                "gpu_launches": launches, "clocks": sampler.summary(), "loss": float(out["loss"])}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eager-baseline", action="store_true", help="skip the torch_eager_b200 sub-record")
    ap.add_argument("--no-train-record", action="store_true", help="skip the training-step sub-record")
    ap.add_argument("--no-ar-record", action="store_true", help="skip the Transformer-TTS autoregressive decode sub-record")
    ap.add_argument("--no-train-graph", action="store_true", help="training sub-record with eager steps (no CUDA graph)")
    ap.add_argument("--step-only", action="store_true",
                    help="only the timed steps (no per-kernel rooflines, latency, other modes, baselines): for ncu launch lists")
    ap.add_argument("--duration-precision", default="fp32", choices=["fp32", "bf16"],
                    help="fp32 (default): durations equal the fp32 reference's; bf16: the faster chain (SURVEY F9)")
    ap.add_argument("--variance-precision", default="fp32", choices=["fp32", "bf16"],
                    help="fp32 (default): pitch / energy buckets equal the fp32 reference's too (needs fp32 durations)")
    ap.add_argument("--workload", default="infer", choices=["infer", "train", "ar"])
    ap.add_argument("--train-batch", type=int, default=32)
    ap.add_argument("--overlap", action="store_true",
                    help="train: overlapped tail/head gradient exchange instead of one all-reduce after the backward")
    ap.add_argument("--optimizer", default="radam", choices=["radam", "adam"],
                    help="train workload: radam (BASELINE.json configs[2], radam.py) or adam + Noam rate (the "
                         "reference's working default, SURVEY F5)")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    elif args.workload == "train":
        train(args)
    elif args.workload == "ar":   # BASELINE.json configs[4] alone (one GPU)
        print(json.dumps(ar_decode_record(cpu=not args.no_cpu_baseline)), flush=True)
    else:
        ours(args)


if __name__ == "__main__":
    main()
