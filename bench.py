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

`--impl reference` times the reference algorithm's CPU implementation (the oracle port, fp32, all host
threads) on a bounded sample of the same workload — the unmodified reference is Python that cannot
travel to the GPU box (it lives outside the repo), the oracle is its line-by-line restatement pinned
by tests/golden.
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
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.stop_flag = threading.Event()
        self.samples = []

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([s.strip() for s in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx = max(mx, float(s[1]))
                for n, v in zip(names, s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_batch(rank: int):
    from transformer_tts_b200 import synth

    text, pos, lens = synth.synth_text_batch(B_PER_GPU, L_PHON, seed=1234 + rank)
    return text, pos, lens


def cpu_oracle_run(steps: int, warmup: int, sample_b: int):
    """Times the oracle (reference algorithm, fp32, torch CPU) on `sample_b` utterances of the workload."""
    import torch

    from oracle import fastspeech2_oracle as O
    from transformer_tts_b200 import synth

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    text, pos, _ = make_batch(0)
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
    r = cpu_oracle_run(max(1, min(args.steps, 20)), 1, sample_b=8)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "CPU, bounded sample"},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def kernel_rooflines(model, peaks, tmax):
    """Per-kernel rooflines at the decoder shape of the workload (B x Tmax rows): every hot kernel is timed
    alone with CUDA events around a captured graph of 12 launches, rotating over three buffer sets so that
    consecutive launches never find their operands in the 126 MB L2.  Algorithmic work per launch as in DESIGN.md."""
    import torch

    from transformer_tts_b200 import lib
    from transformer_tts_b200.Models import _engine as E

    B, T, D, H = B_PER_GPU, int(tmax), 384, 4
    dev = "cuda"
    lc = E.layer_cache(model.decoder.layers[0])
    k = lc["k"]
    NS = 3
    g = torch.Generator(device="cpu").manual_seed(7)
    lens = torch.randint(int(0.8 * T), T + 1, (B,), generator=g)
    lens[0] = T
    mask = (torch.arange(T)[None, :] < lens[:, None]).to(torch.uint8).to(dev)
    kv = lib.mask_lengths(mask)
    rows_valid = float(lens.sum())
    xs = [(torch.randn(B, T, D, device=dev) * 0.5).to(torch.bfloat16) for _ in range(NS)]
    x32 = [torch.randn(B, T, D, device=dev) for _ in range(NS)]
    qkvs = [(torch.randn(B, T, 3 * D, device=dev) * 0.5).to(torch.bfloat16) for _ in range(NS)]
    hs = [torch.relu(torch.randn(B, T, 4 * D, device=dev)).to(torch.bfloat16) for _ in range(NS)]
    ob = [torch.empty(B, T, D, device=dev, dtype=torch.bfloat16) for _ in range(NS)]
    o32 = [torch.empty(B, T, D, device=dev) for _ in range(NS)]
    L = 128
    xe = [(torch.randn(B, L, D, device=dev)).to(torch.bfloat16) for _ in range(NS)]
    dur = torch.full((B, L), T // L, dtype=torch.int32, device=dev)
    cum, mel_len = lib.lr_scan(dur)
    Tlr = int(mel_len.max().item())
    va = model.variance_adaptor
    pb, eb = va._bins(torch.device(dev))
    tabs = va._tables()
    pv = [torch.rand(B, T, device=dev) * 700 + 80 for _ in range(NS)]
    ev_ = [torch.rand(B, T, device=dev) * 400 for _ in range(NS)]
    ln = lc["ln1"]

    def t_qkv(i):
        lib.conv_gemm(xs[i], lc["wqkv"], lc["bqkv"], block_n=192, out_bf16=qkvs[i])

    def t_att(i):
        lib.attention_fwd(qkvs[i], mask, kv, ob[i], H)

    def t_out(i):
        lib.conv_gemm(xs[i], lc["wo"], lc["bo"], block_n=D, epi=lib.EPI_RES_LN, res_f32=x32[i], res_bstride=T * D,
                      res_tstride=D, ln1=lc["ln2"], out_f32=o32[i], out_bf16=ob[i])

    def t_ffn1(i):
        lib.conv_gemm(xs[i], lc["w1"], lc["b1"], taps=k, pad_left=k // 2, block_n=256, act=lib.ACT_RELU, out_bf16=hs[i])

    def t_ffn2(i):
        lib.conv_gemm(hs[i], lc["w2"], lc["b2"], taps=k, pad_left=k // 2, block_n=D, epi=lib.EPI_FFN2, res_bf16=xs[i],
                      res_f32=x32[i], res_bstride=T * D, res_tstride=D, ln1=lc["ln_ff"], ln2=ln, out_f32=o32[i],
                      out_bf16=ob[i])

    def t_lr(i):
        lib.lr_expand(xe[i], cum, mel_len, Tlr, want_aux=True)

    def t_bucket(i):
        lib.bucket_embed_add(xs[i], pv[i], ev_[i], pb, eb, tabs["ptab"], tabs["etab"])

    def t_ln(i):
        lib.layernorm(x32[i], ln[0], ln[1], E.LN_EPS, y_bf16=ob[i])

    M = float(B * T)
    att_flops = float((4.0 * lens.double() ** 2 * D).sum())
    s2 = 2  # bf16 bytes
    # (name, launcher, algorithmic FLOPs, algorithmic HBM bytes): the kernel is reported against whichever roof
    # gives the LARGER floor time (DESIGN.md section 4 lists both figures)
    specs0 = [
        ("conv_gemm<192,AFFINE,A-stationary> decoder q|k|v linears", t_qkv, 2 * M * D * 3 * D, M * D * s2 * 4),
        ("attention_fwd decoder masked self-attention", t_att, att_flops, M * D * s2 * 4),
        ("conv_gemm<384,RES_LN> decoder out-proj + residual + LayerNorm", t_out, 2 * M * D * D, M * D * (2 * s2 + 8)),
        ("conv_gemm<256,AFFINE,A-stationary> decoder FFN conv1 + ReLU", t_ffn1, 2 * M * D * 4 * D * k, M * D * s2 * 5),
        ("conv_gemm<384,FFN2> decoder FFN conv2 + LN + residual + LN", t_ffn2, 2 * M * D * 4 * D * k,
         M * D * (s2 * 6 + 8)),
        ("lr_expand LengthRegulator frame gather", t_lr, 0,
         B * L * D * s2 + B * L * 4 + B * Tlr * D * s2 + B * Tlr * (4 + 8 + 1) + B * 4),
        ("bucket_embed_add pitch/energy bucketize + embedding add", t_bucket, 0, M * D * s2 * 2 + M * 8),
        ("layernorm standalone fp32 -> bf16", t_ln, 0, M * D * (4 + s2)),
    ]
    specs = []
    for name, fn, flops, nbytes in specs0:
        t_tc = flops / (peaks["tc_burst"] * 1e12)
        t_hbm = nbytes / (peaks["hbm"] * 1e9)
        specs.append((name, fn, "tensor", flops) if t_tc >= t_hbm else (name, fn, "hbm", nbytes))
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from ONE `ncu --set full` capture of the same shapes
    # (profiles/r01_hot_kernels_v5_full_summary.csv, tools/run_hot_kernels.py).  Writes that are still resident in
    # the 126 MB write-back L2 when the kernel ends are not counted by the DRAM counters.
    traffic_mb = [55.2 + 109.0, 154.9 + 36.8, 163.4 + 112.9, 56.3 + 162.8, 389.4 + 128.1, 6.3 + 2.4, 55.5 + 7.1,
                  108.4 + 8.0]
    out = []
    reps = 4
    for (name, fn, bound, work), tr_mb in zip(specs, traffic_mb):
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
        if bound == "tensor":
            ach, peak, unit = work / (ms * 1e-3) / 1e12, peaks["tc_burst"], "TFLOP/s"
        else:
            ach, peak, unit = work / (ms * 1e-3) / 1e9, peaks["hbm"], "GB/s"
        out.append({"kernel": name, "bound": bound, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak,
                    "traffic": tr_mb * 1e6, "ms_per_launch": ms, "algorithmic_work_per_launch": work,
                    "peak_source": peaks["src"] + (" burst bf16 (kernel timed alone)" if bound == "tensor" else " copy bandwidth")})
    return out


def ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl")
    from transformer_tts_b200 import build_fastspeech2, lib, synth

    lib.load()
    hp = synth.make_hp()
    model = build_fastspeech2(hp).eval()
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda()
    text_h, pos_h, _ = make_batch(rank)
    text_h, mask_h = text_h.pin_memory(), (pos_h != 0).unsqueeze(-2).pin_memory()
    text_d, mask_d = text_h.cuda(), mask_h.cuda()
    peaks = read_peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    frames = 0
    with torch.no_grad():
        for _ in range(max(3, args.warmup)):
            model(text_d, mask_d)
        frames = int(model.last_aux["mel_len"].sum().item())
        tmax = int(model.last_aux["mel_len"].max().item())
        # ---- device-resident throughput ----
        sampler = ClockSampler(local_rank)
        sampler.start()
        barrier()
        l0 = lib.LAUNCH_COUNT
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            model(text_d, mask_d)
        e1.record()
        barrier()
        launches = lib.LAUNCH_COUNT - l0
        my_ms = e0.elapsed_time(e1)
        dev_ms = max_over_ranks(my_ms)
        rank_ms = [my_ms / args.steps]
        if world > 1:  # the step is the max over ranks: report every rank's own time next to it
            g = [torch.zeros(1, device="cuda", dtype=torch.float64) for _ in range(world)]
            dist.all_gather(g, torch.tensor([my_ms / args.steps], device="cuda", dtype=torch.float64))
            rank_ms = [round(float(x.item()), 4) for x in g]
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
        t0 = time.perf_counter()
        e0.record()
        for i in range(args.steps):
            e2e_step(i)
        main_s.wait_stream(copy_s)
        e1.record()
        barrier()
        e2e_ms = max_over_ranks(e0.elapsed_time(e1))
        sampler.stop_flag.set()
        sampler.join(timeout=2)
        roofs = kernel_rooflines(model, peaks, tmax) if rank == 0 else None
    total_frames = sum_over_ranks(float(frames))
    value = total_frames * args.steps / (dev_ms * 1e-3)
    e2e_value = total_frames * args.steps / (e2e_ms * 1e-3)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_oracle_run(10, 1, sample_b=8)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": WORKLOAD, "batch_per_gpu": B_PER_GPU, "phonemes": L_PHON,
                           "frames_per_step_per_gpu": frames, "t_max": tmax, "sharding": f"utterances x{world}, no collective",
                           "ms_per_step_by_rank": rank_ms,
                           "l2": "working set per step (>1 GB of activations) exceeds the 126 MB L2; no explicit flush"},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(text_h.numel() * 8 + mask_h.numel()),
                        "d2h_bytes_per_step": int(out_h[0].numel() * 4), "ms_per_step": e2e_ms / args.steps},
                "gpu_launches": launches, "clocks": sampler.summary(),
                # dominant kernel = largest share of the step (profiles/): the decoder attention
                "roofline": roofs[1], "kernel_rooflines": roofs, "cpu_baseline": cpu}
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
        dist.init_process_group("nccl")
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
    ap.add_argument("--workload", default="infer", choices=["infer", "train"])
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
    else:
        ours(args)


if __name__ == "__main__":
    main()
