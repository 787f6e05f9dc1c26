"""Kernel-time sum vs wall time of one FastSpeech2 Trainer.train_step at the bench shape (torch.profiler, CUPTI)."""
import sys

import torch

sys.path.insert(0, ".")
import bench  # noqa: E402


def main():
    from torch.profiler import ProfilerActivity, profile

    from transformer_tts_b200 import build_fastspeech2, synth
    from transformer_tts_b200.training import Trainer

    hp = synth.make_hp()
    if "--dropout" in sys.argv:   # the reference's training construction (train_fastspeech2.py:381-389)
        model = build_fastspeech2(hp, dropout=0.1, dropout_postnet=0.5, dropout_variance_adaptor=0.5)
    else:
        model = build_fastspeech2(hp)
    model.load_state_dict(synth.synth_state_dict(hp))
    model = model.cuda().train()
    tr = Trainer(model, optimizer="radam")
    print("dropout active:", tr.drop.active, tr.drop.p_stack)
    host, frames = bench.make_train_batch(0, 32)
    dev = {k: v.cuda() for k, v in host.items()}
    for _ in range(3):
        tr.train_step(**dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        tr.train_step(**dev)
    e1.record()
    torch.cuda.synchronize()
    print(f"train_step: {e0.elapsed_time(e1) / 5:.3f} ms")
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            tr.train_step(**dev)
        torch.cuda.synchronize()
    ka = prof.key_averages()
    tot = sum(k.self_device_time_total for k in ka) / 3e3
    n = sum(k.count for k in ka if k.self_device_time_total > 0) / 3
    print(f"kernel-time sum per step: {tot:.3f} ms over {n:.0f} device activities")
    print(ka.table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=70))


if __name__ == "__main__":
    main()
