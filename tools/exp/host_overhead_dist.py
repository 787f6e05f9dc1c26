"""Per-rank wall time per free-running forward and the time the host spends blocked in the Tmax read-back
(≈0 => the host, not the GPU, paces the step).  torchrun --nproc-per-node N tools/exp/host_overhead_dist.py"""
import os, sys, time
sys.path.insert(0, '.')
import torch
lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
from transformer_tts_b200 import build_fastspeech2, lib, synth
lib.load()
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
text, pos, lens = synth.synth_text_batch(64, 128, seed=1234 + lr)
text, src_mask = text.cuda(), (pos != 0).unsqueeze(-2).cuda()
blocked = [0.0]
orig = torch.Tensor.item
def item(self):
    t = time.perf_counter(); r = orig(self); blocked[0] += time.perf_counter() - t; return r
torch.Tensor.item = item
with torch.no_grad():
    for _ in range(10): model(text, src_mask)
    torch.cuda.synchronize()
    time.sleep(0.5 + 0.0 * lr)
    blocked[0] = 0.0
    n = 100
    t0 = time.perf_counter()
    for _ in range(n): model(text, src_mask)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
print(f"rank {lr}: wall/step {(t2 - t0) / n * 1e3:.3f} ms, host blocked in item() {blocked[0] / n * 1e3:.3f} ms/step, "
      f"final drain {(t2 - t1) * 1e3:.2f} ms, cpus {len(os.sched_getaffinity(0))}", flush=True)
