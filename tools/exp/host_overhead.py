"""Host-side cost of one free-running forward: wall time of the Python launch sequence vs the GPU step time."""
import cProfile, pstats, sys, time
sys.path.insert(0, '.')
import torch
from transformer_tts_b200 import build_fastspeech2, lib, synth
lib.load()
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
text, pos, lens = synth.synth_text_batch(64, 128, seed=1234)
text, src_mask = text.cuda(), (pos != 0).unsqueeze(-2).cuda()
with torch.no_grad():
    for _ in range(5): model(text, src_mask)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20): model(text, src_mask)
    torch.cuda.synchronize()
    print(f"wall per step {(time.perf_counter() - t0) / 20 * 1e3:.3f} ms")
    pr = cProfile.Profile(); pr.enable()
    for _ in range(20): model(text, src_mask)
    torch.cuda.synchronize()
    pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(18)
