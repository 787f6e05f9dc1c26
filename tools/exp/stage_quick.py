"""Whole free-running step with different conv1 -> conv2 group sizes (Models/_engine.FFN_GROUP_BYTES)."""
import sys; sys.path.insert(0, '.')
import torch
from transformer_tts_b200 import build_fastspeech2, lib, synth
from transformer_tts_b200.Models import _engine as E
lib.load()
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
text, pos, lens = synth.synth_text_batch(64, 128, seed=1234)
text, src_mask = text.cuda(), (pos != 0).unsqueeze(-2).cuda()
def run(n=20):
    with torch.no_grad():
        for _ in range(3): o = model(text, src_mask)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): o = model(text, src_mask)
        e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, o
import time
for mb in (0, 1, 2, 0, 1, 2, 3):
    E.FFN_GROUP_WAVES = mb
    time.sleep(1.0)
    ms, o = run()
    print(f"FFN_GROUP_WAVES={mb} (group {E.ffn_chunk_utterances(64, o[0].shape[1], 384)} utt): {ms:.3f} ms/step  mel checksum {float(o[1].double().abs().sum()):.8e}")
