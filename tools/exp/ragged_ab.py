"""Free-running cfg2 step with and without the ragged tile list (Models/_engine.RAGGED_TILES): time and parity."""
import sys, time; sys.path.insert(0, '.')
import torch
from transformer_tts_b200 import build_fastspeech2, lib, synth
from transformer_tts_b200.Models import _engine as E
lib.load()
hp = synth.make_hp()
model = build_fastspeech2(hp).eval(); model.load_state_dict(synth.synth_state_dict(hp)); model = model.cuda()
outs = {}
for seed in (1234, 1236):
    text, pos, lens = synth.synth_text_batch(64, 128, seed=seed)
    text, src_mask = text.cuda(), (pos != 0).unsqueeze(-2).cuda()
    for rag in (False, True, False, True):
        E.RAGGED_TILES = rag
        with torch.no_grad():
            for _ in range(3): o = model(text, src_mask)
            torch.cuda.synchronize(); time.sleep(0.5)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20): o = model(text, src_mask)
            e1.record(); torch.cuda.synchronize()
        ml = model.last_aux["mel_len"].long()
        outs[rag] = (o[1].clone(), o[0].clone(), ml.clone())
        print(f"seed {seed} ragged={rag}: {e0.elapsed_time(e1) / 20:.3f} ms/step  Tmax {o[0].shape[1]} frames {int(ml.sum())}")
    a, b, ml = outs[False][0], outs[True][0], outs[True][2]
    assert torch.equal(outs[False][2], outs[True][2])
    worst = 0.0
    for i in range(64):
        n = int(ml[i])
        worst = max(worst, float((a[i, :n] - b[i, :n]).abs().max()))
        assert torch.equal(outs[False][1][i, :n], outs[True][1][i, :n]), f"mel_pre differs for utterance {i}"
    print(f"seed {seed}: valid-frame max |diff| mel_post {worst:.3e} (mel_pre bit-identical); padded finite: {bool(torch.isfinite(b).all())}")
