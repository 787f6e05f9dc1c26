"""In-situ stage times of the inference forward (CUDA events between the stages, weights warm in L2)."""
import sys, torch
sys.path.insert(0, '.')
import bench
from transformer_tts_b200 import build_fastspeech2, lib, synth
from transformer_tts_b200.Models import _engine as E
lib.load()
hp = synth.make_hp()
model = build_fastspeech2(hp).eval()
model.load_state_dict(synth.synth_state_dict(hp))
model = model.cuda()
text_h, pos_h, _ = bench.make_batch(0)
src, src_mask = text_h.cuda(), (pos_h != 0).unsqueeze(-2).cuda()
N = 20
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(N)]
with torch.no_grad():
    for it in range(N + 3):
        i = max(it - 3, 0)
        B, L = src.shape
        ev[i][0].record()
        sm = E.mask_to_u8(src_mask, B, L, src.device)
        kv_src = lib.mask_lengths(sm)
        e_bf, _, _ = model.encoder.run(src, sm, kv_src, want_f32=False, return_attn=False)
        ev[i][1].record()
        va = model.variance_adaptor.run(e_bf, sm, None, None, None, None, want_aux=True)
        ev[i][2].record()
        mel_u8 = va["mel_mask"]
        kv_mel = lib.mask_lengths(mel_u8)
        d_bf, _, _ = model.decoder.run(va["x"], mel_u8, kv_mel, want_f32=False, return_attn=False)
        ev[i][3].record()
        mel_pre, mel_post = model.postnet.run(d_bf)
        ev[i][4].record()
torch.cuda.synchronize()
names = ["encoder", "variance adaptor", "decoder", "mel linear + postnet"]
tot = 0
for k, n in enumerate(names):
    t = sum(ev[i][k].elapsed_time(ev[i][k + 1]) for i in range(N)) / N
    tot += t
    print(f"{n:22s} {t*1e3:8.1f} us")
print(f"{'sum':22s} {tot*1e3:8.1f} us")
