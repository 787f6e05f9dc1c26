"""A short generate() for the ncu launch list of the decode kernels (B=16, 128 phonemes, 40 positions eager + graph)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from transformer_tts_b200 import build_transformer_tts, synth
hp = synth.make_tts_hp()
model = build_transformer_tts(hp).eval(); model.load_state_dict(synth.synth_tts_state_dict(hp)); model = model.cuda()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
text, pos, _ = synth.synth_text_batch(B, 128, ragged=True, seed=77, min_len=64)
model.use_cuda_graphs = False
out = model.generate(text.cuda(), (pos != 0).unsqueeze(-2).cuda(), max_steps=int(sys.argv[2]) if len(sys.argv) > 2 else 12)
torch.cuda.synchronize()
print("ok", out["mel"].shape)
