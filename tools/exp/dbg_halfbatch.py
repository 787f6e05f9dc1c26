import sys; sys.path.insert(0, '.')
import torch
from transformer_tts_b200 import build_fastspeech2, lib, synth
lib.load()
hp = synth.make_hp(); sd = synth.synth_state_dict(hp)
model = build_fastspeech2(hp).eval(); model.load_state_dict(sd); model = model.cuda()
B, L = 64, 128
text, pos, lens = synth.synth_text_batch(B, L, ragged=True, seed=1234, min_len=L // 2)
text, pos = text.cuda(), pos.cuda()
src_mask = (pos != 0).unsqueeze(-2)
with torch.no_grad():
    out = model(text, src_mask); aux = dict(model.last_aux)
    out_h = model(text[16:48], src_mask[16:48]); aux_h = dict(model.last_aux)
ml = aux["mel_len"].long()
print("Tmax full", out[0].shape[1], "Tmax half", out_h[0].shape[1], "lens", lens[16:48].tolist())
names = ["mel_pre", "mel_post", "log_d", "p", "e", "va_x", "text_dur"]
for j, i in enumerate(range(16, 48)):
    n = int(ml[i])
    d = (out_h[1][j, :n] - out[1][i, :n]).abs().amax(-1)
    if float(d.max()) > 0:
        nz = torch.nonzero(d > 0).flatten()
        per = {}
        for k, nm in enumerate(names):
            a, b = out_h[k][j], out[k][i]
            m = min(a.shape[0], b.shape[0]) if a.dim() > 0 else 0
            lim = n if k not in (2,) else int(lens[i])
            per[nm] = float((a[:lim].float() - b[:lim].float()).abs().max())
        eo = float((aux_h["encoder_output"][j].float() - aux["encoder_output"][i].float()).abs().max())
        do_ = (aux_h["decoder_output"][j, :n].float() - aux["decoder_output"][i, :n].float()).abs().amax(-1)
        print(f"utt {i}: n={n} L={int(lens[i])} first diff frame {int(nz[0])} last {int(nz[-1])} count {len(nz)} maxdiff {float(d.max()):.3e} enc_out diff {eo:.2e} "
              f"dec_out first diff {int(torch.nonzero(do_ > 0).flatten()[0]) if (do_>0).any() else -1}", per)
