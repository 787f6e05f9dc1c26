"""Kernel-time breakdown of one TTSTrainer.train_step at BASELINE.json configs[3]'s shape (torch.profiler, CUPTI)."""
import sys

import torch

sys.path.insert(0, ".")
import bench  # noqa: E402


def main():
    from torch.profiler import ProfilerActivity, profile

    from oracle import transformer_tts_oracle as T
    from transformer_tts_b200 import build_transformer_tts, synth
    from transformer_tts_b200.tts_training import TTSTrainer

    hp = synth.make_tts_hp()
    r = hp.reduction_rate
    if "--dropout" in sys.argv:   # the reference's training construction (train.py:117-126)
        from transformer_tts_b200.Models.transformer import Transformer

        model = Transformer(hp=hp, src_vocab=hp.vocab_size, trg_vocab=hp.mel_dim, d_model_encoder=hp.d_model_encoder,
                            N_e=hp.n_layer_encoder, n_head_encoder=hp.n_head_encoder,
                            ff_conv_kernel_size_encoder=hp.ff_conv_kernel_size_encoder,
                            concat_after_encoder=hp.concat_after_encoder, d_model_decoder=hp.d_model_decoder,
                            N_d=hp.n_layer_decoder, n_head_decoder=hp.n_head_decoder,
                            ff_conv_kernel_size_decoder=hp.ff_conv_kernel_size_decoder,
                            concat_after_decoder=hp.concat_after_decoder, reduction_rate=r, dropout=0.1, dropout_prenet=0.5,
                            dropout_postnet=0.5)
    else:
        model = build_transformer_tts(hp)
    model.load_state_dict(synth.synth_tts_state_dict(hp))
    model = model.cuda().train()
    batch, phonemes, mel_len = 32, 128, 800
    text, pos, _ = synth.synth_text_batch(batch, phonemes, ragged=True, seed=81, min_len=phonemes // 2)
    g = torch.Generator().manual_seed(82)
    lens = torch.randint(mel_len // 2 // r, mel_len // r + 1, (batch,), generator=g) * r
    lens[0] = mel_len
    pos_mel = torch.arange(1, mel_len + 1).unsqueeze(0).repeat(batch, 1)
    pos_mel[pos_mel > lens.unsqueeze(1)] = 0
    mel = torch.randn(batch, mel_len, hp.mel_dim, generator=g) * (pos_mel != 0).unsqueeze(-1)
    stop = (pos_mel == 0).float()
    mel_in, pos_in = mel[:, :-r:r, :].contiguous(), pos_mel[:, :-r:r].contiguous()
    src_mask, _ = T.create_masks(pos, pos_in)
    a = (text.cuda(), mel_in.cuda(), src_mask.cuda(), (pos_in != 0).to(torch.uint8).cuda(), mel[:, r:, :].contiguous().cuda(),
         stop[:, r:].contiguous().cuda())
    tr = TTSTrainer(model, positive_weight=5.0)
    for _ in range(3):
        tr.train_step(*a)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for name, fn in (("forward_train", lambda: tr.forward_train(*a[:4])), ("train_step", lambda: tr.train_step(*a))):
        fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{name}: {e0.elapsed_time(e1) / 5:.3f} ms")
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            tr.train_step(*a)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))


if __name__ == "__main__":
    main()
