"""ttsb200_length_regulate_static: the LengthRegulator entry point that allocates nothing (its scratch is a static
device array inside the library).  No counterpart in the reference.

Called through the C-ABI and compared BIT-EXACTLY (integer / copy work) with the existing ttsb200_lr_scan +
ttsb200_lr_expand pair, on two calls in a row with different shapes (the second call sees the first call's leftovers
in the shared scratch), plus the size-limit error.  Collected last on purpose.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_BATCH = 1024       # TTSB200_LR_STATIC_MAX_BATCH
MAX_PHONEMES = 65536   # TTSB200_LR_STATIC_MAX_PHONEMES


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _inputs(B, L, D, seed, max_dur):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, L, D, generator=g).to(torch.bfloat16).cuda()
    dur = torch.randint(-1, max_dur + 1, (B, L), generator=g, dtype=torch.int32)  # negatives count as 0
    lens = torch.randint(1, L + 1, (B,), generator=g)
    dur[torch.arange(L)[None, :] >= lens[:, None]] = 0   # ragged batch: padded phonemes have duration 0
    dur[0, : lens[0]] = max_dur                            # one utterance at the maximum
    return x, dur.cuda()


def _static_call(dll, x, dur, Tmax, want_f32):
    B, L, D = x.shape
    dev = x.device
    out = torch.full((B, Tmax, D), float("nan"), device=dev, dtype=torch.bfloat16)
    frame_idx = torch.full((B, Tmax), -7, device=dev, dtype=torch.int32)
    mel_pos = torch.full((B, Tmax), -7, device=dev, dtype=torch.int64)
    mel_mask = torch.full((B, Tmax), 7, device=dev, dtype=torch.uint8)
    out32 = torch.full((B, Tmax, D), float("nan"), device=dev, dtype=torch.float32) if want_f32 else None
    mel_len = torch.full((B,), -7, device=dev, dtype=torch.int32)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = dll.ttsb200_length_regulate_static(_p(x), _p(dur), _p(out), _p(frame_idx), _p(mel_pos), _p(mel_mask),
                                            _p(out32), _p(mel_len), B, L, Tmax, D, stream)
    assert rc == 0, dll.ttsb200_last_error_string().decode()
    torch.cuda.synchronize()
    return out, frame_idx, mel_pos, mel_mask, out32, mel_len


def _existing_call(lib, x, dur, Tmax, want_f32):
    B, _, D = x.shape
    cum, mel_len = lib.lr_scan(dur)
    out32 = torch.full((B, Tmax, D), float("nan"), device=x.device, dtype=torch.float32) if want_f32 else None
    out, frame_idx, mel_pos, mel_mask = lib.lr_expand(x, cum, mel_len, Tmax, want_aux=True, out_f32=out32)
    torch.cuda.synchronize()
    return out, frame_idx, mel_pos, mel_mask, out32, mel_len


def test_length_regulate_static_matches_scan_expand_on_two_calls_in_a_row():
    from transformer_tts_b200 import lib

    dll = lib.load()
    # call 1: the cfg2 shape; call 2: a smaller, differently shaped batch that reuses (part of) the same scratch
    for (B, L, D, max_dur, seed, want_f32) in [(64, 128, 384, 12, 1, True), (3, 37, 256, 9, 2, False)]:
        x, dur = _inputs(B, L, D, seed, max_dur)
        want_len = dur.clamp(min=0).sum(1)
        Tmax = int(want_len.max())
        got = _static_call(dll, x, dur, Tmax, want_f32)
        ref = _existing_call(lib, x, dur, Tmax, want_f32)
        assert torch.equal(got[5].cpu(), want_len.to(torch.int32).cpu())
        names = ["out_bf16", "frame_idx", "mel_pos", "mel_mask", "out_f32", "mel_len"]
        for name, a, b in zip(names, got, ref):
            if a is None:
                assert b is None
                continue
            if a.dtype.is_floating_point:  # bit-exact: the entry point copies rows, it does no arithmetic
                a, b = a.view(torch.int16 if a.dtype == torch.bfloat16 else torch.int32), b.view(
                    torch.int16 if b.dtype == torch.bfloat16 else torch.int32)
            assert torch.equal(a, b), f"{name} differs at shape {(B, L, D)}"


def test_length_regulate_static_rejects_sizes_beyond_the_scratch():
    from transformer_tts_b200 import lib

    dll = lib.load()
    x = torch.zeros(8, dtype=torch.bfloat16, device="cuda")     # never dereferenced: the size check comes first
    d = torch.zeros(8, dtype=torch.int32, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for B, L in [(MAX_BATCH + 1, 1), (64, MAX_PHONEMES // 64 + 1)]:
        rc = dll.ttsb200_length_regulate_static(_p(x), _p(d), _p(x), None, None, None, None, None, B, L, 16, 384, stream)
        assert rc == -1
        msg = dll.ttsb200_last_error_string().decode()
        assert "exceed the static scratch" in msg and str(MAX_PHONEMES) in msg, msg
    torch.cuda.synchronize()
