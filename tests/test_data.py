"""Data path: `transformer_tts_b200.data` against golden vectors made by the UNMODIFIED reference
`collate_fn` / `LengthsBatchSampler._batch_indices` (oracle/make_golden_data.py), and against the live reference when
/root/reference is present.  Integer fields and padding values: bit-exact; float fields: bit-exact too (same casts)."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden" / "fs2_collate.npz"

from oracle import make_golden_data as MG  # noqa: E402  (the item generator; test infrastructure)
from transformer_tts_b200 import data as D  # noqa: E402


def _check_tuple(name, tup, g):
    none_fields = set(g[f"{name}.none_fields"].tolist())
    assert len(tup) == 16
    for f, v in zip(MG.FIELDS, tup):
        if f == "mel_name":
            assert all(isinstance(x, str) for x in v)
            continue
        if f in none_fields:
            assert v is None or all(x is None for x in v), f"{name}.{f} should be None"
            continue
        ref = g[f"{name}.{f}"]
        got = v.numpy() if hasattr(v, "numpy") else np.asarray(v)
        assert got.dtype == ref.dtype and got.shape == ref.shape, f"{name}.{f}: {got.dtype}{got.shape} vs {ref.dtype}{ref.shape}"
        assert np.array_equal(got, ref), f"{name}.{f} differs"


def test_collate_matches_reference_golden():
    g = np.load(GOLDEN)
    for name, kw in MG.CASES:
        _check_tuple(name, D.collate_fn(MG.synth_items(**kw)), g)


def test_collate_padding_rules():
    items = MG.synth_items(n=4, seed=21, reduction_rate=4)
    t = D.collate_fn(items)
    text, mel, pos_text, pos_mel, tl, ml, stop = t[:7]
    T = int(ml.max())
    assert mel.shape[1] == D._round_up(T, 4) and mel.shape[1] % 4 == 0 and mel.dtype == torch.float32
    for i, d in enumerate(items):
        n, L = d["mel_length"], d["text_length"]
        assert (mel[i, n:] == -0.5).all() and (stop[i, n:] == 1.0).all()
        assert (text[i, L:] == 0).all() and (pos_text[i, L:] == 0).all() and (pos_mel[i, n:] == 0).all()
        assert torch.equal(pos_mel[i, :n], torch.arange(1, n + 1))
        assert (t[8][i, n:] == 0).all() and (t[10][i, L:] == 0).all()   # f0, alignment
    src_mask, mel_mask = D.create_masks(pos_text, pos_mel)
    assert src_mask.shape == (4, 1, text.shape[1]) and mel_mask.dtype == torch.bool
    assert torch.equal(mel_mask.sum(-1).flatten(), ml)
    with pytest.raises(TypeError):
        D.collate_fn([[1, 2, 3]])
    bad = MG.synth_items(n=2, seed=1)
    for d in bad:
        d["f0"] = None
    with pytest.raises(TypeError):   # the reference dies the same way (FloatTensor(None))
        D.collate_fn(bad)


def test_frame_budget_batching_matches_reference_golden():
    g = np.load(GOLDEN)
    k = 0
    while f"sampler{k}.lengths" in g.files:
        lengths, budget = g[f"sampler{k}.lengths"], int(g[f"sampler{k}.budget"])
        batches = D.frame_budget_batches(lengths, budget)
        assert len(batches) == int(g[f"sampler{k}.n_batches"])
        assert np.array_equal(np.array([i for b in batches for i in b]), g[f"sampler{k}.index"])
        assert np.array_equal(np.array([b for b, idx in enumerate(batches) for _ in idx]), g[f"sampler{k}.batch_of"])
        for idx in batches:  # the budget holds on the PADDED size
            assert max(lengths[i] for i in idx) * len(idx) <= budget
        k += 1
    assert k >= 4
    with pytest.raises(ValueError):  # the reference loops forever on an utterance longer than the budget
        D.frame_budget_batches([100, 5000, 100], 1000)
    assert D.frame_budget_batches([], 100) == [] and D.frame_budget_batches([50], 100) == []  # reference: < 2 left


def test_samplers(tmp_path):
    items = MG.synth_items(n=12, seed=8)

    class DS(torch.utils.data.Dataset):
        def __len__(self):
            return len(items)

        def __getitem__(self, i):
            return items[i]

    class HP:
        lengths_file = str(tmp_path / "lengths.npy")

    s = D.LengthsBatchSampler(DS(), 400, HP, lengths_file=None, shuffle=False)
    assert os.path.exists(HP.lengths_file)
    assert np.array_equal(np.load(HP.lengths_file), np.array([d["mel_length"] for d in items]))
    s2 = D.LengthsBatchSampler(DS(), 400, HP, lengths_file=HP.lengths_file, shuffle=False, reverse=True)
    assert list(s) == list(reversed(list(s2))) and len(s) == len(s2)
    flat = [i for b in s for i in b]
    assert flat == sorted(flat) and len(set(flat)) == len(flat)
    loader = torch.utils.data.DataLoader(DS(), batch_sampler=s, collate_fn=D.collate_fn)
    n = 0
    for batch in loader:
        assert batch[1].shape[0] * batch[1].shape[1] <= 400
        n += batch[0].shape[0]
    assert n == len(flat)
    nb = D.NumBatchSampler(DS(), 5, shuffle=False)
    assert list(nb) == [[0, 1, 2, 3, 4], [5, 6, 7, 8, 9]] and len(nb) == 2
    with pytest.raises(AssertionError):
        D.LengthsBatchSampler(DS(), 400, HP, shuffle=True, reverse=True)


@pytest.mark.skipif(not os.path.isdir("/root/reference/datasets"), reason="live reference only in the build container")
def test_collate_and_batching_match_live_reference():
    R = MG.load_reference_datasets()
    for seed, rr, multi in [(31, 1, ""), (32, 2, ""), (33, 5, "x_vector")]:
        items = MG.synth_items(n=6, seed=seed, reduction_rate=rr, multi=multi)
        a, b = D.collate_fn(items), R.collate_fn(items)
        assert len(a) == len(b)
        for x, y in zip(a, b):
            if y is None:
                assert x is None
            elif torch.is_tensor(y):
                assert x.dtype == y.dtype and torch.equal(x, y)
            else:
                assert list(x) == list(y)
    rng = np.random.RandomState(2)
    for _ in range(20):
        lengths = rng.randint(50, 1200, size=int(rng.randint(2, 60)))
        budget = int(rng.randint(1200, 8000))
        s = R.LengthsBatchSampler.__new__(R.LengthsBatchSampler)
        s.lengths_np, s.n_lengths = lengths, budget
        assert D.frame_budget_batches(lengths, budget) == s._batch_indices()


@pytest.mark.gpu
def test_device_prefetcher_feeds_training_step():
    """The prefetcher's batches equal the collated host batches, arrive on the device with their masks, and a
    Trainer step runs on them."""
    from transformer_tts_b200 import build_fastspeech2, synth
    from transformer_tts_b200.training import Trainer

    items = MG.synth_items(n=8, seed=13)

    class DS(torch.utils.data.Dataset):
        def __len__(self):
            return len(items)

        def __getitem__(self, i):
            return items[i]

    loader = torch.utils.data.DataLoader(DS(), batch_size=4, collate_fn=D.collate_fn)
    host = list(loader)
    hp = synth.make_hp()
    model = build_fastspeech2(hp)
    model.load_state_dict(synth.synth_state_dict(hp))
    tr = Trainer(model.cuda().train())
    seen = 0
    for hb, db in zip(host, D.DevicePrefetcher(loader)):
        assert db["text"].is_cuda and torch.equal(db["text"].cpu(), hb[0]) and torch.equal(db["mel"].cpu(), hb[1])
        assert torch.equal(db["alignment"].cpu(), hb[10]) and torch.equal(db["f0"].cpu(), hb[8])
        assert torch.equal(db["mel_mask"].cpu(), (hb[3] != 0).unsqueeze(-2))
        out = tr.train_step(db["text"], db["src_mask"], db["mel_mask"], db["alignment"], db["f0"], db["energy"],
                            db["mel"])
        assert torch.isfinite(out["loss"]).item()
        seen += 1
    assert seen == 2
