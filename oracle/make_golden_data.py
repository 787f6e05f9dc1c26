"""ORACLE support — test infrastructure, NOT the product.

Golden vectors for the data path: runs the UNMODIFIED reference `collate_fn` and
`LengthsBatchSampler._batch_indices` (datasets/datasets_fastspeech2.py:521-616, :781-798) on seeded synthetic items
and commits the outputs as tests/golden/fs2_collate.npz.  The module imports `librosa` at the top (unused by these
two functions; not installed here), which is stubbed.  Run in the build container only:

    python oracle/make_golden_data.py
"""
import importlib.util
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("TTSB_REFERENCE_ROOT", "/root/reference")


def load_reference_datasets():
    for name in ("librosa",):
        if name not in sys.modules:
            try:
                __import__(name)
            except ImportError:
                sys.modules[name] = types.ModuleType(name)
    sys.path.insert(0, REF)
    spec = importlib.util.spec_from_file_location("ref_datasets_fastspeech2",
                                                  os.path.join(REF, "datasets", "datasets_fastspeech2.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def synth_items(n: int, seed: int, reduction_rate: int = 1, multi: str = ""):
    """Item dicts shaped like TrainDatasets.__getitem__'s (datasets_fastspeech2.py:300-420)."""
    rng = np.random.RandomState(seed)
    items = []
    for i in range(n):
        L = int(rng.randint(5, 24))
        dur = rng.randint(0, 9, size=L)
        dur[0] = max(dur[0], 1)
        T = int(dur.sum())
        d = dict(text=rng.randint(1, 152, size=L).astype(np.int64), mel_input=rng.randn(T, 80).astype(np.float32),
                 text_length=L, mel_length=T, pos_mel=np.arange(1, T + 1), pos_text=np.arange(1, L + 1),
                 stop_token=np.concatenate([np.zeros(T - 1), np.ones(1)]).astype(np.float32),
                 spk_emb=None, f0=rng.rand(T) * 700 + 71, energy=rng.rand(T) * 400, alignment=dur.astype(np.int64),
                 accent=None, gender=None, mel_name=f"utt{seed}_{i}.npy", spk_emb_postprocess=None, hop_size=None,
                 reduction_rate=reduction_rate, spk_emb_type=None, is_multi_speaker=False,
                 spk_emb_postprocess_type=None)
        if multi == "x_vector":
            d.update(is_multi_speaker=True, spk_emb_type="x_vector", spk_emb=rng.randn(16).astype(np.float32),
                     accent=rng.randint(0, 3, size=L), gender=int(rng.randint(0, 2)), hop_size=int(256),
                     spk_emb_postprocess_type="speaker_id", spk_emb_postprocess=int(rng.randint(0, 10)))
        items.append(d)
    return items


CASES = [("single_r1", dict(n=5, seed=3)), ("single_r2", dict(n=4, seed=4, reduction_rate=2)),
         ("xvec_r3", dict(n=3, seed=5, reduction_rate=3, multi="x_vector"))]
FIELDS = ["text", "mel_input", "pos_text", "pos_mel", "text_length", "mel_length", "stop_token", "spk_emb", "f0",
          "energy", "alignment", "accent", "gender", "spk_emb_postprocess", "mel_name", "hop_size"]


def main():
    D = load_reference_datasets()
    out = {}
    for name, kw in CASES:
        tup = D.collate_fn(synth_items(**kw))
        assert len(tup) == 16
        is_none = lambda v: v is None or (isinstance(v, list) and all(x is None for x in v))  # noqa: E731
        for f, v in zip(FIELDS, tup):
            if is_none(v) or f == "mel_name":
                continue
            out[f"{name}.{f}"] = v.numpy() if hasattr(v, "numpy") else np.asarray(v)
        out[f"{name}.none_fields"] = np.array([f for f, v in zip(FIELDS, tup) if is_none(v)], dtype="U32")
    # frame-budget batching: the sampler's _batch_indices on fixed length lists (bypassing __init__, which needs files)
    rng = np.random.RandomState(9)
    for k, (n, budget) in enumerate([(40, 3000), (17, 1500), (2, 5000), (9, 900)]):
        lengths = rng.randint(100, 900, size=n)
        s = D.LengthsBatchSampler.__new__(D.LengthsBatchSampler)
        s.lengths_np, s.n_lengths = lengths, budget
        batches = s._batch_indices()
        out[f"sampler{k}.lengths"] = lengths
        out[f"sampler{k}.budget"] = np.array(budget)
        out[f"sampler{k}.batch_of"] = np.array([b for b, idx in enumerate(batches) for _ in idx])
        out[f"sampler{k}.index"] = np.array([i for idx in batches for i in idx])
        out[f"sampler{k}.n_batches"] = np.array(len(batches))
    path = os.path.join(ROOT, "tests", "golden", "fs2_collate.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) // 1024} KiB")


if __name__ == "__main__":
    main()
