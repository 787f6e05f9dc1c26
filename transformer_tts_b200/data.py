"""Host-side data path either side of the model: batching and padding rules of the reference's
loader, and the pinned-memory host -> device prefetch the training loop feeds the kernels from.

Mirrors (names, argument meaning, return layout) of
  * ``collate_fn`` and its padding helpers      datasets/datasets_fastspeech2.py:521-616, :702-745
  * ``LengthsBatchSampler`` / ``NumBatchSampler``  datasets/datasets_fastspeech2.py:747-845
  * ``create_masks`` (FastSpeech2 branch)        train_fastspeech2.py:55-82
plus ``DevicePrefetcher`` (no reference counterpart: the reference copies every tensor synchronously with
``.to(DEVICE)`` inside the step, train_fastspeech2.py:139-165).

The file readers (``.npy`` / HTK mel files, SentencePiece text) stay the reference's: they are I/O, not arithmetic, and
produce the item dicts ``collate_fn`` takes.
"""
from __future__ import annotations

import os
import random
from collections.abc import Mapping
from typing import Dict, Iterable, Iterator, List, Optional, Sequence

import numpy as np
import torch

MEL_PAD = -0.5        # datasets_fastspeech2.py:728-733: the (normalised) mel is padded with -0.5
STOP_PAD = 1.0        # :741: the stop token of a padded frame is 1


def _round_up(x: int, multiple: int) -> int:
    """datasets_fastspeech2.py:723-726."""
    r = x % multiple
    return x if r == 0 else x + multiple - r


def _pad_rows(seqs: Sequence[np.ndarray], length: int, value, dtype=None) -> np.ndarray:
    """Right-pads 1-D (or (T, C)) arrays with `value` up to `length` along axis 0 and stacks them."""
    first = np.asarray(seqs[0])
    out = np.full((len(seqs), length) + first.shape[1:], value, dtype=dtype or first.dtype)
    for i, x in enumerate(seqs):
        x = np.asarray(x)
        out[i, :x.shape[0]] = x
    return out


def _prepare_data(inputs: Sequence[np.ndarray]) -> np.ndarray:
    """Zero-pad to the longest item (datasets_fastspeech2.py:702-708)."""
    return _pad_rows(inputs, max(len(x) for x in inputs), 0)


def _pad_mel(inputs: Sequence[np.ndarray], reduction_rate: int = 1, _pad: Optional[float] = None) -> np.ndarray:
    """(T_i, n_mel) items -> (B, round_up(max T, reduction_rate), n_mel), padded with -0.5 (:728-739)."""
    if not _pad:  # the reference's test: a pad of 0 (or None) means "use -0.5"
        _pad = MEL_PAD
    T = _round_up(max(x.shape[0] for x in inputs), reduction_rate)
    return _pad_rows(inputs, T, _pad)


def _pad_stop_token(inputs: Sequence[np.ndarray], reduction_rate: int = 1, _pad: float = STOP_PAD) -> np.ndarray:
    """:741-748."""
    T = _round_up(max(x.shape[0] for x in inputs), reduction_rate)
    return _pad_rows(inputs, T, _pad)


def collate_fn(batch: List[Mapping]):
    """Item dicts -> the reference's 16-tuple (15 for ``speaker_id`` embeddings), datasets_fastspeech2.py:521-616:
    ``text, mel_input, pos_text, pos_mel, text_length, mel_length, stop_token, spk_emb, f0, energy, alignment, accent,
    gender, spk_emb_postprocess, mel_name, hop_size``.  ids / positions / alignment int64 zero-padded, mel fp32 padded
    with -0.5 up to a multiple of ``reduction_rate``, f0 / energy fp32 zero-padded, stop token fp32 padded with 1."""
    if not isinstance(batch[0], Mapping):
        raise TypeError(f"batch must contain tensors, numbers, dicts or lists; found {type(batch[0])}")
    first = batch[0]
    col = lambda k: [d[k] for d in batch]  # noqa: E731
    reduction_rate = first["reduction_rate"]

    def padded(key, np_dtype, make):
        vals = col(key)
        if vals[0] is None:
            return None
        return make(_prepare_data(vals).astype(np_dtype))

    text = torch.LongTensor(_prepare_data(col("text")).astype(np.int32))
    mel_input = torch.FloatTensor(_pad_mel(col("mel_input"), reduction_rate))
    pos_mel = torch.LongTensor(_prepare_data(col("pos_mel")).astype(np.int32))
    pos_text = torch.LongTensor(_prepare_data(col("pos_text")).astype(np.int32))
    f0 = padded("f0", np.float64, torch.FloatTensor)
    energy = padded("energy", np.float64, torch.FloatTensor)
    alignment = padded("alignment", np.int32, torch.LongTensor)
    accent = padded("accent", np.int32, torch.LongTensor)
    gender = torch.LongTensor(col("gender")) if first["gender"] is not None else None
    hop_size = col("hop_size")
    if hop_size[0] is not None:
        hop_size = torch.LongTensor(hop_size)
    post_type = first["spk_emb_postprocess_type"]
    if post_type == "speaker_id":
        spk_post = torch.LongTensor(col("spk_emb_postprocess"))
    elif post_type == "x_vector":
        spk_post = torch.FloatTensor(np.asarray(col("spk_emb_postprocess")))
    else:
        spk_post = None
    stop_token = torch.FloatTensor(_pad_stop_token(col("stop_token"), reduction_rate, _pad=STOP_PAD))
    text_length = torch.LongTensor(col("text_length"))
    mel_length = torch.LongTensor(col("mel_length"))
    mel_name = col("mel_name")

    if first["is_multi_speaker"]:
        kind = first["spk_emb_type"]
        if kind == "x_vector":
            return (text, mel_input, pos_text, pos_mel, text_length, mel_length, stop_token,
                    torch.FloatTensor(np.asarray(col("spk_emb"))), f0, energy, alignment, accent, gender, spk_post,
                    mel_name, hop_size)
        if kind == "speaker_id":  # the reference returns 15 fields here (no hop_size)
            return (text, mel_input, pos_text, pos_mel, text_length, mel_length, stop_token,
                    torch.LongTensor(col("spk_emb")), f0, energy, alignment, accent, gender, spk_post, mel_name)
        raise AttributeError(f"{kind} is unknown in spk_emb_type")
    if f0 is None:  # the reference calls torch.FloatTensor(None) here and dies with a TypeError
        raise TypeError("collate_fn: single-speaker batches need f0 (the reference wraps it in FloatTensor again)")
    return (text, mel_input, pos_text, pos_mel, text_length, mel_length, stop_token, None, f0, energy, alignment,
            accent, gender, spk_post, mel_name, hop_size)


def frame_budget_batches(lengths: Sequence[int], n_lengths: int) -> List[List[int]]:
    """The batching rule of ``LengthsBatchSampler._batch_indices`` (datasets_fastspeech2.py:781-798): walk the
    utterances in file order and close a batch when (longest mel in it) x (items in it) would exceed ``n_lengths``
    padded frames.  Like the reference, the walk stops once fewer than two utterances remain un-batched at a batch
    boundary, so a trailing utterance that would open its own batch is dropped."""
    n = len(lengths)
    out: List[List[int]] = []
    i = 0
    while i + 1 < n:
        idx: List[int] = []
        longest = 0
        while i < n:
            cur = int(lengths[i])
            if max(longest, cur) * (len(idx) + 1) > n_lengths:
                break
            longest = max(longest, cur)
            idx.append(i)
            i += 1
        if not idx:  # the reference spins forever here
            raise ValueError(f"utterance {i} has {int(lengths[i])} frames, more than the budget n_lengths={n_lengths}")
        out.append(idx)
    return out


class LengthsBatchSampler(torch.utils.data.Sampler):
    """Variable-size batches under a padded-frame budget; same constructor as the reference (:754-779).  The lengths
    come from ``lengths_file`` (a ``.npy`` of mel lengths in dataset order) or, when it does not exist, from one pass
    over the dataset, saved to ``hp.lengths_file`` as the reference does."""

    def __init__(self, dataset, n_lengths, hp, lengths_file=None, shuffle=True, shuffle_one_time=False,
                 reverse=False):
        assert not ((shuffle == reverse) and shuffle is True), "shuffle and reverse cannot set True at the same time."
        if lengths_file is None or not os.path.exists(lengths_file):
            # the reference measures collate_fn([item]): the mel length rounded up to the reduction rate
            items = (dataset[i] for i in range(len(dataset)))
            self.lengths_np = np.array([_round_up(int(np.asarray(d["mel_input"]).shape[0]), d["reduction_rate"])
                                        for d in items])
            if getattr(hp, "lengths_file", None):
                np.save(hp.lengths_file, self.lengths_np)
        else:
            self.lengths_np = np.load(lengths_file)
            assert len(dataset) == len(self.lengths_np), \
                f"mismatch the number of lines between dataset and {lengths_file}"
        self.n_lengths = n_lengths
        self.all_indices = frame_budget_batches(self.lengths_np, n_lengths)
        if shuffle_one_time:
            random.shuffle(self.all_indices)
        self.shuffle, self.shuffle_one_time, self.reverse = shuffle, shuffle_one_time, reverse

    def __iter__(self):
        if self.shuffle and not self.shuffle_one_time:
            random.shuffle(self.all_indices)
        if self.reverse:
            self.all_indices.reverse()
        yield from self.all_indices

    def __len__(self):
        return len(self.all_indices)


class NumBatchSampler(torch.utils.data.Sampler):
    """Fixed-size batches in file order, remainder dropped (datasets_fastspeech2.py:815-845)."""

    def __init__(self, dataset, batch_size, drop_last=True, shuffle=True):
        self.batch_size, self.drop_last, self.dataset_len = batch_size, drop_last, len(dataset)
        n = self.dataset_len - self.dataset_len % batch_size
        self.all_indices = np.arange(n).reshape(-1, batch_size).tolist()
        self.shuffle = shuffle
        if shuffle:
            np.random.shuffle(self.all_indices)

    def __iter__(self):
        if self.shuffle:
            np.random.shuffle(self.all_indices)
        yield from self.all_indices

    def __len__(self):
        return len(self.all_indices)


def create_masks(src_pos: torch.Tensor, trg_pos: Optional[torch.Tensor], src_pad: int = 0, trg_pad: int = 0):
    """FastSpeech2 branch of train_fastspeech2.py:55-82: ``(pos != pad).unsqueeze(-2)`` for text and mel."""
    src_mask = (src_pos != src_pad).unsqueeze(-2)
    trg_mask = None if trg_pos is None else (trg_pos != trg_pad).unsqueeze(-2)
    return src_mask, trg_mask


# the fields of the collate tuple the FastSpeech2 step consumes (train_fastspeech2.py:139-165), by position
_STEP_FIELDS = dict(text=0, mel=1, pos_text=2, pos_mel=3, f0=8, energy=9, alignment=10)


class DevicePrefetcher:
    """Wraps a loader of ``collate_fn`` tuples: every batch is staged in pinned host memory and copied to the device on
    a side stream ONE BATCH AHEAD of the step that consumes it, so the H2D copy (mel is ~10 MB per batch at B=32)
    overlaps the previous step's kernels.  Yields dicts of device tensors ready for ``Trainer.train_step``:
    ``text, src_mask, mel_mask, alignment, f0, energy, mel`` (+ ``pos_text``, ``pos_mel``).  There is no CPU path."""

    def __init__(self, loader: Iterable, device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("DevicePrefetcher needs a CUDA device (there is no CPU path)")
        self.loader = loader
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.stream = torch.cuda.Stream(device=self.device)

    def _stage(self, batch) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        with torch.cuda.stream(self.stream):
            for name, k in _STEP_FIELDS.items():
                t = batch[k]
                if t is None:
                    raise ValueError(f"DevicePrefetcher: the batch has no {name} (field {k} of collate_fn's tuple)")
                out[name] = t.pin_memory().to(self.device, non_blocking=True)
            out["src_mask"], out["mel_mask"] = create_masks(out["pos_text"], out["pos_mel"])
        return out

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        it = iter(self.loader)
        try:
            nxt = self._stage(next(it))
        except StopIteration:
            return
        while nxt is not None:
            cur = nxt
            torch.cuda.current_stream(self.device).wait_stream(self.stream)  # cur's copies have landed
            for t in cur.values():
                t.record_stream(torch.cuda.current_stream(self.device))
            try:
                nxt = self._stage(next(it))  # the next batch's copies run under this step's kernels
            except StopIteration:
                nxt = None
            yield cur

    def __len__(self):
        return len(self.loader)
