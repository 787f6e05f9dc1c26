"""Utterance sharding for multi-GPU inference: utterances are independent, so ranks share no
data-path collective.  Utterances are dealt to ranks by descending length in a snake order so that the
frame counts (which grow with the phoneme counts) balance; each rank batches its own shard; results are
gathered on the host by utterance id.  Works with any ``torch.distributed`` backend (NCCL on the GPU box,
gloo in the CPU tests)."""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence

import torch
import torch.distributed as dist


def shard_by_length(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """Returns, per rank, the utterance ids it owns.  Longest first, dealt 0,1,..,W-1,W-1,..,1,0,0,1,...
    Deterministic: every rank computes the same assignment without communicating."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for pos, utt in enumerate(order):
        lap, k = divmod(pos, world_size)
        shards[k if lap % 2 == 0 else world_size - 1 - k].append(utt)
    return shards


def batches_of(ids: Sequence[int], batch_size: int) -> List[List[int]]:
    return [list(ids[i:i + batch_size]) for i in range(0, len(ids), batch_size)]


def pad_text_batch(texts: Sequence[torch.Tensor]):
    """(B, Lmax) int64 ids padded with 0 and the matching positions 1..len (0 on padding) — the batch layout
    of test_fastspeech2.py:146-165 / datasets collate."""
    L = max(int(t.numel()) for t in texts)
    ids = torch.zeros((len(texts), L), dtype=torch.int64)
    pos = torch.zeros((len(texts), L), dtype=torch.int64)
    for i, t in enumerate(texts):
        n = int(t.numel())
        ids[i, :n] = t
        pos[i, :n] = torch.arange(1, n + 1)
    return ids, pos


def run_sharded(texts: Sequence[torch.Tensor], batch_size: int,
                infer: Callable[[torch.Tensor, torch.Tensor], Sequence[torch.Tensor]],
                rank: int | None = None, world_size: int | None = None, gather: bool = True) -> Dict[int, torch.Tensor]:
    """Runs ``infer(ids, src_mask) -> per-utterance tensors`` on this rank's shard and (optionally) gathers
    ``{utterance id: result}`` from all ranks on every rank through ``all_gather_object`` (host side)."""
    if world_size is None:
        world_size = dist.get_world_size() if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    mine = shard_by_length([int(t.numel()) for t in texts], world_size)[rank]
    local: Dict[int, torch.Tensor] = {}
    for ids in batches_of(mine, batch_size):
        text, pos = pad_text_batch([texts[i] for i in ids])
        outs = infer(text, (pos != 0).unsqueeze(-2))
        for i, o in zip(ids, outs):
            local[i] = o.detach().cpu()
    if not gather or world_size == 1:
        return local
    parts: List[Dict[int, torch.Tensor]] = [None] * world_size  # type: ignore[list-item]
    dist.all_gather_object(parts, local)
    merged: Dict[int, torch.Tensor] = {}
    for p in parts:
        merged.update(p)
    return merged
