"""Data-parallel check of the Transformer-TTS training step, run under torchrun on >= 2 GPUs of one box (NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 \
        tests/dist_tts_train_check.py

Each rank trains on its shard of one global batch (same padded lengths on every rank) with the gradient all-reduce of
transformer_tts_b200.tts_training.TTSTrainer (inherited from training.Trainer); rank 0 also runs the whole batch in a single
process.  The losses of train.py:211-217 are unmasked means, so with equal shard shapes the averaged per-rank gradients equal
the whole-batch gradients up to fp32 atomic-order noise — what DistributedDataParallel gives the reference.  Dropout 0 (a
rank-dependent mask stream has no single-process counterpart).  Prints one JSON line; exit code 0 iff the check passes.
"""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from oracle import transformer_tts_oracle as T
    from oracle.make_golden_tts_train import tts_train_batch
    from transformer_tts_b200 import build_transformer_tts, lib, synth
    from transformer_tts_b200.tts_training import TTSTrainer

    lib.load()
    hp = synth.make_tts_hp()
    r = hp.reduction_rate
    sd = synth.synth_tts_state_dict(hp)
    per = 2
    b = tts_train_batch(hp, per * world, 32, 160, seed=17)
    src_mask, _ = T.create_masks(b["pos"], b["pos_in"])
    full = (b["text"], b["mel_input"], src_mask, (b["pos_in"] != 0).to(torch.uint8), b["mel"][:, r:, :].contiguous(),
            b["stop"][:, r:].contiguous())

    def make():
        m = build_transformer_tts(hp)
        m.load_state_dict(sd)
        return TTSTrainer(m.cuda().train(), positive_weight=5.0)

    tr = make()
    tr.broadcast_parameters()
    out = tr.train_step(*[v[rank * per:(rank + 1) * per].contiguous().cuda() for v in full])
    torch.cuda.synchronize()
    ok, info = True, {}
    if rank == 0:
        ref = make()
        ref.world = 1  # single-process reference on the whole batch: no collectives
        o2 = ref.train_step(*[v.cuda() for v in full])
        torch.cuda.synchronize()
        g_dist, g_ref = tr.flat_g.double() / world, ref.flat_g.double()
        e_g = float((g_dist - g_ref).norm() / g_ref.norm())
        e_p = float((tr.flat_p.double() - ref.flat_p.double()).abs().max())
        info = dict(world=world, grad_rel_l2=e_g, param_max_abs_diff=e_p, loss_full=float(o2["loss"]),
                    loss_rank0_shard=float(out["loss"]))
        ok = e_g < 2e-3 and e_p < 1e-5
        info["ok"] = ok
        print(json.dumps(info), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
