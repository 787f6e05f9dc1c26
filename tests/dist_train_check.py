"""Data-parallel training check, run under torchrun on >= 2 GPUs of one box (NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
        tests/dist_train_check.py

Each rank trains on its shard of one global batch (same padded lengths on every rank) with the gradient all-reduce
and SyncBatchNorm statistics of transformer_tts_b200.training.Trainer; rank 0 also runs the whole batch in a single
process.  With equal shard shapes the averaged per-rank L1 means equal the global means and SyncBatchNorm statistics
equal the global-batch statistics, so the two gradient vectors (and the parameters after one optimizer step) must
agree up to fp32 atomic-order noise.  This is what DDP(SyncBatchNorm(model)) gives the reference
(train_fastspeech2.py:421).  Prints one JSON line; exit code 0 iff the check passes.
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
    dist.init_process_group("nccl")
    from transformer_tts_b200 import build_fastspeech2, lib, synth
    from transformer_tts_b200.training import Trainer

    lib.load()
    hp = synth.make_hp()
    sd = synth.synth_state_dict(hp)
    per = 2
    B = per * world
    text, pos, _ = synth.synth_text_batch(B, 48, ragged=True, seed=11)
    dur, pos_mel, f0, en = synth.synth_targets(pos, seed=11, max_dur=10)
    mel = synth.synth_mel(pos_mel, hp.mel_dim, seed=11)
    full = dict(text=text, src_mask=(pos != 0).unsqueeze(-2), mel_mask=(pos_mel != 0).unsqueeze(-2), alignment=dur,
                f0=f0, energy=en, mel=mel)

    def make(sync):
        m = build_fastspeech2(hp)
        m.load_state_dict(sd)
        return Trainer(m.cuda().train(), sync_bn=sync, overlap_allreduce=sync)  # the overlapped exchange is the one checked

    tr = make(True)
    tr.broadcast_parameters()
    shard = {k: v[rank * per:(rank + 1) * per].contiguous().cuda() for k, v in full.items()}
    out = tr.train_step(**shard)
    torch.cuda.synchronize()
    ok, info = True, {}
    if rank == 0:
        ref = make(False)
        ref.world = 1  # single-process reference on the whole batch: no collectives
        ref.sync_bn = False
        o2 = ref.train_step(**{k: v.cuda() for k, v in full.items()})
        torch.cuda.synchronize()
        g_dist = tr.flat_g.double() / world
        g_ref = ref.flat_g.double()
        e_g = float((g_dist - g_ref).norm() / g_ref.norm())
        e_p = float((tr.flat_p.double() - ref.flat_p.double()).abs().max())
        bn = tr.model.postnet.pre_batchnorm.running_var
        e_bn = float((bn - ref.model.postnet.pre_batchnorm.running_var).abs().max())
        info = dict(world=world, grad_rel_l2=e_g, param_max_abs_diff=e_p, bn_running_var_max_abs_diff=e_bn,
                    loss_full=float(o2["loss"]), loss_rank0_shard=float(out["loss"]))
        ok = e_g < 2e-3 and e_p < 1e-5 and e_bn < 1e-4
        info["ok"] = ok
        print(json.dumps(info), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
