"""Multi-rank inference sharding on CPU: world size 2 over gloo (the N>1 host logic of bench.py / SURVEY §8e).
The per-rank "model" is the oracle (the product needs a GPU); what is tested is the sharding, batching, padding
and gather logic: sharded results must equal the single-process results utterance by utterance."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from transformer_tts_b200 import sharding, synth


def test_shard_by_length_balances_and_partitions():
    g = torch.Generator().manual_seed(0)
    lens = torch.randint(20, 200, (37,), generator=g).tolist()
    for world in (1, 2, 3, 8):
        shards = sharding.shard_by_length(lens, world)
        flat = sorted(i for s in shards for i in s)
        assert flat == list(range(len(lens)))  # a partition
        sums = [sum(lens[i] for i in s) for s in shards]
        assert max(sums) - min(sums) <= max(lens)  # snake dealing keeps the load within one utterance
    assert sharding.shard_by_length([], 4) == [[], [], [], []]
    with pytest.raises(ValueError):
        sharding.shard_by_length(lens, 0)


def test_pad_text_batch_layout():
    ids, pos = sharding.pad_text_batch([torch.tensor([5, 6, 7]), torch.tensor([9])])
    assert ids.tolist() == [[5, 6, 7], [9, 0, 0]] and pos.tolist() == [[1, 2, 3], [1, 0, 0]]


def _texts():
    g = torch.Generator().manual_seed(5)
    return [torch.randint(1, 152, (int(n),), generator=g) for n in (9, 14, 6, 11, 7)]


def _small_model():
    hp = synth.make_hp(n_layer_encoder=1, n_layer_decoder=1)
    return hp, synth.synth_state_dict(hp, frames_per_phoneme=2.0)


def _infer_factory():
    from oracle import fastspeech2_oracle as O

    hp, sd = _small_model()

    def infer(text, src_mask):
        # the reference's batched free-running call is broken: one utterance at a time, like its script
        outs = []
        with torch.no_grad():
            for b in range(text.shape[0]):
                n = int(src_mask[b].sum())
                o = O.fastspeech2_forward(sd, vars(hp), text[b:b + 1, :n], src_mask[b:b + 1, :, :n])
                outs.append(o["outputs_postnet"][0])
        return outs

    return infer


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        merged = sharding.run_sharded(_texts(), batch_size=2, infer=_infer_factory())
        local = sharding.run_sharded(_texts(), batch_size=2, infer=_infer_factory(), gather=False)
        q.put((rank, {k: v.numpy() for k, v in merged.items()}, sorted(local)))  # numpy: no shared-memory handles
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_inference_matches_single_process():
    torch.set_num_threads(2)  # same thread count as the workers: same summation order
    single = sharding.run_sharded(_texts(), batch_size=2, infer=_infer_factory(), rank=0, world_size=1)
    assert sorted(single) == [0, 1, 2, 3, 4]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    owned = []
    for rank, merged, local_ids in got:
        assert sorted(merged) == [0, 1, 2, 3, 4]
        for k in single:
            assert merged[k].shape == single[k].shape, k  # same durations -> same frame count
            assert torch.allclose(torch.from_numpy(merged[k]), single[k], atol=1e-5, rtol=1e-5), k
        owned += local_ids
    assert sorted(owned) == [0, 1, 2, 3, 4]  # the two shards partition the utterances, no overlap


# --------------------------------------------------------------------------------------------------
# data-parallel TRAINING host logic (transformer_tts_b200/training.py) under gloo, world size 2
# --------------------------------------------------------------------------------------------------
def _train_dp_worker(rank, world, port, q):
    import os

    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from transformer_tts_b200.training import (allreduce_mean_scale, allreduce_tail_begin, allreduce_tail_finish,
                                               noam_lr, sync_bn_moments)

    g = torch.Generator().manual_seed(5)
    x_all = torch.randn(2 * 37, 8, generator=g) * 3 + 1  # rows of both ranks
    grads_all = torch.randn(2, 1000, generator=g)
    # gradient exchange: SUM all-reduce, mean folded into the returned scale
    flat = grads_all[rank].clone()
    scale = allreduce_mean_scale(flat, world)
    mean_grad = flat * scale
    # the overlapped exchange (tail first, asynchronously; head + join later) gives the same mean
    flat2 = grads_all[rank].clone()
    work = allreduce_tail_begin(flat2, 600)
    scale2 = allreduce_tail_finish(flat2, 600, work, world)
    ok_split = torch.allclose(flat2 * scale2, grads_all.mean(0), atol=1e-6)
    # SyncBatchNorm moments from local [sum | sum^2]
    x = x_all[rank * 37:(rank + 1) * 37]
    stats = torch.stack([x.sum(0), (x * x).sum(0)])
    ms, var, count = sync_bn_moments(stats, x.shape[0], world, True, 1e-5)
    ok = torch.allclose(mean_grad, grads_all.mean(0), atol=1e-6) and ok_split
    ok &= count == 74.0 and torch.allclose(ms[0], x_all.mean(0), atol=1e-5)
    ok &= torch.allclose(var, x_all.var(0, unbiased=False), rtol=1e-4, atol=1e-5)
    ok &= torch.allclose(ms[1], torch.rsqrt(x_all.var(0, unbiased=False) + 1e-5), rtol=1e-4)
    ok &= abs(noam_lr(1, 384, 1.0, 4000) - 4000 ** -1.5 * 384 ** -0.5) < 1e-15
    ok &= abs(noam_lr(10000, 384, 1.0, 4000) - 10000 ** -0.5 * 384 ** -0.5) < 1e-15
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_training_data_parallel_host_logic_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29400 + (os.getpid() % 500)
    procs = [ctx.Process(target=_train_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]
