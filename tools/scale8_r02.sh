# Round 2: the default bench line (inference + train sub-record) at 1 and 8 GPUs of one box, the reference arm, and the
# 2-GPU data-parallel training check.  Run under `gpurun --gpus 8`.
mkdir -p gpurun_out/scale_r02
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tests/dist_train_check.py > gpurun_out/scale_r02/dist_train_check.json 2>gpurun_out/scale_r02/dist_train_check.err
echo "dist_train_check rc=$?"
python bench.py --gpus 1 --steps 50 --no-cpu-baseline > gpurun_out/scale_r02/bench_1.json 2>gpurun_out/scale_r02/bench_1.err
for n in 2 4 8; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2957$n bench.py --gpus $n --steps 50 --no-cpu-baseline > gpurun_out/scale_r02/bench_$n.json 2>gpurun_out/scale_r02/bench_$n.err
done
for f in gpurun_out/scale_r02/bench_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
t=d.get('train') or {}
print(sys.argv[1], d['n_gpus'], 'infer', round(d['value']), round(d['ms_per_step'],3), d.get('ms_per_step_by_rank'), 'train', round(t.get('value',0)), round(t.get('ms_per_step',0),3), 'allreduce', round(t.get('allreduce_ms',0),3))
PY
done
tail -2 gpurun_out/scale_r02/dist_train_check.json
