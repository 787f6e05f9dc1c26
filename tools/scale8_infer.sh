mkdir -p gpurun_out/scale3
python bench.py --gpus 1 --steps 50 --no-cpu-baseline > gpurun_out/scale3/infer_1.json 2>/dev/null
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 50 --no-cpu-baseline > gpurun_out/scale3/infer_8.json 2>/dev/null
for f in gpurun_out/scale3/*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], d['n_gpus'], round(d['value']), round(d['ms_per_step'],3), d['config'].get('ms_per_step_by_rank'))
PY
done
