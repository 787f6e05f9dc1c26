mkdir -p gpurun_out/scale
for N in 1 2 4 8; do
  if [ $N -eq 1 ]; then
    python bench.py --gpus 1 --no-cpu-baseline > gpurun_out/scale/infer_$N.json 2> gpurun_out/scale/infer_$N.err
    python bench.py --gpus 1 --workload train --no-cpu-baseline > gpurun_out/scale/train_$N.json 2> gpurun_out/scale/train_$N.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N bench.py --gpus $N --no-cpu-baseline > gpurun_out/scale/infer_$N.json 2> gpurun_out/scale/infer_$N.err
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --workload train --no-cpu-baseline > gpurun_out/scale/train_$N.json 2> gpurun_out/scale/train_$N.err
  fi
done
for f in gpurun_out/scale/*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d['n_gpus'], d['metric'], round(d['value']), d['ms_per_step'], d.get('e2e',{}).get('value'))
except Exception as e:
    print('ERR', e)
PY
done
