# Round 2, final tree (after the Transformer-TTS training step, the wgrad bias fusion, the batched dgrad-operand kernel and the
# graphed training step): GPU tests, smoke, default bench line, reference arm, ncu launch lists (inference step, training step).
# One GPU.  ncu runs only after the same command exited 0 without it.  The --set full capture of the inference hot kernels is
# r02b's (those kernels are unchanged since).
set -u
O=gpurun_out/r02c; mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -2 $O/gpu_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 $O/smoke.log
S=$(date +%s); python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$? wall=$(( $(date +%s) - S ))s"
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
python bench.py --steps 2 --warmup 3 --step-only > $O/step_only.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --step-only > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
python bench.py --workload train --steps 2 --warmup 3 > $O/train.json 2> $O/train.err && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/train_launches.csv python bench.py --workload train --steps 1 --warmup 3 > $O/ncu_train.log 2>&1
echo "train launch list rc=$?"
python tools/exp/tts_train_profile.py > $O/tts_train_profile.txt 2>&1; echo "tts train profile rc=$?"
python tools/exp/fs2_train_profile.py > $O/fs2_train_profile.txt 2>&1; echo "fs2 train profile rc=$?"
ls -la $O | head -30
