# Round 2, final tree: GPU tests, smoke, default bench line, reference arm, ncu launch lists (inference step, AR decode) and
# the --set full capture of the hot kernels (traffic).  One GPU.  ncu runs only after the same command exited 0 without it.
set -u
O=gpurun_out/r02b; mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -2 $O/gpu_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
S=$(date +%s); python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$? wall=$(( $(date +%s) - S ))s"
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
python bench.py --steps 2 --warmup 3 --step-only > $O/step_only.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --step-only > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
python tools/exp/tts_decode_ncu.py 16 600 > $O/tts_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/tts_decode_launches.csv python tools/exp/tts_decode_ncu.py 16 12 > $O/ncu_tts.log 2>&1
echo "tts launch list rc=$?"
python tools/run_hot_kernels.py > $O/hot_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -c 34 -o $O/hot_kernels_full python tools/run_hot_kernels.py > $O/ncu_hot.log 2>&1
echo "hot kernels rc=$?"
# the report itself is ~90 MB (gpurun_out/ is capped at 64 MiB): keep the per-kernel summary, drop the report
python tools/ncu_key_metrics.py $O/hot_kernels_full.ncu-rep > $O/hot_kernels_full_summary.csv && rm -f $O/hot_kernels_full.ncu-rep
ls -la $O | head -30
