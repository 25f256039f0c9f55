#!/bin/bash
# Round-1 evidence run (one gpurun call): default bench line, ncu launch list of the bench command, ncu --set full of one
# whole model.pt forward.  Each ncu pass runs only after the same command exited 0 without it.
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || { echo "bench failed"; tail -5 gpurun_out/bench_final.err; exit 1; }
cat gpurun_out/bench_final.json
python bench.py --steps 2 --warmup 3 --no-cpu --no-train > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err || { echo "short bench failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-train > gpurun_out/ncu_bench.log 2>&1
echo "ncu launch list: $?"
python tools/gpu_fwd_once.py 3 > gpurun_out/fwd_once.log 2>&1 || { echo "fwd_once failed"; tail -5 gpurun_out/fwd_once.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on -s 24 -c 12 -f -o gpurun_out/fwd_full \
    python tools/gpu_fwd_once.py 3 > gpurun_out/ncu_fwd.log 2>&1
echo "ncu full: $?"
ls -la gpurun_out/fwd_full.ncu-rep
