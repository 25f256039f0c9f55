#!/bin/bash
# Round-1 (second half) evidence run, one gpurun call: final bench line, reference arm, ncu launch lists of the bench
# command and of one eager training step, ncu --set full of the batched weight-gradient kernel and the fused optimizer
# tail.  Each ncu pass runs only after the same command exited 0 without it.
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_final2.json 2> gpurun_out/bench_final2.err || { echo "bench failed"; tail -5 gpurun_out/bench_final2.err; exit 1; }
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref2.json 2> gpurun_out/bench_ref2.err || echo "reference arm failed"
python bench.py --steps 2 --warmup 3 --no-cpu --no-train > /dev/null 2> gpurun_out/bench_short.err || { echo "short bench failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-train > gpurun_out/ncu_bench2.log 2>&1
echo "ncu bench launch list: $?"
python tools/gpu_train_once.py 2 > gpurun_out/train_once.log 2>&1 || { echo "train_once failed"; tail -5 gpurun_out/train_once.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train_final.csv \
    python tools/gpu_train_once.py 2 > gpurun_out/ncu_train.log 2>&1
echo "ncu train launch list: $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wgrad_batch_kernel -c 1 -f -o gpurun_out/wgrad_full \
    python tools/gpu_train_once.py 1 > gpurun_out/ncu_wgrad.log 2>&1
echo "ncu wgrad full: $?"
python tools/gpu_optim_once.py > gpurun_out/optim_once.log 2>&1 || { echo "optim_once failed"; tail -5 gpurun_out/optim_once.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"adam_ema_kernel|grad_stats_kernel" -s 2 -c 2 -f -o gpurun_out/optim_full \
    python tools/gpu_optim_once.py > gpurun_out/ncu_optim.log 2>&1
echo "ncu optim full: $?"
python tools/gpu_train_phases.py flat > gpurun_out/train_phases_flat.log 2>&1
python tools/gpu_train_phases.py fused > gpurun_out/train_phases_torchtail.log 2>&1
python tools/gpu_wgrad_probe.py > gpurun_out/wgrad_probe.log 2>&1
python tools/gpu_train_probe.py > gpurun_out/train_probe.log 2>&1
ls -la gpurun_out/*.ncu-rep
