#!/bin/bash
# closing run of the round on one B200: GPU tests, the bench line, the ncu launch list of the bench command, one `--set full`
# capture of a whole forward (11 launches).  Every ncu pass only after its command has exited 0 without ncu.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu_r2_final.log; tail -3 gpurun_out/pytest_gpu_r2_final.log
timeout 900 python bench.py > gpurun_out/bench_r2_last.json 2> gpurun_out/bench_r2_last.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-gpu-reference --no-sustained > gpurun_out/b_short.json 2> gpurun_out/b_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_r2_last.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-gpu-reference --no-sustained > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python tools/launch_agg.py gpurun_out/launches_r2_last.csv > gpurun_out/launches_r2_last_by_kernel.txt; head -12 gpurun_out/launches_r2_last_by_kernel.txt
timeout 300 python tools/gpu_fwd_once.py 3 > gpurun_out/fwd_once.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -s 20 -c 10 -o gpurun_out/ncu_fwd_r2_last -f python tools/gpu_fwd_once.py 3 > gpurun_out/ncu_fwd.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
