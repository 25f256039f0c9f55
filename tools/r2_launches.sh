#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-gpu-reference --no-sustained > gpurun_out/b_short.json 2> gpurun_out/b_short.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-gpu-reference --no-sustained > gpurun_out/ncu_launches.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/launches_r2.csv
python tools/launch_agg.py gpurun_out/launches_r2.csv | head -40
