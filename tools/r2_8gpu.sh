#!/bin/bash
# one 8-GPU lease: the whole GPU suite (0 skips), configs 3 / 4 from one host process at G = 1, 2, 4, 8, and the bench line at N = 8
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests -m gpu -q -rs > gpurun_out/pytest_8gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_8gpu.log; tail -4 gpurun_out/pytest_8gpu.log
timeout 600 python tools/gpu_scale_rs.py 8 > gpurun_out/scale_rs.log 2>&1; echo "scale rc=$?"; cat gpurun_out/scale_rs.log | tail -22
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err; echo "bench8 rc=$?"
python tools/show_bench.py gpurun_out/bench_8gpu.json | cut -c1-700 | head -14
