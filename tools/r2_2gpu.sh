#!/bin/bash
# two-GPU run of the final tree: the whole GPU suite (0 skips: the two-device cases run), then the bench under torchrun
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/pytest_2gpu_r2_final.log; tail -3 gpurun_out/pytest_2gpu_r2_final.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu --no-gpu-reference --no-sustained > gpurun_out/bench_r2_2gpu_ingest.json 2> gpurun_out/bench_r2_2gpu_ingest.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_r2_2gpu_ingest.json
