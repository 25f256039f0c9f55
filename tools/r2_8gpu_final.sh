#!/bin/bash
# one 8-GPU lease on the final tree: the GPU suite (0 skips), the bench line at N = 8 and at N = 1 on the same box, configs 3 / 4 from one host process
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests -m gpu -q -rs > gpurun_out/pytest_8gpu_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_8gpu_final.log; tail -3 gpurun_out/pytest_8gpu_final.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 --no-sustained --no-gpu-reference > gpurun_out/bench_8gpu_final.json 2> gpurun_out/bench_8gpu_final.err; echo "bench8 rc=$?"
python tools/show_bench.py gpurun_out/bench_8gpu_final.json | grep -E "^value|train_step" | cut -c1-420
timeout 600 python bench.py --no-cpu --no-gpu-reference --no-sustained > gpurun_out/bench_1gpu_samebox_final.json 2> gpurun_out/bench_1gpu_final.err; echo "bench1 rc=$?"
python tools/show_bench.py gpurun_out/bench_1gpu_samebox_final.json | grep -E "^value|train_step" | cut -c1-420
timeout 600 python tools/gpu_scale_rs.py 8 > gpurun_out/scale_rs_final.log 2>&1; echo "scale rc=$?"; tail -24 gpurun_out/scale_rs_final.log
