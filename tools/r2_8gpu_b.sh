#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gpu_scale_rs.py 8 > gpurun_out/scale_rs.log 2>&1; echo "scale rc=$?"; tail -28 gpurun_out/scale_rs.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/ddp_check.py > gpurun_out/ddp_check.log 2>&1; echo "ddp rc=$?"; grep rank gpurun_out/ddp_check.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 --no-sustained > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err; echo "bench8 rc=$?"
python tools/show_bench.py gpurun_out/bench_8gpu.json | grep -E "^value|train_step" | cut -c1-420
timeout 600 python bench.py --no-cpu --no-gpu-reference --no-sustained > gpurun_out/bench_1gpu_samebox.json 2> gpurun_out/bench_1gpu.err; echo "bench1 rc=$?"
python tools/show_bench.py gpurun_out/bench_1gpu_samebox.json | grep -E "^value|train_step" | cut -c1-420
