#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-train > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/bench.json 2>/dev/null | cut -c1-330 | head -12
