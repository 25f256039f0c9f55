#!/bin/bash
# round 2: the whole GPU suite, then the bench line (one gpurun call)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "EResNet\(|1080p|4096|passed|failed|FAILED|rc=" gpurun_out/pytest_gpu.log | cut -c1-400 | head -30
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/bench.json 2>/dev/null | head -40 || tail -c 1500 gpurun_out/bench.json
