#!/bin/bash
# Quick GPU visit while iterating on a kernel: the building-block tests first (bounded), then the full
# parity suite and one bench line.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_attention_gpu.py -x -q > gpurun_out/pytest_blocks.log 2>&1; echo "blocks exit=$?"
tail -5 gpurun_out/pytest_blocks.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
