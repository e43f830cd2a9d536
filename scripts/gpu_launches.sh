#!/bin/bash
# ncu launch list of one bench step (after a plain run of the same command has exited 0)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 405 -c 135 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
