#!/bin/bash
# r1d visit: parity tests, smoke, bench, launch list of one step, full captures of mel + attention kernels.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mel_frames -s 3 -c 1 -f -o gpurun_out/prof_mel $CMD > gpurun_out/ncu_full_mel.log 2>&1
echo "ncu mel exit=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 60 -c 1 -f -o gpurun_out/prof_attn $CMD > gpurun_out/ncu_full_attn.log 2>&1
echo "ncu attn exit=$?"
tail -3 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/smoke.log; cat gpurun_out/bench.json
