#!/bin/bash
# round 2, visit H: final consistency -- full GPU suite, the other BASELINE.json configs + the aligner's device path
# (scripts/gpu_configs.py), hash-stamped launch list of the shipped sources
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu -rf > gpurun_out/r2h_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2h_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2h_pytest.log | head -24; tail -3 gpurun_out/r2h_pytest.log
timeout 900 python scripts/gpu_configs.py > gpurun_out/r2h_configs.txt 2>&1; echo "configs exit=$?"; tail -16 gpurun_out/r2h_configs.txt | cut -c1-260
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus > gpurun_out/r2h_bench_short.json 2>/dev/null; echo "bench short exit=$?"
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
