#!/bin/bash
# Round-end visit: parity tests, smoke, bench (both arms), launch list with DRAM bytes of one step, full captures.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/prof_*.ncu-rep gpurun_out/launches*.csv
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref exit=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || exit 1
# device-resident steps come first: 3 warm-up + 1 timed step of 135 launches each -> skip 405, take 135
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none \
  -s 405 -c 135 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
# gemm_tc_kernel launches of a step, in order: conv2, conv3, conv_out, then per layer QKV, out, up, down (77 per step);
# after 3 warm-up steps: conv2 = #231, layer-0 up-projection = #236, layer-0 down-projection = #237
for spec in "mel_frames:prof_mel:3:1" "attention_kernel:prof_attn:60:1" "gemm_tc_kernel:prof_conv2:231:1" "gemm_tc_kernel:prof_gemm_up:236:2" "conv1_tc:prof_conv1:3:1"; do
  IFS=: read pat name skip cnt <<< "$spec"
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$pat" -s $skip -c $cnt -f -o gpurun_out/$name $CMD > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name exit=$?"
done
tail -3 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/smoke.log; cat gpurun_out/bench.json
