#!/bin/bash
# round 2, visit S: four instead of eight epilogue warps in the residual GEMMs with a 16-bit copy (five ring stages instead
# of four): parity with it on, step A/B, launch list
set -u
mkdir -p gpurun_out
Q3A_DUAL_EPI_WARPS=4 timeout 900 python -m pytest tests/test_encoder_gpu.py tests/test_decoder_gpu.py -q -m gpu -rf -x > gpurun_out/r2s_pytest_part.log 2>&1; echo "encoder + decoder pytest (4 warps) exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/r2s_pytest_part.log | head -12; tail -2 gpurun_out/r2s_pytest_part.log
for rep in 1 2; do
for V in "Q3A_X=1" "Q3A_DUAL_EPI_WARPS=4"; do
  env $V timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2s_bench_ab_${V%%=*}_$rep.json" 2>/dev/null; echo "bench $V exit=$?"
done
done
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
Q3A_DUAL_EPI_WARPS=4 timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
Q3A_DUAL_EPI_WARPS=4 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_dual4.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2s_bench_ab_*.json")):
    d = json.loads([l for l in open(p) if l.startswith("{")][-1])
    print(p.split("ab_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "gemm", round(d["roofline"]["frac"], 4), round(d["roofline"]["ms_per_step"], 3), "clk", d["clocks"]["sm_mhz"])
PY
