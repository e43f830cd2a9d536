#!/bin/bash
# round 2, visit T: four epilogue warps in the dual-output residual GEMMs as the default -- full suite, smoke, bench,
# A/B with the plain residual GEMM at four warps too, launch list of that combination
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu -rf > gpurun_out/r2t_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2t_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2t_pytest.log | head -24; tail -3 gpurun_out/r2t_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2t_smoke.log 2>&1; echo "smoke exit=$?"; tail -4 gpurun_out/r2t_smoke.log
for rep in 1 2; do
for V in "Q3A_X=1" "Q3A_RES_EPI_WARPS=4" "Q3A_DUAL_EPI_WARPS=8"; do
  env $V timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2t_bench_ab_${V%%=*}_${V##*=}_$rep.json" 2>/dev/null; echo "bench $V exit=$?"
done
done
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
Q3A_RES_EPI_WARPS=4 timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
Q3A_RES_EPI_WARPS=4 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_res4.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2t_bench_ab_*.json")):
    d = json.loads([l for l in open(p) if l.startswith("{")][-1])
    print(p.split("ab_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "gemm", round(d["roofline"]["frac"], 4), round(d["roofline"]["ms_per_step"], 3), "clk", d["clocks"]["sm_mhz"])
PY
