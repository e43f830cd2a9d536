#!/bin/bash
# round 2, visit D: full GPU suite with the 12-warp epilogues, A/B bench lines (x2), launch lists of both
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu -rf > gpurun_out/r2d_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2d_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2d_pytest.log | head -24; tail -3 gpurun_out/r2d_pytest.log
for rep in 1 2; do
for V in "Q3A_X=1" "Q3A_EPI_WARPS=8" "Q3A_LN_FOLD_MASK=7" "Q3A_NO_LN_FOLD=1"; do
  env $V timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus > "gpurun_out/r2d_bench_ab_${V%%=*}_${V##*=}_$rep.json" 2>/dev/null; echo "bench $V exit=$?"
done
done
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
Q3A_EPI_WARPS=8 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_epi8.csv $CMD > gpurun_out/ncu_launches_epi8.log 2>&1
echo "ncu launches (8 epilogue warps) exit=$?"
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2d_bench_ab_*.json")):
    try:
        d = json.loads([l for l in open(p) if l.startswith("{")][-1])
        print(p.split("ab_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "parts", round(d["e2e"]["with_intra_call_parts_value"]),
              "sync", round(d["e2e"]["synchronous_call_value"]), "gemm", round(d["roofline"]["frac"], 3), round(d["roofline"]["ms_per_step"], 2), "clk", d["clocks"]["sm_mhz"])
    except Exception as ex:
        print(p, "unreadable", ex)
PY
