#!/bin/bash
# round 2, visit R: final sources under the driver's commands (smoke, bench, reference arm) + A/B of the two defaults the
# cheaper epilogue could have moved (8 vs 12 epilogue warps, fold mask 7 vs 5) + launch list with mask 7
set -u
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2r_smoke.log 2>&1; echo "smoke exit=$?"; tail -4 gpurun_out/r2r_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2r_bench_ref.json 2> gpurun_out/r2r_bench_ref.err; echo "bench reference exit=$?"
timeout 900 python bench.py > gpurun_out/r2r_bench.json 2> gpurun_out/r2r_bench.err; echo "bench exit=$?"
for rep in 1 2; do
for V in "Q3A_X=1" "Q3A_EPI_WARPS=8" "Q3A_LN_FOLD_MASK=7"; do
  env $V timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2r_bench_ab_${V%%=*}_${V##*=}_$rep.json" 2>/dev/null; echo "bench $V exit=$?"
done
done
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
Q3A_LN_FOLD_MASK=7 timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
Q3A_LN_FOLD_MASK=7 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_mask7.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches (mask 7) exit=$?"
python - <<'PY'
import json, glob
for p in ["gpurun_out/r2r_bench.json"] + sorted(glob.glob("gpurun_out/r2r_bench_ab_*.json")):
    d = json.loads([l for l in open(p) if l.startswith("{")][-1])
    print(p.split("r2r_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "gemm", round(d["roofline"]["frac"], 4), round(d["roofline"]["ms_per_step"], 3), "traffic", d["roofline"].get("traffic") is not None, "clk", d["clocks"]["sm_mhz"])
PY
