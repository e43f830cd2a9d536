#!/bin/bash
# round 2, visit Q: GEMM epilogue -- accumulator buffer handed back right after a warp's last tcgen05.ld with a CTA-scope
# arrive (was: end of the tile, cluster-scope release = MEMBAR.ALL.GPU), per-column constants through lanes + shuffle
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu -rf > gpurun_out/r2q_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2q_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2q_pytest.log | head -24; tail -3 gpurun_out/r2q_pytest.log
for rep in 1 2 3; do
  timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2q_bench_$rep.json" 2>/dev/null; echo "bench exit=$?"
done
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2q_bench_*.json")):
    d = json.loads([l for l in open(p) if l.startswith("{")][-1])
    print(p, "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "gemm", round(d["roofline"]["frac"], 4), round(d["roofline"]["ms_per_step"], 3), "clk", d["clocks"]["sm_mhz"])
PY
