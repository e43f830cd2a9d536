#!/bin/bash
# round 2, visit F: final state -- full GPU suite, the driver's own bench commands, MMA probe incl. M = 64, launch list
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu -rf > gpurun_out/r2f_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2f_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2f_pytest.log | head -24; tail -3 gpurun_out/r2f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f_smoke.log 2>&1; echo "smoke exit=$?"; tail -4 gpurun_out/r2f_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2f_bench_ref.json 2> gpurun_out/r2f_bench_ref.err; echo "bench reference exit=$?"
timeout 900 python bench.py > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench exit=$?"
for rep in 1 2; do
  timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus > "gpurun_out/r2f_bench_short_$rep.json" 2>/dev/null; echo "bench short exit=$?"
done
timeout 300 python scripts/gpu_mma_probe.py > gpurun_out/r2f_mma_probe.log 2>&1; echo "probe exit=$?"; grep "M=64" gpurun_out/r2f_mma_probe.log | head -8
for KT in 0 64 128; do   # 0 = the per-segment choice
  for L in 390 780 1024 1196 2048; do
    echo "== Q3A_ATTN_KT=$KT L=$L" >> gpurun_out/r2f_attn_bench.log
    Q3A_ATTN_KT=$KT ATTN_V=1 ATTN_L=$L ATTN_NS=$((24960 / L)) timeout 200 python scripts/gpu_attn_bench.py 2>&1 | tail -2 >> gpurun_out/r2f_attn_bench.log
  done
done
echo "attn bench done"; cat gpurun_out/r2f_attn_bench.log | cut -c1-160
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2f_bench*.json")):
    try:
        d = json.loads([l for l in open(p) if l.startswith("{")][-1])
        e = d.get("e2e", {})
        print(p.split("r2f_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(e.get("value", 0)),
              "link", (e.get("host_link") or {}).get("min_over_ranks"), "roof", (d.get("roofline") or {}).get("frac"), "clk", (d.get("clocks") or {}).get("sm_mhz"))
    except Exception as ex:
        print(p, "unreadable", ex)
PY
