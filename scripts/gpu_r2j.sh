#!/bin/bash
# round 2, visit J: CUDA-core path for query tiles with few live rows -- parity, kernel A/B, step A/B, full suite, launch list
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_attention_gpu.py -q -m gpu -rf -k attention > gpurun_out/r2j_attn_pytest.log 2>&1; echo "attention pytest exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/r2j_attn_pytest.log | head -12; tail -2 gpurun_out/r2j_attn_pytest.log
for V in "Q3A_X=1" "Q3A_ATTN_NO_REM_PATH=1"; do
  echo "== $V" >> gpurun_out/r2j_attn_bench.log
  env $V ATTN_V=1 ATTN_L=390 ATTN_NS=64 timeout 200 python scripts/gpu_attn_bench.py 2>&1 | tail -1 >> gpurun_out/r2j_attn_bench.log
  env $V ATTN_V=1 ATTN_L=1542 ATTN_NS=16 timeout 200 python scripts/gpu_attn_bench.py 2>&1 | tail -1 >> gpurun_out/r2j_attn_bench.log
done
cut -c1-140 gpurun_out/r2j_attn_bench.log
for rep in 1 2; do
for V in "Q3A_X=1" "Q3A_ATTN_NO_REM_PATH=1"; do
  env $V timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2j_bench_ab_${V%%=*}_$rep.json" 2>/dev/null; echo "bench $V exit=$?"
done
done
timeout 2400 python -m pytest tests -x -q -m gpu -rf > gpurun_out/r2j_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2j_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2j_pytest.log | head -24; tail -3 gpurun_out/r2j_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j_smoke.log 2>&1; echo "smoke exit=$?"; tail -4 gpurun_out/r2j_smoke.log
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2j_bench_ab_*.json")):
    try:
        d = json.loads([l for l in open(p) if l.startswith("{")][-1])
        print(p.split("ab_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]),
              "attn", round(d["roofline_attention"]["frac"], 4), round(d["roofline_attention"]["ms_per_step"], 3), "clk", d["clocks"]["sm_mhz"])
    except Exception as ex:
        print(p, "unreadable", ex)
PY
