#!/bin/bash
# visit J, second pass: pipelined CUDA-core path for few-row tiles -- parity + kernel A/B + step A/B
set -u
mkdir -p gpurun_out
rm -f gpurun_out/r2j2_attn_bench.log
timeout 900 python -m pytest tests/test_gemm_attention_gpu.py -q -m gpu -rf -k attention > gpurun_out/r2j2_attn_pytest.log 2>&1; echo "attention pytest exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/r2j2_attn_pytest.log | head -12; tail -2 gpurun_out/r2j2_attn_pytest.log
for V in "Q3A_X=1" "Q3A_ATTN_NO_REM_PATH=1"; do
  echo "== $V" >> gpurun_out/r2j2_attn_bench.log
  env $V ATTN_V=1 ATTN_L=390 ATTN_NS=64 timeout 200 python scripts/gpu_attn_bench.py 2>&1 | tail -1 >> gpurun_out/r2j2_attn_bench.log
  env $V ATTN_V=1 ATTN_L=1542 ATTN_NS=16 timeout 200 python scripts/gpu_attn_bench.py 2>&1 | tail -1 >> gpurun_out/r2j2_attn_bench.log
done
cut -c1-140 gpurun_out/r2j2_attn_bench.log
for V in "Q3A_X=1" "Q3A_ATTN_NO_REM_PATH=1"; do
  env $V timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2j2_bench_ab_${V%%=*}.json" 2>/dev/null; echo "bench $V exit=$?"
done
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2j2_bench_ab_*.json")):
    d = json.loads([l for l in open(p) if l.startswith("{")][-1])
    print(p.split("ab_")[1], "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "attn", round(d["roofline_attention"]["frac"], 4), round(d["roofline_attention"]["ms_per_step"], 3))
PY
