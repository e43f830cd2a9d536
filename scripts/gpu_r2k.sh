#!/bin/bash
# round 2, visit K: packed fp32 (FFMA2 / FADD2 / FMUL2) in the attention softmax -- parity (bit-identical by construction),
# kernel time on the shapes of the earlier sweeps, step time
set -u
mkdir -p gpurun_out
rm -f gpurun_out/r2k_attn_bench.log
timeout 900 python -m pytest tests/test_gemm_attention_gpu.py tests/test_encoder_gpu.py -q -m gpu -rf -x > gpurun_out/r2k_pytest_part.log 2>&1; echo "attention + encoder pytest exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/r2k_pytest_part.log | head -12; tail -2 gpurun_out/r2k_pytest_part.log
for L in 390 1196 2048; do
  ATTN_V=1 ATTN_L=$L ATTN_NS=$((24960 / L)) timeout 200 python scripts/gpu_attn_bench.py 2>&1 | tail -1 >> gpurun_out/r2k_attn_bench.log
done
ATTN_V=1 timeout 300 python scripts/gpu_attn_bench.py 2>&1 | grep "attention v1" >> gpurun_out/r2k_attn_bench.log
cut -c1-140 gpurun_out/r2k_attn_bench.log
for rep in 1 2; do
  timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-corpus --no-link-probe > "gpurun_out/r2k_bench_$rep.json" 2>/dev/null; echo "bench exit=$?"
done
python - <<'PY'
import json, glob
for p in sorted(glob.glob("gpurun_out/r2k_bench_*.json")):
    d = json.loads([l for l in open(p) if l.startswith("{")][-1])
    print(p, "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "attn", round(d["roofline_attention"]["frac"], 4), round(d["roofline_attention"]["ms_per_step"], 3), "clk", d["clocks"]["sm_mhz"])
PY
