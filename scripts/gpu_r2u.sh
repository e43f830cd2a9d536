#!/bin/bash
# round 2, visit U: the shipped defaults under the driver's commands + the hash-stamped launch list
set -u
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference > gpurun_out/r2u_bench_ref.json 2> gpurun_out/r2u_bench_ref.err; echo "bench reference exit=$?"
timeout 900 python bench.py > gpurun_out/r2u_bench.json 2> gpurun_out/r2u_bench.err; echo "bench exit=$?"
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2u_bench.json") if l.startswith("{")][-1])
c = d["corpus"]
print("bench ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "gemm", round(d["roofline"]["frac"], 4), "attn", round(d["roofline_attention"]["frac"], 4), "corpus", round(c["audio_s_per_s"]), c["bit_identical"], "clk", d["clocks"]["sm_mhz"])
PY
