#!/bin/bash
# round 2, visit I: the allocation-free mel entry point alone in a fresh process, then the full GPU suite (it runs last),
# then the hash-stamped launch list of the shipped sources and the driver's bench command
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -rf > gpurun_out/r2i_noalloc_alone.log 2>&1; echo "no-alloc alone exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/r2i_noalloc_alone.log | head -12; tail -2 gpurun_out/r2i_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu -rf > gpurun_out/r2i_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2i_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2i_pytest.log | head -24; tail -3 gpurun_out/r2i_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2i_smoke.log 2>&1; echo "smoke exit=$?"; tail -4 gpurun_out/r2i_smoke.log
timeout 900 python bench.py > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; echo "bench exit=$?"
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2i_bench.json") if l.startswith("{")][-1])
print("bench ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "roof", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"], "clk", d["clocks"])
PY
