#!/bin/bash
# round 2, visit L: final sources (packed fp32 in the attention softmax; corpus leg in pipelined sub-batches) -- full GPU
# suite, smoke, the driver's bench commands, hash-stamped launch list
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu -rf > gpurun_out/r2l_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2l_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2l_pytest.log | head -24; tail -3 gpurun_out/r2l_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2l_smoke.log 2>&1; echo "smoke exit=$?"; tail -4 gpurun_out/r2l_smoke.log
timeout 900 python bench.py > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; echo "bench exit=$?"; tail -3 gpurun_out/r2l_bench.err
timeout 900 python bench.py --impl reference > gpurun_out/r2l_bench_ref.json 2> gpurun_out/r2l_bench_ref.err; echo "bench reference exit=$?"
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2l_bench.json") if l.startswith("{")][-1])
c = d["corpus"]
print("bench ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "roof", round(d["roofline"]["frac"], 4), "attn", round(d["roofline_attention"]["frac"], 4))
print("corpus", round(c["audio_s_per_s"]), "one call", round(c["one_synchronous_call_per_rank_audio_s_per_s"]), "subs", c["sub_batches_this_rank"], "same bytes", c["sub_batches_equal_one_call_bitwise"], "identical", c["bit_identical"])
PY
