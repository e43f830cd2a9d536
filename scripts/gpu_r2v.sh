#!/bin/bash
# round 2, visit V (8 GPUs): the shipped defaults at N = 8 and N = 1 on one box (corpus leg included)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571"
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2v_bench_8gpu.json 2> gpurun_out/r2v_bench_8gpu.err; echo "bench 8 GPUs exit=$?"
timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2v_bench_1gpu.json 2> gpurun_out/r2v_bench_1gpu.err; echo "bench 1 GPU exit=$?"
python - <<'PY'
import json
for n in (1, 8):
    d = json.loads([l for l in open(f"gpurun_out/r2v_bench_{n}gpu.json") if l.startswith("{")][-1])
    e, c = d["e2e"], d["corpus"]
    print(n, "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(e["value"]), "sync", round(e["synchronous_call_value"]), "corpus", round(c["audio_s_per_s"]), round(c["one_synchronous_call_per_rank_audio_s_per_s"]), c["bit_identical"], "cpu", (e.get("host_cpu") or {}).get("pct_of_all_cores"))
PY
