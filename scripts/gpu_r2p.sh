#!/bin/bash
# round 2, visit P (8 GPUs): host CPU utilisation next to the link probe at N = 8 and N = 1 (what bounds e2e on this box)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561"
timeout 600 $TR bench.py --gpus 8 --steps 10 --warmup 3 --no-corpus > gpurun_out/r2p_bench_8gpu.json 2> gpurun_out/r2p_bench_8gpu.err; echo "bench 8 GPUs exit=$?"
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline --no-corpus > gpurun_out/r2p_bench_1gpu.json 2> gpurun_out/r2p_bench_1gpu.err; echo "bench 1 GPU exit=$?"
python - <<'PY'
import json
for n in (1, 8):
    d = json.loads([l for l in open(f"gpurun_out/r2p_bench_{n}gpu.json") if l.startswith("{")][-1])
    e = d["e2e"]
    print(n, "value", round(d["value"]), "e2e", round(e["value"]), "cpu", e.get("host_cpu"), "link", e["host_link"]["min_over_ranks"])
PY
