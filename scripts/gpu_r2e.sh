#!/bin/bash
# round 2, visit E (2 GPUs): the N>1 paths on real hardware -- bench.py under torchrun with NCCL (own arm and the reference
# arm, launched the way the driver launches them), the pool executor and the C++ pool harness on two devices
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | tee gpurun_out/r2e_gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 900 python -m pytest tests/test_pool_gpu.py tests/test_shim_cpp.py -q -m gpu -rf > gpurun_out/r2e_pool_pytest.log 2>&1; echo "pool/shim pytest exit=$?"; tail -3 gpurun_out/r2e_pool_pytest.log
timeout 900 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2e_bench_2gpu.json 2> gpurun_out/r2e_bench_2gpu.err; echo "bench 2 GPUs exit=$?"
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/r2e_bench_1gpu.json 2> gpurun_out/r2e_bench_1gpu.err; echo "bench 1 GPU exit=$?"
timeout 900 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2e_bench_ref_2gpu.json 2> gpurun_out/r2e_bench_ref_2gpu.err; echo "reference arm under torchrun exit=$?"
python - <<'PY'
import json
for p in ("r2e_bench_1gpu", "r2e_bench_2gpu", "r2e_bench_ref_2gpu"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/{p}.json") if l.startswith("{")][-1])
        c = d.get("corpus") or {}
        print(p, "n", d.get("n_gpus"), "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]),
              "corpus", {k: c.get(k) for k in ("value", "bit_identical", "n_ranks", "imbalance")}, "flags", d.get("ranks_share_gpus"))
    except Exception as ex:
        print(p, "unreadable", ex)
PY
tail -5 gpurun_out/r2e_bench_2gpu.err
