#!/bin/bash
# round 2, visit M (final sources) (8 GPUs of one box): the scaling lines the driver takes at round end, on one box back to back, with the
# host link probe in every line (what bounds e2e at N = 8), and the pool executor on all eight devices
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/r2m_gpus.txt; wc -l gpurun_out/r2m_gpus.txt
nproc > gpurun_out/r2m_host.txt; grep -c processor /proc/cpuinfo >> gpurun_out/r2m_host.txt; lscpu | grep -E "NUMA|Model name|Socket" >> gpurun_out/r2m_host.txt; free -g | head -2 >> gpurun_out/r2m_host.txt
nvidia-smi topo -m > gpurun_out/r2m_topo.txt 2>&1
for N in 8 4 2; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540 + N))"
  timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2m_bench_${N}gpu.json 2> gpurun_out/r2m_bench_${N}gpu.err; echo "bench $N GPUs exit=$?"
done
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2m_bench_1gpu.json 2> gpurun_out/r2m_bench_1gpu.err; echo "bench 1 GPU exit=$?"
timeout 600 python -m pytest tests/test_pool_gpu.py tests/test_shim_cpp.py -q -m gpu -rf > gpurun_out/r2m_pool_pytest.log 2>&1; echo "pool pytest exit=$?"; tail -2 gpurun_out/r2m_pool_pytest.log
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads([l for l in open(f"gpurun_out/r2m_bench_{n}gpu.json") if l.startswith("{")][-1])
        e, c = d["e2e"], d.get("corpus") or {}
        print(n, "ms", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(e["value"]), "sync", round(e["synchronous_call_value"]),
              "link min", (e.get("host_link") or {}).get("min_over_ranks"), "copy ms", (e.get("host_link") or {}).get("copy_ms_per_step"),
              "corpus", round(c.get("audio_s_per_s", 0)), "one call", round(c.get("one_synchronous_call_per_rank_audio_s_per_s", 0)), c.get("rank_time_max_over_mean"), c.get("bit_identical"))
    except Exception as ex:
        print(n, "unreadable", ex)
PY
