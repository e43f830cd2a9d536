#!/bin/bash
# round 2, visit A: attention v2 bring-up, the new parity tests, a bench line
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_attention_gpu.py -x -q -m gpu -k attention > gpurun_out/r2a_attn_pytest.log 2>&1
rc=$?; echo "attention pytest exit=$rc" | tee -a gpurun_out/r2a_attn_pytest.log
tail -5 gpurun_out/r2a_attn_pytest.log
if [ $rc -ne 0 ]; then export Q3A_ATTN_V1=1; export ATTN_V=1; echo "FALLING BACK TO ATTENTION V1 FOR THE REST"; fi
timeout 300 python scripts/gpu_attn_bench.py > gpurun_out/r2a_attn_bench.log 2>&1; echo "attn bench exit=$?"; cat gpurun_out/r2a_attn_bench.log | tail -12
timeout 1500 python -m pytest tests/test_encoder_gpu.py -x -q -m gpu -s -k "real_widths or device_entry or pad_rows or full_shape or three_way or tiny_q8 or 30s_clip" > gpurun_out/r2a_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench exit=$?"
tail -15 gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_bench.json; tail -3 gpurun_out/r2a_bench.err
