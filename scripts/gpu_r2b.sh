#!/bin/bash
# round 2, visit B: full GPU test suite, smoke, both bench arms, then the ncu evidence (launch list with DRAM bytes of one
# bench step; --set full captures of the attention and GEMM kernels).  ncu only after the same command exited 0 plain.
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu -rf > gpurun_out/r2b_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2b_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2b_pytest.log | head -24; tail -4 gpurun_out/r2b_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2b_smoke.log 2>&1; echo "smoke exit=$?"; tail -3 gpurun_out/r2b_smoke.log
timeout 900 python bench.py > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench exit=$?"; cat gpurun_out/r2b_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2b_bench_ref.json 2> gpurun_out/r2b_bench_ref.err; echo "bench ref exit=$?"; cat gpurun_out/r2b_bench_ref.json
for V in "Q3A_LN_FOLD_MASK=1" "Q3A_LN_FOLD_MASK=2" "Q3A_NO_LN_FOLD=1" "Q3A_ATTN_V2=1" "Q3A_CONV_TAP_MAJOR=1" "Q3A_NO_RES_PREFETCH=1"; do
  env $V timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-corpus > "gpurun_out/r2b_bench_ab_${V%%=*}_${V##*=}.json" 2>/dev/null; echo "bench $V exit=$?"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_bench_2ranks_1gpu.json 2> gpurun_out/r2b_bench_2ranks_1gpu.err; echo "2-rank functional bench exit=$?"; tail -c 1500 gpurun_out/r2b_bench_2ranks_1gpu.json
timeout 300 python scripts/gpu_mma_probe.py > gpurun_out/r2b_mma_probe.log 2>&1; echo "mma probe exit=$?"; tail -40 gpurun_out/r2b_mma_probe.log
timeout 300 python scripts/gpu_attn_bench.py > gpurun_out/r2b_attn_bench.log 2>&1; echo "attn bench exit=$?"
timeout 1200 python scripts/gpu_configs.py > gpurun_out/r2b_configs.log 2>&1; echo "configs exit=$?"; tail -12 gpurun_out/r2b_configs.log
python qwen3-asr.cpp_b200/build.py --hash > gpurun_out/csrc_sha256.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 2000 --csv --log-file gpurun_out/launches_all.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit=$?"
Q3A_NO_LN_FOLD=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base --clock-control none -k regex:'q3a|mel|conv1|attention|gemm|layernorm|pcm16' -c 700 --csv --log-file gpurun_out/launches_nofold.csv $CMD > gpurun_out/ncu_launches_nofold.log 2>&1
echo "ncu launches (no fold) exit=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 40 -c 2 -f -o gpurun_out/prof_attn1 $CMD > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn v1 exit=$?"
ACMD="python scripts/gpu_attn_bench.py"
ATTN_V=2 ATTN_POLY=0 ATTN_L=390 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention2 -s 4 -c 2 -f -o gpurun_out/prof_attn2 $ACMD > gpurun_out/ncu_attn2.log 2>&1; echo "ncu attn v2 exit=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 300 -c 8 -f -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm exit=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mel_frames -s 3 -c 1 -f -o gpurun_out/prof_mel $CMD > gpurun_out/ncu_mel.log 2>&1; echo "ncu mel exit=$?"
ls -la gpurun_out | tail -20
