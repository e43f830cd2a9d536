#!/bin/bash
# round 2, visit N: one `ncu --set full` capture of the final kernels (stem GEMMs + every launch of layer 0) for the
# per-kernel summary under profiles/ (tensor-active %, warps per scheduler, top stalls)
set -u
mkdir -p gpurun_out
rm -f gpurun_out/prof_*.ncu-rep
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-corpus --no-link-probe"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'attention_kernel|gemm_tc_kernel|conv1_tc_kernel|layernorm_kernel|mel_frames_kernel' --launch-skip 116 -c 16 -o gpurun_out/prof_final $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"; ls -la gpurun_out/prof_final.ncu-rep; tail -3 gpurun_out/ncu_full.log
