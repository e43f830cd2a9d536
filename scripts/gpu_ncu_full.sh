#!/bin/bash
# ncu --set full capture of the first launches matching $1 (regex) in one bench step; $2 = count, $3 = skip
set -u
mkdir -p gpurun_out
PAT="${1:-gemm_tc_kernel}"; CNT="${2:-8}"; SKIP="${3:-0}"; NAME="${4:-prof_sel}"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$PAT" -s "$SKIP" -c "$CNT" -f -o "gpurun_out/$NAME" $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit=$?"
tail -3 gpurun_out/ncu_full.log
