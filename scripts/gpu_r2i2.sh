#!/bin/bash
# visit I, second pass: the allocation-free entry point's test alone in a fresh process, then behind the whole suite
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -rf > gpurun_out/r2i_noalloc_alone.log 2>&1; echo "no-alloc alone exit=$?"; grep -E "^(E  |FAILED)" gpurun_out/r2i_noalloc_alone.log | head -12; tail -2 gpurun_out/r2i_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu -rf > gpurun_out/r2i_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r2i_pytest.log; grep -E "^(E  |FAILED)" gpurun_out/r2i_pytest.log | head -24; tail -3 gpurun_out/r2i_pytest.log
