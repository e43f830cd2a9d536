#!/bin/bash
# round 2, visit A: bring-up of attention v2, LayerNorm fold, fused Q8_0, pool, async call; new parity tests; A/B bench lines.
# Every stage runs in its own process under its own timeout, so that one broken kernel does not take the others down.
set -u
mkdir -p gpurun_out
run() { name=$1; shift; "$@" > gpurun_out/r2a_$name.log 2>&1; rc=$?; echo "== $name exit=$rc"; tail -${TAIL:-4} gpurun_out/r2a_$name.log; return $rc; }
PT="python -m pytest -x -q -m gpu"
run attn_pytest timeout 900 $PT tests/test_gemm_attention_gpu.py -k attention
if [ $? -ne 0 ]; then
  run attn_v1_pytest timeout 600 $PT tests/test_gemm_attention_gpu.py -k "attention and 1-"
  export Q3A_ATTN_V1=1; export ATTN_V=1; echo "FALLING BACK TO ATTENTION V1 FOR THE REST"
fi
TAIL=12 run attn_bench timeout 300 python scripts/gpu_attn_bench.py
run fold_pytest timeout 900 $PT tests/test_encoder_gpu.py -k "tiny_tower or hf_audio_tower or debug_variants or golden"
if [ $? -ne 0 ]; then
  Q3A_NO_LN_FOLD=1 run nofold_pytest timeout 900 $PT tests/test_encoder_gpu.py -k "tiny_tower or hf_audio_tower or debug_variants or golden"
  export Q3A_NO_LN_FOLD=1; echo "FALLING BACK TO UNFUSED LAYERNORM FOR THE REST"
fi
run gemm_pytest timeout 600 $PT tests/test_gemm_attention_gpu.py -k gemm
TAIL=8 run q8_pytest timeout 900 $PT -s tests/test_encoder_gpu.py -k "q8"
TAIL=8 run parity_pytest timeout 1500 $PT -s tests/test_encoder_gpu.py -k "device_entry or pad_rows or full_shape or 30s_clip or batch_invariance or async or merged_copy"
run pool_pytest timeout 600 $PT tests/test_pool_gpu.py
TAIL=8 run decoder_pytest timeout 600 $PT tests/test_decoder_gpu.py
run shim_pytest timeout 600 $PT tests/test_shim_cpp.py
run bench timeout 900 python bench.py --steps 10 --warmup 3; cat gpurun_out/r2a_bench.log | head -c 6000
Q3A_NO_LN_FOLD=1 run bench_nofold timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-corpus
Q3A_ATTN_V1=1 run bench_attnv1 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-corpus
Q3A_CONV_TAP_MAJOR=1 run bench_tapmajor timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-corpus
Q3A_STEM_GROUP=11 run bench_group11 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-corpus
Q3A_STEM_GROUP=22 run bench_group22 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-corpus
python - <<'PY'
import json
for n in ("bench", "bench_nofold", "bench_attnv1", "bench_tapmajor", "bench_group11", "bench_group22"):
    try:
        line = [l for l in open(f"gpurun_out/r2a_{n}.log") if l.startswith("{")][-1]
        d = json.loads(line)
        st = d.get("stage_ms_per_step", {})
        print(n, "ms/step", round(d["ms_per_step"], 3), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "sync", round(d["e2e"].get("synchronous_call_value", 0)),
              "gemm frac", round(d["roofline"]["frac"], 3), "attn frac", round(d["roofline_attention"]["frac"], 3), "attn ms", round(d["roofline_attention"]["ms_per_step"], 3),
              "stem ms", round(st.get("audio_encoding.conv_chunk", 0), 3), "xfmr ms", round(st.get("audio_encoding.transformer", 0), 3), "launches", d["gpu_launches"],
              "corpus", d.get("corpus", {}).get("audio_s_per_s"))
    except Exception as ex:
        print(n, "unreadable", ex)
PY
