"""Throughput of BASELINE.json configs 3, 4 and 5 on one B200 through the host-pointer C ABI (q3a_encode_pcm_batch:
pinned host PCM in, pinned host features out).  Config 2 is bench.py's line; these are parity-test cases whose speed is
recorded for DESIGN.md.  Writes gpurun_out/configs.json."""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np, torch
import q3a_b200  # noqa
from q3a_b200 import binding as B, synth

B.load()
tmp = tempfile.mkdtemp()
res = {}


def run(name, enc, fe, lens, mode, reps=4):
    base = synth.synth_audio(77, max(lens))
    hosts = [torch.from_numpy(np.ascontiguousarray(base[:n])).pin_memory() for n in lens]
    outs = [torch.empty((max(enc.n_tokens(n // 160, mode), 1), enc.hp.out_dim), dtype=torch.float32).pin_memory() for n in lens]
    k = len(lens)
    pp = (C.c_void_p * k)(*[h.data_ptr() for h in hosts]); po = (C.c_void_p * k)(*[o.data_ptr() for o in outs])
    ns = (C.c_int * k)(*lens); nt = (C.c_int * k)()
    def call():
        B.check(enc.lib.q3a_encode_pcm_batch(fe.h, enc.h, pp, ns, k, mode, po, nt))
    call(); call()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); call(); ts.append(time.perf_counter() - t0)
    secs = sum(lens) / 16000.0
    best, med = min(ts), sorted(ts)[len(ts) // 2]
    res[name] = {"utterances": k, "audio_s": secs, "tokens": int(sum(nt)), "ms_median": med * 1e3, "ms_best": best * 1e3,
                 "audio_s_per_s_median": secs / med, "device_bytes": enc.device_bytes()}
    print(name, json.dumps(res[name]), flush=True)


fe = B.Frontend(0)
# config 3: ASR-0.6B, Q8_0 GGUF, 8 x 92-s clips: (a) one 1196-token full-attention sequence per clip, (b) cut into 30-s pieces
shape = synth.ASR_06B
w = synth.random_weights(shape, 0)
q8_path = synth.write_gguf(os.path.join(tmp, "asr_q8.gguf"), shape, w, "q8_0")
# row a12: Q8_0 weights resident in HBM and dequantised inside the GEMM, against expanding them at load (Q3A_Q8_DEQUANT_AT_LOAD):
# weight-arena bytes, single-clip latency, 64-clip throughput
for tag, env in (("fused", "Q3A_Q8_FUSED"), ("at_load", "Q3A_Q8_DEQUANT_AT_LOAD")):
    os.environ.pop("Q3A_Q8_FUSED", None); os.environ.pop("Q3A_Q8_DEQUANT_AT_LOAD", None)
    os.environ[env] = "1"
    enc = B.Encoder(q8_path, 0, B.OP_BF16)
    res[f"q8_{tag}_weight_bytes"] = enc.device_bytes()
    print(f"q8 {tag}: weight arena {enc.device_bytes() / 1e6:.1f} MB", flush=True)
    run(f"q8_{tag}_1x30s_latency", enc, fe, [480000], B.MODE_ASR, reps=8)
    run(f"q8_{tag}_64x30s", enc, fe, [480000] * 64, B.MODE_ASR)
    sfx = "_fused" if tag == "fused" else ""
    run("config3_q8_92s_full_attention" + sfx, enc, fe, [92 * 16000] * 8, B.MODE_ASR)
    run("config3_q8_92s_cut_30s" + sfx, enc, fe, [480000, 480000, 480000, 32000] * 8, B.MODE_ASR)
    enc.close()
os.environ.pop("Q3A_Q8_DEQUANT_AT_LOAD", None); os.environ.pop("Q3A_Q8_FUSED", None)
# config 5 (per-GPU share of the 10-h corpus at 8 GPUs = 4500 s): log-uniform 2-60 s utterances, ASR bf16
enc = B.Encoder(synth.write_gguf(os.path.join(tmp, "asr_f16.gguf"), shape, w, "f16"), 0, B.OP_BF16)
rng = np.random.default_rng(500)
lens, tot = [], 0.0
while tot < 4500.0:
    s = float(np.exp(rng.uniform(np.log(2), np.log(60)))); lens.append(int(s * 16000)); tot += s
run("config5_ragged_4500s_shard", enc, fe, lens, B.MODE_ASR)
enc.close(); del w
# config 4: aligner-0.6B (24 layers, d 1024), 64 utterances U[5,120] s, padded-tail stem, 104-token windows
shape = synth.ALIGNER_06B
w = synth.random_weights(shape, 1)
enc = B.Encoder(synth.write_gguf(os.path.join(tmp, "al_f16.gguf"), shape, w, "f16"), 0, B.OP_BF16)
rng = np.random.default_rng(400)
run("config4_aligner_64x5-120s", enc, fe, [int(rng.uniform(5, 120) * 16000) for _ in range(64)], B.MODE_ALIGNER)
# row f4: the aligner's whole device path for one 120-s utterance -- mel + encoder (1560 tokens) on the device, features
# consumed in place by the NAR decoder prefill (28 layers, 1024 wide, 5000 classes) over a 1560 + 240-token prompt
al_fe_dim = enc.hp.out_dim
dshape = synth.ALIGNER_DECODER_06B
dw = synth.random_decoder_weights(dshape, 2)
dec = B.Decoder(synth.write_decoder_gguf(os.path.join(tmp, "dec_f16.gguf"), dshape, dw, "f16"), 0, B.OP_BF16)
n_smp = 120 * 16000
n_audio = enc.n_tokens(n_smp // 160, B.MODE_ALIGNER)
pcm_dev = torch.from_numpy(synth.synth_audio(9, n_smp)).cuda()
feat_dev = torch.empty((n_audio, al_fe_dim), dtype=torch.float32, device="cuda")
n_prompt = n_audio + 240
tokens = [(i * 7 + 3) % dshape.vocab for i in range(n_prompt)]
logits_dev = torch.empty((n_prompt, dshape.n_classes), dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def align_step():
    enc.encode_pcm_device(fe, pcm_dev.data_ptr(), [0], [n_smp], feat_dev.data_ptr(), B.MODE_ALIGNER, st)
    dec.forward_device(tokens, feat_dev.data_ptr(), n_audio, 8, logits_dev.data_ptr(), st)
for part, fn in (("encoder+decoder", align_step),
                 ("decoder_only", lambda: dec.forward_device(tokens, feat_dev.data_ptr(), n_audio, 8, logits_dev.data_ptr(), st))):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    res[f"f4_aligner_120s_{part}"] = {"audio_s": 120.0, "audio_tokens": n_audio, "prompt_tokens": n_prompt, "ms": ms,
                                      "finite": bool(torch.isfinite(logits_dev).all())}
    print(f"f4_aligner_120s_{part}", json.dumps(res[f"f4_aligner_120s_{part}"]), flush=True)
dec.close()
enc.close()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/configs.json", "w"), indent=1)
