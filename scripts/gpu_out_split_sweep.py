"""Development probe: host-pointer call time (64 x 30 s, pinned buffers) against the download split
(Q3A_OUT_FRAC) and the number of layers run per download part (Q3A_TAIL_LAYERS).  Outputs must not change."""
import os, sys, time, tempfile, hashlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, ctypes as C
import q3a_b200  # noqa
from q3a_b200 import binding as B, synth

B.load()
shape = synth.ASR_06B
path = synth.write_gguf(os.path.join(tempfile.mkdtemp(), "asr.gguf"), shape, synth.random_weights(shape, 0), "f16")
enc = B.Encoder(path, 0, B.OP_BF16); fe = B.Frontend(0)
BATCH, n = 64, 480000
host = torch.from_numpy(np.stack([synth.synth_audio(200 + i, n) for i in range(BATCH)])).pin_memory()
T = enc.n_tokens(n // 160)
out_host = torch.empty((BATCH, T, enc.hp.out_dim), dtype=torch.float32).pin_memory()
outs_ptr = (C.c_void_p * BATCH)(*[out_host[i].data_ptr() for i in range(BATCH)])
ns_arr = (C.c_int * BATCH)(*([n] * BATCH)); ntok = (C.c_int * BATCH)()
pcm_ptrs = (C.c_void_p * BATCH)(*[host[i].data_ptr() for i in range(BATCH)])
def call():
    B.check(enc.lib.q3a_encode_pcm_batch(fe.h, enc.h, pcm_ptrs, ns_arr, BATCH, B.MODE_ASR, outs_ptr, ntok))
def measure(tag, env):
    for k in ("Q3A_OUT_FRAC", "Q3A_TAIL_LAYERS", "Q3A_NO_PART_OVERLAP"): os.environ.pop(k, None)
    os.environ.update(env)
    for _ in range(3): call()
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); call(); ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort()
    digest = hashlib.sha256(out_host.numpy().tobytes()).hexdigest()[:12]
    print(f"{tag:38s} min {ts[0]:6.2f}  med {ts[5]:6.2f} ms   {BATCH*30/ts[5]:6.1f} k audio-s/s   sha {digest}", flush=True)
measure("default (0.5,1 / 3)", {})
measure("no overlap", {"Q3A_NO_PART_OVERLAP": "1"})
for frac in ("0.5,1", "0.6,1", "0.4,0.75,1", "0.34,0.67,1", "0.3,0.55,0.8,1", "0.25,0.5,0.75,1"):
    for tail in (2, 3, 4, 6):
        measure(f"frac {frac} tail {tail}", {"Q3A_OUT_FRAC": frac, "Q3A_TAIL_LAYERS": str(tail)})
measure("default again", {})
