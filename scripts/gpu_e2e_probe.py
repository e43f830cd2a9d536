"""Development probe: e2e call time vs pass split, and raw PCIe copy rates."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, ctypes as C
import q3a_b200  # noqa
from q3a_b200 import binding as B, synth

B.load()
shape = synth.ASR_06B
w = synth.random_weights(shape, 0)
path = synth.write_gguf(os.path.join(tempfile.mkdtemp(), "asr.gguf"), shape, w, "f16")
enc = B.Encoder(path, 0, B.OP_BF16); fe = B.Frontend(0)
BATCH, n = 64, 480000
host = torch.randn((BATCH, n), dtype=torch.float32).mul_(0.1).pin_memory()
T = enc.n_tokens(n // 160)
out_host = [torch.empty((T, enc.hp.out_dim), dtype=torch.float32).pin_memory() for _ in range(BATCH)]
outs_ptr = (C.c_void_p * BATCH)(*[o.data_ptr() for o in out_host])
ns_arr = (C.c_int * BATCH)(*([n] * BATCH)); ntok = (C.c_int * BATCH)()
pcm_ptrs = (C.c_void_p * BATCH)(*[host[i].data_ptr() for i in range(BATCH)])
def call():
    B.check(enc.lib.q3a_encode_pcm_batch(fe.h, enc.h, pcm_ptrs, ns_arr, BATCH, B.MODE_ASR, outs_ptr, ntok))
for _ in range(3): call()
ts = []
for _ in range(8):
    t0 = time.perf_counter(); call(); ts.append((time.perf_counter() - t0) * 1e3)
print(f"e2e call ms: min {min(ts):.2f} med {sorted(ts)[4]:.2f}  (Q3A_PIPELINE_MIN_CHUNKS={os.environ.get('Q3A_PIPELINE_MIN_CHUNKS')}, P3={os.environ.get('Q3A_PIPELINE_3')})")
if os.environ.get("PROBE_COPY"):
    d = torch.empty((BATCH, n), dtype=torch.float32, device="cuda")
    for nm, f in (("h2d 123MB", lambda: d.copy_(host, non_blocking=True)),):
        torch.cuda.synchronize(); t0 = time.perf_counter(); f(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print(nm, f"{dt*1e3:.2f} ms {host.numel()*4/dt/1e9:.1f} GB/s")
    o = torch.empty((T * BATCH, 1024), dtype=torch.float32, device="cuda"); oh = torch.empty((T * BATCH, 1024), dtype=torch.float32).pin_memory()
    torch.cuda.synchronize(); t0 = time.perf_counter(); oh.copy_(o, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("d2h 102MB", f"{dt*1e3:.2f} ms {o.numel()*4/dt/1e9:.1f} GB/s")
    # device-only
    pcm_off = [i * n for i in range(BATCH)]
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2): enc.encode_pcm_device(fe, d.data_ptr(), pcm_off, [n] * BATCH, o.data_ptr(), B.MODE_ASR, st)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    enc.encode_pcm_device(fe, d.data_ptr(), pcm_off, [n] * BATCH, o.data_ptr(), B.MODE_ASR, st)
    torch.cuda.synchronize(); print(f"device-only single call {(time.perf_counter()-t0)*1e3:.2f} ms")
