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
if os.environ.get("PROBE_OVERLAP"):
    d = torch.empty((BATCH, n), dtype=torch.float32, device="cuda"); d.copy_(host)
    o = torch.empty((T * BATCH, 1024), dtype=torch.float32, device="cuda"); oh = torch.empty((T * BATCH, 1024), dtype=torch.float32).pin_memory()
    d2 = torch.empty_like(d)
    pcm_off = [i * n for i in range(BATCH)]
    s1, s2, s3 = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    def ev(): return torch.cuda.Event(enable_timing=True)
    for load in (False, True):
        torch.cuda.synchronize()
        a0, a1, b0, b1, c0, c1 = ev(), ev(), ev(), ev(), ev(), ev()
        if load:
            with torch.cuda.stream(s1):
                a0.record(s1)
                for _ in range(2): enc.encode_pcm_device(fe, d.data_ptr(), pcm_off, [n] * BATCH, o.data_ptr(), B.MODE_ASR, s1.cuda_stream)
                a1.record(s1)
        with torch.cuda.stream(s2):
            b0.record(s2); d2.copy_(host, non_blocking=True); b1.record(s2)
        with torch.cuda.stream(s3):
            c0.record(s3); oh.copy_(o, non_blocking=True); c1.record(s3)
        torch.cuda.synchronize()
        print(f"under load={load}: h2d {b0.elapsed_time(b1):.2f} ms, d2h {c0.elapsed_time(c1):.2f} ms" + (f", compute x2 {a0.elapsed_time(a1):.2f} ms" if load else ""))
if os.environ.get("PROBE_SPANS"):
    enc.set_profiling(True)
    for _ in range(3):
        t0 = time.perf_counter(); call(); dt = (time.perf_counter() - t0) * 1e3
        tm = enc.timings()
        print(f"profiled e2e call {dt:.2f} ms: mel {tm.mel_ms:.2f} stem-span {tm.conv_chunk_ms:.2f} xfmr-span {tm.transformer_ms:.2f} gemm {tm.gemm_ms:.2f} attn {tm.attention_ms:.2f} other {tm.other_ms:.2f}")
    enc.set_profiling(False)
