"""Development probe: mel parity + tcgen05 GEMM parity on the GPU (first bring-up)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import q3a_b200  # noqa
from q3a_b200 import binding as B, synth
from oracle import mel_oracle as MO

B.load(strict=False)
print("devices", B.load().q3a_device_count(), flush=True)
# ---- mel
fe = B.Frontend(0)
filt = B.generate_mel_filters()
filt_o = MO.mel_filters_numpy()
print("filters max diff vs numpy oracle", np.abs(filt - filt_o).max(), "nnz", (filt != 0).sum())
try:
    ref = MO.RefMel()
    print("filters max diff vs reference", np.abs(filt - ref.filters()).max())
except Exception as e:
    ref = None; print("no ref", e)
for n in (16000, 480000, 501234, 100, 161, 3200):
    pcm = synth.synth_audio(100 + n % 7, n)
    got = fe.mel_batch([pcm])[0]
    want = MO.log_mel_numpy(pcm, filt_o)
    d = np.abs(got - want).max() if want.size else 0.0
    print(f"mel n={n} shape={got.shape} maxabs={d:.3e}", flush=True)
clips = [synth.synth_audio(200 + i, 480000 - 37 * i) for i in range(8)]
t0 = time.time(); outs = fe.mel_batch(clips); t1 = time.time()
print("batch8 host api ms", (t1 - t0) * 1e3, "max err", max(np.abs(o - MO.log_mel_numpy(c, filt_o)).max() for o, c in zip(outs, clips)))
# ---- gemm
torch.manual_seed(0)
for opd, dt in ((B.OP_BF16, torch.bfloat16), (B.OP_F16, torch.float16)):
    for (M, N, K, bn) in ((128, 128, 64, 128), (128, 256, 128, 256), (390, 896, 896, 224), (1000, 2688, 896, 224),
                          (777, 3584, 896, 256), (513, 896, 3584, 224), (130, 1024, 896, 256), (260, 480, 4320, 128)):
        a = (torch.randn(M, K, device="cuda") * 0.5).to(dt)
        w = (torch.randn(N, K, device="cuda") * 0.5).to(dt)
        bias = torch.randn(N, device="cuda")
        out = torch.zeros(M, N, device="cuda")
        B.test_gemm(a, w, bias, out, opd, bn)
        torch.cuda.synchronize()
        want = a.float() @ w.float().T + bias
        err = (out - want).abs().max().item() / want.abs().max().item()
        print(f"gemm {dt} M={M} N={N} K={K} bn={bn} rel_err={err:.3e}", flush=True)
# timing of a large GEMM
M, N, K = 24960, 3584, 896
a = (torch.randn(M, K, device="cuda")).to(torch.bfloat16); w = torch.randn(N, K, device="cuda").to(torch.bfloat16)
out = torch.zeros(M, N, device="cuda"); bias = torch.zeros(N, device="cuda")
for _ in range(3): B.test_gemm(a, w, bias, out, B.OP_BF16, 256)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): B.test_gemm(a, w, bias, out, B.OP_BF16, 256)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"gemm {M}x{N}x{K}: {ms:.3f} ms  {2*M*N*K/ms/1e9:.1f} TFLOP/s")
