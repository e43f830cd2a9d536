"""Development probe: the attention kernels alone (both generations) on the bench shape (64 segments x 390 tokens,
14 heads) and a few other shapes; the test hook's table upload is inside the timed region for both, so compare v1 : v2."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import q3a_b200  # noqa
from q3a_b200 import binding as B

lib = B.load()
torch.manual_seed(0)
SHAPES = [(14, 390, 64), (14, 1196, 8), (16, 104, 480), (14, 4096, 4)]
if os.environ.get("ATTN_L"):
    SHAPES = [(14, int(os.environ["ATTN_L"]), int(os.environ.get("ATTN_NS", 64)))]
VERSIONS = [int(v) for v in os.environ.get("ATTN_V", "1,2").split(",")]
POLYS = [int(v) for v in os.environ.get("ATTN_POLY", "0,4,3,2").split(",")]
for H, L, NS in SHAPES:
    cu = [i * L for i in range(NS + 1)]
    T = cu[-1]
    qkv = (torch.randn(T, 3 * H * 64, device="cuda") * 1.5).to(torch.bfloat16)
    q, k, v = [x.float().view(T, H, 64) for x in qkv.split(H * 64, dim=1)]
    s = slice((NS - 1) * L, NS * L)
    p = torch.softmax(torch.einsum("qhd,khd->hqk", q[s], k[s]) * 0.125, dim=-1)
    want = torch.einsum("hqk,khd->qhd", p, v[s]).reshape(L, H * 64)
    cu_arr = B._int_array(cu)
    st = torch.cuda.current_stream().cuda_stream
    flops = 4.0 * 64 * H * NS * L * L
    for ver, poly in [(v, p) for v in VERSIONS for p in (POLYS if v == 2 else [0])]:
        os.environ["Q3A_ATTN2_POLY"] = str(poly)
        out = torch.zeros(T, H * 64, device="cuda", dtype=torch.bfloat16)
        times = []
        for it in range(12):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            B.check(lib.q3a_test_attention_v(qkv.data_ptr(), out.data_ptr(), T, H, cu_arr, NS, B.OP_BF16, ver, st))
            e1.record(); torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1) * 1e3)
        best = min(times[2:])
        err = (out[s].float() - want).abs().max().item()
        print(f"attention v{ver} poly={poly} {NS}x{L} H={H}: best {best:.1f} us ({flops / best / 1e6:.0f} TFLOP/s), last-seg max err {err:.4f}, "
              f"all {[round(t) for t in times]}", flush=True)
