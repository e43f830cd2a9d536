"""Development probe: attention kernel alone on the bench shape (64 segments x 390 tokens, 14 heads)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import q3a_b200  # noqa
from q3a_b200 import binding as B

lib = B.load()
torch.manual_seed(0)
H, L, NS = 14, int(os.environ.get('ATTN_L', 390)), int(os.environ.get('ATTN_NS', 64))
cu = [i * L for i in range(NS + 1)]
T = cu[-1]
qkv = (torch.randn(T, 3 * H * 64, device="cuda") * 1.5).to(torch.bfloat16)
out = torch.zeros(T, H * 64, device="cuda", dtype=torch.bfloat16)
cu_arr = B._int_array(cu)
st = torch.cuda.current_stream().cuda_stream
times = []
for it in range(12):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    B.check(lib.q3a_test_attention(qkv.data_ptr(), out.data_ptr(), T, H, cu_arr, NS, B.OP_BF16, st))
    e1.record(); torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1) * 1e3)
flops = 4.0 * 64 * H * NS * L * L
best = min(times[2:])
print(f"attention {NS}x{L} H=14 ({NS * ((L + 127) // 128) * H} CTAs, {(L + 63) // 64} key blocks): best {best:.1f} us ({flops / best / 1e6:.0f} TFLOP/s), all {[round(t) for t in times]}")
# correctness on one segment
q, k, v = [x.float().view(T, H, 64) for x in qkv.split(H * 64, dim=1)]
s = slice(0, L)
p = torch.softmax(torch.einsum("qhd,khd->hqk", q[s], k[s]) * 0.125, dim=-1)
want = torch.einsum("hqk,khd->qhd", p, v[s]).reshape(L, H * 64)
print("maxabs err seg0:", (out[s].float() - want).abs().max().item())
