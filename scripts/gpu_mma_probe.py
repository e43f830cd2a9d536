"""Development probe: cycles per tcgen05.mma (M = 128, K = 16, bf16, cta_group::1) against N, operand source and accumulator
reuse -- the number that decides how attention at head_dim 64 has to be tiled."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import q3a_b200  # noqa
from q3a_b200 import binding as B

lib = B.load()
n = 4096
for grid in (1, 148):
    out = torch.zeros(grid, dtype=torch.int64, device="cuda")
    for N in (16, 64, 128, 256):                      # M = 64: what a remainder tile of <= 64 query rows would cost
        for ts in (0, 1):
            B.check(lib.q3a_probe_mma_rate(N, ts | 2, n, 0, 4, grid, out.data_ptr(), None))
            torch.cuda.synchronize()
            per = float(out.float().median()) / n
            print(f"grid {grid:3d} M=64 A from {'TMEM' if ts else 'smem'} N={N:3d}: {per:6.1f} clk/MMA", flush=True)
    for ts in (0, 1):
        for alt in (0, 1):
            for N in (16, 32, 64, 112, 128, 192, 256):
                if alt and N > 192:
                    continue
                B.check(lib.q3a_probe_mma_rate(N, ts, n, alt, 4, grid, out.data_ptr(), None))
                torch.cuda.synchronize()
                c = out.float()
                per = float(c.median()) / n
                nominal = 128 * N / 256
                print(f"grid {grid:3d} A from {'TMEM' if ts else 'smem'} {'alternating D' if alt else 'one D       '} N={N:3d}: "
                      f"{per:6.1f} clk/MMA (nominal {nominal:5.1f}) -> {100 * nominal / per:5.1f}% of the math rate", flush=True)
