"""CPU replays of index arithmetic and bit tricks that live in device code (no GPU): each function mirrors a few lines of
a kernel, cited next to it, and is checked exhaustively over the ranges the kernels accept."""
import numpy as np


def _items(L):                      # attention2_items / attention2_item_count (attention2.cu)
    n = (L + 255) // 256
    rows = (((L + n - 1) // n) + 31) & ~31
    out, q = [], 0
    while q < L:
        out.append((q, min(rows, L - q)))
        q += rows
    return n, rows, out


def test_attention2_items_and_key_blocks_cover_every_segment_length():
    for L in range(1, 9000):
        n, rows, it = _items(L)
        assert rows <= 256 and len(it) == n and sum(r for _, r in it) == L      # the encoder's prefix sums rely on len == n
        nkv = (L + 127) // 128                                                     # Geometry (attention2.cu)
        kb = (((L + nkv - 1) // nkv) + 15) & ~15
        assert kb <= 128 and (nkv - 1) * kb < L <= nkv * kb                        # every block holds at least one key


def test_decoder_attention_sees_exactly_the_causal_prefix():
    """attention_dec.cu: blocks walked by a tile, keys fed to the MMAs (nk), keys a row exponentiates (nv_row), fast path."""
    for n in (1, 2, 15, 16, 17, 63, 64, 65, 127, 128, 129, 191, 192, 193, 300, 515, 1000):
        for q0 in range(0, n, 128):
            q_last = min(q0 + 128, n) - 1
            nkv = q_last // 64 + 1
            for tid in range(128):
                row = q0 + tid
                if row >= n:
                    continue
                seen = []
                for j in range(nkv):
                    k0 = j * 64
                    rem = min(n, q_last + 1) - k0
                    nk = 64 if rem >= 64 else (rem + 15) & ~15
                    nv_row = max(0, min(min(64, n - k0), row - k0 + 1))
                    full = (q0 + (tid >> 5) * 32 - k0 + 1) >= 64 and (n - k0) >= 64
                    assert nv_row <= nk and (not full or (nk == 64 and nv_row == 64))
                    seen += list(range(k0, k0 + (64 if full else nv_row)))
                assert seen == list(range(row + 1))


def test_q8_dequant_bit_trick_is_exact():
    """gemm_tc.cuh (kWq8): x = 1 + (q + 128) 2^-15 from a byte permute, fma(x, 2^15 d, -(2^15 + 128) d) == float(q) * d
    for every int8 q and f16 scale d (the product has at most 18 significant bits, so neither path rounds)."""
    rng = np.random.default_rng(0)
    scales = np.concatenate([(rng.standard_normal(3000) * 0.01).astype(np.float16),
                             np.array([6e-8, 1e-5, 65504, -3.1, 0.0, 1.0], np.float16)])
    q = np.arange(-128, 128)
    u8 = (q.astype(np.int8).view(np.uint8) ^ 0x80).astype(np.uint32)
    x = (np.uint32(0x3F800000) | (u8 << np.uint32(8))).astype(np.uint32).view(np.float32).astype(np.float64)
    for d in scales:
        d = np.float32(d)
        mul, add = np.float32(d * np.float32(32768.0)), np.float32(np.float32(-32896.0) * d)
        assert float(mul) == float(d) * 32768.0 and float(add) == -32896.0 * float(d)          # both constants are exact
        got = (x * float(mul) + float(add)).astype(np.float32)                                  # the FMA is exact in f64
        assert np.array_equal(got, q.astype(np.float32) * d)


def test_fma_pipe_exp2_accuracy():
    """attention2.cu ex2_poly, evaluated in float32 exactly as written"""
    f = np.float32
    x = np.linspace(-126.5, 10, 400001).astype(f)
    xc = np.maximum(x, f(-125))
    magic = f(12582912.0)
    t = (xc + magic).astype(f)
    r = (xc - (t - magic)).astype(f)
    p = (r * f(0.00956051) + f(0.05591704)).astype(f)
    for c in (0.24024981, 0.69312197, 0.99999919):
        p = (p * r + f(c)).astype(f)
    y = (p.view(np.uint32) + (t.view(np.uint32) << np.uint32(23))).astype(np.uint32).view(f)
    ref = np.exp2(xc.astype(np.float64))
    assert np.abs(r).max() <= 0.5 and (np.abs(y - ref) / ref).max() <= 5e-6


def test_gemm_shared_memory_plan_and_epilogue_statistics_layout():
    """Restates gemm_tc.cuh's GemmSmem arithmetic: the ring depth every Linear GEMM of the encoder gets for its epilogue
    warp count, and that the statistics partials of a column block (sums over its even / odd 32-column chunks) are the same
    sets of chunks whether one warp (kEW = 4) or two (kEW = 8) own a TMEM lane quadrant."""
    limit, barrier = 232448, 512

    def stages(block_n, out_bytes, dual, residual, kew, staged=True):
        a, b = 128 * 64 * 2, (block_n // 2) * 64 * 2
        buf = (32 * 32 * out_bytes + (32 * 32 * 2 if dual else 0)) if staged else 0
        bufs = (3 if residual and not dual else 2) if staged else 0
        avail = limit - 1024 - barrier - kew * bufs * buf
        assert (a + b) % 1024 == 0
        return min(6, avail // (a + b))

    assert stages(240, 2, False, False, 8, staged=False) == 6        # conv2 / conv3
    assert stages(224, 2, False, False, 12) == 5                      # QKV with attn_norm folded in
    assert stages(256, 2, False, False, 12) == 5                      # FFN-up + GELU
    assert stages(224, 4, False, True, 8) == 4                        # out-projection + residual
    assert stages(224, 4, True, True, 8) == 4                         # FFN-down, eight epilogue warps
    assert stages(224, 4, True, True, 4) == 5                         # ... four: the fifth stage
    for nch in range(1, 9):                                           # chunks of a column block (7 for BLOCK_N = 224)
        eight = [[c for c in range(hf, nch, 2)] for hf in (0, 1)]     # warp hf of a quadrant takes c = hf, hf + 2, ...
        four = [[c for c in range(nch) if c % 2 == p] for p in (0, 1)]
        assert eight == four
