"""tcgen05 building blocks against plain PyTorch fp32 references of the same ops."""
import pytest
import torch

from q3a_b200 import binding as B
from conftest import needs

pytestmark = pytest.mark.gpu
DT = {B.OP_BF16: torch.bfloat16, B.OP_F16: torch.float16}


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
@pytest.mark.parametrize("M,N,K,bn", [
    (128, 128, 64, 128), (1, 8, 8, 128), (129, 136, 72, 128), (390, 896, 896, 224), (1000, 2688, 896, 224),
    (777, 3584, 896, 256), (513, 896, 3584, 224), (130, 1024, 896, 256), (260, 480, 4320, 128),
    (24960, 896, 896, 0), (65, 1024, 1024, 0), (300, 64, 128, 0)])
def test_gemm_matches_fp32_matmul(lib, opd, M, N, K, bn):
    torch.manual_seed(M + N + K)
    a = (torch.randn(M, K, device="cuda") * 0.5).to(DT[opd])
    w = (torch.randn(N, K, device="cuda") * 0.5).to(DT[opd])
    bias = torch.randn(N, device="cuda")
    out = torch.full((M, N), float("nan"), device="cuda")
    B.test_gemm(a, w, bias, out, opd, bn)
    torch.cuda.synchronize()
    want = a.float() @ w.float().T + bias
    # operands are exact in fp32; only the fp32 accumulation order differs
    assert (out - want).abs().max().item() <= 2e-5 * max(want.abs().max().item(), 1.0)


def _attn_ref(qkv, H, cu):
    T, d = qkv.shape[0], H * 64
    q, k, v = qkv[:, :d].float(), qkv[:, d:2 * d].float(), qkv[:, 2 * d:].float()
    out = torch.zeros(T, d, device=qkv.device)
    for s, e in zip(cu[:-1], cu[1:]):
        if e == s:
            continue
        qs, ks, vs = (x[s:e].view(e - s, H, 64).transpose(0, 1) for x in (q, k, v))
        p = torch.softmax(qs @ ks.transpose(1, 2) * 0.125, dim=-1)     # scale 1/sqrt(64), audio_encoder.cpp:476
        out[s:e] = (p @ vs).transpose(0, 1).reshape(e - s, d)
    return out


# outputs reach |o| ~ 5: half an ulp of bf16 there is 1.6e-2 from the output rounding alone, P's rounding adds ~1e-2
@pytest.mark.parametrize("version", [pytest.param(2, marks=needs("attention_v2")), 1])
@pytest.mark.parametrize("opd,tol", [(B.OP_BF16, 3e-2), (B.OP_F16, 4e-3)])
@pytest.mark.parametrize("H,cu", [
    (2, [0, 128]), (2, [0, 16]), (1, [0, 1]), (3, [0, 100]), (14, [0, 390]), (16, [0, 104, 208, 250]),
    (2, [0, 1, 130, 400, 1596]), (4, [0, 129, 129, 257]), (14, [0, 1560]), (16, [104 * i for i in range(16)]),
    (3, [0, 256, 513, 770, 1027, 1028]), (2, [0, 7805]), (14, [390 * i for i in range(24)])])
def test_attention_matches_fp32_softmax(lib, opd, tol, H, cu, version):
    """full (one segment per utterance), windowed (104-token segments), ragged and empty segments; both kernel
    generations (1 = one tile per CTA, what the encoder runs; 2 = persistent CTAs with two ping-ponging query tiles, opt-in).
    The added shapes hit v2's item boundaries (256 / 257 rows, a one-row tile B), 61 key blocks with many items per
    CTA, and more (item, head) units than SMs so that every CTA walks several items."""
    torch.manual_seed(len(cu) * 100 + H)
    T = cu[-1]
    qkv = (torch.randn(T, 3 * H * 64, device="cuda") * 1.5).to(DT[opd])
    out = torch.full((T, H * 64), float("nan"), device="cuda").to(DT[opd])
    B.test_attention(qkv, out, H, cu, opd, version)
    torch.cuda.synchronize()
    want = _attn_ref(qkv, H, cu)
    assert torch.isfinite(out.float()).all()
    # P and the output are rounded to the 16-bit operand type: 2^-8 (bf16) / 2^-11 (f16) relative
    assert (out.float() - want).abs().max().item() <= tol


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
def test_attention_key_block_is_chosen_per_segment(lib, opd):
    """The 64- / 128-key block size follows the segment's own length (128 from 2048 tokens, attention.cu
    attention_key_block), never the batch: a clip's rows are bit-identical whether it runs alone (one launch) or packed
    between longer and shorter clips (two launches, one per block size)."""
    H = 3
    torch.manual_seed(77)
    lens = [390, 2500, 104, 2048, 2047]
    parts = [(torch.randn(n, 3 * H * 64, device="cuda") * 1.5).to(DT[opd]) for n in lens]
    cu = [0]
    for n in lens:
        cu.append(cu[-1] + n)
    qkv = torch.cat(parts)
    out = torch.full((cu[-1], H * 64), float("nan"), device="cuda").to(DT[opd])
    B.test_attention(qkv, out, H, cu, opd, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    for i, n in enumerate(lens):
        alone = torch.full((n, H * 64), float("nan"), device="cuda").to(DT[opd])
        B.test_attention(parts[i], alone, H, [0, n], opd, 1)
        torch.cuda.synchronize()
        assert torch.equal(alone, out[cu[i]:cu[i + 1]]), f"segment {i} ({n} tokens) depends on its batch"
