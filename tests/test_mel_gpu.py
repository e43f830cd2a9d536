"""CUDA log-mel frontend against the reference (golden vectors made by the reference's own TU) and the
numpy oracle.  Tolerance: max-abs-error <= 1e-4 (BASELINE.json north_star; reference run_all_tests.sh:134)."""
import os

import numpy as np
import pytest

from oracle import mel_oracle as MO
from q3a_b200 import binding as B, synth

pytestmark = pytest.mark.gpu
TOL = 1e-4
GOLD = os.path.join(os.path.dirname(__file__), "golden", "mel_ref.npz")


@pytest.fixture(scope="module")
def fe(lib):
    f = B.Frontend(0)
    yield f
    f.close()


@pytest.mark.parametrize("seed,n", [(101, 16000), (102, 5957), (103, 40000)])
def test_matches_reference_golden(fe, seed, n):
    g = np.load(GOLD)
    got = fe.mel_batch([synth.synth_audio(seed, n)])[0]
    want = g[f"mel_{seed}_{n}"]
    assert got.shape == want.shape
    assert np.abs(got - want).max() <= TOL


@pytest.mark.skipif(not os.path.exists(MO.REF_SO), reason="oracle/_ref not built")
def test_matches_live_reference_tu(fe):
    ref = MO.RefMel()
    filt = ref.filters()
    for seed, n in ((1, 8000), (2, 16000 + 159), (3, 161), (4, 400), (5, 24000)):
        pcm = synth.synth_audio(seed, n)
        assert np.abs(fe.mel_batch([pcm])[0] - ref.log_mel(pcm, filt)).max() <= TOL


@pytest.mark.parametrize("n", [0, 1, 159, 160, 161, 200, 201, 399, 400, 401, 5120, 5121, 5279, 480000, 501234])
def test_edge_lengths_against_oracle(fe, filters, n):
    """empty, shorter than the reflect pad (zero-filled edges, mel_spectrogram.cpp:499-515), exactly one tile, ragged"""
    pcm = synth.synth_audio(7, n)
    got = fe.mel_batch([pcm])[0]
    want = MO.log_mel_numpy(pcm, filters)
    assert got.shape == want.shape == (128, n // 160)
    if want.size:
        assert np.abs(got - want).max() <= TOL


def test_silence_and_floor(fe, filters):
    got = fe.mel_batch([np.zeros(16000, np.float32)])[0]
    assert np.allclose(got, (-10.0 + 4.0) / 4.0)                 # log10(1e-10) everywhere, clamp is a no-op
    pcm = synth.synth_audio(9, 48000)
    got = fe.mel_batch([pcm])[0]
    # all-zero filterbank rows (mel 0, 3, 6, 13) sit at the clamp floor max-8
    floor = got.min()
    assert np.all(got[[0, 3, 6, 13]] == floor)
    assert abs((got.max() * 4 - 4) - (floor * 4 - 4) - 8.0) < 1e-5


def test_ragged_batch_equals_single_clips(fe, filters):
    lens = [480000, 31 * 16000 + 5920, 160, 0, 16000 * 5 + 1, 92 * 16000]
    clips = [synth.synth_audio(300 + i, n) for i, n in enumerate(lens)]
    outs = fe.mel_batch(clips)
    for pcm, o in zip(clips, outs):
        single = fe.mel_batch([pcm])[0]
        assert np.array_equal(o, single)                           # batch composition does not change a clip's result
        if o.size:
            assert np.abs(o - MO.log_mel_numpy(pcm, filters)).max() <= TOL


def test_custom_filterbank(fe, lib):
    rng = np.random.default_rng(0)
    filt = (rng.random((80, 201)) * (rng.random((80, 201)) < 0.05)).astype(np.float32)
    f2 = B.Frontend(0, filt)
    pcm = synth.synth_audio(11, 32000)
    got = f2.mel_batch([pcm])[0]
    assert got.shape == (80, 200)
    assert np.abs(got - MO.log_mel_numpy(pcm, filt)).max() <= TOL
    f2.close()


def test_full_size_properties(fe, filters):
    """BASELINE config 2 size (64 x 30 s): range/clamp invariants and gain equivariance, no oracle needed.
    Scaling the PCM by g shifts every unclamped log-mel value by 2*log10(g)/4 — and the clamp moves with the max."""
    clips = [synth.synth_audio(200 + i, 480000) for i in range(64)]
    outs = fe.mel_batch(clips)
    half = fe.mel_batch([c * np.float32(0.5) for c in clips[:8]])
    for i, o in enumerate(outs):
        assert o.shape == (128, 3000) and np.isfinite(o).all()
        assert o.max() - o.min() <= 2.0 + 1e-6                      # clamp to max-8, then /4
    for o, h in zip(outs[:8], half):
        d = o - h
        shift = 2 * np.log10(2.0) / 4
        interior = (o > o.min() + 1e-3) & (h > h.min() + 1e-3)
        assert np.abs(d[interior] - shift).max() < 2e-4
    # spot-check three clips against the oracle at full length
    for i in (0, 31, 63):
        assert np.abs(outs[i] - MO.log_mel_numpy(clips[i], filters)).max() <= TOL


def test_device_resident_and_unfinalized(fe, filters):
    import torch
    clips = [synth.synth_audio(400 + i, n) for i, n in enumerate((48000, 16000 * 7 + 33))]
    offs, cur = [], 0
    for c in clips:
        offs.append(cur); cur += (len(c) + 63) // 64 * 64
    pcm = torch.zeros(cur + 64, device="cuda")
    for c, o in zip(clips, offs):
        pcm[o:o + len(c)] = torch.from_numpy(c).cuda()
    moffs, cur = [], 0
    for c in clips:
        moffs.append(cur); cur += 128 * (len(c) // 160)
    mel = torch.zeros(cur, device="cuda")
    mx = torch.zeros(len(clips), device="cuda")
    fe.mel_batch_device(pcm.data_ptr(), offs, [len(c) for c in clips], mel.data_ptr(), moffs, mx.data_ptr(), finalize=False)
    torch.cuda.synchronize()
    for i, c in enumerate(clips):
        raw = MO.log_mel_raw_numpy(c, filters)[:, :len(c) // 160]
        got = mel[moffs[i]:moffs[i] + raw.size].cpu().numpy().reshape(raw.shape)
        # raw log10 values are 4x the final scale.  Every consumer clamps at (clip max - 8) first (the fused
        # path does it inside conv1's loads, mel_spectrogram.cpp:615-625), so the 1e-4 bound of the final
        # scale applies after that clamp; cells below it sit at the fp32 noise floor of the FFT (relative
        # to the frame's strongest bin) and only get a sanity bound.
        floor = raw.max() - 8.0
        assert np.abs(np.maximum(got, floor) - np.maximum(raw, floor)).max() <= 4e-4
        assert np.abs(got - raw).max() <= 5e-3
        assert abs(float(mx[i]) - raw.max()) <= 4e-4
