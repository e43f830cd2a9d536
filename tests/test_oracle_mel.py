"""The mel oracle against the reference: golden vectors made by the reference's own TU
(tests/golden/make_golden.py) and, where oracle/_ref/libref_mel.so is present, the TU itself."""
import os

import numpy as np
import pytest

from oracle import mel_oracle as MO
from q3a_b200 import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden", "mel_ref.npz")
CASES = ((101, 16000), (102, 5920 + 37), (103, 40000))


def test_filterbank_matches_reference_golden():
    g = np.load(GOLD)
    bank = MO.mel_filters_numpy()
    assert bank.shape == (128, 201)
    # float32 libm vs numpy may differ in the last ulp of pow/log10
    assert np.abs(bank - g["filters"]).max() <= 1e-6
    # facts probed on the reference: 396 non-zeros, 4 all-zero rows
    assert int((g["filters"] != 0).sum()) == 396
    assert [int(i) for i in np.where(~g["filters"].any(axis=1))[0]] == [0, 3, 6, 13]


@pytest.mark.parametrize("seed,n", CASES)
def test_numpy_oracle_matches_reference_golden(seed, n):
    g = np.load(GOLD)
    pcm = synth.synth_audio(seed, n)
    want = g[f"mel_{seed}_{n}"]
    got = MO.log_mel_numpy(pcm, g["filters"])
    assert got.shape == want.shape == (128, n // 160)
    assert np.abs(got - want).max() <= 1e-5      # tests/test_mel.cpp:9 default tolerance


@pytest.mark.skipif(not os.path.exists(MO.REF_SO), reason="oracle/_ref not built (needs /root/reference)")
def test_numpy_oracle_matches_live_reference():
    ref = MO.RefMel()
    filt = ref.filters()
    for seed, n in ((7, 8000), (8, 16000 + 159), (9, 161), (10, 400)):
        pcm = synth.synth_audio(seed, n)
        assert np.abs(MO.log_mel_numpy(pcm, filt) - ref.log_mel(pcm, filt)).max() <= 1e-5


def test_edge_lengths():
    filt = MO.mel_filters_numpy()
    assert MO.log_mel_numpy(np.zeros(0, np.float32), filt).shape == (128, 0)
    assert MO.log_mel_numpy(np.zeros(159, np.float32), filt).shape == (128, 0)
    m = MO.log_mel_numpy(np.zeros(1600, np.float32), filt)          # silence: log10(1e-10) = -10 everywhere
    assert m.shape == (128, 10) and np.allclose(m, (-10.0 + 4.0) / 4.0)
