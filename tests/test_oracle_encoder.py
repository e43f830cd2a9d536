"""The encoder restatement against the fixture-free facts the reference's tests state, its own
golden vectors (drift guard) and an f64 'ideal' evaluation that bounds ggml's rounding choices."""
import os

import numpy as np

from oracle import encoder_oracle as EO, mel_oracle as MO
from q3a_b200 import synth
from conftest import compare

GOLD = os.path.join(os.path.dirname(__file__), "golden", "encoder_tiny.npz")


def test_conv2d_known_answer():
    """reference tests/test_conv2d.cpp:77-100: 4x4 ramp 1..16 (*) 3x3 kernel 1..9, stride 1, pad 1 -> out[0,0] = 111.
    The oracle's stride-2 conv shares im2col with it: out_s2[0,0] == out_s1[0,0]."""
    x = np.arange(1, 17, dtype=np.float32).reshape(1, 1, 4, 4)
    w = np.arange(1, 10, dtype=np.float32).reshape(1, 1, 3, 3)
    y = EO._conv2d_s2(x, w, "f32", EO.Numerics(act_round="none", gelu_table=False))
    assert y.shape == (1, 1, 2, 2)
    assert y[0, 0, 0, 0] == 111.0
    # hand-computed second output: window rows 0..2 (one padded), cols 1..3
    xp = np.pad(x[0, 0], 1)
    assert y[0, 0, 0, 1] == float((xp[0:3, 2:5] * w[0, 0]).sum())


def test_kernel_element_order():
    """reference tests/test_kernel_load.cpp:32-33: element (kw, kh, oc) lives at kw + kh*3 + oc*9,
    i.e. PyTorch [OC][IC][KH][KW] memory order with KH on the mel axis."""
    w = np.zeros((2, 1, 3, 3), np.float32)
    w.reshape(-1)[1 + 2 * 3 + 1 * 9] = 1.0           # kw=1, kh=2, oc=1
    assert w[1, 0, 2, 1] == 1.0
    x = np.zeros((1, 1, 8, 8), np.float32)
    x[0, 0, 3, 2] = 5.0                               # H (mel) = 3, W (time) = 2
    y = EO._conv2d_s2(x, w, "f32", EO.Numerics(act_round="none", gelu_table=False))
    # output (oh, ow) sees input (2*oh + kh - 1, 2*ow + kw - 1) = (3, 2) for oh=1, ow=1
    assert y[0, 1, 1, 1] == 5.0 and np.count_nonzero(y) == 1


def test_token_counts():
    # 30 s -> 390, 92 s -> 1196, 120 s -> 1560, 5 s -> 65
    assert [synth.chunk_tokens(n) for n in (3000, 9200, 12000, 500)] == [390, 1196, 1560, 65]
    assert EO.chunk_out_len(100) == 13 and EO.chunk_out_len(1) == 1 and EO.chunk_out_len(37) == 5
    assert EO.window_segments(250, 104) == [(0, 104), (104, 208), (208, 250)]


def test_sinusoidal_pe():
    pe = EO.sinusoidal_pe(13, 896)
    assert pe.shape == (13, 896)
    assert np.all(pe[0, :448] == 0) and np.all(pe[0, 448:] == 1)
    assert abs(pe[1, 0] - np.sin(1.0)) < 1e-6 and abs(pe[1, 447] - np.sin(1e-4)) < 1e-7


def _tiny(ftype):
    shape = synth.tiny_shape()
    w = synth.random_weights(shape, 7)
    stored = {k: v for k, v in synth.stored_weights(w, ftype).items() if "#" not in k}
    pcm = synth.synth_audio(104, 16000 * 2 + 4321)
    filt = np.load(os.path.join(os.path.dirname(GOLD), "mel_ref.npz"))["filters"]   # the reference's bank
    mel = MO.log_mel_numpy(pcm, filt)
    return shape, EO.Weights(stored, ftype), mel


def test_golden_drift_guard():
    g = np.load(GOLD)
    for ftype in ("f16", "q8_0"):
        shape, W, mel = _tiny(ftype)
        for variant in ("asr", "aligner"):
            y = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant=variant)
            want = g[f"{ftype}_{variant}"]
            assert y.shape == want.shape == (synth.chunk_tokens(mel.shape[1]), shape.out_dim)
            assert np.abs(y - want).max() <= 2e-4 * np.abs(want).max()     # BLAS summation order may differ


def test_ggml_rounding_is_small_against_f64_ideal():
    shape, W, mel = _tiny("f16")
    y = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant="asr")
    ideal = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant="asr",
                      nm=EO.IDEAL)
    cos, rel = compare(y, ideal)
    assert cos > 0.99999 and rel < 5e-3


def test_variants_differ_only_where_the_reference_differs():
    """ASR and aligner stems agree on full chunks and differ on a ragged tail."""
    shape, W, mel = _tiny("f16")
    nm = EO.Numerics()
    full = mel[:, :200]
    a = EO.stem_asr(full, W, shape.d_model, nm)
    b = EO.stem_aligner(full, W, shape.d_model, nm)
    assert np.array_equal(a, b)
    ragged = mel[:, :237]
    a = EO.stem_asr(ragged, W, shape.d_model, nm)
    b = EO.stem_aligner(ragged, W, shape.d_model, nm)
    assert a.shape == b.shape and np.array_equal(a[:26], b[:26]) and not np.array_equal(a[26:], b[26:])


HF_GOLD = os.path.join(os.path.dirname(__file__), "golden", "hf_audio_tower_tiny.npz")


def test_graph_matches_upstream_hf_audio_tower():
    """tests/golden/hf_audio_tower_tiny.npz holds outputs of transformers' Qwen3OmniMoeAudioEncoder -- the class whose
    parameters the reference's converter maps onto the GGUF tensors (convert_hf_to_gguf.py:48-103) -- on the tiny tower
    in float64 (make_hf_golden.py).  The restatement in its ideal mode, with the exact erf GELU the HF code uses, must
    reproduce them: this pins weight layouts, feature order, chunk padding, per-chunk positions, windows and layer
    order to an implementation that is not ours.  ggml's tanh-form GELU then moves the result by ~3e-4."""
    g = np.load(HF_GOLD)
    shape = synth.tiny_shape()
    w = synth.random_weights(shape, 7)
    W = EO.Weights({k: v for k, v in synth.stored_weights(w, "f32").items() if "#" not in k}, "f32")
    filt = np.load(os.path.join(os.path.dirname(GOLD), "mel_ref.npz"))["filters"]
    erf = EO.Numerics(act_round="none", gelu_table=False, gelu_erf=True, dtype=np.float64)
    for key in [k for k in g.files if k.startswith("win_")]:
        _, seed, n = key.split("_")
        mel = MO.log_mel_numpy(synth.synth_audio(int(seed), int(n)), filt)
        kw = dict(n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads)
        # flash-attention reading: 104-token windows = ForcedAligner::encode_audio
        win = EO.encode(mel, W, variant="aligner", nm=erf, **kw)
        assert win.shape == g[key].shape
        assert np.abs(win - g[key]).max() <= 2e-6 * np.abs(g[key]).max()
        assert [int(c) for c in g[f"cu_{seed}_{n}"]] == [0] + [b for _, b in EO.window_segments(win.shape[0], 104)]
        # eager reading: padded stem, every token attends to every token
        h = EO.stem_aligner(mel, W, shape.d_model, erf)
        full = EO.transformer(h, W, shape.n_layers, shape.n_heads, 1e-5, EO.window_segments(h.shape[0], 0), erf)
        want = g[f"full_{seed}_{n}"]
        assert np.abs(full - want).max() <= 2e-6 * np.abs(want).max()
        if mel.shape[1] % 100 == 0:                      # no ragged tail: AudioEncoder::encode is that same graph
            asr = EO.encode(mel, W, variant="asr", nm=erf, **kw)
            assert np.abs(asr - want).max() <= 2e-6 * np.abs(want).max()
        # the price of ggml's tanh-form GELU (audio_encoder.cpp:112 -> ggml_gelu) against the upstream erf form
        tanh = EO.encode(mel, W, variant="aligner", nm=EO.IDEAL, **kw)
        assert np.abs(tanh - g[key]).max() <= 1e-3 * np.abs(g[key]).max()


HF_WIDE = os.path.join(os.path.dirname(__file__), "golden", "hf_audio_tower_wide.npz")
WIDE = {"asr": (synth.EncoderShape("asr-2l", 2, 896, 14, 3584), 21),
        "aligner": (synth.EncoderShape("aligner-2l", 2, 1024, 16, 4096), 22)}


def test_graph_matches_upstream_hf_audio_tower_at_real_widths():
    """Same pin at the production widths (2-layer towers, d = 896 / 14 heads / ffn 3584 and d = 1024 / 16 heads / ffn 4096,
    conv 480; make_hf_golden.py WIDE): the restatement in ideal mode reproduces upstream's float64 outputs."""
    g = np.load(HF_WIDE)
    filt = np.load(os.path.join(os.path.dirname(GOLD), "mel_ref.npz"))["filters"]
    erf = EO.Numerics(act_round="none", gelu_table=False, gelu_erf=True, dtype=np.float64)
    for name, (shape, wseed) in WIDE.items():
        w = synth.random_weights(shape, wseed)
        W = EO.Weights({k: v for k, v in synth.stored_weights(w, "f32").items() if "#" not in k}, "f32")
        kw = dict(n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads)
        for key in [k for k in g.files if k.startswith(name + ".win_")]:
            _, seed, n = key.split("_")
            mel = MO.log_mel_numpy(synth.synth_audio(int(seed), int(n)), filt)
            win = EO.encode(mel, W, variant="aligner", nm=erf, **kw)
            assert win.shape == g[key].shape
            assert np.abs(win - g[key]).max() <= 2e-6 * np.abs(g[key]).max()
            if mel.shape[1] % 100 == 0:
                asr = EO.encode(mel, W, variant="asr", nm=erf, **kw)
                want = g[f"{name}.full_{seed}_{n}"]
                assert np.abs(asr - want).max() <= 2e-6 * np.abs(want).max()
