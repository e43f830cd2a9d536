"""The C++ shim headers (include/q3a/*.h) compile with plain g++ -std=c++17 against libq3a.so and behave like
the reference's entry points: CPU part checks the build and the loud failure without a GPU; GPU part runs the
harness the way Qwen3ASR::transcribe_internal would and compares with the oracle."""
import os
import subprocess

import numpy as np
import pytest

from conftest import needs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "qwen3-asr.cpp_b200")


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("shim") / "shim_harness")
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "shim_harness.cpp"), "-o", out, "-L", PKG, "-lq3a", f"-Wl,-rpath,{PKG}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "warning" not in r.stderr, r.stderr
    return out


def test_shim_compiles_and_fails_loudly_without_gpu(lib, harness, tmp_path):
    from q3a_b200 import synth
    pcm = synth.synth_audio(5, 16000)
    p = tmp_path / "a.f32"
    pcm.tofile(p)
    r = subprocess.run([harness, str(p), "-", str(tmp_path / "mel.f32"), str(tmp_path / "feat.f32")],
                       capture_output=True, text=True)
    if lib.q3a_device_count() == 0:
        assert r.returncode == 2 and "no CUDA device" in r.stderr       # no CPU fallback
    else:
        assert r.returncode == 0, r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["asr", "aligner"])
def test_shim_matches_oracle(lib, harness, tmp_path, mode):
    from q3a_b200 import synth
    from oracle import encoder_oracle as EO, mel_oracle as MO
    shape = synth.tiny_shape()
    w = synth.random_weights(shape, 11)
    gguf = synth.write_gguf(str(tmp_path / "tiny.gguf"), shape, w, "f16")
    pcm = synth.synth_audio(6, 16000 * 3 + 1234)
    p = tmp_path / "a.f32"
    pcm.tofile(p)
    r = subprocess.run([harness, str(p), gguf, str(tmp_path / "mel.f32"), str(tmp_path / "feat.f32"), mode],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "Mel bins mismatch: expected 128, got 64" in r.stdout        # audio_encoder.cpp:321-325
    filt = MO.mel_filters_numpy()
    want_mel = MO.log_mel_numpy(pcm, filt)
    mel = np.fromfile(tmp_path / "mel.f32", dtype=np.float32).reshape(want_mel.shape)
    assert np.abs(mel - want_mel).max() <= 1e-4
    W = EO.Weights(synth.stored_weights(w, "f16"), "f16")
    want = EO.encode(want_mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads,
                     variant="asr" if mode == "asr" else "aligner")
    got = np.fromfile(tmp_path / "feat.f32", dtype=np.float32).reshape(want.shape)
    cos = float((got * want).sum() / (np.linalg.norm(got) * np.linalg.norm(want)))
    rel = float(np.abs(got - want).max() / np.abs(want).max())
    assert cos >= 0.999 and rel <= 1e-2, (cos, rel)


@pytest.mark.gpu
@needs("decoder")
def test_shim_decoder_consumes_encoder_features(lib, harness, tmp_path):
    """mel -> encode_aligner -> forward_decoder through the C++ shims, the order ForcedAligner::align runs them
    (forced_aligner.cpp:1653-1690), against the two CPU restatements chained the same way."""
    from q3a_b200 import synth
    from oracle import decoder_oracle as DO, encoder_oracle as EO, mel_oracle as MO
    dshape = synth.tiny_decoder_shape()
    dw = synth.random_decoder_weights(dshape, 3)
    dgguf = synth.write_decoder_gguf(str(tmp_path / "dec.gguf"), dshape, dw, "f32")
    shape = synth.tiny_shape(out_dim=dshape.hidden)
    w = synth.random_weights(shape, 11)
    gguf = synth.write_gguf(str(tmp_path / "tiny.gguf"), shape, w, "f16")
    pcm = synth.synth_audio(6, 16000 * 3 + 1234)
    p = tmp_path / "a.f32"
    pcm.tofile(p)
    r = subprocess.run([harness, str(p), gguf, str(tmp_path / "mel.f32"), str(tmp_path / "feat.f32"), "aligner", dgguf,
                        str(tmp_path / "logits.f32")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    W = EO.Weights(synth.stored_weights(w, "f16"), "f16")
    feats = EO.encode(MO.log_mel_numpy(pcm, MO.mel_filters_numpy()), W, n_layers=shape.n_layers, d_model=shape.d_model,
                      n_heads=shape.n_heads, variant="aligner")
    n = feats.shape[0] + 12
    tokens = np.array([(i * 7 + 3) % dshape.vocab for i in range(n)], dtype=np.int32)
    got = np.fromfile(tmp_path / "logits.f32", dtype=np.float32)

    def compare(want):
        g = got.reshape(want.shape)
        return (float((g * want).sum() / (np.linalg.norm(g) * np.linalg.norm(want))),
                float(np.abs(g - want).max() / np.abs(want).max()))
    # (1) the decoder stage on its own: the oracle fed the features the encoder shim actually produced -- the per-stage
    #     bf16 bound (1e-2 of the peak), and the proof that forward_decoder injected those features at audio_start
    got_feats = np.fromfile(tmp_path / "feat.f32", dtype=np.float32).reshape(feats.shape)
    cos, rel = compare(DO.forward(tokens, dw, dshape, got_feats.astype(np.float64), 1))
    assert cos >= 0.9999 and rel <= 1e-2, (cos, rel)
    # (2) the whole chain against the two f64 restatements chained: two bf16 towers in series, each within 1e-2 of its own
    #     oracle (this file and test_decoder_gpu.py), so the chain is held to the sum of the two bounds
    cos, rel = compare(DO.forward(tokens, dw, dshape, feats, 1))
    assert cos >= 0.9999 and rel <= 2e-2, (cos, rel)


@pytest.fixture(scope="module")
def pool_harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("pool") / "pool_harness")
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "pool_harness.cpp"), "-o", out, "-L", PKG, "-lq3a", f"-Wl,-rpath,{PKG}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0 and "warning" not in r.stderr, r.stderr
    return out


def test_pool_harness_compiles_and_fails_loudly_without_gpu(lib, pool_harness, tmp_path):
    if lib.q3a_device_count() == 0:
        r = subprocess.run([pool_harness, "none.gguf", "2", str(tmp_path / "o.f32")], capture_output=True, text=True)
        assert r.returncode == 2 and "no CUDA device" in r.stderr


@pytest.mark.gpu
@needs("pool")
@needs("async_call")
def test_cpp_caller_uses_the_pool_and_the_async_call(lib, pool_harness, tmp_path):
    """tests/pool_harness.cpp: a single-process C++ caller (the reference's shape, main.cpp:294-504) drives three
    replicas through q3a_pool_encode_pcm with pinned memory from q3a_host_alloc, then the same clips through two
    asynchronous calls in flight on one encoder; both must agree bit for bit, and with the Python-side single encoder."""
    from q3a_b200 import binding as B, synth
    shape = synth.tiny_shape()
    gguf = synth.write_gguf(str(tmp_path / "tiny.gguf"), shape, synth.random_weights(shape, 11), "f16")
    outp = tmp_path / "feat.f32"
    r = subprocess.run([pool_harness, gguf, "3", str(outp)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "pool == async" in r.stdout
    got = np.fromfile(outp, dtype=np.float32).reshape(-1, shape.out_dim)
    assert got.shape[0] > 300 and np.isfinite(got).all()
