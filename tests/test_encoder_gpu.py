"""CUDA encoder forward against the CPU oracle (oracle/encoder_oracle.py: restatement of the ggml graph with
ggml's CPU rounding).  Tolerance (BASELINE.json north_star): cosine >= 0.999 and
max|a-b| / max|b| <= 1e-2 per utterance."""
import os

import numpy as np
import pytest

from oracle import encoder_oracle as EO, mel_oracle as MO
from q3a_b200 import binding as B, synth
from conftest import compare, needs

pytestmark = pytest.mark.gpu
COS, REL = 0.999, 1e-2
GOLD = os.path.join(os.path.dirname(__file__), "golden")
LENS = (16000 * 3 + 777, 16000 + 5, 100 * 160, 237 * 160, 160, 37 * 160 + 3)


def _oracle(mel, W, shape, variant):
    return EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant=variant)


def _mels(filters, lens=LENS):
    return [MO.log_mel_numpy(synth.synth_audio(11 + i, n), filters) for i, n in enumerate(lens)]


@pytest.mark.parametrize("ftype", ["f32", "f16"])
@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
@pytest.mark.parametrize("mode,variant", [(B.MODE_ASR, "asr"), (B.MODE_ALIGNER, "aligner")])
def test_tiny_tower_matches_oracle(lib, tiny_models, filters, ftype, opd, mode, variant):
    shape, models = tiny_models
    path, W = models[ftype]
    enc = B.Encoder(path, 0, opd)
    assert (enc.hp.n_encoder_layers, enc.hp.d_model, enc.hp.n_attention_heads, enc.hp.ffn_dim, enc.hp.conv_channels,
            enc.hp.out_dim) == (shape.n_layers, shape.d_model, shape.n_heads, shape.ffn_dim, shape.conv_ch, shape.out_dim)
    mels = _mels(filters)
    outs = enc.encode_batch(mels, mode)                       # ragged batch, packed
    for mel, got in zip(mels, outs):
        want = _oracle(mel, W, shape, variant)
        assert got.shape == want.shape == (synth.chunk_tokens(mel.shape[1]), shape.out_dim)
        cos, rel = compare(got, want)
        assert cos >= COS and rel <= REL, (cos, rel)
    enc.close()


@pytest.mark.parametrize("mode,variant", [(B.MODE_ASR, "asr"), (B.MODE_ALIGNER, "aligner")])
def test_ten_minute_clip(lib, tiny_models, filters, mode, variant):
    """One 10-minute utterance (600 chunks + a 37-frame tail, 7805 tokens): in ASR mode a single full-attention
    sequence of 61 query tiles x 122 key blocks per head (the K/V ring and the lazy rescale wrap many times), in aligner
    mode 76 windows of 104 tokens.  The reference has no length limit on this path (audio_encoder.cpp:312-354)."""
    shape, models = tiny_models
    path, W = models["f16"]
    enc = B.Encoder(path, 0, B.OP_F16)
    fe = B.Frontend(0)
    pcm = synth.synth_audio(77, 16000 * 600 + 37 * 160 + 11)
    mel = MO.log_mel_numpy(pcm, filters)
    want = _oracle(mel, W, shape, variant)
    got = enc.encode_batch([mel], mode)[0]
    assert got.shape == want.shape == (synth.chunk_tokens(mel.shape[1]), shape.out_dim)
    cos, rel = compare(got, want)
    assert cos >= COS and rel <= REL, (cos, rel)
    fused = enc.encode_pcm_batch(fe, [pcm], mode)[0]           # PCM in: mel on the device, clamp fused into conv1
    cos, rel = compare(fused, want)
    assert cos >= COS and rel <= REL, (cos, rel)
    enc.close(); fe.close()


def test_tiny_q8_file_matches_oracle(lib, tiny_models, filters):
    """Q8_0 file: conv kernels stay F16, Linear weights are Q8_0 blocks (convert_hf_to_gguf.py:303-308).  Held at the
    stated tolerance against the oracle reading that matches what the product computes (ggml's Q8_0 weight values,
    f16-rounded activations); ggml's additional Q8_0 *activation* quantisation is a separate, documented term
    (test_q8_file_three_way_at_real_width)."""
    shape, models = tiny_models
    path, W = models["q8_0"]
    enc = B.Encoder(path, 0, B.OP_F16)
    assert enc.hp.weight_ftype == 8
    for mel in _mels(filters)[:3]:
        want = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant="asr",
                         nm=EO.Numerics(act_round="f16"))
        cos, rel = compare(enc.encode_batch([mel], B.MODE_ASR)[0], want)
        assert cos >= COS and rel <= REL, (cos, rel)
    enc.close()


def test_golden_vectors(lib, tiny_models):
    shape, models = tiny_models
    g = np.load(os.path.join(GOLD, "encoder_tiny.npz"))
    filt = np.load(os.path.join(GOLD, "mel_ref.npz"))["filters"]
    mel = MO.log_mel_numpy(synth.synth_audio(104, 16000 * 2 + 4321), filt)
    enc = B.Encoder(models["f16"][0], 0, B.OP_F16)
    for mode, variant in ((B.MODE_ASR, "asr"), (B.MODE_ALIGNER, "aligner")):
        cos, rel = compare(enc.encode_batch([mel], mode)[0], g[f"f16_{variant}"])
        assert cos >= COS and rel <= REL, (variant, cos, rel)
    enc.close()


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
def test_matches_upstream_hf_audio_tower(lib, tiny_models, opd):
    """The CUDA path against outputs of the upstream implementation itself (transformers' Qwen3OmniMoeAudioEncoder in
    float64, tests/golden/make_hf_golden.py): aligner mode = its 104-token-window reading, ASR mode = its eager
    full-attention reading (identical graphs when the clip has no ragged tail chunk).  Same tolerance as against the
    oracle; the upstream erf GELU vs ggml's tanh form accounts for ~3e-4 of it."""
    g = np.load(os.path.join(GOLD, "hf_audio_tower_tiny.npz"))
    shape, models = tiny_models
    path, _ = models["f32"]
    enc = B.Encoder(path, 0, opd)
    filt = np.load(os.path.join(GOLD, "mel_ref.npz"))["filters"]
    for key in [k for k in g.files if k.startswith("win_")]:
        _, seed, n = key.split("_")
        mel = MO.log_mel_numpy(synth.synth_audio(int(seed), int(n)), filt)
        got = enc.encode_batch([mel], B.MODE_ALIGNER)[0]
        cos, rel = compare(got, g[key])
        assert got.shape == g[key].shape and cos >= COS and rel <= REL, (key, cos, rel)
        if mel.shape[1] % 100 == 0:
            got = enc.encode_batch([mel], B.MODE_ASR)[0]
            cos, rel = compare(got, g[f"full_{seed}_{n}"])
            assert cos >= COS and rel <= REL, (key, "full", cos, rel)
    enc.close()


def test_debug_variants(lib, tiny_models, filters):
    """encode_conv_only (stem, no chunking, no PE: audio_encoder.cpp:854-913) and encode_no_chunk (:603-852)"""
    shape, models = tiny_models
    path, W = models["f16"]
    enc = B.Encoder(path, 0, B.OP_F16)
    nm = EO.Numerics()
    for mel in _mels(filters, (16000 * 2 + 999, 16000)):
        got = enc.encode_batch([mel], B.MODE_CONV_ONLY)[0]
        want = EO.stem(mel[None], W, nm)[0]
        assert got.shape == want.shape
        cos, rel = compare(got, want)
        assert cos >= COS and rel <= REL, (cos, rel)
        got = enc.encode_batch([mel], B.MODE_NO_CHUNK)[0]
        h = want + EO.sinusoidal_pe(want.shape[0], shape.d_model)
        want2 = EO.transformer(h, W, shape.n_layers, shape.n_heads, 1e-5, [(0, h.shape[0])], nm)
        cos, rel = compare(got, want2)
        assert cos >= COS and rel <= REL, (cos, rel)
    enc.close()


def test_empty_and_degenerate_inputs(lib, tiny_models):
    shape, models = tiny_models
    enc = B.Encoder(models["f16"][0], 0, B.OP_BF16)
    assert enc.encode_batch([], B.MODE_ASR) == []
    outs = enc.encode_batch([np.zeros((128, 0), np.float32), np.zeros((128, 1), np.float32)], B.MODE_ASR)
    assert outs[0].shape == (0, shape.out_dim) and outs[1].shape == (1, shape.out_dim) and np.isfinite(outs[1]).all()
    with pytest.raises(B.Q3AError, match="Mel bins mismatch: expected 128, got 80"):    # audio_encoder.cpp:321-325
        enc.encode_batch([np.zeros((80, 100), np.float32)], B.MODE_ASR)
    enc.close()


@pytest.fixture(scope="module")
def asr_model(tmp_path_factory):
    d = tmp_path_factory.mktemp("asr")
    shape = synth.ASR_06B
    w = synth.random_weights(shape, 0)
    paths = {ft: synth.write_gguf(str(d / f"asr_{ft}.gguf"), shape, w, ft) for ft in ("f16", "q8_0")}
    Ws = {ft: EO.Weights({k: v for k, v in synth.stored_weights(w, ft).items() if "#" not in k}, ft) for ft in paths}
    return shape, paths, Ws


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
def test_asr_06b_30s_clip_matches_oracle(lib, asr_model, filters, opd):
    """BASELINE config 1: one 30-s clip, F16 GGUF, full attention -> [390][1024]"""
    shape, paths, Ws = asr_model
    enc = B.Encoder(paths["f16"], 0, opd)
    mel = MO.log_mel_numpy(synth.synth_audio(100, 480000), filters)
    got = enc.encode_batch([mel], B.MODE_ASR)[0]
    want = _oracle(mel, Ws["f16"], shape, "asr")
    assert got.shape == want.shape == (390, 1024)
    cos, rel = compare(got, want)
    assert cos >= COS and rel <= REL, (cos, rel)
    enc.close()


def test_asr_06b_92s_clip_full_attention(lib, asr_model, filters):
    """BASELINE config 3's length: one 92-s clip = 1196 tokens attended as ONE sequence, the way the reference runs
    it (audio_encoder.cpp:466-486 has no mask): ten query tiles, nineteen key blocks with the lazily rescaled running
    maximum; and the same audio cut into 30 + 30 + 30 + 2 s utterances (the north star's "chunked" reading), each equal
    to that piece encoded on its own."""
    shape, paths, Ws = asr_model
    enc = B.Encoder(paths["f16"], 0, B.OP_BF16)
    fe = B.Frontend(0)
    pcm = synth.synth_audio(301, 16000 * 92)
    mel = MO.log_mel_numpy(pcm, filters)
    got = enc.encode_batch([mel], B.MODE_ASR)[0]
    want = _oracle(mel, Ws["f16"], shape, "asr")
    assert got.shape == want.shape == (1196, 1024)
    cos, rel = compare(got, want)
    assert cos >= COS and rel <= REL, (cos, rel)
    pieces = [pcm[:480000], pcm[480000:960000], pcm[960000:1440000], pcm[1440000:]]
    outs = enc.encode_pcm_batch(fe, pieces, B.MODE_ASR)
    assert [o.shape[0] for o in outs] == [390, 390, 390, 26]
    for piece, o in zip(pieces, outs):
        assert np.array_equal(o, enc.encode_pcm_batch(fe, [piece], B.MODE_ASR)[0])
    enc.close(); fe.close()


def test_pcm_path_batch_invariance_and_full_size(lib, asr_model, filters):
    """BASELINE config 2 shape (64 x 30 s, bf16): the fused PCM->features call; a clip's features do not depend on
    what else is in the batch (bit-identical), and agree with the oracle on a spot-checked clip."""
    shape, paths, Ws = asr_model
    enc = B.Encoder(paths["f16"], 0, B.OP_BF16)
    fe = B.Frontend(0)
    clips = [synth.synth_audio(200 + i, 480000) for i in range(64)]
    outs = enc.encode_pcm_batch(fe, clips, B.MODE_ASR)
    assert all(o.shape == (390, 1024) and np.isfinite(o).all() for o in outs)
    alone = enc.encode_pcm_batch(fe, clips[5:6], B.MODE_ASR)[0]
    assert np.array_equal(alone, outs[5])
    pair = enc.encode_pcm_batch(fe, [clips[63], clips[5]], B.MODE_ASR)
    assert np.array_equal(pair[1], outs[5]) and np.array_equal(pair[0], outs[63])
    want = _oracle(MO.log_mel_numpy(clips[5], filters), Ws["f16"], shape, "asr")
    cos, rel = compare(outs[5], want)
    assert cos >= COS and rel <= REL, (cos, rel)
    tm = enc.timings()
    assert tm.total_launches > 90          # 116 by default, 98 with every LayerNorm folded, 135 with none
    enc.close(); fe.close()


def test_copy_overlap_partitions_do_not_change_results(lib, asr_model, monkeypatch):
    """The host-pointer call overlaps PCIe copies with kernels by running the stem per upload part and the last
    layers per download part (and, when forced, by cutting the call into two passes).  Ragged 2-40 s clips, all
    three schedules: every clip's features are bit-identical."""
    shape, paths, Ws = asr_model
    enc = B.Encoder(paths["f16"], 0, B.OP_BF16)
    fe = B.Frontend(0)
    rng = np.random.default_rng(7)
    clips = [synth.synth_audio(700 + i, int(rng.integers(2 * 16000, 40 * 16000))) for i in range(48)]
    clips[17] = clips[17][:0]                                 # an empty clip in the middle of a part
    parts = enc.encode_pcm_batch(fe, clips, B.MODE_ASR)
    monkeypatch.setenv("Q3A_NO_PART_OVERLAP", "1")
    plain = enc.encode_pcm_batch(fe, clips, B.MODE_ASR)
    monkeypatch.delenv("Q3A_NO_PART_OVERLAP")
    for a, b, c in zip(parts, plain, clips):
        assert a.shape == (synth.chunk_tokens(len(c) // 160), 1024)
        assert np.array_equal(a, b)
    mels = fe.mel_batch(clips)
    via_mel = enc.encode_batch(mels, B.MODE_ASR)               # host mel in, same partitioning machinery
    for a, b in zip(parts, via_mel):
        cos, rel = compare(a, b) if a.size else (1.0, 0.0)
        assert cos >= 0.99999 and rel <= 2e-3                   # mel finalised on the host path vs clamp fused in conv1
    enc.close(); fe.close()


def test_no_kernel_writes_outside_its_buffers(lib, tiny_models, asr_model):
    """The library's own bounds check (compute-sanitizer is not available on this pool): in a child process with
    Q3A_GUARD=1 every device buffer is exact-size between two pattern bands; tests/guard_battery.py drives every
    entry point over ragged / empty / multi-part inputs and no band may change."""
    import json, subprocess, sys
    _, models = tiny_models
    _, paths, _ = asr_model
    env = dict(os.environ, Q3A_GUARD="1")
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "guard_battery.py"),
                        models["f16"][0], models["q8_0"][0], paths["f16"]], env=env, capture_output=True, text=True, timeout=900)
    assert r.stdout.strip(), r.stderr[-2000:]
    rep = json.loads(r.stdout.strip().splitlines()[-1])
    assert rep["damaged"] == 0, rep["msg"]
    assert r.returncode == 0, r.stderr[-2000:]
    assert rep["bands"] >= 100 and rep["cases"] >= 50, rep
    assert rep["selftest_damaged"] == 1, rep                   # the detector fires on a deliberate one-byte overrun


def test_int16_ingest_matches_load_wav_conversion(lib, tiny_models):
    """q3a_encode_pcm16_batch = load_wav's int16 -> float conversion (mel_spectrogram.cpp:195-210) + the float path:
    mono and interleaved stereo, bit-identical to converting on the host the way the reference does."""
    shape, models = tiny_models
    path, _ = models["f16"]
    enc = B.Encoder(path, 0, B.OP_F16)
    fe = B.Frontend(0)
    rng = np.random.default_rng(3)
    lens = [16000 * 3 + 77, 160 * 7, 0, 16000 + 1, 159]
    mono16 = [np.round(synth.synth_audio(900 + i, n) * 32767).astype(np.int16) for i, n in enumerate(lens)]
    want = enc.encode_pcm_batch(fe, [c.astype(np.float32) / np.float32(32768.0) for c in mono16], B.MODE_ASR)
    got = enc.encode_pcm16_batch(fe, mono16, 1, B.MODE_ASR)
    for a, b in zip(got, want):
        assert a.shape == b.shape and np.array_equal(a, b)
    stereo16 = [np.stack([c, rng.integers(-3000, 3000, len(c)).astype(np.int16)], axis=1) for c in mono16]
    host = []
    for c in stereo16:                                        # (sum / num_channels) / 32768.0f in float, as load_wav does
        ssum = c[:, 0].astype(np.float32) + c[:, 1].astype(np.float32)
        host.append((ssum / np.float32(2)) / np.float32(32768.0))
    want = enc.encode_pcm_batch(fe, host, B.MODE_ASR)
    got = enc.encode_pcm16_batch(fe, [c.reshape(-1) for c in stereo16], 2, B.MODE_ASR)
    for a, b in zip(got, want):
        assert a.shape == b.shape and np.array_equal(a, b)
    enc.close(); fe.close()


def test_aligner_shape_variable_length_batch(lib, filters, tmp_path):
    """BASELINE config 4 (reduced: 4 layers of the aligner width, 5-40 s utterances): padded-tail stem, 104-token windows"""
    shape = synth.EncoderShape("aligner-4l", 4, 1024, 16, 4096)
    w = synth.random_weights(shape, 1)
    path = synth.write_gguf(str(tmp_path / "al.gguf"), shape, w, "f16")
    W = EO.Weights(synth.stored_weights(w, "f16"), "f16")
    enc = B.Encoder(path, 0, B.OP_BF16)
    rng = np.random.default_rng(400)
    lens = [int(rng.uniform(5, 40) * 16000) for _ in range(5)]
    mels = _mels(filters, lens)
    outs = enc.encode_batch(mels, B.MODE_ALIGNER)
    for mel, got in zip(mels, outs):
        cos, rel = compare(got, _oracle(mel, W, shape, "aligner"))
        assert cos >= COS and rel <= REL, (cos, rel)
    enc.close()


# ---- round 2: the parity holes VERDICT r1 names -----------------------------------------------------------------
WIDE = {"asr": (synth.EncoderShape("asr-2l", 2, 896, 14, 3584), 21),
        "aligner": (synth.EncoderShape("aligner-2l", 2, 1024, 16, 4096), 22)}


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
@pytest.mark.parametrize("tower", ["asr", "aligner"])
def test_matches_upstream_hf_audio_tower_at_real_widths(lib, tmp_path, opd, tower):
    """The production tile paths (N = 896 / 2688 -> BLOCK_N 224, N = 3584 / 1024 / 4096 / 3072 -> 256, conv N = 480 ->
    240, d = 896 and 1024, 14 and 16 heads) against float64 outputs of upstream's Qwen3OmniMoeAudioEncoder on 2-layer
    towers of those widths (tests/golden/make_hf_golden.py WIDE -> hf_audio_tower_wide.npz): not our restatement."""
    g = np.load(os.path.join(GOLD, "hf_audio_tower_wide.npz"))
    shape, wseed = WIDE[tower]
    path = synth.write_gguf(str(tmp_path / f"{tower}.gguf"), shape, synth.random_weights(shape, wseed), "f32")
    enc = B.Encoder(path, 0, opd)
    filt = np.load(os.path.join(GOLD, "mel_ref.npz"))["filters"]
    keys = [k for k in g.files if k.startswith(tower + ".win_")]
    assert len(keys) == 3
    for key in keys:
        _, seed, n = key.split("_")
        mel = MO.log_mel_numpy(synth.synth_audio(int(seed), int(n)), filt)
        got = enc.encode_batch([mel], B.MODE_ALIGNER)[0]
        cos, rel = compare(got, g[key])
        assert got.shape == g[key].shape and cos >= COS and rel <= REL, (key, cos, rel)
        if mel.shape[1] % 100 == 0:
            got = enc.encode_batch([mel], B.MODE_ASR)[0]
            cos, rel = compare(got, g[f"{tower}.full_{seed}_{n}"])
            assert cos >= COS and rel <= REL, (key, "full", cos, rel)
    enc.close()


def _pack(arrs, align):
    offs, cur = [], 0
    for a in arrs:
        offs.append(cur)
        cur += (a.size + align - 1) // align * align
    return offs, cur


def test_device_entry_points(lib, asr_model, filters):
    """q3a_encode_pcm_device and q3a_encoder_encode_device (finalised mel with max_dev = NULL, and raw log-mel + per-clip
    maxima) on a caller-supplied stream: rows land at tok_off, are bit-identical to the host-pointer twins (same
    kernels, batch-invariant) and agree with the oracle.  These are the calls bench.py's `value` is timed through and
    the hand-off a GPU-resident decoder would consume (text_decoder.cpp:422-456,650-655 reads the same [T][1024] rows)."""
    import torch
    shape, paths, Ws = asr_model
    enc = B.Encoder(paths["f16"], 0, B.OP_BF16)
    fe = B.Frontend(0)
    clips = [synth.synth_audio(800 + i, n) for i, n in enumerate((16000 * 7 + 333, 0, 16000 * 2, 16000 * 31 + 5999, 150))]
    ns = [len(c) for c in clips]
    offs, total = _pack(clips, 64)
    pcm = torch.zeros(total + 64, device="cuda")
    for c, o in zip(clips, offs):
        pcm[o:o + len(c)] = torch.from_numpy(c).cuda()
    T = sum(enc.n_tokens(n // 160) for n in ns)
    st = torch.cuda.Stream()
    host = enc.encode_pcm_batch(fe, clips, B.MODE_ASR)
    # (1) PCM on the device -> features on the device
    out = torch.full((T + 1, 1024), float("nan"), device="cuda")
    with torch.cuda.stream(st):
        pcm.record_stream(st)
        tok = enc.encode_pcm_device(fe, pcm.data_ptr(), offs, ns, out.data_ptr(), B.MODE_ASR, st.cuda_stream)
    st.synchronize()
    assert tok[-1] == T and [b - a for a, b in zip(tok, tok[1:])] == [h.shape[0] for h in host]
    assert torch.isnan(out[T]).all()                                        # nothing written past the last row
    got = out[:T].cpu().numpy()
    for i, h in enumerate(host):
        assert np.array_equal(got[tok[i]:tok[i + 1]], h), i
    want = _oracle(MO.log_mel_numpy(clips[0], filters), Ws["f16"], shape, "asr")
    cos, rel = compare(got[tok[0]:tok[1]], want)
    assert cos >= COS and rel <= REL, (cos, rel)
    # (2) mel on the device, both hand-offs from q3a_mel_batch_device
    nf = [n // 160 for n in ns]
    moffs, mtotal = _pack([np.empty(128 * f) for f in nf], 64)
    for finalize in (True, False):
        mel = torch.zeros(mtotal + 64, device="cuda")
        mx = torch.zeros(len(clips), device="cuda")
        out2 = torch.full((T + 1, 1024), float("nan"), device="cuda")
        with torch.cuda.stream(st):
            fe.mel_batch_device(pcm.data_ptr(), offs, ns, mel.data_ptr(), moffs, mx.data_ptr(), finalize=finalize,
                                stream=st.cuda_stream)
            tok2 = enc.encode_device(mel.data_ptr(), moffs, nf, out2.data_ptr(), B.MODE_ASR,
                                     max_ptr=0 if finalize else mx.data_ptr(), stream=st.cuda_stream)
        st.synchronize()
        assert tok2 == tok and torch.isnan(out2[T]).all()
        got2 = out2[:T].cpu().numpy()
        if not finalize:                      # raw log-mel + maxima: exactly what the fused PCM call does internally
            assert np.array_equal(got2, got)
        else:                                 # finalised mel: same as the host-mel entry point on the same mel
            via_host = enc.encode_batch(fe.mel_batch(clips), B.MODE_ASR)
            for i, h in enumerate(via_host):
                assert np.array_equal(got2[tok[i]:tok[i + 1]], h), i
    enc.close(); fe.close()


def test_plane_pad_rows_do_not_leak_between_calls(lib, tiny_models, filters):
    """conv2 never writes the last odd-time row of a chunk when its row count is even, and conv3 reads that
    row as zero padding; a previous call with an odd row count (same P3) used to leave data there.  96-frame clip, then
    a 92-frame clip on the same handle, against a fresh handle; and the NO_CHUNK (101..104 frames) -> ASR sequence."""
    shape, models = tiny_models
    path, _ = models["f16"]
    mel96, mel92, mel103, mel300 = (MO.log_mel_numpy(synth.synth_audio(950 + i, 160 * f), filters)
                                    for i, f in enumerate((96, 92, 103, 300)))
    for first, fmode, second, smode in ((mel96, B.MODE_ASR, mel92, B.MODE_ASR),
                                        (mel103, B.MODE_NO_CHUNK, mel300, B.MODE_ASR),
                                        (mel103, B.MODE_CONV_ONLY, mel96, B.MODE_ALIGNER)):
        fresh = B.Encoder(path, 0, B.OP_F16)
        want = fresh.encode_batch([second], smode)[0]
        fresh.close()
        used = B.Encoder(path, 0, B.OP_F16)
        used.encode_batch([first], fmode)
        got = used.encode_batch([second], smode)[0]
        used.close()
        assert np.array_equal(got, want)


def test_aligner_06b_full_shape_5_to_120_s(lib, filters, tmp_path):
    """BASELINE configs[3] at its full shape: the 24-layer d = 1024 forced-aligner tower (forced_aligner.h:36-48), one
    packed batch holding a 5-s, a 61.3-s and a 120-s utterance (1 / 8 / 15 windows of 104 tokens),
    (65 / 797 / 1560 tokens) each against the oracle's restatement of ForcedAligner::encode_audio (forced_aligner.cpp:591-924)."""
    shape = synth.ALIGNER_06B
    w = synth.random_weights(shape, 1)
    path = synth.write_gguf(str(tmp_path / "aligner06b.gguf"), shape, w, "f16")
    W = EO.Weights({k: v for k, v in synth.stored_weights(w, "f16").items() if "#" not in k}, "f16")
    enc = B.Encoder(path, 0, B.OP_BF16)
    assert (enc.hp.n_encoder_layers, enc.hp.d_model, enc.hp.n_attention_heads, enc.hp.ffn_dim) == (24, 1024, 16, 4096)
    fe = B.Frontend(0)
    clips = [synth.synth_audio(410 + i, n) for i, n in enumerate((16000 * 5, 16000 * 61 + 4800 + 77, 16000 * 120))]
    outs = enc.encode_pcm_batch(fe, clips, B.MODE_ALIGNER)
    assert [o.shape[0] for o in outs] == [synth.chunk_tokens(len(c) // 160) for c in clips] == [65, 797, 1560]
    for clip, got in zip(clips, outs):
        want = _oracle(MO.log_mel_numpy(clip, filters), W, shape, "aligner")
        cos, rel = compare(got, want)
        assert got.shape == want.shape and cos >= COS and rel <= REL, (len(clip), cos, rel)
    enc.close(); fe.close()


def test_q8_file_three_way_at_real_width(lib, asr_model, filters):
    """BASELINE configs[2]: Q8_0 GGUF, 92-s clip (1196 tokens, one full-attention sequence), at the stated tolerance.
    The product keeps ggml's Q8_0 WEIGHT quantisation exactly (block scale x int8) but feeds 16-bit activations to the
    tensor cores, where ggml's CPU kernel also quantises the ACTIVATIONS to Q8_0 (vec_dot_type).  Three readings:
      A  CUDA (Q8_0 file)
      B  oracle, dequantised Q8_0 weights, f16 activations      -> A vs B must meet cos >= 0.999, rel <= 1e-2
      C  oracle, dequantised Q8_0 weights, Q8_0 activations     -> C vs B is ggml's activation-quantisation noise,
                                                                   reported and bounded; A vs C is held at A-B + B-C"""
    shape, paths, Ws = asr_model
    mel = MO.log_mel_numpy(synth.synth_audio(300, 16000 * 92), filters)
    kw = dict(n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant="asr")
    ideal = EO.encode(mel, Ws["q8_0"], nm=EO.Numerics(act_round="f16"), **kw)          # B
    q8act = EO.encode(mel, Ws["q8_0"], nm=EO.Numerics(act_round="q8"), **kw)           # C
    for opd in (B.OP_F16, B.OP_BF16):
        enc = B.Encoder(paths["q8_0"], 0, opd)
        assert enc.hp.weight_ftype == 8
        got = enc.encode_batch([mel], B.MODE_ASR)[0]                                     # A
        enc.close()
        assert got.shape == ideal.shape == (1196, 1024)
        cos_ab, rel_ab = compare(got, ideal)
        cos_cb, rel_cb = compare(q8act, ideal)
        cos_ac, rel_ac = compare(got, q8act)
        print(f"q8 three-way opd={opd}: A-B cos {cos_ab:.6f} rel {rel_ab:.2e} | C-B (ggml act. quant.) cos {cos_cb:.6f} "
              f"rel {rel_cb:.2e} | A-C cos {cos_ac:.6f} rel {rel_ac:.2e}")
        assert cos_ab >= COS and rel_ab <= REL, (cos_ab, rel_ab)
        assert cos_cb >= COS and rel_cb <= 3e-2, (cos_cb, rel_cb)
        assert cos_ac >= COS and rel_ac <= rel_ab + rel_cb + 1e-3, (cos_ac, rel_ac)


@needs("q8_fused")
def test_q8_fused_dequant_is_bit_identical_to_dequant_at_load(lib, tiny_models, asr_model, filters, monkeypatch):
    """Row a12: with a Q8_0 file the Linear weights stay quantised in device memory (int8 plane + f16 block scales, 34 bytes
    per 32 weights as in the GGUF) and the GEMM's dequant warps expand them into the operand stage.  The value they
    produce is d * q rounded once to the operand type -- exactly what expanding at load time stores -- so the two paths
    must agree bit for bit (same LayerNorm kernel path on both sides); and the resident weights must be about half."""
    _, models = tiny_models
    shape, paths, _ = asr_model
    mels_tiny = _mels(filters)[:4]
    mel_big = MO.log_mel_numpy(synth.synth_audio(303, 16000 * 7 + 123), filters)
    for path, mels, opd in ((models["q8_0"][0], mels_tiny, B.OP_F16), (models["q8_0"][0], mels_tiny, B.OP_BF16),
                            (paths["q8_0"], [mel_big], B.OP_BF16)):
        monkeypatch.setenv("Q3A_Q8_FUSED", "1")
        fused = B.Encoder(path, 0, opd)
        monkeypatch.delenv("Q3A_Q8_FUSED")
        bytes_fused = fused.device_bytes()
        got = fused.encode_batch(mels, B.MODE_ASR)
        fused.close()
        monkeypatch.setenv("Q3A_Q8_DEQUANT_AT_LOAD", "1")
        monkeypatch.setenv("Q3A_NO_LN_FOLD", "1")
        plain = B.Encoder(path, 0, opd)
        monkeypatch.delenv("Q3A_Q8_DEQUANT_AT_LOAD"); monkeypatch.delenv("Q3A_NO_LN_FOLD")
        bytes_plain = plain.device_bytes()
        want = plain.encode_batch(mels, B.MODE_ASR)
        plain.close()
        for a, b in zip(got, want):
            assert a.shape == b.shape and np.array_equal(a, b)
        if path == paths["q8_0"]:
            print(f"ASR-0.6B weight arena: Q8_0 resident {bytes_fused / 1e6:.1f} MB vs expanded to 16 bit {bytes_plain / 1e6:.1f} MB")
            assert bytes_fused < 0.62 * bytes_plain


@needs("async_call")
def test_async_submit_wait_matches_synchronous_call(lib, asr_model):
    """q3a_encode_pcm_submit / q3a_encode_wait: four calls over different ragged batches with two in flight at any time
    (call k+1 is submitted before call k is waited for, so its upload and stem overlap call k's tail); every call's
    features equal the synchronous q3a_encode_pcm_batch result bit for bit, and an unknown ticket is refused."""
    import ctypes as C
    shape, paths, _ = asr_model
    enc = B.Encoder(paths["f16"], 0, B.OP_BF16)
    fe = B.Frontend(0)
    rng = np.random.default_rng(21)
    batches = [[synth.synth_audio(1200 + 10 * b + i, int(rng.integers(16000, 16000 * 25))) for i in range(5 + b)] for b in range(4)]
    batches[2][1] = batches[2][1][:0]
    want = [enc.encode_pcm_batch(fe, clips, B.MODE_ASR) for clips in batches]
    keep, tickets, outs_all = [], [], []
    for k, clips in enumerate(batches):
        n = [len(c) for c in clips]
        outs = [np.full((enc.n_tokens(x // 160), 1024), np.nan, np.float32) for x in n]
        ntok = (C.c_int * len(clips))()
        tk = C.c_int64(-1)
        args = (B._ptr_array(clips), B._int_array(n), B._ptr_array(outs))
        keep.append((args, ntok))                              # host arrays stay alive until the wait
        B.check(enc.lib.q3a_encode_pcm_submit(fe.h, enc.h, args[0], args[1], len(clips), B.MODE_ASR, args[2], ntok, C.byref(tk)))
        assert list(ntok) == [o.shape[0] for o in outs] and tk.value == (tickets[-1] + 1 if tickets else tk.value)
        tickets.append(tk.value); outs_all.append(outs)
        if k >= 1:
            B.check(enc.lib.q3a_encode_wait(enc.h, tickets[k - 1]))
            for a, b in zip(outs_all[k - 1], want[k - 1]):
                assert np.array_equal(a, b)
    B.check(enc.lib.q3a_encode_wait(enc.h, tickets[-1]))
    for a, b in zip(outs_all[-1], want[-1]):
        assert np.array_equal(a, b)
    assert enc.lib.q3a_encode_wait(enc.h, tickets[-1] + 5) != 0
    assert np.array_equal(enc.encode_pcm_batch(fe, batches[0], B.MODE_ASR)[1], want[0][1])     # the synchronous call still works after
    enc.close(); fe.close()


def test_contiguous_host_blocks_take_the_merged_copy_path(lib, tiny_models):
    """Clips that sit back to back in one host block (lengths multiples of 64 samples) and feature buffers that sit back
    to back in another are moved with one copy per run instead of one per clip; the features must not change."""
    import ctypes as C
    shape, models = tiny_models
    enc = B.Encoder(models["f16"][0], 0, B.OP_F16)
    fe = B.Frontend(0)
    lens = [64 * 700, 64 * 130, 0, 64 * 999, 64 * 250 + 17, 64 * 40]        # one empty clip, one that breaks the run
    src = [synth.synth_audio(1300 + i, n) for i, n in enumerate(lens)]
    want = enc.encode_pcm_batch(fe, src, B.MODE_ASR)
    block = np.concatenate(src).astype(np.float32)
    offs = np.cumsum([0] + lens)
    clips = [block[offs[i]:offs[i + 1]] for i in range(len(lens))]
    toks = [w.shape[0] for w in want]
    toff = np.cumsum([0] + toks)
    out_block = np.full((toff[-1], shape.out_dim), np.nan, np.float32)
    outs = [out_block[toff[i]:toff[i + 1]] for i in range(len(lens))]
    ntok = (C.c_int * len(lens))()
    B.check(enc.lib.q3a_encode_pcm_batch(fe.h, enc.h, B._ptr_array(clips), B._int_array(lens), len(lens), B.MODE_ASR,
                                         B._ptr_array(outs), ntok))
    assert list(ntok) == toks
    for a, b in zip(outs, want):
        assert np.array_equal(a, b)
    enc.close(); fe.close()


@needs("ln_fold")
@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
def test_layernorm_fold_matches_the_separate_kernel(lib, tiny_models, asr_model, filters, monkeypatch, opd):
    """LayerNorm folded into the QKV / FFN-up / proj1 GEMMs (epilogues.cuh) against the separate layernorm_kernel path:
    both must meet the tolerance against the oracle, agree with each other far inside it, and the folded path must stay
    batch-invariant (its row statistics are partial sums added in a fixed order)."""
    _, models = tiny_models
    shape, paths, Ws = asr_model
    mel = MO.log_mel_numpy(synth.synth_audio(305, 16000 * 9 + 4321), filters)
    others = [MO.log_mel_numpy(synth.synth_audio(306 + i, 16000 * (2 + 3 * i)), filters) for i in range(3)]
    want = _oracle(mel, Ws["f16"], shape, "asr")
    outs = {}
    monkeypatch.setenv("Q3A_LN_FOLD_MASK", "7")               # every site folded (the default folds attn_norm and ln_post)
    for name, env in (("fold", "Q3A_LN_FOLD"), ("separate", "Q3A_NO_LN_FOLD")):
        monkeypatch.setenv(env, "1")
        enc = B.Encoder(paths["f16"], 0, opd)
        monkeypatch.delenv(env)
        outs[name] = enc.encode_batch([mel], B.MODE_ASR)[0]
        if name == "fold":
            assert np.array_equal(enc.encode_batch(others + [mel], B.MODE_ASR)[-1], outs[name])      # batch-invariant
            assert enc.timings().total_launches < 18 * 6 + 12                                          # no LayerNorm launches
        enc.close()
        cos, rel = compare(outs[name], want)
        assert cos >= COS and rel <= REL, (name, cos, rel)
    # the two paths round different intermediates to 16 bits (x against LN(x)), so they differ from each other by about as
    # much as each differs from the oracle (measured r2b, bf16: cos 0.999977, rel 8.7e-3)
    cos, rel = compare(outs["fold"], outs["separate"])
    assert (cos >= 0.9999 and rel <= 1.5e-2) if opd == B.OP_BF16 else (cos >= 0.99999 and rel <= 3e-3), (cos, rel)


@needs("conv_cb_major")
def test_conv_k_order_only_changes_the_summation_order(lib, asr_model, filters, monkeypatch):
    """conv2 / conv3 walk K channel-block-major (L2 reuse of the nine taps) or tap-major (round 1): same products, another
    fp32 summation order, so the features agree to rounding noise and both meet the tolerance."""
    shape, paths, Ws = asr_model
    mel = MO.log_mel_numpy(synth.synth_audio(307, 16000 * 6 + 77), filters)
    want = _oracle(mel, Ws["f16"], shape, "asr")
    outs = []
    for env in ("Q3A_CONV_CB_MAJOR", "Q3A_CONV_TAP_MAJOR"):
        monkeypatch.setenv(env, "1")
        enc = B.Encoder(paths["f16"], 0, B.OP_F16)
        outs.append(enc.encode_batch([mel], B.MODE_ASR)[0])
        enc.close()
        monkeypatch.delenv(env)
        cos, rel = compare(outs[-1], want)
        assert cos >= COS and rel <= REL, (env, cos, rel)
    cos, rel = compare(outs[0], outs[1])
    assert cos >= 0.999999 and rel <= 1e-3, (cos, rel)


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
def test_residual_gemm_epilogue_warp_count_does_not_change_a_bit(lib, asr_model, filters, monkeypatch, opd):
    """The residual GEMMs that also emit the 16-bit copy and the LayerNorm partials run four or eight epilogue warps
    (Q3A_DUAL_EPI_WARPS; four leave room for a fifth operand ring stage).  The partials keep their layout -- sums over the
    even and over the odd 32-column chunks of a column block -- so the features are bit-identical for either count."""
    shape, paths, Ws = asr_model
    mels = [MO.log_mel_numpy(synth.synth_audio(311 + i, n), filters) for i, n in enumerate((16000 * 6 + 77, 16000 * 31 + 5, 4000))]
    outs = []
    for ew in ("8", "4"):
        monkeypatch.setenv("Q3A_DUAL_EPI_WARPS", ew)
        enc = B.Encoder(paths["f16"], 0, opd)
        outs.append(enc.encode_batch(mels, B.MODE_ASR))
        enc.close()
    for a, b in zip(*outs):
        assert np.isfinite(a).all() and np.array_equal(a, b)
