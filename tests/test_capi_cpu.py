"""The C-ABI library on a box without a GPU: it loads, exports every symbol include/q3a.h declares, its
host-only entry points agree with the reference/oracle, and compute entry points fail loudly
(there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from q3a_b200 import binding as B, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_every_declared_symbol(lib):
    with open(os.path.join(ROOT, "include", "q3a.h")) as f:
        header = f.read()
    declared = set(re.findall(r"\b(q3a_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 20
    assert declared == set(B.SYMBOLS), declared ^ set(B.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.q3a_version() >= 100


def test_generate_mel_filters_matches_reference_golden(lib):
    g = np.load(os.path.join(ROOT, "tests", "golden", "mel_ref.npz"))
    bank = B.generate_mel_filters()
    assert np.abs(bank - g["filters"]).max() <= 1e-7
    assert int((bank != 0).sum()) == 396


def test_frame_and_token_counts(lib):
    for n in (0, 1, 159, 160, 16000, 480000, 480001, 1472000):
        assert lib.q3a_mel_n_len(n) == n // 160 == synth.mel_frames(n)
    for nf in (0, 1, 99, 100, 101, 237, 3000, 9200, 12000):
        assert lib.q3a_encoder_n_tokens(nf, B.MODE_ASR) == synth.chunk_tokens(nf)
        assert lib.q3a_encoder_n_tokens(nf, B.MODE_ALIGNER) == synth.chunk_tokens(nf)
    assert lib.q3a_encoder_n_tokens(3000, B.MODE_NO_CHUNK) == 375      # one 3000-frame chunk: 1500 -> 750 -> 375


def test_plan_shards_is_balanced_and_deterministic(lib):
    rng = np.random.default_rng(500)
    frames = [int(x) for x in np.exp(rng.uniform(np.log(200), np.log(6000), 400))]
    a = B.plan_shards(frames, B.MODE_ASR, 8)
    assert a == B.plan_shards(frames, B.MODE_ASR, 8)
    assert set(a) == set(range(8))
    cost = np.array([synth.chunk_tokens(f) for f in frames], dtype=np.float64)
    cost = 0.71e9 * cost + 4 * 64 * 14 * 18 * cost ** 2
    loads = np.array([cost[np.array(a) == s].sum() for s in range(8)])
    assert loads.max() / loads.mean() < 1.02
    assert B.plan_shards([], B.MODE_ASR, 4) == []
    with pytest.raises(B.Q3AError):
        B.plan_shards(frames, B.MODE_ASR, 0)


def test_loader_error_codes(lib, tmp_path):
    h = C.c_void_p()
    rc = lib.q3a_encoder_load(str(tmp_path / "missing.gguf").encode(), 0, B.OP_BF16, C.byref(h))
    assert rc == 3 and b"cannot open" in lib.q3a_last_error()          # Q3A_ERR_IO
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"NOPE" + bytes(64))
    rc = lib.q3a_encoder_load(str(bad).encode(), 0, B.OP_BF16, C.byref(h))
    assert rc == 4 and b"magic" in lib.q3a_last_error()                 # Q3A_ERR_FORMAT
    rc = lib.q3a_encoder_load(str(bad).encode(), 0, 7, C.byref(h))
    assert rc == 1                                                       # Q3A_ERR_ARG
    assert not h.value


def test_null_handles_report_model_not_loaded(lib):
    hp = B.HParams()
    assert lib.q3a_encoder_get_hparams(None, C.byref(hp)) == 5           # Q3A_ERR_STATE
    assert b"Model not loaded" in lib.q3a_last_error()                   # audio_encoder.cpp:316-319
    assert lib.q3a_encoder_encode_batch(None, None, 128, None, 0, 0, None, None) == 5
    lib.q3a_encoder_free(None)
    lib.q3a_frontend_free(None)


def test_compute_fails_loudly_without_gpu(lib, tiny_models):
    if lib.q3a_device_count() > 0:
        pytest.skip("a B200 is present")
    fe = C.c_void_p()
    assert lib.q3a_frontend_create(0, None, 0, 0, C.byref(fe)) == 2      # Q3A_ERR_CUDA
    assert b"no CPU fallback" in lib.q3a_last_error()
    _, models = tiny_models
    h = C.c_void_p()
    assert lib.q3a_encoder_load(models["f16"][0].encode(), 0, B.OP_BF16, C.byref(h)) == 2
    assert not h.value


def test_host_call_schedule(lib):
    """The host-pointer entry points cut a call into passes and, inside a pass, into upload / download parts so that
    PCIe copies hide under kernels.  Pure host logic: contiguous, monotone, bounded, and small calls stay whole."""
    n30 = 480000
    # BASELINE config 2: one pass, 4 upload parts (the first one small: only its copy is exposed), 2 download halves
    ps, ip, op = B.plan_host_call([n30] * 64, B.MODE_ASR)
    assert set(ps) == {0}
    assert ip == sorted(ip) and sorted(set(ip)) == [0, 1, 2, 3]
    assert 4 <= ip.count(0) <= 8 and all(16 <= ip.count(k) <= 22 for k in (1, 2, 3))
    assert op == sorted(op) and op.count(0) == 32 and op.count(1) == 32
    # a small call is not partitioned at all
    ps, ip, op = B.plan_host_call([n30] * 3, B.MODE_ASR)
    assert set(ps) == set(ip) == set(op) == {0}
    # conv-only output is never split (no transformer tail to hide a download under)
    assert set(B.plan_host_call([n30] * 64, B.MODE_CONV_ONLY, 896)[2]) == {0}
    # a 10-h ragged corpus: passes bounded by 4096 stem chunks, everything contiguous and monotone
    rng = np.random.default_rng(500)
    lens = [int(16000 * np.exp(rng.uniform(np.log(2), np.log(60)))) for _ in range(1700)]
    ps, ip, op = B.plan_host_call(lens, B.MODE_ASR)
    assert ps == sorted(ps) and ps[0] == 0 and len(set(ps)) >= 2
    for p in set(ps):
        idx = [i for i, q in enumerate(ps) if q == p]
        assert idx == list(range(idx[0], idx[-1] + 1))
        assert sum(-(-(lens[i] // 160) // 100) for i in idx) <= 4096
        sub_in, sub_out = [ip[i] for i in idx], [op[i] for i in idx]
        assert sub_in == sorted(sub_in) and sub_in[0] == 0 and max(sub_in) <= 3
        assert sub_out == sorted(sub_out) and sub_out[0] == 0 and max(sub_out) <= 1
    assert B.plan_host_call([], B.MODE_ASR) == ([], [], [])
    with pytest.raises(B.Q3AError):
        B.plan_host_call([-1], B.MODE_ASR)


def test_host_binding_never_raises_and_never_widens():
    """hostbind.bind_to_gpu is an optimisation for multi-socket hosts: without NVML (this container) it reports why it
    did nothing, and it can only ever shrink the set of CPUs the process may run on."""
    import os
    from q3a_b200 import hostbind
    before = os.sched_getaffinity(0)
    info = hostbind.bind_to_gpu(0)
    after = os.sched_getaffinity(0)
    assert after <= before and len(after) > 0
    assert info["bound"] == (after != before)
    assert info["cpus"] == len(after)
    os.sched_setaffinity(0, before)


def test_round2_entry_points_validate_their_arguments(lib, tmp_path):
    """pool, asynchronous call, decoder and host helpers: argument and state errors are reported before any CUDA work, so
    they behave the same on a box without a GPU."""
    h = C.c_void_p()
    devs = (C.c_int * 2)(0, 0)
    assert lib.q3a_pool_create(None, devs, 2, B.OP_BF16, C.byref(h)) == 1 and not h.value            # Q3A_ERR_ARG
    assert lib.q3a_pool_create(b"x.gguf", devs, 0, B.OP_BF16, C.byref(h)) == 1
    assert lib.q3a_pool_create(b"x.gguf", devs, 2, B.OP_BF16, None) == 1
    assert lib.q3a_pool_n_devices(None) == 0
    assert lib.q3a_pool_encode_pcm(None, None, None, 0, 0, None, None, None, None) == 5              # Q3A_ERR_STATE
    lib.q3a_pool_free(None)
    tk = C.c_int64(-1)
    assert lib.q3a_encode_pcm_submit(None, None, None, None, 0, 0, None, None, C.byref(tk)) == 5
    assert lib.q3a_encode_wait(None, 0) == 5
    rc = lib.q3a_decoder_load(str(tmp_path / "missing.gguf").encode(), 0, B.OP_BF16, C.byref(h))
    assert rc == 3 and b"cannot open" in lib.q3a_last_error()                                         # Q3A_ERR_IO
    assert lib.q3a_decoder_load(None, 0, B.OP_BF16, C.byref(h)) == 1
    assert lib.q3a_decoder_load(b"x.gguf", 0, 9, C.byref(h)) == 1
    assert lib.q3a_decoder_forward(None, None, 0, None, 0, -1, None) == 5
    assert lib.q3a_decoder_get_hparams(None, None) == 5
    lib.q3a_decoder_free(None)
    p = C.c_void_p(1)
    assert lib.q3a_host_alloc(0, C.byref(p)) == 0 and not p.value                                      # zero bytes: OK, null
    assert lib.q3a_host_alloc(16, None) == 1
    lib.q3a_host_free(None)
    # an encoder-only GGUF is not an aligner decoder file
    shape = synth.tiny_shape()
    path = synth.write_gguf(str(tmp_path / "enc.gguf"), shape, synth.random_weights(shape, 7), "f16")
    rc = lib.q3a_decoder_load(path.encode(), 0, B.OP_BF16, C.byref(h))
    assert rc in (2, 4) and not h.value        # FORMAT on a GPU box; no device is reported first where there is none


def test_decoder_gguf_round_trip_keys(lib, tmp_path):
    """the decoder GGUF the tests write carries the keys ForcedAligner::parse_hparams reads (forced_aligner.cpp:158-168)"""
    import gguf
    shape = synth.tiny_decoder_shape()
    path = synth.write_decoder_gguf(str(tmp_path / "dec.gguf"), shape, synth.random_decoder_weights(shape, 3), "f16")
    r = gguf.GGUFReader(path)
    keys = set(r.fields)
    for k in ("block_count", "embedding_length", "attention.head_count", "attention.head_count_kv", "feed_forward_length",
              "attention.key_length", "attention.layer_norm_rms_epsilon", "rope.freq_base", "vocab_size", "classify_num"):
        assert f"qwen3-asr.{k}" in keys
    names = {t.name for t in r.tensors}
    assert {"token_embd.weight", "output_norm.weight", "output.weight", "blk.0.attn_q_norm.weight", "blk.2.ffn_gate.weight"} <= names
