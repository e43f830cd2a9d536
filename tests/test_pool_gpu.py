"""The in-process multi-GPU executor (csrc/pool.cu, include/q3a.h q3a_pool_*): a corpus in, features out in utterance
order, bit-identical to a single encoder.  With one visible GPU the pool runs two replicas on it (two worker threads, two
stream sets), which exercises the same executor; with more GPUs every device gets a replica."""
import numpy as np
import pytest

from q3a_b200 import binding as B, synth

from conftest import needs  # noqa: E402

pytestmark = [pytest.mark.gpu, needs("pool")]


def _devices(lib):
    n = lib.q3a_device_count()
    return list(range(n)) if n >= 2 else [0, 0]


def test_pool_matches_single_encoder(lib, tiny_models):
    shape, models = tiny_models
    path, _ = models["f16"]
    rng = np.random.default_rng(11)
    clips = [synth.synth_audio(600 + i, int(n)) for i, n in enumerate(rng.integers(0, 16000 * 9, 37))]
    clips[5] = clips[5][:0]
    clips[6] = clips[6][:159]
    pool = B.Pool(path, _devices(lib), B.OP_F16)
    for mode in (B.MODE_ASR, B.MODE_ALIGNER):
        outs, shard, ms = pool.encode_pcm(clips, mode, shape.out_dim)
        assert sorted(set(shard)) == list(range(len(pool.devices))) and len(ms) == len(pool.devices) and all(m >= 0 for m in ms)
        assert shard == B.plan_shards([len(c) // 160 for c in clips], mode, len(pool.devices))
        enc, fe = B.Encoder(path, 0, B.OP_F16), B.Frontend(0)
        want = enc.encode_pcm_batch(fe, clips, mode)
        enc.close(); fe.close()
        for a, b in zip(outs, want):
            assert a.shape == b.shape and np.array_equal(a, b)
    outs, shard, _ = pool.encode_pcm([], B.MODE_ASR, shape.out_dim)        # empty corpus
    assert outs == [] and shard == []
    pool.close()


def test_pool_asr_width_and_errors(lib, tmp_path):
    shape = synth.ASR_06B
    path = synth.write_gguf(str(tmp_path / "asr.gguf"), shape, synth.random_weights(shape, 0), "f16")
    pool = B.Pool(path, _devices(lib), B.OP_BF16)
    clips = [synth.synth_audio(650 + i, 16000 * (3 + 5 * i) + 77 * i) for i in range(6)]
    outs, shard, ms = pool.encode_pcm(clips, B.MODE_ASR)
    enc, fe = B.Encoder(path, 0, B.OP_BF16), B.Frontend(0)
    for i, c in enumerate(clips):                                            # one by one on a single encoder
        assert np.array_equal(outs[i], enc.encode_pcm_batch(fe, [c], B.MODE_ASR)[0])
    enc.close(); fe.close()
    with pytest.raises(B.Q3AError):                                          # a worker's error reaches the caller's thread
        pool.encode_pcm(clips, 17)
    pool.close()
    with pytest.raises(B.Q3AError, match="cannot open"):
        B.Pool(str(tmp_path / "missing.gguf"), [0])
    assert B.load().q3a_bind_thread_to_device(0) >= 0
