"""The NAR-decoder restatement (oracle/decoder_oracle.py) pinned to upstream: transformers' Qwen3Model in float64 on a tiny
tower (tests/golden/make_hf_decoder_golden.py), audio rows spliced into the embeddings as forced_aligner.cpp:964-993 does."""
import os
import sys

import numpy as np

from oracle import decoder_oracle as DO
from q3a_b200 import synth
from conftest import compare

GOLD = os.path.join(os.path.dirname(__file__), "golden")
sys.path.insert(0, GOLD)


def test_decoder_graph_matches_upstream_qwen3():
    import make_hf_decoder_golden as G
    g = np.load(os.path.join(GOLD, "hf_decoder_tiny.npz"))
    shape = synth.tiny_decoder_shape()
    w = synth.random_decoder_weights(shape, 3)
    for seed, n, n_audio, audio_start in G.CASES:
        tokens, audio = G.case_inputs(shape, seed, n, n_audio, audio_start)
        want = g[f"logits_{seed}"]
        got = DO.forward(tokens, w, shape, audio, audio_start, ideal=True)
        assert got.shape == want.shape == (n, shape.n_classes)
        assert np.abs(got - want).max() <= 2e-6 * np.abs(want).max()
        # ggml's f16 activation rounding in the weight matmuls moves the logits by well under the encoder's tolerance
        cos, rel = compare(DO.forward(tokens, w, shape, audio, audio_start), want)
        assert cos >= 0.99999 and rel <= 5e-3


def test_decoder_is_causal_and_splices_audio():
    shape = synth.tiny_decoder_shape()
    w = synth.random_decoder_weights(shape, 3)
    rng = np.random.default_rng(1)
    tok = rng.integers(0, shape.vocab, 40).astype(np.int32)
    aud = rng.standard_normal((10, shape.hidden)).astype(np.float32)
    y = DO.forward(tok, w, shape, aud, 5)
    tok2 = tok.copy(); tok2[30] = (tok2[30] + 1) % shape.vocab
    y2 = DO.forward(tok2, w, shape, aud, 5)
    assert np.array_equal(y[:30], y2[:30]) and not np.array_equal(y[30:], y2[30:])
    tok3 = tok.copy(); tok3[7] = (tok3[7] + 1) % shape.vocab          # under the audio span: the token is ignored
    assert np.array_equal(DO.forward(tok3, w, shape, aud, 5), y)
    # a span that does not fit is not spliced at all (:966)
    assert np.array_equal(DO.forward(tok, w, shape, aud, 35), DO.forward(tok, w, shape, None, -1))
