"""Row f4: the forced aligner's NAR decoder prefill on the GPU (csrc/decoder.cu, attention_dec.cu) against the CPU
restatement of ForcedAligner::build_decoder_graph (oracle/decoder_oracle.py) and against upstream Qwen3Model outputs
(tests/golden/hf_decoder_tiny.npz).  Tolerance: the encoder's (cosine >= 0.999, max|a-b| / max|b| <= 1e-2), plus agreement
of the timestamp classes the caller extracts (argmax over the 5000 classes, forced_aligner.cpp:1279-1308) wherever the
oracle's top-2 margin is not a rounding-level tie."""
import os
import sys

import numpy as np
import pytest

from oracle import decoder_oracle as DO, encoder_oracle as EO, mel_oracle as MO
from q3a_b200 import binding as B, synth
from conftest import compare

from conftest import needs  # noqa: E402

pytestmark = [pytest.mark.gpu, needs("decoder")]
GOLD = os.path.join(os.path.dirname(__file__), "golden")
sys.path.insert(0, GOLD)
COS, REL = 0.999, 1e-2


def _argmax_agrees(got, want):
    top2 = np.sort(want, axis=1)[:, -2:]
    clear = (top2[:, 1] - top2[:, 0]) > 0.05 * np.abs(want).max()
    return np.array_equal(got.argmax(1)[clear], want.argmax(1)[clear]) and clear.mean() > 0.5


@pytest.fixture(scope="module")
def tiny_decoder(tmp_path_factory):
    d = tmp_path_factory.mktemp("dec")
    shape = synth.tiny_decoder_shape()
    w = synth.random_decoder_weights(shape, 3)
    paths = {ft: synth.write_decoder_gguf(str(d / f"dec_{ft}.gguf"), shape, w, ft) for ft in ("f32", "f16")}
    w16 = {k: (v.astype(np.float16).astype(np.float32) if v.ndim > 1 else v) for k, v in w.items()}
    return shape, paths, {"f32": w, "f16": w16}


@pytest.mark.parametrize("opd", [B.OP_BF16, B.OP_F16])
def test_tiny_decoder_matches_oracle_and_upstream(lib, tiny_decoder, opd):
    import make_hf_decoder_golden as G
    shape, paths, W = tiny_decoder
    g = np.load(os.path.join(GOLD, "hf_decoder_tiny.npz"))
    dec = B.Decoder(paths["f32"], 0, opd)
    assert (dec.hp.n_layers, dec.hp.hidden, dec.hp.n_heads, dec.hp.n_kv_heads, dec.hp.head_dim, dec.hp.intermediate, dec.hp.vocab,
            dec.hp.n_classes) == (shape.n_layers, shape.hidden, shape.n_heads, shape.n_kv_heads, 128, shape.intermediate, shape.vocab,
                                  shape.n_classes)
    for seed, n, n_audio, audio_start in G.CASES:
        tokens, audio = G.case_inputs(shape, seed, n, n_audio, audio_start)
        got = dec.forward(tokens, audio, audio_start)
        for want in (g[f"logits_{seed}"], DO.forward(tokens, W["f32"], shape, audio, audio_start)):
            cos, rel = compare(got, want)
            assert got.shape == want.shape and cos >= COS and rel <= REL, (seed, cos, rel)
            assert _argmax_agrees(got, want)
    dec.close()


def test_decoder_edge_lengths_and_errors(lib, tiny_decoder):
    """1 token, a tile boundary (128 / 129 rows), several query tiles (causal block skipping), audio span that does not
    fit (ignored, forced_aligner.cpp:966), out-of-range token ids (clamped), missing file."""
    shape, paths, W = tiny_decoder
    dec = B.Decoder(paths["f16"], 0, B.OP_F16)
    rng = np.random.default_rng(9)
    for n in (1, 2, 64, 65, 128, 129, 300, 515):
        tok = rng.integers(0, shape.vocab, n).astype(np.int32)
        na = min(n // 2, 40)
        aud = (rng.standard_normal((na, shape.hidden)) * 0.7).astype(np.float32) if na else None
        got = dec.forward(tok, aud, 1 if na else -1)
        want = DO.forward(tok, W["f16"], shape, aud, 1 if na else -1)
        cos, rel = compare(got, want)
        assert cos >= COS and rel <= REL, (n, cos, rel)
    tok = rng.integers(0, shape.vocab, 20).astype(np.int32)
    aud = rng.standard_normal((15, shape.hidden)).astype(np.float32)
    assert np.array_equal(dec.forward(tok, aud, 10), dec.forward(tok, None, -1))           # 10 + 15 > 20: no splice
    assert dec.forward(np.zeros(0, np.int32)).shape == (0, shape.n_classes)
    dec.close()
    with pytest.raises(B.Q3AError, match="cannot open"):
        B.Decoder(paths["f16"] + ".missing", 0)


def test_encoder_features_feed_the_decoder_on_the_device(lib, tiny_decoder, tmp_path, filters):
    """Rows f1 + f4 together: q3a_encode_pcm_device leaves [T][hidden] features in HBM, q3a_decoder_forward_device splices
    them in on the same stream -- no host round trip (the reference bounces them through a std::vector,
    forced_aligner.cpp:1661-1667 -> :1143-1146).  Same logits as going through the host."""
    import torch
    shape, paths, W = tiny_decoder
    eshape = synth.tiny_shape(out_dim=shape.hidden)
    ew = synth.random_weights(eshape, 7)
    epath = synth.write_gguf(str(tmp_path / "enc.gguf"), eshape, ew, "f16")
    enc, fe, dec = B.Encoder(epath, 0, B.OP_F16), B.Frontend(0), B.Decoder(paths["f16"], 0, B.OP_F16)
    pcm = synth.synth_audio(77, 16000 * 4 + 800)
    T = enc.n_tokens(len(pcm) // 160, B.MODE_ALIGNER)
    rng = np.random.default_rng(4)
    tokens = rng.integers(0, shape.vocab, T + 25).astype(np.int32)
    dpcm = torch.zeros(len(pcm) + 64, device="cuda"); dpcm[:len(pcm)] = torch.from_numpy(pcm).cuda()
    feats = torch.empty((T, shape.hidden), device="cuda")
    logits = torch.full((len(tokens), shape.n_classes), float("nan"), device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        enc.encode_pcm_device(fe, dpcm.data_ptr(), [0], [len(pcm)], feats.data_ptr(), B.MODE_ALIGNER, st.cuda_stream)
        dec.forward_device(tokens, feats.data_ptr(), T, 7, logits.data_ptr(), st.cuda_stream)
    st.synchronize()
    host_feats = enc.encode_pcm_batch(fe, [pcm], B.MODE_ALIGNER)[0]
    assert np.array_equal(feats.cpu().numpy(), host_feats)
    via_host = dec.forward(tokens, host_feats, 7)
    assert np.array_equal(logits.cpu().numpy(), via_host)
    want = DO.forward(tokens, W["f16"], shape, host_feats, 7)
    cos, rel = compare(via_host, want)
    assert cos >= COS and rel <= REL, (cos, rel)
    enc.close(); fe.close(); dec.close()
