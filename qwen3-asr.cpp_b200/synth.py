"""Synthetic 16 kHz audio and random-init encoder weights (GGUF) for tests and bench.

No real checkpoints or corpora are available offline, so every measurement and parity
check runs on data made here from fixed seeds.

The GGUF files follow the naming and dtype policy of the reference converter
(/root/reference/scripts/convert_hf_to_gguf.py:50-103 names, :254-311 dtypes,
:441-447 metadata keys): 1-D tensors are always F32; >=2-D tensors take the file
type (F32 / F16 / Q8_0), except that conv kernels stay F16 in a Q8_0 file because
their innermost dimension (3) is not a multiple of the 32-element quant block.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

SAMPLE_RATE = 16000


@dataclass(frozen=True)
class EncoderShape:
    """Architecture numbers of one audio tower (configs/*.json in the reference)."""
    name: str
    n_layers: int
    d_model: int
    n_heads: int
    ffn_dim: int
    conv_ch: int = 480
    out_dim: int = 1024
    n_mels: int = 128
    eps: float = 1e-5

    @property
    def conv_feat(self) -> int:            # 16 mel rows left after three stride-2 convs
        return self.conv_ch * (self.n_mels // 8)


# gguf_loader.h:15-25 (ASR defaults), forced_aligner.h:36-44 (aligner defaults)
ASR_06B = EncoderShape("qwen3-asr-0.6b", 18, 896, 14, 3584)
ALIGNER_06B = EncoderShape("qwen3-forced-aligner-0.6b", 24, 1024, 16, 4096)


def tiny_shape(n_layers: int = 2, d_model: int = 128, n_heads: int = 2, ffn_dim: int = 256,
               conv_ch: int = 32, out_dim: int = 64) -> EncoderShape:
    """A small tower with the same topology, for fast CPU-side tests."""
    return EncoderShape("tiny", n_layers, d_model, n_heads, ffn_dim, conv_ch, out_dim)


def synth_audio(seed: int, n_samples: int) -> np.ndarray:
    """Speech-like test signal: AM-modulated tones + noise + silent gaps, int16-quantised.

    Returns float32 in [-1, 1) exactly representable as int16/32768, like the reference's
    WAV loader produces (mel_spectrogram.cpp:202).
    """
    rng = np.random.default_rng(seed)
    t = np.arange(n_samples, dtype=np.float64) / SAMPLE_RATE
    x = np.zeros(n_samples, dtype=np.float64)
    for _ in range(int(rng.integers(3, 7))):
        f = rng.uniform(80.0, 7000.0)
        ph = rng.uniform(0, 2 * np.pi)
        am_f = rng.uniform(0.5, 4.0)
        am_ph = rng.uniform(0, 2 * np.pi)
        amp = rng.uniform(0.2, 1.0)
        x += amp * (0.5 + 0.5 * np.sin(2 * np.pi * am_f * t + am_ph)) * np.sin(2 * np.pi * f * t + ph)
    peak = np.max(np.abs(x)) if n_samples else 1.0
    x *= 0.5 / max(peak, 1e-9)
    x += rng.standard_normal(n_samples) * (10.0 ** (-30.0 / 20.0)) * 0.5
    dur = n_samples / SAMPLE_RATE
    if dur > 2.0:                          # zeroed gaps exercise the 1e-10 floor and the max-8 clamp
        for _ in range(int(rng.integers(2, 4))):
            g = rng.uniform(0.2, min(1.0, dur / 4))
            s = rng.uniform(0, dur - g)
            x[int(s * SAMPLE_RATE):int((s + g) * SAMPLE_RATE)] = 0.0
    q = np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)
    return (q.astype(np.float32) / np.float32(32768.0)).astype(np.float32)


def random_weights(shape: EncoderShape, seed: int) -> dict[str, np.ndarray]:
    """Random-init tower in f32, keyed by GGUF tensor name (PyTorch memory order)."""
    rng = np.random.default_rng(seed)
    w: dict[str, np.ndarray] = {}

    def lin(name, n_out, n_in, bias=True):
        w[f"{name}.weight"] = (rng.standard_normal((n_out, n_in)) / np.sqrt(n_in)).astype(np.float32)
        if bias:
            w[f"{name}.bias"] = (rng.standard_normal(n_out) * 0.02).astype(np.float32)

    def norm(name, n):
        w[f"{name}.weight"] = (1.0 + rng.standard_normal(n) * 0.02).astype(np.float32)
        w[f"{name}.bias"] = (rng.standard_normal(n) * 0.02).astype(np.float32)

    c = shape.conv_ch
    for i, cin in ((1, 1), (2, c), (3, c)):
        fan_in = cin * 9
        w[f"audio.encoder.conv{i}.weight"] = (
            rng.standard_normal((c, cin, 3, 3)) * np.sqrt(2.0 / fan_in)).astype(np.float32)
        w[f"audio.encoder.conv{i}.bias"] = (rng.standard_normal(c) * 0.02).astype(np.float32)
    lin("audio.encoder.conv_out", shape.d_model, shape.conv_feat, bias=False)
    for l in range(shape.n_layers):
        p = f"audio.encoder.blk.{l}"
        norm(f"{p}.attn_norm", shape.d_model)
        for nm in ("attn_q", "attn_k", "attn_v", "attn_out"):
            lin(f"{p}.{nm}", shape.d_model, shape.d_model)
        norm(f"{p}.ffn_norm", shape.d_model)
        lin(f"{p}.ffn_up", shape.ffn_dim, shape.d_model)
        lin(f"{p}.ffn_down", shape.d_model, shape.ffn_dim)
    norm("audio.encoder.ln_post", shape.d_model)
    lin("audio.encoder.proj1", shape.d_model, shape.d_model)
    lin("audio.encoder.proj2", shape.out_dim, shape.d_model)
    return w


def write_gguf(path: str, shape: EncoderShape, weights: dict[str, np.ndarray], ftype: str = "f16") -> str:
    """Write the tower the way the reference converter would (ftype: f32 | f16 | q8_0)."""
    import gguf
    from gguf import GGMLQuantizationType as QT

    assert ftype in ("f32", "f16", "q8_0")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    wr = gguf.GGUFWriter(path, "qwen3-asr")
    wr.add_name(shape.name)
    arch = "qwen3-asr"
    wr.add_uint32(f"{arch}.audio.encoder.layer_count", shape.n_layers)
    wr.add_uint32(f"{arch}.audio.encoder.embedding_length", shape.d_model)
    wr.add_uint32(f"{arch}.audio.encoder.attention.head_count", shape.n_heads)
    wr.add_uint32(f"{arch}.audio.encoder.feed_forward_length", shape.ffn_dim)
    wr.add_uint32(f"{arch}.audio.num_mel_bins", shape.n_mels)
    wr.add_uint32(f"{arch}.audio.conv_channels", shape.conv_ch)
    for name, t in weights.items():
        if t.ndim <= 1 or ftype == "f32":
            wr.add_tensor(name, np.ascontiguousarray(t, dtype=np.float32), raw_dtype=QT.F32)
        elif ftype == "f16" or (t.shape[-1] % 32) != 0:
            wr.add_tensor(name, np.ascontiguousarray(t.astype(np.float16)), raw_dtype=QT.F16)
        else:
            q = gguf.quants.quantize(np.ascontiguousarray(t, dtype=np.float32), QT.Q8_0)
            wr.add_tensor(name, q, raw_dtype=QT.Q8_0)
    wr.write_header_to_file()
    wr.write_kv_data_to_file()
    wr.write_tensors_to_file()
    wr.close()
    return path


def stored_weights(weights: dict[str, np.ndarray], ftype: str) -> dict[str, np.ndarray]:
    """What the file actually holds, dequantised back to f32 (the oracle's view of the file).

    Q8_0 tensors are returned as the (int8 q, f16 d) pair under name+'#q8' as well, because
    the ggml CPU path does an integer dot with them rather than dequantising.
    """
    import gguf
    from gguf import GGMLQuantizationType as QT

    out: dict[str, np.ndarray] = {}
    for name, t in weights.items():
        if t.ndim <= 1 or ftype == "f32":
            out[name] = t.astype(np.float32)
        elif ftype == "f16" or (t.shape[-1] % 32) != 0:
            out[name] = t.astype(np.float16).astype(np.float32)
        else:
            q = gguf.quants.quantize(np.ascontiguousarray(t, dtype=np.float32), QT.Q8_0)
            out[name] = gguf.quants.dequantize(q, QT.Q8_0).astype(np.float32)
            out[name + "#q8"] = q
    return out


def chunk_tokens(n_frames: int) -> int:
    """Tokens the stem emits for n_frames mel frames (audio_encoder.cpp:304-310, :327-345)."""
    total = 0
    for s in range(0, n_frames, 100):
        ln = min(100, n_frames - s)
        for _ in range(3):
            ln = (ln - 1) // 2 + 1
        total += ln
    return total


def mel_frames(n_samples: int) -> int:
    """n_len of log_mel_spectrogram (mel_spectrogram.cpp:517-520)."""
    return n_samples // 160


# ---- forced aligner: NAR decoder (forced_aligner.h:50-61; tensor names forced_aligner.cpp:267-314,381-410) -----------
@dataclass(frozen=True)
class DecoderShape:
    name: str
    n_layers: int
    hidden: int
    n_heads: int
    n_kv_heads: int
    intermediate: int
    vocab: int
    n_classes: int
    head_dim: int = 128
    rms_eps: float = 1e-6
    rope_theta: float = 1000000.0


ALIGNER_DECODER_06B = DecoderShape("qwen3-forced-aligner-0.6b-decoder", 28, 1024, 16, 8, 3072, 152064, 5000)


def tiny_decoder_shape() -> DecoderShape:
    """Same topology (head_dim 128, 2 query heads per kv head), small enough for a CPU oracle in milliseconds."""
    return DecoderShape("tiny-decoder", 3, 128, 4, 2, 256, 320, 48)


def random_decoder_weights(shape: DecoderShape, seed: int) -> dict[str, np.ndarray]:
    rng = np.random.default_rng(seed)
    w: dict[str, np.ndarray] = {}

    def lin(name, n_out, n_in):
        w[name] = (rng.standard_normal((n_out, n_in)) / np.sqrt(n_in)).astype(np.float32)

    def norm(name, n):
        w[name] = (1.0 + rng.standard_normal(n) * 0.05).astype(np.float32)

    dq, dkv = shape.n_heads * shape.head_dim, shape.n_kv_heads * shape.head_dim
    w["token_embd.weight"] = (rng.standard_normal((shape.vocab, shape.hidden)) * 0.5).astype(np.float32)
    for l in range(shape.n_layers):
        p = f"blk.{l}"
        norm(f"{p}.attn_norm.weight", shape.hidden)
        lin(f"{p}.attn_q.weight", dq, shape.hidden)
        lin(f"{p}.attn_k.weight", dkv, shape.hidden)
        lin(f"{p}.attn_v.weight", dkv, shape.hidden)
        norm(f"{p}.attn_q_norm.weight", shape.head_dim)
        norm(f"{p}.attn_k_norm.weight", shape.head_dim)
        lin(f"{p}.attn_output.weight", shape.hidden, dq)
        norm(f"{p}.ffn_norm.weight", shape.hidden)
        lin(f"{p}.ffn_gate.weight", shape.intermediate, shape.hidden)
        lin(f"{p}.ffn_up.weight", shape.intermediate, shape.hidden)
        lin(f"{p}.ffn_down.weight", shape.hidden, shape.intermediate)
    norm("output_norm.weight", shape.hidden)
    lin("output.weight", shape.n_classes, shape.hidden)
    return w


def write_decoder_gguf(path: str, shape: DecoderShape, weights: dict[str, np.ndarray], ftype: str = "f16") -> str:
    """The decoder half of an aligner GGUF: the keys ForcedAligner::parse_hparams reads (forced_aligner.cpp:158-168)."""
    import gguf
    from gguf import GGMLQuantizationType as QT
    assert ftype in ("f32", "f16")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    wr = gguf.GGUFWriter(path, "qwen3-asr")
    wr.add_name(shape.name)
    a = "qwen3-asr"
    wr.add_uint32(f"{a}.block_count", shape.n_layers)
    wr.add_uint32(f"{a}.embedding_length", shape.hidden)
    wr.add_uint32(f"{a}.attention.head_count", shape.n_heads)
    wr.add_uint32(f"{a}.attention.head_count_kv", shape.n_kv_heads)
    wr.add_uint32(f"{a}.feed_forward_length", shape.intermediate)
    wr.add_uint32(f"{a}.attention.key_length", shape.head_dim)
    wr.add_float32(f"{a}.attention.layer_norm_rms_epsilon", shape.rms_eps)
    wr.add_float32(f"{a}.rope.freq_base", shape.rope_theta)
    wr.add_uint32(f"{a}.vocab_size", shape.vocab)
    wr.add_uint32(f"{a}.classify_num", shape.n_classes)
    for name, t in weights.items():
        if t.ndim <= 1 or ftype == "f32":
            wr.add_tensor(name, np.ascontiguousarray(t, dtype=np.float32), raw_dtype=QT.F32)
        else:
            wr.add_tensor(name, np.ascontiguousarray(t.astype(np.float16)), raw_dtype=QT.F16)
    wr.write_header_to_file()
    wr.write_kv_data_to_file()
    wr.write_tensors_to_file()
    wr.close()
    return path
