"""ctypes binding to libq3a.so (include/q3a.h) used by tests/, bench.py and __graft_entry__.py.

This is harness code: the product is the C-ABI library.  There is no CPU fallback — load()
raises if the library is missing, and every compute call raises Q3AError when the library
reports an error (e.g. no sm_100 device).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libq3a.so")

OK = 0
OP_BF16, OP_F16 = 0, 1
MODE_ASR, MODE_ALIGNER, MODE_NO_CHUNK, MODE_CONV_ONLY = 0, 1, 2, 3


class Q3AError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libq3a error {code}: {msg}")
        self.code = code


class HParams(C.Structure):
    _fields_ = [("n_encoder_layers", C.c_int32), ("d_model", C.c_int32), ("n_attention_heads", C.c_int32),
                ("ffn_dim", C.c_int32), ("conv_channels", C.c_int32), ("conv_out_dim", C.c_int32),
                ("n_mel_bins", C.c_int32), ("n_window_infer", C.c_int32), ("layer_norm_eps", C.c_float),
                ("out_dim", C.c_int32), ("weight_ftype", C.c_int32)]


class Timings(C.Structure):
    _fields_ = [("mel_ms", C.c_float), ("conv_chunk_ms", C.c_float), ("transformer_ms", C.c_float),
                ("gemm_ms", C.c_float), ("attention_ms", C.c_float), ("other_ms", C.c_float),
                ("gemm_flops", C.c_double), ("attention_flops", C.c_double), ("gemm_launches", C.c_int32),
                ("attention_launches", C.c_int32), ("total_launches", C.c_int32)]


# every symbol include/q3a.h declares: name -> (restype, argtypes)
_P, _I, _I64 = C.c_void_p, C.c_int, C.c_int64
SYMBOLS = {
    "q3a_last_error": (C.c_char_p, []),
    "q3a_version": (_I, []),
    "q3a_device_count": (_I, []),
    "q3a_generate_mel_filters": (_I, [_P, _I, _I, _I]),
    "q3a_mel_n_len": (_I, [_I]),
    "q3a_frontend_create": (_I, [_I, _P, _I, _I, C.POINTER(_P)]),
    "q3a_frontend_free": (None, [_P]),
    "q3a_mel_batch": (_I, [_P, _P, _P, _I, _P, _P]),
    "q3a_mel_batch_device": (_I, [_P, _P, _P, _P, _I, _P, _P, _P, _I, _P]),
    "q3a_mel_batch_device_noalloc": (_I, [_P, _P, _P, _P, _I, _P, _P, _P, _I, _P]),
    "q3a_mel_noalloc_max_clips": (_I, []),
    "q3a_encoder_load": (_I, [C.c_char_p, _I, _I, C.POINTER(_P)]),
    "q3a_encoder_free": (None, [_P]),
    "q3a_encoder_get_hparams": (_I, [_P, C.POINTER(HParams)]),
    "q3a_encoder_n_tokens": (_I, [_I, _I]),
    "q3a_encoder_encode_batch": (_I, [_P, _P, _I, _P, _I, _I, _P, _P]),
    "q3a_encoder_encode_device": (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P]),
    "q3a_encode_pcm_batch": (_I, [_P, _P, _P, _P, _I, _I, _P, _P]),
    "q3a_encode_pcm_submit": (_I, [_P, _P, _P, _P, _I, _I, _P, _P, C.POINTER(C.c_int64)]),
    "q3a_encode_wait": (_I, [_P, C.c_int64]),
    "q3a_encode_pcm16_batch": (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P]),
    "q3a_encode_pcm_device": (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P]),
    "q3a_encoder_set_profiling": (_I, [_P, _I]),
    "q3a_encoder_get_timings": (_I, [_P, C.POINTER(Timings)]),
    "q3a_encoder_device_bytes": (_I64, [_P]),
    "q3a_decoder_load": (_I, [C.c_char_p, _I, _I, C.POINTER(_P)]),
    "q3a_decoder_free": (None, [_P]),
    "q3a_decoder_get_hparams": (_I, [_P, _P]),
    "q3a_decoder_forward": (_I, [_P, _P, _I, _P, _I, _I, _P]),
    "q3a_decoder_forward_device": (_I, [_P, _P, _I, _P, _I, _I, _P, _P]),
    "q3a_plan_shards": (_I, [_P, _I, _I, _I, _P]),
    "q3a_pool_create": (_I, [C.c_char_p, _P, _I, _I, C.POINTER(_P)]),
    "q3a_pool_free": (None, [_P]),
    "q3a_pool_n_devices": (_I, [_P]),
    "q3a_pool_encode_pcm": (_I, [_P, _P, _P, _I, _I, _P, _P, _P, _P]),
    "q3a_host_alloc": (_I, [C.c_size_t, C.POINTER(_P)]),
    "q3a_host_free": (None, [_P]),
    "q3a_bind_thread_to_device": (_I, [_I]),
    "q3a_plan_host_call": (_I, [_P, _I, _I, _I, _P, _P, _P]),
    "q3a_test_gemm": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "q3a_test_attention": (_I, [_P, _P, _I, _I, _P, _I, _I, _P]),
    "q3a_test_attention_v": (_I, [_P, _P, _I, _I, _P, _I, _I, _I, _P]),
    "q3a_probe_mma_rate": (_I, [_I, _I, _I, _I, _I, _I, _P, _P]),
    "q3a_debug_check_guards": (_I, [_P, _P]),
    "q3a_debug_damage_guard": (_I, []),
}

_lib = None


def load(path: str = LIB_PATH, strict: bool = True) -> C.CDLL:
    """Load libq3a.so and declare every prototype.  Raises if the library or a symbol is missing
    (strict=False tolerates missing symbols; development only)."""
    global _lib
    if _lib is not None and path == LIB_PATH:
        return _lib
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} not built: run `python qwen3-asr.cpp_b200/build.py` (no CPU fallback)")
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        if not strict and not hasattr(lib, name):
            continue
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if path == LIB_PATH:
        _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != OK:
        raise Q3AError(rc, load().q3a_last_error().decode("utf-8", "replace"))


def _ptr_array(arrs) -> C.Array:
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def _int_array(vals) -> C.Array:
    return (C.c_int * len(vals))(*[int(v) for v in vals])


def _i64_array(vals) -> C.Array:
    return (C.c_int64 * len(vals))(*[int(v) for v in vals])


def generate_mel_filters(n_mels: int = 128, n_fft: int = 400, sr: int = 16000) -> np.ndarray:
    out = np.zeros((n_mels, n_fft // 2 + 1), dtype=np.float32)
    check(load().q3a_generate_mel_filters(out.ctypes.data, n_mels, n_fft, sr))
    return out


class Frontend:
    """q3a_frontend: the log-mel frontend resident on one GPU."""

    def __init__(self, device: int = 0, filters: np.ndarray | None = None):
        self.lib = load()
        self.h = C.c_void_p()
        if filters is not None:
            filters = np.ascontiguousarray(filters, dtype=np.float32)
            check(self.lib.q3a_frontend_create(device, filters.ctypes.data, filters.shape[0], filters.shape[1],
                                               C.byref(self.h)))
            self.n_mel = filters.shape[0]
        else:
            check(self.lib.q3a_frontend_create(device, None, 0, 0, C.byref(self.h)))
            self.n_mel = 128

    def close(self):
        if self.h:
            self.lib.q3a_frontend_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def mel_batch(self, clips: list[np.ndarray]) -> list[np.ndarray]:
        """Host PCM clips -> host [n_mel][n_len] arrays (log_mel_spectrogram semantics)."""
        clips = [np.ascontiguousarray(c, dtype=np.float32) for c in clips]
        n = [len(c) for c in clips]
        outs = [np.zeros((self.n_mel, k // 160), dtype=np.float32) for k in n]
        n_len = (C.c_int * len(clips))()
        check(self.lib.q3a_mel_batch(self.h, _ptr_array(clips), _int_array(n), len(clips), _ptr_array(outs), n_len))
        for o, k in zip(outs, n_len):
            assert o.shape[1] == k
        return outs

    def mel_batch_device(self, pcm_ptr: int, pcm_off, n_samples, mel_ptr: int, mel_off, max_ptr: int = 0,
                         finalize: bool = True, stream: int = 0) -> None:
        check(self.lib.q3a_mel_batch_device(self.h, pcm_ptr, _i64_array(pcm_off), _int_array(n_samples),
                                            len(n_samples), mel_ptr, _i64_array(mel_off), max_ptr or None,
                                            1 if finalize else 0, stream or None))


class Encoder:
    """q3a_encoder: resident weights + workspace of one audio tower on one GPU."""

    def __init__(self, gguf_path: str, device: int = 0, operand: int = OP_BF16):
        self.lib = load()
        self.h = C.c_void_p()
        check(self.lib.q3a_encoder_load(gguf_path.encode(), device, operand, C.byref(self.h)))
        self.hp = HParams()
        check(self.lib.q3a_encoder_get_hparams(self.h, C.byref(self.hp)))

    def close(self):
        if self.h:
            self.lib.q3a_encoder_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def n_tokens(self, n_frames: int, mode: int = MODE_ASR) -> int:
        return self.lib.q3a_encoder_n_tokens(n_frames, mode)

    def out_width(self, mode: int) -> int:
        return self.hp.d_model if mode == MODE_CONV_ONLY else self.hp.out_dim

    def encode_batch(self, mels: list[np.ndarray], mode: int = MODE_ASR) -> list[np.ndarray]:
        mels = [np.ascontiguousarray(m, dtype=np.float32) for m in mels]
        n_mel = mels[0].shape[0] if mels else self.hp.n_mel_bins
        nf = [m.shape[1] for m in mels]
        outs = [np.zeros((max(self.n_tokens(f, mode), 0), self.out_width(mode)), dtype=np.float32) for f in nf]
        n_tok = (C.c_int * len(mels))()
        check(self.lib.q3a_encoder_encode_batch(self.h, _ptr_array(mels), n_mel, _int_array(nf), len(mels), mode,
                                                _ptr_array(outs), n_tok))
        for o, k in zip(outs, n_tok):
            assert o.shape[0] == k, (o.shape, k)
        return outs

    def encode_pcm_batch(self, fe: Frontend, clips: list[np.ndarray], mode: int = MODE_ASR) -> list[np.ndarray]:
        clips = [np.ascontiguousarray(c, dtype=np.float32) for c in clips]
        n = [len(c) for c in clips]
        outs = [np.zeros((max(self.n_tokens(k // 160, mode), 0), self.out_width(mode)), dtype=np.float32) for k in n]
        n_tok = (C.c_int * len(clips))()
        check(self.lib.q3a_encode_pcm_batch(fe.h, self.h, _ptr_array(clips), _int_array(n), len(clips), mode,
                                            _ptr_array(outs), n_tok))
        return outs

    def encode_pcm16_batch(self, fe: Frontend, clips: list[np.ndarray], n_channels: int = 1,
                           mode: int = MODE_ASR) -> list[np.ndarray]:
        """clips: int16 arrays of n_frames * n_channels interleaved samples (a WAV file's data chunk)"""
        clips = [np.ascontiguousarray(c, dtype=np.int16).reshape(-1) for c in clips]
        n = [len(c) // n_channels for c in clips]
        outs = [np.zeros((max(self.n_tokens(k // 160, mode), 0), self.out_width(mode)), dtype=np.float32) for k in n]
        n_tok = (C.c_int * len(clips))()
        check(self.lib.q3a_encode_pcm16_batch(fe.h, self.h, _ptr_array(clips), _int_array(n), n_channels, len(clips), mode,
                                              _ptr_array(outs), n_tok))
        return outs

    def encode_pcm_device(self, fe: Frontend, pcm_ptr: int, pcm_off, n_samples, out_ptr: int, mode: int = MODE_ASR,
                          stream: int = 0) -> list[int]:
        tok_off = (C.c_int64 * (len(n_samples) + 1))()
        check(self.lib.q3a_encode_pcm_device(fe.h, self.h, pcm_ptr, _i64_array(pcm_off), _int_array(n_samples),
                                             len(n_samples), mode, out_ptr, tok_off, stream or None))
        return list(tok_off)

    def encode_device(self, mel_ptr: int, mel_off, n_frames, out_ptr: int, mode: int = MODE_ASR, max_ptr: int = 0,
                      stream: int = 0) -> list[int]:
        tok_off = (C.c_int64 * (len(n_frames) + 1))()
        check(self.lib.q3a_encoder_encode_device(self.h, mel_ptr, _i64_array(mel_off), max_ptr or None,
                                                 _int_array(n_frames), len(n_frames), mode, out_ptr, tok_off,
                                                 stream or None))
        return list(tok_off)

    def set_profiling(self, on: bool) -> None:
        check(self.lib.q3a_encoder_set_profiling(self.h, 1 if on else 0))

    def timings(self) -> Timings:
        t = Timings()
        check(self.lib.q3a_encoder_get_timings(self.h, C.byref(t)))
        return t

    def device_bytes(self) -> int:
        return int(self.lib.q3a_encoder_device_bytes(self.h))


class DecoderHParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_layers", "hidden", "n_heads", "n_kv_heads", "head_dim", "intermediate", "vocab",
                                         "n_classes")] + [("rms_eps", C.c_float), ("rope_theta", C.c_float)]


class Decoder:
    """q3a_decoder: the forced aligner's NAR decoder (prefill over the whole prompt, classification head)."""

    def __init__(self, gguf_path: str, device: int = 0, operand: int = OP_BF16):
        self.lib = load()
        self.h = C.c_void_p()
        check(self.lib.q3a_decoder_load(gguf_path.encode(), device, operand, C.byref(self.h)))
        self.hp = DecoderHParams()
        check(self.lib.q3a_decoder_get_hparams(self.h, C.byref(self.hp)))

    def close(self):
        if self.h:
            self.lib.q3a_decoder_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def forward(self, tokens, audio: np.ndarray | None = None, audio_start: int = -1) -> np.ndarray:
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.zeros((len(tok), self.hp.n_classes), dtype=np.float32)
        if audio is not None:
            audio = np.ascontiguousarray(audio, dtype=np.float32)
        check(self.lib.q3a_decoder_forward(self.h, tok.ctypes.data, len(tok), audio.ctypes.data if audio is not None else None,
                                           len(audio) if audio is not None else 0, audio_start, out.ctypes.data))
        return out

    def forward_device(self, tokens, audio_ptr: int, n_audio: int, audio_start: int, logits_ptr: int, stream: int = 0) -> None:
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        check(self.lib.q3a_decoder_forward_device(self.h, tok.ctypes.data, len(tok), audio_ptr or None, n_audio, audio_start,
                                                  logits_ptr, stream or None))


class Pool:
    """q3a_pool: one replica + one worker thread per listed device (a device may repeat); corpus in, features out in
    utterance order."""

    def __init__(self, gguf_path: str, devices, operand: int = OP_BF16):
        self.lib = load()
        self.h = C.c_void_p()
        self.devices = list(devices)
        check(self.lib.q3a_pool_create(gguf_path.encode(), _int_array(self.devices), len(self.devices), operand, C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.q3a_pool_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode_pcm(self, clips, mode: int = MODE_ASR, out_dim: int = 1024, outs=None):
        """-> (features per clip, shard of every clip, wall ms of every replica).  `outs`: preallocated arrays to fill."""
        clips = [np.ascontiguousarray(c, dtype=np.float32) for c in clips]
        n = [len(c) for c in clips]
        lib = self.lib
        if outs is None:
            outs = [np.zeros((max(lib.q3a_encoder_n_tokens(k // 160, mode), 0), out_dim), dtype=np.float32) for k in n]
        n_tok = (C.c_int * max(len(clips), 1))()
        shard = (C.c_int * max(len(clips), 1))()
        ms = (C.c_float * len(self.devices))()
        check(lib.q3a_pool_encode_pcm(self.h, _ptr_array(clips), _int_array(n), len(clips), mode, _ptr_array(outs), n_tok, shard, ms))
        for o, k in zip(outs, n_tok):
            assert o.shape[0] == k, (o.shape, k)
        return outs, list(shard)[:len(clips)], list(ms)


def plan_host_call(n_samples, mode: int, out_width: int = 1024) -> tuple[list[int], list[int], list[int]]:
    """(pass, upload part, download part) of every clip of a q3a_encode_pcm_batch call"""
    n = len(n_samples)
    a, b, c = (C.c_int * max(n, 1))(), (C.c_int * max(n, 1))(), (C.c_int * max(n, 1))()
    check(load().q3a_plan_host_call(_int_array(n_samples), n, mode, out_width, a, b, c))
    return list(a)[:n], list(b)[:n], list(c)[:n]


def plan_shards(n_frames, mode: int, n_shards: int) -> list[int]:
    out = (C.c_int * len(n_frames))()
    check(load().q3a_plan_shards(_int_array(n_frames), len(n_frames), mode, n_shards, out))
    return list(out)


def test_gemm(a, w, bias, out, operand: int, block_n: int = 0) -> None:
    """torch CUDA tensors: out[M][N] f32 = a[M][K] @ w[N][K]^T + bias."""
    M, K = a.shape
    N = w.shape[0]
    check(load().q3a_test_gemm(a.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None,
                               out.data_ptr(), M, N, K, operand, block_n, None))


def test_attention(qkv, out, n_heads: int, cu_seqlens, operand: int, version: int = 0, stream: int = 0) -> None:
    """version: 0 = what the encoder runs, 1 / 2 = kernel generation (include/q3a.h)"""
    T = qkv.shape[0]
    cu = _int_array(cu_seqlens)
    check(load().q3a_test_attention_v(qkv.data_ptr(), out.data_ptr(), T, n_heads, cu, len(cu_seqlens) - 1, operand,
                                      version, stream or None))


def check_guards() -> tuple[int, int]:
    """(bands, damaged) over every live device buffer of the library; needs Q3A_GUARD=1 in the environment before the
    first allocation (include/q3a.h: q3a_debug_check_guards)."""
    n, bad = C.c_int(0), C.c_int(0)
    check(load().q3a_debug_check_guards(C.byref(n), C.byref(bad)))
    return n.value, bad.value
