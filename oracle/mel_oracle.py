"""TEST INFRASTRUCTURE — CPU oracle for the log-mel frontend.  Never on the product path.

Two checkers:

* `RefMel` — ctypes binding to oracle/_ref/libref_mel.so, i.e. the reference's own
  translation unit /root/reference/src/mel_spectrogram.cpp compiled where it lies
  (oracle/Makefile).  Slow (f64 O(N^2) DFT, mel_spectrogram.cpp:569-600) but it *is*
  the reference.
* `log_mel_numpy` / `mel_filters_numpy` — a numpy restatement (f64 rfft instead of the
  naive DFT) used for bulk checks; pinned against `RefMel` by tests/test_oracle_mel.py
  and against tests/golden/mel_*.npz (made by tests/golden/make_golden.py from RefMel).

Parity status: PINNED (against the reference code itself, run in this container).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

N_FFT = 400          # mel_spectrogram.h:11
HOP = 160            # mel_spectrogram.h:12
N_MELS = 128         # mel_spectrogram.h:13
N_BINS = 201         # mel_spectrogram.h:15
SR = 16000

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libref_mel.so")


class RefMel:
    """The compiled reference TU behind a C shim (oracle/ref_mel_shim.cpp)."""

    def __init__(self, path: str = REF_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
        self.lib = ctypes.CDLL(path)
        self.lib.ref_generate_mel_filters.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        self.lib.ref_log_mel.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
        self.lib.ref_log_mel.restype = ctypes.c_int

    def filters(self, n_mels: int = N_MELS, n_fft: int = N_FFT, sr: int = SR) -> np.ndarray:
        out = np.zeros((n_mels, n_fft // 2 + 1), dtype=np.float32)
        self.lib.ref_generate_mel_filters(out.ctypes.data, n_mels, n_fft, sr)
        return out

    def log_mel(self, pcm: np.ndarray, filters: np.ndarray) -> np.ndarray:
        pcm = np.ascontiguousarray(pcm, dtype=np.float32)
        filters = np.ascontiguousarray(filters, dtype=np.float32)
        n_len = len(pcm) // HOP
        out = np.zeros((filters.shape[0], max(n_len, 1)), dtype=np.float32)
        got = self.lib.ref_log_mel(pcm.ctypes.data, len(pcm), filters.ctypes.data, filters.shape[0],
                                   filters.shape[1], out.ctypes.data, out.size)
        if got != n_len:
            raise RuntimeError(f"reference returned n_len={got}, expected {n_len}")
        return out[:, :n_len].copy()


def mel_filters_numpy(n_mels: int = N_MELS, n_fft: int = N_FFT, sr: int = SR) -> np.ndarray:
    """Triangular HTK-scale / Slaney-normalised bank in *float32* arithmetic.

    Follows generate_mel_filters (mel_spectrogram.cpp:361-415): hz->mel by
    2595*log10(1+hz/700) in float (:353-355), inverse (:357-359), n_mels+2 equally spaced
    mel points (:377-380), bin position (n_fft+1)*hz/sr (:388-391), inclusive <= on both
    sides of the triangle (:399-403), Slaney normalisation 2/(hz[m+2]-hz[m]) (:409-414).
    """
    f32 = np.float32
    n_bins = 1 + n_fft // 2
    mel_min = f32(2595.0) * np.log10(f32(1.0) + f32(0.0) / f32(700.0), dtype=f32)
    mel_max = f32(2595.0) * np.log10(f32(1.0) + f32(sr / 2.0) / f32(700.0), dtype=f32)
    i = np.arange(n_mels + 2, dtype=np.int32)
    # C: mel_min + (mel_max - mel_min) * i / (n_mels + 1), all float
    mel_pts = (mel_min + ((mel_max - mel_min) * i.astype(f32)) / f32(n_mels + 1)).astype(f32)
    hz_pts = (f32(700.0) * (np.power(f32(10.0), mel_pts / f32(2595.0), dtype=f32) - f32(1.0))).astype(f32)
    bin_pts = (f32(n_fft + 1) * hz_pts / f32(sr)).astype(f32)
    k = np.arange(n_bins, dtype=f32)
    bank = np.zeros((n_mels, n_bins), dtype=f32)
    for m in range(n_mels):
        left, center, right = bin_pts[m], bin_pts[m + 1], bin_pts[m + 2]
        up = (k >= left) & (k <= center)
        dn = (~up) & (k >= center) & (k <= right)
        with np.errstate(divide="ignore", invalid="ignore"):
            wu = ((k - left) / (center - left)).astype(f32)
            wd = ((right - k) / (right - center)).astype(f32)
        row = np.where(up, wu, np.where(dn, wd, f32(0.0))).astype(f32)
        enorm = f32(2.0) / (hz_pts[m + 2] - hz_pts[m])
        bank[m] = (row * enorm).astype(f32)
    return bank


def hann_window() -> np.ndarray:
    """Periodic Hann, f64 (GlobalCache, mel_spectrogram.cpp:28-58)."""
    n = np.arange(N_FFT, dtype=np.float64)
    return 0.5 * (1.0 - np.cos(2.0 * np.pi * n / N_FFT))


def pad_reflect(pcm: np.ndarray) -> np.ndarray:
    """mel_spectrogram.cpp:493-515: left s[200..1], right s[n-2..n-201], zero where out of range."""
    n = len(pcm)
    pad = N_FFT // 2
    out = np.zeros(n + 2 * pad, dtype=np.float64)
    out[pad:pad + n] = pcm
    i = np.arange(pad)
    src = pad - i
    ok = src < n
    out[i[ok]] = pcm[src[ok]]
    src = n - 2 - i
    ok = src >= 0
    out[n + pad + i[ok]] = pcm[src[ok]]
    return out


def log_mel_raw_numpy(pcm: np.ndarray, filters: np.ndarray) -> np.ndarray:
    """log10(max(mel energy, 1e-10)) for all total_frames = n//160 + 1 frames, f64 [n_mel][frames]."""
    x = pad_reflect(np.asarray(pcm, dtype=np.float64))
    total = (len(x) - N_FFT) // HOP + 1
    idx = np.arange(N_FFT)[None, :] + HOP * np.arange(total)[:, None]
    frames = x[idx] * hann_window()[None, :]
    spec = np.fft.rfft(frames, axis=1)
    power = spec.real ** 2 + spec.imag ** 2                       # [frames][201]
    melE = power @ filters.astype(np.float64).T                    # [frames][n_mel]
    return np.log10(np.maximum(melE, 1e-10)).T


def log_mel_numpy(pcm: np.ndarray, filters: np.ndarray) -> np.ndarray:
    """Restatement of log_mel_spectrogram (mel_spectrogram.cpp:484-628) → f32 [n_mel][n_len].

    n_len = total_frames-1 (:517-520); the max for the clamp is taken over [0,n_len) only
    (:604-613); clamp to max-8 then (v+4)/4 (:615-625).
    """
    raw = log_mel_raw_numpy(pcm, filters)
    n_len = raw.shape[1] - 1
    v = raw[:, :n_len]
    if n_len == 0:
        return np.zeros((raw.shape[0], 0), dtype=np.float32)
    mmax = v.max() - 8.0
    return ((np.maximum(v, mmax) + 4.0) / 4.0).astype(np.float32)
