"""TEST INFRASTRUCTURE — CPU restatement of the forced aligner's NAR decoder forward.  Never on the product path.

Parity status: UNPINNED against ggml (the submodule is absent, see encoder_oracle.py); the graph follows
ForcedAligner::build_decoder_graph (reference forced_aligner.cpp:926-1085) line by line:
  token embedding rows with the audio features spliced in (:964-993), per layer RMSNorm * attn_norm (:1012-1013),
  q/k/v matmuls (:1015-1017), per-head RMSNorm * q_norm / k_norm (:1023-1030), NEOX RoPE with position = index
  (:1032-1038, :1119-1125), causal attention k <= q with scale 1/sqrt(head_dim), f32 accumulation, query heads mapped
  onto kv heads in groups (:1040-1046, mask :1128-1139), output projection + residual (:1048-1050), RMSNorm * ffn_norm,
  silu(gate) * up, down projection + residual (:1052-1064), final RMSNorm * output_norm and the classification head
  (:1067-1076).  ggml's CPU matmul rounds the f32 activation row to f16 when the weight is F16 (same switch as the
  encoder oracle); everything else is f32.
"""
from __future__ import annotations

import numpy as np


def _rms(x, w, eps, dt):
    x = x.astype(dt)
    r = x / np.sqrt((x.astype(np.float64) ** 2).mean(axis=-1, keepdims=True).astype(dt) + dt(eps))
    return (r * w.astype(dt)).astype(dt)


def _mm(x, w, round_act, dt):
    if round_act:
        x = x.astype(np.float16)
    return (x.astype(dt) @ w.astype(dt).T).astype(dt)


def _rope_neox(x, theta):
    """x: [n][heads][hd]; pairs (i, i + hd/2), angle = pos * theta^(-2i/hd) (ggml_rope_ext, GGML_ROPE_TYPE_NEOX)"""
    n, _, hd = x.shape
    half = hd // 2
    inv = theta ** (-2.0 * np.arange(half, dtype=np.float64) / hd)
    ang = np.arange(n, dtype=np.float64)[:, None] * inv[None, :]
    c, s = np.cos(ang)[:, None, :].astype(x.dtype), np.sin(ang)[:, None, :].astype(x.dtype)
    a, b = x[..., :half], x[..., half:]
    return np.concatenate([a * c - b * s, a * s + b * c], axis=-1)


def forward(tokens, W: dict, shape, audio=None, audio_start: int = -1, ideal: bool = False) -> np.ndarray:
    """tokens int32[n]; audio f32[n_audio][hidden] or None -> logits f32[n][n_classes]"""
    dt = np.float64 if ideal else np.float32
    ra = not ideal
    n = len(tokens)
    x = W["token_embd.weight"][np.asarray(tokens)].astype(dt)
    if audio is not None and len(audio) > 0 and audio_start >= 0 and audio_start + len(audio) <= n:
        x[audio_start:audio_start + len(audio)] = audio.astype(dt)
    H, Hkv, hd = shape.n_heads, shape.n_kv_heads, shape.head_dim
    mask = np.where(np.arange(n)[None, :] <= np.arange(n)[:, None], 0.0, -np.inf).astype(dt)
    for l in range(shape.n_layers):
        p = f"blk.{l}"
        cur = _rms(x, W[f"{p}.attn_norm.weight"], shape.rms_eps, dt)
        q = _mm(cur, W[f"{p}.attn_q.weight"], ra, dt).reshape(n, H, hd)
        k = _mm(cur, W[f"{p}.attn_k.weight"], ra, dt).reshape(n, Hkv, hd)
        v = _mm(cur, W[f"{p}.attn_v.weight"], ra, dt).reshape(n, Hkv, hd)
        q = _rope_neox(_rms(q, W[f"{p}.attn_q_norm.weight"], shape.rms_eps, dt), shape.rope_theta)
        k = _rope_neox(_rms(k, W[f"{p}.attn_k_norm.weight"], shape.rms_eps, dt), shape.rope_theta)
        grp = H // Hkv
        out = np.empty((n, H, hd), dt)
        for h in range(H):
            s = (q[:, h] @ k[:, h // grp].T) * dt(1.0 / np.sqrt(hd)) + mask
            s = s - s.max(axis=-1, keepdims=True)
            pr = np.exp(s)
            pr = pr / pr.sum(axis=-1, keepdims=True)
            out[:, h] = pr @ v[:, h // grp]
        x = _mm(out.reshape(n, H * hd), W[f"{p}.attn_output.weight"], ra, dt) + x
        cur = _rms(x, W[f"{p}.ffn_norm.weight"], shape.rms_eps, dt)
        g = _mm(cur, W[f"{p}.ffn_gate.weight"], ra, dt)
        u = _mm(cur, W[f"{p}.ffn_up.weight"], ra, dt)
        act = (g / (1.0 + np.exp(-g))) * u
        x = _mm(act, W[f"{p}.ffn_down.weight"], ra, dt) + x
    cur = _rms(x, W["output_norm.weight"], shape.rms_eps, dt)
    y = _mm(cur, W["output.weight"], ra, dt)
    if "output.bias" in W:
        y = y + W["output.bias"].astype(dt)
    return y.astype(np.float32)
