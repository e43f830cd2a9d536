"""Golden vectors for the forced aligner's NAR decoder from the UPSTREAM implementation of its layers: transformers'
Qwen3Model -- the class behind `thinker.model.*`, whose parameters the reference's converter maps one-to-one onto
token_embd / blk.N.{attn_norm, attn_q, attn_k, attn_v, attn_q_norm, attn_k_norm, attn_output, ffn_norm, ffn_gate, ffn_up,
ffn_down} / output_norm (reference scripts/convert_hf_to_gguf.py:73-75,105-125) -- evaluated in float64 on a tiny tower,
causal, with the audio rows spliced into the input embeddings the way ForcedAligner::build_decoder_graph does
(forced_aligner.cpp:964-993), followed by the classification head (output.weight, :1073).

This pins the GRAPH of oracle/decoder_oracle.py (weight layouts, RMSNorm placement, per-head q/k norm, NeoX RoPE, kv-head
grouping, SwiGLU, residuals) to an implementation that is not ours; it does not pin ggml's f16 roundings.

    python tests/golden/make_hf_decoder_golden.py     # needs transformers and torch, CPU only; writes hf_decoder_tiny.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import q3a_b200  # noqa: E402,F401
from q3a_b200 import synth  # noqa: E402

LAYER = {"attn_norm": "input_layernorm", "attn_q": "self_attn.q_proj", "attn_k": "self_attn.k_proj", "attn_v": "self_attn.v_proj",
         "attn_q_norm": "self_attn.q_norm", "attn_k_norm": "self_attn.k_norm", "attn_output": "self_attn.o_proj",
         "ffn_norm": "post_attention_layernorm", "ffn_gate": "mlp.gate_proj", "ffn_up": "mlp.up_proj", "ffn_down": "mlp.down_proj"}
CASES = ((5, 61, 23, 9), (6, 200, 130, 31), (7, 17, 0, -1))      # (seed, n_tokens, n_audio, audio_start)


def case_inputs(shape, seed, n, n_audio, audio_start):
    rng = np.random.default_rng(seed)
    tokens = rng.integers(0, shape.vocab, n).astype(np.int32)
    audio = (rng.standard_normal((n_audio, shape.hidden)) * 0.7).astype(np.float32) if n_audio else None
    return tokens, audio


def main() -> None:
    from transformers.models.qwen3.configuration_qwen3 import Qwen3Config
    from transformers.models.qwen3.modeling_qwen3 import Qwen3Model
    shape = synth.tiny_decoder_shape()
    w = synth.random_decoder_weights(shape, 3)
    cfg = Qwen3Config(vocab_size=shape.vocab, hidden_size=shape.hidden, intermediate_size=shape.intermediate,
                      num_hidden_layers=shape.n_layers, num_attention_heads=shape.n_heads, num_key_value_heads=shape.n_kv_heads,
                      head_dim=shape.head_dim, hidden_act="silu", max_position_embeddings=4096, rms_norm_eps=shape.rms_eps,
                      rope_theta=shape.rope_theta, attention_bias=False, attention_dropout=0.0, tie_word_embeddings=False,
                      use_sliding_window=False)
    cfg._attn_implementation = "eager"
    model = Qwen3Model(cfg).double().eval()
    sd = {}
    for k, v in w.items():
        t = torch.from_numpy(v.astype(np.float64))
        if k == "token_embd.weight":
            sd["embed_tokens.weight"] = t
        elif k == "output_norm.weight":
            sd["norm.weight"] = t
        elif k.startswith("blk."):
            _, l, name, kind = k.split(".")
            sd[f"layers.{l}.{LAYER[name]}.{kind}"] = t
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected and all("rotary" in m or "inv_freq" in m for m in missing), (missing, unexpected)
    head = torch.from_numpy(w["output.weight"].astype(np.float64))
    out = {}
    for seed, n, n_audio, audio_start in CASES:
        tokens, audio = case_inputs(shape, seed, n, n_audio, audio_start)
        with torch.no_grad():
            x = model.embed_tokens(torch.from_numpy(tokens.astype(np.int64)))[None]
            if audio is not None:
                x[0, audio_start:audio_start + n_audio] = torch.from_numpy(audio.astype(np.float64))
            h = model(inputs_embeds=x).last_hidden_state[0]
            logits = h @ head.T
        out[f"logits_{seed}"] = logits.numpy().astype(np.float32)
        print(seed, n, n_audio, audio_start, out[f"logits_{seed}"].shape, float(np.abs(out[f"logits_{seed}"]).max()))
    np.savez_compressed(os.path.join(HERE, "hf_decoder_tiny.npz"), **out)


if __name__ == "__main__":
    main()
