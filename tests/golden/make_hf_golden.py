"""Golden vectors from the UPSTREAM implementation of the audio tower: transformers' Qwen3OmniMoeAudioEncoder, the class
whose parameters the reference's converter maps one-to-one onto the GGUF tensors (reference
scripts/convert_hf_to_gguf.py:48-103: thinker.audio_tower.conv2d{1,2,3}, conv_out, layers.N.self_attn.{q,k,v,out}_proj,
self_attn_layer_norm, fc1, fc2, final_layer_norm, ln_post, proj1, proj2).  Its forward is the graph
ForcedAligner::encode_audio restates (zero-padded 100-frame chunks, per-chunk sinusoidal positions, 104-token attention
windows, forced_aligner.cpp:591-924); AudioEncoder::encode differs from it only in the unpadded tail chunk and the
missing window mask (audio_encoder.cpp:312-601), which tests/test_oracle_encoder.py covers separately.

This pins the GRAPH of oracle/encoder_oracle.py (weight layouts, feature order, chunking, positions, windows, layer
order) to an implementation that is not ours.  It does not pin ggml's roundings: the tower is evaluated in float64 with
the exact erf GELU that F.gelu computes, and the oracle is compared in its ideal mode with the same GELU.

Two attention readings exist upstream and both are recorded: with eager attention this transformers version passes no
mask to the layers (every token attends to every token -- what AudioEncoder::encode does, audio_encoder.cpp:466-486),
with flash attention the cu_seqlens of 104-token windows apply (what ForcedAligner::encode_audio does,
forced_aligner.cpp:737-766).  "full_*" is the former, "win_*" the latter (through the block mask the class itself builds).

Real widths (hf_audio_tower_wide.npz): the same generator on 2-layer towers with the production widths -- ASR
(d = 896, 14 heads, ffn 3584, conv 480: GEMM tiles BLOCK_N 224 / 256 / 240) and aligner (d = 1024, 16 heads, ffn 4096) --
so that the CUDA tile paths those widths select are pinned to upstream outputs too, not only to our own restatement.

    python tests/golden/make_hf_golden.py        # needs transformers (>= 4.57) and torch, CPU only; writes both .npz files
"""
from __future__ import annotations

import os
import re
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import q3a_b200  # noqa: E402,F401
from q3a_b200 import synth  # noqa: E402
from oracle import mel_oracle as MO  # noqa: E402

# inverse of the converter's TENSOR_MAP / AUDIO_LAYER_PATTERNS (convert_hf_to_gguf.py:48-103), audio tower only
FIXED = {
    "audio.encoder.conv1": "conv2d1", "audio.encoder.conv2": "conv2d2", "audio.encoder.conv3": "conv2d3",
    "audio.encoder.conv_out": "conv_out", "audio.encoder.ln_post": "ln_post",
    "audio.encoder.proj1": "proj1", "audio.encoder.proj2": "proj2",
}
LAYER = {
    "attn_q": "self_attn.q_proj", "attn_k": "self_attn.k_proj", "attn_v": "self_attn.v_proj", "attn_out": "self_attn.out_proj",
    "attn_norm": "self_attn_layer_norm", "ffn_norm": "final_layer_norm", "ffn_up": "fc1", "ffn_down": "fc2",
}


def hf_name(gguf_name: str) -> str:
    base, kind = gguf_name.rsplit(".", 1)
    if base in FIXED:
        return f"{FIXED[base]}.{kind}"
    m = re.fullmatch(r"audio\.encoder\.blk\.(\d+)\.(\w+)", base)
    return f"layers.{m.group(1)}.{LAYER[m.group(2)]}.{kind}"


def build_tower(shape: synth.EncoderShape, weights: dict[str, np.ndarray]):
    from transformers.models.qwen3_omni_moe.configuration_qwen3_omni_moe import Qwen3OmniMoeAudioEncoderConfig
    from transformers.models.qwen3_omni_moe.modeling_qwen3_omni_moe import Qwen3OmniMoeAudioEncoder
    cfg = Qwen3OmniMoeAudioEncoderConfig(
        num_mel_bins=128, encoder_layers=shape.n_layers, encoder_attention_heads=shape.n_heads,
        encoder_ffn_dim=shape.ffn_dim, d_model=shape.d_model, dropout=0.0, attention_dropout=0.0,
        activation_function="gelu", activation_dropout=0.0, scale_embedding=False, max_source_positions=1500,
        n_window=50, output_dim=shape.out_dim, n_window_infer=800, conv_chunksize=500,
        downsample_hidden_size=shape.conv_ch)
    cfg._attn_implementation = "eager"
    tower = Qwen3OmniMoeAudioEncoder(cfg).double().eval()
    sd = {hf_name(k): torch.from_numpy(v.astype(np.float64)) for k, v in weights.items()}
    missing, unexpected = tower.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("positional_embedding" in m for m in missing), missing      # the sinusoid table is a buffer, not a weight
    return tower


# 2-layer towers at the production widths (tests/test_encoder_gpu.py builds the same weights from the same seeds)
WIDE = {
    "asr": (synth.EncoderShape("asr-2l", 2, 896, 14, 3584), 21),
    "aligner": (synth.EncoderShape("aligner-2l", 2, 1024, 16, 4096), 22),
}
WIDE_CLIPS = ((114, 16000 * 2 + 4321), (115, 16000 * 11 + 77), (116, 100 * 160 * 3))


def run_tower(shape, weights, clips, prefix=""):
    tower = build_tower(shape, weights)
    filt = np.load(os.path.join(HERE, "mel_ref.npz"))["filters"]           # the reference's own filterbank
    out = {}
    for seed, n in clips:
        mel = MO.log_mel_numpy(synth.synth_audio(seed, n), filt)
        captured = {}
        hook = tower.layers[0].register_forward_pre_hook(
            lambda mod, args, kwargs: captured.update(h=args[0].clone(), cu=args[1].clone()), with_kwargs=True)
        with torch.no_grad():
            # (1) as shipped with eager attention: no mask is passed to the layers, every token attends to every token
            y = tower(torch.from_numpy(mel.astype(np.float64)), feature_lens=torch.tensor([mel.shape[1]]))
            hook.remove()
            # (2) the cu_seqlens (104-token windows) honoured, as the flash-attention varlen path does, through the
            #     block mask the class builds for exactly this purpose (_prepare_attention_mask)
            h, cu = captured["h"], captured["cu"]
            mask = tower._prepare_attention_mask(h, cu)
            assert mask is not None
            for layer in tower.layers:
                h = layer(h, cu, attention_mask=mask)[0]
            h = tower.proj2(tower.act(tower.proj1(tower.ln_post(h))))
        if not prefix or mel.shape[1] % 100 == 0:      # wide file: the eager reading only where ASR mode is the same graph
            out[f"{prefix}full_{seed}_{n}"] = y.last_hidden_state.numpy().astype(np.float32)
        out[f"{prefix}win_{seed}_{n}"] = h.numpy().astype(np.float32)
        out[f"{prefix}cu_{seed}_{n}"] = cu.numpy().astype(np.int64)
        print(prefix, seed, n, mel.shape, out[f"{prefix}win_{seed}_{n}"].shape, cu.tolist())
    return out


def main() -> None:
    shape = synth.tiny_shape()
    out = run_tower(shape, synth.random_weights(shape, 7), ((104, 16000 * 2 + 4321), (105, 16000 * 11 + 77), (106, 100 * 160 * 3)))
    np.savez_compressed(os.path.join(HERE, "hf_audio_tower_tiny.npz"), **out)
    wide = {}
    for name, (shape, seed) in WIDE.items():
        wide.update(run_tower(shape, synth.random_weights(shape, seed), WIDE_CLIPS, prefix=name + "."))
    np.savez_compressed(os.path.join(HERE, "hf_audio_tower_wide.npz"), **wide)


if __name__ == "__main__":
    main()
