"""Generates the committed golden vectors.  Run in the build container (needs /root/reference):

    make -C oracle ref && python tests/golden/make_golden.py

* mel_ref.npz      — outputs of the REFERENCE's own translation unit (mel_spectrogram.cpp compiled
                     where it lies into oracle/_ref/libref_mel.so) on seeded synthetic clips: the
                     generated filterbank and log-mel of a 1.0 s, a 0.37 s (ragged) and a 2.5 s clip
                     with silent gaps.  These pin the numpy oracle and the CUDA path to the reference.
* encoder_tiny.npz — outputs of oracle/encoder_oracle.py (the CPU restatement; ggml is absent, so this
                     guards drift of the restatement, it does not pin ggml) for the tiny tower in both
                     stem/attention variants.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import q3a_b200  # noqa: E402,F401
from q3a_b200 import synth  # noqa: E402
from oracle import mel_oracle as MO, encoder_oracle as EO  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
MEL_CASES = ((101, 16000), (102, 5920 + 37), (103, 40000))


def main():
    ref = MO.RefMel()
    filt = ref.filters()
    out = {"filters": filt}
    for seed, n in MEL_CASES:
        pcm = synth.synth_audio(seed, n)
        out[f"mel_{seed}_{n}"] = ref.log_mel(pcm, filt)
    np.savez_compressed(os.path.join(HERE, "mel_ref.npz"), **out)

    shape = synth.tiny_shape()
    w = synth.random_weights(shape, 7)
    enc = {}
    for ftype in ("f16", "q8_0"):
        stored = {k: v for k, v in synth.stored_weights(w, ftype).items() if "#" not in k}
        W = EO.Weights(stored, ftype)
        pcm = synth.synth_audio(104, 16000 * 2 + 4321)
        mel = MO.log_mel_numpy(pcm, filt)
        for variant in ("asr", "aligner"):
            y = EO.encode(mel, W, n_layers=shape.n_layers, d_model=shape.d_model, n_heads=shape.n_heads, variant=variant)
            enc[f"{ftype}_{variant}"] = y.astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "encoder_tiny.npz"), **enc)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
