"""Run in a child process with Q3A_GUARD=1 (tests/test_encoder_gpu.py::test_no_kernel_writes_outside_its_buffers):
every device buffer the library owns is then exact-size between two 4-KiB pattern bands (include/q3a.h:
q3a_debug_check_guards).  The battery drives every entry point over ragged, empty, tail-chunk and multi-part inputs,
re-using the same objects so buffers are both grown and re-used, and prints one JSON line with the band count and
the number of damaged bands.

    Q3A_GUARD=1 python tests/guard_battery.py tiny_f16.gguf tiny_q8.gguf [asr_06b_f16.gguf]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import q3a_b200  # noqa: E402,F401
from q3a_b200 import binding as B, synth  # noqa: E402


def main(argv):
    assert os.environ.get("Q3A_GUARD") == "1"
    B.load()
    fe = B.Frontend(0)
    cases = 0
    tally = {"bands": 0, "damaged": 0, "msg": ""}

    def chk():                                       # before an object goes away: its bands are forgotten on release
        bands, bad = B.check_guards()
        tally["bands"] += bands
        tally["damaged"] += bad
        if bad:
            tally["msg"] += B.load().q3a_last_error().decode()
    # mel alone: every edge length of tests/test_mel_gpu.py in one ragged batch, then one by one (buffers re-used)
    lens = [0, 1, 159, 160, 161, 200, 201, 399, 400, 401, 5120, 5121, 5279, 48000 + 77, 480000, 501234]
    clips = [synth.synth_audio(40 + i, n) for i, n in enumerate(lens)]
    fe.mel_batch(clips)
    for c in clips:
        fe.mel_batch([c]); cases += 1
    # tiny tower: both files, both operand types, every mode, ragged clips with tail chunks of 1..99 frames
    ragged = [16000 * 3 + 777, 16000 + 5, 100 * 160, 237 * 160, 160, 37 * 160 + 3, 0, 159, 16000 * 12 + 1]
    pcm = [synth.synth_audio(60 + i, n) for i, n in enumerate(ragged)]
    mels = fe.mel_batch(pcm)
    for path in argv[:2]:
        for opd in (B.OP_BF16, B.OP_F16):
            enc = B.Encoder(path, 0, opd)
            for mode in (B.MODE_ASR, B.MODE_ALIGNER):
                enc.encode_pcm_batch(fe, pcm, mode)
                enc.encode_batch(mels, mode)
                enc.encode_pcm_batch(fe, pcm[:2], mode)              # smaller call on grown buffers
                cases += 3
            for mode in (B.MODE_CONV_ONLY, B.MODE_NO_CHUNK):
                for m in mels:
                    if m.shape[1] > 0:
                        enc.encode_batch([m], mode); cases += 1
            pcm16 = [np.round(c * 32767).astype(np.int16) for c in pcm]
            enc.encode_pcm16_batch(fe, pcm16, 1, B.MODE_ASR)
            enc.encode_pcm16_batch(fe, [np.repeat(c, 2) for c in pcm16], 2, B.MODE_ALIGNER)
            cases += 2
            chk()
            enc.close()
    # the real shape: a call big enough for the copy/compute partitioning (4 upload parts, 2 download parts), a
    # 92-s single sequence (10 attention tiles per head, 19 key blocks), and the windowed mode on the same clips
    if len(argv) > 2:
        enc = B.Encoder(argv[2], 0, B.OP_BF16)
        rng = np.random.default_rng(7)
        big = [synth.synth_audio(700 + i, int(rng.integers(2 * 16000, 40 * 16000))) for i in range(40)]
        big[11] = big[11][:0]
        enc.encode_pcm_batch(fe, big, B.MODE_ASR)
        enc.encode_pcm_batch(fe, big[:7], B.MODE_ALIGNER)
        enc.encode_pcm_batch(fe, [synth.synth_audio(300, 1472000)], B.MODE_ASR)
        enc.encode_batch(fe.mel_batch(big[:5]), B.MODE_ASR)
        cases += 4
        chk()
        enc.close()
    chk()
    # the detector itself: one byte past the end of a live buffer must show up as exactly one damaged band
    B.check(B.load().q3a_debug_damage_guard())
    _, tally["selftest_damaged"] = B.check_guards()
    fe.close()
    print(json.dumps({"cases": cases, **tally}))
    return 1 if tally["damaged"] else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
