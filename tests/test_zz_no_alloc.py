"""q3a_mel_batch_device_noalloc -- the allocation-free entry point (static __device__ scratch, csrc/mel.cu) -- against the
existing q3a_mel_batch_device, through the C ABI.  Both run the same kernels on the same inputs, so the bar is the one
the repository's other same-kernel comparisons use: bit-identical.  The file sorts last on purpose;
it holds no state of its own, so it passes alone in a fresh process and after every other test."""
import numpy as np
import pytest
import torch

from q3a_b200 import binding as B, synth

pytestmark = pytest.mark.gpu


def _pack(clips):
    """clips packed the way every device entry point wants them: clip starts at a multiple of 64 samples"""
    offs, cur = [], 0
    for c in clips:
        offs.append(cur); cur += (len(c) + 63) // 64 * 64
    pcm = torch.zeros(cur + 64, device="cuda")
    for c, o in zip(clips, offs):
        if len(c):
            pcm[o:o + len(c)] = torch.from_numpy(c).cuda()
    moffs, cur = [], 0
    for c in clips:
        moffs.append(cur); cur += 128 * (len(c) // 160)
    return pcm, offs, moffs, cur


def _noalloc(lib, fe, pcm, offs, ns, mel, moffs, mx, finalize, stream=None):
    return lib.q3a_mel_batch_device_noalloc(fe.h, pcm.data_ptr(), B._i64_array(offs), B._int_array(ns), len(ns),
                                            mel.data_ptr(), B._i64_array(moffs), mx.data_ptr() if mx is not None else None,
                                            1 if finalize else 0, stream)


def test_noalloc_mel_matches_the_allocating_entry_point_on_two_calls_in_a_row(lib):
    fe = B.Frontend(0)
    # call 1: 5 loud ragged clips (one shorter than a frame, one empty); call 2: 3 much quieter clips of other lengths --
    # if the shared scratch kept call 1's per-clip maxima or descriptors, call 2's clamp (max - 8) and layout would be wrong
    batches = [
        [synth.synth_audio(40 + i, n) for i, n in enumerate((16000 * 3 + 1234, 16000, 100, 0, 16000 * 7 + 5))],
        [synth.synth_audio(60 + i, n) * np.float32(1e-3) for i, n in enumerate((16000 * 2 + 77, 16000 * 5, 4000))],
    ]
    staged = []
    for clips in batches:                                    # every torch allocation happens before the calls under test
        pcm, offs, moffs, total = _pack(clips)
        bufs = {k: (torch.full((total + 4,), float("nan"), device="cuda"), torch.full((len(clips),), float("nan"), device="cuda"))
                for k in ("ref_final", "ref_raw", "new_final", "new_raw")}
        staged.append((clips, pcm, offs, moffs, total, bufs))
    for clips, pcm, offs, moffs, total, bufs in staged:      # the existing entry point
        ns = [len(c) for c in clips]
        fe.mel_batch_device(pcm.data_ptr(), offs, ns, bufs["ref_final"][0].data_ptr(), moffs, bufs["ref_final"][1].data_ptr(), finalize=True)
        fe.mel_batch_device(pcm.data_ptr(), offs, ns, bufs["ref_raw"][0].data_ptr(), moffs, bufs["ref_raw"][1].data_ptr(), finalize=False)
    torch.cuda.synchronize()
    # one warm-up call resolves the scratch's address on this device and loads the set-up kernel; from here on the free
    # device memory must not move (nothing in the process allocates between the two readings)
    clips, pcm, offs, moffs, total, bufs = staged[0]
    B.check(_noalloc(lib, fe, pcm, offs, [len(c) for c in clips], bufs["new_final"][0], moffs, bufs["new_final"][1], True))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0 = torch.cuda.memory_allocated()
    for clips, pcm, offs, moffs, total, bufs in staged:      # two calls in a row (x 2 output modes), same stream
        ns = [len(c) for c in clips]
        B.check(_noalloc(lib, fe, pcm, offs, ns, bufs["new_final"][0], moffs, bufs["new_final"][1], True))
        B.check(_noalloc(lib, fe, pcm, offs, ns, bufs["new_raw"][0], moffs, bufs["new_raw"][1], False))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0 and torch.cuda.memory_allocated() == alloc0, "the allocation-free entry point changed the device's free memory"
    for clips, pcm, offs, moffs, total, bufs in staged:
        for mode in ("final", "raw"):
            want, want_mx = bufs[f"ref_{mode}"]
            got, got_mx = bufs[f"new_{mode}"]
            assert torch.isfinite(want[:total]).all()
            assert torch.equal(got[:total], want[:total]), f"{mode}: mel differs from q3a_mel_batch_device"
            # maxima by bit pattern: a clip without a single frame (100 and 0 samples here) has no maximum and both
            # entry points report the same NaN for it
            assert torch.isfinite(want_mx[[0, 1, -1]]).all()
            assert torch.equal(got_mx.view(torch.int32), want_mx.view(torch.int32)), f"{mode}: per-clip maxima differ"
            assert torch.isnan(got[total:]).all(), "wrote past the last clip's block"
    fe.close()


def test_noalloc_mel_rejects_more_clips_than_its_scratch_holds(lib):
    fe = B.Frontend(0)
    limit = lib.q3a_mel_noalloc_max_clips()
    assert limit == 128
    n = limit + 1
    pcm = torch.zeros(64 * n + 64, device="cuda")
    mel = torch.zeros(16, device="cuda")
    rc = _noalloc(lib, fe, pcm, [64 * i for i in range(n)], [0] * n, mel, [0] * n, None, True)
    assert rc == 1                                          # Q3A_ERR_ARG (include/q3a.h)
    msg = lib.q3a_last_error().decode()
    assert "128" in msg and "exceeds" in msg, msg
    # exactly the maximum is accepted: 128 one-second clips, checked against the allocating call
    clips = [synth.synth_audio(90, 16000) * np.float32(0.5 + 0.003 * i) for i in range(limit)]
    pcm, offs, moffs, total = _pack(clips)
    ns = [len(c) for c in clips]
    want, got = torch.zeros(total, device="cuda"), torch.zeros(total, device="cuda")
    fe.mel_batch_device(pcm.data_ptr(), offs, ns, want.data_ptr(), moffs, 0, finalize=True)
    B.check(_noalloc(lib, fe, pcm, offs, ns, got, moffs, None, True))
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    fe.close()
