#!/usr/bin/env python
"""bench.py — audio-seconds encoded per second through the log-mel frontend + audio encoder.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (q3a_encode_pcm_*: mel -> conv stem -> transformer layers ->
projection) over one batch of 64 synthetic 30-s clips with random-init Qwen3-ASR-0.6B encoder weights
(BASELINE.json configs[1]).  Every rank owns its own batch (utterance sharding, weak scaling, no
collective on the compute path).  `value` is measured with the PCM resident in HBM (CUDA events on the
launch stream); `e2e` goes through the host-pointer C-ABI call with pinned host buffers, H2D and D2H
copies inside the timed region.  --impl reference times the CPU path (reference mel TU + the numpy
restatement of the ggml encoder) on the box's host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "audio-sec encoded/sec (mel+encoder)"
UNIT = "audio-s/s"
CLIP_SECONDS = 30
BATCH = 64
SR = 16000
ASR_GFLOP_PER_30S = 278.02


def headline_config():
    """The workload both arms are quoted on (BASELINE.json configs[1]); the CPU arm times a bounded sample of it."""
    T = 390 * BATCH
    return {"workload": "qwen3-asr-0.6b mel+encoder, batch of 64 x 30-s synthetic 16 kHz clips per GPU "
                        "(BASELINE.json configs[1]), random-init weights via GGUF F16",
            "batch_clips": BATCH, "clip_seconds": CLIP_SECONDS, "tokens_per_step": T, "mode": "asr (full attention)",
            "l2": "two PCM input sets alternated (246 MB > L2); a step streams ~8 GB of workspace",
            "algorithmic_gflop_per_step": ASR_GFLOP_PER_30S * BATCH}


CPU_SAMPLE_CLIPS = 2          # 30-s clips of the headline workload per CPU step (both the cpu_baseline leg and --impl reference)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "tflops_burst": float(d["bf16_tflops"]), "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tflops_sustained": 1400.0, "tflops_burst": 1590.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.stop_ev, self.th = index, [], threading.Event(), None
        self.source = "nvidia-smi, 0.1-s period"

    def _run_nvml(self) -> bool:
        """NVML in-process (a sample every 5 ms: a 10-step timed region is only ~0.2 s long); False when unavailable"""
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
            get_reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
            N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM); get_reasons(h)
            self.source = "nvml, 5-ms period"
        except Exception:
            return False
        bits = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))
        while not self.stop_ev.is_set():
            try:
                sm, r = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM), int(get_reasons(h))
                self.rows.append([str(sm), str(mx)] + ["Active" if r & b else "Not Active" for b, _ in bits])
            except Exception:
                pass
            self.stop_ev.wait(0.005)
        return True

    def _run(self):
        if self._run_nvml():
            return
        while not self.stop_ev.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop_ev.wait(0.1)

    def start(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def stop(self) -> dict:
        self.stop_ev.set()
        if self.th:
            self.th.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm), "sampler": self.source}


def ensure_gguf(tmpdir: str, ftype: str = "f16") -> str:
    """Random-init ASR-0.6B encoder written the way the reference converter would (no checkpoints offline)."""
    from q3a_b200 import synth
    path = os.path.join(tmpdir, f"asr06b_seed0_{ftype}.gguf")
    if not os.path.exists(path):
        synth.write_gguf(path, synth.ASR_06B, synth.random_weights(synth.ASR_06B, 0), ftype)
    return path


# ------------------------------------------------------------------------------------------------ CPU arm
_CPU_STATE = {}


def _cpu_group(clip_s: float, seeds):
    """len(seeds) clips through the CPU implementation of the path, using every host core the code can use: the mel
    through the reference's own translation unit when oracle/_ref is present -- one clip per thread, because the TU
    itself is single-threaded (it ignores n_threads, mel_spectrogram.cpp:484-628) -- then the encoder through the numpy
    restatement of the ggml graph (BLAS threads = all host cores), clip after clip.  Returns (seconds total, seconds
    in mel)."""
    from concurrent.futures import ThreadPoolExecutor
    from q3a_b200 import synth
    from oracle import mel_oracle as MO, encoder_oracle as EO
    st = _CPU_STATE
    if not st:
        shape = synth.ASR_06B
        st["shape"] = shape
        st["W"] = EO.Weights(synth.stored_weights(synth.random_weights(shape, 0), "f16"), "f16")
        try:
            st["ref"] = MO.RefMel(); st["filt"] = st["ref"].filters()
            st["mel_kind"] = "reference TU mel_spectrogram.cpp (single-threaded code, one clip per host thread)"
        except Exception:
            st["ref"] = None; st["filt"] = MO.mel_filters_numpy()
            st["mel_kind"] = "numpy restatement (oracle/_ref missing)"
    clips = [synth.synth_audio(sd, int(clip_s * SR)) for sd in seeds]
    one = (lambda pcm: st["ref"].log_mel(pcm, st["filt"])) if st["ref"] else (lambda pcm: MO.log_mel_numpy(pcm, st["filt"]))
    t0 = time.perf_counter()
    if len(clips) > 1:
        with ThreadPoolExecutor(max_workers=len(clips)) as ex:       # ctypes calls release the GIL
            mels = list(ex.map(one, clips))
    else:
        mels = [one(clips[0])]
    t1 = time.perf_counter()
    sh = st["shape"]
    for mel in mels:
        EO.encode(mel, st["W"], n_layers=sh.n_layers, d_model=sh.d_model, n_heads=sh.n_heads, variant="asr")
    t2 = time.perf_counter()
    return t2 - t0, t1 - t0


def _cpu_desc(n_clips, clip_s, mel, enc):
    cores = os.cpu_count() or 1
    return {"kind": "port", "cores": cores,
            "sample": f"{n_clips} x {clip_s:g}-s clips; mel: {_CPU_STATE['mel_kind']} ({mel:.1f} s wall); encoder: numpy "
                      f"restatement of the ggml graph, f16 weights, BLAS on all {cores} cores ({enc:.1f} s)"}


def cpu_path_rate(clip_seconds: float, n_clips: int, seed0: int = 100):
    """Bounded CPU sample -> (audio-s/s, description for the cpu_baseline object)."""
    tot, mel = _cpu_group(clip_seconds, [seed0 + i for i in range(n_clips)])
    return n_clips * clip_seconds / tot, _cpu_desc(n_clips, clip_seconds, mel, tot - mel)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    total = args.steps + args.warmup
    # One stated sample for both CPU figures: CPU_SAMPLE_CLIPS 30-s clips of the headline workload per step (a clip
    # costs ~3 s of mel on one thread + ~3 s of encoder on all cores); 1 clip per step for long runs so that the whole
    # --steps K --warmup W run ends within a few minutes.  The per-clip rate does not depend on the batch size: the
    # reference encodes one utterance at a time (qwen3_asr.cpp:81-149).
    clip_s = float(CLIP_SECONDS)
    per_step = CPU_SAMPLE_CLIPS if total <= 20 else 1
    t_all = time.perf_counter()
    secs = mel = 0.0
    for i in range(total):
        a, b = _cpu_group(clip_s, [100 + per_step * i + k for k in range(per_step)])
        if i >= args.warmup:
            secs += a; mel += b
    value = args.steps * per_step * clip_s / secs
    desc = _cpu_desc(args.steps * per_step, clip_s, mel, secs - mel)
    desc["sample"] = f"{per_step} of the workload's 64 x 30-s clips per step; " + desc["sample"]
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f16 weights, f32/f64 arithmetic", "data": "synthetic",
            "config": headline_config(),
            "cpu_baseline": {"value": value, "unit": UNIT, **desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t_all}
    print(json.dumps(line), file=_OUT, flush=True)


# ------------------------------------------------------------------------------------------------ sharded corpus
CORPUS_SECONDS = 36000.0          # BASELINE.json configs[4]: 10 h, utterances log-uniform 2-60 s, seed 500
CORPUS_SEED = 500


def corpus_lengths(total_s: float = CORPUS_SECONDS, seed: int = CORPUS_SEED):
    """Sample counts of the synthetic long-form corpus: log-uniform 2-60 s until `total_s` is reached; not multiples of 160."""
    import numpy as np
    rng = np.random.default_rng(seed)
    out, tot = [], 0.0
    while tot < total_s:
        sec = float(np.exp(rng.uniform(np.log(2.0), np.log(60.0))))
        out.append(int(sec * SR) + int(rng.integers(0, 160)))
        tot += sec
    return out


def corpus_utterance(base, i: int, n: int):
    """Utterance i of the corpus, identical on whichever rank builds it: a shifted, scaled cut of a 60-s synthetic signal."""
    import numpy as np
    rng = np.random.default_rng([CORPUS_SEED, i])
    return (np.roll(base, int(rng.integers(0, len(base))))[:n] * np.float32(rng.uniform(0.5, 1.0))).astype(np.float32)


def corpus_plan(B, rank: int, world: int):
    """(sample counts, utterances this rank owns, utterances of the next rank it re-encodes as a cross-check)"""
    lens = corpus_lengths()
    plan = B.plan_shards([n // 160 for n in lens], B.MODE_ASR, world)
    mine = [i for i, s in enumerate(plan) if s == rank]
    theirs = [i for i, s in enumerate(plan) if s == (rank + 1) % world][:8] if world > 1 else mine[:8]
    return lens, mine, theirs


def corpus_verdict(gathered):
    """gathered: per rank ({utterance: digest} of its own shard, {utterance: digest} of its cross-checks) ->
    (every cross-check equals the owner's digest, number of cross-checks, utterances owned exactly once)"""
    owner, owned = {}, 0
    for o, _ in gathered:
        owned += len(o)
        owner.update(o)
    checked = sum(len(c) for _, c in gathered)
    identical = all(owner.get(i) == h for _, c in gathered for i, h in c.items())
    return identical, checked, owned == len(owner)


def run_corpus(enc, fe, B, synth, rank: int, world: int, dist, torch):
    """configs[4] on hardware: every rank derives the same LPT plan (q3a_plan_shards), encodes only its own shard through
    the host-pointer call, and the job's time is the slowest rank's.  Determinism: every rank also encodes 8 utterances
    owned by the next rank, in a different batch; their SHA-256 must equal the owner's."""
    import ctypes as C
    import hashlib
    import numpy as np
    lens, mine, theirs = corpus_plan(B, rank, world)
    base = synth.synth_audio(CORPUS_SEED, 60 * SR + 159)
    width = enc.hp.out_dim

    def stage(ids):
        ns = [lens[i] for i in ids]
        toks = [enc.n_tokens(n // 160) for n in ns]
        pcm = torch.empty(sum(ns), dtype=torch.float32).pin_memory()
        out = torch.empty((sum(toks), width), dtype=torch.float32).pin_memory()
        pv, po, pp, op = pcm.numpy(), 0, [], []
        ro = 0
        for i, n, t in zip(ids, ns, toks):
            pv[po:po + n] = corpus_utterance(base, i, n)
            pp.append(pcm.data_ptr() + 4 * po); op.append(out.data_ptr() + 4 * width * ro)
            po += n; ro += t
        return pcm, out, ns, toks, (C.c_void_p * len(ids))(*pp), (C.c_void_p * len(ids))(*op), (C.c_int * len(ids))(*ns)

    def encode(st):
        pcm, out, ns, toks, pp, op, na = st
        ntok = (C.c_int * len(ns))()
        B.check(enc.lib.q3a_encode_pcm_batch(fe.h, enc.h, pp, na, len(ns), B.MODE_ASR, op, ntok))
        assert list(ntok) == toks

    def sub_batches(st, target_s: float = 2000.0):
        """the shard cut into consecutive sub-batches of about `target_s` audio-seconds (one 64 x 30-s step is 1920 s), each
        with its own argument arrays over the SAME staged buffers"""
        pcm, out, ns, toks, pp, op, na = st
        cuts, acc = [0], 0.0
        for i, n in enumerate(ns):
            acc += n / SR
            if acc >= target_s and i + 1 < len(ns):
                cuts.append(i + 1); acc = 0.0
        cuts.append(len(ns))
        subs = []
        for a, b in zip(cuts[:-1], cuts[1:]):
            k = b - a
            subs.append(((C.c_void_p * k)(*pp[a:b]), (C.c_void_p * k)(*op[a:b]), (C.c_int * k)(*ns[a:b]), (C.c_int * k)(), k, toks[a:b]))
        return subs

    def encode_pipelined(subs):
        """q3a_encode_pcm_submit / q3a_encode_wait over the sub-batches, two in flight: sub-batch k+1 uploads under k's
        kernels and k-1 downloads under them -- how a caller with a queue of work uses the library"""
        prev = None
        for pp, op, na, ntok, k, toks in subs:
            tk = C.c_int64(-1)
            B.check(enc.lib.q3a_encode_pcm_submit(fe.h, enc.h, pp, na, k, B.MODE_ASR, op, ntok, C.byref(tk)))
            if prev is not None:
                B.check(enc.lib.q3a_encode_wait(enc.h, prev))
            prev = tk.value
        if prev is not None:
            B.check(enc.lib.q3a_encode_wait(enc.h, prev))
        for pp, op, na, ntok, k, toks in subs:
            assert list(ntok) == toks

    def digests(st, ids):
        out, toks = st[1].numpy(), st[3]
        d, r = {}, 0
        for i, t in zip(ids, toks):
            d[i] = hashlib.sha256(out[r:r + t].tobytes()).hexdigest()
            r += t
        return d

    st_mine, st_theirs = stage(mine), stage(theirs)
    encode(st_mine)                                   # warm-up: workspace growth, first-touch of the pinned pages
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    encode(st_mine)
    dt_one_call = time.perf_counter() - t0
    one_call_digest = hashlib.sha256(st_mine[1].numpy().tobytes()).hexdigest()
    subs = sub_batches(st_mine)
    st_mine[1].zero_()
    encode_pipelined(subs)                            # warm-up of this call pattern
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    encode_pipelined(subs)
    dt = time.perf_counter() - t0
    same_bytes = hashlib.sha256(st_mine[1].numpy().tobytes()).hexdigest() == one_call_digest
    encode(st_theirs)
    own, chk = digests(st_mine, mine), digests(st_theirs, theirs)
    t1 = torch.tensor([dt_one_call], device="cuda")
    if world > 1:
        dist.all_reduce(t1, op=dist.ReduceOp.MAX)
    flag = torch.tensor([1 if same_bytes else 0], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    t = torch.tensor([dt], device="cuda")
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        gathered = [None] * world
        dist.all_gather_object(gathered, (own, chk))
    else:
        gathered = [(own, chk)]
    identical, checked, partition_ok = corpus_verdict(gathered)
    identical = identical and partition_ok and sum(len(o) for o, _ in gathered) == len(lens) and bool(flag.item())
    makespan = float(tmax.item())
    audio_s = sum(lens) / SR
    return {"workload": "BASELINE.json configs[4]: 10-h synthetic corpus, utterances log-uniform 2-60 s (seed 500), ASR-0.6B bf16, "
                        "sharded by utterance with q3a_plan_shards (LPT); every rank walks its shard in sub-batches of ~2000 "
                        "audio-seconds through q3a_encode_pcm_submit / q3a_encode_wait (two in flight), pinned host PCM in / "
                        "pinned host features out",
            "utterances": len(lens), "audio_seconds": audio_s, "n_gpus": world, "makespan_s": makespan,
            "audio_s_per_s": audio_s / makespan, "rank_time_max_over_mean": makespan / (float(tsum.item()) / world),
            "sub_batches_this_rank": len(subs),
            "one_synchronous_call_per_rank_audio_s_per_s": audio_s / float(t1.item()),
            "sub_batches_equal_one_call_bitwise": bool(flag.item()),
            "utterances_this_rank": len(mine), "bit_identical": bool(identical), "cross_checked_utterances": checked,
            "check": "SHA-256 of 8 utterances per rank re-encoded by the next rank in a different batch vs the owner's; "
                     "SHA-256 of the whole shard encoded in sub-batches vs in one call"}


# ------------------------------------------------------------------------------------------------ GPU arm
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import q3a_b200  # noqa: F401
    from q3a_b200 import binding as B, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU arm)")
    from q3a_b200 import hostbind
    # before any pinned allocation: buffers land on the GPU's NUMA node (Q3A_BENCH_NO_BIND=1 for an A/B run)
    # Functional check of the multi-rank path on a box with fewer GPUs than ranks (development only, flagged in the line):
    # ranks share GPUs and the bookkeeping collectives go over gloo; there is no collective on the compute path either way.
    n_dev = torch.cuda.device_count()
    emulated = world > n_dev
    local = local % n_dev
    binding_info = {"bound": False, "why_not": "Q3A_BENCH_NO_BIND"} if os.environ.get("Q3A_BENCH_NO_BIND") else hostbind.bind_to_gpu(local)
    torch.cuda.set_device(local)
    if world > 1:
        if emulated:
            dist.init_process_group("gloo")
        else:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B.load()

    tmpdir = os.path.join(tempfile.gettempdir(), "q3a_bench")
    os.makedirs(tmpdir, exist_ok=True)
    if rank == 0:
        ensure_gguf(tmpdir)
    if world > 1:
        dist.barrier()
    gguf_path = ensure_gguf(tmpdir)
    fe = B.Frontend(local)
    enc = B.Encoder(gguf_path, local, B.OP_BF16)

    n = CLIP_SECONDS * SR
    # two input sets per rank, alternated between steps (2 x 123 MB PCM > the 126 MB L2; the ~8 GB of
    # workspace a step streams through evicts everything anyway)
    sets = []
    for s in range(2):
        base = synth.synth_audio(1000 + 17 * rank + s, n)
        host = torch.empty((BATCH, n), dtype=torch.float32).pin_memory()
        rng = np.random.default_rng(5000 + 131 * rank + s)
        hv = host.numpy()
        for i in range(BATCH):                       # distinct clips: circular shift + gain of a synthetic utterance
            hv[i] = np.roll(base, int(rng.integers(0, n))) * np.float32(rng.uniform(0.5, 1.0))
        sets.append(host)
    dev_sets = [h.cuda() for h in sets]
    pcm_off = [i * n for i in range(BATCH)]
    n_samples = [n] * BATCH
    T = enc.n_tokens(n // 160) * BATCH
    out_dev = torch.empty((T, enc.hp.out_dim), dtype=torch.float32, device="cuda")
    # two sets of pinned feature buffers: the asynchronous call keeps two steps in flight
    # (each set is one pinned block, a clip's rows at its offset: the library then moves a download part with one copy)
    out_blocks = [torch.empty((T, enc.hp.out_dim), dtype=torch.float32).pin_memory() for _ in range(2)]
    rows_per_clip = T // BATCH
    out_hosts = [[blk[i * rows_per_clip:(i + 1) * rows_per_clip] for i in range(BATCH)] for blk in out_blocks]
    out_host = out_hosts[0]
    stream = torch.cuda.current_stream().cuda_stream

    def step_dev(i):
        enc.encode_pcm_device(fe, dev_sets[i & 1].data_ptr(), pcm_off, n_samples, out_dev.data_ptr(), B.MODE_ASR, stream)

    import ctypes as C
    outs_ptrs = [(C.c_void_p * BATCH)(*[o.data_ptr() for o in oh]) for oh in out_hosts]
    outs_ptr = outs_ptrs[0]
    ns_arr = (C.c_int * BATCH)(*n_samples)
    ntok = (C.c_int * BATCH)()
    pcm_ptrs = [(C.c_void_p * BATCH)(*[h[i].data_ptr() for i in range(BATCH)]) for h in sets]

    def step_e2e(i):
        B.check(enc.lib.q3a_encode_pcm_batch(fe.h, enc.h, pcm_ptrs[i & 1], ns_arr, BATCH, B.MODE_ASR, outs_ptr, ntok))

    def submit_e2e(i):
        tk = C.c_int64(-1)
        B.check(enc.lib.q3a_encode_pcm_submit(fe.h, enc.h, pcm_ptrs[i & 1], ns_arr, BATCH, B.MODE_ASR, outs_ptrs[i & 1], ntok, C.byref(tk)))
        return tk.value

    def run_e2e_pipelined(k):
        """k steps through the asynchronous public call, two in flight: step i+1 is submitted before step i is waited for"""
        prev = None
        for i in range(k):
            tk = submit_e2e(i)
            if prev is not None:
                B.check(enc.lib.q3a_encode_wait(enc.h, prev))
            prev = tk
        if prev is not None:
            B.check(enc.lib.q3a_encode_wait(enc.h, prev))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing ----------------------------------------------------------------------
    for i in range(args.warmup):
        step_dev(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step_dev(i)
        launches0 += enc.timings().total_launches
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    audio_s = BATCH * CLIP_SECONDS * args.steps * world
    value = audio_s / (ms_total / 1e3)

    # ---- end-to-end through the host-pointer C ABI -------------------------------------------------------
    for i in range(max(1, min(args.warmup, 3))):
        step_e2e(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step_e2e(i)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_sync_value = audio_s / float(dt.item())
    # the same K steps through q3a_encode_pcm_submit / q3a_encode_wait (two calls in flight): every step still uploads its
    # PCM from pinned host memory and downloads its features inside the timed region
    run_e2e_pipelined(2)
    barrier()
    host_cpu = None
    try:                                              # whole-box CPU utilisation over the timed e2e region (all ranks' processes)
        import psutil
        psutil.cpu_percent(interval=None)
    except Exception:
        psutil = None
    t0 = time.perf_counter()
    run_e2e_pipelined(args.steps)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if psutil is not None:
        try:
            host_cpu = {"pct_of_all_cores": psutil.cpu_percent(interval=None), "cores": psutil.cpu_count(),
                        "what": "box-wide CPU utilisation between the start and the end of this rank's timed e2e region"}
        except Exception:
            host_cpu = None
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = audio_s / float(dt.item())
    # A/B of the library's policy for calls in flight: with the intra-call upload / download parts kept (Q3A_ASYNC_PARTS=1)
    os.environ["Q3A_ASYNC_PARTS"] = "1"
    run_e2e_pipelined(2)
    barrier()
    t0 = time.perf_counter()
    run_e2e_pipelined(args.steps)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    del os.environ["Q3A_ASYNC_PARTS"]
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_parts_value = audio_s / float(dt.item())

    # ---- host link probe: every rank moves one step's worth of PCM up and features down at the same moment, with no
    #      kernels running -- the copy time an e2e step has to hide, and the evidence for what bounds e2e at N > 1
    #      (all ranks share the host's memory controllers and PCIe root ports)
    link = None
    if not args.no_link_probe:
        cs_in, cs_out = torch.cuda.Stream(), torch.cuda.Stream()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        res = []
        for both in (False, True):                    # each direction alone, then both at once (what the pipeline does)
            barrier()
            reps = 4
            with torch.cuda.stream(cs_in):
                ev[0].record()
                for r in range(reps):
                    dev_sets[r & 1].copy_(sets[r & 1], non_blocking=True)
                ev[1].record()
            if not both:
                torch.cuda.synchronize()
                barrier()
            with torch.cuda.stream(cs_out):
                ev[2].record()
                for r in range(reps):
                    out_blocks[r & 1].copy_(out_dev, non_blocking=True)
                ev[3].record()
            torch.cuda.synchronize()
            res += [reps * sets[0].numel() * 4 / ev[0].elapsed_time(ev[1]) / 1e6, reps * out_dev.numel() * 4 / ev[2].elapsed_time(ev[3]) / 1e6]
        lt = torch.tensor(res, device="cuda")
        lo, hi = lt.clone(), lt.clone()
        if world > 1:
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        keys = ("h2d_alone", "d2h_alone", "h2d_with_d2h", "d2h_with_h2d")
        link = {"unit": "GB/s per rank, all ranks copying at once, pinned host memory, no kernels running",
                "min_over_ranks": {k: round(float(v), 2) for k, v in zip(keys, lo.tolist())},
                "max_over_ranks": {k: round(float(v), 2) for k, v in zip(keys, hi.tolist())}}
        h2d_ms = sets[0].numel() * 4 / (link["min_over_ranks"]["h2d_with_d2h"] * 1e6)
        d2h_ms = out_dev.numel() * 4 / (link["min_over_ranks"]["d2h_with_h2d"] * 1e6)
        link["copy_ms_per_step"] = {"h2d": round(h2d_ms, 3), "d2h": round(d2h_ms, 3)}

    # ---- the timed device-resident call is checked, not just timed: its output must equal, bit for bit, what the
    #      (tested) host-pointer call returns for the same input set (batch-invariant kernels, same code underneath)
    step_e2e(0)
    out_dev.fill_(float("nan"))
    step_dev(0)
    torch.cuda.synchronize()
    got = out_dev.cpu()
    want = out_blocks[0]
    finite = bool(torch.isfinite(got).all())
    max_diff = float((got - want).abs().max()) if finite else float("nan")
    ok = torch.tensor([1 if (finite and torch.equal(got, want)) else 0], device="cuda")
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    value_checked = bool(ok.item())
    if not value_checked:
        raise SystemExit(f"bench.py: q3a_encode_pcm_device output differs from q3a_encode_pcm_batch on rank {rank} "
                         f"(finite={finite}, max |diff| = {max_diff}) -- the timed value is void")

    corpus = None
    if not args.no_corpus:
        corpus = run_corpus(enc, fe, B, synth, rank, world, dist, torch)
        if not corpus["bit_identical"]:
            raise SystemExit("bench.py: sharded-corpus determinism check failed: an utterance's features depend on the rank / batch")

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- rank 0: per-kernel roofline from a profiled pass of the same steps (events inside the library,
    #      on the launch stream), parity smoke, CPU baseline ---------------------------------------------
    pk = peaks()
    enc.set_profiling(True)
    g_ms = g_fl = a_ms = a_fl = mel_ms = stem_ms = x_ms = 0.0
    g_n = 0
    for i in range(args.steps):
        step_dev(i)
        tm = enc.timings()
        g_ms += tm.gemm_ms; g_fl += tm.gemm_flops; a_ms += tm.attention_ms; a_fl += tm.attention_flops
        mel_ms += tm.mel_ms; stem_ms += tm.conv_chunk_ms; x_ms += tm.transformer_ms; g_n += tm.gemm_launches
    enc.set_profiling(False)
    gemm_tf = g_fl / g_ms / 1e9
    mel_bytes = BATCH * (4 * n + 4 * 128 * (n // 160)) * args.steps
    mel_gbs = mel_bytes / (mel_ms / 1e3) / 1e9
    # DRAM bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum), from the committed launch list of this
    # same workload (profiles/roofline_traffic.json, written by scripts/summarize_profile.py); null without it
    traffic = traffic_attn = traffic_mel = traffic_note = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        try:
            with open(tp) as f:
                tj = json.load(f)
            import importlib.util
            spec = importlib.util.spec_from_file_location("q3a_build", os.path.join(ROOT, "qwen3-asr.cpp_b200", "build.py"))
            bmod = importlib.util.module_from_spec(spec); spec.loader.exec_module(bmod)
            if tj.get("csrc_sha256") == bmod.source_hash():      # the ncu capture was taken from these very sources
                traffic = tj.get("gemm_tc_kernel_dram_bytes_per_launch")
                traffic_attn = tj.get("attention_kernel_dram_bytes_per_launch")
                traffic_mel = tj.get("mel_frames_kernel_dram_bytes_per_launch")
            else:
                traffic_note = "profiles/roofline_traffic.json is from other sources (csrc hash mismatch): traffic not reported"
        except Exception:
            pass

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, desc = cpu_path_rate(float(CLIP_SECONDS), CPU_SAMPLE_CLIPS)
        desc["sample"] = f"{CPU_SAMPLE_CLIPS} of the workload's 64 x 30-s clips; " + desc["sample"]
        cpu = {"value": v, "unit": UNIT, **desc}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 operands, fp32 accumulate (mel: fp32)", "data": "synthetic",
        "config": headline_config(),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": BATCH * n * 4,
                "d2h_bytes_per_step": T * enc.hp.out_dim * 4, "timer": "host wall clock around K steps, max over ranks",
                "api": "q3a_encode_pcm_submit + q3a_encode_wait (pinned host PCM -> pinned host features), two calls in flight",
                "with_intra_call_parts_value": e2e_parts_value,
                "synchronous_call_value": e2e_sync_value,
                "synchronous_api": "q3a_encode_pcm_batch, one call at a time",
                "host_link": link, "host_cpu": host_cpu},
        "gpu_launches": launches0,
        "ranks_share_gpus": emulated,
        "value_checked": value_checked, "value_check": {"against": "q3a_encode_pcm_batch on the same 64 clips (host result)",
                                                         "max_abs_diff": max_diff, "rows": int(got.shape[0])},
        "clocks": clocks,
        "host_binding": binding_info,
        "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 GEMM: conv2/conv3 implicit GEMM + all Linear layers)",
                     "achieved": gemm_tf, "peak": pk["tflops_sustained"], "unit": "TFLOP/s", "frac": gemm_tf / pk["tflops_sustained"],
                     "traffic": traffic, "traffic_note": traffic_note, "peak_source": pk["source"] + ", sustained bf16 (kernel timed inside a long step)",
                     "launches_per_step": g_n // max(args.steps, 1), "ms_per_step": g_ms / args.steps,
                     "how": "algorithmic FLOPs (2*M*N*K, no padding) / CUDA-event time of every launch, separate profiled pass of the same K steps"},
        "roofline_attention": {"bound": "tensor", "kernel": "attention_kernel (tcgen05 varlen attention)",
                               "achieved": a_fl / a_ms / 1e9, "peak": pk["tflops_sustained"], "unit": "TFLOP/s",
                               "frac": a_fl / a_ms / 1e9 / pk["tflops_sustained"], "ms_per_step": a_ms / args.steps,
                               "traffic": traffic_attn},
        "roofline_mel": {"bound": "hbm", "kernel": "mel_frames_kernel (+ max decode)", "achieved": mel_gbs, "peak": pk["hbm_gbs"],
                         "unit": "GB/s", "frac": mel_gbs / pk["hbm_gbs"], "ms_per_step": mel_ms / args.steps, "traffic": traffic_mel,
                         "how": "(4*n_samples + 4*128*n_len) per clip / CUDA-event time"},
        "stage_ms_per_step": {"mel_spectrogram": mel_ms / args.steps, "audio_encoding.conv_chunk": stem_ms / args.steps,
                              "audio_encoding.transformer": x_ms / args.steps},
        "whole_step_tflops": ASR_GFLOP_PER_30S * BATCH * args.steps * world / (ms_total / 1e3) / 1e3,
        "device_bytes": enc.device_bytes(),
    }
    if corpus is not None:
        line["corpus"] = corpus
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), file=_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_OUT = sys.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the ~20 s CPU baseline leg (profiling runs)")
    ap.add_argument("--no-corpus", action="store_true", help="skip the sharded 10-h corpus leg (profiling runs)")
    ap.add_argument("--no-link-probe", action="store_true", help="skip the host<->device copy-rate probe of the e2e leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner) go to stderr
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
