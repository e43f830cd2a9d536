"""Utterance sharding across ranks (world_size 2, gloo, CPU): every rank derives the same plan from the
same corpus description, the shards partition the corpus, and results gathered per shard come back in
utterance order.  The per-rank 'encoder' here is a stand-in that returns the token count — the data path
has no collective, only this host-side bookkeeping does."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from q3a_b200 import binding as B, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    plan = B.plan_shards(frames, B.MODE_ASR, world)
    mine = [i for i, s in enumerate(plan) if s == rank]
    tokens = torch.zeros(len(frames), dtype=torch.int64)
    for i in mine:
        tokens[i] = B.load().q3a_encoder_n_tokens(frames[i], B.MODE_ASR)
    dist.all_reduce(tokens)                     # bookkeeping only: each slot is written by exactly one rank
    plans = [None] * world
    dist.all_gather_object(plans, plan)
    if rank == 0:
        q.put((plans, tokens.tolist(), mine))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding(lib):
    frames = [3000, 120, 777, 6000, 45, 2999, 100, 1500, 0, 9200]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    plans, tokens, mine0 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert plans[0] == plans[1]
    assert sorted(set(plans[0])) == [0, 1]
    assert tokens == [synth.chunk_tokens(f) for f in frames]
    assert mine0 == [i for i, s in enumerate(plans[0]) if s == 0]


def _corpus_worker(rank, world, port, corrupt, q):
    """bench.py's sharded-corpus bookkeeping (configs[4]) with a stand-in encoder: digest of the utterance's samples"""
    import hashlib
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lens, mine, theirs = bench.corpus_plan(B, rank, world)
    base = synth.synth_audio(bench.CORPUS_SEED, 2 * 16000)
    dig = lambda i: hashlib.sha256(bench.corpus_utterance(base, i, min(lens[i], 1000)).tobytes()).hexdigest()
    own = {i: dig(i) for i in mine[:40]}
    own.update({i: dig(i) for i in [j for j, s in enumerate(B.plan_shards([n // 160 for n in lens], B.MODE_ASR, world)) if s == rank][:8]})
    chk = {i: dig(i) for i in theirs}
    if corrupt and rank == 1:
        chk[theirs[0]] = "0" * 64
    gathered = [None] * world
    dist.all_gather_object(gathered, (own, chk))
    if rank == 0:
        q.put((bench.corpus_verdict(gathered), len(lens), len(mine), theirs))
    dist.barrier()
    dist.destroy_process_group()


def test_corpus_bookkeeping_two_ranks(lib):
    ctx = mp.get_context("spawn")
    for corrupt in (False, True):
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_corpus_worker, args=(r, 2, port, corrupt, q)) for r in range(2)]
        for p in procs:
            p.start()
        (identical, checked, _), n_utts, n_mine, theirs = q.get(timeout=180)
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        assert n_utts > 1500 and abs(n_mine - n_utts / 2) < 0.2 * n_utts and len(theirs) == 8
        assert checked == 16 and identical == (not corrupt)
