"""Model check of attention2.cu's barrier protocol (no GPU): the four roles of a CTA -- TMA loader, MMA issuer, softmax
warpgroups A and B -- are replayed as coroutines over mbarriers with the hardware's phase-parity semantics, under a random
scheduler and random completion delays of the asynchronous operations (TMA loads, tcgen05.commit arrivals).  For item
sequences that mix one- and two-tile items and 1..N key blocks it checks that nothing deadlocks and that no buffer is
overwritten while a consumer may still read it: Q slots, K/V ring stages, S/P columns and O columns.

The role bodies below mirror the control flow of the kernel line by line (same waits, same parities, same commits); when
the kernel's protocol changes, this file changes with it."""
import random

import pytest

STAGES = 3


class Bar:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0
        if self.pending == 0:
            self.pending, self.phase = self.count, self.phase + 1

    def passed(self, parity):                 # mbarrier.try_wait.parity
        return (self.phase & 1) != parity


class Sim:
    def __init__(self, items, seed):
        self.items = items                    # list of (ntile, nkv)
        self.rng = random.Random(seed)
        b = lambda c: Bar(c)
        self.qfull, self.qempty = [b(1), b(1)], [b(1), b(1)]
        self.kfull, self.vfull, self.kvempty = [b(1) for _ in range(STAGES)], [b(1) for _ in range(STAGES)], [b(1) for _ in range(STAGES)]
        self.sfull, self.pfull, self.ofull, self.ofree = [b(1), b(1)], [b(4), b(4)], [b(1), b(1)], [b(4), b(4)]
        self.pending = []                     # asynchronous arrivals: (due step, bar, action)
        self.step = 0
        # resource state for the hazard checks
        self.q_busy = [False, False]          # Q slot holds data the tensor pipe may still read
        self.kv_busy = [False] * STAGES
        self.s_state = ["free", "free"]       # free -> S (scores written) -> P (probabilities written) -> free (P.V read it)
        self.o_busy = [False, False]          # O holds a finished item the softmax warps have not read yet
        self.mma_queue = []                   # in-order tensor pipe: ops complete in issue order

    # ---- asynchronous machinery ----
    def later(self, fn):
        self.pending.append((self.step + self.rng.randint(1, 12), fn))

    def mma(self, fn):                        # an MMA or a commit: executes in order, some time after issue
        self.mma_queue.append(fn)

    def pump(self):
        due = [p for p in self.pending if p[0] <= self.step]
        self.pending = [p for p in self.pending if p[0] > self.step]
        for _, fn in due:
            fn()
        if self.mma_queue and self.rng.random() < 0.5:
            self.mma_queue.pop(0)()

    # ---- roles (generators yield ("wait", bar, parity) or None) ----
    def loader(self):
        kv_it, q_ph = 0, 0
        for ntile, nkv in self.items:
            for x in range(ntile):
                yield ("wait", self.qempty[x], ((q_ph >> x) & 1) ^ 1)
                q_ph ^= 1 << x
                assert not self.q_busy[x], "Q slot overwritten while in use"
                self.q_busy[x] = True
                self.later(self.qfull[x].arrive)
            for _ in range(nkv):
                s, ph = kv_it % STAGES, (kv_it // STAGES) & 1
                yield ("wait", self.kvempty[s], ph ^ 1)
                assert not self.kv_busy[s], "K/V stage overwritten while in use"
                self.kv_busy[s] = True
                self.later(self.kfull[s].arrive)
                self.later(self.vfull[s].arrive)
                kv_it += 1

    def issuer(self):
        kv_c, item_ph, blk_ph = 0, 0, 0
        for ntile, nkv in self.items:
            def issue_qk(x, j):
                def run():
                    assert self.s_state[x] == "free", f"S_{x} overwritten in state {self.s_state[x]}"
                    self.s_state[x] = "S"
                self.mma(run)
                self.mma(lambda: self.later(self.sfull[x].arrive))
                if j == nkv - 1:
                    def rel():
                        self.q_busy[x] = False
                        self.later(self.qempty[x].arrive)
                    self.mma(rel)
            s0 = kv_c % STAGES
            yield ("wait", self.kfull[s0], (kv_c // STAGES) & 1)
            for x in range(ntile):
                yield ("wait", self.qfull[x], (item_ph >> x) & 1)
                issue_qk(x, 0)
            for j in range(nkv):
                c = kv_c + j
                s, ph, sn, phn = c % STAGES, (c // STAGES) & 1, (c + 1) % STAGES, ((c + 1) // STAGES) & 1
                for x in range(ntile):
                    yield ("wait", self.pfull[x], (blk_ph >> x) & 1)
                    blk_ph ^= 1 << x
                    if j == 0:
                        yield ("wait", self.ofree[x], ((item_ph >> x) & 1) ^ 1)
                    yield ("wait", self.vfull[s], ph)

                    def pv(x=x, j=j, s=s):
                        assert self.s_state[x] == "P", f"P.V on S_{x} in state {self.s_state[x]}"
                        assert self.kv_busy[s]
                        if j == 0:
                            assert not self.o_busy[x], "O overwritten before the previous item was read"
                        self.s_state[x] = "free"
                    self.mma(pv)
                    if j == nkv - 1:
                        def done(x=x):
                            self.o_busy[x] = True
                            self.later(self.ofull[x].arrive)
                        self.mma(done)
                    if x == ntile - 1:
                        def rel(s=s):
                            self.kv_busy[s] = False
                            self.later(self.kvempty[s].arrive)
                        self.mma(rel)
                    if j + 1 < nkv:
                        yield ("wait", self.kfull[sn], phn)
                        issue_qk(x, j + 1)
            kv_c += nkv
            item_ph ^= 3 if ntile == 2 else 1
        if kv_c:
            yield ("wait", self.kvempty[(kv_c - 1) % STAGES], ((kv_c - 1) // STAGES) & 1)

    def softmax_warp(self, x):
        blk_n, item_n = 0, 0
        for ntile, nkv in self.items:
            if x >= ntile:
                continue
            for _ in range(nkv):
                yield ("wait", self.sfull[x], blk_n & 1)
                assert self.s_state[x] in ("S", "P"), f"softmax {x} reads S in state {self.s_state[x]}"
                for _ in range(self.rng.randint(0, 6)):       # the exponentials take a while
                    yield None
                self.s_state[x] = "P"                         # (all four warps write their rows; the last arrival releases)
                self.pfull[x].arrive()
                blk_n += 1
            yield ("wait", self.ofull[x], item_n & 1)
            item_n += 1
            assert self.o_busy[x]
            for _ in range(self.rng.randint(0, 40)):          # reading O back can be slow (a warp that was descheduled)
                yield None
            self.ofree[x].arrive()

    def run(self):
        roles = [self.loader(), self.issuer()] + [self.softmax_warp(x) for x in (0, 1) for _ in range(4)]
        waiting = [None] * len(roles)
        alive = [True] * len(roles)
        idle = 0
        while any(alive):
            self.step += 1
            self.pump()
            order = [i for i in range(len(roles)) if alive[i]]
            self.rng.shuffle(order)
            progressed = False
            for i in order[: self.rng.randint(1, len(order))]:
                if waiting[i] is not None:
                    _, bar, parity = waiting[i]
                    if not bar.passed(parity):
                        continue
                    waiting[i] = None
                try:
                    ev = next(roles[i])
                    progressed = True
                    if ev is not None:
                        waiting[i] = ev
                except StopIteration:
                    alive[i] = False
                    progressed = True
            idle = 0 if (progressed or self.pending or self.mma_queue) else idle + 1
            assert idle < 50, f"deadlock at step {self.step}: waiting = {[w is not None for w in waiting]}"
        # the o_busy flag of four warps is cleared by the model once all four have arrived: emulate at the end
        return self.step


class SimWithO(Sim):
    """O is free again when all four warps of the tile have read it (the fourth arrival completes o_free)."""
    def softmax_warp(self, x):
        gen = super().softmax_warp(x)
        for ev in gen:
            yield ev

    def __init__(self, items, seed):
        super().__init__(items, seed)
        for x in (0, 1):
            bar = self.ofree[x]
            orig = bar.arrive

            def arrive(bar=bar, orig=orig, x=x):
                before = bar.phase
                orig()
                if bar.phase != before:
                    self.o_busy[x] = False
            bar.arrive = arrive


@pytest.mark.parametrize("seed", range(12))
def test_protocol_is_deadlock_free_and_hazard_free(seed):
    rng = random.Random(1000 + seed)
    items = [(rng.choice((1, 2)), rng.choice((1, 1, 2, 3, 4, 7))) for _ in range(rng.randint(1, 9))]
    SimWithO(items, seed).run()


def test_bench_shape_and_window_shape():
    SimWithO([(2, 4)] * 13, 1).run()          # 390-token clips: two-tile items, four key blocks, 13 items per CTA
    SimWithO([(1, 1)] * 40, 2).run()          # 104-token windows: one tile, one block
    SimWithO([(2, 61)], 3).run()              # a 7805-token segment
