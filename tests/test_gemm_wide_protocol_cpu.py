"""CPU model of the synchronisation protocol of `gemm_wide_skew_kernel` (csrc/gemm_wide_sm100.cu): two MMA streams that
share a 6-slot A ring, a 3-slot B ring each, one TMEM accumulator each, a CTA pair whose two TMA producers fill every
slot together, and epilogue warps that drain accumulator 0, then accumulator 1, of every tile.

The roles are transcribed as Python generators that perform the kernel's barrier operations in the kernel's order (same
barriers, counts, parity bookkeeping); TMA loads land and tensor-core operations retire asynchronously after random
delays (each MMA stream's operations in its own issue order); a random scheduler interleaves everything.  Checked:
termination, that no parity wait is satisfied by the wrong phase, that a ring slot is never overwritten while an MMA
that reads it is in flight, that every MMA multiplies the k-block it means to, and that an accumulator is neither
overwritten before every epilogue warp has read it nor read before its last MMA has retired.
(compute-sanitizer's racecheck is not available on the GPU pool; the arithmetic is checked in tests/test_ops_gpu.py.)
"""
import random

import pytest

NA, NB = 6, 3
EPI_WARPS = 8            # four per CTA, both CTAs arrive on the leader's tmem_empty barriers


class Barrier:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0
        if self.pending == 0:
            self.phase += 1
            self.pending = self.count

    def passes(self, parity):
        return (self.phase & 1) != parity


class Wait:
    def __init__(self, bar, parity, means):
        self.bar, self.parity, self.means = bar, parity, means


class Pair:
    """One CTA pair.  `full` barriers live in the leader and count 1 arrive.expect_tx + the bytes of both CTAs' loads
    (modelled as three arrivals); `empty` barriers exist in both CTAs and receive the same multicast commits."""

    def __init__(self, n_tiles, k_blocks, rng):
        self.n_tiles, self.kb, self.rng = n_tiles, k_blocks, rng
        self.a_full = [Barrier(3) for _ in range(NA)]
        self.a_empty = [[Barrier(2) for _ in range(NA)] for _ in range(2)]           # [cta][slot], one commit per stream
        self.b_full = [[Barrier(3) for _ in range(NB)] for _ in range(2)]            # [stream][slot]
        self.b_empty = [[[Barrier(1) for _ in range(NB)] for _ in range(2)] for _ in range(2)]   # [cta][stream][slot]
        self.tmem_full = [Barrier(1), Barrier(1)]
        self.tmem_empty = [Barrier(EPI_WARPS), Barrier(EPI_WARPS)]
        self.a = [dict(holds=[None, None], readers=0) for _ in range(NA)]            # holds per CTA half: (tile, kb)
        self.b = [[dict(holds=[None, None], readers=0) for _ in range(NB)] for _ in range(2)]
        self.acc = [dict(tile=None, blocks=0, unread=0) for _ in range(2)]
        self.queues = [[], []]           # per MMA stream, in issue order
        self.loads = []

    def tma(self, buf, cta, what, bar):
        assert buf["readers"] == 0, f"TMA of CTA {cta} overwrites a slot an MMA still reads: {buf}"
        buf["holds"][cta] = ("loading",)

        def land():
            buf["holds"][cta] = what
            bar.arrive()
        self.loads.append(land)

    def retire(self, s):
        kind, payload, commits = self.queues[s].pop(0)
        if kind == "mma":
            tile, kb, ia, ib = payload
            self.a[ia]["readers"] -= 1
            self.b[s][ib]["readers"] -= 1
            acc = self.acc[s]
            if kb == 0:
                assert acc["unread"] == 0, f"stream {s} overwrites an accumulator that has not been drained"
                acc.update(tile=tile, blocks=1)
            else:
                assert acc["tile"] == tile and acc["blocks"] == kb
                acc["blocks"] += 1
            if kb == self.kb - 1:
                acc["unread"] = EPI_WARPS
        for group in commits:
            for b in group:
                b.arrive()

    # ------------------------------------------------------------------ roles
    def producer(self, cta, s):
        """warp 0 (s = 0: A ring + stream 0's B ring) or warp 2 (s = 1: stream 1's B ring) of CTA `cta`."""
        ia = ib = 0
        pa = pb = 0
        ua = ub = 0                       # uses so far (for the intended phase of a wait)
        for tile in range(self.n_tiles):
            for kb in range(self.kb):
                if s == 0:
                    yield Wait(self.a_empty[cta][ia], pa ^ 1, ua // NA - 1 if ua >= NA else None)
                    if cta == 0:
                        self.a_full[ia].arrive()                  # arrive.expect_tx by the leader
                    self.tma(self.a[ia], cta, (tile, kb), self.a_full[ia])
                    ia += 1
                    ua += 1
                    if ia == NA:
                        ia, pa = 0, pa ^ 1
                yield Wait(self.b_empty[cta][s][ib], pb ^ 1, ub // NB - 1 if ub >= NB else None)
                if cta == 0:
                    self.b_full[s][ib].arrive()
                self.tma(self.b[s][ib], cta, (tile, kb), self.b_full[s][ib])
                ib += 1
                ub += 1
                if ib == NB:
                    ib, pb = 0, pb ^ 1

    def mma(self, s):
        ia = ib = 0
        pa = pb = tphase = 0
        ua = ub = 0
        for tile in range(self.n_tiles):
            yield Wait(self.tmem_empty[s], tphase ^ 1, tile - 1 if tile > 0 else None)
            for kb in range(self.kb):
                yield Wait(self.a_full[ia], pa, ua // NA)
                yield Wait(self.b_full[s][ib], pb, ub // NB)
                assert self.a[ia]["holds"] == [(tile, kb), (tile, kb)], (s, self.a[ia]["holds"], tile, kb)
                assert self.b[s][ib]["holds"] == [(tile, kb), (tile, kb)], (s, self.b[s][ib]["holds"], tile, kb)
                self.a[ia]["readers"] += 1
                self.b[s][ib]["readers"] += 1
                commits = [[self.a_empty[0][ia], self.a_empty[1][ia]], [self.b_empty[0][s][ib], self.b_empty[1][s][ib]]]
                if kb == self.kb - 1:
                    commits.append([self.tmem_full[s]])
                self.queues[s].append(("mma", (tile, kb, ia, ib), commits))
                ia += 1
                ib += 1
                ua += 1
                ub += 1
                if ia == NA:
                    ia, pa = 0, pa ^ 1
                if ib == NB:
                    ib, pb = 0, pb ^ 1
            tphase ^= 1

    def epilogue(self, w):
        tphase = 0
        for tile in range(self.n_tiles):
            for t in range(2):
                yield Wait(self.tmem_full[t], tphase, tile)
                acc = self.acc[t]
                assert acc["tile"] == tile and acc["blocks"] == self.kb and acc["unread"] > 0, (acc, tile)
                yield None                                        # tcgen05.ld, convert, stage, store
                acc["unread"] -= 1
                self.tmem_empty[t].arrive()
            tphase ^= 1

    # ------------------------------------------------------------------ scheduler
    def run(self):
        agents = {}
        for cta in range(2):
            for s in range(2):
                agents[f"producer{s}@cta{cta}"] = self.producer(cta, s)
        for s in range(2):
            agents[f"mma{s}"] = self.mma(s)
        for w in range(EPI_WARPS):
            agents[f"epilogue{w}"] = self.epilogue(w)
        blocked = {k: None for k in agents}
        steps = 0
        while agents:
            steps += 1
            assert steps < 5_000_000
            choices = [("agent", n) for n in agents if blocked[n] is None or blocked[n].bar.passes(blocked[n].parity)]
            if self.loads:
                choices.append(("load", None))
            for s in range(2):
                if self.queues[s]:
                    choices.append(("tc", s))
            assert choices, f"deadlock: {[(n, w.bar.phase, w.bar.pending, w.parity, w.means) for n, w in blocked.items() if w]}"
            kind, name = self.rng.choice(choices)
            if kind == "load":
                self.loads.pop(self.rng.randrange(len(self.loads)))()
            elif kind == "tc":
                self.retire(name)
            else:
                wt = blocked[name]
                if wt is not None and wt.means is not None:
                    assert wt.bar.phase > wt.means, f"{name}: a parity wait passed on phase {wt.bar.phase}, meant {wt.means}"
                try:
                    blocked[name] = next(agents[name])
                except StopIteration:
                    del agents[name]
                    del blocked[name]
        assert not self.loads and not self.queues[0] and not self.queues[1]
        assert all(a["unread"] == 0 for a in self.acc)


@pytest.mark.parametrize("seed", range(24))
def test_wide_skew_gemm_protocol(seed):
    rng = random.Random(seed)
    n_tiles = rng.randint(1, 6)
    k_blocks = rng.choice((1, 2, 3, 5, 6, 7, 13, 64))       # fewer k-blocks than ring slots up to K = 4096
    Pair(n_tiles, k_blocks, rng).run()
