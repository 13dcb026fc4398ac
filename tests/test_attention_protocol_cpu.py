"""CPU model of the synchronisation protocol of `attention_persistent_kernel` (csrc/attention_sm100.cu).

The three roles of a CTA - TMA producer, MMA issuer, four softmax warps - are transcribed below as Python generators
that perform the kernel's barrier operations in the kernel's order (same barriers, same counts, same parity
expressions); TMA loads and tensor-core operations complete asynchronously, after random delays, the tensor-core
ones in issue order.  A random scheduler interleaves everything.  For random item sequences (1..9 key tiles per item,
incl. runs of one-tile items) the model checks that

  * the CTA terminates (no wait is left hanging, and no parity wait is satisfied by the wrong phase),
  * a buffer is never overwritten while someone still reads it (Q, the K/P ring, the V ring, S, O, the item slots),
  * everybody works on the data it means to (K(g) under Q K^T of tile g, P(g) and V(g) under P V of tile g, ...).

It is a model of the protocol, not of the arithmetic; the arithmetic is checked on the GPU (tests/test_ops_gpu.py).
"""
import random

import pytest

NK, NV = 3, 2


class Barrier:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0
        if self.pending == 0:
            self.phase += 1
            self.pending = self.count

    def passes(self, parity):            # mbarrier.try_wait.parity
        return (self.phase & 1) != parity


class Wait:
    """`mbar_wait(bar, parity)`; `means` = index of the phase the waiter wants completed (None: vacuous first pass)."""

    def __init__(self, bar, parity, means):
        self.bar, self.parity, self.means = bar, parity, means


class Cta:
    def __init__(self, items, rng):
        self.items, self.rng = items, rng            # items: list of key-tile counts
        self.q_full, self.q_empty = Barrier(1), Barrier(1)
        self.k_full, self.k_empty = [Barrier(1) for _ in range(NK)], [Barrier(1) for _ in range(NK)]
        self.v_full, self.v_empty = [Barrier(1) for _ in range(NV)], [Barrier(1) for _ in range(NV)]
        self.s_full = [Barrier(1), Barrier(1)]
        self.p_ready, self.pv_done = Barrier(4), Barrier(1)
        self.item_full = [Barrier(1), Barrier(1)]
        self.epi_count = 0
        self.slots = [None, None]                    # item index (or -1) published in each slot
        # resources: what they hold, and who still reads them
        self.q = dict(item=None, readers=0)          # readers = Q K^T operations issued and not yet retired
        self.k = [dict(holds=None, readers=0) for _ in range(NK)]   # holds: ("K", g) / ("P", g, n_warps) / ("stage", m, n)
        self.v = [dict(holds=None, readers=0) for _ in range(NV)]
        self.s = [dict(tile=None, unread=0) for _ in range(2)]      # unread = softmax warps that still have to read it
        self.o = dict(item=None, tiles=0, unread=0)  # unread = warps that still have to read the finished O
        self.tc_queue = []                           # tensor-core operations in issue order: (kind, payload, commits)
        self.async_loads = []                        # TMA loads in flight: callables
        self.done_items = [0, 0, 0, 0]

    # ---------------------------------------------------------------- asynchronous engines
    def tma_load(self, what, buf, bar):
        assert buf["readers"] == 0, f"TMA overwrites {what} while it is being read"
        if isinstance(buf.get("holds"), tuple) and buf["holds"][0] == "stage":
            assert buf["holds"][2] == 4, f"TMA overwrites a staging buffer before all warps are done: {buf['holds']}"
        buf["holds" if "holds" in buf else "item"] = ("loading",)

        def land():
            buf["holds" if "holds" in buf else "item"] = what
            bar.arrive()
        self.async_loads.append(land)

    def tc_issue(self, kind, payload, commits):
        self.tc_queue.append([kind, payload, list(commits)])

    def tc_commit(self, bar):
        if self.tc_queue:
            self.tc_queue[-1][2].append(bar)
        else:
            bar.arrive()

    def tc_retire_one(self):
        kind, p, commits = self.tc_queue.pop(0)
        if kind == "qk":
            g, ks = p
            self.q["readers"] -= 1
            self.k[ks]["readers"] -= 1
            self.s[g & 1] = dict(tile=g, unread=4)
        else:
            g, ks, vs, first, last, item = p
            self.k[ks]["readers"] -= 1
            self.v[vs]["readers"] -= 1
            if first:
                assert self.o["unread"] == 0, "the first P V of an item overwrites an O somebody has not read yet"
                self.o = dict(item=item, tiles=1, unread=0)
            else:
                assert self.o["item"] == item
                self.o["tiles"] += 1
            if last:
                self.o["unread"] = 4                         # every softmax warp has to read the finished O
        for b in commits:
            b.arrive()

    # ---------------------------------------------------------------- the three roles
    def producer(self):
        gt, m, nxt = 0, 0, 0
        while True:
            have = nxt < len(self.items)
            item = nxt if have else -1
            nxt += 1
            if m > 0:
                yield Wait(self.q_empty, (m - 1) & 1, m - 1)
            # everybody must have read what the slot held (item m - 2)
            self.slots[m & 1] = item
            self.item_full[m & 1].arrive()
            if not have:
                return
            self.tma_load(("Q", item), self.q, self.q_full)
            for j in range(self.items[item]):
                ks, vs = gt % NK, gt % NV
                yield Wait(self.k_empty[ks], ((gt // NK) & 1) ^ 1, gt // NK - 1 if gt >= NK else None)
                self.tma_load(("K", gt), self.k[ks], self.k_full[ks])
                yield Wait(self.v_empty[vs], ((gt // NV) & 1) ^ 1, gt // NV - 1 if gt >= NV else None)
                self.tma_load(("V", gt), self.v[vs], self.v_full[vs])
                gt += 1
            m += 1

    def mma(self):
        gt, m = 0, 0
        while True:
            yield Wait(self.item_full[m & 1], (m >> 1) & 1, m >> 1)
            item = self.slots[m & 1]
            if item < 0:
                return
            assert item == m, f"MMA warp reads item {item} from the slot, expected {m}"
            n = self.items[item]
            yield Wait(self.q_full, m & 1, m)
            assert self.q["item"] == ("Q", item)

            def issue_qk(j):
                g = gt + j
                ks = g % NK
                yield Wait(self.k_full[ks], (g // NK) & 1, g // NK)
                assert self.k[ks]["holds"] == ("K", g), (self.k[ks]["holds"], g)
                assert self.s[g & 1]["unread"] == 0, f"Q K^T of tile {g} overwrites an S buffer that is still unread"
                self.q["readers"] += 1
                self.k[ks]["readers"] += 1
                self.tc_issue("qk", (g, ks), [self.s_full[g & 1]])
                if j == n - 1:
                    self.tc_commit(self.q_empty)

            yield from issue_qk(0)
            for j in range(n):
                if j + 1 < n:
                    yield from issue_qk(j + 1)
                g = gt + j
                ks, vs = g % NK, g % NV
                yield Wait(self.v_full[vs], (g // NV) & 1, g // NV)
                yield Wait(self.p_ready, g & 1, g)
                assert self.k[ks]["holds"] == ("P", g, 4), (self.k[ks]["holds"], g)
                assert self.v[vs]["holds"] == ("V", g)
                if j == 0:
                    assert self.o["unread"] == 0, "the first P V of an item is issued over an O somebody has not read yet"
                self.k[ks]["readers"] += 1
                self.v[vs]["readers"] += 1
                commits = ([self.k_empty[ks]] if j + 1 < n else []) + [self.v_empty[vs], self.pv_done]
                self.tc_issue("pv", (g, ks, vs, j == 0, j == n - 1, item), commits)
            gt += n
            m += 1

    def softmax(self, w):
        gt, m = 0, 0
        while True:
            yield Wait(self.item_full[m & 1], (m >> 1) & 1, m >> 1)
            item = self.slots[m & 1]
            if item < 0:
                return
            assert item == m, f"softmax warp {w} reads item {item} from the slot, expected {m}"
            n = self.items[item]
            for j in range(n):
                g = gt + j
                ks = g % NK
                yield Wait(self.s_full[g & 1], (g >> 1) & 1, g >> 1)
                assert self.s[g & 1]["tile"] == g, (self.s[g & 1], g)
                self.s[g & 1]["unread"] -= 1
                yield None                                   # exponentials
                # P over the dead K tile
                h = self.k[ks]["holds"]
                assert h == ("K", g) or (h[0] == "P" and h[1] == g), (h, g)
                assert self.k[ks]["readers"] == 0, "P is written while the tensor core still reads the buffer"
                self.k[ks]["holds"] = ("P", g, (h[2] if h[0] == "P" else 0) + 1)
                if j > 0:
                    yield Wait(self.pv_done, (g - 1) & 1, g - 1)
                    assert self.o["item"] == item and self.o["tiles"] >= j
                yield None
                self.p_ready.arrive()
            g_last = gt + n - 1
            gt += n
            yield Wait(self.pv_done, g_last & 1, g_last)
            assert self.o["item"] == item and self.o["tiles"] == n, (self.o, item, n)
            yield None
            ks = g_last % NK
            h = self.k[ks]["holds"]
            assert self.k[ks]["readers"] == 0 and (h == ("P", g_last, 4) or h[0] == "stage"), h
            self.k[ks]["holds"] = ("stage", item, (h[2] if h[0] == "stage" else 0))
            yield None                                       # O read, staged, written out
            assert self.o["item"] == item and self.o["unread"] > 0, "O was overwritten before this warp had read it"
            self.o["unread"] -= 1
            st = self.k[ks]["holds"]
            self.k[ks]["holds"] = ("stage", item, st[2] + 1)
            self.epi_count += 1
            if (self.epi_count - 1) & 3 == 3:
                self.k_empty[ks].arrive()
            self.done_items[w] += 1
            m += 1

    # ---------------------------------------------------------------- scheduler
    def run(self):
        agents = {"producer": self.producer(), "mma": self.mma()}
        agents.update({f"softmax{w}": self.softmax(w) for w in range(4)})
        blocked = {k: None for k in agents}          # None = runnable, else the Wait it sits on
        steps = 0
        while agents:
            steps += 1
            assert steps < 2_000_000
            choices = []
            for name in agents:
                wt = blocked[name]
                if wt is None or wt.bar.passes(wt.parity):
                    choices.append(("agent", name))
            if self.async_loads:
                choices.append(("load", None))
            if self.tc_queue:
                choices.append(("tc", None))
            assert choices, f"deadlock: {[(n, vars(w.bar), w.parity, w.means) for n, w in blocked.items() if w]}"
            kind, name = self.rng.choice(choices)
            if kind == "load":
                self.async_loads.pop(self.rng.randrange(len(self.async_loads)))()
                continue
            if kind == "tc":
                self.tc_retire_one()
                continue
            wt = blocked[name]
            if wt is not None and wt.means is not None:
                assert wt.bar.phase > wt.means, f"{name}: a parity wait passed on phase {wt.bar.phase}, meant {wt.means}"
            try:
                blocked[name] = next(agents[name])
            except StopIteration:
                del agents[name]
                del blocked[name]
        assert not self.async_loads and not self.tc_queue
        assert self.done_items == [len(self.items)] * 4


@pytest.mark.parametrize("seed", range(40))
def test_persistent_attention_protocol(seed):
    rng = random.Random(seed)
    n_items = rng.randint(1, 14)
    style = seed % 4
    if style == 0:
        items = [1] * n_items                                   # runs of one-tile items: the tightest case for the rings
    elif style == 1:
        items = [rng.choice((1, 2)) for _ in range(n_items)]
    elif style == 2:
        items = [rng.choice((6, 8)) for _ in range(n_items)]    # the benchmark's prompt shape
    else:
        items = [rng.randint(1, 9) for _ in range(n_items)]
    Cta(items, rng).run()
