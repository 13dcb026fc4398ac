"""Model-parallel layer pipeline with ONE PROCESS PER GPU (torchrun), hand-off over NVLink.

Reference behaviour (utils.py:150-157,159-213; main.py:71-73): G threads of one process, shard j on GPU j mod G
(shard count rounded up to a multiple of G), a shared dict carrying every prompt's activations from stage to
stage, readiness found by polling `prompt2layer` in sleep(1) loops, run-ahead bounded by
`max_activation_in_cpu`; only GPU 0 tokenises, only the owner of `lm_head` returns scores.

Here the same ownership, with the ranks in separate processes (`pipeline.py` keeps the reference's
threads-of-one-process form):

* rank 0 tokenises once and broadcasts the token ids (a host-side object broadcast; besides the exchange of the
  64-byte buffer handles it is the only collective of a pass); every rank packs the same rows from them;
* every rank exports an inbound RING of micro-batch slots (`fls_ipc_alloc`); its upstream neighbour maps it
  (`fls_ipc_open`) and, after finishing a shard on a micro-batch, copies the rows straight into the next free
  slot over NVLink on a hand-off stream (`fls_peer_copy`), then bumps the consumer's `ready` counter with a
  system-scope release store (`fls_flag_signal`).  The consumer's compute stream waits for that counter on the
  device (`fls_flag_wait`) - no host polling, no host round trip, no Python object per micro-batch;
* back-pressure is the same mechanism in the other direction: the consumer bumps the producer's `free` counter
  once a slot's rows have been used and handed on, and the producer's hand-off stream waits for it before
  re-using the slot.  Ring depth: 3 slots everywhere except rank 0, whose inbound link is the wrap-around of the
  pipeline (shard G-1 -> shard G): rank G-1 delivers micro-batch j about n_mb - G micro-batches before rank 0 has
  finished shard 0 and can take it, so rank 0's ring holds every micro-batch of the pass (what EVERY rank used
  to allocate); all other links run in lock-step;
* a wait that is not satisfied within FLS_HANDOFF_TIMEOUT_MS (default 600 s) gives up and the pass fails on
  that rank instead of hanging the GPU.

The per-rank engine (weight slots, kernels, prefix K/V cache, outputs) is ShardedLlama's own.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from time import perf_counter

from . import _lib, planner
from .pipeline import global_shards

FLAG_BYTES = 256              # word 0: `ready` counter (written by the upstream rank); word 16: `free` counter
FREE_FLAG_WORD = 16           # (written by the downstream rank) - 64 bytes apart
INNER_RING_DEPTH = 3


def ring_depths(G: int, n_mb: int) -> list[int]:
    """Inbound ring depth per rank (see module docstring)."""
    return [max(1, n_mb) if r == 0 else max(1, min(n_mb, INNER_RING_DEPTH)) for r in range(G)]


@dataclass
class Item:
    """One (shard, micro-batch) unit of a rank's pass, in execution order, with its hand-off bookkeeping.
    Counters are relative to the start of the pass; `None` = no such step."""
    g: int                    # global shard index
    shard_i: int              # index among this rank's non-empty shards
    bi: int                   # micro-batch
    k_in: int | None          # ordinal of the inbound item to wait for (ready >= k_in + 1), read from slot k_in % depth
    k_out: int | None         # ordinal of the outbound item, written to the downstream slot k_out % depth_next
    wait_free: int | None     # before writing: downstream `free` counter must be >= this (slot's previous tenant gone)


def item_plan(G: int, rank: int, n_mb: int, my_globals: list[int], last_g: int, depths: list[int]) -> list[Item]:
    """The hand-off schedule of rank `rank`: pure arithmetic, shared by the CUDA executor below and by the
    CPU simulation in tests/test_pipeline_dist_cpu.py."""
    d_next = depths[(rank + 1) % G]
    items, k_in, k_out = [], 0, 0
    for i, g in enumerate(my_globals):
        for bi in range(n_mb):
            ki = ko = wf = None
            if g > 0:
                ki, k_in = k_in, k_in + 1
            if g != last_g:
                ko, k_out = k_out, k_out + 1
                if ko >= d_next:
                    wf = ko - d_next + 1
            items.append(Item(g, i, bi, ki, ko, wf))
    return items


class RawRows:
    """`data_ptr()` of device memory that torch does not own (a ring slot)."""

    def __init__(self, ptr: int):
        self._ptr = int(ptr)

    def data_ptr(self) -> int:
        return self._ptr


class Collectives:
    """The few host-side object exchanges of a pass, over whatever torch.distributed group is up."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)

    def broadcast(self, obj, src=0):
        box = [obj if self.rank == src else None]
        self.dist.broadcast_object_list(box, src=src, group=self.group)
        return box[0]

    def all_gather(self, obj):
        out = [None] * self.world
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def barrier(self):
        self.dist.barrier(group=self.group)


def agree_on_prefix_kv(coll: Collectives, mine_says: str) -> str:
    """A pass may load the prefix K/V cache only if every rank holds it, and fill one only if every rank has
    room (otherwise the ranks would pack different rows)."""
    says = coll.all_gather(mine_says)
    if all(v == "load" for v in says):
        return "load"
    if all(v == "save" for v in says):
        return "save"
    return "no"


class IpcRing:
    """This rank's exported flag block + inbound ring, and its mappings of the neighbours'."""

    def __init__(self, ctx, coll: Collectives):
        self.ctx, self.coll = ctx, coll
        self.rank, self.G = coll.rank, coll.world
        self.timeout_ms = int(os.environ.get("FLS_HANDOFF_TIMEOUT_MS", "600000"))
        self.timeouts_seen = 0
        self.n_in = self.n_out = 0         # items taken from / delivered through the rings since creation
        # flag block: allocated once, never moved, so the counters simply keep growing
        self.flags, self.flags_handle = self._alloc(FLAG_BYTES)
        handles = coll.all_gather(self.flags_handle)
        self.next_flags = self._open(handles[(self.rank + 1) % self.G])
        self.prev_flags = (self.next_flags if (self.rank - 1) % self.G == (self.rank + 1) % self.G
                           else self._open(handles[(self.rank - 1) % self.G]))
        self.slots, self.slots_handle, self.capacity, self.version = None, None, 0, 0
        self.next_slots, self.next_version = None, 0
        self.depths, self.slot_bytes = None, 0

    def _alloc(self, nbytes):
        p, h = C.c_void_p(), (C.c_ubyte * _lib.IPC_HANDLE_BYTES)()
        self.ctx.call("fls_ipc_alloc", C.c_size_t(nbytes), C.byref(p), h)
        return p.value, bytes(h)

    def _open(self, handle: bytes) -> int:
        p = C.c_void_p()
        self.ctx.call("fls_ipc_open", (C.c_ubyte * _lib.IPC_HANDLE_BYTES).from_buffer_copy(handle), C.byref(p))
        return p.value

    def ensure(self, depths: list[int], slot_bytes: int, sync_device):
        """Collective, once per pass: (re)allocate this rank's ring if the pass needs a larger one and (re)map the
        downstream rank's.  Every rank has finished the previous pass before it gets here, so nobody is using
        a ring that is being replaced; the barrier keeps the old one alive until its mapper has let go."""
        slot_bytes = (slot_bytes + 255) & ~255
        if slot_bytes < self.slot_bytes:
            slot_bytes = self.slot_bytes                     # never shrink: the downstream mapping stays valid
        need = max(depths[self.rank] * slot_bytes, self.capacity)
        retired, failed = None, None
        if need > self.capacity or self.slots is None:
            try:
                fresh = self._alloc(need)
            except _lib.FlsError as e:                       # reported to every rank below: nobody is left waiting
                failed = (f"rank {self.rank}: no room in HBM for an inbound ring of {depths[self.rank]} micro-batches "
                          f"({need / 1e9:.2f} GB): {e}")
            else:
                retired = self.slots
                self.slots, self.slots_handle = fresh
                self.capacity, self.version = need, self.version + 1
        infos = self.coll.all_gather((self.version, self.slots_handle, self.capacity, failed))
        problems = [f for _, _, _, f in infos if f]
        if problems:
            raise RuntimeError("; ".join(problems) + ".  The layer pipeline hands activations from GPU to GPU and keeps "
                               "them in HBM; score fewer prompts per call (--num_batch), use more GPUs, or run "
                               "--data_parallel True, whose activations can live in host memory or on disk "
                               "(--storage_location cpu|disk)")
        infos = [i[:3] for i in infos]
        nv, nh, ncap = infos[(self.rank + 1) % self.G]
        remap = nv != self.next_version
        if remap:
            if self.next_slots is not None:
                self.ctx.call("fls_ipc_close", C.c_void_p(self.next_slots))
            self.next_slots, self.next_version = self._open(nh), nv
        if any(v != seen for (v, _, _), seen in zip(infos, getattr(self, "_seen", [0] * self.G))):
            sync_device()
            self.coll.barrier()
            if retired is not None:
                self.ctx.call("fls_ipc_free", C.c_void_p(retired))
        self._seen = [v for v, _, _ in infos]
        self.depths, self.slot_bytes = depths, slot_bytes
        if depths[(self.rank + 1) % self.G] * slot_bytes > ncap:
            raise RuntimeError("downstream ring is smaller than this pass needs (ranks disagree on the plan)")

    # addresses
    def slot_in(self, k: int) -> int:
        return self.slots + (k % self.depths[self.rank]) * self.slot_bytes

    def slot_out(self, k: int) -> int:
        return self.next_slots + (k % self.depths[(self.rank + 1) % self.G]) * self.slot_bytes

    ready_in = property(lambda self: C.c_void_p(self.flags))                              # upstream writes it
    free_out = property(lambda self: C.c_void_p(self.flags + 4 * FREE_FLAG_WORD))         # downstream writes it
    next_ready = property(lambda self: C.c_void_p(self.next_flags))
    prev_free = property(lambda self: C.c_void_p(self.prev_flags + 4 * FREE_FLAG_WORD))

    def close(self):
        """Collective.  An exported buffer may be freed only after every process that mapped it has let go, so:
        drop the mappings, meet, then free."""
        for p in {self.next_slots, self.next_flags, self.prev_flags} - {None}:
            self.ctx.call("fls_ipc_close", C.c_void_p(p))
        self.next_slots = self.next_flags = self.prev_flags = None
        self.coll.barrier()
        for p in (self.slots, self.flags):
            if p is not None:
                self.ctx.call("fls_ipc_free", C.c_void_p(p))
        self.slots = self.flags = None

    def check_timeouts(self):
        n = C.c_uint32()
        self.ctx.call("fls_flag_timeouts", C.byref(n))
        if n.value != self.timeouts_seen:
            self.timeouts_seen = n.value
            raise TimeoutError(f"rank {self.rank}: a pipeline hand-off was not signalled within "
                               f"{self.timeout_ms / 1e3:.0f}s (a peer rank has failed or stalled)")


def _u32(v: int):
    return C.c_uint32(v & 0xFFFFFFFF)


def run_model_parallel_distributed(model, inputs, shards, G, rank, want_probs, answer_ids, return_device=False):
    """One scoring pass of rank `rank` of `G` (ShardedLlama.run calls it on every rank with the same prompt
    count; only rank 0 needs the prompts themselves, main.py:16-17)."""
    import torch
    dm = model.device_manager
    coll = getattr(dm, "collectives", None)
    if coll is None:
        coll = dm.collectives = Collectives()
    if coll.world != G or coll.rank != rank:
        raise RuntimeError(f"device list says rank {rank} of {G}, process group says {coll.rank} of {coll.world}")
    t0 = perf_counter()
    my_layers = [li for sh in shards for li in sh]

    # ---- prompts: rank 0 tokenises (utils.py:156-157,246-259); every rank packs the same rows
    if rank == 0:
        tok = planner.tokenize_prompts(model.tokenizer, inputs)
        for i in range(len(inputs)):
            inputs[i] = None                                          # utils.py:259
        dm.suffix_eos = [t.suffix_eos for t in tok]
        ask = None
        if tok and model._kv_policy() != "off":
            ask = dict(sig=model._kv_signature(tok), policy=model._kv_policy(), prefix_rows=sum(t.len_p for t in tok),
                       act_rows=sum(t.rows for t in tok), head_rows=sum(t.n_options for t in tok))
        tok, ask = coll.broadcast((tok, ask))
    else:
        tok, ask = coll.broadcast(None)
    if not tok:
        return None, None
    kv_mode, kv_cache, kv_note = _lib.KV_NONE, None, "off"
    if ask is not None:
        model._configure(shards)
        if ask["sig"] in model._kv:
            mine_says = "load"
        elif ask["policy"] == "on" and model._kv_need(my_layers, ask["prefix_rows"], ask["act_rows"], ask["head_rows"])[1]:
            mine_says = "save"
        else:
            mine_says = "no"
        verdict = agree_on_prefix_kv(coll, mine_says)
        if verdict == "load":
            kv_mode, kv_note = _lib.KV_LOAD, "load"
            kv_cache = model._kv[ask["sig"]] if ask["policy"] == "on" else model._kv.pop(ask["sig"])
        else:
            model._kv.pop(ask["sig"], None)                           # a partial set of ranks is of no use
            if verdict == "save":
                kv_mode, kv_note = _lib.KV_SAVE, "save"
                kv_cache = model._kv_alloc(my_layers, ask["prefix_rows"])
            else:
                kv_note = "off (last pass)" if ask["policy"] == "last" else "skipped: no room on every rank"
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, model.prefix_causal,
                                  cached_prefix=kv_mode == _lib.KV_LOAD,
                                  bucketed=model.pack_order == "length" and model._kv_policy() == "off")
    t_tok = perf_counter() - t0

    model._configure(shards)
    dev = model.torch_device
    n_total = len(model.layer_names)
    all_shards = global_shards(n_total, model.args.layer_num_per_shard, G)
    last_g = max(g for g, s in enumerate(all_shards) if len(s))
    mine = [(rank + i * G, s) for i, s in enumerate(shards) if len(s)]
    owns_head = any(n_total - 1 in s for _, s in mine)
    call = model.ctx.call
    H = model.hidden
    mbs = packed.micro_batches

    with torch.cuda.device(dev), model._on_compute_stream(dev) as stream:
        ring = getattr(model, "_ipc_ring", None)
        if ring is None:
            ring = model._ipc_ring = IpcRing(model.ctx, coll)
        depths = ring_depths(G, len(mbs))
        ring.ensure(depths, max(max(mb.n_rows, mb.n_head) for mb in mbs) * H * 2,
                    lambda: torch.cuda.synchronize(dev))
        items = item_plan(G, rank, len(mbs), [g for g, _ in mine], last_g, depths)

        sp = C.c_void_p(stream.cuda_stream)
        if getattr(model, "_p2p_stream", None) is None:
            model._p2p_stream = torch.cuda.Stream(dev)
        p2p = model._p2p_stream
        p2p_sp = C.c_void_p(p2p.cuda_stream)
        dv = model._device_inputs(packed, dev, kv_mode)
        heads = torch.empty((packed.total_head_rows, H), dtype=torch.float16, device=dev)
        out = model._alloc_outputs(packed, dev, want_probs and owns_head, answer_ids if owns_head else None)
        # rows of the shard that has no upstream (global shard 0, the embedding's owner): three rotating buffers
        work, work_free = None, [None] * 3
        if any(g == 0 for g, _ in mine):
            work = [torch.empty((max(mb.n_rows for mb in mbs), H), dtype=torch.float16, device=dev) for _ in range(3)]
        base_in, base_out, T = ring.n_in, ring.n_out, ring.timeout_ms
        p2p_bytes = 0

        my_live = [s for _, s in mine]
        if mine:
            model._stage_shard(my_live, 0)
        for it in items:
            shard, mb = my_live[it.shard_i], mbs[it.bi]
            slot = it.shard_i % model._n_slots
            if it.bi == 0:
                if it.shard_i + 1 < len(my_live):
                    model._stage_shard(my_live, it.shard_i + 1)       # next shard's H2D overlaps this shard's compute
                call("fls_acquire_shard", slot, sp)
            hv = heads[mb.head_lo:mb.head_lo + mb.n_head]
            if it.k_in is not None:
                call("fls_flag_wait", ring.ready_in, _u32(base_in + it.k_in + 1), T, sp)
                x = RawRows(ring.slot_in(it.k_in))                    # token rows, or scored rows after model.norm
            else:
                j = it.bi % 3
                if work_free[j] is not None:
                    stream.wait_event(work_free[j])                   # its previous rows have been handed on
                x = work[j][:mb.n_rows]
            y, still_tokens = model._run_layers(slot, shard, it.bi, mb, x, hv, dv, out, sp, kv_mode, kv_cache)
            if it.k_out is not None or it.k_in is not None:
                # Everything that signals rides the hand-off stream, in item order: the counters only ever grow, and
                # the next micro-batch's kernels start while these rows cross NVLink.
                computed = torch.cuda.Event()
                computed.record(stream)
                p2p.wait_event(computed)
            if it.k_out is not None:
                nbytes = (mb.n_rows if still_tokens else mb.n_head) * H * 2
                if it.wait_free is not None:   # the slot's previous tenant has been used and handed on
                    call("fls_flag_wait", ring.free_out, _u32(base_out + it.wait_free), T, p2p_sp)
                call("fls_peer_copy", C.c_void_p(ring.slot_out(it.k_out)), -1, C.c_void_p(y.data_ptr()),
                     C.c_size_t(nbytes), p2p_sp)
                call("fls_flag_signal", ring.next_ready, _u32(base_out + it.k_out + 1), p2p_sp)
                p2p_bytes += nbytes
                if it.k_in is None:
                    work_free[it.bi % 3] = torch.cuda.Event()
                    work_free[it.bi % 3].record(p2p)
            if it.k_in is not None:            # inbound slot: computed in place (and read by the copy), now free
                call("fls_flag_signal", ring.prev_free, _u32(base_in + it.k_in + 1), p2p_sp)
            if it.bi == len(mbs) - 1:
                call("fls_release_shard", slot, sp)
        ring.n_in = base_in + sum(1 for it in items if it.k_in is not None)
        ring.n_out = base_out + sum(1 for it in items if it.k_out is not None)

        result, out_bytes = None, 0
        if owns_head:
            result, out_bytes = model._fetch_result(out, want_probs, stream, return_device)
        stream.wait_stream(p2p)
        stream.synchronize()
        ring.check_timeouts()
        if kv_mode == _lib.KV_SAVE and all(coll.all_gather(True)):    # every rank finished: the caches are complete
            model._kv[ask["sig"]] = kv_cache
        model.last_timing = dict(tokenize_s=t_tok, total_s=perf_counter() - t0, h2d_input_bytes=dv["bytes"],
                                 d2h_output_bytes=out_bytes, activation_tier="gpu", p2p_bytes=p2p_bytes,
                                 prefix_kv=kv_note, rows=packed.total_rows, ring_bytes=ring.capacity,
                                 ring_depth=depths[rank])
        model._keepalive = (dv, out, heads, work)
    return (result, packed) if owns_head else (None, packed)
