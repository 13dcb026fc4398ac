"""Model-parallel layer pipeline across the GPUs of one process.

Reference behaviour (utils.py:150-157,185-195; main.py:71-73): G host threads, one per GPU; the shard
count is rounded up to a multiple of G and GPU r owns shards r::G; a Python dict shared through the
DeviceManager carries every prompt's activations from one stage to the next (`.to(self.device)`),
with readiness found by polling `prompt2layer` in `sleep(1)` loops; only GPU 0 tokenises and the GPU
that owns `lm_head` returns scores.

Here the same ownership and thread model, but: rank 0 tokenises/packs once and publishes the packed
prompts; every rank keeps the full `[total_rows, H]` row matrix in its own HBM; after finishing shard j
on a micro-batch the owner writes the rows straight into the next owner's matrix with one peer copy over
NVLink (`fls_peer_copy` = cudaMemcpyPeerAsync on the producer's stream) and publishes a CUDA event; the
consumer's stream waits on that event (no host polling, no sleep, no host round trip).  Micro-batches
form the wave-front: rank r works on micro-batch b of shard j while rank r+1 works on micro-batch b-1 of
shard j+1.  A failing rank aborts the others instead of leaving them waiting.
"""
from __future__ import annotations

import ctypes as C
from time import perf_counter

import torch

from . import _lib, planner

_WAIT_S = 600.0


def _board_post(dm, key, value):
    with dm.condition:
        dm.handoff[key] = value
        dm.condition.notify_all()


def _board_wait(dm, key):
    t0 = perf_counter()
    with dm.condition:
        while key not in dm.handoff:
            if getattr(dm, "failed", None) is not None:
                raise RuntimeError(f"pipeline peer failed: {dm.failed!r}")
            if perf_counter() - t0 > _WAIT_S:
                raise TimeoutError(f"pipeline hand-off {key} not published within {_WAIT_S:.0f}s")
            dm.condition.wait(timeout=1.0)
        return dm.handoff[key]


def _barrier(dm, serial, rank, G):
    _board_post(dm, ("done", serial, rank), True)
    for r in range(G):
        _board_wait(dm, ("done", serial, r))


def global_shards(n_layers_total: int, k: int, G: int):
    """All shards in pipeline order: global shard g is owned by rank g % G (utils.py:151-153)."""
    per_rank = [planner.shard_plan(n_layers_total, k, False, G, r) for r in range(G)]
    n = len(per_rank[0])
    return [per_rank[g % G][g // G] for g in range(n * G)]


def run_model_parallel(model, inputs, shards, G, rank, want_probs, answer_ids, return_device=False):
    dm = model.device_manager
    serial = getattr(model, "_mp_serial", 0)
    model._mp_serial = serial + 1
    try:
        return _run(model, dm, inputs, shards, G, rank, serial, want_probs, answer_ids, return_device)
    except BaseException as e:       # wake the peers up instead of letting them wait for a dead rank
        with dm.condition:
            dm.failed = e
            dm.condition.notify_all()
        raise


def _run(model, dm, inputs, shards, G, rank, serial, want_probs, answer_ids, return_device):
    t0 = perf_counter()
    my_layers = [li for sh in shards for li in sh]
    if rank == 0:
        tok = planner.tokenize_prompts(model.tokenizer, inputs)
        for i in range(len(inputs)):
            inputs[i] = None
        dm.suffix_eos = [t.suffix_eos for t in tok]                  # utils.py:157,258 (kept for inspection)
        ask = None
        if tok and model._kv_policy() != "off":
            ask = dict(sig=model._kv_signature(tok), policy=model._kv_policy(), prefix_rows=sum(t.len_p for t in tok),
                       act_rows=sum(t.rows for t in tok), head_rows=sum(t.n_options for t in tok))
        _board_post(dm, ("kv_ask", serial), ask)
    else:
        tok = None
        ask = _board_wait(dm, ("kv_ask", serial))
    # Prefix K/V cache (utils.py:_kv_plan): every rank caches the layers it owns; the pass loads only if every
    # rank holds the prefixes and saves only if every rank has room, so all ranks pack the same rows.
    kv_mode, kv_cache, kv_note = _lib.KV_NONE, None, "off"
    if ask is not None:
        model._configure(shards)
        if ask["sig"] in model._kv:
            mine_says = "load"
        elif ask["policy"] == "on" and model._kv_need(my_layers, ask["prefix_rows"], ask["act_rows"], ask["head_rows"])[1]:
            mine_says = "save"
        else:
            mine_says = "no"
        _board_post(dm, ("kv_say", serial, rank), mine_says)
        says = [_board_wait(dm, ("kv_say", serial, r)) for r in range(G)]
        if all(v == "load" for v in says):
            kv_mode, kv_note = _lib.KV_LOAD, "load"
            kv_cache = model._kv[ask["sig"]] if ask["policy"] == "on" else model._kv.pop(ask["sig"])
        else:
            model._kv.pop(ask["sig"], None)                          # a partial set of ranks is of no use
            if all(v == "save" for v in says):
                kv_mode, kv_note = _lib.KV_SAVE, "save"
                kv_cache = model._kv_alloc(my_layers, ask["prefix_rows"])
            else:
                kv_note = "off (last pass)" if ask["policy"] == "last" else "skipped: no room on every rank"
    if rank == 0:
        packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, model.prefix_causal,
                                      cached_prefix=kv_mode == _lib.KV_LOAD,
                                      bucketed=model.pack_order == "length" and model._kv_policy() == "off") if tok else None
        _board_post(dm, ("packed", serial), packed)
    else:
        packed = _board_wait(dm, ("packed", serial))
    if packed is None:
        return None, None
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

    with torch.cuda.device(dev), model._on_compute_stream(dev) as stream:
        sp = C.c_void_p(stream.cuda_stream)
        if getattr(model, "_p2p_stream", None) is None:
            model._p2p_stream = torch.cuda.Stream(dev)
        p2p = model._p2p_stream
        p2p_sp = C.c_void_p(p2p.cuda_stream)
        dv = model._device_inputs(packed, dev, kv_mode)
        x_all = torch.empty((packed.total_rows, H), dtype=torch.float16, device=dev)
        heads = torch.empty((packed.total_head_rows, H), dtype=torch.float16, device=dev)
        out = model._alloc_outputs(packed, dev, want_probs and owns_head, answer_ids if owns_head else None)
        _board_post(dm, ("buffers", serial, rank), (x_all, heads, dev.index or 0))
        nxt = None
        p2p_bytes = 0

        my_live = [s for _, s in mine]
        if mine:
            model._stage_shard(my_live, 0)
        for i, (g, shard) in enumerate(mine):
            slot = i % model._n_slots
            if i + 1 < len(mine):
                model._stage_shard(my_live, i + 1)
            call("fls_acquire_shard", slot, sp)
            tokens_in = shard[0] <= n_total - 2
            for bi, mb in enumerate(packed.micro_batches):
                hv = heads[mb.head_lo:mb.head_lo + mb.n_head]
                x = x_all[mb.row_lo:mb.row_lo + mb.n_rows] if tokens_in else hv
                if g > 0:
                    ev = _board_wait(dm, ("ev", serial, g - 1, bi))      # producer has enqueued its copy
                    stream.wait_event(ev)                                 # device-side wait, no host sync
                y, still_tokens = model._run_layers(slot, shard, bi, mb, x, hv, dv, out, sp, kv_mode, kv_cache)
                if g != last_g:
                    if nxt is None:
                        nxt = _board_wait(dm, ("buffers", serial, (rank + 1) % G))
                    nx_all, nheads, ndev = nxt
                    dst = (nx_all[mb.row_lo:mb.row_lo + mb.n_rows] if still_tokens
                           else nheads[mb.head_lo:mb.head_lo + mb.n_head])
                    nbytes = y.numel() * 2
                    # the hand-off rides its own stream: the next micro-batch's kernels start while the rows
                    # cross NVLink (nobody writes these rows again before the consumer has used them)
                    computed = torch.cuda.Event()
                    computed.record(stream)
                    p2p.wait_event(computed)
                    call("fls_peer_copy", dst.data_ptr(), ndev, y.data_ptr(), C.c_size_t(nbytes), p2p_sp)
                    p2p_bytes += nbytes
                    done = torch.cuda.Event()
                    done.record(p2p)
                    _board_post(dm, ("ev", serial, g, bi), done)
            call("fls_release_shard", slot, sp)

        result, out_bytes = None, 0
        if owns_head:
            result, out_bytes = model._fetch_result(out, want_probs, stream, return_device)
        stream.wait_stream(p2p)
        stream.synchronize()
        _barrier(dm, serial, rank, G)          # peers may still be writing into / reading from our buffers
        if kv_mode == _lib.KV_SAVE:
            model._kv[ask["sig"]] = kv_cache               # every rank has passed the barrier: the pass succeeded
        model.last_timing = dict(tokenize_s=t_tok, total_s=perf_counter() - t0, h2d_input_bytes=dv["bytes"],
                                 d2h_output_bytes=out_bytes, activation_tier="gpu", p2p_bytes=p2p_bytes,
                                 prefix_kv=kv_note, rows=packed.total_rows)
        if rank == 0:
            # Entries of call serial-2 and older can go: whoever posted done(serial-1) has left barrier(serial-2).
            with dm.condition:
                for key in [k for k in dm.handoff if len(k) > 1 and k[1] <= serial - 2]:
                    del dm.handoff[key]
    return (result, packed) if owns_head else (None, packed)
