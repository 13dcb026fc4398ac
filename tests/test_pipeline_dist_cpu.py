"""CPU tests of the cross-process layer pipeline's host logic (flexible-llm-sharding_b200/pipeline_dist.py):

* the hand-off schedule (`item_plan`: which ring slot, which counter values) is executed here by a small
  discrete-event simulation of G ranks with a compute queue and a hand-off queue each - the same ordering the
  CUDA streams, `fls_flag_wait` and `fls_flag_signal` give on the device - and must never deadlock, never
  overwrite a slot whose rows have not been used, and always deliver shard g-1's rows of micro-batch b to shard g;
* the host-side agreement steps run in two gloo processes (world_size 2)."""
import os
import socket

import pytest
import torch.multiprocessing as mp

from flexible_llm_sharding_b200 import pipeline, pipeline_dist as PD, planner


def test_ring_depths():
    assert PD.ring_depths(4, 32) == [32, 3, 3, 3]        # only the wrap-around link buffers a whole pass
    assert PD.ring_depths(2, 2) == [2, 2]
    assert PD.ring_depths(8, 1) == [1] * 8


def _simulate(G, n_total, k, n_mb):
    all_shards = pipeline.global_shards(n_total, k, G)
    last_g = max(g for g, s in enumerate(all_shards) if len(s))
    depths = PD.ring_depths(G, n_mb)
    plans = []
    for r in range(G):
        shards = planner.shard_plan(n_total, k, False, G, r)
        my_globals = [r + i * G for i, s in enumerate(shards) if len(s)]
        plans.append(PD.item_plan(G, r, n_mb, my_globals, last_g, depths))
    ready = [0] * G            # rank r's `ready` counter (written by r-1)
    free = [0] * G             # rank r's `free` counter (written by r+1)
    rings = [[None] * depths[r] for r in range(G)]        # slot -> (g, bi) tag of the rows lying in it
    released = [set() for _ in range(G)]                  # inbound ordinals rank r has released
    cq = [0] * G               # next item index on the compute queue
    pq = [0] * G               # next item index on the hand-off queue (items with k_out)
    computed = [set() for _ in range(G)]
    outs = [[i for i, it in enumerate(p) if it.k_out is not None or it.k_in is not None] for p in plans]
    delivered = 0
    progress = True
    while progress:
        progress = False
        for r in range(G):
            # compute queue: wait(ready) -> kernels
            if cq[r] < len(plans[r]):
                it = plans[r][cq[r]]
                if it.k_in is None or ready[r] >= it.k_in + 1:
                    if it.k_in is not None:
                        tag = rings[r][it.k_in % depths[r]]
                        assert tag == (it.g - 1, it.bi), (r, it, tag)     # the right rows, not yet overwritten
                    computed[r].add(cq[r])
                    cq[r] += 1
                    progress = True
            # hand-off queue: wait(computed) -> [wait(free) -> copy -> signal(ready)] -> release the inbound slot
            if pq[r] < len(outs[r]):
                idx = outs[r][pq[r]]
                it = plans[r][idx]
                nxt = (r + 1) % G
                if idx in computed[r] and (it.wait_free is None or free[r] >= it.wait_free):
                    if it.k_out is not None:
                        slot = it.k_out % depths[nxt]
                        prev_tenant = it.k_out - depths[nxt]
                        assert prev_tenant < 0 or prev_tenant in released[nxt], (r, it)   # never clobber unread rows
                        assert (it.wait_free is None) == (prev_tenant < 0)
                        rings[nxt][slot] = (it.g, it.bi)
                        assert ready[nxt] == it.k_out                                    # counters only grow, by one
                        ready[nxt] = it.k_out + 1
                        delivered += 1
                    if it.k_in is not None:
                        released[r].add(it.k_in)
                        assert free[(r - 1) % G] == it.k_in
                        free[(r - 1) % G] = it.k_in + 1
                    pq[r] += 1
                    progress = True
    for r in range(G):
        assert cq[r] == len(plans[r]) and pq[r] == len(outs[r]), f"rank {r} is stuck (deadlock)"
    n_live = sum(1 for s in all_shards if len(s))
    assert sum(len(p) for p in plans) == n_live * n_mb and delivered == (n_live - 1) * n_mb


@pytest.mark.parametrize("G,n_total,k,n_mb", [(2, 7, 1, 5), (2, 7, 3, 1), (4, 35, 1, 32), (8, 83, 1, 32),
                                              (8, 83, 4, 3), (4, 5, 1, 2), (3, 43, 4, 7), (8, 83, 10, 40)])
def test_handoff_schedule_is_deadlock_free_and_never_clobbers(G, n_total, k, n_mb):
    _simulate(G, n_total, k, n_mb)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import types
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from flexible_llm_sharding_b200 import main as M
    coll = PD.Collectives()
    got = coll.broadcast({"tok": [1, 2, 3]} if rank == 0 else None)
    verdicts = [PD.agree_on_prefix_kv(coll, says[rank]) for says in (("load", "load"), ("save", "save"),
                                                                     ("load", "save"), ("save", "no"))]
    args = types.SimpleNamespace(model_path=".", data_parallel=False)
    dm = M.distributed_device_manager(args)
    dp = M.distributed_device_manager(types.SimpleNamespace(model_path=".", data_parallel=True))
    # An inbound ring that does not fit on ONE rank is an error on EVERY rank (nobody is left waiting in a collective).
    import ctypes as C
    from flexible_llm_sharding_b200 import _lib

    class FakeCtx:                      # stands in for the CUDA side: hands out addresses, rank 1 has no room for 1 GB
        def __init__(self):
            self.next = 0x10000
        def call(self, name, *a):
            if name == "fls_ipc_alloc":
                if rank == 1 and a[0].value > (1 << 30):
                    raise _lib.FlsError(3, "cudaMalloc failed: out of memory")
                a[1]._obj.value = self.next
                self.next += 0x10000
                a[2][0] = rank + 1
            elif name == "fls_ipc_open":
                a[1]._obj.value = 0x900000 + a[0][0]
    ring = PD.IpcRing(FakeCtx(), coll)
    ring.ensure([4, 3], 1 << 20, lambda: None)
    small_ok = ring.capacity == 4 << 20 if rank == 0 else ring.capacity == 3 << 20
    try:
        ring.ensure([4, 3], 1 << 29, lambda: None)        # rank 0 needs 2 GB (granted), rank 1 1.5 GB (refused)
        big = "no error"
    except RuntimeError as e:
        big = str(e)
    q.put((rank, got, verdicts, dm.devices, dm.distributed, dp.devices, dp.distributed, small_ok, big))
    coll.barrier()
    dist.destroy_process_group()


def test_two_rank_host_agreement_over_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, got, verdicts, devices, distributed, dp_devices, dp_distributed, small_ok, big in res:
        assert small_ok
        assert "rank 1: no room in HBM" in big and "--data_parallel True" in big, big
        assert got == {"tok": [1, 2, 3]}
        assert verdicts == ["load", "save", "no", "no"]       # all-or-nothing, identical on both ranks
        assert devices == ["cuda:0", "cuda:1"] and distributed is True        # every rank's GPU, in rank order
        assert dp_devices == [f"cuda:{rank}"] and dp_distributed is False
