"""world_size-2 gloo test (CPU) of the N>1 host logic: under torchrun every rank scores its
np.array_split part of the prompts and rank 0 gathers the scores in prompt order
(reference main.py:67-70,74: executor.map over np.array_split, then sum(outputs, []))."""
import os
import socket
import types

import numpy as np
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_prompts, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import flexible_llm_sharding_b200.main as M

    def fake_run(device, prompt_list, device_manager, args):
        # stand-in for the GPU scorer: encodes (rank, prompt id) so ordering can be checked
        return [np.full((2, 1, 4), float(p[0]), dtype=np.float16) + np.float16(0.25 * rank) for p in prompt_list]

    M.run_sharded_llama = fake_run
    args = types.SimpleNamespace(model_path=".", data_parallel=True, num_batch=1, layer_num_per_shard=1,
                                 storage_location="gpu")
    prompts = [(str(i), ("a", "b")) for i in range(n_prompts)]
    out = M.score_all_distributed(args, prompts)
    if rank == 0:
        q.put([float(o[0, 0, 0]) for o in out])
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


def test_torchrun_data_parallel_partition_and_gather():
    world, n = 2, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # np.array_split(7, 2) -> [0..3], [4..6]; gathered back in prompt order
    want = [i + (0.0 if i < 4 else 0.25) for i in range(n)]
    assert got == want


def test_pipeline_global_shard_order():
    from flexible_llm_sharding_b200 import pipeline, planner
    for n_total, k, G in [(7, 1, 2), (35, 1, 4), (83, 4, 8), (5, 1, 4)]:
        g = pipeline.global_shards(n_total, k, G)
        assert [i for s in g for i in s] == list(range(n_total))      # pipeline order == layer order
        for r in range(G):
            assert g[r::G] == planner.shard_plan(n_total, k, False, G, r)
