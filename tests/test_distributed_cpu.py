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


def _fake_scores(opts):
    """Stand-in scorer: the argmax token of an option is 3 + its length in characters modulo 50."""
    out = np.zeros((len(opts), 1, 64), dtype=np.float16)
    for j, o in enumerate(opts):
        out[j, 0, 3 + len(o) % 50] = 1
    return out


def _gen_worker(rank, world, port, model_path, pp, outp, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import flexible_llm_sharding_b200.main as M
    seen = []

    def fake_run(device, prompt_list, device_manager, args):
        seen.append([(pre, tuple(opts)) for pre, opts in prompt_list])
        return [_fake_scores(opts) for _, opts in prompt_list]

    M.run_sharded_llama = fake_run
    res = M.main(["--model_path", model_path, "--prompt_pickle", pp, "--output_file", outp, "--data_parallel", "True",
                  "--num_gen_token", "3"])
    q.put((rank, seen, None if res is None else [r.shape for r in res]))
    dist.barrier()
    dist.destroy_process_group()


def test_torchrun_generation_loop_broadcasts_the_grown_prompts(tiny, tmp_path):
    """main.py under torchrun with --num_gen_token 3 (reference main.py:65-98): rank 0 appends the decoded greedy
    tokens to every ORIGINAL option (main.py:86-90) and every rank scores its part of the grown prompts in the
    next pass; outputs are [S, 3, V]."""
    import pickle
    from flexible_llm_sharding_b200.utils import load_tokenizer
    prompts = [("w5 w6", ("w7", "w8 w9")), ("w10", ("w11 w12 w13",)), ("w14 w15", ("w16", "w17", "w18"))]
    pp, outp = str(tmp_path / "p.pkl"), str(tmp_path / "s.pkl")
    with open(pp, "wb") as f:
        pickle.dump(prompts, f)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gen_worker, args=(r, world, port, tiny["layer_dir"], pp, outp, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict()
    for _ in range(world):
        rank, seen, shapes = q.get(timeout=180)
        got[rank] = (seen, shapes)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # replay the loop here: np.array_split(3 prompts, 2 ranks) -> rank 0 gets prompts 0-1, rank 1 prompt 2
    tok = load_tokenizer(tiny["layer_dir"])
    cur, tokens = list(prompts), [np.zeros((len(o), 0), dtype=np.int64) for _, o in prompts]
    for step in range(3):
        assert got[0][0][step] == cur[:2] and got[1][0][step] == cur[2:], step
        for i, (pre, opts) in enumerate(cur):
            tokens[i] = np.concatenate([tokens[i], _fake_scores(opts).argmax(-1)], axis=1)
        cur = [(pre, tuple(o + tok.decode(t) for o, t in zip(opts0, tokens[i])))
               for i, (pre, opts0) in enumerate(prompts)]
    assert got[0][1] == [(2, 3, 64), (1, 3, 64), (3, 3, 64)] and got[1][1] is None
    with open(pp.replace(".pkl", "_updated.pkl"), "rb") as f:
        assert pickle.load(f) == cur
    assert all(len(n) > len(o) for (_, new), (_, old) in zip(cur, prompts) for n, o in zip(new, old))
