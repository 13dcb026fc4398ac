"""Two-GPU runs (skipped on a one-GPU box): the in-process model-parallel layer pipeline with
peer-to-peer hand-off and the data-parallel thread fan-out must return exactly the single-GPU scores."""
import copy
import os

import numpy as np
import pytest
import torch

from tests.test_scoring_gpu import make_args, run_model

pytestmark = pytest.mark.gpu


def _need_two():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")


@pytest.mark.parametrize("k,max_rows", [(1, 16384), (2, 200), (3, 100)])
def test_model_parallel_pipeline_matches_single_gpu(tiny, k, max_rows):
    _need_two()
    from flexible_llm_sharding_b200.main import score_all
    _, base = run_model(tiny)
    args = make_args(tiny["layer_dir"], layer_num_per_shard=k, max_rows=max_rows, num_batch=2)
    out = score_all(args, copy.deepcopy(tiny["prompts"]), ["cuda:0", "cuda:1"])
    assert len(out) == len(base)
    for a, b in zip(out, base):
        assert np.array_equal(a, b)


def test_data_parallel_threads_match_single_gpu(tiny):
    _need_two()
    from flexible_llm_sharding_b200.main import score_all
    _, base = run_model(tiny)
    args = make_args(tiny["layer_dir"], data_parallel=True)
    out = score_all(args, copy.deepcopy(tiny["prompts"]), ["cuda:0", "cuda:1"])
    assert len(out) == len(base)
    for a, b in zip(out, base):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("data_parallel", [False, True])
def test_generation_loop_with_prefix_kv_cache_on_two_gpus(tiny_gqa, data_parallel):
    """The prefix K/V cache in both multi-GPU modes: every rank caches the layers (MP) or the prompts (DP) it
    owns; three passes with growing options give exactly the scores of a single GPU recomputing everything."""
    _need_two()
    from flexible_llm_sharding_b200.main import score_all
    from flexible_llm_sharding_b200.utils import DeviceManager
    from tests.test_scoring_gpu import _extend_options
    p1 = copy.deepcopy(tiny_gqa["prompts"])
    p2 = _extend_options(p1, ["w7", "w11", "w5"])
    p3 = _extend_options(p2, ["w9", "w4"])
    want = [run_model(tiny_gqa, prompts=copy.deepcopy(p))[1] for p in (p1, p2, p3)]
    devices = ["cuda:0", "cuda:1"]
    args = make_args(tiny_gqa["layer_dir"], layer_num_per_shard=2, max_rows=300, data_parallel=data_parallel,
                     prefix_kv_cache="on")
    dm = DeviceManager(args, devices)
    notes = []
    for step, p in enumerate((p1, p2, p3)):
        if step == 2:
            args.prefix_kv_cache = "last"
        out = score_all(args, copy.deepcopy(p), devices, dm)
        notes.append([dm.models[d].last_timing["prefix_kv"] for d in devices])
        assert len(out) == len(want[step])
        for a, b in zip(out, want[step]):
            assert np.array_equal(a, b), (step, notes)
    assert notes == [["save", "save"], ["load", "load"], ["load", "load"]]
    assert not any(dm.models[d]._kv for d in devices)


# ------------------------------------------------------------------ one PROCESS per GPU (torchrun form)
def _dist_worker(rank, world, port, layer_dir, prompts, kw, passes, q):
    import os
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), FLS_HANDOFF_TIMEOUT_MS="60000")
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from flexible_llm_sharding_b200 import main as M
    args = make_args(layer_dir, **kw)
    dm = M.distributed_device_manager(args)
    outs, notes = [], []
    for step, p in enumerate(passes):
        if "prefix_kv_cache" in kw:
            args.prefix_kv_cache = "last" if step == len(passes) - 1 else "on"
        outs.append(M.score_all_distributed(args, copy.deepcopy(p), dm))
        model = dm.models[f"cuda:{rank}"]
        notes.append((model.last_timing.get("prefix_kv"), model.last_timing.get("ring_depth"),
                      model.last_timing.get("p2p_bytes", 0)))
    q.put((rank, outs if rank == 0 else None, notes))
    for m in dm.models.values():
        m.close()                      # collective: rings are unmapped everywhere before their owners free them
    dist.barrier()
    dist.destroy_process_group()


def _run_distributed(world, layer_dir, passes, **kw):
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_dist_worker, args=(r, world, port, layer_dir, None, kw, passes, q))
             for r in range(world)]
    for p in procs:
        p.start()
    res = dict()
    for _ in range(world):
        rank, outs, notes = q.get(timeout=600)
        res[rank] = (outs, notes)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return res


@pytest.mark.parametrize("k,max_rows,num_batch", [(1, 16384, 1), (2, 200, 2), (3, 100, 1)])
def test_cross_process_pipeline_matches_single_gpu(tiny, k, max_rows, num_batch):
    """torchrun form of the reference's default mode: one process per GPU, shard j on rank j mod G, rows handed
    rank to rank through IPC-mapped rings with device-side flags.  Bit-identical to one GPU."""
    _need_two()
    _, base = run_model(tiny)
    res = _run_distributed(2, tiny["layer_dir"], [tiny["prompts"]], layer_num_per_shard=k, max_rows=max_rows,
                           num_batch=num_batch)
    out = res[0][0][0]
    assert len(out) == len(base)
    for a, b in zip(out, base):
        assert np.array_equal(a, b)
    assert res[0][1][0][2] > 0 and res[1][1][0][2] >= 0          # rank 0 handed rows on over NVLink


def test_cross_process_pipeline_with_prefix_kv_cache(tiny_gqa):
    _need_two()
    from tests.test_scoring_gpu import _extend_options
    p1 = copy.deepcopy(tiny_gqa["prompts"])
    p2 = _extend_options(p1, ["w7", "w11", "w5"])
    p3 = _extend_options(p2, ["w9", "w4"])
    want = [run_model(tiny_gqa, prompts=copy.deepcopy(p))[1] for p in (p1, p2, p3)]
    res = _run_distributed(2, tiny_gqa["layer_dir"], [p1, p2, p3], layer_num_per_shard=2, max_rows=300,
                           prefix_kv_cache="on")
    for step in range(3):
        for a, b in zip(res[0][0][step], want[step]):
            assert np.array_equal(a, b), step
    assert [n[0] for n in res[0][1]] == ["save", "load", "load"] == [n[0] for n in res[1][1]]


def _cli_worker(rank, world, port, argv, q, env=None):
    import os
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), LOCAL_WORLD_SIZE=str(world), **(env or {}))
    os.environ.pop("FLS_STAGE_DIR", None)
    from flexible_llm_sharding_b200 import main as M
    from flexible_llm_sharding_b200 import _lib
    res = M.main(argv)
    q.put((rank, None if res is None else len(res), _lib.load().fls_version()))
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("data_parallel,storage", [(True, "gpu"), (False, "gpu"), (True, "disk")])
def test_torchrun_cli_on_two_gpus(tiny, tmp_path, data_parallel, storage):
    """`torchrun --nproc-per-node 2 main.py ...` as the user would run it: with --data_parallel the disk-resident
    layer files are staged once on tmpfs (and the staging directory is gone afterwards), without it the two ranks form
    the layer pipeline; the pickles rank 0 writes hold the single-GPU scores, bit for bit.  With the disk tier forced
    both ranks write their per-prompt `.npy` files into ONE folder at the same time: the file tag is the global rank,
    so they do not collide, and the folder is empty afterwards."""
    _need_two()
    import pickle
    import socket
    import torch.multiprocessing as mp
    _, base = run_model(tiny)
    pp, outp = tmp_path / "prompts.pkl", tmp_path / "scores.pkl"
    with open(pp, "wb") as f:
        pickle.dump(tiny["prompts"], f)
    argv = ["--model_path", tiny["layer_dir"], "--prompt_pickle", str(pp), "--output_file", str(outp),
            "--storage_location", storage, "--disk_folder", str(tmp_path / "act"), "--layer_num_per_shard", "2"] \
        + (["--data_parallel", "True"] if data_parallel else [])
    env = dict(FLS_FORCE_STORAGE="1", FLS_MAX_ROWS="200") if storage == "disk" else None
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_cli_worker, args=(r, 2, port, argv, q, env)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict()
    for _ in range(2):
        rank, n, _version = q.get(timeout=600)
        res[rank] = n
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res[0] == len(base) and res[1] is None
    assert not os.path.isdir(f"/dev/shm/fls_stage_{os.getuid()}_{port}")     # the automatic staging directory is gone
    if storage == "disk":
        assert not [f for f in os.listdir(tmp_path / "act") if f.endswith(".npy")]
    with open(outp, "rb") as f:
        scores = pickle.load(f)
    for a, b in zip(scores, base):
        assert np.array_equal(a[:, :1], b)
