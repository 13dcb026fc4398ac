"""Two-GPU runs (skipped on a one-GPU box): the in-process model-parallel layer pipeline with
peer-to-peer hand-off and the data-parallel thread fan-out must return exactly the single-GPU scores."""
import copy

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
