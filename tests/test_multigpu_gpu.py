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
