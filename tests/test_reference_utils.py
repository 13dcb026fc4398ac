"""The golden vectors are the reference's own outputs: the UNMODIFIED /root/reference/utils.py (behind
oracle/reference_shim.py: two `accelerate` stand-ins + the 4.3x decoder-layer call convention) reproduces
tests/golden/*.npz bit for bit.  CPU only.

* always: the committed record of that run (tests/golden/reference_utils.npz + _check.json, written by
  tests/golden/make_golden_from_reference.py) equals the committed vectors the oracle and the CUDA path are
  tested against;
* when /root/reference is present (the build container): the reference is run again, in its default
  model-parallel bookkeeping and in the data-parallel / disk-tier mode, and compared with the committed
  vectors and with the oracle."""
import copy
import json
import os

import numpy as np
import pytest
import torch

from oracle import llama_oracle as O
from oracle import reference_shim as RS

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_committed_reference_run_equals_the_golden_vectors(tiny, tiny_gqa):
    ref = np.load(os.path.join(GOLDEN_DIR, "reference_utils.npz"))
    with open(os.path.join(GOLDEN_DIR, "reference_utils_check.json")) as f:
        check = json.load(f)
    for case in (tiny, tiny_gqa):
        name = case["name"]
        assert np.array_equal(ref[f"{name}/probs32"], case["golden"]["probs32_ref"])
        assert np.array_equal(ref[f"{name}/probs16"], case["golden"]["probs16_ref"])
        for i in range(len(case["edge_prompts"])):
            assert np.array_equal(ref[f"{name}/edge32_{i}"], case["edge_golden"][f"probs32_ref_{i}"])
            assert np.array_equal(ref[f"{name}/edge16_{i}"], case["edge_golden"][f"probs16_ref_{i}"])
    runs = {k: v for k, v in check.items() if not k.startswith("_")}
    assert len(runs) >= 10 and all(v["max_rel_dev_from_committed"] <= 2e-4 for v in runs.values())


@pytest.mark.skipif(not RS.available(), reason="/root/reference is only present in the build container")
@pytest.mark.parametrize("mode", ["mp_cpu_k1", "dp_disk_k3"])
def test_unmodified_reference_loop_reproduces_golden_and_oracle(tiny_gqa, mode, tmp_path):
    case = tiny_gqa
    if mode == "mp_cpu_k1":
        out = RS.reference_scores(case["layer_dir"], case["prompts"], torch.float32, layer_num_per_shard=1,
                                  storage_location="cpu")
    else:
        out = RS.reference_scores(case["layer_dir"], case["prompts"], torch.float32, layer_num_per_shard=3,
                                  storage_location="disk", data_parallel=True, disk_folder=str(tmp_path / "act"),
                                  num_batch=2)
    got = np.stack([s[:, 0] for s in out])
    want = case["golden"]["probs32_ref"]
    np.testing.assert_allclose(got, want, rtol=2e-4, atol=1e-9)
    orc = O.OracleShardedLlama(case["layer_dir"], dtype=torch.float32)
    mine = np.stack([s[:, 0] for s in orc(copy.deepcopy(case["prompts"]))])
    np.testing.assert_allclose(mine, got, rtol=2e-4, atol=1e-9)
    assert (mine.argmax(-1) == got.argmax(-1)).all()


@pytest.mark.skipif(not RS.available(), reason="/root/reference is only present in the build container")
def test_reference_crashes_when_the_whole_model_is_one_shard(tiny_gqa):
    """Reference quirk worth knowing (this build handles the case): with layer_num_per_shard >= layers + 3
    nothing is ever stored, so utils.py:291 deletes a `prompt2layer` key that does not exist."""
    with pytest.raises(KeyError):
        RS.reference_scores(tiny_gqa["layer_dir"], tiny_gqa["prompts"][:1], torch.float32, layer_num_per_shard=7)
