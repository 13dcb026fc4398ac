import hashlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.join(ROOT, "tests") not in sys.path:
    sys.path.insert(0, os.path.join(ROOT, "tests"))      # tests/_parity.py

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def layer_dir_digest(layer_dir: str) -> str:
    h = hashlib.sha256()
    for fn in sorted(os.listdir(layer_dir)):
        if fn.endswith(".safetensors"):
            h.update(fn.encode())
            with open(os.path.join(layer_dir, fn), "rb") as f:
                h.update(f.read())
    return h.hexdigest()


def _golden_cases():
    sys.path.insert(0, GOLDEN_DIR)
    import make_golden
    return make_golden


@pytest.fixture(scope="session")
def golden_mod():
    return _golden_cases()


def _build(case_name, tmp_path_factory):
    """Regenerates the golden model directory with THIS repo's splitter (the reference's
    prepare_weights.py is not present on the GPU box) and loads the committed vectors."""
    from flexible_llm_sharding_b200 import synthetic
    from flexible_llm_sharding_b200.prepare_weights import split_into_layers
    mg = _golden_cases()
    case = mg.CASES[case_name]
    shape = synthetic.LlamaShape.named(case["shape"])
    td = tmp_path_factory.mktemp(case_name)
    bin_dir, layer_dir = str(td / "hf_bin"), str(td / "layers")
    synthetic.write_hf_bin_checkpoint(bin_dir, shape, seed=mg.SEED, init_std=mg.INIT_STD)
    split_into_layers(bin_dir, layer_dir)
    prompts = synthetic.make_prompts(case["n_prompts"], shape.vocab_size, case["prefix_tokens"],
                                     case["n_options"], case["option_tokens"], seed=mg.PROMPT_SEED,
                                     ragged=True)
    golden = dict(np.load(os.path.join(GOLDEN_DIR, f"{case_name}.npz")))
    edge_golden = dict(np.load(os.path.join(GOLDEN_DIR, f"{case_name}-edge.npz")))
    return dict(name=case_name, shape=shape, bin_dir=bin_dir, layer_dir=layer_dir, prompts=prompts,
                golden=golden, digest=layer_dir_digest(layer_dir),
                edge_prompts=synthetic.make_edge_prompts(shape.vocab_size), edge_golden=edge_golden)


@pytest.fixture(scope="session")
def tiny(tmp_path_factory):
    return _build("tiny", tmp_path_factory)


@pytest.fixture(scope="session")
def tiny_gqa(tmp_path_factory):
    return _build("tiny-gqa", tmp_path_factory)
