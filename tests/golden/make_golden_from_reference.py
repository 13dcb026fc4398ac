"""Pins the committed golden vectors to the reference's OWN loop code.

    python tests/golden/make_golden_from_reference.py        (build container only: needs /root/reference)

tests/golden/*.npz were produced by a restated per-prompt loop over the real transformers modules
(oracle/hf_reference_loop.py).  This script closes the remaining gap: it imports the UNMODIFIED
/root/reference/utils.py (oracle/reference_shim.py supplies the two `accelerate` functions it needs and adapts
the installed transformers' decoder-layer call convention to the 4.31-4.35 one the reference was written
for), runs `ShardedLlama(args, DeviceManager(args, ["cpu"]), device="cpu")` on the same layer files (split by
the reference's own prepare_weights.py) and the same prompts, in several of the reference's modes, and
compares with the committed vectors.  It writes
  tests/golden/reference_utils.npz         the reference's own fp32 / fp16 scores (regular + edge prompts)
  tests/golden/reference_utils_check.json  max deviation of every run from the committed vectors
Both are checked by tests/test_reference_utils.py (which re-runs the reference when it is present).
"""
import json
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import make_golden as mg  # noqa: E402
from flexible_llm_sharding_b200 import synthetic  # noqa: E402
from oracle import reference_shim as RS  # noqa: E402

# (tag, dtype, kwargs of reference_scores): model-parallel bookkeeping with one device (the default mode),
# data-parallel DeviceManager path with the disk tier, shard sizes 1 / 3 / 4, two batches
MODES = [
    ("fp32_k1_cpu_mp", torch.float32, dict(layer_num_per_shard=1, storage_location="cpu")),
    ("fp16_k3_disk_dp_2batches", torch.float16, dict(layer_num_per_shard=3, storage_location="disk",
                                                     data_parallel=True, num_batch=2)),
    # (the whole model in ONE shard crashes the reference itself: utils.py:291 deletes a `prompt2layer` entry that
    #  only utils.py:185 creates, and no activation is stored when the shard ends with lm_head - KeyError)
    ("fp16_k4_gpu_dp", torch.float16, dict(layer_num_per_shard=4, storage_location="gpu", data_parallel=True)),
]


def run_case(name):
    case = mg.CASES[name]
    shape = synthetic.LlamaShape.named(case["shape"])
    golden = dict(np.load(os.path.join(mg.GOLDEN_DIR, f"{name}.npz")))
    edge_golden = dict(np.load(os.path.join(mg.GOLDEN_DIR, f"{name}-edge.npz")))
    out, check = {}, {}
    with tempfile.TemporaryDirectory() as td:
        layer_dir = mg.build_model_dir(case, td)
        assert mg.layer_dir_digest(layer_dir) == str(golden["digest"])
        prompts = synthetic.make_prompts(case["n_prompts"], shape.vocab_size, case["prefix_tokens"],
                                         case["n_options"], case["option_tokens"], seed=mg.PROMPT_SEED, ragged=True)
        edge = synthetic.make_edge_prompts(shape.vocab_size)
        for tag, dtype, kw in MODES:
            kw = dict(kw, disk_folder=os.path.join(td, "act_" + tag))
            bits = "32" if dtype == torch.float32 else "16"
            got = np.stack([s[:, 0] for s in RS.reference_scores(layer_dir, prompts, dtype, **kw)])
            want = golden[f"probs{bits}_ref"]
            assert got.dtype == want.dtype and got.shape == want.shape
            rel = float(np.abs(got.astype(np.float64) / np.maximum(want.astype(np.float64), 1e-30) - 1).max())
            check[f"{name}/{tag}"] = dict(max_rel_dev_from_committed=rel, bit_identical=bool(np.array_equal(got, want)))
            out[f"{name}/probs{bits}"] = got
            if kw.get("num_batch", 1) == 1:
                worst, same = 0.0, True
                for i, s in enumerate(RS.reference_scores(layer_dir, edge, dtype, **kw)):
                    w = edge_golden[f"probs{bits}_ref_{i}"]
                    g = s[:, 0]
                    worst = max(worst, float(np.abs(g.astype(np.float64) / np.maximum(w.astype(np.float64), 1e-30)
                                                    - 1).max()))
                    same &= bool(np.array_equal(g, w))
                    out[f"{name}/edge{bits}_{i}"] = g
                check[f"{name}/{tag}/edge"] = dict(max_rel_dev_from_committed=worst, bit_identical=same)
    return out, check


def main():
    if not RS.available():
        raise SystemExit("needs /root/reference (build container)")
    torch.set_num_threads(os.cpu_count())
    out, check = {}, {}
    for name in mg.CASES:
        o, c = run_case(name)
        out.update(o)
        check.update(c)
    import transformers
    check["_versions"] = dict(torch=torch.__version__, transformers=transformers.__version__)
    for k, v in check.items():
        print(k, v)
    bad = [k for k, v in check.items() if not k.startswith("_") and v["max_rel_dev_from_committed"] > 2e-4]
    assert not bad, bad
    np.savez_compressed(os.path.join(mg.GOLDEN_DIR, "reference_utils.npz"), **out)
    with open(os.path.join(mg.GOLDEN_DIR, "reference_utils_check.json"), "w") as f:
        json.dump(check, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
