"""What length-bucketed packing (planner.balanced_bins) does on RAGGED prompts: the same prompts are
scored with `pack_order=input` (micro-batches filled in the order given) and `pack_order=length` (the default), through
the public call `model(prompts)`; prints micro-batch statistics, seconds per pass (wall clock around the call, which
ends with the device->host copy of the scores) and whether the scores are bit-identical.

    python tools/ragged_probe.py [--model 7b] [--layers 8] [--prompts 1024] [--passes 3] [--out gpurun_out/ragged.json]
"""
import argparse
import copy
import json
import os
import shutil
import sys
import time
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flexible_llm_sharding_b200 import planner, synthetic
from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="7b")
ap.add_argument("--layers", type=int, default=8)
ap.add_argument("--prompts", type=int, default=1024)
ap.add_argument("--passes", type=int, default=3)
ap.add_argument("--max_prefix", type=int, default=1500)
ap.add_argument("--out", default="")
a = ap.parse_args()

shape = synthetic.LlamaShape.named(a.model, num_hidden_layers=a.layers)
base = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
d = os.path.join(base, f"fls_ragged_{a.model}_L{a.layers}")
if not os.path.exists(os.path.join(d, ".complete")):
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    synthetic.write_layer_files(d, shape, seed=0, init_std=0.02, device="cuda", fast=True)
    open(os.path.join(d, ".complete"), "w").close()

rng = np.random.default_rng(7)
prompts = []
for i in range(a.prompts):
    pre = " ".join(f"w{j}" for j in rng.integers(3, shape.vocab_size, int(rng.integers(16, a.max_prefix))))
    n_opt = int(rng.integers(2, 6))
    prompts.append((pre, tuple(" ".join(f"w{j}" for j in rng.integers(3, shape.vocab_size, int(rng.integers(1, 64))))
                               for _ in range(n_opt))))

res = {"model": f"{a.model} shapes, {a.layers} decoder layers", "prompts": a.prompts}
scores = {}
for order in ("input", "length"):
    args = types.SimpleNamespace(model_path=d, num_batch=1, layer_num_per_shard=1, storage_location="gpu",
                                 max_activation_in_cpu=100, data_parallel=True, disk_folder="./temp", num_gen_token=1,
                                 resident_weights="never", pack_order=order)
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    tok = planner.tokenize_prompts(model.tokenizer, copy.deepcopy(prompts))
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, bucketed=order == "length")
    sizes = [mb.n_rows for mb in packed.micro_batches]
    lp = model.score_tokens(copy.deepcopy(prompts), [5, 6, 7])            # warm-up: page-locks the files
    times, toks = [], []
    for _ in range(a.passes):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        lp = model.score_tokens(copy.deepcopy(prompts), [5, 6, 7])
        times.append(time.perf_counter() - t0)
        toks.append(model.last_timing["tokenize_s"])
    scores[order] = lp
    res[order] = dict(micro_batches=len(sizes), rows_min=min(sizes), rows_max=max(sizes), total_rows=packed.total_rows,
                      s_per_pass=round(float(np.median(times)), 4), s_all=[round(t, 4) for t in times],
                      tokenize_s=round(float(np.median(toks)), 4))
    print(order, res[order], flush=True)
    model.close()
res["scores_bit_identical"] = all(np.array_equal(x, y) for x, y in zip(scores["input"], scores["length"]))
res["speedup_length_over_input"] = round(res["input"]["s_per_pass"] / res["length"]["s_per_pass"], 4)
print(json.dumps(res))
if a.out:
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
