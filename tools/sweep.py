"""Shard-size x activation-tier sweep (BASELINE.json configs[4] in miniature).

    python tools/sweep.py [--workload 70b] [--prompts 64] [--prefix 1792 --options 4 --option_tokens 64]
                          [--k 1 4 10] [--storage gpu cpu disk] [--out gpurun_out/sweep.json]

For every (layer_num_per_shard, storage_location) pair one warm pass and one timed pass through the public
API (`ShardedLlama.__call__`, strings in / numpy probabilities out, weights streamed from pinned host memory
every pass, activations forced to stay in the named tier).  Prints one JSON object with a row per pair:
prompts/s, rows/s, seconds per pass, activation bytes moved host<->device and the weight-stream rate.
The scores of every pair are checked to be bit-identical to the first pair's (scheduling must not change
results).  70b uses 4 distinct decoder layers with the rest aliased (see bench.py).
"""
import argparse
import copy
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from flexible_llm_sharding_b200 import synthetic
from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="70b")
    ap.add_argument("--prompts", type=int, default=64)
    ap.add_argument("--prefix", type=int, default=1792)
    ap.add_argument("--options", type=int, default=4)
    ap.add_argument("--option_tokens", type=int, default=64)
    ap.add_argument("--k", type=int, nargs="+", default=[1, 4, 10])
    ap.add_argument("--storage", nargs="+", default=["gpu", "cpu", "disk"])
    ap.add_argument("--disk_min_k", type=int, default=4,
                    help="skip the disk tier for smaller shards (every shard boundary writes and reads all rows)")
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    import __graft_entry__ as ge
    ge.build()
    shape = synthetic.LlamaShape.named(a.workload)
    alias = 4 if a.workload == "70b" else 0
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    md = bench.model_dir_for(a.workload, 0, 1, dev, alias)
    prompts = synthetic.make_prompts(a.prompts, shape.vocab_size, a.prefix, a.options, a.option_tokens, seed=1234)
    rows_per_prompt = a.prefix + a.options * a.option_tokens
    rows, first = [], None
    disk_dir = tempfile.mkdtemp(prefix="fls_sweep_")
    for k in a.k:
        for storage in a.storage:
            if storage == "disk" and k < a.disk_min_k:
                continue
            args = bench.make_args(md, layer_num_per_shard=k, storage_location=storage, force_storage_tier=True,
                                   disk_folder=disk_dir, max_rows=16384)
            model = ShardedLlama(args, DeviceManager(args, [str(dev)]), device=str(dev))
            model(copy.deepcopy(prompts))                              # warm pass (reads layer files, warms allocators)
            torch.cuda.synchronize()
            model.ctx.call("fls_reset_stats")
            model.ctx.call("fls_set_profiling", 1)
            s0 = model.ctx.stats()
            t0 = time.perf_counter()
            out = model(copy.deepcopy(prompts))
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            s1 = model.ctx.stats()
            prof = model.ctx.profile()
            model.ctx.call("fls_set_profiling", 0)
            kernel_s = sum(v["ms"] for v in prof.values()) / 1e3
            sig = np.stack([o[:, 0] for o in out])
            if first is None:
                first = sig
            same = bool(np.array_equal(sig, first))
            lt = model.last_timing
            rows.append({"layer_num_per_shard": k, "storage_location": storage, "tier_used": lt["activation_tier"],
                         "prompts_per_s": a.prompts / dt, "rows_per_s": a.prompts * rows_per_prompt / dt,
                         "s_per_pass": dt, "kernel_s": kernel_s, "tokenize_s": lt["tokenize_s"], "activation_h2d_GB": lt["act_h2d_bytes"] / 1e9,
                         "activation_d2h_GB": lt["act_d2h_bytes"] / 1e9,
                         "weight_stream_GBps": ((s1.h2d_bytes - s0.h2d_bytes) / max(s1.h2d_ms - s0.h2d_ms, 1e-9) / 1e6),
                         "scores_identical_to_first": same})
            print(json.dumps(rows[-1]), file=sys.stderr, flush=True)
            model.ctx.close()
            del model
            torch.cuda.empty_cache()
    res = {"workload": f"Llama-2-{a.workload} arch" + (f" [{alias} distinct decoder layers, others aliased]" if alias else ""),
           "prompts": a.prompts, "rows_per_prompt": rows_per_prompt,
           "prompt_shape": f"prefix {a.prefix} + {a.options} options x {a.option_tokens}", "rows": rows}
    print(json.dumps(res))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
