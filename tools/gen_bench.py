"""Generation loop with and without the prefix K/V cache.

    python tools/gen_bench.py [--workload 7b] [--prompts 256] [--tokens 3] [--out gpurun_out/gen.json]

Runs the reference's `--num_gen_token` loop (score, append the greedy token to every option, score again,
main.py:65-90) through the public API twice: recomputing everything each token like the reference, and with
`prefix_kv_cache` on (first pass keeps every layer's prefix K/V, later passes push only the option rows).
Prints seconds per pass for both and checks that the scores are bit-identical.  Weights are streamed from
pinned host memory on every pass in both arms."""
import argparse
import copy
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from flexible_llm_sharding_b200 import synthetic
from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama


def gen_loop(model, args, prompts, n_tokens, use_cache):
    times, notes, scores, rows = [], [], [], []
    cur = copy.deepcopy(prompts)
    for t in range(n_tokens):
        args.prefix_kv_cache = ("last" if t == n_tokens - 1 else "on") if use_cache else "off"
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = model(copy.deepcopy(cur))
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
        notes.append(model.last_timing["prefix_kv"])
        rows.append(model.last_timing["rows"])
        scores.append(np.stack([o[:, 0] for o in out]))
        best = [np.argmax(o[:, 0], axis=-1) for o in out]
        cur = [(pre, tuple(s + " " + model.tokenizer.decode(int(b)) for s, b in zip(opts, bs)))
               for (pre, opts), bs in zip(cur, best)]
    return times, notes, scores, rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="7b")
    ap.add_argument("--prompts", type=int, default=256)
    ap.add_argument("--prefix", type=int, default=384)
    ap.add_argument("--options", type=int, default=4)
    ap.add_argument("--option_tokens", type=int, default=32)
    ap.add_argument("--tokens", type=int, default=3)
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
    args = bench.make_args(md, layer_num_per_shard=1, storage_location="gpu", max_rows=16384)
    model = ShardedLlama(args, DeviceManager(args, [str(dev)]), device=str(dev))
    gen_loop(model, args, prompts, 1, False)                      # warm: layer files -> pinned cache
    res = {"workload": f"Llama-2-{a.workload} arch" + (f" [{alias} distinct decoder layers, others aliased]" if alias else ""),
           "prompts": a.prompts, "prompt_shape": f"prefix {a.prefix} + {a.options} options x {a.option_tokens}",
           "tokens": a.tokens}
    t_off, n_off, s_off, r_off = gen_loop(model, args, prompts, a.tokens, False)
    t_on, n_on, s_on, r_on = gen_loop(model, args, prompts, a.tokens, True)
    res["recompute"] = {"s_per_pass": t_off, "rows_per_pass": r_off, "prefix_kv": n_off}
    res["prefix_kv_cache"] = {"s_per_pass": t_on, "rows_per_pass": r_on, "prefix_kv": n_on}
    res["scores_bit_identical"] = all(np.array_equal(x, y) for x, y in zip(s_off, s_on))
    res["speedup_later_tokens"] = float(np.mean(t_off[1:]) / np.mean(t_on[1:])) if a.tokens > 1 else None
    res["speedup_whole_loop"] = float(sum(t_off) / sum(t_on))
    print(json.dumps(res))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
