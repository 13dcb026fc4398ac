"""Per-kernel-class rates of whole decoder layers at a model's shapes under different kernel options, without the
cost of a full pass: a 2-layer random-init model with resident weights, 128 prompts x 512 rows (4 micro-batches of
16384 rows), `--passes` passes per setting, CUDA events around every launch (fls_set_profiling).

    python tools/layer_probe.py --model 70b --set gemm_wide_tiles=2,fuse_rope=1 --set gemm_wide_tiles=2,fuse_rope=0
"""
import argparse
import json
import os
import shutil
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from flexible_llm_sharding_b200 import planner, synthetic
from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="70b")
ap.add_argument("--prompts", type=int, default=128)
ap.add_argument("--passes", type=int, default=6)
ap.add_argument("--set", action="append", default=[], help="comma-separated option=value list for fls_set_option")
a = ap.parse_args()

shape = synthetic.LlamaShape.named(a.model, num_hidden_layers=2)
d = tempfile.mkdtemp(prefix="fls_layer_probe_", dir=bench.bench_base_dir())
try:
    synthetic.write_layer_files(d, shape, seed=0, init_std=0.02, device="cuda:0", fast=True)
    prompts = synthetic.make_prompts(a.prompts, shape.vocab_size, 384, 4, 32, seed=1)
    args = bench.make_args(d, resident_weights="always", pinned_limit_bytes=0)
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    tok = planner.tokenize_prompts(model.tokenizer, prompts)
    packed = planner.pack_prompts(tok, model.max_rows, model.max_head_rows, False)
    model.run(None, packed=packed, return_device=True)
    for setting in a.set or [""]:
        opts = dict(kv.split("=") for kv in setting.split(",") if kv)
        for k, v in opts.items():
            model.ctx.call("fls_set_option", k.encode(), int(v))
        for _ in range(2):
            model.run(None, packed=packed, return_device=True)
        torch.cuda.synchronize()
        model.ctx.call("fls_reset_stats")
        model.ctx.call("fls_set_profiling", 1)
        for _ in range(a.passes):
            model.run(None, packed=packed, return_device=True)
        torch.cuda.synchronize()
        prof = model.ctx.profile()
        model.ctx.call("fls_set_profiling", 0)
        row = {}
        for c, v in prof.items():
            if v["launches"] and v["ms"] > 0:
                row[c] = round(v["work"] / (v["ms"] * 1e-3) / (1e12 if c.startswith("gemm") else 1e9), 1) if v["work"] else \
                    round(v["ms"] / v["launches"] * 1e3, 1)
        row["ms_per_pass"] = round(sum(v["ms"] for v in prof.values()) / a.passes, 2)
        print(json.dumps({"model": a.model, "options": opts, **row}), flush=True)
    model.close()
finally:
    shutil.rmtree(d, ignore_errors=True)
