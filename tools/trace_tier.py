"""Timeline of one pass with the host activation tier (torch.profiler / CUPTI): where does the compute stream idle?

    python tools/trace_tier.py [--workload 70b --prompts 32 --k 10]

Prints the largest gaps between consecutive kernels and which memcpys overlap each gap."""
import argparse
import copy
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

import bench
from flexible_llm_sharding_b200 import synthetic
from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="70b")
ap.add_argument("--prompts", type=int, default=32)
ap.add_argument("--k", type=int, default=10)
ap.add_argument("--storage", default="cpu")
a = ap.parse_args()

import __graft_entry__ as ge
ge.build()
shape = synthetic.LlamaShape.named(a.workload)
alias = 4 if a.workload == "70b" else 0
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
md = bench.model_dir_for(a.workload, 0, 1, dev, alias)
prompts = synthetic.make_prompts(a.prompts, shape.vocab_size, 1792, 4, 64, seed=1234)
args = bench.make_args(md, layer_num_per_shard=a.k, storage_location=a.storage, force_storage_tier=True,
                       disk_folder=tempfile.mkdtemp(), max_rows=16384)
model = ShardedLlama(args, DeviceManager(args, [str(dev)]), device=str(dev))
model(copy.deepcopy(prompts))
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    model(copy.deepcopy(prompts))
    torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy")]
kern = sorted([e for e in ev if e["cat"] == "kernel"], key=lambda e: e["ts"])
cpy = sorted([e for e in ev if e["cat"] == "gpu_memcpy"], key=lambda e: e["ts"])
t0 = kern[0]["ts"]
print("kernels", len(kern), "memcpys", len(cpy), "span_s", (kern[-1]["ts"] + kern[-1]["dur"] - t0) / 1e6,
      "kernel_busy_s", sum(e["dur"] for e in kern) / 1e6)
gaps = []
for x, y in zip(kern, kern[1:]):
    g = y["ts"] - (x["ts"] + x["dur"])
    if g > 2000:
        gaps.append((g, x, y))
print("gaps > 2 ms:", len(gaps), "total_s", sum(g for g, _, _ in gaps) / 1e6)
for g, x, y in sorted(gaps, key=lambda t: -t[0])[:12]:
    lo, hi = x["ts"] + x["dur"], y["ts"]
    over = [c for c in cpy if c["ts"] < hi and c["ts"] + c["dur"] > lo]
    desc = [(c["name"][:22], round(c["dur"] / 1e3, 1), c.get("args", {}).get("stream"), round(c.get("args", {}).get("bytes", 0) / 1e6))
            for c in over[:6]]
    print(f"gap {g / 1e3:8.1f} ms at {(lo - t0) / 1e6:7.3f}s  after {x['name'][:40]} before {y['name'][:40]}  copies in gap: {len(over)} {desc}")
by = {}
for c in cpy:
    k = (c["name"][:24], c.get("args", {}).get("stream"))
    b = by.setdefault(k, [0, 0.0, 0])
    b[0] += 1; b[1] += c["dur"] / 1e3; b[2] += c.get("args", {}).get("bytes", 0)
for k, (n, ms, byts) in by.items():
    print("memcpy", k, "n", n, "ms", round(ms), "GB", round(byts / 1e9, 1), "GB/s", round(byts / max(ms, 1e-9) / 1e6, 1))
