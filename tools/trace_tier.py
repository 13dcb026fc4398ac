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
ap.add_argument("--own_stream", type=int, default=0, help="1: run the pass on a non-default torch stream")
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
side = torch.cuda.Stream(dev) if a.own_stream else torch.cuda.current_stream(dev)
with torch.cuda.stream(side):
    model(copy.deepcopy(prompts))
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        model(copy.deepcopy(prompts))
        torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
allev = [e for e in json.load(open(path))["traceEvents"] if e.get("ph") == "X"]
ev = [e for e in allev if e.get("cat") in ("kernel", "gpu_memcpy")]
rt = sorted([e for e in allev if e.get("cat") in ("cuda_runtime", "cuda_driver")], key=lambda e: e["ts"])
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

# host side: which runtime calls blocked, and when the weight chunks of each shard were ENQUEUED vs EXECUTED
print("host runtime calls > 5 ms:")
for e in sorted(rt, key=lambda e: -e["dur"])[:12]:
    if e["dur"] > 5000:
        print(f"  {e['name'][:40]:40s} {e['dur'] / 1e3:8.1f} ms at {(e['ts'] - t0) / 1e6:7.3f}s")
wstream = max(by, key=lambda k: by[k][2])[1]
w = [c for c in cpy if c.get("args", {}).get("stream") == wstream and c["name"].startswith("Memcpy HtoD")]
corr = {e.get("args", {}).get("correlation"): e for e in rt}
groups, cur = [], []
for c in w:
    if cur and c["ts"] - (cur[-1]["ts"] + cur[-1]["dur"]) > 20000:
        groups.append(cur); cur = []
    cur.append(c)
if cur:
    groups.append(cur)
print("weight-copy bursts (device time) and when their first/last chunk was enqueued on the host:")
for g in groups:
    f, l = g[0], g[-1]
    ef, el = corr.get(f.get("args", {}).get("correlation")), corr.get(l.get("args", {}).get("correlation"))
    print(f"  n {len(g):4d} dev {(f['ts'] - t0) / 1e6:7.3f}..{(l['ts'] + l['dur'] - t0) / 1e6:7.3f}s  "
          f"host enqueue {((ef['ts'] - t0) / 1e6) if ef else -1:7.3f}..{((el['ts'] - t0) / 1e6) if el else -1:7.3f}s")
