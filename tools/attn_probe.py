"""Micro-benchmark of the attention kernels alone on one 16384-row micro-batch of the C2 prompt shape
(32 prompts x (384 + 4x32) rows): the tcgen05 kernel with 3 and with 2 K buffers per CTA, and the mma.sync kernel, at the
7B (32/32 heads) and 70B (64/8 heads) head counts.  Also checks that the variants agree bit for bit.

    python tools/attn_probe.py [--out gpurun_out/attn_probe.json]
"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import _lib, planner

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="")
ap.add_argument("--reps", type=int, default=50)

ap.add_argument("--long", action="store_true", help="8 prompts x (1792 + 4 x 64) rows instead of 32 x (384 + 4 x 32): "
                "29-30 key tiles per CTA instead of 6-8, so the per-CTA prologue and epilogue weigh a quarter as much")
a = ap.parse_args()
d = 128
n_prompts, len_p, n_opt, len_s = (8, 1792, 4, 64) if a.long else (32, 384, 4, 32)
tiles, base = [], 0
for _ in range(n_prompts):
    tiles += planner.prompt_tiles(base, len_p, n_opt, len_s, False)
    base += len_p + n_opt * len_s
rows = base
res = {}
for name, nq, nkv in (("7b", 32, 32), ("70b", 64, 8)):
    ctx = _lib.Context(_lib.ModelDesc(nq * d, 11008, nq, nkv, d, 32000, 2, 1e-5), 0)
    ctx.call("fls_configure", 1, C.c_size_t(256), 256, 256)
    ld = (nq + 2 * nkv) * d
    torch.manual_seed(0)
    qkv = (torch.randn((rows, ld), device="cuda") * 1.28).half()
    t = torch.tensor(tiles, dtype=torch.int32, device="cuda")
    out = torch.empty((rows, nq * d), dtype=torch.float16, device="cuda")
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run(n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            ctx.call("fls_op_attention", qkv.data_ptr(), ld, rows, nq, nkv, d, t.data_ptr(), len(tiles), out.data_ptr(),
                     nq * d, sp)
        e1.record()
        e1.synchronize()
        return e0.elapsed_time(e1) / n * 1e3

    flops = n_prompts * nq * (4 * len_p * len_p * d + 4 * n_opt * len_s * (len_p + len_s) * d)
    r, outs, best = {}, {}, {}
    variants = (("tcgen05, persistent CTAs", 2, 0), ("tcgen05, 3 K buffers", 1, 3), ("tcgen05, 2 K buffers", 1, 2),
                ("mma.sync", 0, 0))
    for rnd in range(3):                        # the variants take turns, so that drift hits them alike
        for label, impl, ks in variants:
            ctx.call("fls_set_option", b"attention_impl", impl)
            ctx.call("fls_set_option", b"attention_k_stages", ks)
            out.fill_(float("nan"))
            run(5)
            best[label] = min(best.get(label, 1e9), run(a.reps))
            outs[label] = out.clone()
    for label, _, _ in variants:
        us = best[label]
        r[label] = dict(us_per_launch=round(us, 1), useful_tflops=round(flops / us / 1e6))
        print(name, label, r[label], flush=True)
    r["3 and 2 K buffers bit-identical"] = bool(torch.equal(outs["tcgen05, 3 K buffers"], outs["tcgen05, 2 K buffers"]))
    r["persistent and grid forms bit-identical"] = bool(torch.equal(outs["tcgen05, persistent CTAs"], outs["tcgen05, 3 K buffers"]))
    r["max |tcgen05 - mma.sync|"] = float((outs["tcgen05, 3 K buffers"].float() - outs["mma.sync"].float()).abs().max())
    print(name, {k: v for k, v in r.items() if not isinstance(v, dict)}, flush=True)
    res[f"{name}: {nq} heads / {nkv} kv heads, {rows} rows, {len(tiles)} tiles"] = r
    ctx.close()
if a.out:
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
