"""Sustained-throughput probe of the tcgen05 GEMM on the decoder shapes of a model (M = 16384 rows), with cuBLAS
(torch.matmul, fp16, plain store) timed in the same process under the same power cap.  Each measurement loops one
launch back to back for ~`--secs` seconds (CUDA events).

    python tools/gemm_probe.py [--models 7b 70b] [--secs 1.5] [--groups 16 8] [--variants tma direct cg1]
    python tools/gemm_probe.py --ncu 7b        # one launch per class in the order qkv, o, gate/up, down (for
                                               # `ncu --set full`; bench.py reads the resulting CSV in that order)
"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import _lib, synthetic

ap = argparse.ArgumentParser()
ap.add_argument("--models", nargs="+", default=["7b", "70b"])
ap.add_argument("--secs", type=float, default=1.5)
ap.add_argument("--groups", type=int, nargs="+", default=[16])
ap.add_argument("--variants", nargs="+", default=["wide", "narrow"],
                help="auto (what a pass uses: wide where K >= 8192) | wide (256x512 tiles, skewed streams) | lockstep (256x512, both accumulators in lock step) | narrow (256x256, TMA store) | direct (256x256, per-thread stores) | cg1")
ap.add_argument("--rows", type=int, default=16384)
ap.add_argument("--rasters", type=int, nargs="+", default=[0])
ap.add_argument("--hints", type=int, nargs="+", default=[0], help="gemm_cache_hints values to sweep")
ap.add_argument("--only", nargs="*", default=[], help="restrict to these classes (gemm_qkv gemm_o gemm_gate_up gemm_down)")
ap.add_argument("--ncu", default="")
ap.add_argument("--out", default="")
a = ap.parse_args()

ctx = _lib.Context(_lib.ModelDesc(4096, 11008, 32, 32, 128, 32000, 2, 1e-5), 0)
ctx.call("fls_configure", 1, C.c_size_t(256), 256, 256)
M = a.rows
dev = "cuda"
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def shapes_of(model):
    s = synthetic.LlamaShape.named(model)
    H, I = s.hidden_size, s.intermediate_size
    qkv = (s.num_attention_heads + 2 * s.num_key_value_heads) * s.head_dim
    # (class, epilogue, N, K): store (the fused-RoPE epilogue needs positions; its cost is the same stores), +residual,
    # SwiGLU, +residual
    return [("gemm_qkv", 0, qkv, H), ("gemm_o", 1, H, s.num_attention_heads * s.head_dim),
            ("gemm_gate_up", 2, I, H), ("gemm_down", 1, H, I)]


def sustained(fn, flops):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); e1.synchronize()
    n = max(3, int(a.secs * 1e3 / max(e0.elapsed_time(e1), 1e-3)))
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); e1.synchronize()
    return flops * n / (e0.elapsed_time(e1) * 1e-3) / 1e12


def operands(epi, N, K):
    A = (torch.randn((M, K), device=dev) * 0.5).half()
    brows = 2 * N if epi == 2 else N
    B = (torch.randn((brows, K), device=dev) * 0.02).half()
    Cm = torch.zeros((M, N), dtype=torch.float16, device=dev)
    return A, B, Cm, brows


def launch(epi, A, B, Cm, brows, N, K):
    ctx.call("fls_op_gemm", epi, A.data_ptr(), K, B.data_ptr(), K, brows, M, N, K, Cm.data_ptr(), N,
             Cm.data_ptr() if epi == 1 else None, N, N if epi == 2 else 0, sp)


if a.ncu:
    variant = a.variants[0]
    ctx.call("fls_set_option", b"gemm_cta_group", 1 if variant == "cg1" else 2)
    ctx.call("fls_set_option", b"gemm_tma_store", 0 if variant == "direct" else 1)
    ctx.call("fls_set_option", b"gemm_wide_tiles", {"wide": 1, "lockstep": 1, "auto": 2}.get(variant, 0))
    ctx.call("fls_set_option", b"gemm_wide_lockstep", 1 if variant == "lockstep" else 0)
    for name, epi, N, K in shapes_of(a.ncu):
        if a.only and name not in a.only:
            continue
        A, B, Cm, brows = operands(epi, N, K)
        for r in a.rasters:
            for g in a.groups:
                ctx.call("fls_set_option", b"gemm_raster", r)
                ctx.call("fls_set_option", b"gemm_group_m", g)
                launch(epi, A, B, Cm, brows, N, K)
                torch.cuda.synchronize()
                print("launched", name, variant, "raster", r, "group", g, flush=True)
        del A, B, Cm
    sys.exit(0)

results = {}
for model in a.models:
    for name, epi, N, K in shapes_of(model):
        if a.only and name not in a.only:
            continue
        A, B, Cm, brows = operands(epi, N, K)
        flops = 2.0 * M * brows * K
        res = {}
        for variant in a.variants:
            ctx.call("fls_set_option", b"gemm_cta_group", 1 if variant == "cg1" else 2)
            ctx.call("fls_set_option", b"gemm_tma_store", 0 if variant == "direct" else 1)
            ctx.call("fls_set_option", b"gemm_wide_tiles", {"wide": 1, "lockstep": 1, "auto": 2}.get(variant, 0))
            ctx.call("fls_set_option", b"gemm_wide_lockstep", 1 if variant == "lockstep" else 0)
            for r in a.rasters:
                ctx.call("fls_set_option", b"gemm_raster", r)
                for g in a.groups:
                    ctx.call("fls_set_option", b"gemm_group_m", g)
                    for h in a.hints:
                        ctx.call("fls_set_option", b"gemm_cache_hints", h)
                        res[f"{variant}/r{r}/group{g}" + (f"/hint{h}" if h else "")] = round(
                            sustained(lambda: launch(epi, A, B, Cm, brows, N, K), flops))
                    ctx.call("fls_set_option", b"gemm_cache_hints", 0)
            ctx.call("fls_set_option", b"gemm_raster", 0)
        ctx.call("fls_set_option", b"gemm_cta_group", 2)
        ctx.call("fls_set_option", b"gemm_tma_store", 1)
        ctx.call("fls_set_option", b"gemm_wide_tiles", 2)
        ctx.call("fls_set_option", b"gemm_group_m", 0)
        Bt = B.t()
        out = torch.empty((M, brows), dtype=torch.float16, device=dev)
        res["cublas(torch.matmul, plain store)"] = round(sustained(lambda: torch.matmul(A, Bt, out=out), flops))
        best = max(v for k, v in res.items() if not k.startswith("cublas"))
        res["best/cublas"] = round(best / res["cublas(torch.matmul, plain store)"], 3)
        results[f"{model}/{name} M={M} N={brows} K={K}"] = res
        print(f"{model}/{name}", f"M={M} N={brows} K={K}", res, flush=True)
        del A, B, Cm, Bt, out
if a.out:
    with open(a.out, "w") as f:
        json.dump(results, f, indent=1)
