"""Sustained-throughput probe of the tcgen05 GEMM on the 7B decoder shapes (M = 16384 rows), with cuBLAS
(torch.matmul, fp16) timed in the same process under the same power cap, and a sweep of the raster
group size.  Each measurement loops one launch back to back for ~1.5 s (CUDA events).
Run on a GPU box: python tools/gemm_probe.py [--secs 1.5]"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import _lib

ap = argparse.ArgumentParser()
ap.add_argument("--secs", type=float, default=1.5)
ap.add_argument("--groups", type=int, nargs="+", default=[16])
ap.add_argument("--hints", type=int, nargs="*", default=[])
ap.add_argument("--hint_groups", type=int, nargs="*", default=[16])
a = ap.parse_args()

ctx = _lib.Context(_lib.ModelDesc(4096, 11008, 32, 32, 128, 32000, 2, 1e-5), 0)
ctx.call("fls_configure", 1, C.c_size_t(256), 256, 256)
M = 16384
dev = "cuda"
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
SHAPES = [("qkv", 0, 12288, 4096), ("o+res", 1, 4096, 4096), ("gate_up+swiglu", 2, 11008, 4096),
          ("down+res", 1, 4096, 11008)]


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


for name, epi, N, K in SHAPES:
    A = (torch.randn((M, K), device=dev) * 0.5).half()
    brows = 2 * N if epi == 2 else N
    B = (torch.randn((brows, K), device=dev) * 0.02).half()
    Cm = torch.zeros((M, N), dtype=torch.float16, device=dev)
    flops = 2.0 * M * brows * K
    res = {}

    def ours():
        ctx.call("fls_op_gemm", epi, A.data_ptr(), K, B.data_ptr(), K, brows, M, N, K, Cm.data_ptr(), N,
                 Cm.data_ptr() if epi == 1 else None, N, N if epi == 2 else 0, sp)

    for cg in (2, 1):
        ctx.call("fls_set_option", b"gemm_cta_group", cg)
        for g in (a.groups if cg == 2 else [16]):
            ctx.call("fls_set_option", b"gemm_group_m", g)
            res[f"cg{cg}/group{g}"] = round(sustained(ours, flops))
    ctx.call("fls_set_option", b"gemm_cta_group", 2)
    ctx.call("fls_set_option", b"gemm_group_m", 0)
    for hint in a.hints:
        ctx.call("fls_set_option", b"gemm_cache_hints", hint)
        for g in a.hint_groups:
            ctx.call("fls_set_option", b"gemm_group_m", g)
            res[f"cg2/group{g}/hint{hint}"] = round(sustained(ours, flops))
    ctx.call("fls_set_option", b"gemm_cache_hints", 0)
    ctx.call("fls_set_option", b"gemm_group_m", 0)
    if epi == 1:      # the same shape with the plain-store epilogue: what does the fused residual cost?
        def plain():
            ctx.call("fls_op_gemm", 0, A.data_ptr(), K, B.data_ptr(), K, brows, M, N, K, Cm.data_ptr(), N, None, 0, 0, sp)
        res["cg2/group16/plain-store"] = round(sustained(plain, flops))
    Bt = B.t()
    out = torch.empty((M, brows), dtype=torch.float16, device=dev)
    res["cublas(torch.matmul, plain store)"] = round(sustained(lambda: torch.matmul(A, Bt, out=out), flops))
    print(name, f"M={M} N={brows} K={K}", res, flush=True)
