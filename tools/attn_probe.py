"""Micro-benchmark of the attention kernels on the 7B C2 shape (32 prompts x (384 + 4x32) rows, 32 heads)
(tcgen05 kernel vs mma.sync kernel).  Run on a GPU box: python tools/attn_probe.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import _lib, planner

nq, nkv, d = 32, 32, 128
ctx = _lib.Context(_lib.ModelDesc(4096, 11008, nq, nkv, d, 32000, 2, 1e-5), 0)
ctx.call("fls_configure", 1, C.c_size_t(256), 256, 256)
tiles, base = [], 0
for _ in range(32):
    tiles += planner.prompt_tiles(base, 384, 4, 32, False)
    base += 512
rows = base
ld = (nq + 2 * nkv) * d
qkv = (torch.randn((rows, ld), device="cuda") * 1.28).half()
t = torch.tensor(tiles, dtype=torch.int32, device="cuda")
out = torch.empty((rows, nq * d), dtype=torch.float16, device="cuda")
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def run(n=5):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        ctx.call("fls_op_attention", qkv.data_ptr(), ld, rows, nq, nkv, d, t.data_ptr(), len(tiles), out.data_ptr(),
                 nq * d, sp)
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / n * 1e3


flops = 32 * nq * (4 * 384 * 384 * d + 4 * 128 * (384 + 32) * d)
for impl in (1, 0):
    ctx.call("fls_set_option", b"attention_impl", impl)
    run(3)
    us = run(20)
    print(f"impl {impl}: {us:.1f} us per launch, {flops / us / 1e6:.0f} TFLOP/s useful")
