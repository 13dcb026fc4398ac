"""Bandwidth probe of the RMSNorm kernel's CTA shapes (fls_set_option "rmsnorm_variant") at decoder sizes.
Each measurement loops the launch for ~0.5 s over a rotating set of buffers larger than L2 (CUDA events);
algorithmic bytes per launch = rows * H * 4 (one fp16 read + one fp16 write per element).
Run on a GPU box: python tools/norm_probe.py [--secs 0.5]"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import _lib

ap = argparse.ArgumentParser()
ap.add_argument("--secs", type=float, default=0.5)
ap.add_argument("--rows", type=int, default=16384)
a = ap.parse_args()

ctx = _lib.Context(_lib.ModelDesc(4096, 11008, 32, 32, 128, 32000, 2, 1e-5), 0)
ctx.call("fls_configure", 1, C.c_size_t(256), 256, 256)
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
out = {}
for H in (4096, 5120, 8192):
    n_buf = max(2, int(600e6 // (a.rows * H * 2)) + 1)          # inputs + outputs together well above the 126 MB L2
    xs = [(torch.randn((a.rows, H), device="cuda") * 0.5).half() for _ in range(n_buf)]
    ys = [torch.empty_like(x) for x in xs]
    w = (1 + 0.1 * torch.randn(H, device="cuda")).half()
    ref = None
    for variant in (0, 1, 2):
        ctx.call("fls_set_option", b"rmsnorm_variant", variant)

        def launch(i):
            ctx.call("fls_op_rmsnorm", xs[i % n_buf].data_ptr(), None, None, w.data_ptr(), C.c_float(1e-5), a.rows, H,
                     None, ys[i % n_buf].data_ptr(), sp)

        for i in range(n_buf):
            launch(i)
        torch.cuda.synchronize()
        if ref is None:
            ref = ys[0].clone()
        # the fp32 sum of squares is split differently per CTA shape: results may differ by one fp16 ulp
        assert torch.allclose(ys[0].float(), ref.float(), rtol=2e-3, atol=1e-4)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); launch(0); e1.record(); e1.synchronize()
        n = max(n_buf, int(a.secs * 1e3 / max(e0.elapsed_time(e1), 1e-3)))
        e0.record()
        for i in range(n):
            launch(i)
        e1.record(); e1.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / n
        out[f"H{H}/variant{variant}"] = {"us": round(us, 1), "GBps": round(a.rows * H * 4 / us / 1e3)}
    ctx.call("fls_set_option", b"rmsnorm_variant", 0)
print(json.dumps(out, indent=1))
