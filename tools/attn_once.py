"""Two launches of each tcgen05 attention form (persistent CTAs; grid form with 3 and with 2 K buffers) on the 7B (or, with
`70b`, the 70B) micro-batch shape - the target of `ncu` captures:
    ncu --set full --import-source on -k regex:attention_ python tools/attn_once.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import _lib, planner

nq, nkv, d = (64, 8, 128) if "70b" in sys.argv else (32, 32, 128)
ctx = _lib.Context(_lib.ModelDesc(nq * d, 11008, nq, nkv, d, 32000, 2, 1e-5), 0)
ctx.call("fls_configure", 1, C.c_size_t(256), 256, 256)
tiles, base = [], 0
for _ in range(32):
    tiles += planner.prompt_tiles(base, 384, 4, 32, False)
    base += 512
ld = (nq + 2 * nkv) * d
qkv = (torch.randn((base, ld), device="cuda") * 1.28).half()
t = torch.tensor(tiles, dtype=torch.int32, device="cuda")
out = torch.empty((base, nq * d), dtype=torch.float16, device="cuda")
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for impl, ks in ((2, 0), (1, 3), (1, 2)):           # persistent CTAs; one CTA per (tile, head) with 3 / 2 K buffers
    ctx.call("fls_set_option", b"attention_impl", impl)
    ctx.call("fls_set_option", b"attention_k_stages", ks)
    for _ in range(2):
        ctx.call("fls_op_attention", qkv.data_ptr(), ld, base, nq, nkv, d, t.data_ptr(), len(tiles), out.data_ptr(), nq * d, sp)
    torch.cuda.synchronize()
    print("launched impl", impl, "with", ks, "K buffers", flush=True)
