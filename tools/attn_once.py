"""One launch of the tcgen05 attention per K-buffer count (3, then 2) on the 7B micro-batch shape - the target of
`ncu` captures:  ncu --set full --import-source on -k regex:attention_tcgen05 python tools/attn_once.py"""
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
for ks in (3, 2):
    ctx.call("fls_set_option", b"attention_k_stages", ks)
    for _ in range(2):
        ctx.call("fls_op_attention", qkv.data_ptr(), ld, base, nq, nkv, d, t.data_ptr(), len(tiles), out.data_ptr(), nq * d, sp)
    torch.cuda.synchronize()
    print("launched with", ks, "K buffers", flush=True)
