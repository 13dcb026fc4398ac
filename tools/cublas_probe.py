"""One torch.matmul (cuBLAS) launch per decoder GEMM shape of a model, for `ncu --set full`: what traffic, tile
shape and cluster shape does the library use on the shapes where the hand-written kernel is compared with it?
    ncu --set full -k regex:gemm\|cutlass\|nvjet -o out python tools/cublas_probe.py 70b"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flexible_llm_sharding_b200 import synthetic

s = synthetic.LlamaShape.named(sys.argv[1] if len(sys.argv) > 1 else "70b")
H, I = s.hidden_size, s.intermediate_size
qkv = (s.num_attention_heads + 2 * s.num_key_value_heads) * s.head_dim
M = 16384
for N, K in [(qkv, H), (H, s.num_attention_heads * s.head_dim), (2 * I, H), (H, I)]:
    A = (torch.randn((M, K), device="cuda") * 0.5).half()
    B = (torch.randn((N, K), device="cuda") * 0.02).half()
    out = torch.empty((M, N), dtype=torch.float16, device="cuda")
    for _ in range(2):
        torch.matmul(A, B.t(), out=out)
    torch.cuda.synchronize()
    del A, B, out
