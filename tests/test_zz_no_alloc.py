"""GPU test of fls_op_head_noalloc, the softmax-head entry point for callers that cannot allocate
device memory (a plugin inside a process whose pool owns all of HBM, or a thread that cannot afford
cudaMalloc's device-wide synchronisation).  Its scratch is a fixed array inside the library, shared
by every call on the device, so the test makes two calls in a row with different sizes (the second
smaller, so stale partials from the first are still lying in the array) and checks each against the
context-based fls_op_head on the same inputs, within the tolerances tests/test_ops_gpu.py uses for the
head.  The file name sorts last so a `-x` run has executed every other test before this one; it
also passes alone in a fresh process (the entry point does its own per-device set-up)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, std, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (std * torch.randn(shape, generator=g)).to(torch.float16).to(dev)


def _run_both(lib, ctx, rows, V, K, seed, dev):
    x, w = _rand((rows, K), 1.0, seed, dev), _rand((V, K), 0.2, seed + 1, dev)
    ids = torch.randint(0, V, (rows, 2), dtype=torch.int32, device=dev)
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for which in ("ctx", "noalloc"):
        probs = torch.full((rows, V), float("nan"), dtype=torch.float16, device=dev)
        lp = torch.full((rows, 2), float("nan"), dtype=torch.float32, device=dev)
        lse = torch.full((rows,), float("nan"), dtype=torch.float32, device=dev)
        if which == "ctx":
            ctx.call("fls_op_head", x.data_ptr(), w.data_ptr(), rows, V, K, probs.data_ptr(), ids.data_ptr(), 2,
                     lp.data_ptr(), lse.data_ptr(), sp)
        else:
            rc = lib.fls_op_head_noalloc(x.data_ptr(), w.data_ptr(), rows, V, K, probs.data_ptr(), ids.data_ptr(), 2,
                                         lp.data_ptr(), lse.data_ptr(), sp)
            assert rc == 0, lib.fls_last_error(None)
        torch.cuda.synchronize()
        outs.append((probs, lp, lse))
    return outs


def test_noalloc_head_matches_context_head_twice():
    from flexible_llm_sharding_b200 import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    ctx = _lib.Context(_lib.ModelDesc(512, 1408, 4, 2, 128, 32000, 2, 1e-5), 0)
    ctx.call("fls_configure", 1, C.c_size_t(256), 256, 1024)
    try:
        # call 1 fills most of the scratch; call 2 is smaller and must not see call 1's leftovers
        for rows, V, K, seed in [(700, 32000, 256, 21), (37, 1000, 512, 23)]:
            (p0, lp0, lse0), (p1, lp1, lse1) = _run_both(lib, ctx, rows, V, K, seed, dev)
            assert torch.isfinite(p1).all() and torch.isfinite(lp1).all() and torch.isfinite(lse1).all()
            normal = p0.float() > 1e-4
            rel = ((p1.float() - p0.float()).abs() / p0.float().clamp_min(1e-7))[normal].max().item()
            assert rel <= 2e-2, rel                                            # tolerance of test_ops_gpu head test
            assert (p1.float() - p0.float()).abs()[~normal].max().item() <= 2e-6
            assert (lse1 - lse0).abs().max().item() <= 2e-3
            assert (lp1 - lp0).abs().max().item() <= 2e-2
    finally:
        ctx.close()


def test_noalloc_head_rejects_sizes_beyond_the_builtin_scratch():
    from flexible_llm_sharding_b200 import _lib
    torch.cuda.set_device(0)
    lib = _lib.load()
    x = torch.zeros((8, 64), dtype=torch.float16, device="cuda:0")
    for rows, V in [(1025, 1000), (8, 32776)]:
        rc = lib.fls_op_head_noalloc(x.data_ptr(), x.data_ptr(), rows, V, 64, None, None, 0, None, None, None)
        assert rc == -5                                                       # FLS_ERR_CAPACITY
        msg = lib.fls_last_error(None).decode()
        assert "exceeds the built-in scratch" in msg and "1024" in msg and "32768" in msg
