"""GPU test of fls_op_head_noalloc, the softmax-head entry point for callers that cannot allocate
device memory (a plugin inside a process whose pool owns all of HBM, or a thread that cannot afford
cudaMalloc's device-wide synchronisation).  Its scratch is a fixed array inside the library, shared
by every call on the device, so the test makes two calls in a row with different sizes (the second
smaller, so stale partials from the first are still lying in the array) and checks each against the
context-based fls_op_head on the same inputs (same kernels: bit-identical) and against a plain torch fp32
softmax / log-softmax of the same fp16 logits, within the tolerances tests/test_ops_gpu.py uses for the
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
    logits = (x.float() @ w.float().t()).half().float()          # lm_head output as the reference holds it (fp16)
    want = (torch.softmax(logits, dim=-1), torch.log_softmax(logits, dim=-1).gather(1, ids.long()),
            torch.logsumexp(logits, dim=-1))
    return outs + [want]


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
            (p0, lp0, lse0), (p1, lp1, lse1), (wp, wlp, wlse) = _run_both(lib, ctx, rows, V, K, seed, dev)
            assert torch.isfinite(p1).all() and torch.isfinite(lp1).all() and torch.isfinite(lse1).all()
            # the context-based entry point runs the same kernels on the same inputs
            assert torch.equal(p1, p0) and torch.equal(lp1, lp0) and torch.equal(lse1, lse0)
            # and both agree with torch's fp32 softmax of the fp16 logits (tolerances of the test_ops_gpu head test)
            normal = wp > 1e-4
            rel = ((p1.float() - wp).abs() / wp.clamp_min(1e-7))[normal].max().item()
            assert rel <= 2e-2, rel
            assert (p1.float() - wp).abs()[~normal].max().item() <= 2e-6
            assert (lse1 - wlse).abs().max().item() <= 2e-3
            assert (lp1 - wlp).abs().max().item() <= 2e-2
    finally:
        ctx.close()


def test_noalloc_head_flags_answer_ids_outside_the_vocabulary():
    """The ids are device data the host cannot check: an id outside [0, n) must come back as NaN, not as a
    read outside the weight matrix."""
    from flexible_llm_sharding_b200 import _lib
    torch.cuda.set_device(0)
    lib = _lib.load()
    x, w = _rand((4, 64), 1.0, 31, "cuda:0"), _rand((1000, 64), 0.2, 32, "cuda:0")
    ids = torch.tensor([[3, 999], [1000, 7], [-1, 0], [2 ** 30, 5]], dtype=torch.int32, device="cuda:0")
    lp = torch.zeros((4, 2), dtype=torch.float32, device="cuda:0")
    rc = lib.fls_op_head_noalloc(x.data_ptr(), w.data_ptr(), 4, 1000, 64, None, ids.data_ptr(), 2, lp.data_ptr(), None,
                                 C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.fls_last_error(None)
    torch.cuda.synchronize()
    bad = torch.tensor([[False, False], [True, False], [True, False], [True, False]], device="cuda:0")
    assert torch.isnan(lp[bad]).all() and torch.isfinite(lp[~bad]).all()
    # misaligned / odd sizes are refused up front
    rc = lib.fls_op_head_noalloc(x.data_ptr(), w.data_ptr(), 4, 1000, 60, None, None, 0, None, None, None)
    assert rc == -1 and "multiples of 8" in lib.fls_last_error(None).decode()


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
