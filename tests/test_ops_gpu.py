"""Kernel-level parity: each CUDA kernel, called through the C ABI (ctypes), against the oracle's
restatement of the same op (oracle/llama_oracle.py) or plain torch math on the same seeded inputs.
Tolerances are stated next to each check."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from flexible_llm_sharding_b200 import _lib, planner
    dev = torch.device("cuda:0")
    desc = _lib.ModelDesc(512, 1408, 4, 2, 128, 1000, 2, 1e-5)
    ctx = _lib.Context(desc, 0)
    ctx.call("fls_configure", 2, C.c_size_t(1 << 20), 4096, 1024)
    cos, sin = planner.rope_tables(4096, 128)
    ctx.call("fls_set_rope_table", cos.data_ptr(), sin.data_ptr(), 4096)
    yield dict(ctx=ctx, dev=dev, lib=_lib, cos=cos, sin=sin, planner=planner)
    ctx.close()


def _sp():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _rand(shape, std, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (std * torch.randn(shape, generator=g)).to(torch.float16).to(dev)


# M, N, K: full tiles, M tail, N tail (vocab-like 1000), tiny K, K tail (not a multiple of 64), multi-wave
GEMM_SHAPES = [(128, 256, 64), (256, 512, 512), (200, 256, 128), (130, 1000, 512), (64, 8, 8), (300, 264, 200),
               (4096, 2304, 512), (1000, 512, 1408), (257, 384, 72), (5000, 4096, 256)]


@pytest.fixture(params=[(2, 1, 1), (2, 1, 3), (2, 1, 0), (1, 1, 0), (2, 0, 0), (1, 0, 0)],
                ids=["cta_group2-wide", "cta_group2-wide-lockstep", "cta_group2", "cta_group1", "cta_group2-direct_store",
                     "cta_group1-direct_store"])
def cta_group(request, env):
    """GEMM variants: CTA pairs (256 x 512 or 256 x 256 tiles) or single CTAs; epilogue through smem + TMA store
    (default) or per-thread stores."""
    cg, tma_store, wide = request.param
    env["ctx"].call("fls_set_option", b"gemm_cta_group", cg)
    env["ctx"].call("fls_set_option", b"gemm_tma_store", tma_store)
    env["ctx"].call("fls_set_option", b"gemm_wide_tiles", 1 if wide else 0)
    env["ctx"].call("fls_set_option", b"gemm_wide_lockstep", 1 if wide == 3 else 0)
    yield cg
    env["ctx"].call("fls_set_option", b"gemm_cta_group", 2)
    env["ctx"].call("fls_set_option", b"gemm_tma_store", 1)
    env["ctx"].call("fls_set_option", b"gemm_wide_tiles", 2)
    env["ctx"].call("fls_set_option", b"gemm_wide_lockstep", 0)


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_store(env, cta_group, M, N, K):
    dev, ctx = env["dev"], env["ctx"]
    a, b = _rand((M, K), 1.0, 1, dev), _rand((N, K), 0.05, 2, dev)
    c = torch.full((M, N), float("nan"), dtype=torch.float16, device=dev)
    ctx.call("fls_op_gemm", 0, a.data_ptr(), K, b.data_ptr(), K, N, M, N, K, c.data_ptr(), N, None, 0, 0, _sp())
    torch.cuda.synchronize()
    ref = (a.float() @ b.float().t())
    # fp32 accumulation on both sides: only summation order and the final fp16 rounding differ
    err = (c.float() - ref).abs().max().item()
    assert torch.isfinite(c).all()
    assert err <= 2e-3 * max(1.0, ref.abs().max().item()), err
    # tighter: at most 1 fp16 ulp off the correctly rounded fp32 result
    assert (c.float() - ref.half().float()).abs().max().item() <= 2 ** -10 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("M,N,K", [(256, 512, 512), (130, 520, 200)])
def test_gemm_residual_in_place(env, cta_group, M, N, K):
    dev, ctx = env["dev"], env["ctx"]
    a, b = _rand((M, K), 1.0, 3, dev), _rand((N, K), 0.05, 4, dev)
    x = _rand((M, N), 1.0, 5, dev)
    want = (x + (a.float() @ b.float().t()).half())          # fp16 add of two fp16 values (HF residual)
    ctx.call("fls_op_gemm", 1, a.data_ptr(), K, b.data_ptr(), K, N, M, N, K, x.data_ptr(), N, x.data_ptr(), N, 0, _sp())
    torch.cuda.synchronize()
    # identical except where the GEMM's fp32 sum rounds to a neighbouring fp16 value
    assert (x.float() - want.float()).abs().max().item() <= 4e-3
    assert ((x != want).float().mean().item()) < 0.05


@pytest.mark.parametrize("M,I,K", [(256, 1408, 512), (100, 136, 64)])
def test_gemm_swiglu_matches_unfused(env, cta_group, M, I, K):
    dev, ctx = env["dev"], env["ctx"]
    a = _rand((M, K), 1.0, 6, dev)
    wgu = _rand((2 * I, K), 0.05, 7, dev)            # gate rows then up rows, as laid out in a weight slot
    out = torch.empty((M, I), dtype=torch.float16, device=dev)
    ctx.call("fls_op_gemm", 2, a.data_ptr(), K, wgu.data_ptr(), K, 2 * I, M, I, K, out.data_ptr(), I, None, 0, I, _sp())
    g = (a.float() @ wgu[:I].float().t()).half()
    u = (a.float() @ wgu[I:].float().t()).half()
    want = torch.nn.functional.silu(g) * u           # HF LlamaMLP in fp16
    torch.cuda.synchronize()
    assert (out.float() - want.float()).abs().max().item() <= 3e-3 * max(1.0, want.abs().max().item())
    # the stand-alone SwiGLU kernel is the same arithmetic
    out2 = torch.empty_like(out)
    ctx.call("fls_op_swiglu", g.data_ptr(), u.data_ptr(), I, M, I, out2.data_ptr(), I, _sp())
    torch.cuda.synchronize()
    assert torch.equal(out2, want)


@pytest.mark.parametrize("H", [512, 4096, 5120, 8192])
def test_rmsnorm_bit_exact(env, H):
    from oracle import llama_oracle as O
    dev, ctx = env["dev"], env["ctx"]
    x = _rand((37, H), 1.5, 8, dev)
    w = (1 + 0.1 * torch.randn(H)).half().to(dev)
    out = torch.empty_like(x)
    ctx.call("fls_op_rmsnorm", x.data_ptr(), None, None, w.data_ptr(), C.c_float(1e-5), 37, H, None, out.data_ptr(), _sp())
    torch.cuda.synchronize()
    want = O.rmsnorm(x.cpu(), w.cpu(), 1e-5)
    diff = (out.cpu().float() - want.float()).abs()
    # same rounding points; the fp32 sum of squares is ordered differently -> at most 1 fp16 ulp, rarely
    assert diff.max().item() <= 2 ** -9 * 4 and (diff > 0).float().mean().item() < 0.01


def test_rmsnorm_residual_and_gather(env):
    from oracle import llama_oracle as O
    dev, ctx = env["dev"], env["ctx"]
    H = 512
    x, d = _rand((50, H), 1.0, 9, dev), _rand((50, H), 1.0, 10, dev)
    w = (1 + 0.1 * torch.randn(H)).half().to(dev)
    xs, out = torch.empty_like(x), torch.empty_like(x)
    ctx.call("fls_op_rmsnorm", x.data_ptr(), d.data_ptr(), None, w.data_ptr(), C.c_float(1e-5), 50, H, xs.data_ptr(),
             out.data_ptr(), _sp())
    torch.cuda.synchronize()
    assert torch.equal(xs, x + d)
    want = O.rmsnorm((x + d).cpu(), w.cpu(), 1e-5)
    assert (out.cpu().float() - want.float()).abs().max().item() <= 2 ** -7
    idx = torch.tensor([49, 0, 7, 7, 13], dtype=torch.int32, device=dev)
    out2 = torch.empty((5, H), dtype=torch.float16, device=dev)
    ctx.call("fls_op_rmsnorm", x.data_ptr(), None, idx.data_ptr(), w.data_ptr(), C.c_float(1e-5), 5, H, None,
             out2.data_ptr(), _sp())
    torch.cuda.synchronize()
    want2 = O.rmsnorm(x[idx.long()].cpu(), w.cpu(), 1e-5)
    assert (out2.cpu().float() - want2.float()).abs().max().item() <= 2 ** -7


def test_rope_bit_exact(env):
    from oracle import llama_oracle as O
    dev, ctx = env["dev"], env["ctx"]
    nq, nkv, d, rows = 4, 2, 128, 77
    ld = (nq + 2 * nkv) * d
    qkv = _rand((rows, ld), 1.0, 11, dev)
    orig = qkv.clone()
    pos = torch.randint(0, 4096, (rows,), dtype=torch.int32, device=dev)
    ctx.call("fls_op_rope", qkv.data_ptr(), ld, rows, nq, nkv, d, pos.data_ptr(), _sp())
    torch.cuda.synchronize()
    # the kernel's contract is "rotate with the table it was given": expected values use that table; that the table
    # is the reference's formula is checked separately below (and on CPU in test_host_logic.py)
    cos, sin = env["cos"][pos.cpu().long()], env["sin"][pos.cpu().long()]
    c2, s2 = O.rope_cos_sin(pos.cpu().long(), d, 10000.0, torch.float16)
    for a, b in ((cos, c2), (sin, s2)):      # fp32 cos/sin may differ by an ulp between two evaluations on some hosts
        assert (a.float() - b.float()).abs().max().item() <= 2 ** -10 and (a != b).float().mean().item() <= 1e-3
    x = orig.cpu()[:, :(nq + nkv) * d].view(rows, nq + nkv, d)
    want = (x * cos[:, None]) + (O.rotate_half(x) * sin[:, None])
    assert torch.equal(qkv.cpu()[:, :(nq + nkv) * d].view(rows, nq + nkv, d), want)     # bit exact
    assert torch.equal(qkv[:, (nq + nkv) * d:], orig[:, (nq + nkv) * d:])                # v untouched


def test_embed_exact(env):
    dev, ctx = env["dev"], env["ctx"]
    table = _rand((1000, 512), 1.0, 12, dev)
    ids = torch.randint(0, 1000, (333,), dtype=torch.int32, device=dev)
    out = torch.empty((333, 512), dtype=torch.float16, device=dev)
    ctx.call("fls_op_embed", table.data_ptr(), ids.data_ptr(), 333, 512, out.data_ptr(), _sp())
    torch.cuda.synchronize()
    assert torch.equal(out, table[ids.long()])


@pytest.mark.parametrize("nq,nkv,d,impl,k_stages", [(4, 4, 128, 1, 0), (4, 2, 128, 1, 3), (4, 2, 128, 1, 2),
                                                    (4, 4, 128, 2, 0), (4, 2, 128, 2, 0),     # persistent CTAs
                                                    (4, 2, 128, 0, 0), (8, 2, 64, 0, 0)])
@pytest.mark.parametrize("prefix_causal", [False, True])
def test_attention_matches_oracle(env, nq, nkv, d, impl, k_stages, prefix_causal):
    """Prefix + shared-prefix suffix attention for several prompts against the oracle's eager
    attention (softmax in fp32, probabilities rounded to fp16, GQA by repeat_kv)."""
    from oracle import llama_oracle as O
    dev, ctx, planner = env["dev"], env["ctx"], env["planner"]
    shapes = [(70, 3, 9), (1, 2, 5), (129, 4, 33), (64, 1, 64), (200, 2, 70), (384, 4, 32), (30, 2, 150)]
    rows = sum(lp + S * ls for lp, S, ls in shapes)                                   # (len_p, S, len_s)
    ctx.call("fls_set_option", b"attention_impl", impl)
    ctx.call("fls_set_option", b"attention_k_stages", k_stages)      # K buffers per CTA: this device's default, 3, 2
    ld = (nq + 2 * nkv) * d
    qkv = _rand((rows, ld), 1.0, 13, dev)
    tiles, base = [], 0
    for lp, S, ls in shapes:
        tiles += planner.prompt_tiles(base, lp, S, ls, prefix_causal)
        base += lp + S * ls
    t = torch.tensor(tiles, dtype=torch.int32, device=dev)
    out = torch.full((rows, nq * d), float("nan"), dtype=torch.float16, device=dev)
    ctx.call("fls_op_attention", qkv.data_ptr(), ld, rows, nq, nkv, d, t.data_ptr(), len(tiles), out.data_ptr(), nq * d,
             _sp())
    torch.cuda.synchronize()
    ctx.call("fls_set_option", b"attention_impl", 2)
    ctx.call("fls_set_option", b"attention_k_stages", 0)
    assert torch.isfinite(out).all()

    q_all, k_all, v_all = qkv.cpu().float().split([nq * d, nkv * d, nkv * d], dim=1)
    mask_full = torch.finfo(torch.float32).min * torch.ones((512, 512)).triu(1)[None, None]
    base = 0
    worst = 0.0
    for lp, S, ls in shapes:
        def heads(x, n, lo, L):
            return x[lo:lo + L].view(1, L, n, d).transpose(1, 2)
        qp, kp, vp = heads(q_all, nq, base, lp), heads(k_all, nkv, base, lp), heads(v_all, nkv, base, lp)
        pm = mask_full[:, :, :lp, :lp] if prefix_causal else None
        want_p = O.attention(qp, kp, vp, pm, nq // nkv).reshape(lp, nq * d)
        worst = max(worst, (out[base:base + lp].cpu().float() - want_p).abs().max().item())
        for s in range(S):
            lo = base + lp + s * ls
            qs, ks, vs = heads(q_all, nq, lo, ls), heads(k_all, nkv, lo, ls), heads(v_all, nkv, lo, ls)
            sm = mask_full[:, :, 512 - ls:, 512 - lp - ls:]
            want_s = O.attention(qs, torch.cat((kp, ks), 2), torch.cat((vp, vs), 2), sm, nq // nkv).reshape(ls, nq * d)
            worst = max(worst, (out[lo:lo + ls].cpu().float() - want_s).abs().max().item())
        base += lp + S * ls
    # fp32 oracle vs fp16 P and fp16 output rounding: values are O(1), one fp16 ulp is ~1e-3
    assert worst <= 6e-3, worst


@pytest.mark.parametrize("rows,V,K", [(12, 1000, 512), (300, 32000, 256)])
def test_head_softmax_and_logprob(env, rows, V, K):
    from flexible_llm_sharding_b200 import _lib
    dev = env["dev"]
    desc = _lib.ModelDesc(K, 1408, 4, 2, 128, V, 2, 1e-5)
    ctx = _lib.Context(desc, 0)
    ctx.call("fls_configure", 1, C.c_size_t(256), 256, 512)
    x, w = _rand((rows, K), 1.0, 14, dev), _rand((V, K), 0.2, 15, dev)
    probs = torch.empty((rows, V), dtype=torch.float16, device=dev)
    ids = torch.randint(0, V, (rows, 3), dtype=torch.int32, device=dev)
    lp = torch.empty((rows, 3), dtype=torch.float32, device=dev)
    lse = torch.empty((rows,), dtype=torch.float32, device=dev)
    ctx.call("fls_op_head", x.data_ptr(), w.data_ptr(), rows, V, K, probs.data_ptr(), ids.data_ptr(), 3, lp.data_ptr(),
             lse.data_ptr(), _sp())
    torch.cuda.synchronize()
    logits = (x.float() @ w.float().t()).half()                 # lm_head output as the reference holds it
    want = torch.softmax(logits.float(), dim=-1)
    # probabilities: relative error of one fp16 rounding (2^-11) + a logit that rounded to its fp16 neighbour
    # (checked on fp16-normal outputs; below 6.1e-5 fp16 is subnormal with absolute steps of 6e-8)
    rel = ((probs.float() - want).abs() / want.clamp_min(1e-7))[want > 1e-4].max().item()
    assert rel <= 2e-2, rel
    assert (probs.float() - want).abs()[want <= 1e-4].max().item() <= 2e-6
    assert abs(probs.float().sum(-1) - 1).max().item() <= 2e-3
    want_lse = torch.logsumexp(logits.float(), dim=-1)
    assert (lse - want_lse).abs().max().item() <= 2e-3
    want_lp = torch.log_softmax(logits.float(), dim=-1).gather(1, ids.long())
    assert (lp - want_lp).abs().max().item() <= 2e-2
    ctx.close()


def test_errors_are_reported_not_fatal(env):
    ctx, lib = env["ctx"], env["lib"]
    with pytest.raises(lib.FlsError) as e:
        ctx.call("fls_op_gemm", 0, None, 7, None, 7, 8, 8, 8, 7, None, 8, None, 0, 0, None)
    assert e.value.code == -1
    with pytest.raises(lib.FlsError) as e:
        ctx.call("fls_acquire_shard", 0, None)
    assert e.value.code == -4
    with pytest.raises(lib.FlsError) as e:
        ctx.call("fls_register_layer_file", 0, 0, b"/nonexistent/x.safetensors", 0)
    assert e.value.code == -3
    assert b"cannot open" in lib.load().fls_last_error(ctx._h)
