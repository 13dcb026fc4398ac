"""Parity at the shapes that are benchmarked (BASELINE.json configs): 2-layer random-init models with the 7B,
13B and 70B layer shapes (hidden 4096 / 5120 / 8192, intermediate 11008 / 13824 / 28672, 32 / 40 / 64 heads,
GQA 8:1 at 70B, vocabulary 32000), ragged 512-row prompts plus one 2048-row prompt (the C5 shape), scored by
the CUDA path (ShardedLlama -> ctypes -> C ABI) and by two checkers on the same layer files:

  * the fp32 CPU oracle (restated reference loop, pinned to the reference by tests/test_reference_utils.py),
  * the fp16 oracle on the GPU - torch eager + cuBLAS, one prompt at a time: the arithmetic path the reference
    itself takes on this box (utils.py:228-297 through transformers eager).

Tolerance as in tests/test_scoring_gpu.py: |d log p| <= 5e-2 on tokens with p >= 1e-4, same top-1 token and
same option ranking wherever the checker's own margin exceeds twice that.  Also here: configs[0] exactly
(tiny model, 32 x 256-row prompts, k = 1, storage_location = cpu) and the GEMM / attention kernels alone at
the full-size operand shapes.
"""
import copy
import ctypes as C
import os
import shutil
import tempfile

import numpy as np
import pytest
import torch

from _parity import LOGP_TOL, P_MIN, check_against_reference, make_args

pytestmark = pytest.mark.gpu


def _scratch_dir(tag):
    base = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
    return tempfile.mkdtemp(prefix=f"fls_test_{tag}_", dir=base)


@pytest.fixture(scope="module", params=["7b", "13b", "70b"])
def real_model(request):
    from flexible_llm_sharding_b200 import synthetic
    name = request.param
    shape = synthetic.LlamaShape.named(name, num_hidden_layers=2)
    d = _scratch_dir(name)
    synthetic.write_layer_files(d, shape, seed=3, init_std=0.02, device="cuda:0")
    # four ragged 512-row prompts (prefix 384 + 4 options of up to 32 tokens) and one 2048-row prompt (1792 + 4 x 64)
    prompts = (synthetic.make_prompts(4, shape.vocab_size, 384, 4, 32, seed=77, ragged=True)
               + synthetic.make_prompts(1, shape.vocab_size, 1792, 4, 64, seed=78, ragged=True))
    yield dict(name=name, shape=shape, layer_dir=d, prompts=prompts)
    shutil.rmtree(d, ignore_errors=True)


def _cuda_scores(case, **kw):
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    args = make_args(case["layer_dir"], **kw)
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    out = model(copy.deepcopy(case["prompts"]))
    launches = model.ctx.stats().kernel_launches
    model.ctx.close()
    return out, launches


def _dlogp(got, want):
    g, w = got.astype(np.float64), want.astype(np.float64)
    m = w >= P_MIN
    assert m.sum() > 100
    return np.abs(np.log(np.maximum(g[m], 1e-30)) - np.log(w[m])).max()


def test_scores_match_both_checkers_at_full_layer_shapes(real_model):
    from oracle import llama_oracle as O
    case = real_model
    out, launches = _cuda_scores(case, max_rows=2304)          # several micro-batches, the 2048-row prompt alone in one
    assert launches > 0 and len(out) == len(case["prompts"])
    V = case["shape"].vocab_size
    assert all(o.shape == (4, 1, V) and o.dtype == np.float16 for o in out)
    torch.set_num_threads(os.cpu_count())
    want32 = O.OracleShardedLlama(case["layer_dir"], dtype=torch.float32)(copy.deepcopy(case["prompts"]))
    want16 = O.OracleShardedLlama(case["layer_dir"], dtype=torch.float16, device="cuda:0")(copy.deepcopy(case["prompts"]))
    w32 = np.stack([s[:, 0] for s in want32])
    w16 = np.stack([s[:, 0] for s in want16])
    worst32 = check_against_reference(out, w32)
    got = np.stack([o[:, 0] for o in out])
    worst16 = _dlogp(got, w16)
    eps0 = _dlogp(w16, w32)                                    # the reference's own fp16-vs-fp32 gap at this shape
    print(f"{case['name']}: max |dlogp| vs fp32 CPU oracle {worst32:.4f}, vs fp16 torch-eager on the GPU {worst16:.4f}; "
          f"fp16 eager vs fp32 oracle {eps0:.4f}")
    assert worst16 <= LOGP_TOL
    # option ranking on an answer token against the fp16 eager run too
    gl, wl = np.log(np.maximum(got[..., 5].astype(np.float64), 1e-30)), np.log(np.maximum(w16[..., 5].astype(np.float64), 1e-30))
    for p in range(got.shape[0]):
        order = np.argsort(-wl[p])
        if (-np.diff(wl[p][order]) > 2 * LOGP_TOL).all():
            assert (np.argsort(-gl[p]) == order).all()


def test_schedule_invariance_at_full_layer_shapes(real_model):
    """One micro-batch per prompt vs everything in one, shard size 1 vs the whole model, host activation tier:
    bit-identical scores at the real operand sizes too."""
    case = real_model
    base, _ = _cuda_scores(case, max_rows=4096)
    other, _ = _cuda_scores(case, max_rows=2048, layer_num_per_shard=5, storage_location="cpu", force_storage_tier=True)
    for a, b in zip(base, other):
        assert np.array_equal(a, b)


def test_config0_exactly(tiny):
    """BASELINE.json configs[0]: tiny model split by prepare_weights.py, layer_num_per_shard = 1,
    storage_location = cpu (forced, so the pinned-host tier really runs), 32 prompts x 256 rows."""
    from flexible_llm_sharding_b200 import synthetic
    from oracle import llama_oracle as O
    prompts = synthetic.make_prompts(32, tiny["shape"].vocab_size, 192, 4, 16, seed=1234)
    case = dict(tiny, prompts=prompts)
    out, _ = _cuda_scores(case, layer_num_per_shard=1, storage_location="cpu", force_storage_tier=True)
    want = O.OracleShardedLlama(tiny["layer_dir"], dtype=torch.float32)(copy.deepcopy(prompts))
    worst = check_against_reference(out, np.stack([s[:, 0] for s in want]))
    print(f"configs[0] (32 x 256 rows): max |dlogp| = {worst:.4f}")


# ------------------------------------------------------------------ kernels alone at full operand sizes
@pytest.fixture(scope="module")
def big_ctx():
    from flexible_llm_sharding_b200 import _lib, planner
    desc = _lib.ModelDesc(8192, 28672, 64, 8, 128, 32000, 2, 1e-5)
    ctx = _lib.Context(desc, 0)
    ctx.call("fls_configure", 1, C.c_size_t(256), 128, 128)
    cos, sin = planner.rope_tables(4096, 128)
    ctx.call("fls_set_rope_table", cos.data_ptr(), sin.data_ptr(), 4096)
    yield ctx
    ctx.close()


def _sp():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _randn(shape, std, seed):
    g = torch.Generator(device="cuda:0").manual_seed(seed)
    return (std * torch.randn(shape, generator=g, device="cuda:0", dtype=torch.float32)).half()


@pytest.mark.parametrize("M,N,K,epi", [
    (16384, 12288, 4096, "store"),       # 7B q|k|v
    (16384, 4096, 11008, "residual"),    # 7B down + residual
    (16384, 10240, 8192, "store"),       # 70B q|k|v (64 + 2 x 8 heads)
    (16384, 8192, 28672, "residual"),    # 70B down + residual
    (16384, 28672, 8192, "swiglu"),      # 70B gate|up (B has 2 x 28672 rows)
    (16384, 13824, 5120, "swiglu"),      # 13B gate|up
    (16000, 5120, 5120, "residual"),     # 13B o + residual, ragged M
])
def test_gemm_at_model_shapes(big_ctx, M, N, K, epi):
    """tcgen05 GEMM vs an fp32 torch.matmul of the same fp16 operands (TF32 off): fp32 accumulation on both
    sides, so only summation order and the final fp16 rounding differ - at most one fp16 ulp of the result."""
    torch.backends.cuda.matmul.allow_tf32 = False
    a = _randn((M, K), 1.0, 1)
    b = _randn(((2 * N if epi == "swiglu" else N), K), 0.02, 2)
    chunk = 2048                                    # fp32 reference in row chunks (keeps its memory small)
    if epi == "swiglu":
        c = torch.empty((M, N), dtype=torch.float16, device="cuda:0")
        big_ctx.call("fls_op_gemm", 2, a.data_ptr(), K, b.data_ptr(), K, 2 * N, M, N, K, c.data_ptr(), N, None, 0, N, _sp())
    elif epi == "residual":
        x0 = _randn((M, N), 1.0, 3)
        c = x0.clone()
        big_ctx.call("fls_op_gemm", 1, a.data_ptr(), K, b.data_ptr(), K, N, M, N, K, c.data_ptr(), N, c.data_ptr(), N, 0, _sp())
    else:
        c = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda:0")
        big_ctx.call("fls_op_gemm", 0, a.data_ptr(), K, b.data_ptr(), K, N, M, N, K, c.data_ptr(), N, None, 0, 0, _sp())
    torch.cuda.synchronize()
    assert torch.isfinite(c).all()
    worst, mismatched = 0.0, 0
    for lo in range(0, M, chunk):
        af = a[lo:lo + chunk].float()
        if epi == "swiglu":
            g = (af @ b[:N].float().t()).half()
            u = (af @ b[N:].float().t()).half()
            want = torch.nn.functional.silu(g) * u
        elif epi == "residual":
            want = x0[lo:lo + chunk] + (af @ b.float().t()).half()
        else:
            want = (af @ b.float().t()).half()
        got = c[lo:lo + chunk]
        scale = max(1.0, want.float().abs().max().item())
        worst = max(worst, (got.float() - want.float()).abs().max().item() / scale)
        mismatched += (got != want).sum().item()
    # one fp16 ulp relative to the largest magnitude (2^-10); two where a second rounding follows the GEMM's own
    # (the SwiGLU product of two rounded values, the fp16 residual add of a rounded projection)
    assert worst <= (2 ** -10 if epi == "store" else 2 ** -9), worst
    # how often two fp32 summation orders land on different sides of an fp16 rounding boundary grows with K
    # (5.3 % of the entries at K = 28672)
    assert mismatched / (M * N) < 0.10


@pytest.mark.parametrize("impl", [1, 2])          # one CTA per (query tile, head) / persistent CTAs
@pytest.mark.parametrize("nq,nkv", [(64, 8), (40, 40), (32, 32)])
def test_attention_at_model_head_counts(big_ctx, nq, nkv, impl):
    """Shared-prefix attention with 70B / 13B / 7B head counts (GQA 8:1 at 70B) on the benchmark's prompt shapes
    (384 + 4 x 32 and 1792 + 4 x 64 rows) against the oracle's eager attention evaluated in fp32 on the GPU."""
    from flexible_llm_sharding_b200 import planner
    from oracle import llama_oracle as O
    d = 128
    shapes = [(384, 4, 32), (1792, 4, 64), (383, 3, 31)]
    rows = sum(lp + S * ls for lp, S, ls in shapes)
    ld = (nq + 2 * nkv) * d
    qkv = _randn((rows, ld), 1.0, 13)
    tiles, base = [], 0
    for lp, S, ls in shapes:
        tiles += planner.prompt_tiles(base, lp, S, ls, False)
        base += lp + S * ls
    t = torch.tensor(tiles, dtype=torch.int32, device="cuda:0")
    out = torch.full((rows, nq * d), float("nan"), dtype=torch.float16, device="cuda:0")
    big_ctx.call("fls_set_option", b"attention_impl", impl)
    big_ctx.call("fls_op_attention", qkv.data_ptr(), ld, rows, nq, nkv, d, t.data_ptr(), len(tiles), out.data_ptr(),
                 nq * d, _sp())
    torch.cuda.synchronize()
    big_ctx.call("fls_set_option", b"attention_impl", 2)
    assert torch.isfinite(out).all()
    q_all, k_all, v_all = qkv.float().split([nq * d, nkv * d, nkv * d], dim=1)
    mask_full = (torch.finfo(torch.float32).min * torch.ones((2048, 2048), device="cuda:0")).triu(1)[None, None]
    base, worst = 0, 0.0

    def heads(x, n, lo, L):
        return x[lo:lo + L].view(1, L, n, d).transpose(1, 2)

    for lp, S, ls in shapes:
        qp, kp, vp = heads(q_all, nq, base, lp), heads(k_all, nkv, base, lp), heads(v_all, nkv, base, lp)
        want_p = O.attention(qp, kp, vp, None, nq // nkv).reshape(lp, nq * d)
        worst = max(worst, (out[base:base + lp].float() - want_p).abs().max().item())
        for s in range(S):
            lo = base + lp + s * ls
            qs, ks, vs = heads(q_all, nq, lo, ls), heads(k_all, nkv, lo, ls), heads(v_all, nkv, lo, ls)
            sm = mask_full[:, :, 2048 - ls:, 2048 - lp - ls:]
            want_s = O.attention(qs, torch.cat((kp, ks), 2), torch.cat((vp, vs), 2), sm, nq // nkv).reshape(ls, nq * d)
            worst = max(worst, (out[lo:lo + ls].float() - want_s).abs().max().item())
        base += lp + S * ls
    assert worst <= 6e-3, worst            # same bound as the small-shape test (fp16 P and output rounding)
