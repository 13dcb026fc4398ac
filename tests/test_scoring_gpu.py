"""End-to-end parity of the CUDA scoring path (through ShardedLlama -> ctypes -> C ABI) with the
committed golden vectors (real transformers modules under the reference call convention) and with
the CPU oracle run on the same inputs.

Tolerance (calibrated on the golden vectors): eps0 = max |log p_fp16 - log p_fp32|
between the fp16 and fp32 reference runs over tokens with p >= 1e-4 is 0.015-0.016; the CUDA path
must stay within LOGP_TOL = max(2*eps0, 5e-2) = 5e-2 of the fp32 reference on those tokens, and
give the same ranking of options / same top-1 token wherever the reference's own margin exceeds
2*LOGP_TOL (ties inside the tolerance are not rankable in fp16 by either implementation).
"""
import copy
import os
import pickle
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from _parity import ANSWER_TOKEN, LOGP_TOL, P_MIN, check_against_reference, make_args  # noqa: E402,F401


def run_model(case, prompts=None, **kw):
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    args = make_args(case["layer_dir"], **kw)
    dm = DeviceManager(args, ["cuda:0"])
    model = ShardedLlama(args, dm, device="cuda:0")
    prompts = copy.deepcopy(case["prompts"] if prompts is None else prompts)
    out = model(prompts)
    assert all(p is None for p in prompts)                    # reference utils.py:259
    return model, out


@pytest.mark.parametrize("which", ["tiny", "tiny_gqa"])
@pytest.mark.parametrize("causal", [False, True])
def test_scores_match_golden(which, causal, request):
    case = request.getfixturevalue(which)
    tag = "causal" if causal else "ref"
    model, out = run_model(case, prefix_causal=causal)
    assert len(out) == len(case["prompts"])
    S = len(case["prompts"][0][1])
    assert out[0].shape == (S, 1, case["shape"].vocab_size)
    worst = check_against_reference(out, case["golden"][f"probs32_{tag}"])
    # and as close to the fp16 reference run as that run is to fp32
    g16 = case["golden"][f"probs16_{tag}"].astype(np.float64)
    got = np.stack([o[:, 0] for o in out]).astype(np.float64)
    m = g16 >= P_MIN
    assert np.abs(np.log(np.maximum(got[m], 1e-30)) - np.log(g16[m])).max() <= LOGP_TOL
    print(f"{which}/{tag}: max |dlogp| vs fp32 reference = {worst:.4f}")


@pytest.mark.parametrize("which", ["tiny", "tiny_gqa"])
@pytest.mark.parametrize("causal", [False, True])
def test_edge_shapes_match_golden(which, causal, request):
    """The CUDA path on the shapes of tests/golden/*-edge.npz (real HF modules through the reference's call
    convention): an empty option scored at its wrapped-around last pad position, an empty prefix, one option,
    five very ragged options, options / a prefix longer than one 128-row attention tile."""
    case = request.getfixturevalue(which)
    tag = "causal" if causal else "ref"
    prompts = case["edge_prompts"]
    _, out = run_model(case, prompts=prompts, prefix_causal=causal, layer_num_per_shard=2, max_rows=400)
    assert len(out) == len(prompts)
    worst = 0.0
    for i, ((_, options), o) in enumerate(zip(prompts, out)):
        assert o.shape == (len(options), 1, case["shape"].vocab_size) and o.dtype == np.float16
        for suffix, ref in (("32", 1.0), ("16", 1.0)):
            w = case["edge_golden"][f"probs{suffix}_{tag}_{i}"].astype(np.float64)
            g = o[:, 0].astype(np.float64)
            m = w >= P_MIN
            d = np.abs(np.log(np.maximum(g[m], 1e-30)) - np.log(w[m])).max()
            assert d <= LOGP_TOL, (i, suffix, d)
            worst = max(worst, d)
        assert np.abs(o[:, 0].astype(np.float64).sum(-1) - 1).max() <= 5e-3
    print(f"{which}/{tag} edge shapes: max |dlogp| = {worst:.4f}")


@pytest.mark.parametrize("k,storage,num_batch,max_rows", [
    (2, "gpu", 1, 16384), (3, "cpu", 2, 16384), (7, "gpu", 1, 200), (1, "cpu", 1, 160), (2, "disk", 1, 250)])
def test_schedule_does_not_change_scores(tiny, k, storage, num_batch, max_rows, tmp_path):
    """Shard size, activation tier (forced), batching and micro-batch size are scheduling choices:
    the scores must be bit-identical to the k=1 / HBM-resident run."""
    from flexible_llm_sharding_b200.main import run_sharded_llama
    from flexible_llm_sharding_b200.utils import DeviceManager
    _, base = run_model(tiny)
    args = make_args(tiny["layer_dir"], layer_num_per_shard=k, storage_location=storage, num_batch=num_batch,
                     max_rows=max_rows, force_storage_tier=True, disk_folder=str(tmp_path / "act"))
    dm = DeviceManager(args, ["cuda:0"])
    out = run_sharded_llama("cuda:0", copy.deepcopy(tiny["prompts"]), dm, args)
    assert len(out) == len(base)
    for a, b in zip(out, base):
        assert np.array_equal(a, b)
    if storage == "disk":
        assert not [f for f in os.listdir(tmp_path / "act") if f.endswith(".npy")]      # cleaned up


@pytest.mark.parametrize("storage,max_rows", [("gpu", 400), ("cpu", 330), ("disk", 450), ("gpu", 16384)])
def test_length_bucketed_packing_does_not_change_scores(tiny_gqa, storage, max_rows, tmp_path):
    """assigning prompts to micro-batches by length (the default) instead of in the order given
    moves rows between launches, nothing else - every prompt's scores are bit for bit the same, in the caller's order."""
    from flexible_llm_sharding_b200 import planner
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    prompts = list(tiny_gqa["prompts"]) + list(tiny_gqa["edge_prompts"])
    outs, moved = {}, None
    for order in ("input", "length"):
        args = make_args(tiny_gqa["layer_dir"], storage_location=storage, max_rows=max_rows, force_storage_tier=True,
                         disk_folder=str(tmp_path / f"act_{order}"), pack_order=order)
        model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
        assert model.pack_order == order
        probs, packed = model.run(copy.deepcopy(prompts))
        outs[order] = [probs[slice(*packed.head_range(i))] for i in range(len(prompts))]
        if order == "length":
            moved = packed.order
            tok = planner.tokenize_prompts(model.tokenizer, copy.deepcopy(prompts))
            assert [packed.shapes[p] for p in range(len(tok))] == \
                   [(tok[packed.prompt_index(p)].len_p, tok[packed.prompt_index(p)].n_options,
                     tok[packed.prompt_index(p)].len_s) for p in range(len(tok))]
        else:
            assert packed.order is None
        model.close()
    assert moved is not None and sorted(moved.tolist()) == list(range(len(prompts)))   # the prompts really moved
    for a, b in zip(outs["input"], outs["length"]):
        assert a.shape == b.shape and np.array_equal(a, b)


def test_disk_tier_writes_the_reference_file_layout(tiny, tmp_path, monkeypatch):
    """`storage_location=disk` keeps the reference's files (utils.py:171-177): `prefix{rank}-{idx:05d}.npy`
    [1, len_p, H] and `suffix…` [S, len_s, H], float16, one pair per prompt, loadable with plain np.load."""
    from flexible_llm_sharding_b200.activation_store import ActivationStore
    from flexible_llm_sharding_b200.main import run_sharded_llama
    from flexible_llm_sharding_b200.utils import DeviceManager
    monkeypatch.setattr(ActivationStore, "keep_disk_files", True)
    folder = tmp_path / "act"
    args = make_args(tiny["layer_dir"], layer_num_per_shard=2, storage_location="disk", max_rows=250,
                     force_storage_tier=True, disk_folder=str(folder), data_parallel=True)
    dm = DeviceManager(args, ["cuda:0"])
    prompts = copy.deepcopy(tiny["prompts"])
    out = run_sharded_llama("cuda:0", copy.deepcopy(prompts), dm, args)
    _, base = run_model(tiny)
    for a, b in zip(out, base):
        assert np.array_equal(a, b)
    files = sorted(os.listdir(folder))
    assert files == sorted([f"prefix0-{i:05d}.npy" for i in range(len(prompts))] +
                           [f"suffix0-{i:05d}.npy" for i in range(len(prompts))])
    for i, (_, options) in enumerate(prompts):
        pre = np.load(folder / f"prefix0-{i:05d}.npy")
        suf = np.load(folder / f"suffix0-{i:05d}.npy")
        assert pre.dtype == np.float16 and suf.dtype == np.float16
        assert pre.ndim == 3 and pre.shape[0] == 1 and suf.ndim == 3 and suf.shape[0] == len(options)
        assert pre.shape[2] == suf.shape[2] and np.isfinite(pre.astype(np.float32)).all()


def test_kernel_variants_give_identical_scores(tiny_gqa):
    """RoPE in the QKV epilogue vs the stand-alone kernel, and cta_group 1 vs 2 GEMMs, are the same
    arithmetic: bit-identical scores.  tcgen05 vs mma.sync attention differ only in summation order."""
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    args = make_args(tiny_gqa["layer_dir"])
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    base = model(copy.deepcopy(tiny_gqa["prompts"]))
    for opt, val in [(b"fuse_rope", 0), (b"gemm_cta_group", 1), (b"gemm_tma_store", 0), (b"gemm_wide_tiles", 1)]:
        model.ctx.call("fls_set_option", opt, val)
        out = model(copy.deepcopy(tiny_gqa["prompts"]))
        model.ctx.call("fls_set_option", opt, 2 if opt in (b"gemm_cta_group", b"gemm_wide_tiles") else 1)
        for a, b in zip(out, base):
            assert np.array_equal(a, b), opt
    model.ctx.call("fls_set_option", b"attention_impl", 0)
    out = model(copy.deepcopy(tiny_gqa["prompts"]))
    check_against_reference(out, tiny_gqa["golden"]["probs32_ref"])


def test_resident_weights_stream_once(tiny):
    """with one HBM slot per shard the weights are streamed on the first pass only; later
    passes (more batches, more generated tokens) reuse them and give the same scores."""
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    _, base = run_model(tiny)
    args = make_args(tiny["layer_dir"], resident_weights="always", layer_num_per_shard=2)
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    first = model(copy.deepcopy(tiny["prompts"]))
    streamed = model.ctx.stats().h2d_bytes
    assert streamed > 0
    second = model(copy.deepcopy(tiny["prompts"][:2]))
    assert model.ctx.stats().h2d_bytes == streamed            # nothing streamed again
    for a, b in zip(first, base):
        assert np.array_equal(a, b)
    for a, b in zip(second, base[:2]):
        assert np.array_equal(a, b)


def test_bounded_pinned_cache_rereads_layers(tiny):
    """With a pinned-host bound smaller than the model (hosts with little RAM: the reference's raison d'etre)
    layer images are evicted after use and read from their files again on the next pass; scores unchanged."""
    from flexible_llm_sharding_b200 import _lib
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    _, base = run_model(tiny)
    args = make_args(tiny["layer_dir"], pinned_limit_bytes=1)
    import gc
    gc.collect()                                              # contexts of earlier tests release their slot pins
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    try:
        others = model.ctx.stats().pinned_bytes               # process-wide figure: images other live contexts hold
        out1 = model(copy.deepcopy(tiny["prompts"]))
        r1 = model.ctx.stats().file_bytes_read
        out2 = model(copy.deepcopy(tiny["prompts"]))
        st = model.ctx.stats()
        assert st.file_bytes_read > r1                        # layers were read again
        assert st.pinned_bytes - others < sum(model._layer_bytes)      # and not all of them are held
        for a, b, c in zip(out1, out2, base):
            assert np.array_equal(a, c) and np.array_equal(b, c)
    finally:
        _lib.load().fls_set_host_cache_limit(0)


def _recast_layer_dir(src, dst, cast):
    """Copy a layer directory, passing every tensor through `cast` (non-weight files are copied as they are)."""
    import shutil
    from safetensors.torch import load_file, save_file
    os.makedirs(dst, exist_ok=True)
    for name in os.listdir(src):
        if name.endswith(".safetensors"):
            save_file({k: cast(v).contiguous() for k, v in load_file(os.path.join(src, name)).items()},
                      os.path.join(dst, name))
        else:
            shutil.copy(os.path.join(src, name), os.path.join(dst, name))
    return str(dst)


def test_fp32_and_bf16_layer_files_are_converted_to_fp16_on_load(tiny, tmp_path):
    """The reference casts whatever dtype the layer files hold to fp16 when it loads them
    (`set_module_tensor_to_device(..., dtype=self.dtype)`, utils.py:130); the pinned-host loader does the same.
    fp32 copies of fp16 weights must score bit-identically; bf16 files must score like fp16 files holding the
    bf16-rounded values (bf16 -> fp16 is exact for weights of this magnitude)."""
    _, base = run_model(tiny)
    d32 = _recast_layer_dir(tiny["layer_dir"], tmp_path / "f32", lambda t: t.float())
    _, out32 = run_model(dict(tiny, layer_dir=d32))
    for a, b in zip(out32, base):
        assert np.array_equal(a, b)
    dbf = _recast_layer_dir(tiny["layer_dir"], tmp_path / "bf16", lambda t: t.bfloat16())
    dbf_as_f16 = _recast_layer_dir(tiny["layer_dir"], tmp_path / "bf16_f16", lambda t: t.bfloat16().half())
    _, out_bf = run_model(dict(tiny, layer_dir=dbf))
    _, out_ref = run_model(dict(tiny, layer_dir=dbf_as_f16))
    for a, b in zip(out_bf, out_ref):
        assert np.array_equal(a, b)
    assert any(not np.array_equal(a, b) for a, b in zip(out_bf, base))      # the rounding was really applied


def test_fused_answer_token_logprobs(tiny):
    """lm_head + log-softmax + gather without the full-vocabulary output."""
    model, out = run_model(tiny)
    lp = model.score_tokens(copy.deepcopy(tiny["prompts"]), [ANSWER_TOKEN, 7])
    want = np.log(tiny["golden"]["probs32_ref"].astype(np.float64))
    for p in range(len(lp)):
        assert lp[p].shape == (want.shape[1], 2) and lp[p].dtype == np.float32
        assert np.abs(lp[p][:, 0] - want[p, :, ANSWER_TOKEN]).max() <= LOGP_TOL
        assert np.abs(lp[p][:, 1] - want[p, :, 7]).max() <= LOGP_TOL
    st = model.ctx.stats()
    assert st.kernel_launches > 0 and st.h2d_bytes > 0


def test_oracle_on_fresh_inputs(tiny_gqa):
    """Not only the committed vectors: new seeded prompts, CUDA vs the CPU oracle (fp32)."""
    from flexible_llm_sharding_b200 import synthetic
    from oracle import llama_oracle as O
    prompts = synthetic.make_prompts(5, tiny_gqa["shape"].vocab_size, 90, 5, 20, seed=77, ragged=True)
    case = dict(tiny_gqa, prompts=prompts)
    _, out = run_model(case, layer_num_per_shard=3)
    orc = O.OracleShardedLlama(tiny_gqa["layer_dir"], dtype=torch.float32)
    want = np.stack([s[:, 0] for s in orc(copy.deepcopy(prompts))])
    check_against_reference(out, want)


@pytest.mark.parametrize("causal", [False, True])
def test_long_prompts_many_key_tiles(tiny_gqa, causal):
    """Long prefix (11+ key tiles of 64, 6 query tiles) and options longer than one 128-row query tile
    (an option is then cut into several tiles that see their own earlier rows), ragged lengths, one
    micro-batch per prompt: CUDA vs the CPU oracle."""
    from flexible_llm_sharding_b200 import synthetic
    from oracle import llama_oracle as O
    V = tiny_gqa["shape"].vocab_size
    prompts = synthetic.make_prompts(2, V, 700, 2, 300, seed=11, ragged=True) + \
        synthetic.make_prompts(1, V, 1, 3, 129, seed=12, ragged=True)          # BOS-only prefix
    case = dict(tiny_gqa, prompts=prompts)
    _, out = run_model(case, layer_num_per_shard=7, max_rows=1400, prefix_causal=causal)
    orc = O.OracleShardedLlama(tiny_gqa["layer_dir"], dtype=torch.float32, prefix_causal=causal)
    want = [s[:, 0] for s in orc(copy.deepcopy(prompts))]
    for o, w in zip(out, want):
        check_against_reference([o], w[None])


def _extend_options(prompts, words):
    """What the generation loop does (main.py:84-90): append a decoded token to every option."""
    return [(pre, tuple(o + " " + words[(i + j) % len(words)] for j, o in enumerate(opts)))
            for i, (pre, opts) in enumerate(prompts)]


@pytest.mark.parametrize("which,storage,max_rows,k", [("tiny", "gpu", 16384, 1), ("tiny_gqa", "gpu", 300, 2),
                                                      ("tiny_gqa", "cpu", 300, 3), ("tiny", "disk", 250, 2)])
def test_prefix_kv_cache_gives_identical_scores(which, storage, max_rows, k, request, tmp_path):
    """later passes over the same prefixes (the generation loop) reuse the prefix K/V of every
    layer and push only the option rows through; the scores must be bit-identical to recomputing everything."""
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    case = request.getfixturevalue(which)
    kw = dict(layer_num_per_shard=k, storage_location=storage, max_rows=max_rows, force_storage_tier=True,
              disk_folder=str(tmp_path / "act"))
    p1 = copy.deepcopy(case["prompts"])
    p2 = _extend_options(p1, ["w7", "w11", "w5"])
    p3 = _extend_options(p2, ["w9", "w4"])
    plain = ShardedLlama(make_args(case["layer_dir"], **kw), DeviceManager(make_args(case["layer_dir"]), ["cuda:0"]))
    want = [plain(copy.deepcopy(p)) for p in (p1, p2, p3)]
    assert plain.last_timing["prefix_kv"] == "off"
    args = make_args(case["layer_dir"], prefix_kv_cache="on", **kw)
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]))
    notes, rows = [], []
    for step, p in enumerate((p1, p2, p3)):
        if step == 2:
            args.prefix_kv_cache = "last"
        got = model(copy.deepcopy(p))
        notes.append(model.last_timing["prefix_kv"])
        rows.append(model.last_timing["rows"])
        for a, b in zip(got, want[step]):
            assert np.array_equal(a, b), (step, notes)
    assert notes == ["save", "load", "load"] and rows[1] < rows[0] and not model._kv
    # a different set of prefixes never hits the cache
    args.prefix_kv_cache = "on"
    model(copy.deepcopy(p1))
    other = [(pre + " w3", opts) for pre, opts in p1]
    model(copy.deepcopy(other))
    assert model.last_timing["prefix_kv"] == "save" and len(model._kv) == 2
    model.drop_prefix_kv_cache()
    assert not model._kv


def test_prefix_kv_cache_with_the_mma_sync_attention_kernel(tiny_gqa):
    """The comparison attention kernel (the one head_dim 64 models use) takes the same tile descriptors: the
    [cached prefix rows | option rows] layout must give it the same bits as recomputing."""
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    p1 = copy.deepcopy(tiny_gqa["prompts"])
    p2 = _extend_options(p1, ["w7", "w11", "w5"])
    outs = {}
    for policy in ("off", "on"):
        args = make_args(tiny_gqa["layer_dir"], prefix_kv_cache=policy, layer_num_per_shard=2, max_rows=300)
        model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]))
        model.ctx.call("fls_set_option", b"attention_impl", 0)
        outs[policy] = [model(copy.deepcopy(p)) for p in (p1, p2)]
        assert model.last_timing["prefix_kv"] == ("load" if policy == "on" else "off")
    for step in range(2):
        for a, b in zip(outs["off"][step], outs["on"][step]):
            assert np.array_equal(a, b), step


def test_cli_generation_with_and_without_prefix_kv_cache(tiny, tmp_path, monkeypatch):
    """main.py --num_gen_token 3: the prefix K/V cache is on by default and must not change a single bit."""
    from flexible_llm_sharding_b200.main import main
    res = {}
    for mode in ("off", "auto"):
        monkeypatch.setenv("FLS_PREFIX_KV_CACHE", mode)
        pp = tmp_path / f"prompts_{mode}.pkl"
        with open(pp, "wb") as f:
            pickle.dump(tiny["prompts"][:5], f)
        outp = tmp_path / f"scores_{mode}.pkl"
        main(["--model_path", tiny["layer_dir"], "--prompt_pickle", str(pp), "--output_file", str(outp),
              "--storage_location", "gpu", "--layer_num_per_shard", "3", "--num_gen_token", "3", "--num_batch", "2"])
        with open(outp, "rb") as f:
            scores = pickle.load(f)
        with open(str(pp).replace(".pkl", "_updated.pkl"), "rb") as f:
            res[mode] = (scores, pickle.load(f))
    assert res["off"][1] == res["auto"][1]
    for a, b in zip(res["off"][0], res["auto"][0]):
        assert a.shape[1] == 3 and np.array_equal(a, b)


def test_prompt_larger_than_a_micro_batch_is_rejected(tiny):
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    args = make_args(tiny["layer_dir"], max_rows=64)
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    with pytest.raises(ValueError, match="raise max_rows"):
        model(copy.deepcopy(tiny["prompts"]))


def test_cli_round_trip(tiny, tmp_path):
    """main.py: pickle in -> pickles out, two generated tokens (reference main.py:55-56,65-98)."""
    from flexible_llm_sharding_b200.main import main
    pp = tmp_path / "prompts.pkl"
    with open(pp, "wb") as f:
        pickle.dump(tiny["prompts"][:3], f)
    outp = tmp_path / "scores.pkl"
    main(["--model_path", tiny["layer_dir"], "--prompt_pickle", str(pp), "--output_file", str(outp),
          "--storage_location", "gpu", "--layer_num_per_shard", "2", "--num_gen_token", "2"])
    with open(outp, "rb") as f:
        scores = pickle.load(f)
    with open(str(pp).replace(".pkl", "_updated.pkl"), "rb") as f:
        updated = pickle.load(f)
    V = tiny["shape"].vocab_size
    assert len(scores) == 3 and scores[0].shape == (4, 2, V) and scores[0].dtype == np.float16
    want = tiny["golden"]["probs32_ref"][:3]
    got0 = np.stack([s[:, 0] for s in scores]).astype(np.float64)
    m = want >= P_MIN
    assert np.abs(np.log(got0[m]) - np.log(want[m].astype(np.float64))).max() <= LOGP_TOL
    for (pre, suf), (pre0, suf0) in zip(updated, tiny["prompts"][:3]):
        assert pre == pre0 and all(s.startswith(s0) and len(s) > len(s0) for s, s0 in zip(suf, suf0))


def test_unexpected_tensors_and_max_rows_env(tiny, tmp_path, monkeypatch):
    """A layer file with a tensor the kernels would ignore (a bias) is refused at registration; FLS_MAX_ROWS
    raises the micro-batch limit for CLI users."""
    from safetensors.torch import load_file, save_file
    from flexible_llm_sharding_b200 import _lib
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    d = tmp_path / "with_bias"
    shutil_copy = __import__("shutil").copytree
    shutil_copy(tiny["layer_dir"], d)
    sd = load_file(str(d / "model.layers.1.safetensors"))
    sd["model.layers.1.self_attn.q_proj.bias"] = torch.zeros(sd["model.layers.1.self_attn.q_proj.weight"].shape[0],
                                                             dtype=torch.float16)
    save_file(sd, str(d / "model.layers.1.safetensors"))
    args = make_args(str(d))
    with pytest.raises(_lib.FlsError, match="unexpected tensor"):
        ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    monkeypatch.setenv("FLS_MAX_ROWS", "96")
    args = make_args(tiny["layer_dir"])
    model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    assert model.max_rows == 96


def test_layer_files_are_locked_in_place_not_copied(tiny):
    """SURVEY.md 8f-2 / DP read-once sharing: F16 layer files are mmap'ed and their own pages page-locked
    (cudaHostRegister), so a model costs its size in host RAM once per BOX, not once per process plus the page
    cache.  Same scores as the copying path (which BF16/F32 files still take)."""
    import gc
    from flexible_llm_sharding_b200 import _lib
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    import shutil
    import tempfile
    if not os.path.isdir("/dev/shm"):
        pytest.skip("no tmpfs to put the layer files on")
    lib = _lib.load()
    gc.collect()
    lib.fls_trim_host_cache(0)
    outs = {}
    shm_dir = os.path.join(tempfile.mkdtemp(prefix="fls_test_map_", dir="/dev/shm"), "layers")
    shutil.copytree(tiny["layer_dir"], shm_dir)
    try:
        # files on a disk file system cannot be locked in place: "map or fail" must say so, "auto" must copy
        assert lib.fls_set_host_image_mode(2) == 0
        if os.stat(tiny["layer_dir"]).st_dev != os.stat(shm_dir).st_dev:
            rc = lib.fls_preload_layer_file(os.path.join(tiny["layer_dir"], "model.norm.safetensors").encode())
            assert rc == -3 and b"tmpfs" in lib.fls_last_error(None)
        for mode, name in ((2, "map"), (1, "copy")):
            assert lib.fls_set_host_image_mode(mode) == 0
            args = make_args(shm_dir)
            model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
            before = model.ctx.stats()
            outs[name] = model(copy.deepcopy(tiny["prompts"]))
            st = model.ctx.stats()
            grew = st.mapped_bytes - before.mapped_bytes
            if name == "map":
                assert grew >= sum(model._layer_bytes) * 0.9      # every layer file, locked where it lies
                assert st.pinned_bytes - before.pinned_bytes == grew
            else:
                assert grew == 0 and st.pinned_bytes > before.pinned_bytes
            model.ctx.close()
            del model
            gc.collect()
            lib.fls_trim_host_cache(0)
    finally:
        lib.fls_set_host_image_mode(0)
        shutil.rmtree(os.path.dirname(shm_dir), ignore_errors=True)
    _, base = run_model(tiny)                                  # auto mode on the disk copy of the same files
    for a, b, c in zip(outs["map"], outs["copy"], base):
        assert np.array_equal(a, b) and np.array_equal(a, c)


@pytest.mark.parametrize("where", ["disk", "tmpfs"])
def test_packed_layer_files_give_identical_scores(tiny_gqa, tmp_path, where):
    """a directory that also holds `<layer>.flspack` files scores bit-identically, reads exactly the
    packed bytes (one contiguous range per layer) and needs fewer, larger weight copies."""
    import shutil
    import tempfile
    from flexible_llm_sharding_b200 import _lib, prepare_weights as PW
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    if where == "tmpfs" and not os.path.isdir("/dev/shm"):
        pytest.skip("no tmpfs")
    base_dir = tempfile.mkdtemp(prefix="fls_test_pack_", dir="/dev/shm") if where == "tmpfs" else str(tmp_path)
    d = os.path.join(base_dir, "layers")
    shutil.copytree(tiny_gqa["layer_dir"], d)
    try:
        args = make_args(d, layer_num_per_shard=2)
        plain = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
        want = plain(copy.deepcopy(tiny_gqa["prompts"]))
        copies_plain = plain.ctx.stats().h2d_copies
        assert all(f.endswith(".safetensors") for f in plain.layer_files)
        plain.close()
        PW.pack_layer_dir(d)
        model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
        assert all(f.endswith(".flspack") for f in model.layer_files)
        got = model(copy.deepcopy(tiny_gqa["prompts"]))
        st = model.ctx.stats()
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
        assert st.h2d_copies < copies_plain                    # 9 tensors per decoder layer became one range
        if where == "tmpfs":
            assert st.mapped_bytes > 0                         # locked in place, like the safetensors files
        model.close()
    finally:
        if where == "tmpfs":
            shutil.rmtree(base_dir, ignore_errors=True)


def test_disk_resident_layers_are_staged_once_and_locked_in_place(tiny, tmp_path):
    """Layer files on a disk file system cannot be page-locked in place; with a staging directory on tmpfs they are
    copied there ONCE (the copy every process of the box then maps), and a later load finds the staged copy."""
    import gc
    import shutil
    import tempfile
    from flexible_llm_sharding_b200 import _lib
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama, _on_tmpfs
    if not os.path.isdir("/dev/shm") or _on_tmpfs(tiny["layer_dir"]):
        pytest.skip("needs a disk-resident layer directory and a tmpfs")
    lib = _lib.load()
    stage = tempfile.mkdtemp(prefix="fls_test_stage_", dir="/dev/shm")
    gc.collect()
    lib.fls_trim_host_cache(0)
    first, base = run_model(tiny)                              # no staging: private pinned copies
    first.close()                                              # (its weight slots hold on to the last images)
    del first
    gc.collect()
    lib.fls_trim_host_cache(0)
    try:
        assert lib.fls_set_stage_dir(stage.encode()) == 0
        args = make_args(tiny["layer_dir"])
        model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
        s0 = model.ctx.stats()
        out = model(copy.deepcopy(tiny["prompts"]))
        s1 = model.ctx.stats()
        assert s1.mapped_bytes - s0.mapped_bytes >= 0.9 * sum(model._layer_bytes)
        staged = [f for f in os.listdir(stage) if f.endswith(".safetensors")]
        assert len(staged) == len(model.layer_names) and not [f for f in os.listdir(stage) if f.endswith((".tmp", ".lock"))]
        for a, b in zip(out, base):
            assert np.array_equal(a, b)
        model.close()
        del model
        gc.collect()
        lib.fls_trim_host_cache(0)
        model = ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
        r0 = model.ctx.stats().file_bytes_read
        out2 = model(copy.deepcopy(tiny["prompts"]))
        assert model.ctx.stats().file_bytes_read == r0         # the staged copies were found, nothing read again
        for a, b in zip(out2, base):
            assert np.array_equal(a, b)
        model.close()
    finally:
        lib.fls_set_stage_dir(None)
        gc.collect()
        lib.fls_trim_host_cache(0)
        shutil.rmtree(stage, ignore_errors=True)
