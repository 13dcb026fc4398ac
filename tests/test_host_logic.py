"""CPU-only checks of the host side: plans against the oracle's restatement of the reference
(which uses numpy's own array_split), row packing invariants, and the C-ABI library's exports."""
import ctypes
import os
import re

import numpy as np
import pytest

from flexible_llm_sharding_b200 import planner
from oracle import llama_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_total", [7, 35, 43, 83])
@pytest.mark.parametrize("k", [1, 2, 3, 4, 10])
def test_shard_plan_matches_reference_semantics(n_total, k):
    assert planner.shard_plan(n_total, k, True) == O.shard_plan(n_total, k, True)
    for G in (1, 2, 4, 8):
        seen = []
        for r in range(G):
            mine = planner.shard_plan(n_total, k, False, G, r)
            assert mine == O.shard_plan(n_total, k, False, G, r)
            seen += [i for s in mine for i in s]
        assert sorted(seen) == list(range(n_total))        # every layer owned exactly once


def test_batch_and_partition_match_numpy():
    for n in (0, 1, 5, 32, 1024):
        for b in (1, 2, 3, 7):
            assert planner.batch_ranges(n, b) == O.batch_ranges(n, b)
        for w in (1, 2, 3, 8):
            parts = planner.partition_prompts(n, w)
            want = np.array_split(np.arange(n), w)
            assert [hi - lo for lo, hi in parts] == [len(x) for x in want]
            assert all(lo == (x[0] if len(x) else lo) for (lo, hi), x in zip(parts, want))


def _tok(shapes, seed=0):
    rng = np.random.default_rng(seed)
    out = []
    for lp, S, ls in shapes:
        eos = rng.integers(-1, ls, S)
        out.append(planner.TokenizedPrompt(rng.integers(3, 100, lp).astype(np.int32),
                                           rng.integers(3, 100, (S, ls)).astype(np.int32), eos.astype(np.int64)))
    return out


def test_pack_prompts_layout_and_tiles():
    shapes = [(70, 3, 9), (1, 2, 5), (129, 4, 33), (64, 1, 64), (200, 2, 70), (30, 2, 150)]
    tok = _tok(shapes)
    packed = planner.pack_prompts(tok, max_rows=400, max_head_rows=6, prefix_causal=False)
    assert packed.total_rows == sum(lp + S * ls for lp, S, ls in shapes)
    assert packed.total_head_rows == sum(S for _, S, _ in shapes)
    # micro-batches tile the prompt list in order and respect both limits
    assert [mb.prompt_lo for mb in packed.micro_batches][0] == 0
    assert packed.micro_batches[-1].prompt_hi == len(shapes)
    for a, b in zip(packed.micro_batches, packed.micro_batches[1:]):
        assert a.prompt_hi == b.prompt_lo and a.row_lo + a.n_rows == b.row_lo
    covered = np.zeros(packed.total_rows, dtype=int)
    for mb in packed.micro_batches:
        assert mb.n_rows <= 400 and mb.n_head <= 6
        for q_row, n_q, q_seq0, mode, s0r, s0l, s1r, s1l in mb.tiles:
            assert 1 <= n_q <= planner.ATTN_TILE_Q
            covered[mb.row_lo + q_row: mb.row_lo + q_row + n_q] += 1
            if mode == planner.ATTN_SUFFIX:
                L = q_seq0                         # option length; whole options start at s1r
                assert (q_row - s1r) >= 0 and s1r >= s0r + s0l and (s1r - (s0r + s0l)) % L == 0
                assert s1l >= (q_row - s1r) + n_q  # every query's own position is among the visible rows
                for i in (0, n_q - 1):             # rotary position == len_p + position inside the option
                    assert packed.row_pos[mb.row_lo + q_row + i] == s0l + ((q_row - s1r) + i) % L
            else:
                assert mode == planner.ATTN_PREFIX and s0r <= q_row < s0r + s0l and s1l == 0
                assert packed.row_pos[mb.row_lo + q_row] == q_seq0
        for j, g in enumerate(mb.gather):
            assert 0 <= g < mb.n_rows
    assert (covered == 1).all()                              # every row is a query exactly once
    # gather picks suffix[arange(S), eos] with python negative-index wrap (utils.py:286)
    mb0 = packed.micro_batches[0]
    lp, S, ls = shapes[0]
    for s in range(S):
        e = int(tok[0].suffix_eos[s]) % ls
        assert mb0.gather[s] == lp + s * ls + e
    with pytest.raises(ValueError):
        planner.pack_prompts(tok, max_rows=100, max_head_rows=6)
    with pytest.raises(ValueError):
        planner.pack_prompts(_tok([(4000, 1, 200)]), max_rows=10000, max_head_rows=6)     # > 4096 (utils.py:276)


def test_cached_prefix_packing_is_the_normal_packing_remapped():
    """Prefix K/V cache: a pass over the option rows only must use the same micro-batches, rotary positions,
    gathers and option tiles as the full pass, with rows renumbered to [cached prefix rows | option rows]."""
    shapes = [(70, 3, 9), (1, 2, 5), (129, 4, 33), (64, 1, 64), (200, 2, 70), (30, 2, 150)]
    tok = _tok(shapes)
    full = planner.pack_prompts(tok, max_rows=400, max_head_rows=6)
    opt = planner.pack_prompts(tok, max_rows=400, max_head_rows=6, cached_prefix=True)
    assert opt.cached_prefix and not full.cached_prefix
    assert opt.prefix_len == [lp for lp, _, _ in shapes] and opt.shapes == [(0, S, ls) for _, S, ls in shapes]
    assert opt.total_rows == sum(S * ls for _, S, ls in shapes) and len(opt.micro_batches) == len(full.micro_batches)
    kv_seen = 0
    for f, o in zip(full.micro_batches, opt.micro_batches):
        assert (f.prompt_lo, f.prompt_hi, f.head_lo, f.n_head) == (o.prompt_lo, o.prompt_hi, o.head_lo, o.n_head)
        assert (f.kv_lo, f.n_kv) == (o.kv_lo, o.n_kv) == (kv_seen, sum(lp for lp, _, _ in shapes[f.prompt_lo:f.prompt_hi]))
        kv_seen += f.n_kv
        assert o.kv_rows is None and o.n_kv + o.n_rows == f.n_rows <= 400
        # row map of the full micro-batch -> workspace rows of the option-only pass
        remap = np.full(f.n_rows, -1)
        kv, x = 0, 0
        for i in range(f.prompt_lo, f.prompt_hi):
            lp, S, ls = shapes[i]
            b = int(full.prompt_row[i]) - f.row_lo
            remap[b:b + lp] = kv + np.arange(lp)
            remap[b + lp:b + lp + S * ls] = o.n_kv + x + np.arange(S * ls)
            assert np.array_equal(f.kv_rows[kv:kv + lp], b + np.arange(lp))
            assert np.array_equal(opt.ids[o.row_lo + x:o.row_lo + x + S * ls], full.ids[f.row_lo + b + lp:f.row_lo + b + lp + S * ls])
            assert np.array_equal(opt.row_pos[o.row_lo + x:o.row_lo + x + S * ls],
                                  full.row_pos[f.row_lo + b + lp:f.row_lo + b + lp + S * ls])
            kv, x = kv + lp, x + S * ls
        assert len(f.kv_rows) == f.n_kv
        want = []
        for q_row, n_q, q_seq0, mode, s0r, s0l, s1r, s1l in f.tiles:
            if mode == planner.ATTN_SUFFIX:
                want.append([remap[q_row], n_q, q_seq0, mode, remap[s0r], s0l, remap[s1r], s1l])
        assert np.array_equal(o.tiles, np.asarray(want, dtype=np.int32))
        assert (o.tiles[:, 0] >= o.n_kv).all()                     # queries are never cached rows
        assert np.array_equal(o.gather, remap[f.gather] - o.n_kv)  # gathers index the option-only rows


def test_prefix_signature_depends_on_the_prefixes_only():
    """The prefix K/V cache is keyed by the prefixes' token ids in order: growing options (the generation loop)
    keep the key, any change to a prefix, to the order or to the number of prompts changes it."""
    from flexible_llm_sharding_b200.utils import ShardedLlama
    sig = ShardedLlama._kv_signature
    shapes = [(70, 3, 9), (12, 2, 5), (129, 4, 33)]
    a = _tok(shapes, seed=1)
    b = _tok([(lp, S, ls + 1) for lp, S, ls in shapes], seed=2)       # other options ...
    for x, y in zip(a, b):
        y.prefix_ids = x.prefix_ids.copy()                             # ... same prefixes
    assert sig(a) == sig(b)
    c = _tok(shapes, seed=1)
    c[1].prefix_ids[3] += 1
    assert sig(c) != sig(a)
    assert sig(a[::-1]) != sig(a) and sig(a[:2]) != sig(a)
    # lengths are part of the key: moving a token from one prefix to the next must not collide
    d = _tok(shapes, seed=1)
    d[0].prefix_ids, d[1].prefix_ids = (np.concatenate([a[0].prefix_ids, a[1].prefix_ids[:1]]), a[1].prefix_ids[1:])
    assert sig(d) != sig(a)


def test_rope_table_matches_oracle():
    import torch
    cos, sin = planner.rope_tables(300, 128)
    c2, s2 = O.rope_cos_sin(torch.arange(300), 128, 10000.0, torch.float16)
    assert cos.dtype == torch.float16 and cos.shape == (300, 128) and cos.is_contiguous() and sin.is_contiguous()
    # same formula; two evaluations of the vectorised fp32 cos/sin may differ by an ulp on some hosts, which can
    # move a rare entry by one fp16 ulp
    for a, b in ((cos, c2), (sin, s2)):
        assert (a.float() - b.float()).abs().max().item() <= 2 ** -10 and (a != b).float().mean().item() <= 1e-3
    big_c, big_s = planner.rope_tables(4096, 128)                 # block-wise evaluation: a prefix of a longer table
    assert torch.equal(big_c[:300], cos) and torch.equal(big_s[:300], sin)


def test_tokenize_matches_oracle(tiny):
    from transformers import AutoTokenizer
    tok = AutoTokenizer.from_pretrained(tiny["layer_dir"])
    tok.pad_token = tok.eos_token
    tok.padding_side = "right"
    prompts = tiny["prompts"] + tiny["edge_prompts"]          # incl. empty option, empty prefix, one option
    got = planner.tokenize_prompts(tok, prompts)
    for g, (pre, suf) in zip(got, prompts):
        p, s, e = O.tokenize_prompt(tok, pre, suf)
        assert g.prefix_ids.tolist() == p[0].tolist() and g.suffix_ids.tolist() == s.tolist()
        assert g.suffix_eos.tolist() == e.tolist()
    # the edge prompts pack: a 150-token option spans two attention tiles, a 140-token prefix two query tiles
    packed = planner.pack_prompts(got[len(tiny["prompts"]):], max_rows=400, max_head_rows=64)
    assert packed.total_head_rows == sum(len(o) for _, o in tiny["edge_prompts"])


def test_batched_tokenisation_equals_per_prompt_tokenisation(tiny):
    from transformers import AutoTokenizer
    from flexible_llm_sharding_b200 import synthetic
    tok = AutoTokenizer.from_pretrained(tiny["layer_dir"])
    tok.pad_token = tok.eos_token
    tok.padding_side = "right"
    prompts = synthetic.make_prompts(9, 1000, 30, 3, 11, seed=3, ragged=True) + [("w5", ("w6", "w7 w8 w9", ""))]
    a = planner.tokenize_prompts(tok, prompts)
    b = planner.tokenize_prompts_serial(tok, prompts)
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert x.prefix_ids.tolist() == y.prefix_ids.tolist()
        assert x.suffix_ids.tolist() == y.suffix_ids.tolist() and x.suffix_ids.dtype == np.int32
        assert x.suffix_eos.tolist() == y.suffix_eos.tolist()
    assert a[-1].suffix_eos.tolist() == [0, 2, -1]          # an empty option: index -1, wraps like the reference


def test_cabi_library_exports_every_declared_symbol():
    """No compute: the library must load on a CPU box and export exactly what the header declares."""
    import __graft_entry__ as ge
    from flexible_llm_sharding_b200 import _lib
    ge.build()
    header = open(os.path.join(ROOT, "include", "fls_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = sorted(set(re.findall(r"\b(fls_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/fls_b200.h but not exported"
    assert sorted(_lib.exported_symbols()) == declared       # the ctypes binding covers the whole header
    assert lib.fls_version() == 1
    typed = _lib.load()
    assert typed.fls_preload_layer_file(b"/nonexistent.safetensors") == -3
    assert b"cannot open" in typed.fls_last_error(None)


def test_safetensors_header_parse_rejects_garbage(tmp_path):
    from flexible_llm_sharding_b200 import _lib
    bad = tmp_path / "bad.safetensors"
    bad.write_bytes(b"\x10\x00\x00\x00\x00\x00\x00\x00{not json.......")
    lib = _lib.load()
    assert lib.fls_preload_layer_file(str(bad).encode()) == -3


def test_disk_tier_reads_reference_npy_files_in_place(tmp_path):
    """The disk tier reads `.npy` payloads straight into rows of a (pinned) buffer: files written by plain
    `np.save` (what the reference writes, utils.py:171-177) must land bit-exactly, and a wrong shape must raise."""
    from flexible_llm_sharding_b200.activation_store import ActivationStore
    rng = np.random.default_rng(0)
    H = 64
    pre = rng.standard_normal((1, 37, H)).astype(np.float16)
    suf = rng.standard_normal((3, 5, H)).astype(np.float16)
    np.save(tmp_path / "prefix-00000.npy", pre)
    np.save(tmp_path / "suffix-00000.npy", suf)
    rows = np.zeros((60, H), np.float16)
    ActivationStore._read_npy_into(str(tmp_path / "prefix-00000.npy"), rows[2:39], (1, 37, H))
    ActivationStore._read_npy_into(str(tmp_path / "suffix-00000.npy"), rows[39:54], (3, 5, H))
    assert np.array_equal(rows[2:39], pre[0]) and np.array_equal(rows[39:54], suf.reshape(15, H))
    assert not rows[:2].any() and not rows[54:].any()
    with pytest.raises(ValueError):
        ActivationStore._read_npy_into(str(tmp_path / "suffix-00000.npy"), rows[39:54], (5, 3, H))
    np.save(tmp_path / "f32.npy", suf.astype(np.float32))
    with pytest.raises(ValueError):
        ActivationStore._read_npy_into(str(tmp_path / "f32.npy"), rows[39:54], (3, 5, H))
    with open(tmp_path / "short.npy", "wb") as f:
        f.write(open(tmp_path / "suffix-00000.npy", "rb").read()[:-10])
    with pytest.raises(IOError):
        ActivationStore._read_npy_into(str(tmp_path / "short.npy"), rows[39:54], (3, 5, H))


def test_product_path_fails_loudly_without_its_cuda_library(tiny, monkeypatch, tmp_path):
    """No CPU fallback anywhere in the product: a missing shared object, a missing CUDA device and an import of
    the test-only oracle from the package are all errors."""
    import torch
    from flexible_llm_sharding_b200 import _lib
    from flexible_llm_sharding_b200.utils import DeviceManager, ShardedLlama
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libfls_b200.so"))
    with pytest.raises(FileNotFoundError, match="no non-CUDA fallback"):
        _lib.load()
    monkeypatch.undo()
    if not torch.cuda.is_available():
        import types
        args = types.SimpleNamespace(model_path=tiny["layer_dir"], layer_num_per_shard=1, storage_location="gpu",
                                     data_parallel=False, num_batch=1, max_activation_in_cpu=100, disk_folder="./temp",
                                     num_gen_token=1)
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            ShardedLlama(args, DeviceManager(args, ["cuda:0"]), device="cuda:0")
    # nothing under the package imports the oracle (test infrastructure only)
    pkg = os.path.dirname(_lib.__file__)
    for root, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith(".py"):
                src = open(os.path.join(root, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), fn


def test_config_fields_that_change_the_arithmetic_are_refused():
    """the reference honours every config field through AutoModelForCausalLM.from_config (utils.py:112);
    this build implements Llama-2 only and must say so instead of scoring wrongly."""
    from flexible_llm_sharding_b200 import synthetic
    from flexible_llm_sharding_b200.utils import ShardedLlama
    ok = synthetic.LlamaShape.named("tiny").hf_config()
    ShardedLlama._check_config(ok)
    ShardedLlama._check_config(dict(ok, rope_scaling=None))
    for bad in (dict(rope_scaling={"rope_type": "llama3", "factor": 8.0}), dict(hidden_act="gelu"),
                dict(attention_bias=True), dict(mlp_bias=True), dict(model_type="mistral"),
                dict(tie_word_embeddings=True), dict(pretraining_tp=2)):
        with pytest.raises(ValueError, match="unsupported model configuration"):
            ShardedLlama._check_config(dict(ok, **bad))


def test_disk_tier_file_tag_is_the_global_rank_under_torchrun(monkeypatch):
    """with one process per GPU every process has device index 0; the files of the disk tier must
    still be kept apart in a shared --disk_folder (reference names: utils.py:172,176)."""
    import types
    from flexible_llm_sharding_b200.utils import ShardedLlama
    dp = types.SimpleNamespace(args=types.SimpleNamespace(data_parallel=True))
    mp = types.SimpleNamespace(args=types.SimpleNamespace(data_parallel=False))
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    assert ShardedLlama._rank_tag(dp, 3) == "3" and ShardedLlama._rank_tag(mp, 3) == ""
    monkeypatch.setenv("WORLD_SIZE", "4")
    monkeypatch.setenv("RANK", "2")
    assert ShardedLlama._rank_tag(dp, 0) == "2" and ShardedLlama._rank_tag(mp, 0) == "2"


def test_raw_safetensors_writer_is_byte_identical_to_save_file(tmp_path):
    """The benchmark-sized models are written with a streaming writer; same bytes as safetensors' own save_file
    (the call prepare_weights.py makes, reference prepare_weights.py:40-43)."""
    from flexible_llm_sharding_b200 import synthetic
    shape = synthetic.LlamaShape.named("tiny", num_hidden_layers=2)
    a, b = tmp_path / "a", tmp_path / "b"
    synthetic.write_layer_files(str(a), shape)
    synthetic.write_layer_files(str(b), shape, fast=True, only_layers={"model.layers.1", "lm_head"})
    assert sorted(f for f in os.listdir(b) if f.endswith(".safetensors")) == ["lm_head.safetensors",
                                                                             "model.layers.1.safetensors"]
    for fn in ("lm_head.safetensors", "model.layers.1.safetensors"):
        assert (a / fn).read_bytes() == (b / fn).read_bytes()


def test_packed_layer_files_keep_the_reference_layout_and_slot_order(tmp_path):
    """prepare_weights.py --packed: the reference layout is still written byte for byte; the
    packed twin holds the same tensors as fp16, 256-byte aligned, in HBM slot order, in whole 4096-byte pages."""
    import torch
    from safetensors.torch import load_file
    from flexible_llm_sharding_b200 import prepare_weights as PW, synthetic
    shape = synthetic.LlamaShape.named("tiny-gqa", num_hidden_layers=2)
    hf = tmp_path / "hf"
    synthetic.write_hf_bin_checkpoint(str(hf), shape, seed=0, init_std=0.05)
    plain, packed = tmp_path / "plain", tmp_path / "packed"
    PW.split_into_layers(str(hf), str(plain))
    PW.main([str(hf), str(packed), "--packed"])
    for fn in sorted(os.listdir(plain)):
        assert (plain / fn).read_bytes() == (packed / fn).read_bytes()
    for layer in shape.layer_names():
        sd = load_file(str(packed / f"{layer}.safetensors"))
        entries = PW.read_packed_header(str(packed / f"{layer}.flspack"))
        assert [e[0] for e in entries] == PW.pack_order(layer, sd)
        size = os.path.getsize(packed / f"{layer}.flspack")
        assert size % 4096 == 0
        raw = (packed / f"{layer}.flspack").read_bytes()
        prev_end = 0
        for name, rows, cols, off, nb in entries:
            assert off % 256 == 0 and off >= prev_end
            want = sd[name].to(torch.float16).contiguous().numpy().tobytes()
            assert raw[4096 + off:4096 + off + nb] == want and tuple(sd[name].shape) == ((rows, cols) if cols else (rows,))
            prev_end = off + nb
    # q|k|v and gate|up are adjacent: one GEMM operand each without repacking
    e = {n.split(".", 3)[-1]: (off, nb) for n, _, _, off, nb in PW.read_packed_header(str(packed / "model.layers.1.flspack"))}
    assert e["self_attn.k_proj.weight"][0] == sum(e["self_attn.q_proj.weight"])
    assert e["mlp.up_proj.weight"][0] == sum(e["mlp.gate_proj.weight"])
    with pytest.raises(ValueError, match="no place in the packed layout"):
        PW.pack_order("model.layers.0", list(load_file(str(packed / "model.layers.0.safetensors"))) + ["model.layers.0.self_attn.q_proj.bias"])


def _ragged_shapes(n, seed):
    rng = np.random.default_rng(seed)
    return [(int(rng.integers(1, 900)), int(rng.integers(1, 6)), int(rng.integers(1, 60))) for _ in range(n)]


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_length_bucketed_packing_is_a_balanced_permutation(seed):
    """same prompts, same rows per prompt, fewer and more even micro-batches, long prompts first."""
    shapes = _ragged_shapes(200, seed)
    tok = _tok(shapes, seed)
    plain = planner.pack_prompts(tok, max_rows=4096, max_head_rows=64)
    buck = planner.pack_prompts(tok, max_rows=4096, max_head_rows=64, bucketed=True)
    assert plain.order is None and buck.order is not None
    assert sorted(buck.order.tolist()) == list(range(len(tok)))
    assert buck.total_rows == plain.total_rows and buck.total_head_rows == plain.total_head_rows
    # every prompt's token rows, positions and shape are the ones the in-order packing holds for it
    for p, i in enumerate(buck.order.tolist()):
        assert buck.shapes[p] == plain.shapes[i] and buck.prompt_index(p) == i
        a, b = int(buck.prompt_row[p]), int(buck.prompt_row[p + 1])
        c, d = int(plain.prompt_row[i]), int(plain.prompt_row[i + 1])
        assert np.array_equal(buck.ids[a:b], plain.ids[c:d]) and np.array_equal(buck.row_pos[a:b], plain.row_pos[c:d])
        lo, hi = buck.head_range(i)
        assert (lo, hi) == (int(buck.head_row[p]), int(buck.head_row[p + 1])) and hi - lo == shapes[i][1]
    # the gather rows point at the same tokens
    def gathered(pk):
        out = {}
        for mb in pk.micro_batches:
            k = 0
            for p in range(mb.prompt_lo, mb.prompt_hi):
                S = pk.shapes[p][1]
                rows = mb.gather[k:k + S] + mb.row_lo - int(pk.prompt_row[p])
                out[pk.prompt_index(p)] = rows.tolist()
                k += S
        return out
    assert gathered(buck) == gathered(plain)
    # caps hold, no more micro-batches than in order, and they are even: the spread is at most one prompt
    sizes = [mb.n_rows for mb in buck.micro_batches]
    assert all(mb.n_rows <= 4096 and mb.n_head <= 64 for mb in buck.micro_batches)
    assert len(buck.micro_batches) <= len(plain.micro_batches)
    assert max(sizes) - min(sizes) <= max(lp + S * ls for lp, S, ls in shapes)
    plain_sizes = [mb.n_rows for mb in plain.micro_batches]
    assert max(sizes) - min(sizes) <= max(plain_sizes) - min(plain_sizes)
    for mb in buck.micro_batches:                      # longest first inside a micro-batch
        rows = [sum(x * y for x, y in [(1, buck.shapes[p][0]), (buck.shapes[p][1], buck.shapes[p][2])])
                for p in range(mb.prompt_lo, mb.prompt_hi)]
        assert rows == sorted(rows, reverse=True)
    # micro-batches still tile the packed rows
    for a, b in zip(buck.micro_batches, buck.micro_batches[1:]):
        assert a.prompt_hi == b.prompt_lo and a.row_lo + a.n_rows == b.row_lo and a.head_lo + a.n_head == b.head_lo


def test_length_bucketed_packing_leaves_uniform_prompts_in_order():
    tok = _tok([(384, 4, 32)] * 72)
    buck = planner.pack_prompts(tok, max_rows=16384, max_head_rows=4096, bucketed=True)
    assert buck.order is None                                        # nothing moved
    assert [mb.prompt_hi - mb.prompt_lo for mb in buck.micro_batches] == [24, 24, 24]   # in order it would be 32, 32, 8
    full = planner.pack_prompts(_tok([(384, 4, 32)] * 64), max_rows=16384, max_head_rows=4096, bucketed=True)
    plain = planner.pack_prompts(_tok([(384, 4, 32)] * 64), max_rows=16384, max_head_rows=4096)
    assert [(m.prompt_lo, m.prompt_hi) for m in full.micro_batches] == [(m.prompt_lo, m.prompt_hi) for m in plain.micro_batches]


def test_length_bucketed_packing_limits_and_refusals():
    # the head-row cap, not the row cap, decides the micro-batch count here
    tok = _tok([(10, 5, 2)] * 7 + [(300, 1, 3)])
    buck = planner.pack_prompts(tok, max_rows=1000, max_head_rows=10, bucketed=True)
    assert all(mb.n_head <= 10 for mb in buck.micro_batches) and len(buck.micro_batches) == 4
    with pytest.raises(ValueError, match="prompt 2 has"):
        planner.pack_prompts(_tok([(10, 1, 2), (10, 1, 2), (500, 2, 10)]), max_rows=100, max_head_rows=6, bucketed=True)
    # with a cached prefix the packing stays in the order of the pass that filled the cache
    tok = _tok(_ragged_shapes(20, 5), 5)
    assert planner.pack_prompts(tok, max_rows=4096, max_head_rows=64, cached_prefix=True, bucketed=True).order is None
    # one prompt, and no prompts
    assert planner.pack_prompts(tok[:1], max_rows=4096, max_head_rows=64, bucketed=True).order is None


def test_every_tuning_option_is_documented_in_the_public_header():
    """`fls_set_option` names accepted by csrc/api.cu all appear, quoted, in include/fls_b200.h."""
    with open(os.path.join(ROOT, "flexible-llm-sharding_b200", "csrc", "api.cu")) as f:
        accepted = set(re.findall(r'strcmp\(name, "([a-z_]+)"\)', f.read()))
    with open(os.path.join(ROOT, "include", "fls_b200.h")) as f:
        header = f.read()
    assert len(accepted) >= 10
    assert not [o for o in sorted(accepted) if f'"{o}"' not in header]


def test_length_bucketed_packing_never_makes_more_micro_batches_than_in_order():
    """Longest-first into the emptiest micro-batch is a heuristic: rows 5,5,4,3,3 under a cap of 10 would take three
    micro-batches where the given order takes two.  Such a mix is packed in order."""
    tok = _tok([(1, 1, 4), (1, 1, 4), (1, 1, 3), (1, 1, 2), (1, 1, 2)])          # rows 5, 5, 4, 3, 3
    assert [t.rows for t in tok] == [5, 5, 4, 3, 3]
    assert len(planner.balanced_bins(tok, 10, 64)) == 3
    buck = planner.pack_prompts(tok, max_rows=10, max_head_rows=64, bucketed=True)
    assert buck.order is None and [mb.n_rows for mb in buck.micro_batches] == [10, 10]
    rng = np.random.default_rng(11)
    for _ in range(300):
        shapes = [(int(rng.integers(1, 40)), int(rng.integers(1, 4)), int(rng.integers(1, 12))) for _ in range(int(rng.integers(2, 30)))]
        tok = _tok(shapes, 3)
        cap = max(t.rows for t in tok) + int(rng.integers(0, 120))
        heads = int(rng.integers(3, 20))
        plain = planner.pack_prompts(tok, max_rows=cap, max_head_rows=heads)
        buck = planner.pack_prompts(tok, max_rows=cap, max_head_rows=heads, bucketed=True)
        assert len(buck.micro_batches) <= len(plain.micro_batches)
        assert planner._in_order_count(tok, cap, heads) == len(plain.micro_batches)
        assert all(mb.n_rows <= cap and mb.n_head <= heads for mb in buck.micro_batches)
        assert sorted(buck.prompt_index(p) for p in range(len(tok))) == list(range(len(tok)))
