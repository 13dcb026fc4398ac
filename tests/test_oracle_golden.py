"""Pins oracle/llama_oracle.py to the committed golden vectors (tests/golden/*.npz), which were
produced by the real `transformers` Llama modules driven through the reference's call
convention (oracle/hf_reference_loop.py), on layer files split by the reference's own
prepare_weights.py (tests/golden/make_golden.py).  CPU only."""
import copy
import os

import numpy as np
import pytest
import torch

from oracle import llama_oracle as O

# fp32 restatement vs fp32 HF modules: same ops in the same order -> differences are only
# matmul blocking / summation order.
FP32_PROB_RTOL = 2e-4
FP32_HIDDEN_ATOL = 2e-4


def test_splitter_is_byte_identical_to_reference(tiny, tiny_gqa):
    # golden digest = sha256 over the files the REFERENCE's prepare_weights.py wrote
    for case in (tiny, tiny_gqa):
        assert case["digest"] == str(case["golden"]["digest"])
        names = sorted(f for f in os.listdir(case["layer_dir"]))
        assert "pytorch_model.bin.index.json" not in names      # reference prepare_weights.py:15
        assert "config.json" in names and "tokenizer.json" in names
        assert len([n for n in names if n.endswith(".safetensors")]) == case["shape"].num_hidden_layers + 3


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("which", ["tiny", "tiny_gqa"])
def test_oracle_fp32_matches_hf_modules(which, causal, request):
    case = request.getfixturevalue(which)
    tag = "causal" if causal else "ref"
    L = case["shape"].num_hidden_layers
    orc = O.OracleShardedLlama(case["layer_dir"], dtype=torch.float32, layer_num_per_shard=2,
                               prefix_causal=causal)
    scores = orc(copy.deepcopy(case["prompts"]), trace_layers=(0, 1, L))
    got = np.stack([s[:, 0] for s in scores])
    want = case["golden"][f"probs32_{tag}"]
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=FP32_PROB_RTOL, atol=1e-9)
    assert (got.argmax(-1) == want.argmax(-1)).all()
    for li in (0, 1, L):
        p, s = orc.hidden_trace[li][0]
        np.testing.assert_allclose(p.numpy(), case["golden"][f"h{li}_prefix_{tag}"], atol=FP32_HIDDEN_ATOL)
        np.testing.assert_allclose(s.numpy(), case["golden"][f"h{li}_suffix_{tag}"], atol=FP32_HIDDEN_ATOL)


def test_oracle_fp16_matches_hf_modules_fp16(tiny):
    """The fp16 restatement must round where HF eager rounds: compare with the fp16 HF run."""
    orc = O.OracleShardedLlama(tiny["layer_dir"], dtype=torch.float16)
    prompts = copy.deepcopy(tiny["prompts"][:3])
    got = np.stack([s[:, 0] for s in orc(prompts)]).astype(np.float32)
    want = tiny["golden"]["probs16_ref"][:3].astype(np.float32)
    assert got.dtype == np.float32 and orc.dtype == torch.float16
    # same rounding points -> equal up to a few fp16 ulps of the probabilities
    np.testing.assert_allclose(got, want, rtol=2e-2, atol=2e-6)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("which", ["tiny", "tiny_gqa"])
def test_oracle_matches_hf_modules_on_edge_shapes(which, causal, request):
    """Empty option (all pad: `suffix_eos = -1` wraps to the last pad position, utils.py:258,286), empty prefix
    (BOS only), a single option, five very ragged options, options and a prefix longer than 128 rows."""
    case = request.getfixturevalue(which)
    tag = "causal" if causal else "ref"
    prompts = case["edge_prompts"]
    assert str(case["edge_golden"]["digest"]) == case["digest"]
    orc = O.OracleShardedLlama(case["layer_dir"], dtype=torch.float32, layer_num_per_shard=3, prefix_causal=causal)
    scores = orc(copy.deepcopy(prompts))
    assert len(scores) == len(prompts)
    for i, ((_, options), s) in enumerate(zip(prompts, scores)):
        want = case["edge_golden"][f"probs32_{tag}_{i}"]
        assert s.shape == (len(options), 1, want.shape[-1])
        np.testing.assert_allclose(s[:, 0], want, rtol=FP32_PROB_RTOL, atol=1e-9, err_msg=f"prompt {i}")
    # the empty option really has no token: index -1, i.e. the last (pad) position, is what gets scored
    assert int(orc.suffix_eos[0][1]) == -1


def test_prefix_mask_fork_matters(tiny):
    """The unmasked-prefix semantics (utils.py:272) is not a detail: it moves the scores."""
    g = tiny["golden"]
    d = np.abs(np.log(g["probs32_ref"]) - np.log(g["probs32_causal"])).max()
    assert d > 1e-2


def test_shard_plan_matches_numpy_semantics():
    # DP: ceil(n/k) shards via array_split (utils.py:145-146)
    assert O.shard_plan(7, 2, True) == [(0, 1), (2, 3), (4, 5), (6,)]
    assert O.shard_plan(35, 1, True) == [(i,) for i in range(35)]
    # MP: shard count rounded up to a multiple of num_gpu, rank r takes r::G (utils.py:151-153)
    all_r = [O.shard_plan(7, 3, False, num_gpu=2, gpu_rank=r) for r in range(2)]
    assert all_r == [[(0, 1), (4, 5)], [(2, 3), (6,)]]
    # may create empty shards
    plans = [O.shard_plan(5, 1, False, num_gpu=4, gpu_rank=r) for r in range(4)]
    assert sum(len(s) for p in plans for s in p) == 5
    assert any(len(s) == 0 for p in plans for s in p)
    assert O.batch_ranges(10, 3) == [(0, 3), (3, 6), (6, 10)]
    assert O.batch_ranges(5, 1) == [(0, 5)]


def test_tokenize_matches_reference_convention(tiny):
    from transformers import AutoTokenizer
    tok = AutoTokenizer.from_pretrained(tiny["layer_dir"])
    tok.pad_token = tok.eos_token
    tok.padding_side = "right"
    p, s, eos = O.tokenize_prompt(tok, "w3 w4 w5", ("w6 w7", "w8", "w9 w10 w11"))
    assert p.tolist() == [[1, 3, 4, 5]]                          # BOS kept on the prefix
    assert s.tolist() == [[6, 7, 2], [8, 2, 2], [9, 10, 11]]     # BOS stripped, EOS right padding
    assert eos.tolist() == [1, 0, 2]
