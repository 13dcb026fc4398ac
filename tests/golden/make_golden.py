"""Regenerates tests/golden/*.npz.  Run in the build container (needs /root/reference for the
reference's own prepare_weights.py; falls back to this repo's splitter if it is absent):

    python tests/golden/make_golden.py

For each tiny model (MHA and GQA) it
  1. writes a random-init legacy HF checkpoint (flexible-llm-sharding_b200/synthetic.py),
  2. splits it into per-layer safetensors with the REFERENCE's prepare_weights.py, unmodified,
  3. scores seeded synthetic prompts (ragged option lengths) with the real `transformers`
     modules driven through the reference's call convention (oracle/hf_reference_loop.py),
     in fp32 and fp16, with the reference-faithful unmasked prefix and with a causal prefix,
  4. stores: a sha256 over the layer files (so tests know the regenerated weights are
     bit-identical), full-vocab fp32 probabilities of the fp32 run, fp16 probabilities of the
     fp16 run, and hidden states after the embedding and after decoder layers 1 and L for
     prompt 0.
The vectors are small (a few hundred kB) and are the known answers oracle/llama_oracle.py
and the CUDA path are checked against.
"""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from flexible_llm_sharding_b200 import synthetic  # noqa: E402
from oracle.hf_reference_loop import hf_reference_scores  # noqa: E402

GOLDEN_DIR = os.path.dirname(os.path.abspath(__file__))
# tiny config of BASELINE.json configs[0], scaled down in prompt count for the fixture
CASES = {
    "tiny": dict(shape="tiny", n_prompts=6, prefix_tokens=48, n_options=4, option_tokens=12),
    "tiny-gqa": dict(shape="tiny-gqa", n_prompts=4, prefix_tokens=40, n_options=3, option_tokens=10),
}
INIT_STD = 0.05
SEED = 0
PROMPT_SEED = 1234


def build_model_dir(case: dict, out_dir: str) -> str:
    """HF .bin checkpoint -> layer files via the reference's prepare_weights.py."""
    shape = synthetic.LlamaShape.named(case["shape"])
    bin_dir = os.path.join(out_dir, "hf_bin")
    layer_dir = os.path.join(out_dir, "layers")
    synthetic.write_hf_bin_checkpoint(bin_dir, shape, seed=SEED, init_std=INIT_STD)
    ref = "/root/reference/prepare_weights.py"
    if os.path.exists(ref):
        subprocess.run([sys.executable, ref, bin_dir, layer_dir], check=True, stdout=subprocess.DEVNULL,
                       stderr=subprocess.DEVNULL)
    else:
        from flexible_llm_sharding_b200.prepare_weights import split_into_layers
        split_into_layers(bin_dir, layer_dir)
    return layer_dir


def layer_dir_digest(layer_dir: str) -> str:
    h = hashlib.sha256()
    for fn in sorted(os.listdir(layer_dir)):
        if fn.endswith(".safetensors"):
            h.update(fn.encode())
            with open(os.path.join(layer_dir, fn), "rb") as f:
                h.update(f.read())
    return h.hexdigest()


def main():
    from transformers import AutoTokenizer
    torch.set_num_threads(os.cpu_count())
    for name, case in CASES.items():
        with tempfile.TemporaryDirectory() as td:
            layer_dir = build_model_dir(case, td)
            shape = synthetic.LlamaShape.named(case["shape"])
            prompts = synthetic.make_prompts(case["n_prompts"], shape.vocab_size, case["prefix_tokens"],
                                             case["n_options"], case["option_tokens"], seed=PROMPT_SEED,
                                             ragged=True)
            tok = AutoTokenizer.from_pretrained(layer_dir)
            L = shape.num_hidden_layers
            out = {"digest": np.array(layer_dir_digest(layer_dir))}
            for causal in (False, True):
                tag = "causal" if causal else "ref"
                s32, tr = hf_reference_scores(layer_dir, prompts, tok, torch.float32, causal,
                                              trace_layers=(0, 1, L))
                s16, _ = hf_reference_scores(layer_dir, prompts, tok, torch.float16, causal)
                out[f"probs32_{tag}"] = np.stack([s[:, 0] for s in s32]).astype(np.float32)
                out[f"probs16_{tag}"] = np.stack([s[:, 0] for s in s16]).astype(np.float16)
                for li in (0, 1, L):
                    p, s = tr[li][0]
                    out[f"h{li}_prefix_{tag}"] = p.numpy().astype(np.float32)
                    out[f"h{li}_suffix_{tag}"] = s.numpy().astype(np.float32)
            np.savez_compressed(os.path.join(GOLDEN_DIR, f"{name}.npz"), **out)
            print(name, {k: v.shape for k, v in out.items()})
            # edge shapes (empty option / empty prefix / one option / very ragged / longer than a tile): the number
            # of options differs per prompt, so the probabilities are stored per prompt
            edge = synthetic.make_edge_prompts(shape.vocab_size)
            eout = {"digest": out["digest"]}
            for causal in (False, True):
                tag = "causal" if causal else "ref"
                s32, _ = hf_reference_scores(layer_dir, edge, tok, torch.float32, causal)
                s16, _ = hf_reference_scores(layer_dir, edge, tok, torch.float16, causal)
                for i, (a, b) in enumerate(zip(s32, s16)):
                    eout[f"probs32_{tag}_{i}"] = a[:, 0].astype(np.float32)
                    eout[f"probs16_{tag}_{i}"] = b[:, 0].astype(np.float16)
            np.savez_compressed(os.path.join(GOLDEN_DIR, f"{name}-edge.npz"), **eout)
            print(name + "-edge", len(edge), "prompts")


if __name__ == "__main__":
    main()
