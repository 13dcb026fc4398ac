"""Synthetic Llama-2-architecture checkpoints, tokenizers and prompts.

There is no network for real checkpoints, so every test and benchmark runs on
random-init weights of the named architecture (BASELINE.json `configs`).  Two
writers produce the on-disk formats the reference consumes:

* :func:`write_hf_bin_checkpoint` — a legacy HuggingFace directory
  (`pytorch_model-0000X-of-0000N.bin` + `pytorch_model.bin.index.json` +
  `config.json` + tokenizer files), i.e. the *input* of the reference's
  `prepare_weights.py` (reference `prepare_weights.py:12-49`).
* :func:`write_layer_files` — the *output* layout of `prepare_weights.py`
  directly (one `<layer>.safetensors` per top-level layer, original key
  names), used for the 7B/13B/70B-sized benchmarks where a `.bin` round trip
  would only cost time.

Both draw every tensor from its own generator seeded by (seed, tensor name), so
the two routes give bit-identical weights and a layer can be produced alone.
"""
from __future__ import annotations

import json
import os
import zlib
from dataclasses import dataclass, asdict

import numpy as np
import torch

# Llama-2 shapes (dims line).
LLAMA_SHAPES = {
    "tiny": dict(num_hidden_layers=4, hidden_size=512, intermediate_size=1408,
                 num_attention_heads=4, num_key_value_heads=4, vocab_size=1000),
    "tiny-gqa": dict(num_hidden_layers=4, hidden_size=512, intermediate_size=1408,
                     num_attention_heads=4, num_key_value_heads=2, vocab_size=1000),
    "7b": dict(num_hidden_layers=32, hidden_size=4096, intermediate_size=11008,
               num_attention_heads=32, num_key_value_heads=32, vocab_size=32000),
    "13b": dict(num_hidden_layers=40, hidden_size=5120, intermediate_size=13824,
                num_attention_heads=40, num_key_value_heads=40, vocab_size=32000),
    "70b": dict(num_hidden_layers=80, hidden_size=8192, intermediate_size=28672,
                num_attention_heads=64, num_key_value_heads=8, vocab_size=32000),
}


@dataclass
class LlamaShape:
    num_hidden_layers: int
    hidden_size: int
    intermediate_size: int
    num_attention_heads: int
    num_key_value_heads: int
    vocab_size: int
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    max_position_embeddings: int = 4096

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    @classmethod
    def named(cls, name: str, **over) -> "LlamaShape":
        d = dict(LLAMA_SHAPES[name])
        d.update(over)
        return cls(**d)

    def hf_config(self) -> dict:
        d = asdict(self)
        d.update(
            architectures=["LlamaForCausalLM"], model_type="llama", hidden_act="silu",
            bos_token_id=1, eos_token_id=2, pad_token_id=None, tie_word_embeddings=False,
            torch_dtype="float16", attention_bias=False, mlp_bias=False,
            initializer_range=0.02, use_cache=True,
        )
        return d

    def layer_names(self) -> list[str]:
        # reference utils.py:106-107
        return (["model.embed_tokens"]
                + [f"model.layers.{i}" for i in range(self.num_hidden_layers)]
                + ["model.norm", "lm_head"])

    def layer_tensor_shapes(self, layer_name: str) -> dict[str, tuple[int, ...]]:
        H, I, V = self.hidden_size, self.intermediate_size, self.vocab_size
        d = self.head_dim
        nq, nkv = self.num_attention_heads * d, self.num_key_value_heads * d
        if layer_name == "model.embed_tokens":
            return {"model.embed_tokens.weight": (V, H)}
        if layer_name == "model.norm":
            return {"model.norm.weight": (H,)}
        if layer_name == "lm_head":
            return {"lm_head.weight": (V, H)}
        p = layer_name
        return {
            f"{p}.self_attn.q_proj.weight": (nq, H),
            f"{p}.self_attn.k_proj.weight": (nkv, H),
            f"{p}.self_attn.v_proj.weight": (nkv, H),
            f"{p}.self_attn.o_proj.weight": (H, nq),
            f"{p}.mlp.gate_proj.weight": (I, H),
            f"{p}.mlp.up_proj.weight": (I, H),
            f"{p}.mlp.down_proj.weight": (H, I),
            f"{p}.input_layernorm.weight": (H,),
            f"{p}.post_attention_layernorm.weight": (H,),
        }


def _tensor_seed(seed: int, name: str) -> int:
    return (seed * 1000003 + zlib.crc32(name.encode())) % (2**63 - 1)


def make_tensor(name: str, shape, seed: int, init_std: float, device="cpu",
                dtype=torch.float16) -> torch.Tensor:
    """N(0, init_std) for matrices, 1 + 0.1·N(0,1) for RMSNorm gains (so the gain
    multiply is exercised).  Drawn in fp32 then cast."""
    g = torch.Generator(device=device)
    g.manual_seed(_tensor_seed(seed, name))
    if len(shape) == 1:
        t = 1.0 + 0.1 * torch.randn(shape, generator=g, device=device, dtype=torch.float32)
    else:
        t = init_std * torch.randn(shape, generator=g, device=device, dtype=torch.float32)
    return t.to(dtype)


def make_layer_state_dict(shape: LlamaShape, layer_name: str, seed=0, init_std=0.02,
                          device="cpu", dtype=torch.float16) -> dict[str, torch.Tensor]:
    return {k: make_tensor(k, s, seed, init_std, device, dtype)
            for k, s in shape.layer_tensor_shapes(layer_name).items()}


# --------------------------------------------------------------------------------------
# tokenizer + config
# --------------------------------------------------------------------------------------
def write_tokenizer(out_dir: str, vocab_size: int) -> None:
    """WordLevel tokenizer: ids 0/1/2 = <unk>/<s>/</s>, word `w{id}` for id >= 3, BOS added
    by template.  Loads through `AutoTokenizer.from_pretrained` and behaves like the Llama
    tokenizer at the reference's call sites (utils.py:246-258: BOS first, right padding with
    EOS once `pad_token = eos_token`)."""
    from tokenizers import Tokenizer, models, pre_tokenizers, processors

    vocab = {"<unk>": 0, "<s>": 1, "</s>": 2}
    for i in range(3, vocab_size):
        vocab[f"w{i}"] = i
    tok = Tokenizer(models.WordLevel(vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    tok.post_processor = processors.TemplateProcessing(single="<s> $A", special_tokens=[("<s>", 1)])
    os.makedirs(out_dir, exist_ok=True)
    tok.save(os.path.join(out_dir, "tokenizer.json"))
    with open(os.path.join(out_dir, "tokenizer_config.json"), "w") as f:
        json.dump({"tokenizer_class": "PreTrainedTokenizerFast", "bos_token": "<s>",
                   "eos_token": "</s>", "unk_token": "<unk>", "model_max_length": 1 << 20}, f)


def write_config(out_dir: str, shape: LlamaShape) -> None:
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(shape.hf_config(), f, indent=1)


# --------------------------------------------------------------------------------------
# checkpoint writers
# --------------------------------------------------------------------------------------
def write_hf_bin_checkpoint(out_dir: str, shape: LlamaShape, seed=0, init_std=0.02,
                            n_bin_shards=2) -> None:
    """Legacy HF layout = the input of prepare_weights.py (reference prepare_weights.py:18-37)."""
    write_config(out_dir, shape)
    write_tokenizer(out_dir, shape.vocab_size)
    names = shape.layer_names()
    per = -(-len(names) // n_bin_shards)
    weight_map = {}
    for s in range(n_bin_shards):
        fn = f"pytorch_model-{s + 1:05d}-of-{n_bin_shards:05d}.bin"
        sd = {}
        for ln in names[s * per:(s + 1) * per]:
            sd.update(make_layer_state_dict(shape, ln, seed, init_std))
        for k in sd:
            weight_map[k] = fn
        torch.save(sd, os.path.join(out_dir, fn))
    with open(os.path.join(out_dir, "pytorch_model.bin.index.json"), "w") as f:
        json.dump({"metadata": {}, "weight_map": weight_map}, f)


def save_safetensors_raw(path: str, tensors: dict[str, torch.Tensor]) -> None:
    """Writes the safetensors container (8-byte little-endian header length, JSON header padded with spaces to a
    multiple of 8, raw tensor bytes in header order; the layout prepare_weights.py's `save_file` produces) straight
    from the tensors' memory - no intermediate `bytes` copy of the whole file, which is what makes writing the
    138 GB of a 70B-sized model take tens of seconds instead of minutes.  fp16 CPU tensors only."""
    names = sorted(tensors)                    # one dtype -> safetensors orders by name
    header, off = {}, 0
    for n in names:
        t = tensors[n]
        assert t.dtype == torch.float16 and t.device.type == "cpu" and t.is_contiguous(), n
        nbytes = t.numel() * 2
        header[n] = {"dtype": "F16", "shape": list(t.shape), "data_offsets": [off, off + nbytes]}
        off += nbytes
    blob = json.dumps(header, separators=(",", ":")).encode()
    blob += b" " * (-len(blob) % 8)
    with open(path, "wb", buffering=0) as f:
        f.write(len(blob).to_bytes(8, "little"))
        f.write(blob)
        for n in names:
            f.write(memoryview(tensors[n].numpy()).cast("B"))


def write_layer_files(out_dir: str, shape: LlamaShape, seed=0, init_std=0.02, device="cpu",
                      alias_decoder_layers: int = 0, only_layers=None, fast: bool = False) -> None:
    """prepare_weights.py's output layout written directly (reference prepare_weights.py:40-43).

    `alias_decoder_layers=n>0` writes only n distinct decoder layers and symlinks the rest to them
    (70B-sized runs on a host whose RAM cannot hold 138 GB of pinned weights plus the files); the
    caller must say so in its results.  `only_layers` (a set of layer names) restricts the call to those files, so
    that several processes can write one model together; `fast` uses `save_safetensors_raw` with a few writer
    threads (benchmark-sized models) instead of safetensors' own `save_file`."""
    from concurrent.futures import ThreadPoolExecutor

    from safetensors.torch import save_file

    write_config(out_dir, shape)
    write_tokenizer(out_dir, shape.vocab_size)
    pool = ThreadPoolExecutor(max_workers=4) if fast else None
    pending = []
    for ln in shape.layer_names():
        if only_layers is not None and ln not in only_layers:
            continue
        dst = os.path.join(out_dir, f"{ln}.safetensors")
        if alias_decoder_layers and ln.startswith("model.layers."):
            i = int(ln.rsplit(".", 1)[1])
            if i >= alias_decoder_layers:
                # Same bytes as layer i % n: a symlink.  The tensor names inside then carry the source
                # layer's prefix; the CUDA loader matches tensors by suffix and the pinned cache is keyed
                # by the resolved path, so aliased layers share one host image.  (Not loadable by name-
                # strict readers such as the reference or the oracle - benchmark use only.)
                if os.path.lexists(dst):
                    os.remove(dst)
                os.symlink(f"model.layers.{i % alias_decoder_layers}.safetensors", dst)
                continue
        sd = make_layer_state_dict(shape, ln, seed, init_std, device)
        host = {k: v.cpu().contiguous() for k, v in sd.items()}
        if pool is None:
            save_file(host, dst)
        else:
            pending.append(pool.submit(save_safetensors_raw, dst, host))
            while len(pending) > 4:                      # bound the host copies in flight
                pending.pop(0).result()
    for f in pending:
        f.result()
    if pool is not None:
        pool.shutdown()


# --------------------------------------------------------------------------------------
# prompts (SURVEY.md §8(d) "Synthetic prompts")
# --------------------------------------------------------------------------------------
def make_prompts(n_prompts: int, vocab_size: int, prefix_tokens: int, n_options: int,
                 option_tokens: int, seed=1234, ragged=False) -> list[tuple[str, tuple[str, ...]]]:
    """`prefix_tokens` counts the BOS the tokenizer adds (so the prefix has prefix_tokens-1
    words).  With `ragged`, option lengths are U[ceil(option_tokens/2), option_tokens]."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n_prompts):
        pre = " ".join(f"w{i}" for i in rng.integers(3, vocab_size, prefix_tokens - 1))
        opts = []
        for _ in range(n_options):
            n = int(rng.integers((option_tokens + 1) // 2, option_tokens + 1)) if ragged else option_tokens
            opts.append(" ".join(f"w{i}" for i in rng.integers(3, vocab_size, n)))
        if ragged:  # make sure the padded width is `option_tokens`
            opts[0] = " ".join(f"w{i}" for i in rng.integers(3, vocab_size, option_tokens))
        out.append((pre, tuple(opts)))
    return out


def make_edge_prompts(vocab_size: int, seed=4321) -> list[tuple[str, tuple[str, ...]]]:
    """Prompt shapes the reference's loop accepts (utils.py:246-258,284-286) that `make_prompts` never produces."""
    rng = np.random.default_rng(seed)

    def words(n):
        return " ".join(f"w{i}" for i in rng.integers(3, vocab_size, n))

    return [
        (words(20), (words(6), "", words(3))),        # an empty option: all pad, suffix_eos = -1 wraps to the last pad
        (words(15), (words(9),)),                     # a single option
        ("", (words(5), words(2))),                   # an empty prefix: BOS only
        (words(33), (words(1), words(24), words(12), words(24), words(2))),      # five very ragged options
        (words(40), tuple(words(7) for _ in range(3))),                          # equal lengths: no padding at all
        (words(25), (words(150), words(131))),        # options longer than one 128-row attention tile
        (words(140), (words(4), words(4))),           # a prefix longer than one query tile
    ]
