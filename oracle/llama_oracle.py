"""CPU oracle for the layer-sharded prompt-scoring path.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module, and only as the checker / reported CPU baseline.  The product
(`flexible-llm-sharding_b200/`) never imports it and has no CPU fallback.

What it restates
----------------
* the reference's scoring loop: `ShardedLlama.__call__`, /root/reference/utils.py:133-305
  (shard plan :143-157, mask/positions :218-221, tokenisation :246-259, the five layer
  branches :266-291, activation store/fetch :159-213 in their `cpu` form);
* the layer arithmetic the reference delegates to the third-party `transformers` package
  (unpinned in requirements.txt:2; the calling convention at utils.py:272,277-279 - tuple KV
  cache, `past_key_value=`, decoder layer returning `(hidden, (k, v))` - fixes the usable window
  to transformers 4.31-4.35).  That package is not vendored in /root/reference and the installed
  transformers 5.5.0 is API-incompatible with the reference loop, so the published HF Llama
  *eager* algorithm is restated here, op by op and with the same rounding points:
    RMSNorm      x32 = x.float(); x32 * rsqrt(mean(x32^2) + eps) -> cast to dtype -> * weight
    RoPE         inv_freq = theta^(-2i/d) (fp32); cos/sin computed in fp32, cast to dtype;
                 q*cos + rotate_half(q)*sin with the half-split convention, in dtype
    attention    scores = (q @ k^T) * d^-0.5 (+ additive mask) in dtype; softmax in fp32,
                 cast to dtype; @ v.  GQA by repeating each kv head n_rep times
    MLP          down(silu(gate(x)) * up(x)), every op rounded to dtype
    residuals    added in dtype
    head         softmax(lm_head(x)) evaluated on the dtype tensor (utils.py:289)

Parity pinning: the reference has no tests or golden vectors for this path.
The oracle is pinned instead against the *real* transformers modules (5.5.0, the version in
this image) driven through the reference's call convention - see oracle/hf_reference_loop.py
and tests/golden/make_golden.py - and the resulting vectors are committed under tests/golden/.

Semantic fork: the reference runs the prefix pass with `attention_mask=None` (utils.py:272),
which in HF eager attention means *no mask*: prefix tokens attend bidirectionally; only the
suffix pass is causal (utils.py:276).  `prefix_causal=False` is therefore the reference-faithful
default; `prefix_causal=True` gives ordinary causal scoring of `prefix + suffix_i`.
"""
from __future__ import annotations

import json
import math
import os
from collections import defaultdict

import numpy as np
import torch

MAX_TOKEN_LEN = 4096  # utils.py:14


# --------------------------------------------------------------------------------------
# host-side pieces
# --------------------------------------------------------------------------------------
def shard_plan(n_layers_total: int, layer_num_per_shard: int, data_parallel: bool,
               num_gpu: int = 1, gpu_rank: int = 0) -> list[tuple[int, ...]]:
    """utils.py:143-157.  `n_layers_total` = decoder layers + 3."""
    if data_parallel:
        num_shards = np.ceil(n_layers_total / layer_num_per_shard)
        return [tuple(int(i) for i in s) for s in np.array_split(np.arange(n_layers_total), num_shards)]
    num_shards = np.ceil(np.ceil(n_layers_total / layer_num_per_shard) / num_gpu) * num_gpu
    all_shards = np.array_split(np.arange(n_layers_total), num_shards)
    return [tuple(int(i) for i in s) for s in all_shards[gpu_rank::num_gpu]]


def batch_ranges(n_prompts: int, num_batch: int) -> list[tuple[int, int]]:
    """main.py:19-20."""
    ends = [n_prompts // num_batch * i for i in range(1, num_batch)] + [n_prompts]
    return list(zip([0] + ends[:-1], ends))


def tokenize_prompt(tokenizer, prefix: str, suffixes) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """utils.py:246-258.  Returns prefix ids [1,len_p], suffix ids [S,len_s] (BOS stripped,
    right-padded with pad=eos) and suffix_eos [S] = index of each option's last real token."""
    p = tokenizer(prefix, return_tensors="pt", return_attention_mask=False,
                  truncation=True, max_length=MAX_TOKEN_LEN)["input_ids"]
    s = tokenizer(list(suffixes), return_tensors="pt", return_attention_mask=False,
                  truncation=True, max_length=MAX_TOKEN_LEN, padding=True)["input_ids"][:, 1:]
    eos = (s != tokenizer.pad_token_id).sum(1) - 1
    return p, s, eos


# --------------------------------------------------------------------------------------
# layer arithmetic (HF Llama eager, restated)
# --------------------------------------------------------------------------------------
def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    dt = x.dtype
    x32 = x.to(torch.float32)
    var = x32.pow(2).mean(-1, keepdim=True)
    x32 = x32 * torch.rsqrt(var + eps)
    return w * x32.to(dt)


def rope_cos_sin(positions: torch.Tensor, head_dim: int, theta: float, dtype) -> tuple[torch.Tensor, torch.Tensor]:
    """positions [L] int64 -> cos, sin [L, head_dim] in `dtype` (fp32 math, then cast)."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).to(torch.float32) / head_dim))
    # evaluated on the host like the reference's rotary buffers (built at module construction on the CPU, then
    # moved and cast by utils.py:117-119), whatever device the run is on
    freqs = positions.cpu().to(torch.float32)[:, None] * inv_freq[None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype).to(positions.device), emb.sin().to(dtype).to(positions.device)


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x [B, heads, L, d]; cos/sin [L, d]."""
    return (x * cos[None, None]) + (rotate_half(x) * sin[None, None])


def repeat_kv(x: torch.Tensor, n_rep: int) -> torch.Tensor:
    if n_rep == 1:
        return x
    b, kv, L, d = x.shape
    return x[:, :, None].expand(b, kv, n_rep, L, d).reshape(b, kv * n_rep, L, d)


def attention(q, k, v, mask, n_rep: int) -> torch.Tensor:
    """q [B,h,Lq,d], k/v [B,kv,Lk,d], additive mask broadcastable to [B,1,Lq,Lk] or None."""
    k = repeat_kv(k, n_rep)
    v = repeat_kv(v, n_rep)
    w = torch.matmul(q, k.transpose(2, 3)) * (q.shape[-1] ** -0.5)
    if mask is not None:
        w = w + mask
    w = torch.softmax(w, dim=-1, dtype=torch.float32).to(q.dtype)
    o = torch.matmul(w, v)                       # [B,h,Lq,d]
    return o.transpose(1, 2).contiguous()        # [B,Lq,h,d]


def decoder_layer(x: torch.Tensor, W: dict, prefix: str, cfg: dict, positions: torch.Tensor,
                  mask, past_kv=None):
    """One LlamaDecoderLayer.  x [B,L,H].  Returns (y, (k, v)) with k/v = post-RoPE,
    pre-repeat, *own* keys/values [B,kv,L,d] (what tf 4.3x returns as present_key_value
    before concatenation is irrelevant here: the caller keeps prefix KV itself)."""
    B, L, H = x.shape
    nh, nkv = cfg["num_attention_heads"], cfg["num_key_value_heads"]
    d = H // nh
    eps = cfg["rms_norm_eps"]
    lin = torch.nn.functional.linear

    h = rmsnorm(x, W[f"{prefix}.input_layernorm.weight"], eps)
    q = lin(h, W[f"{prefix}.self_attn.q_proj.weight"]).view(B, L, nh, d).transpose(1, 2)
    k = lin(h, W[f"{prefix}.self_attn.k_proj.weight"]).view(B, L, nkv, d).transpose(1, 2)
    v = lin(h, W[f"{prefix}.self_attn.v_proj.weight"]).view(B, L, nkv, d).transpose(1, 2)
    cos, sin = rope_cos_sin(positions, d, cfg.get("rope_theta", 10000.0), x.dtype)
    q = apply_rope(q, cos, sin)
    k = apply_rope(k, cos, sin)
    if past_kv is not None:
        k_all = torch.cat((past_kv[0], k), dim=2)
        v_all = torch.cat((past_kv[1], v), dim=2)
    else:
        k_all, v_all = k, v
    a = attention(q, k_all, v_all, mask, nh // nkv).reshape(B, L, H)
    x = x + lin(a, W[f"{prefix}.self_attn.o_proj.weight"])

    h = rmsnorm(x, W[f"{prefix}.post_attention_layernorm.weight"], eps)
    g = lin(h, W[f"{prefix}.mlp.gate_proj.weight"])
    u = lin(h, W[f"{prefix}.mlp.up_proj.weight"])
    x = x + lin(torch.nn.functional.silu(g) * u, W[f"{prefix}.mlp.down_proj.weight"])
    return x, (k, v)


# --------------------------------------------------------------------------------------
# the scoring loop (utils.py:133-305 restated)
# --------------------------------------------------------------------------------------
def load_layer(model_path: str, layer_name: str, dtype, device="cpu") -> dict:
    """utils.py:121-131 (weights moved to the device and cast to the run dtype, as set_module_tensor_to_device does)."""
    from safetensors.torch import load_file
    sd = load_file(os.path.join(model_path, f"{layer_name}.safetensors"))
    return {k: v.to(device).to(dtype) for k, v in sd.items()}


class OracleShardedLlama:
    """Restated `ShardedLlama` (single device, `storage_location='cpu'`).  `device="cuda:0"` with
    `dtype=torch.float16` is the reference's own arithmetic path on this box (torch eager + cuBLAS, one prompt
    at a time, utils.py:228-297): the like-for-like comparator of the GPU tests and bench.py's
    `gpu_eager_baseline`."""

    def __init__(self, model_path: str, tokenizer=None, dtype=torch.float32,
                 layer_num_per_shard: int = 1, prefix_causal: bool = False,
                 max_decoder_layers: int | None = None, device="cpu"):
        self.model_path = model_path
        self.device = torch.device(device)
        with open(os.path.join(model_path, "config.json")) as f:
            self.cfg = json.load(f)
        if tokenizer is None:
            from transformers import AutoTokenizer
            tokenizer = AutoTokenizer.from_pretrained(model_path)
        tokenizer.pad_token = tokenizer.eos_token       # utils.py:103
        tokenizer.padding_side = "right"                # utils.py:104
        self.tokenizer = tokenizer
        self.dtype = dtype
        self.k = layer_num_per_shard
        self.prefix_causal = prefix_causal
        L = self.cfg["num_hidden_layers"]
        if max_decoder_layers is not None:              # bounded cpu_baseline sample only
            L = min(L, max_decoder_layers)
        self.layer_names = (["model.embed_tokens"] + [f"model.layers.{i}" for i in range(L)]
                            + ["model.norm", "lm_head"])
        self.hidden_trace = None   # optional: dict layer_idx -> list of (prefix, suffix) copies

    def __call__(self, inputs: list, trace_layers=()) -> list[np.ndarray]:
        n = len(inputs)
        dt = self.dtype
        shards = shard_plan(len(self.layer_names), self.k, True)
        activation = {}
        prompt2layer = defaultdict(lambda: -1)
        suffix_eos: list[torch.Tensor] = []
        out: list[np.ndarray] = []
        trace = {li: [] for li in trace_layers}

        # utils.py:219-221
        dev = self.device
        mask_full = torch.finfo(dt).min * torch.ones((MAX_TOKEN_LEN, MAX_TOKEN_LEN), dtype=dt)
        mask_full = mask_full.triu(diagonal=1)[None, None].to(dev)
        position_ids = torch.arange(MAX_TOKEN_LEN, dtype=torch.long, device=dev)

        with torch.inference_mode():
            for shard in shards:
                W = {}
                for li in shard:
                    W.update(load_layer(self.model_path, self.layer_names[li], dt, dev))
                for pi in range(n):
                    for li in shard:
                        name = self.layer_names[li]
                        if li == 0:
                            prefix, suffix, eos = tokenize_prompt(self.tokenizer, inputs[pi][0], inputs[pi][1])
                            prefix, suffix = prefix.to(dev), suffix.to(dev)
                            suffix_eos.append(eos)
                        elif li == shard[0]:
                            assert prompt2layer[pi] == li - 1
                            prefix, suffix = activation[pi]

                        if name == "model.embed_tokens":
                            E = W["model.embed_tokens.weight"]
                            prefix, suffix = E[prefix], E[suffix]
                        elif name.startswith("model.layers."):
                            len_p, len_s, S = prefix.shape[1], suffix.shape[1], suffix.shape[0]
                            pmask = mask_full[:, :, :len_p, :len_p] if self.prefix_causal else None
                            prefix, (kc, vc) = decoder_layer(prefix, W, name, self.cfg,
                                                             position_ids[:len_p], pmask)
                            smask = mask_full[:, :, -len_s:, -len_p - len_s:]      # utils.py:276
                            kv = (kc.expand(S, -1, -1, -1), vc.expand(S, -1, -1, -1))
                            suffix, _ = decoder_layer(suffix, W, name, self.cfg,
                                                      position_ids[len_p:len_p + len_s], smask, kv)
                        elif name == "model.norm":
                            prefix = None
                            pos = suffix_eos[pi]
                            suffix = rmsnorm(suffix[torch.arange(suffix.shape[0], device=dev), pos.to(dev)][:, None],
                                             W["model.norm.weight"], self.cfg["rms_norm_eps"])
                        else:
                            logits = torch.nn.functional.linear(suffix, W["lm_head.weight"])[:, 0]
                            scores = torch.softmax(logits, dim=-1).cpu().numpy()
                            out.append(np.expand_dims(scores, axis=1))
                        if li in trace:
                            trace[li].append((None if prefix is None else prefix.clone(), suffix.clone()))
                        if li == shard[-1] and li != len(self.layer_names) - 1:
                            activation[pi] = (prefix, suffix)
                            prompt2layer[pi] = li
        self.hidden_trace = trace
        self.suffix_eos = suffix_eos
        return out


def option_logprobs(scores: list[np.ndarray], token_id: int) -> np.ndarray:
    """Per-option log-likelihood of `token_id` as the next token: [n_prompts, S] (float64)."""
    return np.stack([np.log(np.maximum(s[:, 0, token_id].astype(np.float64), 1e-300)) for s in scores])
