"""Golden-vector generator: the reference's scoring loop driven through the REAL `transformers`
Llama modules.  TEST INFRASTRUCTURE ONLY (see oracle/llama_oracle.py header for the import rule).

Why this exists: the reference's arithmetic lives in the third-party `transformers` package
(requirements.txt:2, unpinned; effective window 4.31-4.35), not in /root/reference, and
`import utils` cannot run in this image (`accelerate` is absent, and transformers 5.5.0 changed
the decoder-layer call convention used at utils.py:272,277-279).  What *can* be run is the
installed package's own modules - `LlamaDecoderLayer`, `LlamaRMSNorm`, `LlamaRotaryEmbedding`,
`nn.Embedding`, `lm_head` - called layer by layer exactly the way the reference calls them:

  prefix pass   layer(prefix, attention_mask=None, use_cache=True)            utils.py:272
  suffix pass   layer(suffix, past_key_value=(k.expand(S), v.expand(S)),
                      position_ids=len_p.., attention_mask=mask[-len_s:, -len_p-len_s:])
                                                                               utils.py:275-279
  norm          norm(suffix[arange(S), suffix_eos][:, None])                   utils.py:281-286
  head          softmax(lm_head(x)[:, 0])                                      utils.py:287-291

The only adaptation is mechanical: transformers 5.5 wants `position_embeddings=(cos, sin)` from
the model-level rotary module and a `Cache` object instead of a `(k, v)` tuple.  The vectors this
produces pin oracle/llama_oracle.py (tests/test_oracle_golden.py) and, through it, the CUDA path.
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from .llama_oracle import MAX_TOKEN_LEN, tokenize_prompt


def build_hf_model(model_path: str, dtype):
    from safetensors.torch import load_file
    from transformers import LlamaConfig, LlamaForCausalLM

    with open(os.path.join(model_path, "config.json")) as f:
        cfgd = json.load(f)
    cfg = LlamaConfig(**{k: v for k, v in cfgd.items() if k not in ("architectures", "model_type", "torch_dtype")})
    cfg._attn_implementation = "eager"
    model = LlamaForCausalLM(cfg)
    names = (["model.embed_tokens"] + [f"model.layers.{i}" for i in range(cfg.num_hidden_layers)]
             + ["model.norm", "lm_head"])
    sd = {}
    for n in names:
        sd.update(load_file(os.path.join(model_path, f"{n}.safetensors")))
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected and all("rotary" in m or "inv_freq" in m for m in missing), (missing, unexpected)
    # Keep inv_freq in fp32.  In transformers 4.31-4.35 the cos/sin tables are built at module
    # construction from the fp32 inv_freq and only then cast to the run dtype by the reference
    # (utils.py:117-119), so for sequences within max_position_embeddings the reference sees
    # fp16(cos(fp32 angle)).  `model.to(fp16)` would instead round inv_freq itself to fp16.
    inv_freq32 = model.model.rotary_emb.inv_freq.detach().clone().to(torch.float32)
    model = model.to(dtype).eval()
    model.model.rotary_emb.inv_freq = inv_freq32
    return model


def hf_reference_scores(model_path: str, prompts, tokenizer, dtype=torch.float32,
                        prefix_causal: bool = False, trace_layers=()):
    """Returns (scores list[np.ndarray [S,1,V]], trace {layer_idx: [(prefix, suffix), ...]})."""
    from transformers import DynamicCache

    model = build_hf_model(model_path, dtype)
    tokenizer.pad_token = tokenizer.eos_token
    tokenizer.padding_side = "right"
    base = model.model
    mask_full = torch.finfo(dtype).min * torch.ones((MAX_TOKEN_LEN, MAX_TOKEN_LEN), dtype=dtype)
    mask_full = mask_full.triu(diagonal=1)[None, None]
    position_ids = torch.arange(MAX_TOKEN_LEN, dtype=torch.long)[None, :]
    out, trace = [], {li: [] for li in trace_layers}

    with torch.inference_mode():
        for prefix_s, suffixes in prompts:
            p_ids, s_ids, eos = tokenize_prompt(tokenizer, prefix_s, suffixes)
            prefix, suffix = base.embed_tokens(p_ids), base.embed_tokens(s_ids)
            if 0 in trace:
                trace[0].append((prefix.clone(), suffix.clone()))
            len_p, len_s, S = prefix.shape[1], suffix.shape[1], suffix.shape[0]
            for i, layer in enumerate(base.layers):
                pos_p = position_ids[:, :len_p]
                cache = DynamicCache()
                pe = base.rotary_emb(prefix, pos_p)
                pmask = mask_full[:, :, :len_p, :len_p] if prefix_causal else None
                prefix = layer(prefix, attention_mask=pmask, position_ids=pos_p, past_key_values=cache,
                               use_cache=True, position_embeddings=pe)
                if isinstance(prefix, tuple):
                    prefix = prefix[0]
                k, v = cache.layers[i].keys, cache.layers[i].values

                pos_s = position_ids[:, len_p:len_p + len_s].expand(S, -1)
                attn = mask_full[:, :, -len_s:, -len_p - len_s:].expand(S, -1, -1, -1)
                cache2 = DynamicCache()
                cache2.update(k.expand(S, -1, -1, -1).contiguous(), v.expand(S, -1, -1, -1).contiguous(), i)
                pe = base.rotary_emb(suffix, pos_s)
                suffix = layer(suffix, attention_mask=attn, position_ids=pos_s, past_key_values=cache2,
                               use_cache=True, position_embeddings=pe)
                if isinstance(suffix, tuple):
                    suffix = suffix[0]
                if i + 1 in trace:
                    trace[i + 1].append((prefix.clone(), suffix.clone()))
            x = base.norm(suffix[torch.arange(S), eos][:, None])
            logits = model.lm_head(x)[:, 0]
            scores = torch.softmax(logits, dim=-1).cpu().numpy()
            out.append(np.expand_dims(scores, axis=1))
    return out, trace
