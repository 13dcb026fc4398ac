"""Runs the UNMODIFIED reference `utils.py` (imported from /root/reference, never copied) on the CPU.
TEST INFRASTRUCTURE ONLY (see oracle/llama_oracle.py header for the import rule); needs /root/reference,
so it works in the build container only - its outputs travel as fixtures (tests/golden/).

Why a shim is needed: `import utils` fails here on two counts.
  1. `accelerate` is not installed.  The reference uses exactly two of its functions
     (utils.py:10-11): `init_empty_weights` (utils.py:111) and `set_module_tensor_to_device`
     (utils.py:119,130).  Both are restated below from their published behaviour: parameters created inside
     the context land on the meta device while buffers stay real, and a named parameter / buffer is replaced
     by `value` moved to `device` and cast to `dtype` when it is a floating-point tensor.
  2. The installed transformers (5.5) changed `LlamaDecoderLayer.forward`: it returns a bare tensor, wants
     `position_embeddings=(cos, sin)` from the model-level rotary module and a `Cache` object, and would
     swallow the reference's `past_key_value=` tuple into **kwargs.  The reference was written against the
     4.31-4.35 convention (utils.py:272,277-279): `layer(x, use_cache=True, attention_mask=None)` returns
     `(hidden, (k, v))`, `layer(x, past_key_value=(k, v), position_ids=..., attention_mask=...)` returns
     `(hidden,)`.  `_forward_43x` adapts the call and delegates ALL arithmetic to the installed module
     (RMSNorm, projections, rotary, eager attention, MLP, residuals run in transformers' own code).
     Rotary tables: in 4.31-4.35 every attention module held cos/sin tables precomputed in fp32 at
     construction, which utils.py:117-119 then casts to the run dtype; so the adapter evaluates cos/sin from
     an fp32 inv_freq and casts the tables (not inv_freq) to the run dtype.  With `position_ids=None` the
     4.3x code indexes the table with `cos[None]`, i.e. positions 0..L-1.

Everything else - DeviceManager, the shard-outer / prompt-inner loop, tokenisation, store / fetch of
activations in all three tiers, the mask slice, the last-token gather, the fp16 softmax - is the reference's
own code, executed as it stands.
"""
from __future__ import annotations

import contextlib
import importlib.util
import os
import sys
import types

import torch
from torch import nn

REFERENCE_DIR = "/root/reference"


def available() -> bool:
    return os.path.exists(os.path.join(REFERENCE_DIR, "utils.py"))


# ------------------------------------------------------------------ accelerate stand-ins
@contextlib.contextmanager
def init_empty_weights(include_buffers: bool = False):
    old_register = nn.Module.register_parameter

    def register_on_meta(module, name, param):
        old_register(module, name, param)
        if param is not None:
            p = module._parameters[name]
            module._parameters[name] = type(p)(p.to("meta"), requires_grad=p.requires_grad)

    nn.Module.register_parameter = register_on_meta
    try:
        yield
    finally:
        nn.Module.register_parameter = old_register


def set_module_tensor_to_device(module, tensor_name, device, value=None, dtype=None, **_unused):
    *path, leaf = tensor_name.split(".")
    for part in path:
        module = getattr(module, part)
    is_buffer = leaf in module._buffers
    old = module._buffers[leaf] if is_buffer else module._parameters[leaf]
    if value is None:
        value = old
    if dtype is None:
        value = value.to(old.dtype)
    elif value.is_floating_point():
        value = value.to(dtype)
    value = value.to(device)
    if is_buffer:
        module._buffers[leaf] = value
    else:
        module._parameters[leaf] = nn.Parameter(value, requires_grad=False)


def _install_accelerate_stub():
    if "accelerate" in sys.modules and not getattr(sys.modules["accelerate"], "_fls_stub", False):
        return                                     # a real accelerate is importable: use it
    acc = types.ModuleType("accelerate")
    acc._fls_stub = True
    acc.init_empty_weights = init_empty_weights
    acc_utils = types.ModuleType("accelerate.utils")
    acc_modeling = types.ModuleType("accelerate.utils.modeling")
    acc_modeling.set_module_tensor_to_device = set_module_tensor_to_device
    acc_utils.modeling = acc_modeling
    acc.utils = acc_utils
    sys.modules["accelerate"] = acc
    sys.modules["accelerate.utils"] = acc_utils
    sys.modules["accelerate.utils.modeling"] = acc_modeling


# ------------------------------------------------------------------ decoder-layer call convention of tf 4.31-4.35
def _forward_43x(self, hidden_states, attention_mask=None, position_ids=None, past_key_value=None,
                 output_attentions=False, use_cache=False, **kwargs):
    from transformers import DynamicCache
    cfg = self.self_attn.config
    cfg._attn_implementation = "eager"              # the only implementation the 4.31-4.35 window had by default
    B, L, _ = hidden_states.shape
    past_len = 0 if past_key_value is None else past_key_value[0].shape[2]
    if position_ids is None:
        if past_len:
            raise ValueError("tf 4.3x needs position_ids when a KV cache is passed")
        position_ids = torch.arange(L, device=hidden_states.device)[None, :]
    d = self.self_attn.head_dim
    theta = getattr(cfg, "rope_theta", None) or (getattr(cfg, "rope_parameters", None) or {}).get("rope_theta", 10000.0)
    inv_freq = 1.0 / (theta ** (torch.arange(0, d, 2, dtype=torch.int64).to(torch.float32) / d))
    freqs = position_ids.to(torch.float32)[:, :, None] * inv_freq[None, None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    cos, sin = emb.cos().to(hidden_states.dtype), emb.sin().to(hidden_states.dtype)
    cache = DynamicCache()
    idx = self.self_attn.layer_idx
    if past_key_value is not None:
        cache.update(past_key_value[0].contiguous(), past_key_value[1].contiguous(), idx)
    out = self._fls_forward_installed(hidden_states, attention_mask=attention_mask, position_ids=position_ids,
                                      past_key_values=cache, use_cache=True, position_embeddings=(cos, sin))
    if isinstance(out, tuple):
        out = out[0]
    if use_cache:
        return out, (cache.layers[idx].keys, cache.layers[idx].values)
    return (out,)


@contextlib.contextmanager
def reference_environment():
    """Context in which `import utils` of the reference works and its layer calls reach transformers 5.x."""
    from transformers.models.llama.modeling_llama import LlamaDecoderLayer
    _install_accelerate_stub()
    installed = LlamaDecoderLayer.forward
    LlamaDecoderLayer._fls_forward_installed = installed
    LlamaDecoderLayer.forward = _forward_43x
    try:
        yield
    finally:
        LlamaDecoderLayer.forward = installed
        del LlamaDecoderLayer._fls_forward_installed


def load_reference_utils():
    """The reference's utils.py as a module object, loaded from where it lies."""
    _install_accelerate_stub()
    spec = importlib.util.spec_from_file_location("fls_reference_utils", os.path.join(REFERENCE_DIR, "utils.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_scores(model_path: str, prompts, dtype=torch.float32, layer_num_per_shard: int = 1,
                     storage_location: str = "cpu", data_parallel: bool = False, disk_folder: str = "./temp",
                     num_batch: int = 1):
    """`ShardedLlama(args, DeviceManager(args, ["cpu"]), device="cpu", dtype)(prompts)` of the reference,
    batched like main.py:19-23.  Returns list[np.ndarray [S, 1, V]]."""
    import copy
    with reference_environment():
        ref = load_reference_utils()
        args = types.SimpleNamespace(model_path=model_path, num_batch=num_batch,
                                     layer_num_per_shard=layer_num_per_shard, storage_location=storage_location,
                                     max_activation_in_cpu=100, data_parallel=data_parallel, disk_folder=disk_folder,
                                     num_gen_token=1)
        if storage_location == "disk":
            os.makedirs(disk_folder, exist_ok=True)
        dm = ref.DeviceManager(args, ["cpu"])
        model = ref.ShardedLlama(args, dm, device="cpu", dtype=dtype)
        prompts = copy.deepcopy(list(prompts))
        ends = [len(prompts) // num_batch * i for i in range(1, num_batch)] + [len(prompts)]
        out = []
        for lo, hi in zip([0] + ends[:-1], ends):
            out += model(prompts[lo:hi])
        return out
