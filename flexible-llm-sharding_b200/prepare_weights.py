"""Split a HuggingFace Llama checkpoint into one safetensors file per top-level layer.

Drop-in for the reference's `prepare_weights.py` (reference prepare_weights.py:12-62): same CLI
(`python prepare_weights.py <bin_dir> <new_file_dir>`), same output layout, byte-identical files:

* every file of `bin_dir` whose name does not contain ".bin" is copied (config, tokenizer; the
  `pytorch_model.bin.index.json` is therefore *not* copied - reference :14-16),
* tensors are grouped by the first three dotted components of their name with ".weight"
  removed (`model.layers.7`, `model.embed_tokens`, `model.norm`, `lm_head` - reference :21),
* each group is written to `<new_file_dir>/<layer>.safetensors` under the original key names and
  dtypes (reference :40-43).

Beyond the reference: a `model.safetensors.index.json` (+ `.safetensors` shards) input is
accepted too, and `.bin` shards are released as soon as no later layer needs them.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
from collections import OrderedDict


def layer_of(tensor_name: str) -> str:
    return ".".join(tensor_name.replace(".weight", "").split(".")[:3])


def _load_shard(path: str) -> dict:
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file
        return load_file(path)
    import torch
    return torch.load(path, map_location="cpu")


def split_into_layers(bin_dir: str, new_file_dir: str) -> list[str]:
    from safetensors.torch import save_file

    os.makedirs(new_file_dir, exist_ok=True)
    for entry in sorted(os.listdir(bin_dir)):
        src = os.path.join(bin_dir, entry)
        if ".bin" not in src and os.path.isfile(src) and not entry.endswith(".safetensors") \
                and entry != "model.safetensors.index.json":
            shutil.copy(src, os.path.join(new_file_dir, entry))

    index_path = os.path.join(bin_dir, "pytorch_model.bin.index.json")
    if not os.path.exists(index_path):
        index_path = os.path.join(bin_dir, "model.safetensors.index.json")
    with open(index_path) as f:
        weight_map: dict[str, str] = json.load(f)["weight_map"]

    members: "OrderedDict[str, list[str]]" = OrderedDict()
    for tensor_name in weight_map:
        members.setdefault(layer_of(tensor_name), []).append(tensor_name)
    # process layers in the order their first shard appears, so each shard is opened once
    order = sorted(members, key=lambda l: (min(weight_map[t] for t in members[l]), l))
    last_use = {}
    for pos, layer in enumerate(order):
        for t in members[layer]:
            last_use[weight_map[t]] = pos

    open_shards: dict[str, dict] = {}
    written = []
    for pos, layer in enumerate(order):
        tensors = {}
        for t in members[layer]:
            shard = weight_map[t]
            if shard not in open_shards:
                open_shards[shard] = _load_shard(os.path.join(bin_dir, shard))
            tensors[t] = open_shards[shard][t].contiguous()
        if len(tensors) != len(members[layer]):
            raise RuntimeError(f"{layer}: expected {len(members[layer])} tensors, found {len(tensors)}")
        save_file(tensors, os.path.join(new_file_dir, f"{layer}.safetensors"))
        written.append(layer)
        for shard in [s for s, p in last_use.items() if p == pos]:
            open_shards.pop(shard, None)
    return written


def main(argv=None):
    parser = argparse.ArgumentParser(description="Split model weights into layers.")
    parser.add_argument("bin_dir", type=str, help="Path to the HF weights directory")
    parser.add_argument("new_file_dir", type=str, help="Path to the new layer-wise file directory")
    args = parser.parse_args(argv)
    split_into_layers(args.bin_dir, args.new_file_dir)


if __name__ == "__main__":
    main()
