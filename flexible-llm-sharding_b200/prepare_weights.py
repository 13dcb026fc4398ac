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

`--packed` additionally writes `<layer>.flspack` next to every `<layer>.safetensors`: the same
tensors as fp16, each 256-byte aligned and in the order the HBM weight slots hold them (q | k | v, o, gate | up, down,
the two norm gains), behind a 4096-byte header.  The loader then reads ONE contiguous, page-aligned range per layer
(O_DIRECT from a disk, or locked in place on tmpfs) and uploads it as one run of 32 MiB pieces instead of nine
tensors cut up separately.  The reference layout is still written, byte for byte; `pack_layer_dir` adds the packed
files to a directory that already holds it.
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


# ---- packed layer files -------------------------------------------------------------------------------------------
PACK_MAGIC = b"FLSPACK1"
PACK_HEADER_BYTES = 4096
PACK_ENTRY_BYTES = 128          # char name[96]; u64 rows, cols (0 = 1-D), offset, nbytes
PACK_ALIGN = 256
# tensor-name suffixes in HBM slot order (csrc/api.cu fls_prefetch_shard)
PACK_ORDER = {
    "decoder": ["self_attn.q_proj.weight", "self_attn.k_proj.weight", "self_attn.v_proj.weight",
                "self_attn.o_proj.weight", "mlp.gate_proj.weight", "mlp.up_proj.weight", "mlp.down_proj.weight",
                ".input_layernorm.weight", "post_attention_layernorm.weight"],
    "model.embed_tokens": ["embed_tokens.weight"], "model.norm": ["norm.weight"], "lm_head": ["lm_head.weight"],
}


def pack_order(layer: str, names) -> list[str]:
    """The layer's tensor names in slot order; raises if the layer holds anything else (biases, adapters)."""
    suffixes = PACK_ORDER.get(layer, PACK_ORDER["decoder"])
    names = list(names)
    out = []
    for suf in suffixes:
        hit = [n for n in names if n.endswith(suf)]
        if len(hit) != 1:
            raise ValueError(f"{layer}: expected exactly one tensor *{suf}, found {hit}")
        out.append(hit[0])
    extra = [n for n in names if n not in out and not n.endswith("rotary_emb.inv_freq")]
    if extra:
        raise ValueError(f"{layer}: tensors {extra} have no place in the packed layout")
    return out


def write_packed_layer(path: str, layer: str, tensors: dict) -> None:
    """tensors: name -> torch tensor (any float dtype; stored as fp16, as the loader would convert them)."""
    import struct
    import torch
    order = pack_order(layer, tensors)
    entries, off = [], 0
    for n in order:
        t = tensors[n]
        nbytes = t.numel() * 2
        rows, cols = (t.shape[0], t.shape[1]) if t.dim() == 2 else (t.shape[0], 0)
        entries.append((n, rows, cols, off, nbytes))
        off = (off + nbytes + PACK_ALIGN - 1) // PACK_ALIGN * PACK_ALIGN
    data_bytes = (off + PACK_HEADER_BYTES - 1) // PACK_HEADER_BYTES * PACK_HEADER_BYTES     # whole pages: O_DIRECT
    head = bytearray(PACK_HEADER_BYTES)
    struct.pack_into("<8sIIQQ", head, 0, PACK_MAGIC, 0, len(entries), PACK_HEADER_BYTES, data_bytes)
    for i, (n, rows, cols, o, nb) in enumerate(entries):
        name = n.encode()
        if len(name) > 95:
            raise ValueError(f"tensor name too long for the packed header: {n}")
        struct.pack_into("<96sQQQQ", head, 32 + i * PACK_ENTRY_BYTES, name, rows, cols, o, nb)
    with open(path, "wb", buffering=0) as f:
        f.write(head)
        pos = 0
        for (n, rows, cols, o, nb) in entries:
            if o > pos:
                f.write(bytes(o - pos))
            t16 = tensors[n].detach().to("cpu").to(torch.float16).contiguous()
            f.write(memoryview(t16.numpy()).cast("B"))
            pos = o + nb
        if data_bytes > pos:
            f.write(bytes(data_bytes - pos))


def read_packed_header(path: str) -> list[tuple]:
    """[(name, rows, cols, offset, nbytes)] of a packed layer file (tests, inspection)."""
    import struct
    with open(path, "rb") as f:
        head = f.read(PACK_HEADER_BYTES)
    magic, _kind, n, data_off, _data_bytes = struct.unpack_from("<8sIIQQ", head, 0)
    if magic != PACK_MAGIC or data_off != PACK_HEADER_BYTES:
        raise ValueError(f"{path}: not a packed layer file")
    out = []
    for i in range(n):
        name, rows, cols, o, nb = struct.unpack_from("<96sQQQQ", head, 32 + i * PACK_ENTRY_BYTES)
        out.append((name.rstrip(b"\0").decode(), rows, cols, o, nb))
    return out


def pack_layer_dir(layer_dir: str) -> list[str]:
    """Adds `<layer>.flspack` for every `<layer>.safetensors` of a directory in the reference layout."""
    from safetensors.torch import load_file
    done = []
    for fn in sorted(os.listdir(layer_dir)):
        if fn.endswith(".safetensors"):
            layer = fn[:-len(".safetensors")]
            write_packed_layer(os.path.join(layer_dir, layer + ".flspack"), layer,
                               load_file(os.path.join(layer_dir, fn)))
            done.append(layer)
    return done


def split_into_layers(bin_dir: str, new_file_dir: str, packed: bool = False) -> list[str]:
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
        if packed:
            write_packed_layer(os.path.join(new_file_dir, f"{layer}.flspack"), layer, tensors)
        written.append(layer)
        for shard in [s for s, p in last_use.items() if p == pos]:
            open_shards.pop(shard, None)
    return written


def main(argv=None):
    parser = argparse.ArgumentParser(description="Split model weights into layers.")
    parser.add_argument("bin_dir", type=str, help="Path to the HF weights directory")
    parser.add_argument("new_file_dir", type=str, help="Path to the new layer-wise file directory")
    parser.add_argument("--packed", action="store_true",
                        help="also write <layer>.flspack files (one aligned, slot-ordered fp16 blob per layer)")
    args = parser.parse_args(argv)
    split_into_layers(args.bin_dir, args.new_file_dir, packed=args.packed)


if __name__ == "__main__":
    main()
