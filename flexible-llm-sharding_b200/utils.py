"""DeviceManager and ShardedLlama - same names, constructor arguments and call convention as the
reference's `utils.py` (reference utils.py:24-75 and :78-305), re-built for B200.

What stays: `ShardedLlama(args, device_manager, device="cuda:0", dtype=torch.float16)`;
`model(inputs)` takes a list of `(prefix, (suffix, ...))`, sets every consumed entry of `inputs`
to None (utils.py:259) and returns one `np.ndarray [n_suffixes, 1, vocab]` of float16 next-token
probabilities per prompt (utils.py:289-290), scored layer-shard by layer-shard.

What changes: all prompts are tokenised up front and packed into one row matrix;
each shard is streamed from pinned host memory into one of two HBM slots on a copy stream while the
previous shard computes; every layer runs as hand-written sm_100a kernels behind the C ABI
(`include/fls_b200.h`); activations stay in HBM when they fit.  There is no CPU path.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import hashlib
import json
import os
import threading
import warnings
from collections import defaultdict
from time import perf_counter

import numpy as np
import torch

from . import _lib, planner
from .activation_store import ActivationStore

max_token_len = planner.MAX_TOKEN_LEN  # reference utils.py:14

_tokenizer_lock = threading.Lock()


def load_tokenizer(model_path):
    """AutoTokenizer.from_pretrained under a lock: the per-GPU worker threads construct their models at
    the same time and transformers' lazy module import is not thread-safe."""
    with _tokenizer_lock:
        from transformers import AutoTokenizer
        return AutoTokenizer.from_pretrained(model_path)


def _on_tmpfs(path) -> bool:
    """True when `path` lies on a RAM file system (its F16 layer files are then page-locked in place)."""
    try:
        real = os.path.realpath(path)
        best, kind = "", ""
        with open("/proc/mounts") as f:
            for line in f:
                _, mnt, fstype = line.split()[:3]
                if (real == mnt or real.startswith(mnt.rstrip("/") + "/")) and len(mnt) > len(best):
                    best, kind = mnt, fstype
        return kind in ("tmpfs", "ramfs")
    except (OSError, ValueError):
        return False


def clean_memory():
    """Kept for API compatibility (reference utils.py:18-21).  Nothing to do: weight slots and
    workspace are allocated once per context instead of per shard."""
    return None


class DeviceManager:
    """Shared state of the per-GPU workers of one process (reference utils.py:24-75).

    data_parallel: the reference keeps at most `max_layer_num` layer state-dicts in host RAM and
    hands the same dict to every GPU thread.  Here that cache lives inside the C library (one
    pinned image per layer file per process, `csrc/host_cache.cpp`); `send_device_request`
    pre-loads a layer file into it and `get_state_dict` exposes the tensors for inspection.
    model parallel: `shared_activation_dict`, `prompt2layer` and `suffix_eos` are the hand-off
    state between pipeline stages, as in the reference (utils.py:42-44,157).
    """

    def __init__(self, args, devices, max_layer_num=1):
        assert max_layer_num >= 1, "Must set max_layer_num >= 1 to allow loading a layer"
        self.args = args
        self.devices = list(devices)
        self.max_layer_num = max_layer_num
        self.lock = threading.Lock()
        self.condition = threading.Condition()
        self.waiting_queue = []
        # model-parallel hand-off state
        self.shared_activation_dict = dict()
        self.prompt2layer = defaultdict(lambda: -1)
        self.suffix_eos = []
        self.handoff = {}          # pipeline hand-off board: packed prompts, peer buffers, events
        self.distributed = False   # True: `devices` lists one GPU per PROCESS of the torch.distributed group
        self.models = {}           # device -> ShardedLlama, reused across batches / generated tokens
        self.failed = None

    def send_device_request(self, device, layer):
        """Ask for `layer` to be resident in host memory (pinned)."""
        path = f"{self.args.model_path}/{layer}.safetensors"
        with self.lock:
            if layer not in self.waiting_queue:
                self.waiting_queue.append(layer)
        rc = _lib.load().fls_preload_layer_file(path.encode())
        with self.lock:
            if layer in self.waiting_queue:
                self.waiting_queue.remove(layer)
        if rc != 0:
            raise _lib.FlsError(rc, _lib.load().fls_last_error(None).decode())

    def get_state_dict(self, device, layer):
        from safetensors.torch import load_file
        return load_file(f"{self.args.model_path}/{layer}.safetensors")

    def clean_waiting_queue(self):
        _lib.load().fls_trim_host_cache(C.c_size_t(0))


class ShardedLlama:
    def __init__(self, args, device_manager, device="cuda:0", dtype=torch.float16):
        if dtype != torch.float16:
            raise ValueError("the B200 path computes in fp16 like the reference's default (utils.py:80)")
        if not torch.cuda.is_available():
            raise RuntimeError("ShardedLlama needs a CUDA device: there is no CPU fallback")
        self.args = args
        self.device = device
        self.torch_device = torch.device(device)
        self.dtype = dtype
        self.device_manager = device_manager

        with open(os.path.join(args.model_path, "config.json")) as f:
            self.config = json.load(f)
        self._check_config(self.config)
        self.tokenizer = load_tokenizer(args.model_path)
        self.tokenizer.pad_token = self.tokenizer.eos_token      # utils.py:103
        self.tokenizer.padding_side = "right"                    # utils.py:104
        cfg = self.config
        L = cfg["num_hidden_layers"]
        self.layer_names = ["model.embed_tokens"] + [f"model.layers.{i}" for i in range(L)] + ["model.norm", "lm_head"]
        self.layer_kinds = [_lib.LAYER_EMBED] + [_lib.LAYER_DECODER] * L + [_lib.LAYER_NORM, _lib.LAYER_HEAD]

        # extras with reference-compatible defaults (not CLI flags in the reference)
        self.prefix_causal = bool(getattr(args, "prefix_causal", False))     # False = reference (utils.py:272)
        self.max_rows = int(getattr(args, "max_rows", None) or os.environ.get("FLS_MAX_ROWS", 16384))
        self.max_head_rows = int(getattr(args, "max_head_rows", 4096))
        self.force_tier = bool(getattr(args, "force_storage_tier", False) or os.environ.get("FLS_FORCE_STORAGE"))
        self.hbm_activation_budget = getattr(args, "hbm_activation_budget", None)
        # "length" = prompts go to micro-batches by planner.balanced_bins; "input" = in the
        # caller's order.  Scores come back in the caller's order either way, bit for bit the same.
        self.pack_order = str(getattr(args, "pack_order", None) or os.environ.get("FLS_PACK_ORDER", "length"))
        if self.pack_order not in ("length", "input"):
            raise ValueError(f"pack_order must be 'length' or 'input', not {self.pack_order!r}")
        # Weight residency: "auto" keeps every shard of this rank in HBM across calls when they
        # fit `resident_weight_fraction` of the free HBM (one slot per shard, streamed once); "never" always
        # streams through two slots like the reference re-loads per call; "always" insists.
        self.resident_weights = str(getattr(args, "resident_weights", os.environ.get("FLS_RESIDENT_WEIGHTS", "auto")))
        self.resident_weight_fraction = float(getattr(args, "resident_weight_fraction", 0.5))

        H = cfg["hidden_size"]
        nh = cfg["num_attention_heads"]
        self.hidden = H
        self.vocab = cfg["vocab_size"]
        self.head_dim = cfg.get("head_dim") or H // nh
        desc = _lib.ModelDesc(H, cfg["intermediate_size"], nh, cfg.get("num_key_value_heads", nh), self.head_dim,
                              self.vocab, L, float(cfg.get("rms_norm_eps", 1e-6)))
        idx = self.torch_device.index if self.torch_device.index is not None else 0
        self.ctx = _lib.Context(desc, idx)
        # `<layer>.flspack` (prepare_weights.py --packed) is preferred where it exists next to the reference layout
        use_packed = str(getattr(args, "use_packed", os.environ.get("FLS_USE_PACKED", "1"))) not in ("0", "False")
        self.layer_files = []
        for i, (name, kind) in enumerate(zip(self.layer_names, self.layer_kinds)):
            path = os.path.join(args.model_path, f"{name}.safetensors")
            packed = os.path.join(args.model_path, f"{name}.flspack")
            if use_packed and os.path.exists(packed):
                path = packed
            self.layer_files.append(path)
            self.ctx.call("fls_register_layer_file", i, kind, path.encode(), 0)
        self._layer_bytes = []
        for i in range(len(self.layer_names)):
            b = C.c_size_t()
            self.ctx.call("fls_layer_slot_bytes", i, C.byref(b))
            self._layer_bytes.append(b.value)
        cos, sin = planner.rope_tables(cfg.get("max_position_embeddings", max_token_len) if
                                       cfg.get("max_position_embeddings", max_token_len) >= max_token_len
                                       else max_token_len, self.head_dim, float(cfg.get("rope_theta", 10000.0)))
        self.ctx.call("fls_set_rope_table", cos.data_ptr(), sin.data_ptr(), cos.shape[0])
        self._set_pinned_limit(getattr(args, "pinned_limit_bytes", os.environ.get("FLS_PINNED_LIMIT_BYTES", "auto")))
        self._configured = None
        self._pinned_pool = {}
        self._stream = None
        self._kv = {}            # prefix signature -> per-layer prefix K/V of an earlier pass (see _kv_plan)
        self.kv_dim = cfg.get("num_key_value_heads", nh) * self.head_dim
        self.last_timing = {}

    @staticmethod
    def _check_config(cfg):
        """The reference builds whatever `AutoModelForCausalLM.from_config` returns (utils.py:112) and so honours
        every config field; the kernels here implement the Llama-2 architecture only.  Anything that would change
        the arithmetic is refused up front instead of giving silently different scores."""
        bad = []
        if cfg.get("model_type", "llama") != "llama":
            bad.append(f"model_type={cfg['model_type']!r} (only 'llama')")
        if cfg.get("hidden_act", "silu") != "silu":
            bad.append(f"hidden_act={cfg['hidden_act']!r} (only 'silu')")
        scaling = cfg.get("rope_scaling")
        if scaling and (scaling.get("rope_type", scaling.get("type")) not in (None, "default")):
            bad.append(f"rope_scaling={scaling!r} (not implemented)")
        for key in ("attention_bias", "mlp_bias", "tie_word_embeddings"):
            if cfg.get(key):
                bad.append(f"{key}=true (not implemented)")
        if cfg.get("pretraining_tp", 1) not in (None, 1):
            bad.append(f"pretraining_tp={cfg['pretraining_tp']} (sliced projections are not implemented)")
        if bad:
            raise ValueError("unsupported model configuration: " + "; ".join(bad))

    def _rank_tag(self, rank: int) -> str:
        """Suffix of the disk tier's file names (utils.py:172,176: the GPU rank with --data_parallel, else none).
        With one process per GPU (torchrun) the index within this process's device list is always 0, so the
        global rank is what keeps the processes' files apart in a shared --disk_folder."""
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            return os.environ.get("RANK", str(rank))
        return str(rank) if self.args.data_parallel else ""

    def close(self):
        """Release the GPU side of this model now (HBM slots, workspace, streams) instead of at garbage collection.
        In the cross-process pipeline this is a collective: every rank must call it (the hand-off rings are unmapped
        by all ranks before their owners free them)."""
        ring = getattr(self, "_ipc_ring", None)
        if ring is not None:
            torch.cuda.synchronize(self.torch_device)
            ring.close()
            self._ipc_ring = None
        self._kv.clear()
        self._keepalive = None
        self.ctx.close()

    def _set_pinned_limit(self, limit):
        """Bound the pinned host copies of the layer files (the reference's `max_layer_num`, utils.py:31, in bytes).
        "auto": unlimited when the whole model fits 60 % of this process's share of the available host RAM,
        otherwise that share - evicted layers are re-read from their files on the next pass."""
        if limit == "auto" and _on_tmpfs(self.args.model_path) and os.environ.get("FLS_WEIGHT_IMAGE", "auto") != "copy":
            limit = 0              # the files' own pages get locked in place: holding them all costs no extra RAM
        if limit == "auto":
            limit = 0
            try:
                with open("/proc/meminfo") as f:
                    avail_kb = next(int(l.split()[1]) for l in f if l.startswith("MemAvailable"))
                share = int(avail_kb * 1024 * 0.6 / max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1"))))
                if sum(self._layer_bytes) > share:
                    limit = share
            except (OSError, StopIteration, ValueError):
                limit = 0
        self.pinned_limit_bytes = int(limit)
        rc = self.ctx.lib.fls_set_host_cache_limit(C.c_size_t(self.pinned_limit_bytes))
        if rc != 0:
            raise _lib.FlsError(rc, "fls_set_host_cache_limit failed")

    # ------------------------------------------------------------------ plan
    def _plan(self):
        num_gpu = len(self.device_manager.devices)
        rank = self.device_manager.devices.index(self.device)
        dp = bool(self.args.data_parallel) or num_gpu == 1
        shards = planner.shard_plan(len(self.layer_names), self.args.layer_num_per_shard, dp, num_gpu, rank)
        return shards, num_gpu, rank, dp

    def _configure(self, shards):
        live = [s for s in shards if len(s)]
        slot_bytes = max([sum(self._layer_bytes[i] for i in s) for s in live] + [256])
        n_slots = 2
        if self.resident_weights in ("auto", "always") and len(live) > 2:
            need = slot_bytes * len(live)
            free = torch.cuda.mem_get_info(self.torch_device)[0] + (self._configured[0] * self._configured[3]
                                                                    if self._configured else 0)
            if self.resident_weights == "always" or need <= free * self.resident_weight_fraction:
                n_slots = len(live)
        key = (slot_bytes, self.max_rows, self.max_head_rows, n_slots)
        if self._configured != key:
            self.ctx.call("fls_configure", n_slots, C.c_size_t(slot_bytes), self.max_rows, self.max_head_rows)
            self._configured = key
            self._slot_holds = {}            # slot -> shard currently held there
        self._n_slots = n_slots

    @contextlib.contextmanager
    def _on_compute_stream(self, dev):
        """Run a pass on this model's own (non-blocking) CUDA stream, ordered after the caller's current stream
        at entry and before it at exit.  The legacy default stream is avoided on purpose: with the host activation
        tier the weight stream only made progress at shard boundaries when compute sat on stream 0
        (tools/trace_tier.py)."""
        caller = torch.cuda.current_stream(dev)
        if self._stream is None:
            self._stream = torch.cuda.Stream(dev)
        self._stream.wait_stream(caller)
        try:
            with torch.cuda.stream(self._stream):
                yield self._stream
        finally:
            caller.wait_stream(self._stream)

    def _stage_shard(self, live, si):
        """Start the H2D stream of shard `si` unless its slot already holds it (resident weights)."""
        slot = si % self._n_slots
        if self._slot_holds.get(slot) != live[si]:
            arr = (C.c_int32 * len(live[si]))(*live[si])
            self.ctx.call("fls_prefetch_shard", slot, arr, len(live[si]))
            self._slot_holds[slot] = live[si]
        return slot

    # ------------------------------------------------------------------ public call
    def __call__(self, inputs):
        """Reference-compatible scoring: list of fp16 probability arrays [S, 1, vocab]."""
        probs, packed = self.run(inputs, want_probs=True)
        if probs is None:
            return []
        out = []
        for i in range(len(packed.shapes)):
            lo, hi = packed.head_range(i)
            out.append(np.expand_dims(probs[lo:hi], axis=1))
        return out

    def score_tokens(self, inputs, answer_token_ids):
        """B200-native extra: log-probabilities [n_prompts][S, n_ids] (fp32) of the given next-token
        ids without materialising full-vocabulary probabilities (fused lm_head + log-softmax + gather)."""
        ids = np.asarray(answer_token_ids, dtype=np.int32).reshape(-1)
        lp, packed = self.run(inputs, want_probs=False, answer_ids=ids)
        if lp is None:
            return []
        return [lp[slice(*packed.head_range(i))] for i in range(len(packed.shapes))]

    # ------------------------------------------------------------------ engine
    def _device_inputs(self, packed, dev, kv_mode=_lib.KV_NONE):
        """Token ids, rotary positions, attention tiles and gather lists on the device (pinned -> async H2D)."""
        up = lambda a: torch.from_numpy(a).pin_memory().to(dev, non_blocking=True)
        mbs = packed.micro_batches
        d = dict(ids=up(packed.ids), pos=up(packed.row_pos), tiles=[up(mb.tiles) for mb in mbs],
                 gather=[up(mb.gather) for mb in mbs])
        d["bytes"] = packed.ids.nbytes + packed.row_pos.nbytes + sum(mb.tiles.nbytes + mb.gather.nbytes for mb in mbs)
        if kv_mode == _lib.KV_SAVE:
            d["kv_rows"] = [up(mb.kv_rows) for mb in mbs]
            d["bytes"] += sum(mb.kv_rows.nbytes for mb in mbs)
        return d

    # ------------------------------------------------------------------ prefix K/V cache
    def _kv_policy(self) -> str:
        p = str(getattr(self.args, "prefix_kv_cache", "off"))
        return p if p in ("on", "last") else "off"

    @staticmethod
    def _kv_signature(tok) -> bytes:
        """Identifies a list of prefixes (token ids and lengths, in order)."""
        h = hashlib.blake2b(digest_size=16)
        h.update(np.asarray([t.len_p for t in tok], dtype=np.int64).tobytes())
        for t in tok:
            h.update(t.prefix_ids.tobytes())
        return h.digest()

    def _kv_need(self, layers, prefix_rows: int, act_rows: int, head_rows: int):
        """(bytes the prefix K/V of `layers` would take, whether they fit 80 % of the free HBM once the
        HBM-resident activations and the scores are accounted for)."""
        n_dec = sum(1 for li in layers if self.layer_kinds[li] == _lib.LAYER_DECODER)
        need = n_dec * prefix_rows * 2 * self.kv_dim * 2
        # what the driver reports free plus what torch's allocator holds but has handed back to us
        free = torch.cuda.mem_get_info(self.torch_device)[0] + (torch.cuda.memory_reserved(self.torch_device)
                                                                 - torch.cuda.memory_allocated(self.torch_device))
        act = act_rows * self.hidden * 2 if self.args.storage_location == "gpu" else 0
        return need, need <= 0.8 * (free - act - head_rows * self.vocab * 2)

    def _kv_alloc(self, layers, prefix_rows: int):
        """{layer index: [prefix_rows, 2 * kv_dim] fp16} for the decoder layers among `layers` (one allocation)."""
        dec = [li for li in layers if self.layer_kinds[li] == _lib.LAYER_DECODER]
        big = torch.empty((len(dec), prefix_rows, 2 * self.kv_dim), dtype=torch.float16, device=self.torch_device)
        return {li: big[j] for j, li in enumerate(dec)}

    def _kv_plan(self, tok, layers):
        """Decide what this pass does with the prefix K/V cache.  The reference recomputes prefix and options for
        every generated token (main.py:65-90) although a prompt's prefix rows never change; with
        `args.prefix_kv_cache` = "on" the first pass over a set of prefixes keeps every decoder layer's post-RoPE
        K|V of the prefix rows in HBM (if they fit) and later passes over the same prefixes push only the option
        rows through the layers.  "last" = use a cache if there is one, then drop it, and never start one.
        Returns (kv_mode, {layer: cache rows} or None, note, signature); a cache being filled is registered
        under its signature by `run` only once the pass has succeeded."""
        policy = self._kv_policy()
        if policy == "off":
            return _lib.KV_NONE, None, "off", None
        sig = self._kv_signature(tok)
        if sig in self._kv:
            cache = self._kv[sig] if policy == "on" else self._kv.pop(sig)
            return _lib.KV_LOAD, cache, "load", sig
        if policy == "last":
            return _lib.KV_NONE, None, "off (last pass)", sig
        rows = sum(t.len_p for t in tok)
        sizes = (sum(t.rows for t in tok), sum(t.n_options for t in tok))
        need, fits = self._kv_need(layers, rows, *sizes)
        while not fits and self._kv:                      # make room: the least recently started cache goes first
            self._kv.pop(next(iter(self._kv)))
            need, fits = self._kv_need(layers, rows, *sizes)
        if not fits:
            return _lib.KV_NONE, None, f"skipped: {need / 1e9:.1f} GB of prefix K/V do not fit", sig
        return _lib.KV_SAVE, self._kv_alloc(layers, rows), "save", sig

    def drop_prefix_kv_cache(self):
        self._kv.clear()

    def _run_layers(self, slot, shard, bi, mb, x, heads, dv, out, sp, kv_mode=_lib.KV_NONE, kv_cache=None):
        """Runs the layers of `shard` on micro-batch `bi`.  `x` = token-stage rows (or scored rows when
        the shard starts after model.norm).  Returns (tensor holding the shard's output, is_token_stage)."""
        call = self.ctx.call
        n_total = len(self.layer_names)
        stage_tokens = shard[0] <= n_total - 2
        for li in shard:
            kind = self.layer_kinds[li]
            if kind == _lib.LAYER_EMBED:
                call("fls_embed", slot, li, dv["ids"][mb.row_lo:].data_ptr(), mb.n_rows, x.data_ptr(), sp)
            elif kind == _lib.LAYER_DECODER and kv_mode != _lib.KV_NONE:
                rows = dv["kv_rows"][bi].data_ptr() if kv_mode == _lib.KV_SAVE else None
                call("fls_decoder_layer_kv", slot, li, x.data_ptr(), mb.n_rows, dv["pos"][mb.row_lo:].data_ptr(),
                     dv["tiles"][bi].data_ptr(), int(mb.tiles.shape[0]), kv_mode, rows, mb.n_kv,
                     kv_cache[li][mb.kv_lo:].data_ptr(), sp)
            elif kind == _lib.LAYER_DECODER:
                call("fls_decoder_layer", slot, li, x.data_ptr(), mb.n_rows, dv["pos"][mb.row_lo:].data_ptr(),
                     dv["tiles"][bi].data_ptr(), int(mb.tiles.shape[0]), sp)
            elif kind == _lib.LAYER_NORM:
                call("fls_final_norm", slot, li, x.data_ptr(), dv["gather"][bi].data_ptr(), mb.n_head,
                     heads.data_ptr(), sp)
                x, stage_tokens = heads, False
            else:
                pd = out["probs"][mb.head_lo:].data_ptr() if out["probs"] is not None else None
                ad = out["aid"][mb.head_lo:].data_ptr() if out["n_ids"] else None
                ld = out["lp"][mb.head_lo:].data_ptr() if out["n_ids"] else None
                call("fls_lm_head", slot, li, x.data_ptr(), mb.n_head, pd, ad, out["n_ids"], ld, None, sp)
        return x, stage_tokens

    def _alloc_outputs(self, packed, dev, want_probs, answer_ids):
        n_head = packed.total_head_rows
        out = dict(probs=None, aid=None, lp=None, n_ids=0 if answer_ids is None else int(len(answer_ids)))
        if want_probs:
            out["probs"] = torch.empty((n_head, self.vocab), dtype=torch.float16, device=dev)
        if out["n_ids"]:
            aid = np.tile(np.asarray(answer_ids, dtype=np.int32), (n_head, 1))
            out["aid"] = torch.from_numpy(aid).pin_memory().to(dev, non_blocking=True)
            out["lp"] = torch.empty((n_head, out["n_ids"]), dtype=torch.float32, device=dev)
        return out

    def _fetch_result(self, out, want_probs, stream, return_device):
        src = out["probs"] if want_probs else out["lp"]
        if return_device:
            return src, 0
        # page-locked result buffer, kept across calls (pinning 262 MB costs ~0.1 s)
        key = (tuple(src.shape), src.dtype)
        if getattr(self, "_host_out_key", None) != key:
            self._host_out = torch.empty(src.shape, dtype=src.dtype).pin_memory()
            self._host_out_key = key
        self._host_out.copy_(src, non_blocking=True)
        stream.synchronize()
        return self._host_out.numpy().copy(), src.numel() * src.element_size()

    def run(self, inputs, want_probs=True, answer_ids=None, packed=None, return_device=False):
        shards, num_gpu, rank, dp = self._plan()
        if not dp:
            if self.args.storage_location != "gpu" and not getattr(self, "_mp_tier_noted", False):
                self._mp_tier_noted = True
                warnings.warn(f"--storage_location {self.args.storage_location}: the layer pipeline across GPUs hands "
                              "activations from GPU to GPU and keeps them in HBM; the host and disk tiers apply to "
                              "data-parallel passes (--data_parallel True)", stacklevel=3)
            if getattr(self.device_manager, "distributed", False):    # one process per GPU (torchrun)
                from .pipeline_dist import run_model_parallel_distributed
                return run_model_parallel_distributed(self, inputs, shards, num_gpu, rank, want_probs, answer_ids,
                                                      return_device)
            from .pipeline import run_model_parallel
            return run_model_parallel(self, inputs, shards, num_gpu, rank, want_probs, answer_ids, return_device)
        t0 = perf_counter()
        live = [s for s in shards if len(s)]
        if packed is None:
            if len(inputs) and live:
                # The first shard's weights (the embedding table) start their way to the GPU before the prompts are
                # tokenised: the copy - and on a first pass the file read behind it - runs under the tokeniser.
                with torch.cuda.device(self.torch_device):
                    self._configure(shards)
                    self._stage_shard(live, 0)
            tok = planner.tokenize_prompts(self.tokenizer, inputs)
            for i in range(len(inputs)):
                inputs[i] = None                                  # utils.py:259
            if not tok:
                return None, None
            self._configure(shards)
            kv_mode, kv_cache, kv_note, kv_sig = self._kv_plan(tok, range(len(self.layer_names)))
            packed = planner.pack_prompts(tok, self.max_rows, self.max_head_rows, self.prefix_causal,
                                          cached_prefix=kv_mode == _lib.KV_LOAD,
                                          bucketed=self.pack_order == "length" and self._kv_policy() == "off")
        elif packed.cached_prefix:
            raise ValueError("a pre-packed pass cannot use the prefix K/V cache")
        else:
            kv_mode, kv_cache, kv_note, kv_sig = _lib.KV_NONE, None, "off", None
        t_tok = perf_counter() - t0
        self._configure(shards)
        dev = self.torch_device
        n_layers_total = len(self.layer_names)
        call = self.ctx.call
        with torch.cuda.device(dev), self._on_compute_stream(dev) as stream:
            sp = C.c_void_p(stream.cuda_stream)
            dv = self._device_inputs(packed, dev, kv_mode)
            mbs = packed.micro_batches
            rank_tag = self._rank_tag(rank)
            store = ActivationStore(self.args.storage_location, packed, self.hidden, dev, stream,
                                    disk_folder=getattr(self.args, "disk_folder", "./temp"), rank_tag=rank_tag,
                                    hbm_budget_bytes=self.hbm_activation_budget, force_tier=self.force_tier,
                                    pinned_pool=self._pinned_pool)
            out = self._alloc_outputs(packed, dev, want_probs, answer_ids)
            # Host/disk activation tiers copy rows host->device on the same copy engine as the weight stream, which
            # serves requests in arrival order: pace the stream (a slice per micro-batch) so those copies interleave.
            paced = store.tier != "gpu"
            call("fls_set_option", b"paced_weight_stream", int(paced))
            try:
                if live:
                    self._stage_shard(live, 0)
                for si, shard in enumerate(live):
                    slot = si % self._n_slots
                    nxt_slot, nxt_slice = None, 0
                    if si + 1 < len(live):
                        nxt_slot = self._stage_shard(live, si + 1)   # H2D of shard si+1 overlaps this shard's compute
                        nxt_slice = sum(self._layer_bytes[i] for i in live[si + 1]) // max(1, len(mbs)) + (1 << 20)
                    call("fls_acquire_shard", slot, sp)
                    has_input = shard[0] != 0
                    tokens_in = shard[0] <= n_layers_total - 2      # input is token-stage rows (incl. norm layer)
                    nxt_tokens_in = si + 1 < len(live) and live[si + 1][0] <= n_layers_total - 2
                    for bi, mb in enumerate(mbs):
                        if tokens_in:
                            if bi + 1 < len(mbs):
                                store.prefetch(bi + 1, has_input)
                            elif nxt_tokens_in:
                                store.prefetch(len(mbs), True)       # next shard's first micro-batch
                            x = store.fetch(bi, has_input)
                        else:
                            x = store.heads_view(bi)
                        if paced and nxt_slot is not None:
                            call("fls_prefetch_pump", nxt_slot, C.c_size_t(nxt_slice), sp)
                        y, still_tokens = self._run_layers(slot, shard, bi, mb, x, store.heads_view(bi), dv, out, sp,
                                                           kv_mode, kv_cache)
                        if tokens_in:
                            store.store(bi, x, keep_tokens=still_tokens)
                    store.end_shard()
                    call("fls_release_shard", slot, sp)
            except BaseException:
                store.cleanup(abort=True)
                raise
            finally:
                call("fls_set_option", b"paced_weight_stream", 0)

            result, out_bytes = None, 0
            if any(n_layers_total - 1 in s for s in live):
                result, out_bytes = self._fetch_result(out, want_probs, stream, return_device)
            else:
                stream.synchronize()
            store.cleanup()
            if kv_mode == _lib.KV_SAVE:
                self._kv[kv_sig] = kv_cache              # filled by a pass that ran to its end
            self.last_timing = dict(tokenize_s=t_tok, total_s=perf_counter() - t0, h2d_input_bytes=dv["bytes"],
                                    d2h_output_bytes=out_bytes, activation_tier=store.tier,
                                    act_h2d_bytes=store.h2d_bytes, act_d2h_bytes=store.d2h_bytes,
                                    prefix_kv=kv_note, rows=packed.total_rows)
            self._keepalive = (dv, out, store)
        return result, packed
