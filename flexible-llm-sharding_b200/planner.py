"""Host-side scheduling: shard plan, batching, prompt partitioning, row packing, attention tiles.

Pure numpy/Python (no CUDA) so it is testable on a CPU box.  Mirrors the reference's plans:
  shard_plan          utils.py:143-157   (np.array_split; MP rounds the shard count up to a multiple
                                          of the GPU count and rank r takes shards r::G)
  batch_ranges        main.py:19-20
  partition_prompts   main.py:69-70      (np.array_split of the prompt list across GPUs)
and adds what the packed B200 path needs: every prompt's `len_p + S*len_s` token rows are laid out
contiguously so that one GEMM covers a whole micro-batch, with per-row rotary positions and
per-tile attention descriptors replacing the reference's per-prompt mask/KV-expand logic
(utils.py:219-221,275-277).
"""
from __future__ import annotations

import heapq
from dataclasses import dataclass, field

import numpy as np

MAX_TOKEN_LEN = 4096     # reference utils.py:14
ATTN_TILE_Q = 128        # max queries per attention tile (csrc/attention_sm100.cu BQ)


def shard_plan(n_layers_total: int, layer_num_per_shard: int, data_parallel: bool,
               num_gpu: int = 1, gpu_rank: int = 0) -> list[tuple[int, ...]]:
    """Layer indices per shard for this rank; `n_layers_total` = decoder layers + 3."""
    n_min = -(-n_layers_total // layer_num_per_shard)
    if data_parallel:
        n_shards = n_min
        mine = slice(None)
    else:
        n_shards = -(-n_min // num_gpu) * num_gpu
        mine = slice(gpu_rank, None, num_gpu)
    # np.array_split semantics: the first (n % k) parts get one extra element; parts may be empty
    base, extra = divmod(n_layers_total, n_shards)
    shards, start = [], 0
    for i in range(n_shards):
        size = base + (1 if i < extra else 0)
        shards.append(tuple(range(start, start + size)))
        start += size
    return shards[mine]


def batch_ranges(n_prompts: int, num_batch: int) -> list[tuple[int, int]]:
    ends = [n_prompts // num_batch * i for i in range(1, num_batch)] + [n_prompts]
    return list(zip([0] + ends[:-1], ends))


def partition_prompts(n_prompts: int, world: int) -> list[tuple[int, int]]:
    """[start, end) of each rank's prompts, np.array_split order (main.py:69-70)."""
    base, extra = divmod(n_prompts, world)
    out, start = [], 0
    for r in range(world):
        size = base + (1 if r < extra else 0)
        out.append((start, start + size))
        start += size
    return out


@dataclass
class TokenizedPrompt:
    prefix_ids: np.ndarray   # [len_p] int32, BOS first
    suffix_ids: np.ndarray   # [S, len_s] int32, BOS stripped, right-padded with pad(=eos)
    suffix_eos: np.ndarray   # [S] index of the last non-pad token (may be -1: wraps like torch indexing)

    @property
    def len_p(self): return int(self.prefix_ids.shape[0])
    @property
    def n_options(self): return int(self.suffix_ids.shape[0])
    @property
    def len_s(self): return int(self.suffix_ids.shape[1])
    @property
    def rows(self): return self.len_p + self.n_options * self.len_s


def tokenize_prompts_serial(tokenizer, prompts) -> list[TokenizedPrompt]:
    """utils.py:246-258 literally: two tokenizer calls per prompt (kept as the checked-against form)."""
    out = []
    pad = tokenizer.pad_token_id
    for prefix, suffixes in prompts:
        p = tokenizer(prefix, return_tensors="np", return_attention_mask=False, truncation=True,
                      max_length=MAX_TOKEN_LEN)["input_ids"][0]
        s = tokenizer(list(suffixes), return_tensors="np", return_attention_mask=False, truncation=True,
                      max_length=MAX_TOKEN_LEN, padding=True)["input_ids"][:, 1:]
        eos = (s != pad).sum(1) - 1
        out.append(TokenizedPrompt(np.ascontiguousarray(p, dtype=np.int32),
                                   np.ascontiguousarray(s, dtype=np.int32), eos.astype(np.int64)))
    return out


def tokenize_prompts(tokenizer, prompts) -> list[TokenizedPrompt]:
    """Same result as `tokenize_prompts_serial`, but every prefix and every option goes through the
    tokenizer in two batched calls (the Rust tokenizers parallelise a batch over the host cores; the
    reference tokenises one prompt at a time inside its GPU loop, utils.py:246-258).  Padding is
    redone per prompt: options are right-padded with the pad id to the prompt's longest option and
    the leading (BOS) column is dropped, exactly what `padding=True` followed by `[:, 1:]` yields."""
    prompts = list(prompts)
    if not prompts:
        return []
    pad = tokenizer.pad_token_id
    pre_txt, flat_txt = [p for p, _ in prompts], [s for _, ss in prompts for s in ss]
    backend = getattr(tokenizer, "backend_tokenizer", None) if getattr(tokenizer, "is_fast", False) else None
    if backend is not None:
        # what `tokenizer(texts, truncation=True, max_length=...)` does underneath, minus the per-item Python
        # wrapping of the results (a fifth of the tokenisation time at 1024 prompts)
        backend.enable_truncation(max_length=MAX_TOKEN_LEN, stride=0, strategy="longest_first",
                                  direction=getattr(tokenizer, "truncation_side", "right"))
        backend.no_padding()
        # `encode_batch_fast` (tokenizers >= 0.20) skips the character offsets nobody reads here: same ids, a third faster
        encode = getattr(backend, "encode_batch_fast", None) or backend.encode_batch
        pre = [e.ids for e in encode(pre_txt, add_special_tokens=True)]
        flat = [e.ids for e in encode(flat_txt, add_special_tokens=True)]
    else:
        kw = dict(return_attention_mask=False, truncation=True, max_length=MAX_TOKEN_LEN)
        pre = tokenizer(pre_txt, **kw)["input_ids"]
        flat = tokenizer(flat_txt, **kw)["input_ids"]
    out, k = [], 0
    for i, (_, ss) in enumerate(prompts):
        S = len(ss)
        opts = flat[k:k + S]
        k += S
        width = max(len(o) for o in opts)
        s = np.full((S, width), pad, dtype=np.int32)
        for j, o in enumerate(opts):
            s[j, :len(o)] = o
        s = np.ascontiguousarray(s[:, 1:])
        eos = (s != pad).sum(1) - 1
        out.append(TokenizedPrompt(np.asarray(pre[i], dtype=np.int32), s, eos.astype(np.int64)))
    return out


@dataclass
class MicroBatch:
    prompt_lo: int
    prompt_hi: int
    row_lo: int              # first packed row (global)
    n_rows: int
    head_lo: int             # first scored row (global) = sum of S over earlier prompts
    n_head: int
    tiles: np.ndarray        # [n_tiles, 8] int32, rows relative to row_lo
    gather: np.ndarray       # [n_head] int32 rows (relative to row_lo) of each option's last real token
    kv_lo: int = 0           # first prefix row of its prompts in the pass-wide prefix K/V cache (prompt order)
    n_kv: int = 0            # prefix rows of its prompts
    kv_rows: np.ndarray | None = None   # [n_kv] int32: those rows relative to row_lo (None when the prefix is cached)


@dataclass
class PackedPrompts:
    ids: np.ndarray          # [total_rows] int32 token ids
    row_pos: np.ndarray      # [total_rows] int32 rotary positions
    prompt_row: np.ndarray   # [n_prompts + 1] first row of each prompt
    head_row: np.ndarray     # [n_prompts + 1] first scored row of each prompt
    shapes: list[tuple[int, int, int]]   # (len_p, S, len_s) of the packed ROWS: len_p is 0 when the prefix is cached
    micro_batches: list[MicroBatch] = field(default_factory=list)
    cached_prefix: bool = False          # rows hold the options only; prefix K/V come from a cache (prefix_len rows)
    prefix_len: list[int] = field(default_factory=list)     # real prefix length of every prompt
    # Every per-prompt field above is in PACKED order.  `order[p]` = the caller's index of the prompt packed at
    # position p (None: packed in the caller's order); results are handed back through `head_range`.
    order: np.ndarray | None = None
    _position: np.ndarray | None = None

    @property
    def total_rows(self): return int(self.prompt_row[-1])
    @property
    def total_head_rows(self): return int(self.head_row[-1])

    def prompt_index(self, p: int) -> int:
        """The caller's index of the prompt at packed position `p`."""
        return p if self.order is None else int(self.order[p])

    def head_range(self, i: int) -> tuple[int, int]:
        """Scored rows [lo, hi) of the caller's prompt `i`."""
        if self.order is not None:
            if self._position is None:
                self._position = np.empty(len(self.order), dtype=np.int64)
                self._position[self.order] = np.arange(len(self.order))
            i = int(self._position[i])
        return int(self.head_row[i]), int(self.head_row[i + 1])


ATTN_PREFIX, ATTN_PREFIX_CAUSAL, ATTN_SUFFIX = 0, 1, 2


def prompt_tiles(base: int, len_p: int, S: int, len_s: int, prefix_causal: bool,
                 option_base: int | None = None) -> list[list[int]]:
    """Attention tiles of one prompt whose rows start at `base` (include/fls_b200.h: fls_attn_tile).
    Prefix rows: tiles of <= 128 queries over the prefix keys.  Option rows: as many whole options
    as fit 128 rows share one tile (their common prefix keys are then read once); an option longer
    than 128 rows is cut into 128-row tiles that see the option's rows up to their own last query.
    With `option_base` the prefix is cached: its keys sit at `base`, the option rows start at
    `option_base`, and there are no prefix queries."""
    Q = ATTN_TILE_Q
    t = []
    if option_base is None:
        mode = ATTN_PREFIX_CAUSAL if prefix_causal else ATTN_PREFIX
        for q0 in range(0, len_p, Q):
            t.append([base + q0, min(Q, len_p - q0), q0, mode, base, len_p, 0, 0])
        first = base + len_p
    else:
        first = option_base
    if len_s <= Q:
        per_tile = max(1, Q // len_s)
        for s0 in range(0, S, per_tile):
            n_opt = min(per_tile, S - s0)
            row = first + s0 * len_s
            t.append([row, n_opt * len_s, len_s, ATTN_SUFFIX, base, len_p, row, n_opt * len_s])
    else:
        for s in range(S):
            row = first + s * len_s
            for j0 in range(0, len_s, Q):
                n_q = min(Q, len_s - j0)
                t.append([row + j0, n_q, len_s, ATTN_SUFFIX, base, len_p, row, j0 + n_q])
    return t


def balanced_bins(tok: list[TokenizedPrompt], max_rows: int, max_head_rows: int) -> list[list[int]]:
    """Length-bucketed assignment of prompts to micro-batches.  Packed rows carry no padding
    between prompts, so bucketing by length cannot save rows here; what it buys on ragged data is BALANCE: the
    fewest micro-batches the two caps allow, all about the same size (equal pipeline items across GPUs, no small
    trailing launch), and inside a micro-batch the long prompts first, so the attention grid - dispatched in tile
    order - ends on its cheapest tiles.  Longest-first into the emptiest micro-batch that still has room; another
    micro-batch is opened only when a prompt fits none.  Ties keep the caller's order, so prompts of one common
    shape come out exactly as the in-order packing puts them."""
    n = len(tok)
    rows = [t.rows for t in tok]
    heads = [t.n_options for t in tok]
    n_bins = max(1, -(-sum(rows) // max_rows), -(-sum(heads) // max_head_rows))
    by_len = sorted(range(n), key=lambda i: -rows[i])              # stable: equal lengths stay in caller order
    while True:
        bins = [[] for _ in range(n_bins)]
        n_heads = [0] * n_bins
        heap = [(0, b) for b in range(n_bins)]                     # (rows held, micro-batch): emptiest on top
        ok = True
        for i in by_len:
            aside = []                                             # emptier micro-batches whose head rows are used up
            while heap and heap[0][0] + rows[i] <= max_rows and n_heads[heap[0][1]] + heads[i] > max_head_rows:
                aside.append(heapq.heappop(heap))
            if not heap or heap[0][0] + rows[i] > max_rows:        # the emptiest one left has no room: none has
                ok = False
                break
            filled, b = heapq.heappop(heap)
            bins[b].append(i)
            n_heads[b] += heads[i]
            heapq.heappush(heap, (filled + rows[i], b))
            for item in aside:
                heapq.heappush(heap, item)
        if ok:
            break
        n_bins += 1
        if n_bins > n:
            raise ValueError(f"a prompt does not fit a micro-batch (max_rows {max_rows}, max_head_rows {max_head_rows})")
    bins = [b for b in bins if b]
    if all(rows[i] == rows[0] for i in range(n)):
        # one common shape: the round-robin above deals prompts out like cards; hand them back as contiguous runs
        flat, sizes = list(range(n)), [len(b) for b in bins]
        bins, k = [], 0
        for sz in sizes:
            bins.append(flat[k:k + sz])
            k += sz
    else:
        bins.sort(key=lambda b: min(b))                            # micro-batches in the order of their first prompt
    return bins


def _in_order_count(tok: list[TokenizedPrompt], max_rows: int, max_head_rows: int) -> int:
    """Micro-batches that filling them in the caller's order would make (each takes what fits)."""
    n, rows, heads = 0, max_rows + 1, 0
    for t in tok:
        if rows + t.rows > max_rows or heads + t.n_options > max_head_rows:
            n, rows, heads = n + 1, 0, 0
        rows += t.rows
        heads += t.n_options
    return n


def pack_prompts(tok: list[TokenizedPrompt], max_rows: int, max_head_rows: int,
                 prefix_causal: bool = False, cached_prefix: bool = False, bucketed: bool = False) -> PackedPrompts:
    """Packs the prompts' token rows and cuts them into micro-batches of <= max_rows rows.
    `cached_prefix`: only the option rows are packed (the prefix K/V of every layer come from a cache filled by
    an earlier pass over the same prefixes); a micro-batch's attention then works in the row space
    [its prompts' cached prefix rows | its option rows], which is what max_rows bounds.
    `bucketed`: prompts are assigned to micro-batches by `balanced_bins` instead of in the caller's order
    (`PackedPrompts.order` / `head_range` map results back); not combined with the prefix cache, whose rows are
    laid out in the order of the pass that filled it."""
    order, ranges = None, None
    if bucketed and not cached_prefix and len(tok) > 1:
        for i, t in enumerate(tok):                                # same refusals, with the caller's prompt index
            if t.rows > max_rows:
                raise ValueError(f"prompt {i} has {t.rows} token rows; raise max_rows (currently {max_rows})")
        bins = balanced_bins(tok, max_rows, max_head_rows)
        if len(bins) <= _in_order_count(tok, max_rows, max_head_rows):
            flat = [i for b in bins for i in b]
            if flat != list(range(len(tok))):
                order = np.asarray(flat, dtype=np.int64)
                tok = [tok[i] for i in flat]
            ranges, k = [], 0
            for b in bins:
                ranges.append((k, k + len(b)))
                k += len(b)
        # else: longest-first into the emptiest micro-batch is a heuristic; the rare prompt mix it packs into MORE
        # micro-batches than filling them in order does is packed in order
    n = len(tok)
    prompt_row = np.zeros(n + 1, dtype=np.int64)
    head_row = np.zeros(n + 1, dtype=np.int64)
    kv_row = np.zeros(n + 1, dtype=np.int64)
    for i, t in enumerate(tok):
        who = i if order is None else int(order[i])
        if t.len_p + t.len_s > MAX_TOKEN_LEN:
            # the reference would fail slicing its 4096^2 mask (utils.py:276)
            raise ValueError(f"prompt {who}: prefix + option length {t.len_p + t.len_s} exceeds {MAX_TOKEN_LEN}")
        if t.rows > max_rows:
            raise ValueError(f"prompt {who} has {t.rows} token rows; raise max_rows (currently {max_rows})")
        if t.len_s == 0:
            raise ValueError(f"prompt {who}: every option is empty")
        prompt_row[i + 1] = prompt_row[i] + (t.rows - t.len_p if cached_prefix else t.rows)
        head_row[i + 1] = head_row[i] + t.n_options
        kv_row[i + 1] = kv_row[i] + t.len_p
    total = int(prompt_row[-1])
    ids = np.empty(total, dtype=np.int32)
    row_pos = np.empty(total, dtype=np.int32)
    shapes = []
    for i, t in enumerate(tok):
        so = int(prompt_row[i])
        if not cached_prefix:
            ids[so:so + t.len_p] = t.prefix_ids
            row_pos[so:so + t.len_p] = np.arange(t.len_p)
            so += t.len_p
        ids[so:so + t.n_options * t.len_s] = t.suffix_ids.reshape(-1)
        row_pos[so:so + t.n_options * t.len_s] = np.tile(np.arange(t.len_p, t.len_p + t.len_s), t.n_options)
        shapes.append((0 if cached_prefix else t.len_p, t.n_options, t.len_s))
    packed = PackedPrompts(ids, row_pos, prompt_row, head_row, shapes, cached_prefix=cached_prefix,
                           prefix_len=[t.len_p for t in tok], order=order)

    def space(lo, hi):      # rows of the attention workspace prompts [lo, hi) need
        return int(prompt_row[hi] - prompt_row[lo]) + (int(kv_row[hi] - kv_row[lo]) if cached_prefix else 0)

    if ranges is None:                                             # in order: each micro-batch takes what fits
        ranges, lo = [], 0
        while lo < n:
            hi = lo
            while hi < n and space(lo, hi + 1) <= max_rows and head_row[hi + 1] - head_row[lo] <= max_head_rows:
                hi += 1
            if hi == lo:
                raise ValueError(f"prompt {lo} does not fit a micro-batch (max_rows {max_rows}, max_head_rows {max_head_rows})")
            ranges.append((lo, hi))
            lo = hi
    for lo, hi in ranges:
        row_lo = int(prompt_row[lo])
        kv_lo, n_kv = int(kv_row[lo]), int(kv_row[hi] - kv_row[lo])
        tiles, gather, kv_rows = [], [], []
        for i in range(lo, hi):
            len_p, S, len_s = tok[i].len_p, tok[i].n_options, tok[i].len_s
            base = int(prompt_row[i]) - row_lo
            if cached_prefix:
                first = base                                        # x rows: options only
                tiles += prompt_tiles(int(kv_row[i]) - kv_lo, len_p, S, len_s, prefix_causal, option_base=n_kv + base)
            else:
                first = base + len_p
                tiles += prompt_tiles(base, len_p, S, len_s, prefix_causal)
                kv_rows.append(np.arange(base, base + len_p, dtype=np.int32))
            for s in range(S):
                e = int(tok[i].suffix_eos[s])
                if e < 0:
                    e += len_s          # torch negative indexing at utils.py:286
                gather.append(first + s * len_s + e)
        packed.micro_batches.append(MicroBatch(
            lo, hi, row_lo, int(prompt_row[hi]) - row_lo, int(head_row[lo]), int(head_row[hi] - head_row[lo]),
            np.asarray(tiles, dtype=np.int32).reshape(-1, 8), np.asarray(gather, dtype=np.int32), kv_lo, n_kv,
            None if cached_prefix else (np.concatenate(kv_rows) if kv_rows else np.zeros(0, np.int32))))
    return packed


def rope_tables(n_pos: int, head_dim: int, theta: float = 10000.0):
    """fp16(cos), fp16(sin) [n_pos, head_dim], built the way the reference's rotary buffers are
    (transformers 4.3x `_set_cos_sin_cache` in fp32, cast to fp16 by utils.py:117-119)."""
    import torch
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).to(torch.float32) / head_dim))
    cos = torch.empty((n_pos, head_dim), dtype=torch.float16)
    sin = torch.empty((n_pos, head_dim), dtype=torch.float16)
    # Evaluated in row blocks small enough for torch to stay on one thread: the vectorised fp32 cos/sin pick their
    # range-reduction path per SIMD vector, so a table cut into per-thread chunks at non-row boundaries can differ
    # from a row-aligned evaluation by one fp32 ulp (and, rarely, one fp16 ulp) depending on the host's core count.
    block = max(1, 16384 // head_dim)
    for lo in range(0, n_pos, block):
        pos = torch.arange(lo, min(lo + block, n_pos), dtype=torch.int64).to(torch.float32)
        freqs = pos[:, None] * inv_freq[None, :]
        emb = torch.cat((freqs, freqs), dim=-1)
        cos[lo:lo + emb.shape[0]] = emb.cos().to(torch.float16)
        sin[lo:lo + emb.shape[0]] = emb.sin().to(torch.float16)
    return cos, sin
