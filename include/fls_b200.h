/* fls_b200 - C ABI of the B200-native layer-sharded prompt-scoring path.
 *
 * The reference (eyedealism/flexible-LLM-sharding) is pure Python and has no FFI of its own; this
 * header is the boundary a maintainer binds (ctypes stub in INTEGRATION.md) to replace the hot
 * path of `ShardedLlama.__call__` (reference utils.py:133-305).  Each entry point names the
 * reference code it stands in for.
 *
 * Conventions
 *  - plain C types only: pointers, sizes, ints.  `stream` arguments are `cudaStream_t` passed as
 *    `void*` (NULL = the legacy default stream).  Pointers ending in `_dev` are device pointers,
 *    `_host` host pointers.  All activations are fp16, row-major, packed token rows `[rows, width]`.
 *  - one context per GPU; a context is used by one host thread at a time (the reference runs one
 *    Python thread per GPU, main.py:68-73).  Different contexts may be driven concurrently; the
 *    pinned host copies of layer files are shared between all contexts of the process
 *    (the analogue of DeviceManager's read-once state dict, utils.py:35-41,67-75).
 *  - every function returns 0 on success or a negative FLS_ERR_* code and never throws or exits;
 *    `fls_last_error` returns the message of the last failure on that context (or, with a NULL
 *    context, of the calling thread).
 *  - the caller owns every activation / id / output buffer.  The library owns the pinned layer
 *    images, the HBM weight slots and its workspace, all allocated in `fls_create` /
 *    `fls_configure` / `fls_register_layer_file`; nothing is allocated by the compute calls.
 */
#ifndef FLS_B200_H
#define FLS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLS_OK 0
#define FLS_ERR_INVALID (-1)  /* bad argument */
#define FLS_ERR_CUDA (-2)     /* CUDA runtime / driver error (sticky for the context) */
#define FLS_ERR_IO (-3)       /* file missing, short read, malformed safetensors header */
#define FLS_ERR_STATE (-4)    /* call sequence error (e.g. compute on a slot that holds no such layer) */
#define FLS_ERR_CAPACITY (-5) /* request exceeds what fls_configure reserved */

typedef struct fls_ctx fls_ctx;

/* Llama-2 architecture numbers = the fields of config.json the reference reads through
 * AutoConfig (utils.py:101). */
typedef struct fls_model_desc {
    int32_t hidden_size;
    int32_t intermediate_size;
    int32_t num_attention_heads;
    int32_t num_key_value_heads;
    int32_t head_dim; /* 64 or 128 */
    int32_t vocab_size;
    int32_t num_hidden_layers;
    float rms_norm_eps;
} fls_model_desc;

/* Kind of a layer file, in the reference's layer order (utils.py:106-107,114-115). */
#define FLS_LAYER_EMBED 0   /* model.embed_tokens.safetensors */
#define FLS_LAYER_DECODER 1 /* model.layers.N.safetensors */
#define FLS_LAYER_NORM 2    /* model.norm.safetensors */
#define FLS_LAYER_HEAD 3    /* lm_head.safetensors */

/* One attention query tile (<= 128 consecutive query rows of one prompt); built by the host
 * scheduler.  Replaces the reference's per-prompt kv_cache expand + mask slice (utils.py:275-277).
 * Keys/values are read in place from the packed qkv rows: segment 0 = the prompt's prefix rows
 * (shared by every option), segment 1 = the rows of the options the queries belong to.
 *   mode 0  prefix queries, every prefix key visible   (reference prefix pass: attention_mask=None,
 *           utils.py:272)
 *   mode 1  prefix queries, causal: key t visible iff t <= q_seq0 + i (q_seq0 = index of the first query)
 *   mode 2  option queries: the tile covers rows of whole options of length q_seq0 (= len_s) that
 *           start at seg1_row; a query sees all prefix keys and the keys of its own option at
 *           positions <= its own (== the mask slice of utils.py:276 with position offset len_p).
 * Every tile must see at least one key (seg0_len + seg1_len > 0; the host scheduler never builds another kind): the
 * output rows of a tile without keys are not defined. */
#define FLS_ATTN_PREFIX 0
#define FLS_ATTN_PREFIX_CAUSAL 1
#define FLS_ATTN_SUFFIX 2
typedef struct fls_attn_tile {
    int32_t q_row;    /* first query row */
    int32_t n_q;      /* 1..128 */
    int32_t q_seq0;   /* mode 1: sequence index of the first query; mode 2: option length */
    int32_t mode;     /* FLS_ATTN_* */
    int32_t seg0_row; /* key/value rows of the shared prefix */
    int32_t seg0_len;
    int32_t seg1_row; /* mode 2: first row of the covered options */
    int32_t seg1_len; /*         rows of those options that can be visible to the tile */
} fls_attn_tile;

typedef struct fls_stats {
    uint64_t kernel_launches; /* kernels this context has launched */
    uint64_t h2d_bytes;       /* weight bytes copied pinned host -> HBM slots */
    double h2d_ms;            /* device time of those copies (copy-stream events) */
    uint64_t file_bytes_read; /* bytes read from layer files into pinned memory (process-wide) */
    uint64_t pinned_bytes;    /* pinned host bytes currently held (process-wide) */
    uint64_t workspace_bytes; /* HBM held by this context (slots + workspace) */
    uint64_t p2p_bytes;       /* activation bytes handed to other GPUs (fls_peer_copy) */
    uint64_t h2d_copies;      /* cudaMemcpyAsync calls the weight stream was cut into */
    uint64_t mapped_bytes;    /* part of pinned_bytes that is file pages locked in place, shared with every
                                 other process that maps the same files (process-wide) */
} fls_stats;

/* ---------------------------------------------------------------------------------------------
 * lifecycle - replaces ShardedLlama.__init__/init_model (utils.py:80-119): instead of a meta-device
 * model rebuilt on every call (utils.py:135-137) a context is created once per GPU. */
int fls_version(void);
const char* fls_last_error(const fls_ctx* ctx);
int fls_create(const fls_model_desc* desc, int device, fls_ctx** out);
int fls_destroy(fls_ctx* ctx);

/* Reserve HBM: `n_slots` (1..512) weight slots of `slot_bytes` each (2 = double buffering: shard k+1 is
 * copied while shard k computes; one per shard = resident weights) and workspace for micro-batches of up to `max_rows` token rows
 * and `max_head_rows` scored rows.  May be called again to resize (frees the old reservation). */
int fls_configure(fls_ctx* ctx, int n_slots, size_t slot_bytes, int max_rows, int max_head_rows);

/* Rotary tables [n_pos, head_dim] fp16 (host pointers), built by the host exactly as the
 * reference's rotary buffers are built and cast (utils.py:117-119): fp16(cos(fp32 pos * inv_freq)). */
int fls_set_rope_table(fls_ctx* ctx, const uint16_t* cos_host, const uint16_t* sin_host, int n_pos);

/* ---------------------------------------------------------------------------------------------
 * weight stream - replaces load_layer_to_gpu (utils.py:121-131) and DeviceManager's DP state
 * (utils.py:46-75): file -> pinned host image (read once per process, shared by all contexts)
 * -> asynchronous H2D into an HBM slot on the context's copy stream. */

/* Parse `<layer>.safetensors` written by prepare_weights.py (8-byte LE header length, JSON
 * header, raw tensors) and bind it to `layer_idx`.  The file is read into pinned memory at the
 * first prefetch that needs it (or now, if `load_now`).  F16 tensors are used as stored; BF16/F32
 * are converted to fp16 once on load (the reference casts to fp16 too, utils.py:130); integer
 * dtypes are rejected like utils.py:129.  A path ending in `.flspack` names a packed layer file
 * (prepare_weights.py --packed: the same tensors as fp16 in slot order behind a 4096-byte header); it is read as
 * ONE page-aligned range (O_DIRECT from a disk, locked in place on tmpfs) and uploaded as one run of pieces. */
int fls_register_layer_file(fls_ctx* ctx, int layer_idx, int kind, const char* path, int load_now);
/* HBM bytes layer `layer_idx` needs in a slot. */
int fls_layer_slot_bytes(fls_ctx* ctx, int layer_idx, size_t* out_bytes);
/* Enqueue the H2D copies of `n` layers into `slot` on the copy stream.  Waits (on the device) for
 * the slot's previous user to have released it. */
int fls_prefetch_shard(fls_ctx* ctx, int slot, const int32_t* layer_ids, int n);
/* Paced weight stream (option "paced_weight_stream" = 1): fls_prefetch_shard only prepares the shard's copies and
 * fls_prefetch_pump enqueues the next `max_bytes` of them; fls_acquire_shard enqueues whatever is left.  The copy
 * engine serves requests in arrival order, so a whole shard enqueued at once would make every later host->device
 * copy of the caller (activation tiers) wait for it; pumping a slice per micro-batch interleaves the two.  With
 * `after_stream` non-NULL the slice is additionally held back until everything enqueued on that stream so far has
 * run, so the pacing follows the GPU's progress rather than the host's. */
int fls_prefetch_pump(fls_ctx* ctx, int slot, size_t max_bytes, void* after_stream);
/* Make `stream` wait until the slot's copies have landed (no host blocking).  A slot that was released
 * but not overwritten still holds its layers and may be acquired again without a new prefetch: with one
 * slot per shard the weights stay resident in HBM across passes (the reference re-reads
 * every layer per batch and per generated token, main.py:22-23,65). */
int fls_acquire_shard(fls_ctx* ctx, int slot, void* stream);
/* Mark the slot reusable once everything enqueued on `stream` so far has finished.  Replaces
 * `layer.to("meta")` + clean_memory (utils.py:299-302). */
int fls_release_shard(fls_ctx* ctx, int slot, void* stream);
/* Read a layer file into the process-wide pinned cache without binding it to a context: the
 * analogue of DeviceManager.send_device_request/clean_waiting_queue (utils.py:61-75).  Safe to call
 * from a reader thread running ahead of the GPU loop. */
int fls_preload_layer_file(const char* path);
/* Bound the pinned host memory of the process-wide layer cache (0 = unlimited, the default).  Beyond the bound
 * the least recently used images that no slot is copying from are dropped and re-read from their files when
 * needed again - the analogue of DeviceManager's max_layer_num (utils.py:31,70) for hosts that cannot hold the
 * whole model in RAM. */
int fls_set_host_cache_limit(size_t max_bytes);
/* Drop pinned images no context needs any more (process-wide cache). */
int fls_trim_host_cache(size_t keep_bytes);
/* How a layer file becomes a page-locked host image (reference utils.py:72-73,126-127: `f.read()` into a bytes
 * object, then safetensors parses a second copy).  0 (default, or FLS_WEIGHT_IMAGE=auto): files whose tensors are
 * all stored as F16 and that lie on tmpfs (/dev/shm; writable by the caller) are mmap'ed and their own pages are
 * locked in place with cudaHostRegister - no second copy in RAM, and all processes of the box that map the file
 * share one physical copy (DeviceManager's read-once layer, utils.py:35-41, across processes); BF16/F32 files,
 * files on a disk file system (the driver refuses to lock their pages) take the copying path (parallel pread into cudaHostAlloc memory, converted to fp16,
 * the file's pages then dropped from the page cache).  1 (FLS_WEIGHT_IMAGE=copy): always copy.  2 (=map): map
 * or fail with FLS_ERR_IO. */
int fls_set_host_image_mode(int mode);
/* Staging directory (on tmpfs) for layer files that lie on a disk file system: the first process of the box that
 * needs a file copies it there once (read past the page cache, lock file + atomic rename; the name carries the
 * source's size and modification time), every process then locks THAT copy's pages in place - DeviceManager's "read
 * each layer to RAM once and hand it to every GPU" (utils.py:35-41,67-75) for ranks that are processes.  NULL or ""
 * switches staging off (default; each process then holds its own pinned copy).  The directory and its files are the
 * caller's to remove. */
int fls_set_stage_dir(const char* dir);

/* ---------------------------------------------------------------------------------------------
 * layer compute on a resident shard.  `x_dev` etc. are fp16 [rows, hidden]. */

/* utils.py:266-268  prefix/suffix = embed_tokens(ids), for all packed rows at once. */
int fls_embed(fls_ctx* ctx, int slot, int layer_idx, const int32_t* ids_dev, int n_rows, uint16_t* out_dev,
              void* stream);
/* utils.py:269-279  one LlamaDecoderLayer over every packed row of the micro-batch, in place:
 * RMSNorm -> QKV GEMM -> RoPE -> shared-prefix attention -> O GEMM(+residual) -> RMSNorm ->
 * gate/up GEMM(+SwiGLU) -> down GEMM(+residual).  `row_pos_dev[r]` = rotary position of row r
 * (0..len_p-1 for prefix rows, len_p.. for option rows, utils.py:221,275). */
int fls_decoder_layer(fls_ctx* ctx, int slot, int layer_idx, uint16_t* x_dev, int n_rows, const int32_t* row_pos_dev,
                      const fls_attn_tile* tiles_dev, int n_tiles, void* stream);
/* The same layer with a prefix K/V cache.  No counterpart in the reference: its generation loop (main.py:65-90)
 * re-runs prefix and options from scratch for every new token, although the prefix rows never change
 * (utils.py:272: the prefix attends only to itself).  The cache is caller-owned fp16 memory
 * [n_kv_rows, 2 * kv_heads * head_dim] per layer: post-RoPE keys, then values.
 *   FLS_KV_SAVE  a normal pass over x_dev[n_rows]; afterwards row kv_rows_dev[j] (an index into the rows of this
 *                call, normally the prefix rows in prompt order) has its K|V copied to cache row j.
 *   FLS_KV_LOAD  x_dev holds only the n_rows NEW rows (the options).  The layer works in a row space of
 *                n_kv_rows + n_rows rows: rows [0, n_kv_rows) are the cached rows, whose K|V are restored from
 *                the cache (kv_rows_dev is ignored), rows n_kv_rows.. are x_dev's rows.  Tile descriptors index
 *                that space, and every query tile must start at or after n_kv_rows; row_pos_dev has n_rows
 *                entries.  n_kv_rows + n_rows must not exceed the configured max_rows.
 *   FLS_KV_NONE  identical to fls_decoder_layer.
 * Results are bit-identical to recomputing the cached rows (every kernel treats rows independently). */
#define FLS_KV_NONE 0
#define FLS_KV_SAVE 1
#define FLS_KV_LOAD 2
int fls_decoder_layer_kv(fls_ctx* ctx, int slot, int layer_idx, uint16_t* x_dev, int n_rows,
                         const int32_t* row_pos_dev, const fls_attn_tile* tiles_dev, int n_tiles, int kv_mode,
                         const int32_t* kv_rows_dev, int n_kv_rows, uint16_t* kv_cache_dev, void* stream);
/* utils.py:281-286  out[j] = norm(x[gather_dev[j]]) - the last real token of every option. */
int fls_final_norm(fls_ctx* ctx, int slot, int layer_idx, const uint16_t* x_dev, const int32_t* gather_dev, int n_out,
                   uint16_t* out_dev, void* stream);
/* utils.py:287-291  softmax(lm_head(x)) without storing logits: pass 1 = GEMM with a per-tile
 * (max, sum-exp) epilogue + reduction -> log-sum-exp per row; then, as requested,
 *   probs_dev   [n_rows, vocab] fp16 probabilities (the reference's output format), GEMM pass 2 with
 *               an exp(logit - lse) epilogue; may be NULL;
 *   logprob_dev [n_rows, n_ids] fp32 log-probabilities of `answer_ids_dev[n_rows, n_ids]` (an id outside
 *               [0, vocab) gives NaN in its entry, never a read outside the weight); may be NULL;
 *   lse_dev     [n_rows] fp32; may be NULL. */
int fls_lm_head(fls_ctx* ctx, int slot, int layer_idx, const uint16_t* x_dev, int n_rows, uint16_t* probs_dev,
                const int32_t* answer_ids_dev, int n_ids, float* logprob_dev, float* lse_dev, void* stream);

/* Model-parallel hand-off (utils.py:163,185,193-195: the reference passes activations between GPU
 * threads through a shared dict and `.to(self.device)`): copy `bytes` from this context's GPU
 * straight into a buffer on GPU `dst_device` over NVLink (cudaMemcpyPeerAsync on `stream`, peer
 * access enabled on first use).  Ordering with the consumer is the caller's (events).  `dst_device` < 0:
 * `dst_dev` is a mapping obtained from fls_ipc_open (another process's buffer). */
int fls_peer_copy(fls_ctx* ctx, void* dst_dev, int dst_device, const void* src_dev, size_t bytes, void* stream);

/* Model-parallel hand-off between PROCESSES (one rank per GPU under torchrun).  The reference's pipeline
 * stages are threads of one process that share a Python dict, find out that a prompt is ready by polling
 * `prompt2layer` in sleep(1) loops (utils.py:179-180,189-190) and bound the stages' run-ahead with
 * `max_activation_in_cpu` (utils.py:178-180).  Across processes the same three things are:
 *   - the consumer's inbound ring of micro-batch slots, exported once: fls_ipc_alloc (cudaMalloc, zero-filled,
 *     + cudaIpcGetMemHandle; the 64 handle bytes travel over any host channel) and mapped by its producer:
 *     fls_ipc_open (peer access over NVLink is enabled by the mapping); rows then go across with
 *     fls_peer_copy(dst_device = -1);
 *   - readiness: a 32-bit counter in the WAITING GPU's memory.  fls_flag_signal enqueues, on `stream`, a
 *     system-scope release store of `value` (everything the stream did before - kernels, peer copies - is
 *     complete and visible first); fls_flag_wait holds `stream` on the device until *flag >= target.  No host
 *     polling, no host round trip;
 *   - back-pressure: the same counters in the other direction (the consumer signals how many items it has
 *     released; the producer waits before re-using a ring slot).
 * A wait gives up after `timeout_ms` (a dead peer must not hang the GPU): the stream then continues and
 * fls_flag_timeouts (a device read, call it after synchronising) reports how many waits have given up since
 * the context was created - the host must treat any increase as a failed pass. */
#define FLS_IPC_HANDLE_BYTES 64
int fls_ipc_alloc(fls_ctx* ctx, size_t bytes, void** dev_out, unsigned char* handle_out /* [64] */);
int fls_ipc_free(fls_ctx* ctx, void* dev);
int fls_ipc_open(fls_ctx* ctx, const unsigned char* handle /* [64] */, void** dev_out);
int fls_ipc_close(fls_ctx* ctx, void* dev);
int fls_flag_signal(fls_ctx* ctx, uint32_t* flag_dev, uint32_t value, void* stream);
int fls_flag_wait(fls_ctx* ctx, const uint32_t* flag_dev, uint32_t target, uint32_t timeout_ms, void* stream);
int fls_flag_timeouts(fls_ctx* ctx, uint32_t* count_out);

/* ---------------------------------------------------------------------------------------------
 * single kernels on caller-supplied device buffers (unit parity tests, profiling). */
#define FLS_EPI_STORE 0
#define FLS_EPI_RESIDUAL 1
#define FLS_EPI_SWIGLU 2
/* C[M,N] = A[M,K] * B[N,K]^T (+ epilogue).  SWIGLU: B holds gate rows [0,N) and up rows
 * [up_off, up_off+N); RESIDUAL: C = R + A*B^T (R may alias C). */
int fls_op_gemm(fls_ctx* ctx, int epilogue, const uint16_t* a_dev, int lda, const uint16_t* b_dev, int ldb, int b_rows,
                int m, int n, int k, uint16_t* c_dev, int ldc, const uint16_t* r_dev, int ldr, int up_off,
                void* stream);
int fls_op_rmsnorm(fls_ctx* ctx, const uint16_t* x_dev, const uint16_t* delta_dev, const int32_t* gather_dev,
                   const uint16_t* w_dev, float eps, int n_rows, int hidden, uint16_t* x_sum_dev, uint16_t* out_dev,
                   void* stream);
int fls_op_rope(fls_ctx* ctx, uint16_t* qkv_dev, int ld, int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                const int32_t* row_pos_dev, void* stream);
/* n_rows = rows of the qkv buffer.  head_dim 128 runs the tcgen05 kernel (option "attention_impl" = 2, the default:
 * persistent CTAs; 1: one CTA per query tile and head) or the mma.sync kernel (0); head_dim 64 always the latter.
 * With "attention_impl" = 2 two launches through the same context must not run concurrently. */
int fls_op_attention(fls_ctx* ctx, const uint16_t* qkv_dev, int ld, int n_rows, int n_q_heads, int n_kv_heads,
                     int head_dim, const fls_attn_tile* tiles_dev, int n_tiles, uint16_t* out_dev, int ld_out,
                     void* stream);
int fls_op_swiglu(fls_ctx* ctx, const uint16_t* gate_dev, const uint16_t* up_dev, int ld_in, int n_rows, int inter,
                  uint16_t* out_dev, int ld_out, void* stream);
int fls_op_embed(fls_ctx* ctx, const uint16_t* table_dev, const int32_t* ids_dev, int n_rows, int hidden,
                 uint16_t* out_dev, void* stream);
/* softmax over [n_rows, n] given W[n, k] (lm_head without a registered layer). */
int fls_op_head(fls_ctx* ctx, const uint16_t* x_dev, const uint16_t* w_dev, int n_rows, int n, int k,
                uint16_t* probs_dev, const int32_t* answer_ids_dev, int n_ids, float* logprob_dev, float* lse_dev,
                void* stream);

/* No counterpart in the reference.  The same computation as fls_op_head for callers that cannot
 * allocate device memory: a plugin inside a serving process whose own pool already owns all of HBM,
 * or a latency-critical thread that cannot afford cudaMalloc's device-wide synchronisation.  It
 * takes no context and calls no cudaMalloc/cudaFree: its scratch (the per-tile softmax partials and
 * the per-row log-sum-exp) is a fixed ~1 MiB array inside the library image, held by every GPU
 * that loads the library until the process exits.
 *  - limit: n_rows <= FLS_NOALLOC_MAX_ROWS and n <= FLS_NOALLOC_MAX_VOCAB per call, otherwise
 *    FLS_ERR_CAPACITY (split the rows over several calls, or use fls_op_head);
 *  - runs on the current device; the first call on a device resolves the array's address - make it
 *    before any stream capture begins;
 *  - all calls on a device share that array: calls whose kernels overlap on the GPU race, so keep
 *    them on one stream or order them.  Nothing in the array needs clearing between calls (each
 *    call writes every entry it reads).
 * Inputs/outputs are the caller's buffers, as in fls_op_head: x [n_rows, k] and w [n, k] fp16 row-major and
 * probs [n_rows, n] fp16, each 16-byte aligned, n and k multiples of 8 (FLS_ERR_INVALID otherwise);
 * answer_ids [n_rows, n_ids] int32 on the device - an id outside [0, n) yields NaN in its logprob entry
 * instead of a read outside w.  Errors via fls_last_error(NULL). */
#define FLS_NOALLOC_MAX_ROWS 1024
#define FLS_NOALLOC_MAX_VOCAB 32768
int fls_op_head_noalloc(const uint16_t* x_dev, const uint16_t* w_dev, int n_rows, int n, int k, uint16_t* probs_dev,
                        const int32_t* answer_ids_dev, int n_ids, float* logprob_dev, float* lse_dev, void* stream);

/* Per-kernel device timing (CUDA events recorded around every launch on the caller's stream while
 * profiling is on; a few microseconds per launch).  The reference's only instrumentation is a
 * wall-clock sum of weight-loading time (utils.py:223,230-233,304). */
#define FLS_PROF_GEMM_QKV 0
#define FLS_PROF_GEMM_O 1
#define FLS_PROF_GEMM_GATE_UP 2
#define FLS_PROF_GEMM_DOWN 3
#define FLS_PROF_GEMM_HEAD 4
#define FLS_PROF_ATTENTION 5
#define FLS_PROF_RMSNORM 6
#define FLS_PROF_ROPE 7
#define FLS_PROF_EMBED 8
#define FLS_PROF_OTHER 9
#define FLS_PROF_CLASSES 10
typedef struct fls_kernel_profile {
    double ms;         /* summed device time of the class */
    double work;       /* algorithmic FLOPs (GEMM classes: 2*M*N*K) or bytes (HBM-bound classes) */
    uint64_t launches;
} fls_kernel_profile;
int fls_set_profiling(fls_ctx* ctx, int enabled);
/* Waits for the profiled launches to finish, then fills out[0..FLS_PROF_CLASSES). */
int fls_get_profile(fls_ctx* ctx, fls_kernel_profile* out);

/* Tuning knobs.  "gemm_cta_group": 2 (default; CTA pairs share a 256x256 tile through
 * tcgen05 cta_group::2) or 1 (one CTA per 128x256 tile).  "attention_impl" (head_dim 128): 2 (default; tcgen05 kernel
 * with persistent CTAs that fetch (query tile, head) items from a counter owned by the context - launches of ONE context
 * must therefore not overlap each other), 1 (the same kernel launched as one CTA per item) or 0 (mma.sync kernel).  "attention_k_stages": K buffers per CTA of the tcgen05 attention -
 * 0 (default: 3 where two such CTAs still share an SM, else 2), 2 or 3; the results do not depend on it.  "fuse_rope": 1 (default; rotary embedding applied in the
 * QKV GEMM epilogue when head_dim is 128 and the GEMM runs 256 x 256 tiles), 2 (also with 256 x 512 tiles) or 0
 * (separate kernel).  "paced_weight_stream": see fls_prefetch_pump.
 * "gemm_wide_tiles": with cta_group 2, 256 x 512 tiles per CTA pair (one staged A slice feeds both TMEM
 * accumulators: a quarter less L2->SM traffic, fewer DRAM re-reads at large K; the epilogue no longer overlaps the
 * next tile's MMAs) - 2 (default) where K >= 8192, 1 wherever the epilogue has the variant, 0 never (256 x 256
 * tiles with two accumulator stages); "gemm_wide_lockstep" 1 = both accumulators of a wide tile advance in lock step
 * (comparator; default 0 = two skewed MMA streams so that one accumulator's epilogue overlaps the other's MMAs).
 * "gemm_raster" 0..3: tile order experiments (csrc/gemm_tile.cuh).
 * "gemm_tma_store": 1 (default; GEMM epilogues stage 32 x 64 sub-tiles in shared memory and write them with TMA
 * stores, the residual operand arrives by TMA load) or 0 (every thread stores 16-byte pieces of its own row).
 * "gemm_group_m", "gemm_cache_hints": raster group / L2 hint experiments (tools/gemm_probe.py), default 0;
 * "rmsnorm_variant" 0..2: CTA shape of the RMSNorm kernel (tools/norm_probe.py), default 0. */
int fls_set_option(fls_ctx* ctx, const char* name, int value);

int fls_get_stats(fls_ctx* ctx, fls_stats* out);
int fls_reset_stats(fls_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* FLS_B200_H */
