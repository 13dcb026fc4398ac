// C-ABI entry points (include/fls_b200.h): context, weight slots + copy stream, layer orchestration.
#include <algorithm>
#include <cstdarg>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only; ranges cost a pointer test unless a profiler is attached

#include "../../include/fls_b200.h"
#include "common.cuh"
#include "host_cache.h"

namespace fls {

char* tls_error_buffer() {
    static thread_local char buf[512];
    return buf;
}
int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(tls_error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

namespace {

struct DecoderView {
    const __half *wqkv, *wo, *wgu, *wd, *ln1, *ln2;
};
struct LayerReg {
    int kind = -1;
    std::string path;
    std::shared_ptr<HostLayerImage> img;  // loaded lazily
    std::vector<TensorEntry> header;      // parsed at registration (shapes validated)
    size_t slot_bytes = 0;
};
struct ResidentLayer {
    int layer_idx = -1;
    int kind = -1;
    DecoderView dec{};
    const __half* single = nullptr;  // embed table / norm gain / lm_head weight
};
struct Slot {
    char* dev = nullptr;
    size_t cap = 0;
    cudaEvent_t ready = nullptr, freed = nullptr;
    bool in_use = false;       // copies enqueued, not yet released
    bool freed_valid = false;  // `freed` has been recorded at least once since the last prefetch
    std::vector<ResidentLayer> layers;
    std::vector<std::shared_ptr<HostLayerImage>> pins;  // keep host images alive while copies fly
    // paced weight stream: copies prepared by fls_prefetch_shard but enqueued piecewise by fls_prefetch_pump
    struct CopyOp {
        char* dst;
        const char* src;
        size_t bytes;
        const void* image;   // host image the source lies in: ranges of different page-locked regions never merge
    };
    std::vector<CopyOp> pending_ops;
    size_t next_op = 0;
    int timing_idx = -1;
};

}  // namespace
}  // namespace fls

using namespace fls;

struct fls_ctx {
    int device = 0;
    int num_sms = 148;
    fls_model_desc m{};
    int q_dim = 0, kv_dim = 0, qkv_dim = 0;
    std::vector<LayerReg> layers;
    std::vector<Slot> slots;
    cudaStream_t copy_stream = nullptr;
    // workspace
    int max_rows = 0, max_head_rows = 0;
    __half *xn = nullptr, *qkv = nullptr, *attn = nullptr, *hmid = nullptr;
    float2* lse_partial = nullptr;
    float* lse = nullptr;
    __half *rope_cos = nullptr, *rope_sin = nullptr;
    int rope_n_pos = 0;
    size_t workspace_bytes = 0;
    fls_stats stats{};
    // weight-stream timing: ring of (begin, end) event pairs on the copy stream
    struct CopyTiming {
        cudaEvent_t a = nullptr, b = nullptr;
        size_t bytes = 0;
        bool pending = false;
    };
    std::vector<CopyTiming> copy_ring;
    size_t copy_next = 0;
    // per-kernel event profiling
    bool profiling = false;
    struct ProfSlot {
        cudaEvent_t a = nullptr, b = nullptr;
        int cls = 0;
        bool pending = false;
    };
    std::vector<ProfSlot> prof_ring;
    size_t prof_next = 0;
    fls_kernel_profile prof[FLS_PROF_CLASSES] = {};
    char err[512] = {0};
    bool cuda_poisoned = false;
    unsigned long long peer_enabled_mask = 0;
    int gemm_cg = 2;  // tcgen05 cta_group of the GEMM kernels (2 = CTA pairs, 1 = single CTA)
    int attn_impl = 2;  // head_dim 128: 2 = tcgen05 attention with persistent CTAs (default); 1 = one CTA per (query tile, head); 0 = mma.sync attention
    uint32_t* attn_work = nullptr;  // work counter of the persistent attention (two uint32, zero between launches)
    int attn_k_stages = 0;  // K buffers per CTA of the tcgen05 attention: 0 = chosen per device at init (3 where two CTAs fit an SM), 2, 3
    std::vector<cudaEvent_t> pace_events;
    size_t pace_next = 0;
    int paced_stream = 0;  // 1: fls_prefetch_shard only prepares the copies; fls_prefetch_pump enqueues them
    int gemm_group_m = 0;  // raster group of the GEMM (0 = kernel default)
    int gemm_cache_hints = 0;
    int gemm_raster = 0;
    int rmsnorm_variant = 0;
    int fuse_rope = 1;  // rotary embedding in the QKV GEMM epilogue (head_dim 128): 1 = with 256 x 256 tiles, 2 = always, 0 = never
    int gemm_tma_store = 1;  // 1 = GEMM epilogues write through smem + TMA store, 0 = per-thread stores
    int gemm_wide_lockstep = 0;  // 1: both accumulators of a wide tile advance in lock step (comparator)
    int gemm_wide = 2;       // 256 x 512 tiles per CTA pair: 0 never, 1 wherever the epilogue has the variant, 2 = when K >= 8192
    // cross-process hand-off: buffers this context exported (fls_ipc_alloc) and peers' buffers it mapped (fls_ipc_open)
    std::vector<void*> ipc_owned, ipc_mapped;
    uint32_t* flag_timeouts = nullptr;  // device counter bumped by a flag wait that gave up
};

namespace {

int ctx_fail(fls_ctx* c, int code) {
    if (c) {
        strncpy(c->err, tls_error_buffer(), sizeof(c->err) - 1);
        if (code == kErrCuda) c->cuda_poisoned = true;
    }
    return code;
}
#define CTX_TRY(c, expr)                      \
    do {                                      \
        int _s = (expr);                      \
        if (_s != 0) return ctx_fail((c), _s); \
    } while (0)
#define CTX_CUDA(c, expr)                                                                                  \
    do {                                                                                                   \
        cudaError_t _e = (expr);                                                                           \
        if (_e != cudaSuccess)                                                                             \
            return ctx_fail((c), fail(kErrCuda, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                                      __LINE__));                                                          \
    } while (0)

int enter(fls_ctx* c) {
    if (!c) return fail(kErrInvalid, "null context");
    if (c->cuda_poisoned) return fail(kErrCuda, "context is unusable after an earlier CUDA error: %s", c->err);
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(kErrCuda, "cudaSetDevice(%d): %s", c->device, cudaGetErrorString(e));
    return kOk;
}
#define ENTER(c)                                 \
    do {                                         \
        int _s = enter(c);                       \
        if (_s != 0) return ctx_fail((c), _s);   \
    } while (0)

inline size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

int expect_shape(const LayerReg& L, const char* suffix, long long d0, long long d1, size_t* bytes) {
    for (auto& t : L.header) {
        const size_t n = strlen(suffix);
        if (t.name.size() >= n && t.name.compare(t.name.size() - n, n, suffix) == 0) {
            const bool ok = d1 < 0 ? (t.shape.size() == 1 && t.shape[0] == d0)
                                   : (t.shape.size() == 2 && t.shape[0] == d0 && t.shape[1] == d1);
            if (!ok) return fail(kErrInvalid, "%s: tensor %s has an unexpected shape", L.path.c_str(), t.name.c_str());
            *bytes += align256(t.numel * 2);
            return kOk;
        }
    }
    return fail(kErrIo, "%s: tensor *%s is missing", L.path.c_str(), suffix);
}

void free_workspace(fls_ctx* c) {
    for (auto& s : c->slots) {
        if (s.dev) cudaFree(s.dev);
        if (s.ready) cudaEventDestroy(s.ready);
        if (s.freed) cudaEventDestroy(s.freed);
    }
    c->slots.clear();
    for (void* p : {static_cast<void*>(c->xn), static_cast<void*>(c->qkv), static_cast<void*>(c->attn),
                    static_cast<void*>(c->hmid), static_cast<void*>(c->lse_partial), static_cast<void*>(c->lse)})
        if (p) cudaFree(p);
    c->xn = c->qkv = c->attn = c->hmid = nullptr;
    c->lse_partial = nullptr;
    c->lse = nullptr;
    c->workspace_bytes = 0;
}

void harvest_copy_timing(fls_ctx* c, fls_ctx::CopyTiming& t, bool wait) {
    if (!t.pending) return;
    if (wait) cudaEventSynchronize(t.b);
    if (cudaEventQuery(t.b) == cudaSuccess) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) {
            c->stats.h2d_ms += ms;
            c->stats.h2d_bytes += t.bytes;
        }
        t.pending = false;
    } else {
        cudaGetLastError();
    }
}

const ResidentLayer* find_resident(fls_ctx* c, int slot, int layer_idx, int kind) {
    if (slot < 0 || slot >= static_cast<int>(c->slots.size())) {
        fail(kErrInvalid, "slot %d out of range", slot);
        return nullptr;
    }
    for (auto& r : c->slots[slot].layers)
        if (r.layer_idx == layer_idx) {
            if (r.kind != kind) {
                fail(kErrState, "layer %d is of kind %d, not %d", layer_idx, r.kind, kind);
                return nullptr;
            }
            return &r;
        }
    fail(kErrState, "layer %d is not resident in slot %d (prefetch it first)", layer_idx, slot);
    return nullptr;
}

void prof_harvest(fls_ctx* c, fls_ctx::ProfSlot& p, bool wait) {
    if (!p.pending) return;
    if (wait) cudaEventSynchronize(p.b);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
        c->prof[p.cls].ms += ms;
        p.pending = false;
    } else {
        cudaGetLastError();
    }
}

// Brackets one launch with events on its stream when profiling is on.
struct ProfScope {
    fls_ctx* c;
    fls_ctx::ProfSlot* slot = nullptr;
    cudaStream_t st;
    ProfScope(fls_ctx* ctx, int cls, double work, cudaStream_t stream) : c(ctx), st(stream) {
        static const char* const kNames[FLS_PROF_CLASSES] = {"fls:gemm_qkv", "fls:gemm_o", "fls:gemm_gate_up",
                                                             "fls:gemm_down", "fls:gemm_head", "fls:attention",
                                                             "fls:rmsnorm", "fls:rope", "fls:embed", "fls:other"};
        nvtxRangePushA(kNames[cls]);     // the reference's only trace is one wall-clock print (utils.py:304)
        if (!c->profiling) return;
        if (c->prof_ring.empty()) {
            c->prof_ring.resize(4096);
            for (auto& p : c->prof_ring) {
                cudaEventCreate(&p.a);
                cudaEventCreate(&p.b);
            }
        }
        slot = &c->prof_ring[c->prof_next++ % c->prof_ring.size()];
        prof_harvest(c, *slot, true);
        slot->cls = cls;
        c->prof[cls].work += work;
        c->prof[cls].launches += 1;
        cudaEventRecord(slot->a, st);
    }
    ~ProfScope() {
        if (slot) {
            cudaEventRecord(slot->b, st);
            slot->pending = true;
        }
        nvtxRangePop();
    }
};

struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
inline const __half* H(const uint16_t* p) { return reinterpret_cast<const __half*>(p); }
inline __half* H(uint16_t* p) { return reinterpret_cast<__half*>(p); }

int attention_dispatch(fls_ctx* c, const __half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                       const AttnTile* tiles, int n_tiles, __half* out, int ld_out, cudaStream_t st) {
    if (c->attn_impl == 2 && head_dim == 128)
        return attention_sm100_launch_persistent(qkv, ld, n_rows, n_q_heads, n_kv_heads, tiles, n_tiles, out, ld_out, st,
                                                 c->attn_work, c->num_sms);
    if (c->attn_impl == 1 && head_dim == 128)
        return attention_sm100_launch(qkv, ld, n_rows, n_q_heads, n_kv_heads, tiles, n_tiles, out, ld_out, st,
                                      c->attn_k_stages);
    return attention_launch(qkv, ld, n_q_heads, n_kv_heads, head_dim, tiles, n_tiles, out, ld_out, st);
}

int head_impl(fls_ctx* c, const __half* x, const __half* W, int n_rows, int V, int K, __half* probs,
              const int32_t* answer_ids, int n_ids, float* logprob, float* lse_out, cudaStream_t st) {
    if (n_rows > c->max_head_rows)
        return fail(kErrCapacity, "lm_head: %d rows exceed the configured max_head_rows %d", n_rows, c->max_head_rows);
    if (gemm_lse_blocks(V) > gemm_lse_blocks(c->m.vocab_size))
        return fail(kErrCapacity, "lm_head: vocab %d exceeds the configured %d", V, c->m.vocab_size);
    GemmArgs a{};
    a.A = x, a.lda = K, a.B = W, a.ldb = K, a.b_rows = V, a.M = n_rows, a.N = V, a.K = K;
    a.partial = c->lse_partial;
    const double flops = 2.0 * n_rows * V * K;
    {
        ProfScope ps(c, FLS_PROF_GEMM_HEAD, flops, st);
        a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
        FLS_TRY(gemm_launch(kEpiLsePartial, a, c->num_sms, st, c->gemm_cg));
    }
    {
        ProfScope ps(c, FLS_PROF_OTHER, 0, st);
        FLS_TRY(lse_reduce_launch(c->lse_partial, n_rows, gemm_lse_blocks(V), c->lse, st));
    }
    c->stats.kernel_launches += 2;
    if (probs) {
        a.C = probs, a.ldc = V, a.lse = c->lse;
        ProfScope ps(c, FLS_PROF_GEMM_HEAD, flops, st);
        a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
        FLS_TRY(gemm_launch(kEpiProbs, a, c->num_sms, st, c->gemm_cg));
        c->stats.kernel_launches += 1;
    }
    if (logprob && n_ids > 0) {
        ProfScope ps(c, FLS_PROF_OTHER, 0, st);
        FLS_TRY(gather_logprob_launch(x, K, W, K, K, answer_ids, n_ids, c->lse, n_rows, V, logprob, st));
        c->stats.kernel_launches += 1;
    }
    if (lse_out) FLS_CUDA(cudaMemcpyAsync(lse_out, c->lse, sizeof(float) * n_rows, cudaMemcpyDeviceToDevice, st));
    return kOk;
}

}  // namespace

namespace {
// Enqueue up to `max_bytes` of a slot's prepared weight copies on the copy stream; once the last one is
// in, record the timing and `ready` events.
int pump_slot(fls_ctx* c, Slot& s, size_t max_bytes) {
    if (s.next_op >= s.pending_ops.size() && s.timing_idx < 0) return kOk;
    fls_ctx::CopyTiming& timing = c->copy_ring[s.timing_idx];
    if (s.next_op == 0) FLS_CUDA(cudaEventRecord(timing.a, c->copy_stream));
    size_t done = 0;
    while (s.next_op < s.pending_ops.size() && done < max_bytes) {
        const Slot::CopyOp& op = s.pending_ops[s.next_op++];
        FLS_CUDA(cudaMemcpyAsync(op.dst, op.src, op.bytes, cudaMemcpyHostToDevice, c->copy_stream));
        done += op.bytes;
    }
    if (s.next_op >= s.pending_ops.size()) {
        FLS_CUDA(cudaEventRecord(timing.b, c->copy_stream));
        timing.pending = true;
        FLS_CUDA(cudaEventRecord(s.ready, c->copy_stream));
        s.pending_ops.clear();
        s.next_op = 0;
        s.timing_idx = -1;
    }
    return kOk;
}
}  // namespace

// =============================================================================================
extern "C" {

int fls_version(void) { return 1; }

const char* fls_last_error(const fls_ctx* ctx) { return ctx ? ctx->err : tls_error_buffer(); }

int fls_create(const fls_model_desc* d, int device, fls_ctx** out) {
    if (!d || !out) return fail(kErrInvalid, "fls_create: null argument");
    if (d->hidden_size <= 0 || d->hidden_size % 8 || d->intermediate_size % 8 || d->vocab_size % 8 ||
        d->num_attention_heads <= 0 || d->num_key_value_heads <= 0 ||
        d->num_attention_heads % d->num_key_value_heads || (d->head_dim != 64 && d->head_dim != 128))
        return fail(kErrInvalid, "fls_create: unsupported model shape (hidden %d, inter %d, vocab %d, heads %d/%d, "
                    "head_dim %d)", d->hidden_size, d->intermediate_size, d->vocab_size, d->num_attention_heads,
                    d->num_key_value_heads, d->head_dim);
    FLS_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    FLS_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(kErrCuda, "device %d is sm_%d%d; this library contains sm_100a code only", device, prop.major,
                    prop.minor);
    FLS_TRY(gemm_init());
    FLS_TRY(attention_init());
    FLS_TRY(attention_sm100_init());
    auto c = new fls_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    c->m = *d;
    c->q_dim = d->num_attention_heads * d->head_dim;
    c->kv_dim = d->num_key_value_heads * d->head_dim;
    c->qkv_dim = c->q_dim + 2 * c->kv_dim;
    c->layers.resize(d->num_hidden_layers + 3);
    // the weight stream yields to everything else that uses the copy engines (activation tiers, score D2H)
    int prio_least = 0, prio_greatest = 0;
    cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    cudaError_t e = cudaStreamCreateWithPriority(&c->copy_stream, cudaStreamNonBlocking, prio_least);
    if (e != cudaSuccess) {
        delete c;
        return fail(kErrCuda, "cudaStreamCreate: %s", cudaGetErrorString(e));
    }
    e = cudaMalloc(reinterpret_cast<void**>(&c->attn_work), 2 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(c->attn_work, 0, 2 * sizeof(uint32_t));
    if (e != cudaSuccess) {
        cudaStreamDestroy(c->copy_stream);
        delete c;
        return fail(kErrCuda, "attention work counter: %s", cudaGetErrorString(e));
    }
    *out = c;
    return kOk;
}

int fls_destroy(fls_ctx* c) {
    if (!c) return kOk;
    // may run from a garbage collector on any thread: leave that thread's current device as it was
    int prev = -1;
    cudaGetDevice(&prev);
    struct Restore {
        int d;
        ~Restore() {
            if (d >= 0) cudaSetDevice(d);
        }
    } restore{prev};
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_workspace(c);
    if (c->rope_cos) cudaFree(c->rope_cos);
    if (c->rope_sin) cudaFree(c->rope_sin);
    if (c->attn_work) cudaFree(c->attn_work);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (void* p : c->ipc_mapped) cudaIpcCloseMemHandle(p);
    for (void* p : c->ipc_owned) cudaFree(p);
    if (c->flag_timeouts) cudaFree(c->flag_timeouts);
    for (auto& p : c->prof_ring) {
        cudaEventDestroy(p.a);
        cudaEventDestroy(p.b);
    }
    for (auto& t : c->copy_ring) {
        cudaEventDestroy(t.a);
        cudaEventDestroy(t.b);
    }
    for (auto& e : c->pace_events) cudaEventDestroy(e);
    delete c;
    return kOk;
}

int fls_configure(fls_ctx* c, int n_slots, size_t slot_bytes, int max_rows, int max_head_rows) {
    ENTER(c);
    if (n_slots < 1 || n_slots > 512 || max_rows < 1 || max_head_rows < 1)
        return ctx_fail(c, fail(kErrInvalid, "fls_configure: n_slots %d, max_rows %d, max_head_rows %d", n_slots,
                                max_rows, max_head_rows));
    CTX_CUDA(c, cudaDeviceSynchronize());
    free_workspace(c);
    c->slots.resize(n_slots);
    size_t total = 0;
    for (auto& s : c->slots) {
        s.cap = align256(slot_bytes);
        CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&s.dev), std::max<size_t>(s.cap, 256)));
        CTX_CUDA(c, cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming));
        CTX_CUDA(c, cudaEventCreateWithFlags(&s.freed, cudaEventDisableTiming));
        total += s.cap;
    }
    const size_t R = static_cast<size_t>(max_rows);
    const size_t b_xn = R * c->m.hidden_size * 2, b_qkv = R * c->qkv_dim * 2, b_attn = R * c->q_dim * 2,
                 b_h = R * c->m.intermediate_size * 2;
    const size_t b_part = static_cast<size_t>(max_head_rows) * gemm_lse_blocks(c->m.vocab_size) * sizeof(float2);
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->xn), b_xn));
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->qkv), b_qkv));
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->attn), b_attn));
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->hmid), b_h));
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->lse_partial), b_part));
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->lse), sizeof(float) * max_head_rows));
    c->max_rows = max_rows;
    c->max_head_rows = max_head_rows;
    c->workspace_bytes = total + b_xn + b_qkv + b_attn + b_h + b_part + sizeof(float) * max_head_rows;
    return kOk;
}

int fls_set_rope_table(fls_ctx* c, const uint16_t* cos_host, const uint16_t* sin_host, int n_pos) {
    ENTER(c);
    if (!cos_host || !sin_host || n_pos < 1) return ctx_fail(c, fail(kErrInvalid, "fls_set_rope_table: bad argument"));
    const size_t bytes = static_cast<size_t>(n_pos) * c->m.head_dim * 2;
    CTX_CUDA(c, cudaDeviceSynchronize());
    if (c->rope_cos) cudaFree(c->rope_cos);
    if (c->rope_sin) cudaFree(c->rope_sin);
    c->rope_cos = c->rope_sin = nullptr;
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->rope_cos), bytes));
    CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->rope_sin), bytes));
    CTX_CUDA(c, cudaMemcpy(c->rope_cos, cos_host, bytes, cudaMemcpyHostToDevice));
    CTX_CUDA(c, cudaMemcpy(c->rope_sin, sin_host, bytes, cudaMemcpyHostToDevice));
    c->rope_n_pos = n_pos;
    return kOk;
}

// --------------------------------------------------------------------------------- weight stream
int fls_register_layer_file(fls_ctx* c, int layer_idx, int kind, const char* path, int load_now) {
    ENTER(c);
    if (layer_idx < 0 || layer_idx >= static_cast<int>(c->layers.size()) || !path || kind < 0 || kind > 3)
        return ctx_fail(c, fail(kErrInvalid, "fls_register_layer_file: layer %d kind %d", layer_idx, kind));
    LayerReg L;
    L.kind = kind;
    L.path = path;
    size_t ds = 0, fs = 0;
    CTX_TRY(c, parse_layer_header(L.path, &L.header, &ds, &fs));
    for (auto& t : L.header)
        if (t.dtype != "F16" && t.dtype != "BF16" && t.dtype != "F32")
            return ctx_fail(c, fail(kErrInvalid, "%s: tensor %s has dtype %s; int8/quantised weights are not supported",
                                    path, t.name.c_str(), t.dtype.c_str()));
    const auto& m = c->m;
    size_t bytes = 0;
    if (kind == FLS_LAYER_EMBED) {
        CTX_TRY(c, expect_shape(L, "embed_tokens.weight", m.vocab_size, m.hidden_size, &bytes));
    } else if (kind == FLS_LAYER_NORM) {
        CTX_TRY(c, expect_shape(L, "norm.weight", m.hidden_size, -1, &bytes));
    } else if (kind == FLS_LAYER_HEAD) {
        CTX_TRY(c, expect_shape(L, "lm_head.weight", m.vocab_size, m.hidden_size, &bytes));
    } else {
        CTX_TRY(c, expect_shape(L, "self_attn.q_proj.weight", c->q_dim, m.hidden_size, &bytes));
        CTX_TRY(c, expect_shape(L, "self_attn.k_proj.weight", c->kv_dim, m.hidden_size, &bytes));
        CTX_TRY(c, expect_shape(L, "self_attn.v_proj.weight", c->kv_dim, m.hidden_size, &bytes));
        CTX_TRY(c, expect_shape(L, "self_attn.o_proj.weight", m.hidden_size, c->q_dim, &bytes));
        CTX_TRY(c, expect_shape(L, "mlp.gate_proj.weight", m.intermediate_size, m.hidden_size, &bytes));
        CTX_TRY(c, expect_shape(L, "mlp.up_proj.weight", m.intermediate_size, m.hidden_size, &bytes));
        CTX_TRY(c, expect_shape(L, "mlp.down_proj.weight", m.hidden_size, m.intermediate_size, &bytes));
        CTX_TRY(c, expect_shape(L, ".input_layernorm.weight", m.hidden_size, -1, &bytes));
        CTX_TRY(c, expect_shape(L, "post_attention_layernorm.weight", m.hidden_size, -1, &bytes));
    }
    // Every tensor of the file must be one this architecture uses: a bias or an adapter tensor that the kernels
    // would silently ignore (the reference, which builds the model from the config, would not) is refused.
    {
        static const char* const kDecoder[] = {"self_attn.q_proj.weight", "self_attn.k_proj.weight",
                                               "self_attn.v_proj.weight", "self_attn.o_proj.weight",
                                               "mlp.gate_proj.weight", "mlp.up_proj.weight", "mlp.down_proj.weight",
                                               ".input_layernorm.weight", "post_attention_layernorm.weight",
                                               "rotary_emb.inv_freq"};   // old checkpoints carry this buffer; unused
        static const char* const kEmbed[] = {"embed_tokens.weight"};
        static const char* const kNorm[] = {"norm.weight"};
        static const char* const kHead[] = {"lm_head.weight"};
        const char* const* names = kind == FLS_LAYER_EMBED ? kEmbed : kind == FLS_LAYER_NORM ? kNorm
                                   : kind == FLS_LAYER_HEAD ? kHead : kDecoder;
        const int n_names = kind == FLS_LAYER_DECODER ? 10 : 1;
        for (auto& t : L.header) {
            bool known = false;
            for (int i = 0; i < n_names && !known; ++i) {
                const size_t n = strlen(names[i]);
                known = t.name.size() >= n && t.name.compare(t.name.size() - n, n, names[i]) == 0;
            }
            if (!known)
                return ctx_fail(c, fail(kErrInvalid, "%s: unexpected tensor %s (biases and non-Llama-2 tensors are not "
                                        "supported)", path, t.name.c_str()));
        }
    }
    L.slot_bytes = bytes;
    if (load_now) CTX_TRY(c, acquire_host_image(L.path, &L.img));
    c->layers[layer_idx] = std::move(L);
    return kOk;
}

int fls_layer_slot_bytes(fls_ctx* c, int layer_idx, size_t* out) {
    if (!c || !out || layer_idx < 0 || layer_idx >= static_cast<int>(c->layers.size()) || c->layers[layer_idx].kind < 0)
        return ctx_fail(c, fail(kErrInvalid, "fls_layer_slot_bytes: layer %d is not registered", layer_idx));
    *out = c->layers[layer_idx].slot_bytes;
    return kOk;
}

int fls_prefetch_shard(fls_ctx* c, int slot, const int32_t* layer_ids, int n) {
    ENTER(c);
    if (slot < 0 || slot >= static_cast<int>(c->slots.size()) || n < 0 || (n > 0 && !layer_ids))
        return ctx_fail(c, fail(kErrInvalid, "fls_prefetch_shard: slot %d, n %d", slot, n));
    Slot& s = c->slots[slot];
    if (s.in_use) return ctx_fail(c, fail(kErrState, "slot %d is still in use: release it before prefetching", slot));
    NvtxRange prefetch_range("fls:prefetch_shard");
    size_t need = 0;
    for (int i = 0; i < n; ++i) {
        const int li = layer_ids[i];
        if (li < 0 || li >= static_cast<int>(c->layers.size()) || c->layers[li].kind < 0)
            return ctx_fail(c, fail(kErrInvalid, "fls_prefetch_shard: layer %d is not registered", li));
        need += c->layers[li].slot_bytes;
    }
    if (need > s.cap)
        return ctx_fail(c, fail(kErrCapacity, "shard needs %zu bytes but slots hold %zu", need, s.cap));
    // host images first (may read files), so the copy stream is not stalled half-way through a shard
    for (int i = 0; i < n; ++i) {
        LayerReg& L = c->layers[layer_ids[i]];
        if (!L.img) CTX_TRY(c, acquire_host_image(L.path, &L.img));
    }
    if (s.freed_valid) CTX_CUDA(c, cudaStreamWaitEvent(c->copy_stream, s.freed, 0));
    s.layers.clear();
    s.pins.clear();
    if (c->copy_ring.empty()) {
        c->copy_ring.resize(64);
        for (auto& t : c->copy_ring) {
            CTX_CUDA(c, cudaEventCreate(&t.a));
            CTX_CUDA(c, cudaEventCreate(&t.b));
        }
    }
    for (auto& t : c->copy_ring) harvest_copy_timing(c, t, false);
    s.timing_idx = static_cast<int>(c->copy_next++ % c->copy_ring.size());
    fls_ctx::CopyTiming& timing = c->copy_ring[s.timing_idx];
    harvest_copy_timing(c, timing, true);  // only blocks if 64 shard copies are still in flight
    s.pending_ops.clear();
    s.next_op = 0;
    size_t off = 0;
    auto put = [&](const HostLayerImage& img, const char* suffix, const __half** dst) -> int {
        const TensorEntry* t = img.find(suffix);
        if (!t) return fail(kErrIo, "%s: tensor *%s is missing", img.path.c_str(), suffix);
        // 32 MiB pieces: other streams' copies (activation tiers) can slip in between instead of queueing
        // behind a whole shard
        const size_t total = t->numel * 2, piece = size_t(32) << 20;
        for (size_t o = 0; o < total; o += piece)
            s.pending_ops.push_back({s.dev + off + o, img.pinned + t->host_off + o, std::min(piece, total - o), &img});
        *dst = reinterpret_cast<const __half*>(s.dev + off);
        off += align256(t->numel * 2);
        return kOk;
    };
    for (int i = 0; i < n; ++i) {
        LayerReg& L = c->layers[layer_ids[i]];
        ResidentLayer r;
        r.layer_idx = layer_ids[i];
        r.kind = L.kind;
        const __half* tmp = nullptr;
        if (L.kind == FLS_LAYER_DECODER) {
            // q, k, v rows adjacent -> one [q_dim + 2 kv_dim, H] operand; gate, up adjacent -> [2I, H]
            CTX_TRY(c, put(*L.img, "self_attn.q_proj.weight", &r.dec.wqkv));
            CTX_TRY(c, put(*L.img, "self_attn.k_proj.weight", &tmp));
            CTX_TRY(c, put(*L.img, "self_attn.v_proj.weight", &tmp));
            CTX_TRY(c, put(*L.img, "self_attn.o_proj.weight", &r.dec.wo));
            CTX_TRY(c, put(*L.img, "mlp.gate_proj.weight", &r.dec.wgu));
            CTX_TRY(c, put(*L.img, "mlp.up_proj.weight", &tmp));
            CTX_TRY(c, put(*L.img, "mlp.down_proj.weight", &r.dec.wd));
            CTX_TRY(c, put(*L.img, ".input_layernorm.weight", &r.dec.ln1));
            CTX_TRY(c, put(*L.img, "post_attention_layernorm.weight", &r.dec.ln2));
        } else {
            const char* suffix = L.kind == FLS_LAYER_EMBED ? "embed_tokens.weight"
                                 : L.kind == FLS_LAYER_NORM ? "norm.weight" : "lm_head.weight";
            CTX_TRY(c, put(*L.img, suffix, &r.single));
        }
        s.layers.push_back(r);
        s.pins.push_back(L.img);
        L.img.reset();  // the slot keeps the image alive while its copies fly; afterwards the cache may evict it
    }
    // Runs of copies that are contiguous on both sides (a packed layer file lies in slot order) become one range,
    // re-cut into 32 MiB pieces.
    {
        std::vector<Slot::CopyOp> merged;
        for (const Slot::CopyOp& op : s.pending_ops) {
            if (!merged.empty() && merged.back().image == op.image && merged.back().dst + merged.back().bytes == op.dst &&
                merged.back().src + merged.back().bytes == op.src)
                merged.back().bytes += op.bytes;
            else
                merged.push_back(op);
        }
        s.pending_ops.clear();
        const size_t piece = size_t(32) << 20;
        for (const Slot::CopyOp& op : merged)
            for (size_t o = 0; o < op.bytes; o += piece)
                s.pending_ops.push_back({op.dst + o, op.src + o, std::min(piece, op.bytes - o), op.image});
        c->stats.h2d_copies += s.pending_ops.size();
    }
    timing.bytes = off;
    s.in_use = true;
    s.freed_valid = false;
    if (!c->paced_stream) CTX_TRY(c, pump_slot(c, s, ~size_t(0)));
    return kOk;
}

int fls_prefetch_pump(fls_ctx* c, int slot, size_t max_bytes, void* after_stream) {
    ENTER(c);
    if (slot < 0 || slot >= static_cast<int>(c->slots.size()))
        return ctx_fail(c, fail(kErrInvalid, "fls_prefetch_pump: slot %d", slot));
    if (after_stream != nullptr) {
        // hold this slice back until the work enqueued on `after_stream` so far has run: the pacing then follows
        // the GPU's progress, not the (far ahead) host's
        if (c->pace_events.empty()) {
            c->pace_events.resize(64);
            for (auto& e : c->pace_events) CTX_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        }
        cudaEvent_t e = c->pace_events[c->pace_next++ % c->pace_events.size()];
        CTX_CUDA(c, cudaEventRecord(e, as_stream(after_stream)));
        CTX_CUDA(c, cudaStreamWaitEvent(c->copy_stream, e, 0));
    }
    CTX_TRY(c, pump_slot(c, c->slots[slot], max_bytes));
    return kOk;
}

int fls_acquire_shard(fls_ctx* c, int slot, void* stream) {
    ENTER(c);
    if (slot < 0 || slot >= static_cast<int>(c->slots.size()) ||
        (!c->slots[slot].in_use && c->slots[slot].layers.empty()))
        return ctx_fail(c, fail(kErrState, "fls_acquire_shard: slot %d has no prefetched shard", slot));
    // a released slot that still holds its layers may be acquired again without a new copy (resident weights)
    c->slots[slot].in_use = true;
    CTX_TRY(c, pump_slot(c, c->slots[slot], ~size_t(0)));  // paced stream: whatever is still pending goes in now
    CTX_CUDA(c, cudaStreamWaitEvent(as_stream(stream), c->slots[slot].ready, 0));
    return kOk;
}

int fls_release_shard(fls_ctx* c, int slot, void* stream) {
    ENTER(c);
    if (slot < 0 || slot >= static_cast<int>(c->slots.size()) || !c->slots[slot].in_use)
        return ctx_fail(c, fail(kErrState, "fls_release_shard: slot %d is not in use", slot));
    Slot& s = c->slots[slot];
    CTX_CUDA(c, cudaEventRecord(s.freed, as_stream(stream)));
    s.freed_valid = true;
    s.in_use = false;
    return kOk;
}

int fls_preload_layer_file(const char* path) {
    if (!path) return fail(kErrInvalid, "fls_preload_layer_file: null path");
    std::shared_ptr<HostLayerImage> img;
    return acquire_host_image(path, &img);
}

int fls_set_host_cache_limit(size_t max_bytes) {
    set_host_cache_limit(max_bytes);
    return kOk;
}

int fls_set_host_image_mode(int mode) {
    if (mode < 0 || mode > 2) return fail(kErrInvalid, "fls_set_host_image_mode: mode %d (0 auto, 1 copy, 2 map)", mode);
    set_host_image_mode(mode);
    return kOk;
}

int fls_set_stage_dir(const char* dir) {
    set_stage_dir(dir);
    return kOk;
}

int fls_trim_host_cache(size_t keep_bytes) {
    trim_host_cache(keep_bytes);
    return kOk;
}

// --------------------------------------------------------------------------------- layer compute
int fls_embed(fls_ctx* c, int slot, int layer_idx, const int32_t* ids, int n_rows, uint16_t* out, void* stream) {
    ENTER(c);
    const ResidentLayer* r = find_resident(c, slot, layer_idx, FLS_LAYER_EMBED);
    if (!r) return ctx_fail(c, kErrState);
    {
        ProfScope ps(c, FLS_PROF_EMBED, static_cast<double>(n_rows) * c->m.hidden_size * 4.0, as_stream(stream));
        CTX_TRY(c, embed_launch(r->single, ids, n_rows, c->m.hidden_size, H(out), as_stream(stream)));
    }
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_decoder_layer(fls_ctx* c, int slot, int layer_idx, uint16_t* x_dev, int n_rows, const int32_t* row_pos,
                      const fls_attn_tile* tiles, int n_tiles, void* stream) {
    return fls_decoder_layer_kv(c, slot, layer_idx, x_dev, n_rows, row_pos, tiles, n_tiles, FLS_KV_NONE, nullptr, 0,
                                nullptr, stream);
}

int fls_decoder_layer_kv(fls_ctx* c, int slot, int layer_idx, uint16_t* x_dev, int n_rows, const int32_t* row_pos,
                         const fls_attn_tile* tiles, int n_tiles, int kv_mode, const int32_t* kv_rows, int n_kv_rows,
                         uint16_t* kv_cache, void* stream) {
    ENTER(c);
    const ResidentLayer* r = find_resident(c, slot, layer_idx, FLS_LAYER_DECODER);
    if (!r) return ctx_fail(c, kErrState);
    if (kv_mode != FLS_KV_NONE && kv_mode != FLS_KV_SAVE && kv_mode != FLS_KV_LOAD)
        return ctx_fail(c, fail(kErrInvalid, "decoder_layer: kv_mode %d", kv_mode));
    if (kv_mode == FLS_KV_NONE) n_kv_rows = 0;
    if (n_kv_rows < 0 || (n_kv_rows > 0 && (!kv_cache || (kv_mode == FLS_KV_SAVE && !kv_rows))))
        return ctx_fail(c, fail(kErrInvalid, "decoder_layer: %d cached rows need a cache%s", n_kv_rows,
                                kv_mode == FLS_KV_SAVE ? " and their row indices" : ""));
    // rows of the cache ahead of x's rows in the attention workspace (FLS_KV_LOAD only)
    const int row0 = kv_mode == FLS_KV_LOAD ? n_kv_rows : 0;
    if (row0 + n_rows > c->max_rows)
        return ctx_fail(c, fail(kErrCapacity, "decoder_layer: %d rows (%d of them cached) exceed the configured "
                                "max_rows %d", row0 + n_rows, row0, c->max_rows));
    if (!c->rope_cos) return ctx_fail(c, fail(kErrState, "decoder_layer: call fls_set_rope_table first"));
    if (n_rows <= 0) return kOk;
    NvtxRange layer_range(kv_mode == FLS_KV_LOAD ? "fls:decoder_layer(kv load)" : "fls:decoder_layer");
    const auto& m = c->m;
    const int Hd = m.hidden_size, I = m.intermediate_size;
    cudaStream_t st = as_stream(stream);
    __half* x = H(x_dev);
    static_assert(sizeof(fls_attn_tile) == sizeof(AttnTile), "tile layouts must match");

    const double R = n_rows;
    const double norm_bytes = R * Hd * 4 + Hd * 2.0;
    // hidden = x + o_proj(attn(rope(qkv(norm(x)))))
    {
        ProfScope ps(c, FLS_PROF_RMSNORM, norm_bytes, st);
        CTX_TRY(c, rmsnorm_launch(x, nullptr, nullptr, r->dec.ln1, m.rms_norm_eps, n_rows, Hd, nullptr, c->xn, st,
                                  c->rmsnorm_variant));
    }
    GemmArgs a{};
    a.A = c->xn, a.lda = Hd, a.B = r->dec.wqkv, a.ldb = Hd, a.b_rows = c->qkv_dim;
    __half* qkv_new = c->qkv + static_cast<size_t>(row0) * c->qkv_dim;     // where this call's rows go
    a.M = n_rows, a.N = c->qkv_dim, a.K = Hd, a.C = qkv_new, a.ldc = c->qkv_dim;
    // With 256 x 512 tiles the epilogue is not hidden behind the next tile's MMAs, and the rotary epilogue (a table
    // row per thread) is the longest one: there the plain store + the stand-alone RoPE kernel is faster (70B layer:
    // 184.8 vs 188.0 ms per 65536 rows, tools/layer_probe.py); with 256 x 256 tiles the fused form wins.
    const bool wide_qkv = c->gemm_cg == 2 && c->gemm_tma_store && (c->gemm_wide == 1 || (c->gemm_wide == 2 && Hd >= 8192));
    const bool fused_rope = m.head_dim == 128 && (c->fuse_rope == 2 || (c->fuse_rope == 1 && !wide_qkv));
    {
        ProfScope ps(c, FLS_PROF_GEMM_QKV, 2.0 * R * c->qkv_dim * Hd, st);
        if (fused_rope) {
            a.row_pos = row_pos, a.cos_tab = c->rope_cos, a.sin_tab = c->rope_sin;
            a.rope_cols = c->q_dim + c->kv_dim;
            a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
            CTX_TRY(c, gemm_launch(kEpiStoreRope, a, c->num_sms, st, c->gemm_cg));
        } else {
            a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
            CTX_TRY(c, gemm_launch(kEpiStore, a, c->num_sms, st, c->gemm_cg));
        }
    }
    if (!fused_rope) {
        ProfScope ps(c, FLS_PROF_ROPE, R * (c->q_dim + c->kv_dim) * 4.0, st);
        CTX_TRY(c, rope_launch(qkv_new, c->qkv_dim, n_rows, m.num_attention_heads, m.num_key_value_heads, m.head_dim,
                               row_pos, c->rope_cos, c->rope_sin, c->rope_n_pos, st));
    }
    if (n_kv_rows > 0) {
        ProfScope ps(c, FLS_PROF_OTHER, static_cast<double>(n_kv_rows) * c->kv_dim * 8.0, st);
        CTX_TRY(c, kv_rows_copy_launch(c->qkv, c->qkv_dim, c->q_dim, 2 * c->kv_dim,
                                       kv_mode == FLS_KV_SAVE ? kv_rows : nullptr, n_kv_rows, H(kv_cache),
                                       kv_mode == FLS_KV_SAVE, st));
        c->stats.kernel_launches += 1;
    }
    {
        ProfScope ps(c, FLS_PROF_ATTENTION, 0, st);
        CTX_TRY(c, attention_dispatch(c, c->qkv, c->qkv_dim, row0 + n_rows, m.num_attention_heads,
                                      m.num_key_value_heads, m.head_dim, reinterpret_cast<const AttnTile*>(tiles),
                                      n_tiles, c->attn, c->q_dim, st));
    }
    a = GemmArgs{};
    a.A = c->attn + static_cast<size_t>(row0) * c->q_dim, a.lda = c->q_dim, a.B = r->dec.wo, a.ldb = c->q_dim;
    a.b_rows = Hd;
    a.M = n_rows, a.N = Hd, a.K = c->q_dim, a.C = x, a.ldc = Hd, a.R = x, a.ldr = Hd;
    {
        ProfScope ps(c, FLS_PROF_GEMM_O, 2.0 * R * Hd * c->q_dim, st);
        a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
        CTX_TRY(c, gemm_launch(kEpiResidual, a, c->num_sms, st, c->gemm_cg));
    }

    // hidden = hidden + down_proj(silu(gate_proj(norm(hidden))) * up_proj(norm(hidden)))
    {
        ProfScope ps(c, FLS_PROF_RMSNORM, norm_bytes, st);
        CTX_TRY(c, rmsnorm_launch(x, nullptr, nullptr, r->dec.ln2, m.rms_norm_eps, n_rows, Hd, nullptr, c->xn, st,
                                  c->rmsnorm_variant));
    }
    a = GemmArgs{};
    a.A = c->xn, a.lda = Hd, a.B = r->dec.wgu, a.ldb = Hd, a.b_rows = 2 * I;
    a.M = n_rows, a.N = I, a.K = Hd, a.C = c->hmid, a.ldc = I, a.up_off = I;
    {
        ProfScope ps(c, FLS_PROF_GEMM_GATE_UP, 2.0 * R * (2.0 * I) * Hd, st);
        a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
        CTX_TRY(c, gemm_launch(kEpiSwiGLU, a, c->num_sms, st, c->gemm_cg));
    }
    a = GemmArgs{};
    a.A = c->hmid, a.lda = I, a.B = r->dec.wd, a.ldb = I, a.b_rows = Hd;
    a.M = n_rows, a.N = Hd, a.K = I, a.C = x, a.ldc = Hd, a.R = x, a.ldr = Hd;
    {
        ProfScope ps(c, FLS_PROF_GEMM_DOWN, 2.0 * R * Hd * I, st);
        a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
        CTX_TRY(c, gemm_launch(kEpiResidual, a, c->num_sms, st, c->gemm_cg));
    }
    c->stats.kernel_launches += fused_rope ? 7 : 8;
    return kOk;
}

int fls_final_norm(fls_ctx* c, int slot, int layer_idx, const uint16_t* x, const int32_t* gather, int n_out,
                   uint16_t* out, void* stream) {
    ENTER(c);
    const ResidentLayer* r = find_resident(c, slot, layer_idx, FLS_LAYER_NORM);
    if (!r) return ctx_fail(c, kErrState);
    {
        ProfScope ps(c, FLS_PROF_RMSNORM, static_cast<double>(n_out) * c->m.hidden_size * 4.0, as_stream(stream));
        CTX_TRY(c, rmsnorm_launch(H(x), nullptr, gather, r->single, c->m.rms_norm_eps, n_out, c->m.hidden_size,
                                  nullptr, H(out), as_stream(stream)));
    }
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_lm_head(fls_ctx* c, int slot, int layer_idx, const uint16_t* x, int n_rows, uint16_t* probs,
                const int32_t* answer_ids, int n_ids, float* logprob, float* lse_out, void* stream) {
    ENTER(c);
    const ResidentLayer* r = find_resident(c, slot, layer_idx, FLS_LAYER_HEAD);
    if (!r) return ctx_fail(c, kErrState);
    if (n_rows <= 0) return kOk;
    CTX_TRY(c, head_impl(c, H(x), r->single, n_rows, c->m.vocab_size, c->m.hidden_size, H(probs), answer_ids, n_ids,
                         logprob, lse_out, as_stream(stream)));
    return kOk;
}

int fls_peer_copy(fls_ctx* c, void* dst_dev, int dst_device, const void* src_dev, size_t bytes, void* stream) {
    ENTER(c);
    if (!dst_dev || !src_dev) return ctx_fail(c, fail(kErrInvalid, "fls_peer_copy: null pointer"));
    if (bytes == 0) return kOk;
    if (dst_device < 0) {
        // `dst_dev` is a mapping made by fls_ipc_open: unified addressing knows which GPU it lives on
        CTX_CUDA(c, cudaMemcpyAsync(dst_dev, src_dev, bytes, cudaMemcpyDefault, as_stream(stream)));
        c->stats.p2p_bytes += bytes;
        return kOk;
    }
    if (dst_device != c->device && !(c->peer_enabled_mask & (1ull << (dst_device & 63)))) {
        int can = 0;
        CTX_CUDA(c, cudaDeviceCanAccessPeer(&can, c->device, dst_device));
        if (can) {
            cudaError_t e = cudaDeviceEnablePeerAccess(dst_device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return ctx_fail(c, fail(kErrCuda, "cudaDeviceEnablePeerAccess(%d): %s", dst_device, cudaGetErrorString(e)));
            cudaGetLastError();
        }  // without peer access cudaMemcpyPeerAsync still works (staged by the driver)
        c->peer_enabled_mask |= 1ull << (dst_device & 63);
    }
    CTX_CUDA(c, cudaMemcpyPeerAsync(dst_dev, dst_device, src_dev, c->device, bytes, as_stream(stream)));
    c->stats.p2p_bytes += bytes;
    return kOk;
}

// --------------------------------------------------------------------------------- hand-off across processes
int fls_ipc_alloc(fls_ctx* c, size_t bytes, void** dev_out, unsigned char* handle_out) {
    ENTER(c);
    if (!dev_out || !handle_out || bytes == 0) return ctx_fail(c, fail(kErrInvalid, "fls_ipc_alloc: bad argument"));
    static_assert(sizeof(cudaIpcMemHandle_t) == FLS_IPC_HANDLE_BYTES, "handle size");
    void* p = nullptr;
    CTX_CUDA(c, cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return ctx_fail(c, fail(kErrCuda, "fls_ipc_alloc(%zu bytes): %s", bytes, cudaGetErrorString(e)));
    }
    memcpy(handle_out, &h, sizeof(h));
    c->ipc_owned.push_back(p);
    c->workspace_bytes += bytes;
    *dev_out = p;
    return kOk;
}

int fls_ipc_free(fls_ctx* c, void* dev) {
    ENTER(c);
    auto it = std::find(c->ipc_owned.begin(), c->ipc_owned.end(), dev);
    if (it == c->ipc_owned.end()) return ctx_fail(c, fail(kErrInvalid, "fls_ipc_free: not a buffer of this context"));
    CTX_CUDA(c, cudaDeviceSynchronize());
    c->ipc_owned.erase(it);
    CTX_CUDA(c, cudaFree(dev));
    return kOk;
}

int fls_ipc_open(fls_ctx* c, const unsigned char* handle, void** dev_out) {
    ENTER(c);
    if (!handle || !dev_out) return ctx_fail(c, fail(kErrInvalid, "fls_ipc_open: null argument"));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    // the exporting GPU becomes peer-accessible from this one as part of the mapping (NVLink / NVSwitch)
    CTX_CUDA(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->ipc_mapped.push_back(p);
    *dev_out = p;
    return kOk;
}

int fls_ipc_close(fls_ctx* c, void* dev) {
    ENTER(c);
    auto it = std::find(c->ipc_mapped.begin(), c->ipc_mapped.end(), dev);
    if (it == c->ipc_mapped.end()) return ctx_fail(c, fail(kErrInvalid, "fls_ipc_close: not a mapping of this context"));
    CTX_CUDA(c, cudaDeviceSynchronize());
    c->ipc_mapped.erase(it);
    CTX_CUDA(c, cudaIpcCloseMemHandle(dev));
    return kOk;
}

int fls_flag_signal(fls_ctx* c, uint32_t* flag_dev, uint32_t value, void* stream) {
    ENTER(c);
    if (!flag_dev) return ctx_fail(c, fail(kErrInvalid, "fls_flag_signal: null flag"));
    CTX_TRY(c, flag_signal_launch(flag_dev, value, as_stream(stream)));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_flag_wait(fls_ctx* c, const uint32_t* flag_dev, uint32_t target, uint32_t timeout_ms, void* stream) {
    ENTER(c);
    if (!flag_dev) return ctx_fail(c, fail(kErrInvalid, "fls_flag_wait: null flag"));
    if (!c->flag_timeouts) {
        CTX_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->flag_timeouts), 256));
        CTX_CUDA(c, cudaMemset(c->flag_timeouts, 0, 256));
    }
    CTX_TRY(c, flag_wait_launch(flag_dev, target, timeout_ms, c->flag_timeouts, as_stream(stream)));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_flag_timeouts(fls_ctx* c, uint32_t* count_out) {
    ENTER(c);
    if (!count_out) return ctx_fail(c, fail(kErrInvalid, "fls_flag_timeouts: null output"));
    *count_out = 0;
    if (c->flag_timeouts)
        CTX_CUDA(c, cudaMemcpy(count_out, c->flag_timeouts, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return kOk;
}

// --------------------------------------------------------------------------------- single kernels
int fls_op_gemm(fls_ctx* c, int epilogue, const uint16_t* a_dev, int lda, const uint16_t* b_dev, int ldb, int b_rows,
                int m, int n, int k, uint16_t* c_dev, int ldc, const uint16_t* r_dev, int ldr, int up_off,
                void* stream) {
    ENTER(c);
    if (epilogue < FLS_EPI_STORE || epilogue > FLS_EPI_SWIGLU)
        return ctx_fail(c, fail(kErrInvalid, "fls_op_gemm: epilogue %d", epilogue));
    GemmArgs a{};
    a.A = H(a_dev), a.lda = lda, a.B = H(b_dev), a.ldb = ldb, a.b_rows = b_rows, a.M = m, a.N = n, a.K = k;
    a.C = H(c_dev), a.ldc = ldc, a.R = H(r_dev), a.ldr = ldr, a.up_off = up_off;
    a.group_m = c->gemm_group_m;
    a.cache_hints = c->gemm_cache_hints;
    a.raster = c->gemm_raster;
    a.direct_store = !c->gemm_tma_store, a.wide = (c->gemm_wide == 1 || (c->gemm_wide == 2 && a.K >= 8192)) ? (c->gemm_wide_lockstep ? 3 : 1) : 0;
    CTX_TRY(c, gemm_launch(static_cast<GemmEpilogue>(epilogue), a, c->num_sms, as_stream(stream), c->gemm_cg));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_op_rmsnorm(fls_ctx* c, const uint16_t* x, const uint16_t* delta, const int32_t* gather, const uint16_t* w,
                   float eps, int n_rows, int hidden, uint16_t* x_sum, uint16_t* out, void* stream) {
    ENTER(c);
    CTX_TRY(c, rmsnorm_launch(H(x), H(delta), gather, H(w), eps, n_rows, hidden, H(x_sum), H(out), as_stream(stream),
                              c->rmsnorm_variant));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_op_rope(fls_ctx* c, uint16_t* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                const int32_t* row_pos, void* stream) {
    ENTER(c);
    if (!c->rope_cos || head_dim != c->m.head_dim)
        return ctx_fail(c, fail(kErrState, "fls_op_rope: rope table missing or head_dim mismatch"));
    CTX_TRY(c, rope_launch(H(qkv), ld, n_rows, n_q_heads, n_kv_heads, head_dim, row_pos, c->rope_cos, c->rope_sin,
                           c->rope_n_pos, as_stream(stream)));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_op_attention(fls_ctx* c, const uint16_t* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                     const fls_attn_tile* tiles, int n_tiles, uint16_t* out, int ld_out, void* stream) {
    ENTER(c);
    CTX_TRY(c, attention_dispatch(c, H(qkv), ld, n_rows, n_q_heads, n_kv_heads, head_dim,
                                  reinterpret_cast<const AttnTile*>(tiles), n_tiles, H(out), ld_out,
                                  as_stream(stream)));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_op_swiglu(fls_ctx* c, const uint16_t* gate, const uint16_t* up, int ld_in, int n_rows, int inter,
                  uint16_t* out, int ld_out, void* stream) {
    ENTER(c);
    CTX_TRY(c, swiglu_launch(H(gate), H(up), ld_in, n_rows, inter, H(out), ld_out, as_stream(stream)));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_op_embed(fls_ctx* c, const uint16_t* table, const int32_t* ids, int n_rows, int hidden, uint16_t* out,
                 void* stream) {
    ENTER(c);
    CTX_TRY(c, embed_launch(H(table), ids, n_rows, hidden, H(out), as_stream(stream)));
    c->stats.kernel_launches += 1;
    return kOk;
}

int fls_op_head(fls_ctx* c, const uint16_t* x, const uint16_t* w, int n_rows, int n, int k, uint16_t* probs,
                const int32_t* answer_ids, int n_ids, float* logprob, float* lse_out, void* stream) {
    ENTER(c);
    if (!c->lse_partial) return ctx_fail(c, fail(kErrState, "fls_op_head: call fls_configure first"));
    CTX_TRY(c, head_impl(c, H(x), H(w), n_rows, n, k, H(probs), answer_ids, n_ids, logprob, lse_out,
                         as_stream(stream)));
    return kOk;
}

int fls_set_option(fls_ctx* c, const char* name, int value) {
    if (!c || !name) return fail(kErrInvalid, "fls_set_option: null argument");
    if (strcmp(name, "gemm_cta_group") == 0) {
        if (value != 1 && value != 2) return ctx_fail(c, fail(kErrInvalid, "gemm_cta_group must be 1 or 2"));
        c->gemm_cg = value;
        return kOk;
    }
    if (strcmp(name, "rmsnorm_variant") == 0) {
        if (value < 0 || value > 2) return ctx_fail(c, fail(kErrInvalid, "rmsnorm_variant must be 0..2"));
        c->rmsnorm_variant = value;
        return kOk;
    }
    if (strcmp(name, "gemm_cache_hints") == 0) {
        if (value < 0 || value > 4) return ctx_fail(c, fail(kErrInvalid, "gemm_cache_hints must be 0..4"));
        c->gemm_cache_hints = value;
        return kOk;
    }
    if (strcmp(name, "paced_weight_stream") == 0) {
        c->paced_stream = value != 0;
        return kOk;
    }
    if (strcmp(name, "gemm_raster") == 0) {
        if (value < 0 || value > 3) return ctx_fail(c, fail(kErrInvalid, "gemm_raster must be 0..3"));
        c->gemm_raster = value;
        return kOk;
    }
    if (strcmp(name, "gemm_group_m") == 0) {
        if (value < 0 || value > 1024) return ctx_fail(c, fail(kErrInvalid, "gemm_group_m out of range"));
        c->gemm_group_m = value;
        return kOk;
    }
    if (strcmp(name, "gemm_wide_lockstep") == 0) {
        c->gemm_wide_lockstep = value != 0;
        return kOk;
    }
    if (strcmp(name, "gemm_wide_tiles") == 0) {
        if (value < 0 || value > 2) return ctx_fail(c, fail(kErrInvalid, "gemm_wide_tiles must be 0, 1 or 2"));
        c->gemm_wide = value;
        return kOk;
    }
    if (strcmp(name, "gemm_tma_store") == 0) {
        c->gemm_tma_store = value != 0;
        return kOk;
    }
    if (strcmp(name, "fuse_rope") == 0) {
        if (value < 0 || value > 2) return ctx_fail(c, fail(kErrInvalid, "fuse_rope must be 0, 1 or 2"));
        c->fuse_rope = value;
        return kOk;
    }
    if (strcmp(name, "attention_impl") == 0) {
        if (value < 0 || value > 2) return ctx_fail(c, fail(kErrInvalid, "attention_impl must be 0, 1 or 2"));
        c->attn_impl = value;
        return kOk;
    }
    if (strcmp(name, "attention_k_stages") == 0) {
        if (value != 0 && value != 2 && value != 3)
            return ctx_fail(c, fail(kErrInvalid, "attention_k_stages must be 0 (this device's default), 2 or 3"));
        c->attn_k_stages = value;
        return kOk;
    }
    return ctx_fail(c, fail(kErrInvalid, "fls_set_option: unknown option '%s'", name));
}

int fls_set_profiling(fls_ctx* c, int enabled) {
    if (!c) return fail(kErrInvalid, "fls_set_profiling: null context");
    c->profiling = enabled != 0;
    return kOk;
}

int fls_get_profile(fls_ctx* c, fls_kernel_profile* out) {
    ENTER(c);
    if (!out) return ctx_fail(c, fail(kErrInvalid, "fls_get_profile: null output"));
    for (auto& p : c->prof_ring) prof_harvest(c, p, true);
    for (int i = 0; i < FLS_PROF_CLASSES; ++i) out[i] = c->prof[i];
    return kOk;
}

int fls_get_stats(fls_ctx* c, fls_stats* out) {
    if (!c || !out) return fail(kErrInvalid, "fls_get_stats: null argument");
    cudaSetDevice(c->device);
    for (auto& t : c->copy_ring) harvest_copy_timing(c, t, true);
    c->stats.file_bytes_read = host_cache_file_bytes();
    c->stats.pinned_bytes = host_cache_pinned_bytes();
    c->stats.mapped_bytes = host_cache_mapped_bytes();
    c->stats.workspace_bytes = c->workspace_bytes;
    *out = c->stats;
    return kOk;
}

int fls_reset_stats(fls_ctx* c) {
    if (!c) return fail(kErrInvalid, "fls_reset_stats: null context");
    c->stats = fls_stats{};
    for (auto& t : c->copy_ring) t.pending = false;
    for (auto& p : c->prof_ring) p.pending = false;
    for (auto& p : c->prof) p = fls_kernel_profile{};
    return kOk;
}

}  // extern "C"
