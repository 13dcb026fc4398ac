// Shared declarations for the fls_b200 library (internal; the public surface is include/fls_b200.h).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

struct CUtensorMap_st;  // <cuda.h>

namespace fls {

// ---- error plumbing: every internal launcher returns 0 or a negative fls status and leaves a
// message in a thread-local buffer that the C-ABI layer copies into the context.
constexpr int kOk = 0;
constexpr int kErrInvalid = -1;   // bad argument
constexpr int kErrCuda = -2;      // CUDA runtime/driver error
constexpr int kErrIo = -3;        // file / parse error
constexpr int kErrState = -4;     // call sequence error
constexpr int kErrCapacity = -5;  // exceeds configured workspace / slot capacity

char* tls_error_buffer();  // 512 bytes
int fail(int code, const char* fmt, ...);

#define FLS_CUDA(expr)                                                                           \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess)                                                                   \
            return ::fls::fail(::fls::kErrCuda, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                               __FILE__, __LINE__);                                              \
    } while (0)

#define FLS_TRY(expr)          \
    do {                       \
        int _s = (expr);       \
        if (_s != 0) return _s; \
    } while (0)

// ---- GEMM (gemm_sm100.cu) ---------------------------------------------------------------
// C[M,N] = A[M,K] * B[N,K]^T ; A, B fp16 row-major (K contiguous), fp32 accumulation in TMEM.
enum GemmEpilogue : int {
    kEpiStore = 0,       // C = fp16(acc)
    kEpiResidual = 1,    // C = R + fp16(acc)            (R may alias C)
    kEpiSwiGLU = 2,      // C[M,N] = fp16(silu(fp16(acc_gate))) * fp16(acc_up); B rows [0,N) = gate, [up_off, up_off+N) = up
    kEpiLsePartial = 3,  // partial[row, n_blk] = (max, sum exp) of fp16(acc) over the tile's columns
    kEpiProbs = 4,       // C = fp16(exp(fp16(acc) - lse[row]))
    kEpiStoreRope = 5,   // C = fp16(acc), with HF rotary embedding applied to the 128-wide heads in columns [0, rope_cols)
};

struct GemmArgs {
    const __half* A;
    int lda;
    const __half* B;
    int ldb;
    int b_rows;  // rows of B addressable through the tensor map (>= N, or >= up_off + N for SwiGLU)
    int M, N, K;
    __half* C;
    int ldc;
    const __half* R;
    int ldr;
    float2* partial;  // [M, ceil(N/256)]
    const float* lse;  // [M]
    int up_off;
    int group_m;      // m-blocks per raster group (0 = default 16)
    int cache_hints;  // cta_group 2 only: 0 none, 1 A evict-first / B evict-last, 2 the reverse
    int raster;        // tile order (gemm_tile.cuh tile_coords): 0 m-fastest groups (default), 1 n-fastest, 2/3 serpentine
    int wide;          // 1: 256 x 512 tiles per CTA pair (gemm_wide_sm100.cu) where the epilogue has that variant
    int direct_store;  // 0 (default): epilogue stages tiles in smem and writes them with TMA; 1: per-thread 16-byte stores
    // kEpiStoreRope: rotary position of every row and the fp16 cos/sin tables [n_pos, 128]
    const int32_t* row_pos;
    const __half* cos_tab;
    const __half* sin_tab;
    int rope_cols;
};
int gemm_launch(GemmEpilogue epi, const GemmArgs& a, int num_sms, cudaStream_t stream, int cta_group = 2);
// 256 x 512-tile variant (cta_group::2, staged epilogue): store, store+RoPE, residual and SwiGLU epilogues
int gemm_wide_init();
bool gemm_wide_supports(GemmEpilogue epi);
int gemm_wide_launch(GemmEpilogue epi, const GemmArgs& a, int num_sms, cudaStream_t stream);
int gemm_init();  // resolves cuTensorMapEncodeTiled and opts kernels into large dynamic smem
int gemm_lse_blocks(int N);

// ---- element-wise / reductions (elementwise.cu) ---------------------------------------------
int embed_launch(const __half* table, const int32_t* ids, int n_rows, int H, __half* out, cudaStream_t s);
// Copies `width` columns from `col0` of qkv rows (rows[r], or r when rows == nullptr) to dense cache rows, or back.
int kv_rows_copy_launch(__half* qkv, int ld, int col0, int width, const int32_t* rows, int n, __half* cache,
                        bool to_cache, cudaStream_t s);
// out_norm[r] = w * fp16(x32 * rsqrt(mean(x32^2) + eps)), x = x_in[gather ? gather[r] : r] (+ delta[r]);
// when delta != nullptr the sum is also written back to x_sum (may alias x_in).
int rmsnorm_launch(const __half* x_in, const __half* delta, const int32_t* gather, const __half* w, float eps,
                   int n_rows, int H, __half* x_sum, __half* out_norm, cudaStream_t s, int variant = 0);
// In-place rotary embedding on the q and k column blocks of a packed qkv buffer.
int rope_launch(__half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                const int32_t* row_pos, const __half* cos_tab, const __half* sin_tab, int n_pos, cudaStream_t s);
int swiglu_launch(const __half* gate, const __half* up, int ld_in, int n_rows, int I, __half* out, int ld_out,
                  cudaStream_t s);
int lse_reduce_launch(const float2* partial, int n_rows, int n_blocks, float* lse, cudaStream_t s);
// out[r, j] = float(fp16(dot(x[r], W[ids[r*n_ids + j]]))) - lse[r]; NaN where the id is outside [0, n_vocab)
int gather_logprob_launch(const __half* x, int ldx, const __half* W, int ldw, int H, const int32_t* ids, int n_ids,
                          const float* lse, int n_rows, int n_vocab, float* out, cudaStream_t s);

// flag = value once everything enqueued on `s` before has completed (system-scope release; `flag` may be a peer mapping)
int flag_signal_launch(uint32_t* flag, uint32_t value, cudaStream_t s);
// holds `s` until *flag >= target (wrap-safe); after timeout_ms gives up and increments *timed_out instead
int flag_wait_launch(const uint32_t* flag, uint32_t target, uint32_t timeout_ms, uint32_t* timed_out, cudaStream_t s);

// ---- attention (attention.cu) ------------------------------------------------------------------
// One descriptor per query tile; see include/fls_b200.h fls_attn_tile.
enum AttnMode : int32_t { kAttnPrefix = 0, kAttnPrefixCausal = 1, kAttnSuffix = 2 };
struct AttnTile {
    int32_t q_row;     // first query row (packed row index)
    int32_t n_q;       // queries in this tile (<= 128)
    int32_t q_seq0;    // mode 1: sequence index of the first query; mode 2: option length (len_s)
    int32_t mode;      // AttnMode
    int32_t seg0_row;  // first key/value row of segment 0 (the prompt's prefix rows)
    int32_t seg0_len;
    int32_t seg1_row;  // mode 2: first row of the (whole) options the queries belong to
    int32_t seg1_len;  //         rows of those options that can be visible to the tile
};
int attention_launch(const __half* qkv, int ld, int n_q_heads, int n_kv_heads, int head_dim, const AttnTile* tiles,
                     int n_tiles, __half* out, int ld_out, cudaStream_t s);
int attention_init();
// tcgen05 version (head_dim 128); n_rows = rows of the qkv buffer (bounds the TMA tensor map)
// k_stages: K buffers per CTA - 0 = what attention_sm100_init chose for this device (3 where two CTAs still fit an SM), 2, 3
int attention_sm100_launch(const __half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, const AttnTile* tiles,
                           int n_tiles, __half* out, int ld_out, cudaStream_t s, int k_stages = 0);
// persistent CTAs fetching (tile, head) items from `work_ctr` (two zeroed uint32 of the caller's context)
int attention_sm100_launch_persistent(const __half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads,
                                      const AttnTile* tiles, int n_tiles, __half* out, int ld_out, cudaStream_t s,
                                      uint32_t* work_ctr, int num_sms);
int attention_sm100_init();
int attention_sm100_k_stages();   // the choice made for the current device (0 before init)

// TMA descriptor over a row-major fp16 matrix [rows, cols] (row stride ld), box = 64 columns x box_rows,
// 128-byte swizzle (gemm_sm100.cu).
int encode_tmap_2d(::CUtensorMap_st* tm, const __half* base, int rows, int cols, int ld, int box_rows);

}  // namespace fls
