// HBM-bound row kernels: embedding gather, fused (residual +) RMSNorm (+ row gather), rotary
// embedding, SwiGLU (unit-test twin of the GEMM epilogue), softmax log-sum-exp reduction and the
// answer-token log-prob gather.  All use 128-bit coalesced accesses; one CTA handles one row (or a
// few), grids are sized by rows so they cover the 148 SMs many times over at benchmark sizes.
//
// Rounding follows HF Llama eager in fp16, which is what the reference executes
// (reference utils.py:267-268,272,278,286): see each kernel.
#include "common.cuh"
#include "ptx.cuh"

namespace fls {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---------------------------------------------------------------------------------- embedding
// out[r, :] = table[ids[r], :]                      (reference utils.py:267-268, nn.Embedding)
__global__ void embed_kernel(const __half* __restrict__ table, const int32_t* __restrict__ ids, int n_rows, int H,
                             __half* __restrict__ out) {
    const int vec_per_row = H / 8;
    const long long total = static_cast<long long>(n_rows) * vec_per_row;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / vec_per_row);
        const int v = static_cast<int>(i - static_cast<long long>(r) * vec_per_row);
        const uint4 val = ptx::ld_nc_v4(table + static_cast<size_t>(ids[r]) * H + v * 8);
        *reinterpret_cast<uint4*>(out + static_cast<size_t>(r) * H + v * 8) = val;
    }
}

// ---------------------------------------------------------------------------------- K|V row copy
// Prefix K/V cache (no counterpart in the reference, which recomputes the prefix for every generated token,
// main.py:65): `width` columns starting at `col0` of selected rows of the qkv workspace <-> dense cache rows.
//   to_cache:  cache[r, :] = qkv[rows ? rows[r] : r, col0 : col0 + width]
//   !to_cache: qkv[rows ? rows[r] : r, col0 : col0 + width] = cache[r, :]
__global__ void kv_rows_copy_kernel(__half* __restrict__ qkv, int ld, int col0, int width,
                                    const int32_t* __restrict__ rows, int n, __half* __restrict__ cache,
                                    int to_cache) {
    const int vec_per_row = width / 8;
    const long long total = static_cast<long long>(n) * vec_per_row;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / vec_per_row);
        const int v = static_cast<int>(i - static_cast<long long>(r) * vec_per_row);
        const int row = rows ? rows[r] : r;
        uint4* q = reinterpret_cast<uint4*>(qkv + static_cast<size_t>(row) * ld + col0 + v * 8);
        uint4* k = reinterpret_cast<uint4*>(cache + static_cast<size_t>(r) * width + v * 8);
        if (to_cache) *k = *q;
        else *q = *k;
    }
}

// ---------------------------------------------------------------------------------- RMSNorm
// HF LlamaRMSNorm: x32 = x.float(); y = (x32 * rsqrt(mean(x32^2) + eps)).half(); out = w * y.
// Optional fused residual (x = x_in + delta, rounded to fp16 first, written back to x_sum) and
// optional row gather (x = x_in[gather[r]]: the last-real-token selection of utils.py:284-286).
// One CTA of THREADS per row; each thread keeps VPT 16-byte vectors in registers -> one pass.
template <int THREADS, int VPT>
__global__ void __launch_bounds__(THREADS)
rmsnorm_kernel(const __half* __restrict__ x_in, const __half* __restrict__ delta, const int32_t* __restrict__ gather,
               const __half* __restrict__ w, float eps, int n_rows, int H, __half* __restrict__ x_sum,
               __half* __restrict__ out) {
    __shared__ float red[THREADS / 32];
    const int r = blockIdx.x;
    if (r >= n_rows) return;
    const int src = gather ? gather[r] : r;
    const int nvec = H / 8;
    const __half* xr = x_in + static_cast<size_t>(src) * H;
    uint4 v[VPT];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int vi = threadIdx.x + i * THREADS;
        if (vi < nvec) {
            v[i] = *reinterpret_cast<const uint4*>(xr + vi * 8);
            if (delta) {
                const uint4 d = *reinterpret_cast<const uint4*>(delta + static_cast<size_t>(r) * H + vi * 8);
                __half2* a = reinterpret_cast<__half2*>(&v[i]);
                const __half2* b = reinterpret_cast<const __half2*>(&d);
#pragma unroll
                for (int j = 0; j < 4; ++j) a[j] = __hadd2(a[j], b[j]);
                *reinterpret_cast<uint4*>(x_sum + static_cast<size_t>(r) * H + vi * 8) = v[i];
            }
            const __half2* h = reinterpret_cast<const __half2*>(&v[i]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = __half22float2(h[j]);
                ss += f.x * f.x + f.y * f.y;
            }
        }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < THREADS / 32; ++i) tot += red[i];
    const float rstd = rsqrtf(tot / static_cast<float>(H) + eps);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int vi = threadIdx.x + i * THREADS;
        if (vi < nvec) {
            const uint4 wv = *reinterpret_cast<const uint4*>(w + vi * 8);
            const __half2* h = reinterpret_cast<const __half2*>(&v[i]);
            const __half2* wh = reinterpret_cast<const __half2*>(&wv);
            uint4 o;
            __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = __half22float2(h[j]);
                const __half2 y = __floats2half2_rn(f.x * rstd, f.y * rstd);  // .to(fp16) before the gain
                oh[j] = __hmul2(wh[j], y);
            }
            *reinterpret_cast<uint4*>(out + static_cast<size_t>(r) * H + vi * 8) = o;
        }
    }
}

// ---------------------------------------------------------------------------------- RoPE
// HF apply_rotary_pos_emb, half-split convention, evaluated in fp16 with a rounding after each
// product and after the sum:  out[i]     = x[i]*cos[i]     + (-x[i+d/2])*sin[i]
//                             out[i+d/2] = x[i+d/2]*cos[i+d/2] + x[i]*sin[i+d/2]
// cos/sin tables are [n_pos, d] fp16 = fp16(cos(fp32 angle)), supplied by the host exactly as the
// reference's rotary buffers are built and cast (reference utils.py:117-119).
// One thread rotates 8 consecutive dims of the first half and their partners in the second half.
__global__ void rope_kernel(__half* __restrict__ qkv, int ld, int n_rows, int n_heads_qk, int head_dim,
                            const int32_t* __restrict__ row_pos, const __half* __restrict__ cos_tab,
                            const __half* __restrict__ sin_tab) {
    const int half_vecs = head_dim / 16;  // 16-byte vectors in half a head
    const long long total = static_cast<long long>(n_rows) * n_heads_qk * half_vecs;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int v = static_cast<int>(i % half_vecs);
        const long long t = i / half_vecs;
        const int head = static_cast<int>(t % n_heads_qk);
        const int r = static_cast<int>(t / n_heads_qk);
        const int pos = row_pos[r];
        __half* x = qkv + static_cast<size_t>(r) * ld + head * head_dim + v * 8;
        const __half* c = cos_tab + static_cast<size_t>(pos) * head_dim + v * 8;
        const __half* s = sin_tab + static_cast<size_t>(pos) * head_dim + v * 8;
        const int hd2 = head_dim / 2;
        uint4 x1 = *reinterpret_cast<const uint4*>(x);
        uint4 x2 = *reinterpret_cast<const uint4*>(x + hd2);
        const uint4 c1 = *reinterpret_cast<const uint4*>(c);
        const uint4 c2 = *reinterpret_cast<const uint4*>(c + hd2);
        const uint4 s1 = *reinterpret_cast<const uint4*>(s);
        const uint4 s2 = *reinterpret_cast<const uint4*>(s + hd2);
        const __half2* a1 = reinterpret_cast<const __half2*>(&x1);
        const __half2* a2 = reinterpret_cast<const __half2*>(&x2);
        const __half2* hc1 = reinterpret_cast<const __half2*>(&c1);
        const __half2* hc2 = reinterpret_cast<const __half2*>(&c2);
        const __half2* hs1 = reinterpret_cast<const __half2*>(&s1);
        const __half2* hs2 = reinterpret_cast<const __half2*>(&s2);
        uint4 o1, o2;
        __half2* p1 = reinterpret_cast<__half2*>(&o1);
        __half2* p2 = reinterpret_cast<__half2*>(&o2);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            // *_rn intrinsics: never contracted into an fma, so each product is rounded on its own
            p1[j] = __hadd2_rn(__hmul2_rn(a1[j], hc1[j]), __hmul2_rn(__hneg2(a2[j]), hs1[j]));
            p2[j] = __hadd2_rn(__hmul2_rn(a2[j], hc2[j]), __hmul2_rn(a1[j], hs2[j]));
        }
        *reinterpret_cast<uint4*>(x) = o1;
        *reinterpret_cast<uint4*>(x + hd2) = o2;
    }
}

// ---------------------------------------------------------------------------------- SwiGLU
__global__ void swiglu_kernel(const __half* __restrict__ gate, const __half* __restrict__ up, int ld_in, int n_rows,
                              int I, __half* __restrict__ out, int ld_out) {
    const int vec_per_row = I / 8;
    const long long total = static_cast<long long>(n_rows) * vec_per_row;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / vec_per_row);
        const int v = static_cast<int>(i - static_cast<long long>(r) * vec_per_row);
        const uint4 g = *reinterpret_cast<const uint4*>(gate + static_cast<size_t>(r) * ld_in + v * 8);
        const uint4 u = *reinterpret_cast<const uint4*>(up + static_cast<size_t>(r) * ld_in + v * 8);
        const __half* gh = reinterpret_cast<const __half*>(&g);
        const __half* uh = reinterpret_cast<const __half*>(&u);
        uint4 o;
        __half* oh = reinterpret_cast<__half*>(&o);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gf = __half2float(gh[j]);
            oh[j] = __hmul(__float2half_rn(gf / (1.0f + expf(-gf))), uh[j]);
        }
        *reinterpret_cast<uint4*>(out + static_cast<size_t>(r) * ld_out + v * 8) = o;
    }
}

// ---------------------------------------------------------------------------------- LSE reduce
// lse[r] = log(sum_j exp(logit[r,j])) from the per-tile (max, sum) partials of the lm_head GEMM.
__global__ void lse_reduce_kernel(const float2* __restrict__ partial, int n_rows, int n_blocks,
                                  float* __restrict__ lse) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= n_rows) return;
    const int lane = threadIdx.x & 31;
    float m = -INFINITY;
    for (int j = lane; j < n_blocks; j += 32) m = fmaxf(m, partial[static_cast<size_t>(r) * n_blocks + j].x);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int j = lane; j < n_blocks; j += 32) {
        const float2 p = partial[static_cast<size_t>(r) * n_blocks + j];
        s += p.y * expf(p.x - m);
    }
    s = warp_sum(s);
    if (lane == 0) lse[r] = m + logf(s);
}

// ---------------------------------------------------------------------------------- log-prob gather
// One warp per (row, id): logit = fp16(dot(x[r], W[id])) (the value lm_head's fp16 output would
// hold), out = logit - lse[r].
__global__ void gather_logprob_kernel(const __half* __restrict__ x, int ldx, const __half* __restrict__ W, int ldw,
                                      int H, const int32_t* __restrict__ ids, int n_ids,
                                      const float* __restrict__ lse, int n_rows, int n_vocab,
                                      float* __restrict__ out) {
    const long long wid = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= static_cast<long long>(n_rows) * n_ids) return;
    const int r = static_cast<int>(wid / n_ids);
    const int lane = threadIdx.x & 31;
    const int id = ids[wid];
    if (id < 0 || id >= n_vocab) {  // ids live on the device: an id outside W's rows must not become a wild read
        if (lane == 0) out[wid] = __int_as_float(0x7fc00000);  // NaN marks the entry
        return;
    }
    const __half* xr = x + static_cast<size_t>(r) * ldx;
    const __half* wr = W + static_cast<size_t>(id) * ldw;
    float acc = 0.f;
    for (int k = lane * 8; k < H; k += 256) {
        const uint4 a = *reinterpret_cast<const uint4*>(xr + k);
        const uint4 b = ptx::ld_nc_v4(wr + k);
        const __half2* ah = reinterpret_cast<const __half2*>(&a);
        const __half2* bh = reinterpret_cast<const __half2*>(&b);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 fa = __half22float2(ah[j]), fb = __half22float2(bh[j]);
            acc = fmaf(fa.x, fb.x, acc);
            acc = fmaf(fa.y, fb.y, acc);
        }
    }
    acc = warp_sum(acc);
    if (lane == 0) out[wid] = __half2float(__float2half_rn(acc)) - lse[r];
}

// Hand-off flags between GPUs (model-parallel pipeline across processes).  A flag is a monotonically growing
// 32-bit counter in the WAITING GPU's memory; the signalling GPU writes it through a peer (IPC) mapping.
__global__ void flag_signal_kernel(uint32_t* flag, uint32_t value) {
    // everything this stream did before (kernels, peer copies) is complete and visible before the flag moves
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(flag), "r"(value) : "memory");
}

__global__ void flag_wait_kernel(const uint32_t* flag, uint32_t target, unsigned long long timeout_ns,
                                 uint32_t* timed_out) {
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t0));
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(flag) : "memory");
        if (static_cast<int32_t>(v - target) >= 0) return;      // counters only grow (wrap-safe comparison)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
        if (t - t0 > timeout_ns) {                               // a dead peer must not hang this GPU for ever
            atomicAdd(timed_out, 1u);
            return;
        }
        __nanosleep(256);
    }
}

int grid_for(long long work_items, int threads) {
    long long b = (work_items + threads - 1) / threads;
    const long long cap = 148LL * 16;  // 16 resident CTAs of 256 threads per SM cover latency; loop beyond
    return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace

int embed_launch(const __half* table, const int32_t* ids, int n_rows, int H, __half* out, cudaStream_t s) {
    if (H % 8) return fail(kErrInvalid, "embed: hidden size %d not a multiple of 8", H);
    if (n_rows <= 0) return kOk;
    embed_kernel<<<grid_for(static_cast<long long>(n_rows) * (H / 8), 256), 256, 0, s>>>(table, ids, n_rows, H, out);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int kv_rows_copy_launch(__half* qkv, int ld, int col0, int width, const int32_t* rows, int n, __half* cache,
                        bool to_cache, cudaStream_t s) {
    if (width % 8 || col0 % 8 || ld % 8)
        return fail(kErrInvalid, "kv_rows_copy: ld %d, col0 %d, width %d must be multiples of 8", ld, col0, width);
    if (n <= 0) return kOk;
    kv_rows_copy_kernel<<<grid_for(static_cast<long long>(n) * (width / 8), 256), 256, 0, s>>>(
        qkv, ld, col0, width, rows, n, cache, to_cache ? 1 : 0);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int rmsnorm_launch(const __half* x_in, const __half* delta, const int32_t* gather, const __half* w, float eps,
                   int n_rows, int H, __half* x_sum, __half* out_norm, cudaStream_t s, int variant) {
    if (H % 8) return fail(kErrInvalid, "rmsnorm: hidden size %d not a multiple of 8", H);
    if (delta && !x_sum) return fail(kErrInvalid, "rmsnorm: residual input needs an x_sum output");
    if (n_rows <= 0) return kOk;
    const int nvec = H / 8;
#define FLS_NORM(T, V) rmsnorm_kernel<T, V><<<n_rows, T, 0, s>>>(x_in, delta, gather, w, eps, n_rows, H, x_sum, out_norm)
    // variant 0: the default shapes; 1: half the threads, twice the vectors per thread (more CTAs and more
    // loads in flight per SM); 2: twice the threads (tools/norm_probe.py compares them)
    if (nvec <= 128) {
        if (variant == 1) FLS_NORM(64, 2); else FLS_NORM(128, 1);
    } else if (nvec <= 256) {
        if (variant == 1) FLS_NORM(128, 2); else FLS_NORM(256, 1);
    } else if (nvec <= 512) {
        if (variant == 1) FLS_NORM(128, 4); else if (variant == 2) FLS_NORM(512, 1); else FLS_NORM(256, 2);
    } else if (nvec <= 1024) {
        if (variant == 1) FLS_NORM(128, 8); else if (variant == 2) FLS_NORM(512, 2); else FLS_NORM(256, 4);
    } else if (nvec <= 2048) {
        if (variant == 1) FLS_NORM(256, 8); else FLS_NORM(512, 4);
    } else {
        return fail(kErrInvalid, "rmsnorm: hidden size %d above the supported 16384", H);
    }
#undef FLS_NORM
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int rope_launch(__half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, int head_dim, const int32_t* row_pos,
                const __half* cos_tab, const __half* sin_tab, int n_pos, cudaStream_t s) {
    (void)n_pos;
    if (head_dim % 16 || ld % 8) return fail(kErrInvalid, "rope: head_dim %% 16 and ld %% 8 required");
    if (n_rows <= 0) return kOk;
    // q heads and k heads are adjacent column blocks [0, (nq+nkv)*d) of the packed qkv row
    const int heads = n_q_heads + n_kv_heads;
    const long long items = static_cast<long long>(n_rows) * heads * (head_dim / 16);
    rope_kernel<<<grid_for(items, 256), 256, 0, s>>>(qkv, ld, n_rows, heads, head_dim, row_pos, cos_tab, sin_tab);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int swiglu_launch(const __half* gate, const __half* up, int ld_in, int n_rows, int I, __half* out, int ld_out,
                  cudaStream_t s) {
    if (I % 8 || ld_in % 8 || ld_out % 8) return fail(kErrInvalid, "swiglu: sizes must be multiples of 8");
    if (n_rows <= 0) return kOk;
    swiglu_kernel<<<grid_for(static_cast<long long>(n_rows) * (I / 8), 256), 256, 0, s>>>(gate, up, ld_in, n_rows, I,
                                                                                         out, ld_out);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int lse_reduce_launch(const float2* partial, int n_rows, int n_blocks, float* lse, cudaStream_t s) {
    if (n_rows <= 0) return kOk;
    lse_reduce_kernel<<<(n_rows + 7) / 8, 256, 0, s>>>(partial, n_rows, n_blocks, lse);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int flag_signal_launch(uint32_t* flag, uint32_t value, cudaStream_t s) {
    flag_signal_kernel<<<1, 1, 0, s>>>(flag, value);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int flag_wait_launch(const uint32_t* flag, uint32_t target, uint32_t timeout_ms, uint32_t* timed_out, cudaStream_t s) {
    flag_wait_kernel<<<1, 1, 0, s>>>(flag, target, static_cast<unsigned long long>(timeout_ms) * 1000000ull, timed_out);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

int gather_logprob_launch(const __half* x, int ldx, const __half* W, int ldw, int H, const int32_t* ids, int n_ids,
                          const float* lse, int n_rows, int n_vocab, float* out, cudaStream_t s) {
    if (H % 8 || ldx % 8 || ldw % 8) return fail(kErrInvalid, "gather_logprob: sizes must be multiples of 8");
    const long long warps = static_cast<long long>(n_rows) * n_ids;
    if (warps <= 0) return kOk;
    gather_logprob_kernel<<<static_cast<int>((warps + 7) / 8), 256, 0, s>>>(x, ldx, W, ldw, H, ids, n_ids, lse, n_rows,
                                                                            n_vocab, out);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

}  // namespace fls
