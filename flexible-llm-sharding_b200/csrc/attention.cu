// Shared-prefix prefill attention.
//
// Replaces, per decoder layer, the two eager-attention calls the reference makes through
// transformers (reference utils.py:272 prefix pass, utils.py:275-279 suffix pass): there the prefix
// K/V are `expand`ed across the S options, concatenated with each option's own K/V, repeated
// n_rep times for GQA, and the [B,heads,q,kv] score matrix is materialised in HBM.  Here one kernel
// reads K/V in place from the packed qkv buffer: every query tile names up to two key/value row
// segments (segment 0 = the prompt's prefix rows, shared by all options; segment 1 = the option's
// own rows), GQA is a head-index mapping, and scores never leave the SM (online softmax).
//
//   prefix tile   seg0 = prefix rows; mode 0 reproduces the reference (attention_mask=None =>
//                 bidirectional prefix), mode 1 gives a causal prefix.
//   suffix tile   mode 2: queries are the rows of one or more whole options (<= 128 rows); seg0 =
//                 prefix rows (all visible), seg1 = those options' rows; a query sees the keys of its
//                 own option at positions <= its own == the reference's mask slice
//                 attention_mask[:, :, -len_s:, -len_p-len_s:] applied per option.
// This file is the mma.sync (HMMA) implementation: head_dim 64, and the comparison kernel for the
// tcgen05 version in attention_sm100.cu.
//
// Arithmetic: S = Q K^T in fp32 (tensor cores, fp16 inputs), scaled by d^-1/2, softmax in fp32,
// P rounded to fp16 for the PV product (as HF's softmax(...).to(fp16) @ V), fp32 accumulation,
// output rounded to fp16.  v1 uses mma.sync m16n8k16 (HMMA); the tcgen05 version is future work.
#include "common.cuh"
#include "ptx.cuh"

namespace fls {

namespace {

constexpr int BQ = 64;   // queries per CTA (16 per warp)
constexpr int BK = 64;   // keys per pipeline stage
constexpr int ATT_THREADS = 128;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;  // src-size 0 => zero fill
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// [rows][D] fp16 tile, 16-byte chunks XOR-swizzled by the row so ldmatrix is conflict-free.
template <int D>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
    return static_cast<uint32_t>(row * (D * 2) + ((chunk ^ (row & 7)) << 4));
}

template <int D>
__global__ void __launch_bounds__(ATT_THREADS)
attention_kernel(const __half* __restrict__ qkv, int ld, int n_q_heads, int n_kv_heads,
                 const AttnTile* __restrict__ tiles, __half* __restrict__ out, int ld_out, float scale_log2e) {
    constexpr int CH = D / 8;            // 16-byte chunks per row
    constexpr int TILE_BYTES = BK * D * 2;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sQ = ptx::smem_u32(smem);
    const uint32_t sK = sQ + BQ * D * 2;            // 2 stages
    const uint32_t sV = sK + 2 * TILE_BYTES;        // 2 stages

    AttnTile t = tiles[blockIdx.x];
    // a tile holds up to 128 queries; this kernel takes them 64 at a time (blockIdx.z)
    const int q_off = blockIdx.z * BQ;
    if (q_off >= t.n_q) return;
    t.q_row += q_off;
    t.n_q = min(BQ, t.n_q - q_off);
    const int head = blockIdx.y;
    const int kvh = head / (n_q_heads / n_kv_heads);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tq = lane & 3;

    const __half* q_base = qkv + static_cast<size_t>(head) * D;
    const __half* k_base = qkv + static_cast<size_t>(n_q_heads + kvh) * D;
    const __half* v_base = qkv + static_cast<size_t>(n_q_heads + n_kv_heads + kvh) * D;

    // keys any query of this block can see (concatenated index: segment 0, then segment 1)
    const int Lk = t.mode == kAttnPrefixCausal ? min(t.seg0_len, t.q_seq0 + q_off + t.n_q)
                   : t.mode == kAttnSuffix     ? t.seg0_len + t.seg1_len
                                               : t.seg0_len;
    const int n_kv_tiles = (Lk + BK - 1) / BK;

    auto kv_row = [&](int key) { return key < t.seg0_len ? t.seg0_row + key : t.seg1_row + (key - t.seg0_len); };
    auto load_kv = [&](int tile_idx, int stage) {
#pragma unroll
        for (int i = 0; i < (BK * CH) / ATT_THREADS; ++i) {
            const int idx = threadIdx.x + i * ATT_THREADS;
            const int r = idx / CH, c = idx % CH;
            const int key = tile_idx * BK + r;
            const bool ok = key < Lk;
            const size_t row = ok ? static_cast<size_t>(kv_row(key)) : 0;
            cp_async16(sK + stage * TILE_BYTES + swz<D>(r, c), k_base + row * ld + c * 8, ok);
            cp_async16(sV + stage * TILE_BYTES + swz<D>(r, c), v_base + row * ld + c * 8, ok);
        }
    };

    // ---- prologue: Q tile + first K/V tile
#pragma unroll
    for (int i = 0; i < (BQ * CH) / ATT_THREADS; ++i) {
        const int idx = threadIdx.x + i * ATT_THREADS;
        const int r = idx / CH, c = idx % CH;
        const bool ok = r < t.n_q;
        const size_t row = ok ? static_cast<size_t>(t.q_row + r) : 0;
        cp_async16(sQ + swz<D>(r, c), q_base + row * ld + c * 8, ok);
    }
    if (n_kv_tiles > 0) load_kv(0, 0);
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();

    // Q fragments for this warp's 16 rows
    uint32_t qf[D / 16][4];
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
        const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int c = ks * 2 + (lane >> 4);
        ldmatrix_x4(sQ + swz<D>(r, c), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
    }

    float o[D / 8][4];
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};
    const int qi0 = warp * 16 + g;  // this thread's rows inside the block: qi0 and qi0 + 8
    // visible key range of each of the two rows: [0, lim0) in segment 0, [lo1, hi1] in segment 1
    int lim0[2], lo1[2], hi1[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int qi = q_off + qi0 + r * 8;  // query index inside the whole tile
        lim0[r] = t.mode == kAttnPrefixCausal ? min(t.seg0_len, t.q_seq0 + qi + 1) : t.seg0_len;
        lo1[r] = 0, hi1[r] = -1;
        if (t.mode == kAttnSuffix) {
            const int qa = (t.q_row - q_off - t.seg1_row) + qi, L = t.q_seq0;
            lo1[r] = (qa / L) * L;
            hi1[r] = min(qa, t.seg1_len - 1);
        }
    }

    for (int kt = 0; kt < n_kv_tiles; ++kt) {
        const int stage = kt & 1;
        if (kt + 1 < n_kv_tiles) load_kv(kt + 1, stage ^ 1);
        cp_async_commit();

        const uint32_t kS = sK + stage * TILE_BYTES, vS = sV + stage * TILE_BYTES;
        // ---- S = Q K^T  (16 x 64 per warp)
        float s[BK / 8][4];
#pragma unroll
        for (int nt = 0; nt < BK / 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
        for (int nt = 0; nt < BK / 8; ++nt) {
#pragma unroll
            for (int kp = 0; kp < D / 32; ++kp) {  // two k16 steps per ldmatrix.x4
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4(kS + swz<D>(nt * 8 + (lane & 7), kp * 4 + (lane >> 3)), b0, b1, b2, b3);
                mma_16816(s[nt], qf[kp * 2], b0, b1);
                mma_16816(s[nt], qf[kp * 2 + 1], b2, b3);
            }
        }
        // ---- mask + online softmax (rows qi0, qi0+8; this thread owns cols nt*8 + 2*tq + {0,1})
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < BK / 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int key = kt * BK + nt * 8 + 2 * tq + (e & 1);
                const int r = e >> 1, k1 = key - t.seg0_len;
                const bool vis = key < t.seg0_len ? key < lim0[r] : (k1 >= lo1[r] && k1 <= hi1[r]);
                s[nt][e] = vis ? s[nt][e] * scale_log2e : -INFINITY;
                mx[e >> 1] = fmaxf(mx[e >> 1], s[nt][e]);
            }
        }
        float corr[2], m_use[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
            const float m_new = fmaxf(m_run[r], mx[r]);
            m_use[r] = m_new == -INFINITY ? 0.f : m_new;  // fully masked so far: avoid (-inf) - (-inf)
            corr[r] = exp2f(m_run[r] - m_use[r]);
            m_run[r] = m_new;
        }
        float rs[2] = {0.f, 0.f};
        uint32_t pf[BK / 16][4];
#pragma unroll
        for (int nt = 0; nt < BK / 8; ++nt) {
            const float p0 = exp2f(s[nt][0] - m_use[0]), p1 = exp2f(s[nt][1] - m_use[0]);
            const float p2 = exp2f(s[nt][2] - m_use[1]), p3 = exp2f(s[nt][3] - m_use[1]);
            rs[0] += p0 + p1;
            rs[1] += p2 + p3;
            pf[nt >> 1][(nt & 1) * 2 + 0] = pack2(p0, p1);
            pf[nt >> 1][(nt & 1) * 2 + 1] = pack2(p2, p3);
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) l_run[r] = l_run[r] * corr[r] + rs[r];
#pragma unroll
        for (int i = 0; i < D / 8; ++i) {
            o[i][0] *= corr[0], o[i][1] *= corr[0];
            o[i][2] *= corr[1], o[i][3] *= corr[1];
        }
        // ---- O += P V
#pragma unroll
        for (int ks = 0; ks < BK / 16; ++ks) {
#pragma unroll
            for (int np = 0; np < D / 16; ++np) {  // two d-tiles of 8 per ldmatrix.x4.trans
                uint32_t b0, b1, b2, b3;
                const int r = ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
                const int c = np * 2 + (lane >> 4);
                ldmatrix_x4_trans(vS + swz<D>(r, c), b0, b1, b2, b3);
                mma_16816(o[np * 2], pf[ks], b0, b1);
                mma_16816(o[np * 2 + 1], pf[ks], b2, b3);
            }
        }
        cp_async_wait<0>();
        __syncthreads();
    }

    // ---- normalise, stage through this warp's Q rows, store 16 bytes per thread
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
    const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f;
    const float inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
        const uint32_t lo = pack2(o[i][0] * inv0, o[i][1] * inv0);
        const uint32_t hi = pack2(o[i][2] * inv1, o[i][3] * inv1);
        const int r0 = warp * 16 + g;
        asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(sQ + swz<D>(r0, i) + tq * 4), "r"(lo) : "memory");
        asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(sQ + swz<D>(r0 + 8, i) + tq * 4), "r"(hi) : "memory");
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < (16 * CH) / 32; ++i) {
        const int idx = lane + i * 32;
        const int r = warp * 16 + idx / CH, c = idx % CH;
        if (r < t.n_q) {
            uint4 v;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                         : "r"(sQ + swz<D>(r, c)));
            *reinterpret_cast<uint4*>(out + static_cast<size_t>(t.q_row + r) * ld_out + head * D + c * 8) = v;
        }
    }
}

template <int D>
constexpr int att_smem_bytes() {
    return BQ * D * 2 + 4 * BK * D * 2;
}

}  // namespace

int attention_init() {
    FLS_CUDA(cudaFuncSetAttribute(attention_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  att_smem_bytes<128>()));
    FLS_CUDA(cudaFuncSetAttribute(attention_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  att_smem_bytes<64>()));
    return kOk;
}

int attention_launch(const __half* qkv, int ld, int n_q_heads, int n_kv_heads, int head_dim, const AttnTile* tiles,
                     int n_tiles, __half* out, int ld_out, cudaStream_t s) {
    if (n_tiles <= 0) return kOk;
    if (ld % 8 || ld_out % 8) return fail(kErrInvalid, "attention: row strides must be multiples of 8");
    if (n_kv_heads <= 0 || n_q_heads % n_kv_heads) return fail(kErrInvalid, "attention: heads %d / kv heads %d", n_q_heads, n_kv_heads);
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(head_dim));
    const dim3 grid(n_tiles, n_q_heads, 2);
    if (head_dim == 128)
        attention_kernel<128><<<grid, ATT_THREADS, att_smem_bytes<128>(), s>>>(qkv, ld, n_q_heads, n_kv_heads, tiles,
                                                                              out, ld_out, scale_log2e);
    else if (head_dim == 64)
        attention_kernel<64><<<grid, ATT_THREADS, att_smem_bytes<64>(), s>>>(qkv, ld, n_q_heads, n_kv_heads, tiles, out,
                                                                            ld_out, scale_log2e);
    else
        return fail(kErrInvalid, "attention: head_dim %d not supported (64 or 128)", head_dim);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

}  // namespace fls
