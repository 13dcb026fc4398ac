// Pieces shared by the tcgen05 GEMM kernels (gemm_sm100.cu: 256 x 256 tiles with two accumulator stages;
// gemm_wide_sm100.cu: 256 x 512 tiles): tile constants, kernel parameters, raster order and the staged epilogue
// (registers -> swizzled 32 x 64 sub-tile in shared memory -> TMA store).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace fls {
namespace gemm {

constexpr int BLOCK_M = 128;  // rows per CTA (a CTA pair covers 256)
constexpr int BLOCK_N = 256;  // MMA N (TMEM columns per accumulator stage)
constexpr int BLOCK_K = 64;   // 64 fp16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int ACC_STAGES = 2;
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KiB
constexpr int SMEM_BARRIER_BYTES = 512;
constexpr int EPI_CHUNK_COLS = 64;                      // one swizzle row of fp16
constexpr int EPI_BUF_BYTES = 32 * EPI_CHUNK_COLS * 2;  // 32 rows x 128 B per epilogue warp
constexpr int EPI_STAGE_BYTES = 4 * 2 * EPI_BUF_BYTES;  // 4 warps x 2 buffers = 32 KiB
constexpr int MAX_STAGES = 8;

// CG = tcgen05 cta_group.  CG 1: one CTA per 128x256 tile, B tile 256 rows, 4 stages of 48 KiB.
// CG 2: a CTA pair (cluster of 2) per 256x256 tile; each CTA stages its own 128 A rows and HALF of
// the B rows (the pair's MMA reads both halves), 6 stages of 32 KiB: half the shared-memory operand
// traffic per SM and a deeper TMA look-ahead.
template <int CG>
struct Cfg {
    static constexpr int STAGES = CG == 1 ? 4 : 6;
    static constexpr int B_ROWS = BLOCK_N / CG;
    static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_STAGE_BYTES + SMEM_BARRIER_BYTES + 1024;  // + alignment slack
};
constexpr int NUM_THREADS = 256;
constexpr int EPI_WARP0 = 4;
constexpr int GROUP_M = 16;  // default m-blocks per raster group: A bands of a group stay in L2 while n sweeps

struct KParams {
    int M, N, K;
    __half* C;
    int ldc;
    const __half* R;
    int ldr;
    float2* partial;
    const float* lse;
    int up_off;
    int group_m;
    int cache_hints;
    int raster;
    int num_m_blocks, num_n_blocks;
    const int32_t* row_pos;
    const __half* cos_tab;
    const __half* sin_tab;
    int rope_cols;
};

template <int EPI>
__device__ __host__ __forceinline__ constexpr int out_cols_per_tile() {
    return EPI == kEpiSwiGLU ? BLOCK_N / 2 : BLOCK_N;
}

// Raster order of the persistent tile loop.  A "wave" = the tiles the resident CTA pairs work on at the same time =
// consecutive tile indices, so the order decides which operand bands are hot in L2 together.
//   raster 0: groups of `group` m-blocks; inside a group m runs fastest, then n (the group's A bands stay hot)
//   raster 1: groups of `group` n-blocks; inside a group n runs fastest, then m (the group's B bands stay hot)
//   raster 2 / 3: as 0 / 1 with the slow direction reversed in every other group (the last bands of one group are
//             the first of the next)
__device__ __forceinline__ void tile_coords(int tile, int num_m_blocks, int num_n_blocks, int group, int raster,
                                            int& m_blk, int& n_blk) {
    const bool by_n = raster & 1;
    const int fast_total = by_n ? num_n_blocks : num_m_blocks;   // dimension cut into groups
    const int slow_total = by_n ? num_m_blocks : num_n_blocks;
    const int per_group = group * slow_total;
    const int g = tile / per_group;
    const int first = g * group;
    const int gsize = min(fast_total - first, group);
    const int local = tile - g * per_group;
    const int f = first + local % gsize;
    int sl = local / gsize;
    if ((raster & 2) && (g & 1)) sl = slow_total - 1 - sl;
    m_blk = by_n ? sl : f;
    n_blk = by_n ? f : sl;
}

// silu(x) = x / (1 + e^-x) on the fp32 value of an fp16 gate, for the fused SwiGLU epilogues.  Hardware exp2 and
// reciprocal approximations (2 MUFU operations, ~2 ulp of fp32): the fp16-rounded result equals the correctly rounded
// one except where the fp32 value lies within a few 2^-24 of an fp16 rounding boundary (about one element in a
// thousand, then off by one fp16 ulp) - the accurate expf + IEEE division cost ~100 issue slots per element and, in the
// 256 x 512 kernel whose epilogue is not hidden behind the next tile's MMAs, a fifth of the tensor pipe's time
// (profiles/r02_summary.md).  The stand-alone swiglu_kernel keeps the exact form.
__device__ __forceinline__ float silu_f32(float x) { return __fdividef(x, 1.0f + __expf(-x)); }

__device__ __forceinline__ uint32_t pack_h2(__half a, __half b) {
    __half2 h = __halves2half2(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// 16-byte piece j (0..7) of row r of a 32 x 64 fp16 sub-tile laid out with the 128-byte swizzle TMA expects
__device__ __forceinline__ uint4* stage_ptr(uint8_t* buf, int r, int j) {
    return reinterpret_cast<uint4*>(buf + r * 128 + ((j ^ (r & 7)) << 4));
}


// This warp's first residual chunk of an accumulator: fetched into staging buffer q & 1 (may run while the main loop
// of the tile is still going).
__device__ __forceinline__ void epi_prefetch_residual(const CUtensorMap* tmap_r, uint8_t* buf_base, uint64_t* rbar,
                                                      uint32_t q, int lane, int col0, int row0) {
    if (lane == 0) {
        ptx::bulk_wait_group_read<0>();  // whatever last read this buffer has done so
        ptx::mbar_arrive_expect_tx(&rbar[q & 1], EPI_BUF_BYTES);
        ptx::tma_load_2d(buf_base + (q & 1) * EPI_BUF_BYTES, tmap_r, &rbar[q & 1], col0, row0);
    }
}

// Staged epilogue of ONE accumulator (128 lanes x 256 fp32 columns of TMEM) for one epilogue warp: the warp's 32 rows
// start at row0, the accumulator's first output column is col0, `taddr` = TMEM address of the warp's lanes at the
// accumulator's first column.  buf_base = the warp's two 32 x 64 staging buffers, rbar = its two residual barriers,
// q = 64-column chunks the warp has produced so far (buffer = q & 1, barrier phase = (q >> 1) & 1).  For the
// residual epilogue the first chunk must already have been requested with epi_prefetch_residual.
template <int EPI>
__device__ __forceinline__ void epi_accumulator(const KParams& p, const CUtensorMap* tmap_c_ptr,
                                                const CUtensorMap* tmap_r_ptr, uint32_t taddr, int row0, int col0,
                                                uint8_t* buf_base, uint64_t* rbar, uint32_t& q, int lane) {
    constexpr int OUT_COLS = out_cols_per_tile<EPI>();
    const CUtensorMap& tmap_c = *tmap_c_ptr;
    const CUtensorMap& tmap_r = *tmap_r_ptr;
    auto bufp = [&](uint32_t i) { return buf_base + (i & 1) * EPI_BUF_BYTES; };
    const int row = row0 + lane;
    const bool row_ok = row < p.M;
    if constexpr (EPI == kEpiStoreRope) {
        // Two 128-wide heads per tile.  A thread holds dims i and i+64 of its row, so a head fills
        // both staging buffers at once: buffer 0 <- dims [0,64), buffer 1 <- dims [64,128).
        const int pos = row_ok ? p.row_pos[row] : 0;
#pragma unroll 1
        for (int h = 0; h < BLOCK_N / 128; ++h) {
            const int hcol = col0 + h * 128;
            const bool rotate = hcol < p.rope_cols;
            if (lane == 0) ptx::bulk_wait_group_read<0>();
            __syncwarp();
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t a[32], b[32];  // dims [c*32, +32) and their partners [64 + c*32, +32)
                ptx::tmem_ld_32x32(taddr + h * 128 + c * 32, a);
                ptx::tmem_ld_32x32(taddr + h * 128 + 64 + c * 32, b);
                ptx::tmem_ld_wait();
                const __half* ct = p.cos_tab + static_cast<size_t>(pos) * 128 + c * 32;
                const __half* st = p.sin_tab + static_cast<size_t>(pos) * 128 + c * 32;
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    uint4 c1 = make_uint4(0, 0, 0, 0), s1 = c1, c2 = c1, s2 = c1;
                    if (rotate) {
                        c1 = *reinterpret_cast<const uint4*>(ct + v * 8);
                        s1 = *reinterpret_cast<const uint4*>(st + v * 8);
                        c2 = *reinterpret_cast<const uint4*>(ct + 64 + v * 8);
                        s2 = *reinterpret_cast<const uint4*>(st + 64 + v * 8);
                    }
                    const __half2* hc1 = reinterpret_cast<const __half2*>(&c1);
                    const __half2* hs1 = reinterpret_cast<const __half2*>(&s1);
                    const __half2* hc2 = reinterpret_cast<const __half2*>(&c2);
                    const __half2* hs2 = reinterpret_cast<const __half2*>(&s2);
                    uint32_t o1[4], o2[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int i = v * 8 + j * 2;
                        __half2 x1 = __floats2half2_rn(__uint_as_float(a[i]), __uint_as_float(a[i + 1]));
                        __half2 x2 = __floats2half2_rn(__uint_as_float(b[i]), __uint_as_float(b[i + 1]));
                        if (rotate) {
                            const __half2 r1 = __hadd2_rn(__hmul2_rn(x1, hc1[j]), __hmul2_rn(__hneg2(x2), hs1[j]));
                            const __half2 r2 = __hadd2_rn(__hmul2_rn(x2, hc2[j]), __hmul2_rn(x1, hs2[j]));
                            x1 = r1, x2 = r2;
                        }
                        o1[j] = *reinterpret_cast<const uint32_t*>(&x1);
                        o2[j] = *reinterpret_cast<const uint32_t*>(&x2);
                    }
                    *stage_ptr(bufp(0), lane, c * 4 + v) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
                    *stage_ptr(bufp(1), lane, c * 4 + v) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
                }
            }
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && hcol < p.N) {
                ptx::tma_store_2d(&tmap_c, bufp(0), hcol, row0);
                ptx::tma_store_2d(&tmap_c, bufp(1), hcol + 64, row0);
                ptx::bulk_commit_group();
            }
        }
    } else {
        float lse = 0.f;
        if constexpr (EPI == kEpiProbs) lse = row_ok ? p.lse[row] : 0.f;
        constexpr int N_CHUNKS = OUT_COLS / EPI_CHUNK_COLS;
#pragma unroll 1
        for (int ch = 0; ch < N_CHUNKS; ++ch, ++q) {
            uint8_t* buf = bufp(q & 1);
            uint32_t r0[32], r1[32];  // SwiGLU: gate / up of 32 columns; else 2 x 32 columns
#pragma unroll 1
            for (int half = 0; half < (EPI == kEpiSwiGLU ? 2 : 1); ++half) {
                if constexpr (EPI == kEpiSwiGLU) {
                    ptx::tmem_ld_32x32(taddr + ch * 64 + half * 32, r0);
                    ptx::tmem_ld_32x32(taddr + OUT_COLS + ch * 64 + half * 32, r1);
                } else {
                    ptx::tmem_ld_32x32(taddr + ch * 64, r0);
                    ptx::tmem_ld_32x32(taddr + ch * 64 + 32, r1);
                }
                if (half == 0) {
                    if constexpr (EPI == kEpiResidual) {
                        if (lane == 0 && ch + 1 < N_CHUNKS) {  // next chunk's residual into the other buffer
                            ptx::bulk_wait_group_read<0>();
                            ptx::mbar_arrive_expect_tx(&rbar[(q + 1) & 1], EPI_BUF_BYTES);
                            ptx::tma_load_2d(bufp((q + 1) & 1), &tmap_r, &rbar[(q + 1) & 1],
                                             col0 + (ch + 1) * EPI_CHUNK_COLS, row0);
                        }
                    } else {
                        if (lane == 0) ptx::bulk_wait_group_read<1>();  // the store two chunks back has read `buf`
                    }
                    __syncwarp();
                }
                ptx::tmem_ld_wait();
                if constexpr (EPI == kEpiResidual) ptx::mbar_wait(&rbar[q & 1], (q >> 1) & 1);
                if constexpr (EPI == kEpiSwiGLU) {
                    // HF LlamaMLP: act_fn(gate_proj(x)) * up_proj(x), each op rounded to fp16.
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        uint32_t o[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            __half r2[2];
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int i = v * 8 + j * 2 + e;
                                const float gf = __half2float(__float2half_rn(__uint_as_float(r0[i])));
                                const __half sg = __float2half_rn(silu_f32(gf));
                                const __half uh = __float2half_rn(__uint_as_float(r1[i]));
                                r2[e] = __hmul(sg, uh);
                            }
                            o[j] = pack_h2(r2[0], r2[1]);
                        }
                        *stage_ptr(buf, lane, half * 4 + v) = make_uint4(o[0], o[1], o[2], o[3]);
                    }
                } else {
#pragma unroll
                    for (int part = 0; part < 2; ++part) {
                        const uint32_t* r = part == 0 ? r0 : r1;
#pragma unroll
                        for (int v = 0; v < 4; ++v) {
                            uint4* sp = stage_ptr(buf, lane, part * 4 + v);
                            uint4 res = make_uint4(0, 0, 0, 0);
                            if constexpr (EPI == kEpiResidual) res = *sp;
                            const uint32_t rr[4] = {res.x, res.y, res.z, res.w};
                            uint32_t o[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                __half h0 = __float2half_rn(__uint_as_float(r[v * 8 + j * 2]));
                                __half h1 = __float2half_rn(__uint_as_float(r[v * 8 + j * 2 + 1]));
                                if constexpr (EPI == kEpiResidual) {
                                    // hidden = residual + linear(x): fp16 add of two fp16 values
                                    const __half2 rh = *reinterpret_cast<const __half2*>(&rr[j]);
                                    h0 = __hadd(__low2half(rh), h0);
                                    h1 = __hadd(__high2half(rh), h1);
                                } else if constexpr (EPI == kEpiProbs) {
                                    // softmax evaluated on the fp16 logits (reference utils.py:289)
                                    h0 = __float2half_rn(__expf(__half2float(h0) - lse));
                                    h1 = __float2half_rn(__expf(__half2float(h1) - lse));
                                }
                                o[j] = pack_h2(h0, h1);
                            }
                            *sp = make_uint4(o[0], o[1], o[2], o[3]);
                        }
                    }
                }
            }
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && col0 + ch * EPI_CHUNK_COLS < p.N) {
                ptx::tma_store_2d(&tmap_c, buf, col0 + ch * EPI_CHUNK_COLS, row0);
                ptx::bulk_commit_group();
            }
        }
    }
}

}  // namespace gemm
}  // namespace fls
