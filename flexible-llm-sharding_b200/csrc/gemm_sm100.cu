// tcgen05 / TMEM / TMA GEMM for sm_100a with fused epilogues.
//
//   C[M,N] = A[M,K] * B[N,K]^T      A = packed token rows (fp16, K contiguous)
//                                   B = nn.Linear weight [out,in] exactly as stored in the layer
//                                       file (fp16, K contiguous) -> K-major operand, no transpose
//
// Replaces every `nn.Linear` call the reference issues through transformers (q/k/v/o_proj,
// gate/up/down_proj, lm_head; reference utils.py:272,278,288), which there are cuBLAS HGEMMs
// with M <= a few hundred rows, one prompt at a time.  Here M = all packed rows of a micro-batch.
//
// Structure: persistent CTAs (one per SM), 256 threads, warp-specialised:
//   warp 0      TMA producer (128B swizzle).  Default cta_group::2: a cluster of two CTAs shares one 256x256
//               tile, each CTA stages its own 128 A rows and half of the B rows per stage, 6 stages of 32 KiB;
//               cta_group::1 (option / test comparator): A 128x64 + B 256x64 per stage, 4 stages of 48 KiB
//   warp 1      MMA issuer: one elected thread (of the leader CTA with cta_group::2), tcgen05.mma kind::f16,
//               M=128 per CTA, N=256, K=16, fp32 accumulators in TMEM, two accumulator stages
//               (2 x 256 columns = all 512); tcgen05.commit (multicast to the peer) frees the smem stage
//   warp 2      TMEM allocator
//   warps 4-7   epilogue: tcgen05.ld (32 lanes x 32 columns per warp) -> registers -> fused op -> a 32 x 64
//               sub-tile in shared memory (128-byte swizzle) -> TMA store (full 128-byte lines; rows / columns
//               beyond M / N are clipped by the tensor map).  The residual operand comes in the same way (TMA load
//               into the staging sub-tile, prefetched a chunk ahead).  Overlaps the next tile's main loop through
//               the second accumulator stage.  TS = false keeps the older epilogue (every thread stores 16-byte
//               pieces of its own row straight to global) as option "gemm_tma_store" = 0 / test comparator.
// Rounding in the epilogues follows the reference's eager fp16 order (see each epilogue).
#include <cuda.h>

#include <algorithm>
#include <functional>
#include <mutex>
#include <unordered_map>

#include "gemm_tile.cuh"

namespace fls {

using namespace gemm;

namespace {

template <int EPI, int CG, bool TS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_r,
                    const KParams p) {
    constexpr int STAGES = Cfg<CG>::STAGES;
    constexpr int B_BYTES = Cfg<CG>::B_BYTES;
    constexpr int STAGE_BYTES = Cfg<CG>::STAGE_BYTES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* epi_stage = smem + STAGES * STAGE_BYTES;  // 1024-byte aligned: stages are multiples of 16 KiB
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + EPI_STAGE_BYTES);
    uint64_t* empty_bar = full_bar + MAX_STAGES;
    uint64_t* tmem_full_bar = empty_bar + MAX_STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + ACC_STAGES;
    uint64_t* res_bar = tmem_empty_bar + ACC_STAGES;  // [4 warps][2 buffers]: residual sub-tile has landed
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(res_bar + 8);

    const int warp_idx = threadIdx.x >> 5;  // warp-uniform
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = CG == 2 ? ptx::cluster_ctarank() : 0;  // 0 = leader of the pair
    const bool is_leader = cta_rank == 0;

    if (warp_idx == 0 && ptx::elect_one()) {
        ptx::prefetch_tensormap(&tmap_a);
        ptx::prefetch_tensormap(&tmap_b);
        if constexpr (TS && EPI != kEpiLsePartial) ptx::prefetch_tensormap(&tmap_c);
        if constexpr (TS && EPI == kEpiResidual) ptx::prefetch_tensormap(&tmap_r);
    }
    if (warp_idx == 1 && ptx::elect_one()) {
        for (int i = 0; i < STAGES; ++i) {
            ptx::mbar_init(&full_bar[i], 1);   // the (leader's) producer arrive.expect_tx
            ptx::mbar_init(&empty_bar[i], 1);  // one tcgen05.commit
        }
        for (int i = 0; i < ACC_STAGES; ++i) {
            ptx::mbar_init(&tmem_full_bar[i], 1);
            ptx::mbar_init(&tmem_empty_bar[i], 4 * CG);  // one arrive per epilogue warp of every CTA of the pair
        }
        for (int i = 0; i < 8; ++i) ptx::mbar_init(&res_bar[i], 1);
        ptx::fence_mbar_init();
    }
    if (warp_idx == 2) {
        ptx::tmem_alloc<CG>(tmem_base_slot, 512);
        ptx::tmem_relinquish<CG>();
    }
    ptx::tc_fence_before_sync();
    if constexpr (CG == 2) ptx::cluster_sync_all(); else __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_base_slot;

    const int num_tiles = p.num_m_blocks * p.num_n_blocks;  // m blocks of 128*CG rows
    const int k_blocks = (p.K + BLOCK_K - 1) / BLOCK_K;
    constexpr int OUT_COLS = out_cols_per_tile<EPI>();
    const int worker = blockIdx.x / CG, n_workers = gridDim.x / CG;

    if (warp_idx == 0) {
        // ------------------------------------------------------------------ TMA producer (every CTA)
        if (ptx::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                int m_blk, n_blk;
                tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
                const int a_row = (m_blk * CG + static_cast<int>(cta_rank)) * BLOCK_M;
                int b_row;
                if constexpr (EPI == kEpiSwiGLU)  // CG 2: leader stages the gate rows, its peer the up rows
                    b_row = n_blk * OUT_COLS + (CG == 2 ? static_cast<int>(cta_rank) * p.up_off : 0);
                else
                    b_row = n_blk * BLOCK_N + static_cast<int>(cta_rank) * Cfg<CG>::B_ROWS;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    if constexpr (CG == 1) {
                        ptx::mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
                        ptx::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * BLOCK_K, a_row);
                        if constexpr (EPI == kEpiSwiGLU) {
                            // gate rows -> MMA columns [0,128), up rows -> MMA columns [128,256)
                            ptx::tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * BLOCK_K, b_row);
                            ptx::tma_load_2d(sb + B_BYTES / 2, &tmap_b, &full_bar[stage], kb * BLOCK_K,
                                             p.up_off + b_row);
                        } else {
                            ptx::tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * BLOCK_K, b_row);
                        }
                    } else {
                        // both CTAs' bytes are credited to the leader's barrier (peer bit cleared)
                        if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
                        const uint32_t bar = ptx::smem_u32(&full_bar[stage]) & 0xFEFFFFFFu;
                        if (p.cache_hints == 0) {
                            ptx::tma_load_2d_cta2(sa, &tmap_a, bar, kb * BLOCK_K, a_row);
                            ptx::tma_load_2d_cta2(sb, &tmap_b, bar, kb * BLOCK_K, b_row);
                        } else {  // 1: stream A, keep B in L2;  2: keep A, stream B;  3: A normal, keep B;  4: keep A, B normal
                            const uint64_t ha = p.cache_hints == 1 ? ptx::kL2EvictFirst
                                                : p.cache_hints == 3 ? ptx::kL2EvictNormal : ptx::kL2EvictLast;
                            const uint64_t hb = p.cache_hints == 2 ? ptx::kL2EvictFirst
                                                : p.cache_hints == 4 ? ptx::kL2EvictNormal : ptx::kL2EvictLast;
                            ptx::tma_load_2d_cta2_hint(sa, &tmap_a, bar, kb * BLOCK_K, a_row, ha);
                            ptx::tma_load_2d_cta2_hint(sb, &tmap_b, bar, kb * BLOCK_K, b_row, hb);
                        }
                    }
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA only)
        if (is_leader && ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::umma_idesc_f16_f32(BLOCK_M * CG, BLOCK_N);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                ptx::mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
                ptx::tc_fence_after_sync();
                const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after_sync();
                    const uint32_t sa = ptx::smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t da = ptx::umma_smem_desc_sw128(sa);
                    const uint64_t db = ptx::umma_smem_desc_sw128(sa + A_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        // advance 16 fp16 = 32 B inside the 128 B swizzle row: +2 in the (addr >> 4) field
                        ptx::umma_f16<CG>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    }
                    // frees the smem stage (in both CTAs of a pair) once these MMAs retire
                    if constexpr (CG == 1) ptx::umma_commit(&empty_bar[stage]);
                    else ptx::umma_commit_cta2(&empty_bar[stage], 0b11);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                // accumulator complete -> epilogue warps (of both CTAs)
                if constexpr (CG == 1) ptx::umma_commit(&tmem_full_bar[acc]);
                else ptx::umma_commit_cta2(&tmem_full_bar[acc], 0b11);
                if (++acc == ACC_STAGES) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else if (warp_idx >= EPI_WARP0) {
        // ------------------------------------------------------------------ epilogue
        const int ew = warp_idx - EPI_WARP0;  // == warp_idx % 4 -> TMEM lanes [32*ew, 32*ew+32)
        if constexpr (TS && EPI != kEpiLsePartial) {
            // ---- staged epilogue: registers -> swizzled 32 x 64 sub-tile in smem -> TMA store
            uint8_t* const buf_base = epi_stage + (ew * 2) * EPI_BUF_BYTES;  // this warp's two staging buffers
            uint64_t* rbar = res_bar + ew * 2;
            uint32_t q = 0;  // 64-column chunks this warp has produced so far: buffer = q & 1, barrier phase = (q >> 1) & 1
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                int m_blk, n_blk;
                tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
                const int row0 = (m_blk * CG + static_cast<int>(cta_rank)) * BLOCK_M + ew * 32;  // this warp's first row
                const int col0 = n_blk * OUT_COLS;
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BLOCK_N;
                // the first residual chunk of the tile is fetched while the main loop still runs
                if constexpr (EPI == kEpiResidual) epi_prefetch_residual(&tmap_r, buf_base, rbar, q, lane, col0, row0);
                ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
                ptx::tc_fence_after_sync();

                epi_accumulator<EPI>(p, &tmap_c, &tmap_r, taddr, row0, col0, buf_base, rbar, q, lane);
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (CG == 1) ptx::mbar_arrive(&tmem_empty_bar[acc]);
                    else ptx::mbar_arrive_cluster(&tmem_empty_bar[acc], 0);  // the leader's MMA thread waits on it
                }
                if (++acc == ACC_STAGES) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
            if (lane == 0) ptx::bulk_wait_group_read<0>();  // the staging buffers must outlive the last stores' reads
        } else {
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                int m_blk, n_blk;
                tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
                ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
                ptx::tc_fence_after_sync();
                const int row = (m_blk * CG + static_cast<int>(cta_rank)) * BLOCK_M + ew * 32 + lane;
                const bool row_ok = row < p.M;
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BLOCK_N;
                const int col0 = n_blk * OUT_COLS;

                if constexpr (EPI == kEpiSwiGLU) {
                    // HF LlamaMLP: act_fn(gate_proj(x)) * up_proj(x), each op rounded to fp16.
    #pragma unroll 1
                    for (int c = 0; c < OUT_COLS / 32; ++c) {
                        uint32_t g[32], u[32];
                        ptx::tmem_ld_32x32(taddr + c * 32, g);
                        ptx::tmem_ld_32x32(taddr + OUT_COLS + c * 32, u);
                        ptx::tmem_ld_wait();
                        if (row_ok) {
                            __half* dst = p.C + static_cast<size_t>(row) * p.ldc + col0 + c * 32;
    #pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                if (col0 + c * 32 + v * 8 < p.N) {
                                    uint32_t o[4];
    #pragma unroll
                                    for (int j = 0; j < 4; ++j) {
                                        __half r2[2];
    #pragma unroll
                                        for (int e = 0; e < 2; ++e) {
                                            const int i = v * 8 + j * 2 + e;
                                            const float gf = __half2float(__float2half_rn(__uint_as_float(g[i])));
                                            const __half s = __float2half_rn(silu_f32(gf));
                                            const __half uh = __float2half_rn(__uint_as_float(u[i]));
                                            r2[e] = __hmul(s, uh);
                                        }
                                        o[j] = pack_h2(r2[0], r2[1]);
                                    }
                                    *reinterpret_cast<uint4*>(dst + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
                                }
                            }
                        }
                    }
                } else if constexpr (EPI == kEpiStoreRope) {
                    // q/k heads (128 columns each, two per tile): HF apply_rotary_pos_emb on the fp16-rounded
                    // projection, half-split convention, every product and the sum rounded on their own
                    // (bit-identical to rope_kernel).  Columns >= rope_cols (the v heads) are stored as is.
                    const int pos = row_ok ? p.row_pos[row] : 0;
    #pragma unroll 1
                    for (int h = 0; h < BLOCK_N / 128; ++h) {
                        const int hcol = col0 + h * 128;
                        const bool rotate = hcol < p.rope_cols;
    #pragma unroll 1
                        for (int c = 0; c < 2; ++c) {
                            uint32_t a[32], b[32];  // dims [c*32, +32) and their partners [64 + c*32, +32)
                            ptx::tmem_ld_32x32(taddr + h * 128 + c * 32, a);
                            ptx::tmem_ld_32x32(taddr + h * 128 + 64 + c * 32, b);
                            ptx::tmem_ld_wait();
                            if (row_ok && hcol < p.N) {
                                __half* dst = p.C + static_cast<size_t>(row) * p.ldc + hcol + c * 32;
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
                                    *reinterpret_cast<uint4*>(dst + v * 8) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
                                    *reinterpret_cast<uint4*>(dst + 64 + v * 8) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
                                }
                            }
                        }
                    }
                } else if constexpr (EPI == kEpiLsePartial) {
                    // running (max, sum exp) over this tile's columns of the fp16-rounded logits
                    float m_run = -INFINITY, s_run = 0.f;
    #pragma unroll 1
                    for (int c = 0; c < BLOCK_N / 32; ++c) {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32(taddr + c * 32, r);
                        ptx::tmem_ld_wait();
                        float v[32];
                        float cmax = -INFINITY;
    #pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            v[i] = (col0 + c * 32 + i < p.N) ? __half2float(__float2half_rn(__uint_as_float(r[i])))
                                                             : -INFINITY;
                            cmax = fmaxf(cmax, v[i]);
                        }
                        if (cmax > -INFINITY) {
                            const float m_new = fmaxf(m_run, cmax);
                            float s = 0.f;
    #pragma unroll
                            for (int i = 0; i < 32; ++i) s += __expf(v[i] - m_new);
                            s_run = s_run * __expf(m_run - m_new) + s;
                            m_run = m_new;
                        }
                    }
                    if (row_ok) p.partial[static_cast<size_t>(row) * p.num_n_blocks + n_blk] = make_float2(m_run, s_run);
                } else {
                    float lse = 0.f;
                    if constexpr (EPI == kEpiProbs) lse = row_ok ? p.lse[row] : 0.f;
    #pragma unroll 1
                    for (int c = 0; c < BLOCK_N / 32; ++c) {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32(taddr + c * 32, r);
                        ptx::tmem_ld_wait();
                        if (row_ok) {
                            const size_t off = static_cast<size_t>(row) * p.ldc + col0 + c * 32;
    #pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                if (col0 + c * 32 + v * 8 < p.N) {
                                    uint32_t o[4];
                                    uint4 res = make_uint4(0, 0, 0, 0);
                                    if constexpr (EPI == kEpiResidual)
                                        res = *reinterpret_cast<const uint4*>(p.R + static_cast<size_t>(row) * p.ldr + col0 +
                                                                              c * 32 + v * 8);
                                    const uint32_t rr[4] = {res.x, res.y, res.z, res.w};
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
                                    *reinterpret_cast<uint4*>(p.C + off + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
                                }
                            }
                        }
                    }
                }
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (CG == 1) ptx::mbar_arrive(&tmem_empty_bar[acc]);
                    else ptx::mbar_arrive_cluster(&tmem_empty_bar[acc], 0);  // the leader's MMA thread waits on it
                }
                if (++acc == ACC_STAGES) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    }

    ptx::tc_fence_before_sync();
    __syncwarp();
    // a pair must leave together: the leader's MMAs read the peer's smem and signal its barriers
    if constexpr (CG == 2) ptx::cluster_sync_all(); else __syncthreads();
    if (warp_idx == 2) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc<CG>(tmem_base, 512);
    }
}

// ------------------------------------------------------------------------------------ host side
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

// Encoded descriptors are kept per (pointer, shape, box): a pass launches the same few operand shapes thousands of
// times (weight slots and workspace buffers do not move), and every encode is a driver call.
struct TmapKey {
    uintptr_t base;
    int rows, cols, ld, box_rows;
    bool operator==(const TmapKey& o) const {
        return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows;
    }
};
struct TmapKeyHash {
    size_t operator()(const TmapKey& k) const {
        size_t h = std::hash<uintptr_t>()(k.base);
        for (int v : {k.rows, k.cols, k.ld, k.box_rows}) h = h * 1000003u ^ std::hash<int>()(v);
        return h;
    }
};
std::mutex g_tmap_mu;
std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;

// Box = 64 columns (one 128-byte swizzle row) x box_rows.
int make_tmap(CUtensorMap* tm, const __half* base, int rows, int cols, int ld, int box_rows) {
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld % 8))
        return fail(kErrInvalid, "gemm operand must be 16-byte aligned with a row stride multiple of 8 (ld=%d)", ld);
    const TmapKey key{reinterpret_cast<uintptr_t>(base), rows, cols, ld, box_rows};
    {
        std::lock_guard<std::mutex> lk(g_tmap_mu);
        auto it = g_tmap_cache.find(key);
        if (it != g_tmap_cache.end()) {
            *tm = it->second;
            return kOk;
        }
    }
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * sizeof(__half)};
    const cuuint32_t box[2] = {BLOCK_K, static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(kErrCuda, "cuTensorMapEncodeTiled failed (CUresult %d)", static_cast<int>(r));
    std::lock_guard<std::mutex> lk(g_tmap_mu);
    if (g_tmap_cache.size() >= 4096) g_tmap_cache.clear();  // descriptors are values: dropping them is always safe
    g_tmap_cache.emplace(key, *tm);
    return kOk;
}

template <int EPI, int CG, bool TS>
int launch_t(const GemmArgs& a, int num_sms, cudaStream_t stream) {
    constexpr int OUT_COLS = EPI == kEpiSwiGLU ? BLOCK_N / 2 : BLOCK_N;
    constexpr int B_BOX = (EPI == kEpiSwiGLU || CG == 2) ? BLOCK_N / 2 : BLOCK_N;
    CUtensorMap tma, tmb;
    FLS_TRY(make_tmap(&tma, a.A, a.M, a.K, a.lda, BLOCK_M));
    FLS_TRY(make_tmap(&tmb, a.B, a.b_rows, a.K, a.ldb, B_BOX));
    CUtensorMap tmc = tma, tmr = tma;  // placeholders where the epilogue has no such operand
    if constexpr (TS && EPI != kEpiLsePartial) FLS_TRY(make_tmap(&tmc, a.C, a.M, a.N, a.ldc, 32));
    if constexpr (TS && EPI == kEpiResidual) FLS_TRY(make_tmap(&tmr, a.R, a.M, a.N, a.ldr, 32));
    KParams p;
    p.M = a.M, p.N = a.N, p.K = a.K;
    p.C = a.C, p.ldc = a.ldc, p.R = a.R, p.ldr = a.ldr;
    p.partial = a.partial, p.lse = a.lse, p.up_off = a.up_off;
    p.group_m = a.group_m > 0 ? a.group_m : GROUP_M;
    p.cache_hints = a.cache_hints;
    p.raster = a.raster;
    p.row_pos = a.row_pos, p.cos_tab = a.cos_tab, p.sin_tab = a.sin_tab, p.rope_cols = a.rope_cols;
    p.num_m_blocks = (a.M + BLOCK_M * CG - 1) / (BLOCK_M * CG);
    p.num_n_blocks = (a.N + OUT_COLS - 1) / OUT_COLS;
    const int tiles = p.num_m_blocks * p.num_n_blocks;
    const int workers = std::min(tiles, num_sms / CG);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(workers * CG);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = Cfg<CG>::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    FLS_CUDA(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<EPI, CG, TS>, tma, tmb, tmc, tmr, p));
    return kOk;
}

template <int EPI>
int launch_cg(const GemmArgs& a, int num_sms, int cta_group, cudaStream_t stream) {
    if (a.direct_store)
        return cta_group == 2 ? launch_t<EPI, 2, false>(a, num_sms, stream) : launch_t<EPI, 1, false>(a, num_sms, stream);
    return cta_group == 2 ? launch_t<EPI, 2, true>(a, num_sms, stream) : launch_t<EPI, 1, true>(a, num_sms, stream);
}

template <int EPI>
int set_smem_attr() {
    FLS_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<EPI, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg<1>::SMEM_BYTES));
    FLS_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<EPI, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg<2>::SMEM_BYTES));
    FLS_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<EPI, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg<1>::SMEM_BYTES));
    FLS_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<EPI, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg<2>::SMEM_BYTES));
    return kOk;
}

}  // namespace

int gemm_lse_blocks(int N) { return (N + BLOCK_N - 1) / BLOCK_N; }

int encode_tmap_2d(::CUtensorMap_st* tm, const __half* base, int rows, int cols, int ld, int box_rows) {
    if (!g_encode) return fail(kErrState, "gemm_init() has not run");
    return make_tmap(tm, base, rows, cols, ld, box_rows);
}

int gemm_init() {
    // contexts are created from one host thread per GPU and fls_op_head_noalloc sets itself up on first use: the
    // entry-point lookup and the per-device function attributes are serialised here, not by the callers
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        FLS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (!fn || qres != cudaDriverEntryPointSuccess)
            return fail(kErrCuda, "driver does not export cuTensorMapEncodeTiled");
        g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    }
    FLS_TRY(set_smem_attr<kEpiStore>());
    FLS_TRY(set_smem_attr<kEpiResidual>());
    FLS_TRY(set_smem_attr<kEpiSwiGLU>());
    FLS_TRY(set_smem_attr<kEpiLsePartial>());
    FLS_TRY(set_smem_attr<kEpiProbs>());
    FLS_TRY(set_smem_attr<kEpiStoreRope>());
    FLS_TRY(gemm_wide_init());
    return kOk;
}

int gemm_launch(GemmEpilogue epi, const GemmArgs& a, int num_sms, cudaStream_t stream, int cta_group) {
    if (!g_encode) return fail(kErrState, "gemm_init() has not run");
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return fail(kErrInvalid, "gemm: empty problem %dx%dx%d", a.M, a.N, a.K);
    if (a.K % 8 || a.N % 8) return fail(kErrInvalid, "gemm: N and K must be multiples of 8 (N=%d K=%d)", a.N, a.K);
    if (epi != kEpiLsePartial && (a.ldc % 8 || (reinterpret_cast<uintptr_t>(a.C) & 15)))
        return fail(kErrInvalid, "gemm: C must be 16-byte aligned with ldc %% 8 == 0");
    const bool wide = a.wide && cta_group == 2 && !a.direct_store && gemm_wide_supports(epi);
    switch (epi) {
        case kEpiStore:
            if (wide) return gemm_wide_launch(epi, a, num_sms, stream);
            return launch_cg<kEpiStore>(a, num_sms, cta_group, stream);
        case kEpiResidual:
            if (!a.R || a.ldr % 8) return fail(kErrInvalid, "gemm: residual epilogue needs R with ldr %% 8 == 0");
            if (wide) return gemm_wide_launch(epi, a, num_sms, stream);
            return launch_cg<kEpiResidual>(a, num_sms, cta_group, stream);
        case kEpiSwiGLU:
            if (a.b_rows < a.up_off + a.N) return fail(kErrInvalid, "gemm: SwiGLU needs gate and up rows in B");
            if (wide) return gemm_wide_launch(epi, a, num_sms, stream);
            return launch_cg<kEpiSwiGLU>(a, num_sms, cta_group, stream);
        case kEpiLsePartial:
            if (!a.partial) return fail(kErrInvalid, "gemm: lse epilogue needs a partial buffer");
            return launch_cg<kEpiLsePartial>(a, num_sms, cta_group, stream);
        case kEpiProbs:
            if (!a.lse) return fail(kErrInvalid, "gemm: probs epilogue needs lse");
            return launch_cg<kEpiProbs>(a, num_sms, cta_group, stream);
        case kEpiStoreRope:
            if (!a.row_pos || !a.cos_tab || !a.sin_tab || a.rope_cols % 128 || a.N % 128)
                return fail(kErrInvalid, "gemm: rope epilogue needs positions, tables and 128-wide heads");
            if (wide) return gemm_wide_launch(epi, a, num_sms, stream);
            return launch_cg<kEpiStoreRope>(a, num_sms, cta_group, stream);
    }
    return fail(kErrInvalid, "gemm: unknown epilogue %d", static_cast<int>(epi));
}

}  // namespace fls
