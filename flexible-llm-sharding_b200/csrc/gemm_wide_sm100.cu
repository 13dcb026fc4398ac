// tcgen05 GEMM with 256 x 512 tiles per CTA pair ("wide" tiles): C[M,N] = A[M,K] * B[N,K]^T + fused epilogue.
//
// Why a second tile shape.  With 256 x 256 tiles (gemm_sm100.cu) every k-block moves 256 rows of A and 256 rows of B
// from L2 into the pair's shared memory for 256 x 256 x 64 MACs.  At the Llama-2-70B operand sizes (K = 8192 and
// 28672) `ncu` showed 120 GB of L2->SM traffic and 28.5 GB of DRAM reads for one gate/up launch (algorithmic: 1.2 GB) -
// the tensor pipe is still 99 % busy when the kernel runs alone, but inside a pass the GPU sits at its 1 kW power
// cap and every byte moved is clock taken from the tensor pipe (profiles/r02_summary.md).  Here a pair keeps BOTH
// 256-column TMEM accumulators live at once: one A k-slice (128 rows per CTA) is staged once and multiplied with two
// neighbouring B tiles, so a 256 x 512 tile costs 256 + 512 operand rows per k-block instead of 2 x (256 + 256): a
// quarter less L2->SM traffic, and a wave of 74 pairs touches fewer distinct operand bands, which is what lets L2
// reuse work at K = 8192.  The price: no spare accumulator, so a tile's epilogue does not overlap the next tile's
// MMAs (the TMA producer still runs ahead and refills all stages meanwhile) - a few percent at K = 4096, less at
// larger K.
//
// Structure (as gemm_sm100.cu, cta_group::2 only): warp 0 = TMA producer (per CTA: its 128 A rows + its half of each
// of the two B tiles per stage, 4 stages of 48 KiB), warp 1 = MMA issuer of the leader CTA (tcgen05.mma
// cta_group::2 M256 N256 K16, alternating between the two accumulators), warp 2 = TMEM allocator, warps 4-7 =
// staged epilogue (gemm_tile.cuh: registers -> swizzled smem sub-tile -> TMA store; residual by TMA load).
#include <algorithm>

#include "gemm_tile.cuh"

namespace fls {

using namespace gemm;

namespace {

constexpr int W_STAGES = 4;
constexpr int W_B_HALF_BYTES = (BLOCK_N / 2) * BLOCK_K * 2;          // this CTA's 128 rows of one B tile: 16 KiB
constexpr int W_STAGE_BYTES = A_BYTES + 2 * W_B_HALF_BYTES;          // 48 KiB
constexpr int W_SMEM_BYTES = W_STAGES * W_STAGE_BYTES + EPI_STAGE_BYTES + SMEM_BARRIER_BYTES + 1024;

template <int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_wide_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_r,
                 const KParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* epi_stage = smem + W_STAGES * W_STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + EPI_STAGE_BYTES);
    uint64_t* empty_bar = full_bar + MAX_STAGES;
    uint64_t* tmem_full_bar = empty_bar + MAX_STAGES;   // both accumulators complete
    uint64_t* tmem_empty_bar = tmem_full_bar + 1;       // both accumulators drained by all 8 epilogue warps
    uint64_t* res_bar = tmem_empty_bar + 1;             // [4 warps][2 buffers]
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(res_bar + 8);

    const int warp_idx = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = ptx::cluster_ctarank();
    const bool is_leader = cta_rank == 0;

    if (warp_idx == 0 && ptx::elect_one()) {
        ptx::prefetch_tensormap(&tmap_a);
        ptx::prefetch_tensormap(&tmap_b);
        ptx::prefetch_tensormap(&tmap_c);
        if constexpr (EPI == kEpiResidual) ptx::prefetch_tensormap(&tmap_r);
    }
    if (warp_idx == 1 && ptx::elect_one()) {
        for (int i = 0; i < W_STAGES; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        ptx::mbar_init(tmem_full_bar, 1);
        ptx::mbar_init(tmem_empty_bar, 8);
        for (int i = 0; i < 8; ++i) ptx::mbar_init(&res_bar[i], 1);
        ptx::fence_mbar_init();
    }
    if (warp_idx == 2) {
        ptx::tmem_alloc<2>(tmem_base_slot, 512);
        ptx::tmem_relinquish<2>();
    }
    ptx::tc_fence_before_sync();
    ptx::cluster_sync_all();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_base_slot;

    constexpr int OUT_COLS = out_cols_per_tile<EPI>();      // output columns of ONE accumulator
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;  // m blocks of 256 rows x n blocks of two accumulators
    const int k_blocks = (p.K + BLOCK_K - 1) / BLOCK_K;
    const int worker = blockIdx.x / 2, n_workers = gridDim.x / 2;

    if (warp_idx == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        if (ptx::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t rank = cta_rank;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                int m_blk, n_blk;
                tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
                const int a_row = (m_blk * 2 + static_cast<int>(rank)) * BLOCK_M;
                int b_row[2];
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    const int n256 = n_blk * 2 + t;
                    if constexpr (EPI == kEpiSwiGLU)  // leader stages the gate rows, its peer the up rows
                        b_row[t] = n256 * OUT_COLS + static_cast<int>(rank) * p.up_off;
                    else
                        b_row[t] = n256 * BLOCK_N + static_cast<int>(rank) * (BLOCK_N / 2);
                }
                for (int kb = 0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * W_STAGE_BYTES;
                    // both CTAs' bytes are credited to the leader's barrier (peer bit cleared)
                    if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * W_STAGE_BYTES);
                    const uint32_t bar = ptx::smem_u32(&full_bar[stage]) & 0xFEFFFFFFu;
                    if (p.cache_hints == 0) {
                        ptx::tma_load_2d_cta2(sa, &tmap_a, bar, kb * BLOCK_K, a_row);
                        ptx::tma_load_2d_cta2(sa + A_BYTES, &tmap_b, bar, kb * BLOCK_K, b_row[0]);
                        ptx::tma_load_2d_cta2(sa + A_BYTES + W_B_HALF_BYTES, &tmap_b, bar, kb * BLOCK_K, b_row[1]);
                    } else {  // L2 eviction-priority experiments (same encoding as gemm_sm100.cu)
                        const uint64_t ha = p.cache_hints == 1 ? ptx::kL2EvictFirst
                                            : p.cache_hints == 3 ? ptx::kL2EvictNormal : ptx::kL2EvictLast;
                        const uint64_t hb = p.cache_hints == 2 ? ptx::kL2EvictFirst
                                            : p.cache_hints == 4 ? ptx::kL2EvictNormal : ptx::kL2EvictLast;
                        ptx::tma_load_2d_cta2_hint(sa, &tmap_a, bar, kb * BLOCK_K, a_row, ha);
                        ptx::tma_load_2d_cta2_hint(sa + A_BYTES, &tmap_b, bar, kb * BLOCK_K, b_row[0], hb);
                        ptx::tma_load_2d_cta2_hint(sa + A_BYTES + W_B_HALF_BYTES, &tmap_b, bar, kb * BLOCK_K, b_row[1], hb);
                    }
                    if (++stage == W_STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA only)
        if (is_leader && ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::umma_idesc_f16_f32(BLOCK_M * 2, BLOCK_N);
            int stage = 0;
            uint32_t phase = 0, tphase = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                ptx::mbar_wait(tmem_empty_bar, tphase ^ 1);     // the previous tile's epilogue has read both accumulators
                ptx::tc_fence_after_sync();
                for (int kb = 0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after_sync();
                    const uint32_t sa = ptx::smem_u32(smem + stage * W_STAGE_BYTES);
                    const uint64_t da = ptx::umma_smem_desc_sw128(sa);
                    const uint64_t db0 = ptx::umma_smem_desc_sw128(sa + A_BYTES);
                    const uint64_t db1 = ptx::umma_smem_desc_sw128(sa + A_BYTES + W_B_HALF_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        ptx::umma_f16<2>(tmem_base, da + 2 * k, db0 + 2 * k, idesc, (kb | k) != 0);
                        ptx::umma_f16<2>(tmem_base + BLOCK_N, da + 2 * k, db1 + 2 * k, idesc, (kb | k) != 0);
                    }
                    ptx::umma_commit_cta2(&empty_bar[stage], 0b11);   // frees the stage in both CTAs
                    if (++stage == W_STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                ptx::umma_commit_cta2(tmem_full_bar, 0b11);
                tphase ^= 1;
            }
        }
    } else if (warp_idx >= EPI_WARP0) {
        // ------------------------------------------------------------------ epilogue (both CTAs)
        const int ew = warp_idx - EPI_WARP0;  // TMEM lanes [32*ew, 32*ew+32)
        uint8_t* const buf_base = epi_stage + (ew * 2) * EPI_BUF_BYTES;
        uint64_t* rbar = res_bar + ew * 2;
        uint32_t q = 0;
        uint32_t tphase = 0;
        for (int tile = worker; tile < num_tiles; tile += n_workers) {
            int m_blk, n_blk;
            tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
            const int row0 = (m_blk * 2 + static_cast<int>(cta_rank)) * BLOCK_M + ew * 32;
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
            const int col_first = n_blk * 2 * OUT_COLS;
            if constexpr (EPI == kEpiResidual) epi_prefetch_residual(&tmap_r, buf_base, rbar, q, lane, col_first, row0);
            ptx::mbar_wait(tmem_full_bar, tphase);
            ptx::tc_fence_after_sync();
            tphase ^= 1;
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
                const int col0 = col_first + t * OUT_COLS;
                if (col0 >= p.N) break;       // odd number of 256-column blocks: the second accumulator is padding
                if constexpr (EPI == kEpiResidual)
                    if (t == 1) epi_prefetch_residual(&tmap_r, buf_base, rbar, q, lane, col0, row0);
                epi_accumulator<EPI>(p, &tmap_c, &tmap_r, taddr + t * BLOCK_N, row0, col0, buf_base, rbar, q, lane);
            }
            ptx::tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(tmem_empty_bar, 0);  // the leader's MMA thread waits on it
        }
        if (lane == 0) ptx::bulk_wait_group_read<0>();  // the staging buffers must outlive the last stores' reads
    }

    ptx::tc_fence_before_sync();
    __syncwarp();
    ptx::cluster_sync_all();  // the leader's MMAs read the peer's smem and signal its barriers: leave together
    if (warp_idx == 2) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc<2>(tmem_base, 512);
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// Skewed form of the 256 x 512 tile: the two accumulators are fed by two INDEPENDENT MMA streams that share the A ring.
// In the lock-step kernel above both accumulators finish together and the tensor pipe idles while the epilogue drains
// them one after the other (5-10 % of a K = 8192 tile).  Here each accumulator has its own issuing thread, its own B
// ring and its own full / empty barriers; an A slot is released when BOTH streams have used it.  The epilogue drains
// accumulator 0 first, so stream 0 restarts first and from then on runs a few k-blocks ahead of stream 1 (as far as the
// A ring lets it): at the end of a tile accumulator 0 is drained while stream 1 still multiplies, and accumulator 1
// while stream 0 already works on the next tile - the tensor pipe always has one stream to run.
//   warp 0   TMA producer: A ring (6 slots) + B ring of stream 0 (3 slots)
//   warp 1   MMA stream 0 (leader CTA): accumulator columns [0, 256)
//   warp 2   TMEM allocator, then TMA producer of stream 1's B ring (3 slots)
//   warp 3   MMA stream 1 (leader CTA): accumulator columns [256, 512)
//   warps 4-7 staged epilogue (gemm_tile.cuh), accumulator 0 then accumulator 1 of every tile
constexpr int S_NA = 6;
constexpr int S_NB = 3;
constexpr int S_BLK_BYTES = BLOCK_M * BLOCK_K * 2;  // one 128 x 64 fp16 operand block: 16 KiB
constexpr int S_SMEM_BYTES = (S_NA + 2 * S_NB) * S_BLK_BYTES + EPI_STAGE_BYTES + SMEM_BARRIER_BYTES + 1024;

template <int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_wide_skew_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_r,
                      const KParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_ring = smem;
    uint8_t* b_ring = smem + S_NA * S_BLK_BYTES;  // stream s: b_ring + s * S_NB * S_BLK_BYTES
    uint8_t* epi_stage = b_ring + 2 * S_NB * S_BLK_BYTES;
    uint64_t* a_full = reinterpret_cast<uint64_t*>(epi_stage + EPI_STAGE_BYTES);
    uint64_t* a_empty = a_full + S_NA;
    uint64_t* b_full = a_empty + S_NA;          // [2][S_NB]
    uint64_t* b_empty = b_full + 2 * S_NB;      // [2][S_NB]
    uint64_t* tmem_full_bar = b_empty + 2 * S_NB;   // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;   // [2]
    uint64_t* res_bar = tmem_empty_bar + 2;         // [4 warps][2 buffers]
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(res_bar + 8);

    const int warp_idx = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = ptx::cluster_ctarank();
    const bool is_leader = cta_rank == 0;

    if (warp_idx == 0 && ptx::elect_one()) {
        ptx::prefetch_tensormap(&tmap_a);
        ptx::prefetch_tensormap(&tmap_b);
        ptx::prefetch_tensormap(&tmap_c);
        if constexpr (EPI == kEpiResidual) ptx::prefetch_tensormap(&tmap_r);
    }
    if (warp_idx == 1 && ptx::elect_one()) {
        for (int i = 0; i < S_NA; ++i) {
            ptx::mbar_init(&a_full[i], 1);
            ptx::mbar_init(&a_empty[i], 2);   // one tcgen05.commit from each MMA stream
        }
        for (int i = 0; i < 2 * S_NB; ++i) {
            ptx::mbar_init(&b_full[i], 1);
            ptx::mbar_init(&b_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&tmem_full_bar[i], 1);
            ptx::mbar_init(&tmem_empty_bar[i], 8);   // 4 epilogue warps of both CTAs
        }
        for (int i = 0; i < 8; ++i) ptx::mbar_init(&res_bar[i], 1);
        ptx::fence_mbar_init();
    }
    if (warp_idx == 2) {
        ptx::tmem_alloc<2>(tmem_base_slot, 512);
        ptx::tmem_relinquish<2>();
    }
    ptx::tc_fence_before_sync();
    ptx::cluster_sync_all();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_base_slot;

    constexpr int OUT_COLS = out_cols_per_tile<EPI>();
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;
    const int k_blocks = (p.K + BLOCK_K - 1) / BLOCK_K;
    const int worker = blockIdx.x / 2, n_workers = gridDim.x / 2;

    // first B row this CTA stages for accumulator t of the tile whose n block is n_blk
    auto b_row_of = [&](int n_blk, int t) {
        const int n256 = n_blk * 2 + t;
        if constexpr (EPI == kEpiSwiGLU) return n256 * OUT_COLS + static_cast<int>(cta_rank) * p.up_off;
        else return n256 * BLOCK_N + static_cast<int>(cta_rank) * (BLOCK_N / 2);
    };

    if (warp_idx == 0 || warp_idx == 2) {
        // ------------------------------------------------------------------ TMA producers (both CTAs)
        if (ptx::elect_one()) {
            const int s = warp_idx >> 1;            // warp 0: A ring + stream 0's B ring; warp 2: stream 1's B ring
            uint64_t* bf = b_full + s * S_NB;
            uint64_t* be = b_empty + s * S_NB;
            uint8_t* br = b_ring + s * S_NB * S_BLK_BYTES;
            int ia = 0, ib = 0;
            uint32_t pa = 0, pb = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                int m_blk, n_blk;
                tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
                const int a_row = (m_blk * 2 + static_cast<int>(cta_rank)) * BLOCK_M;
                const int b_row = b_row_of(n_blk, s);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    if (s == 0) {
                        ptx::mbar_wait(&a_empty[ia], pa ^ 1);
                        if (is_leader) ptx::mbar_arrive_expect_tx(&a_full[ia], 2 * S_BLK_BYTES);
                        ptx::tma_load_2d_cta2(a_ring + ia * S_BLK_BYTES, &tmap_a, ptx::smem_u32(&a_full[ia]) & 0xFEFFFFFFu,
                                              kb * BLOCK_K, a_row);
                        if (++ia == S_NA) ia = 0, pa ^= 1;
                    }
                    ptx::mbar_wait(&be[ib], pb ^ 1);
                    if (is_leader) ptx::mbar_arrive_expect_tx(&bf[ib], 2 * S_BLK_BYTES);
                    ptx::tma_load_2d_cta2(br + ib * S_BLK_BYTES, &tmap_b, ptx::smem_u32(&bf[ib]) & 0xFEFFFFFFu, kb * BLOCK_K,
                                          b_row);
                    if (++ib == S_NB) ib = 0, pb ^= 1;
                }
            }
        }
    } else if (warp_idx == 1 || warp_idx == 3) {
        // ------------------------------------------------------------------ MMA streams (leader CTA only)
        if (is_leader && ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::umma_idesc_f16_f32(BLOCK_M * 2, BLOCK_N);
            const int s = warp_idx >> 1;
            uint64_t* bf = b_full + s * S_NB;
            uint64_t* be = b_empty + s * S_NB;
            const uint32_t br = ptx::smem_u32(b_ring + s * S_NB * S_BLK_BYTES);
            const uint32_t ar = ptx::smem_u32(a_ring);
            const uint32_t tmem_d = tmem_base + s * BLOCK_N;
            int ia = 0, ib = 0;
            uint32_t pa = 0, pb = 0, tphase = 0;
            for (int tile = worker; tile < num_tiles; tile += n_workers) {
                ptx::mbar_wait(&tmem_empty_bar[s], tphase ^ 1);   // the epilogue has read this accumulator
                ptx::tc_fence_after_sync();
                for (int kb = 0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&a_full[ia], pa);
                    ptx::mbar_wait(&bf[ib], pb);
                    ptx::tc_fence_after_sync();
                    const uint64_t da = ptx::umma_smem_desc_sw128(ar + ia * S_BLK_BYTES);
                    const uint64_t db = ptx::umma_smem_desc_sw128(br + ib * S_BLK_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                        ptx::umma_f16<2>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    ptx::umma_commit_cta2(&a_empty[ia], 0b11);   // the A slot is free once BOTH streams have said so
                    ptx::umma_commit_cta2(&be[ib], 0b11);
                    if (++ia == S_NA) ia = 0, pa ^= 1;
                    if (++ib == S_NB) ib = 0, pb ^= 1;
                }
                ptx::umma_commit_cta2(&tmem_full_bar[s], 0b11);
                tphase ^= 1;
            }
        }
    } else if (warp_idx >= EPI_WARP0) {
        // ------------------------------------------------------------------ epilogue (both CTAs)
        const int ew = warp_idx - EPI_WARP0;
        uint8_t* const buf_base = epi_stage + (ew * 2) * EPI_BUF_BYTES;
        uint64_t* rbar = res_bar + ew * 2;
        uint32_t q = 0, tphase = 0;
        for (int tile = worker; tile < num_tiles; tile += n_workers) {
            int m_blk, n_blk;
            tile_coords(tile, p.num_m_blocks, p.num_n_blocks, p.group_m, p.raster, m_blk, n_blk);
            const int row0 = (m_blk * 2 + static_cast<int>(cta_rank)) * BLOCK_M + ew * 32;
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
                const int col0 = (n_blk * 2 + t) * OUT_COLS;
                const bool live = col0 < p.N;   // odd number of 256-column blocks: the last second accumulator is padding
                if constexpr (EPI == kEpiResidual)
                    if (live) epi_prefetch_residual(&tmap_r, buf_base, rbar, q, lane, col0, row0);
                ptx::mbar_wait(&tmem_full_bar[t], tphase);
                ptx::tc_fence_after_sync();
                if (live) epi_accumulator<EPI>(p, &tmap_c, &tmap_r, taddr + t * BLOCK_N, row0, col0, buf_base, rbar, q, lane);
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(&tmem_empty_bar[t], 0);
            }
            tphase ^= 1;
        }
        if (lane == 0) ptx::bulk_wait_group_read<0>();
    }

    ptx::tc_fence_before_sync();
    __syncwarp();
    ptx::cluster_sync_all();
    if (warp_idx == 2) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc<2>(tmem_base, 512);
    }
}

template <int EPI, bool SKEW>
int launch_wide_t(const GemmArgs& a, int num_sms, cudaStream_t stream) {
    constexpr int OUT_COLS = EPI == kEpiSwiGLU ? BLOCK_N / 2 : BLOCK_N;
    CUtensorMap tma, tmb, tmc, tmr;
    FLS_TRY(encode_tmap_2d(&tma, a.A, a.M, a.K, a.lda, BLOCK_M));
    FLS_TRY(encode_tmap_2d(&tmb, a.B, a.b_rows, a.K, a.ldb, BLOCK_N / 2));
    FLS_TRY(encode_tmap_2d(&tmc, a.C, a.M, a.N, a.ldc, 32));
    tmr = tmc;
    if constexpr (EPI == kEpiResidual) FLS_TRY(encode_tmap_2d(&tmr, a.R, a.M, a.N, a.ldr, 32));
    KParams p{};
    p.M = a.M, p.N = a.N, p.K = a.K;
    p.C = a.C, p.ldc = a.ldc, p.R = a.R, p.ldr = a.ldr;
    p.partial = a.partial, p.lse = a.lse, p.up_off = a.up_off;
    p.group_m = a.group_m > 0 ? a.group_m : GROUP_M;
    p.cache_hints = a.cache_hints;
    p.raster = a.raster;
    p.row_pos = a.row_pos, p.cos_tab = a.cos_tab, p.sin_tab = a.sin_tab, p.rope_cols = a.rope_cols;
    p.num_m_blocks = (a.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
    p.num_n_blocks = (a.N + 2 * OUT_COLS - 1) / (2 * OUT_COLS);
    const int tiles = p.num_m_blocks * p.num_n_blocks;
    const int workers = std::min(tiles, num_sms / 2);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(workers * 2);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = SKEW ? S_SMEM_BYTES : W_SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if constexpr (SKEW) FLS_CUDA(cudaLaunchKernelEx(&cfg, gemm_wide_skew_kernel<EPI>, tma, tmb, tmc, tmr, p));
    else FLS_CUDA(cudaLaunchKernelEx(&cfg, gemm_wide_kernel<EPI>, tma, tmb, tmc, tmr, p));
    return kOk;
}

template <int EPI>
int set_wide_attr() {
    FLS_CUDA(cudaFuncSetAttribute(gemm_wide_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, W_SMEM_BYTES));
    FLS_CUDA(cudaFuncSetAttribute(gemm_wide_skew_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, S_SMEM_BYTES));
    return kOk;
}

}  // namespace

int gemm_wide_init() {
    FLS_TRY(set_wide_attr<kEpiStore>());
    FLS_TRY(set_wide_attr<kEpiResidual>());
    FLS_TRY(set_wide_attr<kEpiSwiGLU>());
    FLS_TRY(set_wide_attr<kEpiStoreRope>());
    return kOk;
}

bool gemm_wide_supports(GemmEpilogue epi) {
    return epi == kEpiStore || epi == kEpiResidual || epi == kEpiSwiGLU || epi == kEpiStoreRope;
}

int gemm_wide_launch(GemmEpilogue epi, const GemmArgs& a, int num_sms, cudaStream_t stream) {
    const bool skew = a.wide != 3;   // 3 = the lock-step form (comparator)
    switch (epi) {
        case kEpiStore: return skew ? launch_wide_t<kEpiStore, true>(a, num_sms, stream) : launch_wide_t<kEpiStore, false>(a, num_sms, stream);
        case kEpiResidual:
            return skew ? launch_wide_t<kEpiResidual, true>(a, num_sms, stream) : launch_wide_t<kEpiResidual, false>(a, num_sms, stream);
        case kEpiSwiGLU: return skew ? launch_wide_t<kEpiSwiGLU, true>(a, num_sms, stream) : launch_wide_t<kEpiSwiGLU, false>(a, num_sms, stream);
        case kEpiStoreRope:
            return skew ? launch_wide_t<kEpiStoreRope, true>(a, num_sms, stream) : launch_wide_t<kEpiStoreRope, false>(a, num_sms, stream);
        default: return fail(kErrInvalid, "gemm (wide tiles): epilogue %d has no wide variant", static_cast<int>(epi));
    }
}

}  // namespace fls
