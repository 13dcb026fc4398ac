// Shared-prefix prefill attention on tcgen05 / TMEM / TMA (head_dim 128).
//
// Same contract as attention.cu (see there and include/fls_b200.h: fls_attn_tile); replaces the
// reference's two eager attention calls per layer (reference utils.py:272 and :275-279).
//
// Two kernels with the same arithmetic: attention_tcgen05_kernel<NK> launches one CTA per work item (query tile x head);
// attention_persistent_kernel (further down, the default) keeps two CTAs per SM resident and lets them fetch items.
//
// One work item = one query tile (<= 128 rows) x one head, handled by a CTA of 192 threads:
//   warp 0      TMA producer: Q tile once (2 x [128 x 64] sub-tiles, 128B swizzle), then K and V tiles of
//               64 keys (each 2 x [64 x 64] sub-tiles) through two rings: NK = 3 buffers for K (which later
//               hold P), 2 for V.  A key tile never straddles the two key segments (prefix rows / option rows).
//   warp 1      MMA issuer (one elected thread):  S = Q K^T  (tcgen05.mma M128 N64 K16 x 8, K-major
//               operands) into TMEM columns [0,64);  O += P V (M128 N128 K16 x 4, A = P K-major, B = V
//               MN-major exactly as it lies in memory - no transpose) into TMEM columns [128,256).
//   warps 2-5   softmax, one query row per thread: tcgen05.ld S, mask, online softmax in fp32
//               (base-2), P rounded to fp16 (as HF's softmax(..).to(fp16) @ V) and written with the
//               128B swizzle over the K tile it came from (K is dead once S is complete), O row
//               rescaled in TMEM only when the running max grows by more than 2^8; final O / l staged
//               through the (by then dead) Q buffer and written as full 128-byte lines.
// S is double buffered: Q K^T of tile j+1 is issued before P V of tile j, so it runs while the
// softmax warps work on tile j.  `pv_done` tells the softmax warps when O may be rescaled/read.
// Why K and V have rings of their own: a K buffer is busy from its load until P V of its tile has retired (P is
// written over K), so with K and V sharing two stages the load of tile j+2 could only start once softmax j had
// finished, and its latency was exposed: the stall samples of `ncu` showed the softmax warps waiting for S 40 % of
// their time (profiles/r02_summary.md).  A third K buffer lets K(j+2) arrive a whole softmax period earlier; V(j+2)
// is not needed before P V of tile j+2 and does with two.
// Shared memory: Q 32 KiB + 3 x 16 KiB + 2 x 16 KiB = 112 KiB -> still two CTAs per SM (exactly: the launch falls
// back to NK = 2 where the driver says two do not fit), whose softmax and MMA phases overlap; TMEM: 256 columns.
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace fls {

namespace {

constexpr int D = 128;
constexpr int BQ = 128;
constexpr int BKV = 64;
constexpr int V_STAGES = 2;
constexpr int MAX_K_STAGES = 3;
constexpr int Q_BYTES = BQ * D * 2;                  // 32 KiB: two [128 x 64] sub-tiles
constexpr int KSUB_BYTES = BKV * 64 * 2;             // 8 KiB: one [64 keys x 64 d] sub-tile
constexpr int TILE_BYTES = 2 * KSUB_BYTES;           // one K (or P) tile, one V tile: sub0, sub1
constexpr int BARRIER_BYTES = 256;
// Shared memory a CTA uses, from the first 1 KiB boundary (the 128B swizzle needs it) of its dynamic allocation on.
// Two CTAs per SM have 2 x 113 KiB between them: NK = 3 leaves 768 bytes for lining up, so the launch asks for what
// the device's allocations really need (measured once by a probe kernel: 0 on B200) and the kernel traps if the
// buffers would not fit what it was given.
constexpr int smem_used(int nk) { return Q_BYTES + (nk + V_STAGES) * TILE_BYTES + BARRIER_BYTES; }
constexpr int PERSIST_SMEM_USED = Q_BYTES + (3 + V_STAGES) * TILE_BYTES + 512;   // + the two work-item slots
constexpr int THREADS = 192;
constexpr int TMEM_COLS = 256;
constexpr int S_COL = 0, O_COL = 128;  // S: two buffers of 64 columns at 0 and 64; O: 128 columns
constexpr float RESCALE_THRESHOLD = 8.0f;  // log2 units

__device__ __forceinline__ float ex2_approx(float x) {  // 2^x, one MUFU op; 2^-inf = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- pieces shared by the two kernels (one thread = one query row of the tile) ------------------------------------
// Visible keys of query row `i` of tile `t`: segment 0 keys [0, lim0), segment 1 keys [lo1, hi1].
struct RowKeys {
    int lim0, lo1, hi1;
};
__device__ __forceinline__ RowKeys row_keys(const AttnTile& t, int i) {
    RowKeys r;
    r.lim0 = t.mode == kAttnPrefixCausal ? min(t.seg0_len, t.q_seq0 + i + 1) : t.seg0_len;
    r.lo1 = 0;
    r.hi1 = -1;
    if (t.mode == kAttnSuffix) {
        const int qa = (t.q_row - t.seg1_row) + i, L = t.q_seq0;
        r.lo1 = (qa / L) * L;
        r.hi1 = min(qa, t.seg1_len - 1);           // own option, positions <= own
    }
    return r;
}

// One key tile of the online softmax, from "S is complete" to "P is published":
//   s_addr / o_addr  this warp's lanes of the S buffer / of O in TMEM
//   kbase, lo, hi    first key of the tile and this row's visible range, all in the key segment's own numbering
//   m_run, l_run     running reference max (RAW score units; p = exp2(s * c - m_run * c), c = scale * log2 e) and sum
//   p_row            this row's 128 bytes of the P tile (= the dead K tile), `i` = the row (128B swizzle by i & 7)
//   prev_pv          barrier of the previous tile's P V (nullptr for an item's first tile) and its parity: O may be
//                    rescaled only once that has retired
__device__ __forceinline__ void softmax_key_tile(uint32_t s_addr, uint32_t o_addr, int kbase, int lo, int hi,
                                                 float scale_log2e, float thr, float& m_run, float& l_run, uint8_t* p_row,
                                                 int i, uint64_t* prev_pv, uint32_t prev_parity, uint64_t* p_ready,
                                                 int lane) {
    uint32_t sr[2][32];
    ptx::tmem_ld_32x32(s_addr, sr[0]);
    ptx::tmem_ld_32x32(s_addr + 32, sr[1]);
    ptx::tmem_ld_wait();
    // masking only where a row of this warp has an invisible key in the tile (warp-uniform branch)
    if (__any_sync(0xffffffffu, kbase < lo || kbase + BKV - 1 > hi)) {
#pragma unroll
        for (int c = 0; c < 64; ++c) {
            const int key = kbase + c;
            if (key < lo || key > hi) sr[c >> 5][c & 31] = 0xff800000u;  // -inf
        }
    }
    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // four short chains instead of one long one
#pragma unroll
    for (int c = 0; c < 64; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], __uint_as_float(sr[c >> 5][c & 31]));
    const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
    // lazy rescale: keep the old reference max unless the new one is much larger
    float m_new = m_run;
    if (mx > m_run + thr) m_new = mx;
    const bool changed = m_new != m_run;
    const float m_use = m_new == -INFINITY ? 0.f : m_new;
    const float alpha = changed ? ex2_approx((m_run - m_use) * scale_log2e) : 1.f;   // m_run = -inf -> 0
    m_run = m_new;
    const float mc = m_use * scale_log2e;
    float sum = 0.f;
    uint32_t pk[32];
#pragma unroll
    for (int c = 0; c < 64; c += 2) {
        const float p0 = ex2_approx(fmaf(__uint_as_float(sr[c >> 5][c & 31]), scale_log2e, -mc));
        const float p1 = ex2_approx(fmaf(__uint_as_float(sr[(c + 1) >> 5][(c + 1) & 31]), scale_log2e, -mc));
        sum += p0 + p1;
        const __half2 h = __floats2half2_rn(p0, p1);
        pk[c >> 1] = *reinterpret_cast<const uint32_t*>(&h);
    }
    l_run = l_run * alpha + sum;
    // P -> shared memory over the (dead) K tile: [128 rows][64 keys] fp16, 128B swizzle.  Issued before the wait below
    // so that the stores complete under it; the fence that publishes them comes after.
#pragma unroll
    for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(p_row + ((c ^ (i & 7)) << 4)) =
            make_uint4(pk[c * 4], pk[c * 4 + 1], pk[c * 4 + 2], pk[c * 4 + 3]);
    if (prev_pv) {
        // O is touched by the tensor core until P V of the previous tile has retired
        ptx::mbar_wait(prev_pv, prev_parity);
        ptx::tc_fence_after_sync();
        // rescale this row of O (TMEM) if any row of the warp moved its reference max
        if (__any_sync(0xffffffffu, changed)) {
#pragma unroll 1
            for (int c = 0; c < D / 32; ++c) {
                uint32_t o[32];
                ptx::tmem_ld_32x32(o_addr + c * 32, o);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
                ptx::tmem_st_32x32(o_addr + c * 32, o);
            }
            ptx::tmem_st_wait();
        }
    }
    ptx::fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core's async proxy
    ptx::tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(p_ready);
}

// O / l -> fp16 -> global for this warp's 32 query rows, 64 columns at a time, staged in `stage` (4 KiB on a 1 KiB
// boundary; 16-byte pieces XOR-swizzled by row = the TMA 128B swizzle).  A warp whose 32 rows all belong to the tile
// hands the staged piece to one TMA store when it has a tensor map of the output (`tmap_o`, box 64 columns x 32 rows);
// otherwise it copies the valid rows out itself as full 128-byte lines, 4 rows per store instruction.
__device__ __forceinline__ void write_out(uint32_t o_addr, float inv, uint8_t* stage, int rows_here, __half* out_rows,
                                          int ld_out, int lane, const CUtensorMap* tmap_o, int o_col, int o_row) {
    const bool by_tma = tmap_o != nullptr && rows_here == 32;
#pragma unroll 1
    for (int hcol = 0; hcol < D; hcol += 64) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            uint32_t o[32];
            ptx::tmem_ld_32x32(o_addr + hcol + c * 32, o);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                uint32_t wv[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const __half2 h = __floats2half2_rn(__uint_as_float(o[v * 8 + e * 2]) * inv,
                                                        __uint_as_float(o[v * 8 + e * 2 + 1]) * inv);
                    wv[e] = *reinterpret_cast<const uint32_t*>(&h);
                }
                *reinterpret_cast<uint4*>(stage + lane * 128 + (((c * 4 + v) ^ (lane & 7)) << 4)) =
                    make_uint4(wv[0], wv[1], wv[2], wv[3]);
            }
        }
        if (by_tma) {
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                ptx::tma_store_2d(tmap_o, stage, o_col + hcol, o_row);
                ptx::bulk_commit_group();
                ptx::bulk_wait_group_read<0>();      // the piece is written again by the next half / by its next owner
            }
            __syncwarp();
        } else {
            __syncwarp();
#pragma unroll
            for (int rr = 0; rr < 32; rr += 4) {
                const int r = rr + (lane >> 3);                     // row inside this warp's 32
                if (r < rows_here) {
                    const uint4 v4 = *reinterpret_cast<const uint4*>(stage + r * 128 + (((lane & 7) ^ (r & 7)) << 4));
                    *reinterpret_cast<uint4*>(out_rows + static_cast<size_t>(r) * ld_out + hcol + (lane & 7) * 8) = v4;
                }
            }
            __syncwarp();
        }
    }
}

template <int NK>
__global__ void __launch_bounds__(THREADS, 2)
attention_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                         int n_q_heads, int n_kv_heads, const AttnTile* __restrict__ tiles, __half* __restrict__ out,
                         int ld_out, float scale_log2e) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = smem + Q_BYTES;                  // NK tiles: K(j), later P(j), in buffer j % NK
    uint8_t* sV = sK + NK * TILE_BYTES;            // V(j) in buffer j % V_STAGES
    uint64_t* q_full = reinterpret_cast<uint64_t*>(sV + V_STAGES * TILE_BYTES);
    uint64_t* k_full = q_full + 1;
    uint64_t* k_empty = k_full + MAX_K_STAGES;
    uint64_t* v_full = k_empty + MAX_K_STAGES;
    uint64_t* v_empty = v_full + V_STAGES;
    uint64_t* s_full = v_empty + V_STAGES;     // [2]: S buffer b = tile & 1 is complete
    uint64_t* p_ready = s_full + 2;
    uint64_t* pv_done = p_ready + 1;           // O += P V of a tile has retired
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 1);

    uint32_t dyn_bytes;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_bytes));
    if (reinterpret_cast<uint8_t*>(tmem_slot + 1) > smem_raw + dyn_bytes) __trap();   // buffers do not fit
    const int warp_idx = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const AttnTile t = tiles[blockIdx.x];
    const int head = blockIdx.y;
    const int kvh = head / (n_q_heads / n_kv_heads);
    const int q_col = head * D;
    const int k_col = (n_q_heads + kvh) * D;
    const int v_col = (n_q_heads + n_kv_heads + kvh) * D;

    // key tiles: segment 0 first (only up to the diagonal for a causal prefix tile), then segment 1
    const int seg0_keys = t.mode == kAttnPrefixCausal ? min(t.seg0_len, t.q_seq0 + t.n_q) : t.seg0_len;
    const int n0 = (seg0_keys + BKV - 1) / BKV;
    const int n1 = t.mode == kAttnSuffix ? (t.seg1_len + BKV - 1) / BKV : 0;
    const int n_tiles = n0 + n1;

    auto load_k = [&](int j) {      // K(j) -> buffer j % NK (the caller has made sure it is free)
        const int ks = j % NK;
        const int row = j < n0 ? t.seg0_row + j * BKV : t.seg1_row + (j - n0) * BKV;
        uint8_t* kb = sK + ks * TILE_BYTES;
        ptx::mbar_arrive_expect_tx(&k_full[ks], TILE_BYTES);
        ptx::tma_load_2d(kb, &tmap_kv, &k_full[ks], k_col, row);
        ptx::tma_load_2d(kb + KSUB_BYTES, &tmap_kv, &k_full[ks], k_col + 64, row);
    };
    auto load_v = [&](int j) {
        const int vs = j % V_STAGES;
        const int row = j < n0 ? t.seg0_row + j * BKV : t.seg1_row + (j - n0) * BKV;
        uint8_t* vb = sV + vs * TILE_BYTES;
        ptx::mbar_arrive_expect_tx(&v_full[vs], TILE_BYTES);
        ptx::tma_load_2d(vb, &tmap_kv, &v_full[vs], v_col, row);
        ptx::tma_load_2d(vb + KSUB_BYTES, &tmap_kv, &v_full[vs], v_col + 64, row);
    };

    if (warp_idx == 0) {
        if (ptx::elect_one()) {
            ptx::prefetch_tensormap(&tmap_q);
            ptx::prefetch_tensormap(&tmap_kv);
            ptx::mbar_init(q_full, 1);
            for (int i = 0; i < NK; ++i) {
                ptx::mbar_init(&k_full[i], 1);
                ptx::mbar_init(&k_empty[i], 1);
            }
            for (int i = 0; i < V_STAGES; ++i) {
                ptx::mbar_init(&v_full[i], 1);
                ptx::mbar_init(&v_empty[i], 1);
            }
            ptx::mbar_init(&s_full[0], 1);
            ptx::mbar_init(&s_full[1], 1);
            ptx::mbar_init(p_ready, 4);  // one arrive per softmax warp
            ptx::mbar_init(pv_done, 1);
            ptx::fence_mbar_init();
            // Q and the first key tiles (every buffer is free) are requested right here, before the TMEM allocation and
            // the CTA-wide barrier: their latency is the longest thing a short-lived CTA waits for.
            ptx::mbar_arrive_expect_tx(q_full, Q_BYTES);
            ptx::tma_load_2d(sQ, &tmap_q, q_full, q_col, t.q_row);
            ptx::tma_load_2d(sQ + Q_BYTES / 2, &tmap_q, q_full, q_col + 64, t.q_row);
            for (int j = 0; j < min(n_tiles, NK); ++j) {
                load_k(j);
                if (j < V_STAGES) load_v(j);
            }
        }
        __syncwarp();
        ptx::tmem_alloc<1>(tmem_slot, TMEM_COLS);
        ptx::tmem_relinquish<1>();
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp_idx == 0) {
        // ------------------------------------------------------------------ TMA producer (the rest of the key tiles)
        // K(j) and V(j-1) wait for the same event when NK = 3 (P V of tile j-3 retired), K(j) and V(j) when NK = 2, so
        // issuing the loads in this one order never holds a load back behind a later release.
        if (ptx::elect_one()) {
            for (int j = V_STAGES; j < n_tiles; ++j) {
                if (j >= NK) {
                    ptx::mbar_wait(&k_empty[j % NK], ((j / NK) & 1) ^ 1);
                    load_k(j);
                }
                ptx::mbar_wait(&v_empty[j % V_STAGES], ((j / V_STAGES) & 1) ^ 1);
                load_v(j);
            }
        }
    } else if (warp_idx == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (ptx::elect_one()) {
            constexpr uint32_t idesc_qk = ptx::umma_idesc_f16_f32(BQ, BKV);
            constexpr uint32_t idesc_pv = ptx::umma_idesc_f16_f32(BQ, D) | ptx::kUmmaBMajorMN;
            const uint32_t q_addr = ptx::smem_u32(sQ);
            ptx::mbar_wait(q_full, 0);
            // S(j) = Q K(j)^T into S buffer j & 1: d = 128 -> 8 K-steps, 4 per 64-wide sub-tile
            auto issue_qk = [&](int j) {
                const int ks = j % NK;
                const uint32_t st = ptx::smem_u32(sK + ks * TILE_BYTES);
                ptx::mbar_wait(&k_full[ks], (j / NK) & 1);
                ptx::tc_fence_after_sync();
#pragma unroll
                for (int k = 0; k < D / 16; ++k) {
                    const uint64_t da = ptx::umma_smem_desc_sw128(q_addr + (k / 4) * (Q_BYTES / 2)) + 2 * (k % 4);
                    const uint64_t db = ptx::umma_smem_desc_sw128(st + (k / 4) * KSUB_BYTES) + 2 * (k % 4);
                    ptx::umma_f16<1>(tmem_base + S_COL + (j & 1) * BKV, da, db, idesc_qk, k != 0);
                }
                ptx::umma_commit(&s_full[j & 1]);
            };
            if (n_tiles > 0) issue_qk(0);
            for (int j = 0; j < n_tiles; ++j) {
                // the next tile's scores are computed while the softmax warps work on this one
                if (j + 1 < n_tiles) issue_qk(j + 1);
                const int ks = j % NK, vs = j % V_STAGES;
                const uint32_t st = ptx::smem_u32(sK + ks * TILE_BYTES);
                const uint32_t vt = ptx::smem_u32(sV + vs * TILE_BYTES);
                // O += P V : P (fp16, K-major) now lies where K(j) was; V is MN-major as loaded
                ptx::mbar_wait(&v_full[vs], (j / V_STAGES) & 1);
                ptx::mbar_wait(p_ready, j & 1);
                ptx::tc_fence_after_sync();
#pragma unroll
                for (int k = 0; k < BKV / 16; ++k) {
                    const uint64_t da = ptx::umma_smem_desc_sw128(st) + 2 * k;
                    const uint64_t db = ptx::umma_smem_desc_sw128_mn(vt + k * 16 * 128, KSUB_BYTES, 1024);
                    ptx::umma_f16<1>(tmem_base + O_COL, da, db, idesc_pv, (j | k) != 0);
                }
                ptx::umma_commit(&k_empty[ks]);
                ptx::umma_commit(&v_empty[vs]);
                ptx::umma_commit(pv_done);
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax (warps 2..5)
        const int quarter = warp_idx & 3;             // TMEM lanes [32*quarter, +32) belong to this warp
        const int i = quarter * 32 + lane;            // query row inside the tile
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
        const RowKeys rk = row_keys(t, i);
        float m_run = -INFINITY, l_run = 0.f;
        const float thr = RESCALE_THRESHOLD / scale_log2e;
        for (int j = 0; j < n_tiles; ++j) {
            ptx::mbar_wait(&s_full[j & 1], (j >> 1) & 1);
            ptx::tc_fence_after_sync();
            const bool in0 = j < n0;
            softmax_key_tile(lane_addr + S_COL + (j & 1) * BKV, lane_addr + O_COL, in0 ? j * BKV : (j - n0) * BKV,
                             in0 ? 0 : rk.lo1, in0 ? rk.lim0 - 1 : rk.hi1, scale_log2e, thr, m_run, l_run,
                             sK + (j % NK) * TILE_BYTES + i * 128, i, j > 0 ? pv_done : nullptr, (j - 1) & 1, p_ready, lane);
        }
        // ---- write-out.  Every Q K^T has retired by now, so the Q buffer is free: each warp stages its rows there.
        if (n_tiles > 0) ptx::mbar_wait(pv_done, (n_tiles - 1) & 1);
        ptx::tc_fence_after_sync();
        const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
        write_out(lane_addr + O_COL, inv, sQ + quarter * 4096, min(32, t.n_q - quarter * 32),
                  out + static_cast<size_t>(t.q_row + quarter * 32) * ld_out + head * D, ld_out, lane, nullptr, 0, 0);
    }

    ptx::tc_fence_before_sync();
    __syncwarp();
    __syncthreads();
    if (warp_idx == 0) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc<1>(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------ persistent form
// The same arithmetic, instruction for instruction, as attention_tcgen05_kernel<3>; what changes is who does the
// work: 2 x #SMs CTAs stay resident and fetch (query tile, head) items from a counter in global memory, in the
// order the grid above would have launched them.  On the benchmark's prompt shape a CTA of the grid form spends a
// third of its life outside the key-tile loop (barrier set-up, TMEM allocation, the first Q/K round trip, the
// write-out: tools/attn_probe.py --long reaches 1000 useful TFLOP/s with 30 key tiles per CTA against 650 with 6.5).
// Here the set-up happens once per CTA, and while the softmax warps write item i out the producer already has
// item i+1's Q and first key tiles in flight and the MMA warp its first S:
//   * the K/V rings, S buffers and the p_ready / pv_done phases simply keep counting across items (`gt` = key tiles
//     handled so far by this CTA);
//   * Q is free once the last Q K^T of an item has retired (`q_empty`, signalled by tcgen05.commit), because the
//     write-out is staged in the K/P buffer of the item's LAST key tile instead of the Q buffer; that buffer goes
//     back to the producer when the fourth softmax warp has finished reading it (it, not the MMA warp, arrives
//     on k_empty for that tile);
//   * O may be overwritten by the next item's first P V only after every softmax warp has read it - which p_ready
//     of that tile already implies, since a warp arrives there after its own write-out;
//   * the item itself (tile descriptor + head) travels from the producer to the other warps through a two-slot
//     ring in shared memory (`item_full`); a slot is rewritten only after q_empty of the item before, by which time
//     everybody has read the item two back.
struct WorkItem {
    AttnTile t;
    int32_t head;     // < 0: no more work
    int32_t pad[7];
};

__global__ void __launch_bounds__(THREADS, 2)
attention_persistent_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                            const __grid_constant__ CUtensorMap tmap_o, int n_q_heads, int n_kv_heads, const AttnTile* __restrict__ tiles, int n_tile_desc,
                            __half* __restrict__ out, int ld_out, float scale_log2e, uint32_t* __restrict__ work_ctr) {
    constexpr int NK = 3;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = smem + Q_BYTES;
    uint8_t* sV = sK + NK * TILE_BYTES;
    uint64_t* q_full = reinterpret_cast<uint64_t*>(sV + V_STAGES * TILE_BYTES);
    uint64_t* q_empty = q_full + 1;
    uint64_t* k_full = q_empty + 1;
    uint64_t* k_empty = k_full + NK;
    uint64_t* v_full = k_empty + NK;
    uint64_t* v_empty = v_full + V_STAGES;
    uint64_t* s_full = v_empty + V_STAGES;
    uint64_t* p_ready = s_full + 2;
    uint64_t* pv_done = p_ready + 1;
    uint64_t* item_full = pv_done + 1;                        // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(item_full + 2);
    uint32_t* epi_count = tmem_slot + 1;                      // softmax warps that have finished a write-out
    WorkItem* items = reinterpret_cast<WorkItem*>(tmem_slot + 4);   // [2] x 64 bytes (16-byte aligned)
    uint32_t dyn_bytes;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_bytes));
    if (reinterpret_cast<uint8_t*>(items + 2) > smem_raw + dyn_bytes) __trap();

    const int warp_idx = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t n_items = static_cast<uint32_t>(n_tile_desc) * static_cast<uint32_t>(n_q_heads);

    if (warp_idx == 0) {
        if (ptx::elect_one()) {
            ptx::prefetch_tensormap(&tmap_q);
            ptx::prefetch_tensormap(&tmap_kv);
            ptx::prefetch_tensormap(&tmap_o);
            ptx::mbar_init(q_full, 1);
            ptx::mbar_init(q_empty, 1);
            for (int i = 0; i < NK; ++i) {
                ptx::mbar_init(&k_full[i], 1);
                ptx::mbar_init(&k_empty[i], 1);
            }
            for (int i = 0; i < V_STAGES; ++i) {
                ptx::mbar_init(&v_full[i], 1);
                ptx::mbar_init(&v_empty[i], 1);
            }
            ptx::mbar_init(&s_full[0], 1);
            ptx::mbar_init(&s_full[1], 1);
            ptx::mbar_init(p_ready, 4);
            ptx::mbar_init(pv_done, 1);
            ptx::mbar_init(&item_full[0], 1);
            ptx::mbar_init(&item_full[1], 1);
            *epi_count = 0;
            ptx::fence_mbar_init();
        }
        __syncwarp();
        ptx::tmem_alloc<1>(tmem_slot, TMEM_COLS);
        ptx::tmem_relinquish<1>();
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    // key tiles of an item: segment 0 first (only up to the diagonal for a causal prefix tile), then segment 1
    auto tiles_of = [](const AttnTile& t, int& n0, int& n1) {
        const int seg0_keys = t.mode == kAttnPrefixCausal ? min(t.seg0_len, t.q_seq0 + t.n_q) : t.seg0_len;
        n0 = (seg0_keys + BKV - 1) / BKV;
        n1 = t.mode == kAttnSuffix ? (t.seg1_len + BKV - 1) / BKV : 0;
    };

    if (warp_idx == 0) {
        // ------------------------------------------------------------------ producer: work fetch + TMA
        if (ptx::elect_one()) {
            uint32_t gt = 0;                       // key tiles loaded so far
            for (uint32_t m = 0;; ++m) {           // m = items of this CTA so far
                const uint32_t it = atomicAdd(work_ctr, 1u);
                const bool have = it < n_items;
                WorkItem w;
                w.head = -1;
                if (have) {
                    w.t = tiles[it % static_cast<uint32_t>(n_tile_desc)];     // tile fastest, as the grid form launches
                    w.head = static_cast<int32_t>(it / static_cast<uint32_t>(n_tile_desc));
                    // A tile that sees no key at all (the planner never makes one) is not handed on: the other warps
                    // synchronise with each other through an item's key tiles, and it has none.  Its rows stay unwritten.
                    int e0, e1;
                    tiles_of(w.t, e0, e1);
                    if (e0 + e1 == 0) {
                        --m;
                        continue;
                    }
                }
                // the slot (and the Q buffer) of two items / one item ago are free once the last Q K^T of item m-1
                // has retired
                if (m > 0) ptx::mbar_wait(q_empty, (m - 1) & 1);
                WorkItem* slot = &items[m & 1];
                slot->t = w.t;
                slot->head = w.head;
                ptx::mbar_arrive(&item_full[m & 1]);          // release: the slot is written
                if (!have) break;
                const AttnTile& t = w.t;
                const int kvh = w.head / (n_q_heads / n_kv_heads);
                const int q_col = w.head * D;
                const int k_col = (n_q_heads + kvh) * D;
                const int v_col = (n_q_heads + n_kv_heads + kvh) * D;
                ptx::mbar_arrive_expect_tx(q_full, Q_BYTES);
                ptx::tma_load_2d(sQ, &tmap_q, q_full, q_col, t.q_row);
                ptx::tma_load_2d(sQ + Q_BYTES / 2, &tmap_q, q_full, q_col + 64, t.q_row);
                int n0, n1;
                tiles_of(t, n0, n1);
                for (int j = 0; j < n0 + n1; ++j, ++gt) {
                    const int ks = gt % NK, vs = gt % V_STAGES;
                    const int row = j < n0 ? t.seg0_row + j * BKV : t.seg1_row + (j - n0) * BKV;
                    ptx::mbar_wait(&k_empty[ks], ((gt / NK) & 1) ^ 1);
                    uint8_t* kb = sK + ks * TILE_BYTES;
                    ptx::mbar_arrive_expect_tx(&k_full[ks], TILE_BYTES);
                    ptx::tma_load_2d(kb, &tmap_kv, &k_full[ks], k_col, row);
                    ptx::tma_load_2d(kb + KSUB_BYTES, &tmap_kv, &k_full[ks], k_col + 64, row);
                    ptx::mbar_wait(&v_empty[vs], ((gt / V_STAGES) & 1) ^ 1);
                    uint8_t* vb = sV + vs * TILE_BYTES;
                    ptx::mbar_arrive_expect_tx(&v_full[vs], TILE_BYTES);
                    ptx::tma_load_2d(vb, &tmap_kv, &v_full[vs], v_col, row);
                    ptx::tma_load_2d(vb + KSUB_BYTES, &tmap_kv, &v_full[vs], v_col + 64, row);
                }
            }
            // the last CTA to run out of work re-arms the counter for the next launch
            __threadfence();
            if (atomicAdd(work_ctr + 1, 1u) == gridDim.x - 1) {
                work_ctr[0] = 0;
                work_ctr[1] = 0;
                __threadfence();
            }
        }
    } else if (warp_idx == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (ptx::elect_one()) {
            constexpr uint32_t idesc_qk = ptx::umma_idesc_f16_f32(BQ, BKV);
            constexpr uint32_t idesc_pv = ptx::umma_idesc_f16_f32(BQ, D) | ptx::kUmmaBMajorMN;
            const uint32_t q_addr = ptx::smem_u32(sQ);
            uint32_t gt = 0;                       // key tiles of earlier items
            for (uint32_t m = 0;; ++m) {
                ptx::mbar_wait(&item_full[m & 1], (m >> 1) & 1);
                const WorkItem* slot = &items[m & 1];
                if (slot->head < 0) break;
                const AttnTile t = slot->t;
                int n0, n1;
                tiles_of(t, n0, n1);
                const int n_tiles = n0 + n1;
                ptx::mbar_wait(q_full, m & 1);
                // S(j) = Q K(j)^T into S buffer (gt + j) & 1
                auto issue_qk = [&](int j) {
                    const uint32_t g = gt + j;
                    const int ks = g % NK;
                    const uint32_t st = ptx::smem_u32(sK + ks * TILE_BYTES);
                    ptx::mbar_wait(&k_full[ks], (g / NK) & 1);
                    ptx::tc_fence_after_sync();
#pragma unroll
                    for (int k = 0; k < D / 16; ++k) {
                        const uint64_t da = ptx::umma_smem_desc_sw128(q_addr + (k / 4) * (Q_BYTES / 2)) + 2 * (k % 4);
                        const uint64_t db = ptx::umma_smem_desc_sw128(st + (k / 4) * KSUB_BYTES) + 2 * (k % 4);
                        ptx::umma_f16<1>(tmem_base + S_COL + (g & 1) * BKV, da, db, idesc_qk, k != 0);
                    }
                    ptx::umma_commit(&s_full[g & 1]);
                    if (j == n_tiles - 1) ptx::umma_commit(q_empty);     // Q is not read again for this item
                };
                if (n_tiles > 0) issue_qk(0);
                else ptx::umma_commit(q_empty);
                for (int j = 0; j < n_tiles; ++j) {
                    if (j + 1 < n_tiles) issue_qk(j + 1);
                    const uint32_t g = gt + j;
                    const int ks = g % NK, vs = g % V_STAGES;
                    const uint32_t st = ptx::smem_u32(sK + ks * TILE_BYTES);
                    const uint32_t vt = ptx::smem_u32(sV + vs * TILE_BYTES);
                    ptx::mbar_wait(&v_full[vs], (g / V_STAGES) & 1);
                    ptx::mbar_wait(p_ready, g & 1);
                    ptx::tc_fence_after_sync();
#pragma unroll
                    for (int k = 0; k < BKV / 16; ++k) {
                        const uint64_t da = ptx::umma_smem_desc_sw128(st) + 2 * k;
                        const uint64_t db = ptx::umma_smem_desc_sw128_mn(vt + k * 16 * 128, KSUB_BYTES, 1024);
                        ptx::umma_f16<1>(tmem_base + O_COL, da, db, idesc_pv, (j | k) != 0);
                    }
                    // the last tile's K/P buffer is where the write-out is staged: the softmax warps hand it back
                    if (j + 1 < n_tiles) ptx::umma_commit(&k_empty[ks]);
                    ptx::umma_commit(&v_empty[vs]);
                    ptx::umma_commit(pv_done);
                }
                gt += n_tiles;
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax + write-out (warps 2..5)
        const int quarter = warp_idx & 3;
        const int i = quarter * 32 + lane;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
        uint32_t gt = 0;
        for (uint32_t m = 0;; ++m) {
            ptx::mbar_wait(&item_full[m & 1], (m >> 1) & 1);
            const WorkItem* slot = &items[m & 1];
            const int head = slot->head;
            if (head < 0) break;
            const AttnTile t = slot->t;
            int n0, n1;
            tiles_of(t, n0, n1);
            const int n_tiles = n0 + n1;
            const int rows_here = min(32, t.n_q - quarter * 32);      // valid rows of this warp (may be <= 0)
            if (n_tiles == 0) {                                       // no visible key at all: zeros, like O / l with l = 0
                if (lane < rows_here)
                    for (int c = 0; c < D; c += 8)
                        *reinterpret_cast<uint4*>(out + static_cast<size_t>(t.q_row + i) * ld_out + head * D + c) =
                            make_uint4(0, 0, 0, 0);
                continue;
            }
            const RowKeys rk = row_keys(t, i);
            float m_run = -INFINITY, l_run = 0.f;
            const float thr = RESCALE_THRESHOLD / scale_log2e;
            for (int j = 0; j < n_tiles; ++j) {
                const uint32_t g = gt + j;
                ptx::mbar_wait(&s_full[g & 1], (g >> 1) & 1);
                ptx::tc_fence_after_sync();
                const bool in0 = j < n0;
                softmax_key_tile(lane_addr + S_COL + (g & 1) * BKV, lane_addr + O_COL, in0 ? j * BKV : (j - n0) * BKV,
                                 in0 ? 0 : rk.lo1, in0 ? rk.lim0 - 1 : rk.hi1, scale_log2e, thr, m_run, l_run,
                                 sK + (g % NK) * TILE_BYTES + i * 128, i, j > 0 ? pv_done : nullptr, (g - 1) & 1, p_ready,
                                 lane);
            }
            // ---- write-out, staged in this warp's quarter of the K/P buffer of the item's last key tile, whose P V has
            // retired by then
            const uint32_t g_last = gt + n_tiles - 1;
            gt += n_tiles;
            ptx::mbar_wait(pv_done, g_last & 1);
            ptx::tc_fence_after_sync();
            const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
            write_out(lane_addr + O_COL, inv, sK + (g_last % NK) * TILE_BYTES + quarter * 4096, rows_here,
                      out + static_cast<size_t>(t.q_row + quarter * 32) * ld_out + head * D, ld_out, lane, &tmap_o,
                      head * D, t.q_row + quarter * 32);
            // O has been read (the next item's first P V may overwrite it once p_ready says so) and so has the staging
            // buffer: the fourth warp to get here returns it to the producer
            ptx::tc_fence_before_sync();
            if (lane == 0 && (atomicAdd(epi_count, 1u) & 3u) == 3u) ptx::mbar_arrive(&k_empty[g_last % NK]);
        }
    }

    ptx::tc_fence_before_sync();
    __syncwarp();
    __syncthreads();
    if (warp_idx == 0) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc<1>(tmem_base, TMEM_COLS);
    }
}

// Offset of a launch's dynamic shared memory from the next 1 KiB boundary (no static shared memory, as above).
__global__ void smem_base_probe_kernel(uint32_t* misalign) {
    extern __shared__ uint8_t probe_raw[];
    *misalign = (1024u - (ptx::smem_u32(probe_raw) & 1023u)) & 1023u;
}

std::mutex g_attn_mu;
int g_auto_k_stages[64] = {};   // per device: 0 = not decided yet
int g_smem_lead[64] = {};       // per device: bytes from the start of a dynamic allocation to the next 1 KiB boundary

template <int NK>
int launch_nk(const CUtensorMap& tq, const CUtensorMap& tkv, int n_q_heads, int n_kv_heads, const AttnTile* tiles,
              int n_tiles, __half* out, int ld_out, float scale_log2e, int smem_lead, cudaStream_t s) {
    attention_tcgen05_kernel<NK><<<dim3(n_tiles, n_q_heads), THREADS, smem_used(NK) + smem_lead, s>>>(
        tq, tkv, n_q_heads, n_kv_heads, tiles, out, ld_out, scale_log2e);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

}  // namespace

// Runs once per context, on the context's device.  Measures where dynamic shared memory starts relative to a 1 KiB
// boundary and decides how many K buffers a CTA gets there: three if two such CTAs still share an SM, else two.
int attention_sm100_init() {
    std::lock_guard<std::mutex> lock(g_attn_mu);
    constexpr int kMax = smem_used(MAX_K_STAGES) + 1023;
    FLS_CUDA(cudaFuncSetAttribute(attention_tcgen05_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    FLS_CUDA(cudaFuncSetAttribute(attention_tcgen05_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    FLS_CUDA(cudaFuncSetAttribute(attention_tcgen05_kernel<2>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    FLS_CUDA(cudaFuncSetAttribute(attention_tcgen05_kernel<3>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    FLS_CUDA(cudaFuncSetAttribute(attention_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
    FLS_CUDA(cudaFuncSetAttribute(attention_persistent_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    int dev = 0;
    FLS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(kErrInvalid, "attention: device index %d", dev);
    if (g_auto_k_stages[dev] == 0) {
        uint32_t* d_mis = nullptr;
        uint32_t mis = 1023;
        FLS_CUDA(cudaFuncSetAttribute(smem_base_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMax));
        FLS_CUDA(cudaMalloc(&d_mis, sizeof(uint32_t)));
        smem_base_probe_kernel<<<1, 1, smem_used(MAX_K_STAGES)>>>(d_mis);
        cudaError_t e = cudaMemcpy(&mis, d_mis, sizeof(uint32_t), cudaMemcpyDeviceToHost);
        cudaFree(d_mis);
        if (e != cudaSuccess) return fail(kErrCuda, "attention: shared memory probe failed: %s", cudaGetErrorString(e));
        // Two CTAs share an SM when each one's allocation (dynamic + the driver's reserved bytes) fits half of the SM's
        // shared memory.  (cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for this kernel at every size,
        // although `ncu` shows two resident CTAs - launch__occupancy_limit_shared_mem 2 at 114 944 bytes - so the
        // device attributes decide: profiles/r02_occ_probe.txt.)
        int per_sm = 0, reserved = 0;
        FLS_CUDA(cudaDeviceGetAttribute(&per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
        FLS_CUDA(cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev));
        const int need = smem_used(3) + static_cast<int>(mis);
        g_smem_lead[dev] = static_cast<int>(mis);
        g_auto_k_stages[dev] = 2 * (need + reserved) <= per_sm ? 3 : 2;
        if (getenv("FLS_DEBUG_ATTENTION"))
            fprintf(stderr, "[fls] attention on device %d: dynamic shared memory starts %u bytes before a 1 KiB boundary; "
                    "3 K buffers need %d + %d bytes per CTA, the SM has %d -> %d K buffers\n", dev, mis, need, reserved,
                    per_sm, g_auto_k_stages[dev]);
    }
    return kOk;
}

int attention_sm100_k_stages() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    return g_auto_k_stages[dev];
}

int attention_sm100_launch(const __half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads, const AttnTile* tiles,
                           int n_tiles, __half* out, int ld_out, cudaStream_t s, int k_stages) {
    if (n_tiles <= 0) return kOk;
    if (ld_out % 8) return fail(kErrInvalid, "attention: ld_out must be a multiple of 8");
    if (n_kv_heads <= 0 || n_q_heads % n_kv_heads)
        return fail(kErrInvalid, "attention: heads %d / kv heads %d", n_q_heads, n_kv_heads);
    int dev = 0;
    FLS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || g_auto_k_stages[dev] == 0) return fail(kErrState, "attention_sm100_init() has not run");
    if (k_stages == 0) k_stages = g_auto_k_stages[dev];
    const int lead = g_smem_lead[dev];
    if (k_stages != 2 && k_stages != 3) return fail(kErrState, "attention: %d K buffers (init has not run?)", k_stages);
    CUtensorMap tq, tkv;
    const int cols = (n_q_heads + 2 * n_kv_heads) * D;
    FLS_TRY(encode_tmap_2d(&tq, qkv, n_rows, cols, ld, BQ));
    FLS_TRY(encode_tmap_2d(&tkv, qkv, n_rows, cols, ld, BKV));
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    return k_stages == 3 ? launch_nk<3>(tq, tkv, n_q_heads, n_kv_heads, tiles, n_tiles, out, ld_out, scale_log2e, lead, s)
                         : launch_nk<2>(tq, tkv, n_q_heads, n_kv_heads, tiles, n_tiles, out, ld_out, scale_log2e, lead, s);
}

// Persistent form (see attention_persistent_kernel).  `work_ctr`: two zeroed uint32 in device memory that belong to the
// caller's context (the kernel leaves them zeroed); launches that share them must not overlap.  Falls back to the grid
// form where three K buffers do not leave room for two CTAs per SM.
int attention_sm100_launch_persistent(const __half* qkv, int ld, int n_rows, int n_q_heads, int n_kv_heads,
                                      const AttnTile* tiles, int n_tiles, __half* out, int ld_out, cudaStream_t s,
                                      uint32_t* work_ctr, int num_sms) {
    if (n_tiles <= 0) return kOk;
    int dev = 0;
    FLS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || g_auto_k_stages[dev] == 0) return fail(kErrState, "attention_sm100_init() has not run");
    if (!work_ctr || g_auto_k_stages[dev] != 3 || PERSIST_SMEM_USED + g_smem_lead[dev] > smem_used(3) + 768)
        return attention_sm100_launch(qkv, ld, n_rows, n_q_heads, n_kv_heads, tiles, n_tiles, out, ld_out, s, 0);
    if (ld_out % 8) return fail(kErrInvalid, "attention: ld_out must be a multiple of 8");
    if (n_kv_heads <= 0 || n_q_heads % n_kv_heads)
        return fail(kErrInvalid, "attention: heads %d / kv heads %d", n_q_heads, n_kv_heads);
    CUtensorMap tq, tkv;
    const int cols = (n_q_heads + 2 * n_kv_heads) * D;
    FLS_TRY(encode_tmap_2d(&tq, qkv, n_rows, cols, ld, BQ));
    FLS_TRY(encode_tmap_2d(&tkv, qkv, n_rows, cols, ld, BKV));
    CUtensorMap tout;                                       // [n_rows, heads x 128] output, pieces of 32 rows x 64 columns
    FLS_TRY(encode_tmap_2d(&tout, out, n_rows, n_q_heads * D, ld_out, 32));
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    const long long n_items = static_cast<long long>(n_tiles) * n_q_heads;
    const int grid = static_cast<int>(std::min<long long>(n_items, 2LL * (num_sms > 0 ? num_sms : 148)));
    attention_persistent_kernel<<<grid, THREADS, PERSIST_SMEM_USED + g_smem_lead[dev], s>>>(
        tq, tkv, tout, n_q_heads, n_kv_heads, tiles, n_tiles, out, ld_out, scale_log2e, work_ctr);
    FLS_CUDA(cudaGetLastError());
    return kOk;
}

}  // namespace fls
