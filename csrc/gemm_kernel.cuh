// Persistent, warp-specialised bf16 GEMM for sm_100a: TMA -> 4-stage smem ring -> tcgen05.mma (128x256x16,
// fp32 accumulators in TMEM, double-buffered) -> tcgen05.ld epilogue.  One CTA per SM.
//
//   D[M,N] = sum_k A(m,k) * B(n,k)
//
// Each operand is either K-major (row-major [MN, K], `x @ W^T` style) or MN-major (row-major [K, MN], the
// transposed-operand forms the backward pass needs), selected at RUN time: only the TMA box walk, the UMMA
// smem-descriptor strides and two bits of the instruction descriptor depend on it.  The epilogue is a compile-
// time policy class so the FFN's elementwise work (ReLU/GELU, their backward masks, residual add, the SGD
// update) and the peer-memory collectives (fused.cu) run on the accumulator tile before it ever reaches HBM.
//
// Covers the reference's seven GEMM call sites (train_ffns.py:42, :45) - see DESIGN.md for the mapping.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int BM = 128;            // UMMA M (TMEM lanes)
constexpr int BN = 256;            // UMMA N (TMEM fp32 columns per accumulator stage)
constexpr int BK = 64;             // one 128-byte swizzle atom of bf16 along K per stage
constexpr int UMMA_K = 16;
constexpr int STAGES = 4;
constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KiB
constexpr int B_STAGE_BYTES = BN * BK * 2;   // 32 KiB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int EPI_WARPS = 8;       // two warps per TMEM lane quarter, each takes half of the 256 columns
constexpr int GEMM_THREADS = 64 + 32 * EPI_WARPS;  // warp0: TMA, warp1: MMA + TMEM alloc, warps 2..9: epilogue
constexpr int EPI_STAGE_BYTES = 32 * 128;          // per-epilogue-warp transpose buffer (fused push epilogues)
constexpr int GEMM_SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int GEMM_SMEM_BYTES_STAGED = GEMM_SMEM_BYTES + EPI_WARPS * EPI_STAGE_BYTES + 256;

enum Major : int { MAJOR_K = 0, MAJOR_MN = 1 };

struct GemmCore {
    int M, N, K;
    int a_major, b_major;
    int tiles_m, tiles_n, group_m;
    int tile_offset;   // rotation of the tile order (fused kernels stagger ranks so pushes spread over owners)
    unsigned int* err;
};

// tile index -> (m_blk, n_blk) with grouped rasterisation so the ~148 tiles in flight share A/B panels in L2
__device__ __forceinline__ void tile_coords(const GemmCore& g, int tile, int& m_blk, int& n_blk) {
    const int per_group = g.group_m * g.tiles_n;
    const int group = tile / per_group;
    const int first_m = group * g.group_m;
    const int gsize = min(g.tiles_m - first_m, g.group_m);
    const int in_group = tile - group * per_group;
    m_blk = first_m + in_group % gsize;
    n_blk = in_group / gsize;
}

struct SmemView {
    uint8_t* a0;           // stage s of A at a0 + s * A_STAGE_BYTES
    uint8_t* b0;           // stage s of B at b0 + s * B_STAGE_BYTES
    __device__ __forceinline__ uint8_t* a(int s) const { return a0 + s * A_STAGE_BYTES; }
    __device__ __forceinline__ uint8_t* b(int s) const { return b0 + s * B_STAGE_BYTES; }
    uint64_t* full;        // [STAGES]
    uint64_t* empty;       // [STAGES]
    uint64_t* tmem_full;   // [2]
    uint64_t* tmem_empty;  // [2]
    uint32_t* tmem_base;
    uint8_t* epi_stage;    // EPI_WARPS x EPI_STAGE_BYTES, only when the launch asked for GEMM_SMEM_BYTES_STAGED
};

__device__ __forceinline__ SmemView carve_smem(uint8_t* raw) {
    SmemView v;
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    v.a0 = base;
    v.b0 = base + STAGES * A_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + STAGES * STAGE_BYTES);
    v.full = bars;
    v.empty = bars + STAGES;
    v.tmem_full = bars + 2 * STAGES;
    v.tmem_empty = bars + 2 * STAGES + 2;
    v.tmem_base = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    v.epi_stage = base + STAGES * STAGE_BYTES + 256;
    return v;
}

// ---------------------------------------------------------------------------------------------
// The kernel.  `Epi` supplies:
//   struct Params;                                                     (trivially copyable)
//   static constexpr int SMEM_BYTES;                                   (GEMM_SMEM_BYTES or GEMM_SMEM_BYTES_STAGED)
//   struct State;                                                      (per-epilogue-warp registers)
//   static __device__ void init(const Params&, const GemmCore&, const EpiCtx&, State&)
//   static __device__ void process_tile(const Params&, const GemmCore&, const EpiCtx&, uint32_t taddr, State&)
//        one epilogue warp's share of a finished accumulator tile: rows ctx.m0 + q*32 + lane (TMEM lane quarter
//        ctx.q), columns [ctx.n0 + ctx.half*128, +128); taddr already points at (quarter, stage, half)
//   static __device__ void after_tile(const Params&, const GemmCore&, const EpiCtx&, State&)
//        runs after the accumulator stage has been handed back to the MMA warp (signals, opportunistic work)
//   static __device__ void tail(const Params&, const GemmCore&, const EpiCtx&, State&)   (after the last tile)
// ---------------------------------------------------------------------------------------------
struct EpiCtx {
    int tile;        // linear tile index in the scheduler's order
    int m_blk, n_blk;
    int m0, n0;      // first row / column of the tile
    int q;           // TMEM lane quarter (0..3) -> rows m0 + q*32 + lane
    int half;        // column half (0..1)
    int ewarp;       // epilogue warp index 0..EPI_WARPS-1
    int lane;
    uint8_t* stage;  // this warp's EPI_STAGE_BYTES transpose buffer (staged launches only)
    uint32_t* scratch;  // 256 B shared by the epilogue warps, after the stage buffers (staged launches only)
};

template <class Epi>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmCore g, const typename Epi::Params ep) {
    extern __shared__ uint8_t smem_raw[];
    SmemView sm = carve_smem(smem_raw);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_tiles = g.tiles_m * g.tiles_n;
    const int num_kb = (g.K + BK - 1) / BK;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            mbar_init(&sm.tmem_full[s], 1);
            mbar_init(&sm.tmem_empty[s], EPI_WARPS);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_512(sm.tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_base;

    if (threadIdx.x == 0) {
        // ===================== TMA producer (one thread) =====================
        int s = 0;
        uint32_t phase = 0;
        for (int it = blockIdx.x; it < num_tiles; it += gridDim.x) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            int m_blk, n_blk;
            tile_coords(g, tile, m_blk, n_blk);
            const int m0 = m_blk * BM, n0 = n_blk * BN;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&sm.empty[s], phase ^ 1, g.err, 0x100 + s);
                mbar_arrive_expect_tx(&sm.full[s], STAGE_BYTES);
                const int k0 = kb * BK;
                if (g.a_major == MAJOR_K) {
                    tma_load_2d(sm.a(s), &tmA, &sm.full[s], k0, m0);          // box {64 k, 128 m}
                } else {
#pragma unroll
                    for (int j = 0; j < BM / 64; ++j)                          // box {64 m, 64 k} per atom
                        tma_load_2d(sm.a(s) + j * (BK * 128), &tmA, &sm.full[s], m0 + j * 64, k0);
                }
                if (g.b_major == MAJOR_K) {
                    tma_load_2d(sm.b(s), &tmB, &sm.full[s], k0, n0);          // box {64 k, 256 n}
                } else {
#pragma unroll
                    for (int j = 0; j < BN / 64; ++j)
                        tma_load_2d(sm.b(s) + j * (BK * 128), &tmB, &sm.full[s], n0 + j * 64, k0);
                }
                if (++s == STAGES) { s = 0; phase ^= 1; }
            }
        }
    } else if (threadIdx.x == 32) {
        // ===================== MMA issuer (one thread) =====================
        const uint32_t idesc = (1u << 4)                       // D = f32
                               | (1u << 7) | (1u << 10)        // A, B = bf16
                               | ((uint32_t)g.a_major << 15) | ((uint32_t)g.b_major << 16)
                               | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
        // K-major : 8-row x 128B atoms stacked along MN every 1024 B; UMMA_K=16 is 32 B inside the atom.
        // MN-major: 8-k-row x 128B atoms; next k-group 1024 B, next 64-wide MN atom BK*128 B; UMMA_K = 2 groups.
        const uint32_t a_lbo = g.a_major == MAJOR_K ? 16u : (uint32_t)(BK * 128);
        const uint32_t b_lbo = g.b_major == MAJOR_K ? 16u : (uint32_t)(BK * 128);
        const uint32_t a_kadv = g.a_major == MAJOR_K ? 32u : 2048u;
        const uint32_t b_kadv = g.b_major == MAJOR_K ? 32u : 2048u;
        int s = 0;
        uint32_t phase = 0;
        int as = 0;
        uint32_t aphase = 0;
        for (int it = blockIdx.x; it < num_tiles; it += gridDim.x) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            mbar_wait(&sm.tmem_empty[as], aphase ^ 1, g.err, 0x200 + as);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&sm.full[s], phase, g.err, 0x300 + s);
                tc_fence_after();
                const uint32_t a_base = smem_u32(sm.a(s));
                const uint32_t b_base = smem_u32(sm.b(s));
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) {
                    const uint64_t adesc = make_smem_desc_sw128(a_base + k * a_kadv, a_lbo, 1024);
                    const uint64_t bdesc = make_smem_desc_sw128(b_base + k * b_kadv, b_lbo, 1024);
                    umma_bf16(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                }
                umma_commit(&sm.empty[s]);                       // smem slot free once these MMAs retire
                if (kb == num_kb - 1) umma_commit(&sm.tmem_full[as]);
                if (++s == STAGES) { s = 0; phase ^= 1; }
            }
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    } else if (warp >= 2) {
        // ===================== epilogue (8 warps: TMEM lane quarter x column half) =====================
        EpiCtx ctx;
        ctx.ewarp = warp - 2;
        ctx.q = warp & 3;                                        // tcgen05.ld lane window = warp_id % 4
        ctx.half = (warp - 2) >> 2;
        ctx.lane = lane;
        ctx.stage = sm.epi_stage + ctx.ewarp * EPI_STAGE_BYTES;
        ctx.scratch = reinterpret_cast<uint32_t*>(sm.epi_stage + EPI_WARPS * EPI_STAGE_BYTES);
        ctx.tile = -1;
        ctx.m_blk = ctx.n_blk = ctx.m0 = ctx.n0 = 0;
        typename Epi::State est;
        Epi::init(ep, g, ctx, est);
        int as = 0;
        uint32_t aphase = 0;
        for (int it = blockIdx.x; it < num_tiles; it += gridDim.x) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            ctx.tile = tile;
            tile_coords(g, tile, ctx.m_blk, ctx.n_blk);
            ctx.m0 = ctx.m_blk * BM;
            ctx.n0 = ctx.n_blk * BN;
            mbar_wait(&sm.tmem_full[as], aphase, g.err, 0x400 + as);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(ctx.q * 32) << 16) + (uint32_t)(as * BN + ctx.half * (BN / 2));
            Epi::process_tile(ep, g, ctx, taddr, est);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.tmem_empty[as]);
            Epi::after_tile(ep, g, ctx, est);
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
        Epi::tail(ep, g, ctx, est);
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc_512(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// CTA-pair variant: two CTAs of a cluster (one TPC) cooperate on a 256x256 tile with tcgen05.mma.cta_group::2.
// Each CTA stages ITS 128 rows of A and ITS 128 rows of B (half of what the 1-CTA kernel stages for the same math),
// so shared-memory write+read traffic per SM drops by a third and the ring gets 6 stages instead of 4.  The leader
// CTA's MMA thread issues for both; accumulators land in each CTA's own TMEM (rows [rank*128, +128) of the pair
// tile), so the epilogue policies are unchanged.
// ---------------------------------------------------------------------------------------------
constexpr int P_STAGES = 6;
constexpr int P_A_STAGE_BYTES = 128 * BK * 2;   // 16 KiB: this CTA's 128 rows of A
constexpr int P_B_STAGE_BYTES = 128 * BK * 2;   // 16 KiB: this CTA's 128 of the tile's 256 B rows
constexpr int P_STAGE_BYTES = P_A_STAGE_BYTES + P_B_STAGE_BYTES;
static_assert(P_STAGES * P_STAGE_BYTES == STAGES * STAGE_BYTES, "both variants use the same smem carve-up");

template <class Epi>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmCore g, const typename Epi::Params ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = base;
    uint8_t* sB = base + P_STAGES * P_A_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + P_STAGES * P_STAGE_BYTES);
    uint64_t* full = bars;                       // [P_STAGES]  (only the leader's are waited on)
    uint64_t* empty = bars + P_STAGES;           // [P_STAGES]
    uint64_t* tmem_full = bars + 2 * P_STAGES;   // [2]
    uint64_t* tmem_empty = bars + 2 * P_STAGES + 2;   // [2]  (only the leader's are waited on)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * P_STAGES + 4);
    uint8_t* epi_stage = base + P_STAGES * P_STAGE_BYTES + 256;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();     // 0 = leader
    const int pair = blockIdx.x >> 1;
    const int num_pairs = gridDim.x >> 1;
    const int pair_tiles_m = (g.tiles_m + 1) >> 1;
    const int num_tiles = pair_tiles_m * g.tiles_n;
    const int num_kb = (g.K + BK - 1) / BK;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
#pragma unroll
        for (int s = 0; s < P_STAGES; ++s) {
            mbar_init(&full[s], 1);              // the leader's expect_tx arrival; BOTH CTAs' bytes are credited here
            mbar_init(&empty[s], 1);             // one multicast commit
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 2 * EPI_WARPS);   // epilogue warps of BOTH CTAs
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_512_pair(tmem_slot);
    tc_fence_before();
    cluster_sync_all();                          // barriers + TMEM of both CTAs are ready before any cross-CTA traffic
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // pair-tile index -> (pair row, n block), grouped like tile_coords (8 pair rows = 16 m blocks per group)
    auto pair_coords = [&](int tile, int& pm, int& n_blk) {
        const int gm = g.group_m >> 1 > 0 ? g.group_m >> 1 : 1;
        const int per_group = gm * g.tiles_n;
        const int group = tile / per_group;
        const int first = group * gm;
        const int gsize = min(pair_tiles_m - first, gm);
        const int in_group = tile - group * per_group;
        pm = first + in_group % gsize;
        n_blk = in_group / gsize;
    };

    if (threadIdx.x == 0) {
        // ===================== TMA producer (one thread in EACH CTA) =====================
        int s = 0;
        uint32_t phase = 0;
        for (int it = pair; it < num_tiles; it += num_pairs) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            int pm, n_blk;
            pair_coords(tile, pm, n_blk);
            const int m0 = (pm * 2 + (int)rank) * BM;            // my 128 rows of A
            const int n0 = n_blk * BN + (int)rank * 128;         // my 128 rows of B
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&empty[s], phase ^ 1, g.err, 0x100 + s);
                // The peer's bytes may land before this arrival: the transaction count just goes negative for a
                // moment, and the phase cannot complete until the one pending arrival (this one) has happened.
                if (rank == 0) mbar_arrive_expect_tx(&full[s], 2 * P_STAGE_BYTES);
                const int k0 = kb * BK;
                uint8_t* a_dst = sA + s * P_A_STAGE_BYTES;
                uint8_t* b_dst = sB + s * P_B_STAGE_BYTES;
                if (g.a_major == MAJOR_K) {
                    tma_load_2d_pair(a_dst, &tmA, &full[s], k0, m0);                       // box {64 k, 128 m}
                } else {
                    tma_load_2d_pair(a_dst, &tmA, &full[s], m0, k0);                       // 2 x box {64 m, 64 k}
                    tma_load_2d_pair(a_dst + BK * 128, &tmA, &full[s], m0 + 64, k0);
                }
                if (g.b_major == MAJOR_K) {
                    tma_load_2d_pair(b_dst, &tmB, &full[s], k0, n0);                       // box {64 k, 128 n}
                } else {
                    tma_load_2d_pair(b_dst, &tmB, &full[s], n0, k0);
                    tma_load_2d_pair(b_dst + BK * 128, &tmB, &full[s], n0 + 64, k0);
                }
                if (++s == P_STAGES) { s = 0; phase ^= 1; }
            }
        }
    } else if (threadIdx.x == 32 && rank == 0) {
        // ===================== MMA issuer (one thread, leader CTA only) =====================
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10)
                               | ((uint32_t)g.a_major << 15) | ((uint32_t)g.b_major << 16)
                               | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);   // M = 256 across the pair
        const uint32_t a_lbo = g.a_major == MAJOR_K ? 16u : (uint32_t)(BK * 128);
        const uint32_t b_lbo = g.b_major == MAJOR_K ? 16u : (uint32_t)(BK * 128);
        const uint32_t a_kadv = g.a_major == MAJOR_K ? 32u : 2048u;
        const uint32_t b_kadv = g.b_major == MAJOR_K ? 32u : 2048u;
        int s = 0;
        uint32_t phase = 0;
        int as = 0;
        uint32_t aphase = 0;
        for (int it = pair; it < num_tiles; it += num_pairs) {
            mbar_wait(&tmem_empty[as], aphase ^ 1, g.err, 0x200 + as);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&full[s], phase, g.err, 0x300 + s);
                tc_fence_after();
                const uint32_t a_base = smem_u32(sA + s * P_A_STAGE_BYTES);
                const uint32_t b_base = smem_u32(sB + s * P_B_STAGE_BYTES);
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) {
                    const uint64_t adesc = make_smem_desc_sw128(a_base + k * a_kadv, a_lbo, 1024);
                    const uint64_t bdesc = make_smem_desc_sw128(b_base + k * b_kadv, b_lbo, 1024);
                    umma_bf16_pair(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                }
                umma_commit_pair(&empty[s]);                         // frees the slot in BOTH CTAs
                if (kb == num_kb - 1) umma_commit_pair(&tmem_full[as]);
                if (++s == P_STAGES) { s = 0; phase ^= 1; }
            }
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    } else if (warp >= 2) {
        // ===================== epilogue (8 warps per CTA) =====================
        EpiCtx ctx;
        ctx.ewarp = warp - 2;
        ctx.q = warp & 3;
        ctx.half = (warp - 2) >> 2;
        ctx.lane = lane;
        ctx.stage = epi_stage + ctx.ewarp * EPI_STAGE_BYTES;
        ctx.scratch = reinterpret_cast<uint32_t*>(epi_stage + EPI_WARPS * EPI_STAGE_BYTES);
        ctx.tile = -1;
        ctx.m_blk = ctx.n_blk = ctx.m0 = ctx.n0 = 0;
        typename Epi::State est;
        Epi::init(ep, g, ctx, est);
        int as = 0;
        uint32_t aphase = 0;
        for (int it = pair; it < num_tiles; it += num_pairs) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            int pm;
            pair_coords(tile, pm, ctx.n_blk);
            ctx.tile = tile;
            ctx.m_blk = pm * 2 + (int)rank;
            ctx.m0 = ctx.m_blk * BM;
            ctx.n0 = ctx.n_blk * BN;
            mbar_wait(&tmem_full[as], aphase, g.err, 0x400 + as);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(ctx.q * 32) << 16) + (uint32_t)(as * BN + ctx.half * (BN / 2));
            const bool real_tile = ctx.m_blk < g.tiles_m;           // odd tiles_m: the last pair's second half is empty
            if (real_tile) Epi::process_tile(ep, g, ctx, taddr, est);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (rank == 0) mbar_arrive(&tmem_empty[as]);
                else mbar_arrive_remote(&tmem_empty[as], 0);
            }
            if (real_tile) Epi::after_tile(ep, g, ctx, est);
            else Epi::after_skipped_tile(ep, g, ctx, est);
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
        Epi::tail(ep, g, ctx, est);
    }

    tc_fence_before();
    cluster_sync_all();                          // neither CTA may retire while its pair still uses its smem / barriers
    tc_fence_after();
    if (warp == 1) tmem_dealloc_512_pair(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// Local epilogues
// ---------------------------------------------------------------------------------------------
enum EpiKind : int {
    EPI_STORE = 0,      // C = acc (+bias)
    EPI_RELU = 1,       // C = relu(acc + bias)                                   (train_ffns.py:47-48 fused)
    EPI_GELU = 2,       // C = gelu(acc + bias), C2 = acc + bias (pre-activation kept for backward)
    EPI_RELU_BWD = 3,   // C = acc * (aux > 0)        aux = relu output or pre-activation  (train_ffns.py:50-52)
    EPI_GELU_BWD = 4,   // C = acc * gelu'(aux)       aux = pre-activation
    EPI_ADD = 5,        // C = acc + aux (+bias)      residual / gradient accumulation
    EPI_SGD = 6,        // C = C - alpha * acc        in-place optimizer step on the weight tile (:114,:172)
    EPI_STORE_F32 = 7,  // C(fp32) = alpha * acc
    EPI_COUNT = 8,
};

struct LocalEpiParams {
    void* C;
    long long ldc;
    const __nv_bfloat16* aux;
    long long ldaux;
    __nv_bfloat16* C2;
    long long ldc2;
    const float* bias;
    float alpha;
};

__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad_f(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
    return cdf + x * pdf;
}

template <int KIND>
struct LocalEpi {
    using Params = LocalEpiParams;
    static constexpr int SMEM_BYTES = GEMM_SMEM_BYTES;
    static constexpr bool NEEDS_AUX = (KIND == EPI_RELU_BWD || KIND == EPI_GELU_BWD || KIND == EPI_ADD);
    struct State {};
    static __device__ __forceinline__ void init(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_skipped_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void tail(const Params&, const GemmCore&, const EpiCtx&, State&) {}

    static __device__ __forceinline__ void load_aux(const Params& p, const GemmCore& g, int row, int col0, uint4 (&a)[4]) {
        if constexpr (NEEDS_AUX) {
            if (row >= g.M) return;
#pragma unroll
            for (int v = 0; v < 4; ++v)
                if (col0 + v * 8 < g.N) a[v] = ld_nc_u4(p.aux + (long long)row * p.ldaux + col0 + v * 8);
        }
    }

    static __device__ __forceinline__ void process_tile(const Params& p, const GemmCore& g, const EpiCtx& ctx, uint32_t taddr, State&) {
        const int row = ctx.m0 + ctx.q * 32 + ctx.lane;
        const int colbase = ctx.n0 + ctx.half * (BN / 2);
        uint4 aux_next[4] = {};
        load_aux(p, g, row, colbase, aux_next);                 // overlaps the wait-free first TMEM load
#pragma unroll 1
        for (int c = 0; c < BN / 2 / 32; ++c) {
            uint32_t acc[32];
            uint4 aux_cur[4];
#pragma unroll
            for (int v = 0; v < 4; ++v) aux_cur[v] = aux_next[v];
            if (c + 1 < BN / 2 / 32) load_aux(p, g, row, colbase + (c + 1) * 32, aux_next);   // prefetch next chunk
            tmem_ld_32x32(taddr + c * 32, acc);
            tmem_ld_wait();
            chunk(p, g, row, colbase + c * 32, acc, aux_cur);
        }
    }

    static __device__ __forceinline__ void chunk(const Params& p, const GemmCore& g, int row, int col0,
                                                 const uint32_t (&acc)[32], const uint4 (&auxv)[4]) {
        if (row >= g.M) return;
#pragma unroll
        for (int v = 0; v < 4; ++v) {                      // 4 groups of 8 columns = one 16-byte bf16 vector
            const int col = col0 + v * 8;
            if (col >= g.N) break;                         // N % 8 == 0 is a launch precondition
            float f[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(acc[v * 8 + j]);
            if constexpr (KIND == EPI_STORE || KIND == EPI_RELU || KIND == EPI_GELU || KIND == EPI_ADD) {
                if (p.bias != nullptr) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] += __ldg(p.bias + col + j);
                }
            }
            if constexpr (KIND == EPI_STORE_F32) {
                float* out = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col;
                float4 o0 = make_float4(p.alpha * f[0], p.alpha * f[1], p.alpha * f[2], p.alpha * f[3]);
                float4 o1 = make_float4(p.alpha * f[4], p.alpha * f[5], p.alpha * f[6], p.alpha * f[7]);
                *reinterpret_cast<float4*>(out) = o0;
                *reinterpret_cast<float4*>(out + 4) = o1;
                continue;
            }
            __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col;
            if constexpr (KIND == EPI_RELU) {
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = f[j] <= 0.0f ? 0.0f : f[j];   // NaN stays NaN, like torch.where(x<=0,0,x)
            } else if constexpr (KIND == EPI_GELU) {
                uint4 pre;
                pre.x = pack_bf16x2(f[0], f[1]); pre.y = pack_bf16x2(f[2], f[3]);
                pre.z = pack_bf16x2(f[4], f[5]); pre.w = pack_bf16x2(f[6], f[7]);
                *reinterpret_cast<uint4*>(p.C2 + (long long)row * p.ldc2 + col) = pre;
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
            } else if constexpr (KIND == EPI_RELU_BWD || KIND == EPI_GELU_BWD || KIND == EPI_ADD) {
                const uint4 a = auxv[v];
                const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float lo = bf16_lo(aw[j]), hi = bf16_hi(aw[j]);
                    if constexpr (KIND == EPI_RELU_BWD) {
                        f[2 * j] = lo <= 0.0f ? 0.0f : f[2 * j];            // masked_fill_(x <= 0, 0)
                        f[2 * j + 1] = hi <= 0.0f ? 0.0f : f[2 * j + 1];
                    } else if constexpr (KIND == EPI_GELU_BWD) {
                        f[2 * j] *= gelu_grad_f(lo);
                        f[2 * j + 1] *= gelu_grad_f(hi);
                    } else {
                        f[2 * j] += lo;
                        f[2 * j + 1] += hi;
                    }
                }
            } else if constexpr (KIND == EPI_SGD) {
                const uint4 w = *reinterpret_cast<const uint4*>(out);
                const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    f[2 * j] = bf16_lo(ww[j]) - p.alpha * f[2 * j];
                    f[2 * j + 1] = bf16_hi(ww[j]) - p.alpha * f[2 * j + 1];
                }
            }
            uint4 o;
            o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
            o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
            *reinterpret_cast<uint4*>(out) = o;
        }
    }
};

// ---------------------------------------------------------------------------------------------
// host launch helper shared by gemm.cu and fused.cu
// ---------------------------------------------------------------------------------------------
struct GemmOperands {
    const void* A; long long lda; int a_major;
    const void* B; long long ldb; int b_major;
    int M, N, K;
};

int validate_operands(const GemmOperands& o);
int make_core_and_maps(const GemmOperands& o, GemmCore* core, CUtensorMap* tmA, CUtensorMap* tmB, bool pair);

bool use_pair_kernel();   // runtime switch (B200_GEMM_PAIR=0 selects the 1-CTA kernel)

template <class Epi>
int launch_gemm(const GemmOperands& o, const typename Epi::Params& ep, int max_ctas, cudaStream_t stream,
                int tile_offset_num = 0, int tile_offset_den = 1) {
    GemmCore core;
    alignas(64) CUtensorMap tmA, tmB;
    const bool pair = use_pair_kernel();
    int rc = make_core_and_maps(o, &core, &tmA, &tmB, pair);
    if (rc != 0) return rc;
    static unsigned long long attr_mask = 0;   // per template instantiation, one bit per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (!(attr_mask >> (dev & 63) & 1ull)) {
        rc = check_cuda(cudaFuncSetAttribute(gemm_tcgen05_kernel<Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Epi::SMEM_BYTES), "cudaFuncSetAttribute(smem)");
        if (rc != 0) return rc;
        rc = check_cuda(cudaFuncSetAttribute(gemm_tcgen05_pair_kernel<Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Epi::SMEM_BYTES), "cudaFuncSetAttribute(smem, pair)");
        if (rc != 0) return rc;
        attr_mask |= 1ull << (dev & 63);
    }
    int cap = max_ctas > 0 ? max_ctas : num_sms();
    if (!pair) {
        const int tiles = core.tiles_m * core.tiles_n;
        core.tile_offset = (int)((long long)tiles * tile_offset_num / tile_offset_den);
        int grid = tiles < cap ? tiles : cap;
        gemm_tcgen05_kernel<Epi><<<grid, GEMM_THREADS, Epi::SMEM_BYTES, stream>>>(tmA, tmB, core, ep);
        return check_cuda(cudaGetLastError(), "gemm launch");
    }
    const int pair_tiles = ((core.tiles_m + 1) / 2) * core.tiles_n;
    core.tile_offset = (int)((long long)pair_tiles * tile_offset_num / tile_offset_den);
    int pairs = cap / 2;
    if (pairs < 1) pairs = 1;
    if (pairs > pair_tiles) pairs = pair_tiles;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = Epi::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return check_cuda(cudaLaunchKernelEx(&cfg, gemm_tcgen05_pair_kernel<Epi>, tmA, tmB, core, ep), "gemm pair launch");
}

}  // namespace b200
