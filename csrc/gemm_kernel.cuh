// Persistent, warp-specialised bf16 GEMMs for sm_100a: TMA -> 4-stage smem ring -> tcgen05.mma with fp32 accumulators
// in TMEM -> tcgen05.ld epilogue.  One CTA per SM.  Two kernels share the epilogue policies:
//   gemm_tcgen05_pair_kernel  (default) clusters of 2 CTAs, cta_group::2, 512x256 tile per pair = 256x256 per CTA: two
//                             128x256 accumulators (all 512 TMEM columns) share every B stage
//   gemm_tcgen05_kernel       1 CTA, 128x256 tile, TWO accumulator stages: the epilogue of tile i runs under the MMAs of
//                             tile i+1 (used when the epilogue is long: link-bound fused collectives)
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
    // This CTA's index and the size of the grid AS THE KERNEL BODY SEES THEM.  A normal launch fills them from blockIdx /
    // gridDim; the single-launch multi-rank emulation (emu kernels below) gives every emulated rank its own slice of one
    // cooperative grid, so the bodies and the epilogue policies must never read blockIdx.x / gridDim.x themselves.
    int cta, ncta;
    // Split-operand GEMMs (fp32 emulation, Epi::PRODUCTS > 1): each operand is `3` bf16 planes stacked along its storage
    // rows, plane p starting a_plane / b_plane rows after plane p-1 (0 for ordinary single-plane operands).
    int a_plane, b_plane;
};

// fp32 x fp32 from bf16 tensor cores: a = a0 + a1 + a2 (bf16 each, 24 mantissa bits in all), likewise b; the six products
// whose magnitude is >= 2^-16 of the result are accumulated in the fp32 TMEM accumulator, SMALLEST FIRST.
//   (1,1) (0,2) (2,0) (0,1) (1,0) (0,0)
__device__ __forceinline__ int split_plane_a(int pr) { return (0x102010 >> (4 * (5 - pr))) & 0xf; }   // 1,0,2,0,1,0
__device__ __forceinline__ int split_plane_b(int pr) { return (0x120100 >> (4 * (5 - pr))) & 0xf; }   // 1,2,0,1,0,0

// tile index -> (m_blk, n_blk) with grouped rasterisation so the ~148 tiles in flight share A/B panels in L2
__host__ __device__ __forceinline__ void tile_coords(const GemmCore& g, int tile, int& m_blk, int& n_blk) {
    const int per_group = g.group_m * g.tiles_n;
    const int group = tile / per_group;
    const int first_m = group * g.group_m;
    const int gsize = (g.tiles_m - first_m) < g.group_m ? (g.tiles_m - first_m) : g.group_m;
    const int in_group = tile - group * per_group;
    m_blk = first_m + in_group % gsize;
    n_blk = in_group / gsize;
}

// pair-tile index -> (pair row, n block) for the CTA-pair kernel (512-row pair tiles); groups of 4 pair rows (= 16 m
// blocks) keep ~74 live tiles near-square
__host__ __device__ __forceinline__ void pair_tile_coords(const GemmCore& g, int pair_tiles_m, int tile, int& pm, int& n_blk) {
    const int gm = g.group_m >> 2 > 0 ? g.group_m >> 2 : 1;
    const int per_group = gm * g.tiles_n;
    const int group = tile / per_group;
    const int first = group * gm;
    const int gsize = (pair_tiles_m - first) < gm ? (pair_tiles_m - first) : gm;
    const int in_group = tile - group * per_group;
    pm = first + in_group % gsize;
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
//   static constexpr int LAUNCH_BOUND_PAD;   threads added to the launch bound beyond the launch itself: 32 keeps room on
//        every SM (registers) for a one-warp barrier kernel of a concurrent stream (fused.cu); 0 otherwise.  ptxas
//        budgets registers for the bound rounded up to 128 threads: 320 -> 168/thread, 384 -> 168, above 384 -> 128.
//   static constexpr int EXTRA_WARPS;  static __device__ void side_loop(Params, GemmCore, side_warp, lane, scratch)
//        extra warps (between the MMA warp and the epilogue warps) that run side_loop concurrently with the GEMM
//   static __device__ void before_loads(const Params&, const GemmCore&, int m_blk, int n_blks)
//        called by the TMA producer thread before it loads A rows of 128-row blocks [m_blk, m_blk + n_blks) for a
//        tile: a policy whose A operand is still being assembled (gather_gemm.cu) waits here; empty otherwise
//   struct ProducerState;  static __device__ int kblock(const Params&, const GemmCore&, int kb, int num_kb, ProducerState&)
//        the k-block loaded in iteration kb of a tile's k-loop.  Identity by default; a policy whose A operand is sharded
//        along K (gather_gemm.cu, FSDP backward) starts every tile at its own shard's k-range and waits here for the rest
// ---------------------------------------------------------------------------------------------
struct EpiCtx {
    int tile;        // linear tile index in the scheduler's order
    int m_blk, n_blk;
    int m0, n0;      // first row / column of the tile
    int q;           // TMEM lane quarter (0..3) -> rows m0 + q*32 + lane
    int half;        // column half (0..1)
    int h;           // pair kernel: which 128-row accumulator half of the CTA's 256 rows (0, 1); 1-CTA kernel: always 1
                     // (so `h == 1` reads "this CTA has finished a whole scheduler tile")
    int ewarp;       // epilogue warp index 0..EPI_WARPS-1
    int lane;
    uint8_t* stage;  // this warp's EPI_STAGE_BYTES transpose buffer (staged launches only)
    uint32_t* scratch;  // 256 B shared by the epilogue warps, after the stage buffers (staged launches only)
};

template <class Epi>
__device__ __forceinline__ void gemm_1cta_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmCore& g,
                                               const typename Epi::Params& ep) {
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
    if (threadIdx.x >= 64 && threadIdx.x < 128)       // 256-byte scratch shared by epilogue / side warps starts at zero
        reinterpret_cast<uint32_t*>(sm.epi_stage + EPI_WARPS * EPI_STAGE_BYTES)[threadIdx.x - 64] = 0u;
    if (warp == 1) tmem_alloc_512(sm.tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_base;

    if (threadIdx.x == 0) {
        // ===================== TMA producer (one thread) =====================
        int s = 0;
        uint32_t phase = 0;
        [[maybe_unused]] typename Epi::ProducerState pst{};
        for (int it = g.cta; it < num_tiles; it += g.ncta) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            int m_blk, n_blk;
            tile_coords(g, tile, m_blk, n_blk);
            const int m0 = m_blk * BM, n0 = n_blk * BN;
            Epi::before_loads(ep, g, m_blk, 1);
#pragma unroll 1
            for (int pr = 0; pr < Epi::PRODUCTS; ++pr) {                 // 1, or the 6 plane products of the fp32 emulation
                int am = m0, ak = 0, bn = n0, bk = 0;                    // plane offsets along the operand's storage rows
                if constexpr (Epi::PRODUCTS > 1) {
                    const int pa = split_plane_a(pr) * g.a_plane, pb = split_plane_b(pr) * g.b_plane;
                    if (g.a_major == MAJOR_K) am += pa; else ak = pa;
                    if (g.b_major == MAJOR_K) bn += pb; else bk = pb;
                }
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int kbe = Epi::kblock(ep, g, kb, num_kb, pst);     // which k-block this iteration loads (kb itself by default)
                    mbar_wait(&sm.empty[s], phase ^ 1, g.err, 0x100 + s);
                    mbar_arrive_expect_tx(&sm.full[s], STAGE_BYTES);
                    const int k0 = kbe * BK;
                    if (g.a_major == MAJOR_K) {
                        tma_load_2d(sm.a(s), &tmA, &sm.full[s], k0, am);          // box {64 k, 128 m}
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)                          // box {64 m, 64 k} per atom
                            tma_load_2d(sm.a(s) + j * (BK * 128), &tmA, &sm.full[s], am + j * 64, k0 + ak);
                    }
                    if (g.b_major == MAJOR_K) {
                        tma_load_2d(sm.b(s), &tmB, &sm.full[s], k0, bn);          // box {64 k, 256 n}
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(sm.b(s) + j * (BK * 128), &tmB, &sm.full[s], bn + j * 64, k0 + bk);
                    }
                    if (++s == STAGES) { s = 0; phase ^= 1; }
                }
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
        for (int it = g.cta; it < num_tiles; it += g.ncta) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            mbar_wait(&sm.tmem_empty[as], aphase ^ 1, g.err, 0x200 + as);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
            const int kb_total = num_kb * Epi::PRODUCTS;             // all plane products accumulate into the same tile
            for (int kb = 0; kb < kb_total; ++kb) {
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
                if (kb == kb_total - 1) umma_commit(&sm.tmem_full[as]);
                if (++s == STAGES) { s = 0; phase ^= 1; }
            }
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    } else if (warp >= 2 && warp < 2 + Epi::EXTRA_WARPS) {
        // ===================== side warps (policy-defined work that runs under the MMAs) =====================
        Epi::side_loop(ep, g, warp - 2, lane, reinterpret_cast<uint32_t*>(sm.epi_stage + EPI_WARPS * EPI_STAGE_BYTES));
    } else if (warp >= 2 + Epi::EXTRA_WARPS) {
        // ===================== epilogue (8 warps: TMEM lane quarter x column half) =====================
        EpiCtx ctx;
        ctx.ewarp = warp - 2 - Epi::EXTRA_WARPS;
        ctx.q = warp & 3;                                        // tcgen05.ld lane window = warp_id % 4
        ctx.half = ctx.ewarp >> 2;
        ctx.lane = lane;
        ctx.stage = sm.epi_stage + ctx.ewarp * EPI_STAGE_BYTES;
        ctx.scratch = reinterpret_cast<uint32_t*>(sm.epi_stage + EPI_WARPS * EPI_STAGE_BYTES);
        ctx.tile = -1;
        ctx.h = 1;
        ctx.m_blk = ctx.n_blk = ctx.m0 = ctx.n0 = 0;
        typename Epi::State est;
        Epi::init(ep, g, ctx, est);
        int as = 0;
        uint32_t aphase = 0;
        for (int it = g.cta; it < num_tiles; it += g.ncta) {
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

template <class Epi>
__global__ void __launch_bounds__(GEMM_THREADS + 32 * Epi::EXTRA_WARPS + Epi::LAUNCH_BOUND_PAD, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmCore g_in, const typename Epi::Params ep) {
    GemmCore g = g_in;
    g.cta = (int)blockIdx.x;
    g.ncta = (int)gridDim.x;
    gemm_1cta_body<Epi>(tmA, tmB, g, ep);
}

// ---------------------------------------------------------------------------------------------
// CTA-pair variant: two CTAs of a cluster (one TPC) cooperate on a 512x256 tile with tcgen05.mma.cta_group::2.
// Each CTA stages 256 rows of A (two 128-row halves) and 128 of the tile's 256 B rows per k-block, and owns TWO
// 128x256 fp32 accumulators (all 512 TMEM columns): every B stage is used by both halves, so operand bytes per MAC
// drop to (256+128)/(256*256) - 25% below a 128x256-per-CTA pair and half of the 1-CTA kernel.  The leader CTA's MMA
// thread issues for both CTAs.  The epilogue releases the two accumulators separately, so the next tile's first
// MMAs overlap the second half's epilogue; the shared-memory ring keeps prefetching throughout.
// ---------------------------------------------------------------------------------------------
constexpr int P_STAGES = 4;
constexpr int P_A_STAGE_BYTES = 256 * BK * 2;   // 32 KiB: this CTA's 2 x 128 rows of A
constexpr int P_B_STAGE_BYTES = 128 * BK * 2;   // 16 KiB: this CTA's 128 of the tile's 256 B rows
constexpr int P_STAGE_BYTES = P_A_STAGE_BYTES + P_B_STAGE_BYTES;
static_assert(P_STAGES * P_STAGE_BYTES == STAGES * STAGE_BYTES, "both variants use the same smem carve-up");

template <class Epi>
__device__ __forceinline__ void gemm_pair_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmCore& g,
                                               const typename Epi::Params& ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = base;
    uint8_t* sB = base + P_STAGES * P_A_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + P_STAGES * P_STAGE_BYTES);
    uint64_t* full = bars;                            // [P_STAGES]  (only the leader's are waited on)
    uint64_t* empty = bars + P_STAGES;                // [P_STAGES]
    uint64_t* tmem_full = bars + 2 * P_STAGES;        // [1]
    uint64_t* tmem_empty = bars + 2 * P_STAGES + 1;   // [2] one per accumulator half (only the leader's are waited on)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * P_STAGES + 4);
    uint8_t* epi_stage = base + P_STAGES * P_STAGE_BYTES + 256;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();          // 0 = leader
    const int pair = g.cta >> 1;
    const int num_pairs = g.ncta >> 1;
    const int pair_tiles_m = (g.tiles_m + 3) >> 2;    // 512 rows = four 128-row blocks per pair tile
    const int num_tiles = pair_tiles_m * g.tiles_n;
    const int num_kb = (g.K + BK - 1) / BK;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
#pragma unroll
        for (int s = 0; s < P_STAGES; ++s) {
            mbar_init(&full[s], 1);                   // the leader's expect_tx arrival; BOTH CTAs' bytes are credited here
            mbar_init(&empty[s], 1);                  // one multicast commit
        }
        mbar_init(&tmem_full[0], 1);
        mbar_init(&tmem_empty[0], 2 * EPI_WARPS);     // epilogue warps of BOTH CTAs
        mbar_init(&tmem_empty[1], 2 * EPI_WARPS);
        fence_mbar_init();
    }
    if (threadIdx.x >= 64 && threadIdx.x < 128)       // 256-byte scratch shared by epilogue / side warps starts at zero
        reinterpret_cast<uint32_t*>(epi_stage + EPI_WARPS * EPI_STAGE_BYTES)[threadIdx.x - 64] = 0u;
    if (warp == 1) tmem_alloc_512_pair(tmem_slot);
    tc_fence_before();
    cluster_sync_all();                               // barriers + TMEM of both CTAs exist before any cross-CTA traffic
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    auto pair_coords = [&](int tile, int& pm, int& n_blk) { pair_tile_coords(g, pair_tiles_m, tile, pm, n_blk); };

    if (threadIdx.x == 0) {
        // ===================== TMA producer (one thread in EACH CTA) =====================
        int s = 0;
        uint32_t phase = 0;
        [[maybe_unused]] typename Epi::ProducerState pst{};
        for (int it = pair; it < num_tiles; it += num_pairs) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            int pm, n_blk;
            pair_coords(tile, pm, n_blk);
            const int m0 = (pm * 4 + (int)rank * 2) * BM;        // my 256 rows of A
            const int n0 = n_blk * BN + (int)rank * 128;         // my 128 rows of B
            Epi::before_loads(ep, g, pm * 4 + (int)rank * 2, 2);
#pragma unroll 1
            for (int pr = 0; pr < Epi::PRODUCTS; ++pr) {                 // 1, or the 6 plane products of the fp32 emulation
                int am = m0, ak = 0, bn = n0, bk = 0;
                if constexpr (Epi::PRODUCTS > 1) {
                    const int pa = split_plane_a(pr) * g.a_plane, pb = split_plane_b(pr) * g.b_plane;
                    if (g.a_major == MAJOR_K) am += pa; else ak = pa;
                    if (g.b_major == MAJOR_K) bn += pb; else bk = pb;
                }
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int kbe = Epi::kblock(ep, g, kb, num_kb, pst);
                    mbar_wait(&empty[s], phase ^ 1, g.err, 0x100 + s);
                    // The peer's bytes may land before this arrival: the transaction count just goes negative for a
                    // moment, and the phase cannot complete until the one pending arrival (this one) has happened.
                    if (rank == 0) mbar_arrive_expect_tx(&full[s], 2 * P_STAGE_BYTES);
                    const int k0 = kbe * BK;
                    uint8_t* a_dst = sA + s * P_A_STAGE_BYTES;
                    uint8_t* b_dst = sB + s * P_B_STAGE_BYTES;
                    if (g.a_major == MAJOR_K) {
                        tma_load_2d_pair(a_dst, &tmA, &full[s], k0, am);                       // box {64 k, 256 m}
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)                                            // 4 x box {64 m, 64 k}
                            tma_load_2d_pair(a_dst + j * (BK * 128), &tmA, &full[s], am + j * 64, k0 + ak);
                    }
                    if (g.b_major == MAJOR_K) {
                        tma_load_2d_pair(b_dst, &tmB, &full[s], k0, bn);                       // box {64 k, 128 n}
                    } else {
                        tma_load_2d_pair(b_dst, &tmB, &full[s], bn, k0 + bk);
                        tma_load_2d_pair(b_dst + BK * 128, &tmB, &full[s], bn + 64, k0 + bk);
                    }
                    if (++s == P_STAGES) { s = 0; phase ^= 1; }
                }
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
        uint32_t tphase = 0;
        const int kb_total = num_kb * Epi::PRODUCTS;                 // all plane products accumulate into the same tile
        for (int it = pair; it < num_tiles; it += num_pairs) {
            for (int kb = 0; kb < kb_total; ++kb) {
                mbar_wait(&full[s], phase, g.err, 0x300 + s);
                tc_fence_after();
                const uint32_t a_base = smem_u32(sA + s * P_A_STAGE_BYTES);
                const uint32_t b_base = smem_u32(sB + s * P_B_STAGE_BYTES);
#pragma unroll
                for (int h = 0; h < 2; ++h) {                        // the two 128-row halves share this B stage
                    if (kb == 0) {
                        mbar_wait(&tmem_empty[h], tphase ^ 1, g.err, 0x200 + h);   // epilogue has drained this half
                        tc_fence_after();
                    }
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t adesc = make_smem_desc_sw128(a_base + h * (128 * BK * 2) + k * a_kadv, a_lbo, 1024);
                        const uint64_t bdesc = make_smem_desc_sw128(b_base + k * b_kadv, b_lbo, 1024);
                        umma_bf16_pair(tmem_base + (uint32_t)(h * BN), adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit_pair(&empty[s]);                         // frees the slot in BOTH CTAs
                if (kb == kb_total - 1) umma_commit_pair(&tmem_full[0]);
                if (++s == P_STAGES) { s = 0; phase ^= 1; }
            }
            tphase ^= 1;
        }
    } else if (warp >= 2 && warp < 2 + Epi::EXTRA_WARPS) {
        // ===================== side warps (policy-defined work that runs under the MMAs) =====================
        Epi::side_loop(ep, g, warp - 2, lane, reinterpret_cast<uint32_t*>(epi_stage + EPI_WARPS * EPI_STAGE_BYTES));
    } else if (warp >= 2 + Epi::EXTRA_WARPS) {
        // ===================== epilogue (8 warps per CTA) =====================
        EpiCtx ctx;
        ctx.ewarp = warp - 2 - Epi::EXTRA_WARPS;
        ctx.q = warp & 3;
        ctx.half = ctx.ewarp >> 2;
        ctx.lane = lane;
        ctx.stage = epi_stage + ctx.ewarp * EPI_STAGE_BYTES;
        ctx.scratch = reinterpret_cast<uint32_t*>(epi_stage + EPI_WARPS * EPI_STAGE_BYTES);
        ctx.tile = -1;
        ctx.h = 0;
        ctx.m_blk = ctx.n_blk = ctx.m0 = ctx.n0 = 0;
        typename Epi::State est;
        Epi::init(ep, g, ctx, est);
        uint32_t tphase = 0;
        for (int it = pair; it < num_tiles; it += num_pairs) {
            int tile = it + g.tile_offset;
            if (tile >= num_tiles) tile -= num_tiles;
            int pm;
            pair_coords(tile, pm, ctx.n_blk);
            ctx.tile = tile;
            ctx.n0 = ctx.n_blk * BN;
            mbar_wait(&tmem_full[0], tphase, g.err, 0x400);
            tc_fence_after();
#pragma unroll 1
            for (int h = 0; h < 2; ++h) {
                ctx.h = h;
                ctx.m_blk = pm * 4 + (int)rank * 2 + h;
                ctx.m0 = ctx.m_blk * BM;
                const uint32_t taddr = tmem_base + ((uint32_t)(ctx.q * 32) << 16) + (uint32_t)(h * BN + ctx.half * (BN / 2));
                const bool real_tile = ctx.m_blk < g.tiles_m;       // ragged M: trailing 128-row blocks may not exist
                if (real_tile) Epi::process_tile(ep, g, ctx, taddr, est);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (rank == 0) mbar_arrive(&tmem_empty[h]);
                    else mbar_arrive_remote(&tmem_empty[h], 0);
                }
                if (real_tile) Epi::after_tile(ep, g, ctx, est);
                else Epi::after_skipped_tile(ep, g, ctx, est);
            }
            tphase ^= 1;
        }
        Epi::tail(ep, g, ctx, est);
    }

    tc_fence_before();
    cluster_sync_all();                               // neither CTA may retire while its pair still uses its smem / barriers
    tc_fence_after();
    if (warp == 1) tmem_dealloc_512_pair(tmem_base);
}

template <class Epi>
__global__ void __launch_bounds__(GEMM_THREADS + 32 * Epi::EXTRA_WARPS + Epi::LAUNCH_BOUND_PAD, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmCore g_in, const typename Epi::Params ep) {
    GemmCore g = g_in;
    g.cta = (int)blockIdx.x;
    g.ncta = (int)gridDim.x;
    gemm_pair_body<Epi>(tmA, tmB, g, ep);
}

// ---------------------------------------------------------------------------------------------
// Single-launch multi-rank emulation (tests, ncu): ONE cooperative grid in which CTAs [r*cpr, (r+1)*cpr) run the body
// exactly as rank r's kernel would - its own tensor maps, tile walk and policy parameters, read from a table in global
// memory.  Kernels that spin on one another's flags must not be separate launches on one GPU (nothing makes them
// co-resident); this is how the cross-rank protocols of fused.cu / gather_gemm.cu are exercised on a single B200, with
// every "peer" pointer resolving to local memory.  Same code, same protocol; only NVLink is missing.
// ---------------------------------------------------------------------------------------------
template <class Epi>
struct alignas(64) EmuRank {
    CUtensorMap tmA, tmB;
    GemmCore core;
    typename Epi::Params ep;
};

template <class Epi, bool PAIR>
__global__ void __launch_bounds__(GEMM_THREADS + 32 * Epi::EXTRA_WARPS + Epi::LAUNCH_BOUND_PAD, 1)
gemm_tcgen05_emu_kernel(const EmuRank<Epi>* __restrict__ ranks, const int ctas_per_rank) {
    const int r = (int)blockIdx.x / ctas_per_rank;
    const EmuRank<Epi>& R = ranks[r];
    GemmCore g = R.core;
    g.cta = (int)blockIdx.x - r * ctas_per_rank;
    g.ncta = ctas_per_rank;
    if constexpr (PAIR) gemm_pair_body<Epi>(R.tmA, R.tmB, g, R.ep);
    else gemm_1cta_body<Epi>(R.tmA, R.tmB, g, R.ep);
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
    EPI_RELU_MASK = 8,      // RELU, plus a packed sign mask (1 bit per element, C2 as uint8 [M, ldc2 bytes]) for the backward
    EPI_RELU_BWD_MASK = 9,  // C = acc where the mask bit is set, else 0 (aux = the uint8 mask, ldaux in bytes)
    EPI_SGD_SR = 10,        // EPI_SGD with stochastic rounding of the new weight (internal: chosen by b200_set_update_rounding)
    EPI_COUNT = 11,
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
    uint32_t sr_seed;      // EPI_SGD: 0 = round the new weight to nearest, else stochastic rounding with this seed
};

__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad_f(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
    return cdf + x * pdf;
}

// Row-owner registers <-> whole 128-byte lines, through one warp's swizzled 32x128B stage buffer.
// After tcgen05.ld each lane owns a row; a direct store would hit 32 different lines with 16 bytes each (and a
// load would do the same).  Bouncing through shared memory lets 8 lanes cover one 128-byte line: 4 full lines per
// instruction.  Both helpers end with a __syncwarp so the buffer can be reused immediately.
__device__ __forceinline__ void stage_store_rows(uint4* stage, int lane, const uint4 (&rowv)[8], __nv_bfloat16* dst, long long ld,
                                                 int row0, int col0, int M, int N) {
#pragma unroll
    for (int pc = 0; pc < 8; ++pc) stage[lane * 8 + (pc ^ (lane & 7))] = rowv[pc];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = i * 4 + (lane >> 3);
        const int pc = lane & 7;
        const int row = row0 + r, col = col0 + pc * 8;
        if (row < M && col < N) st_u4(dst + (long long)row * ld + col, stage[r * 8 + (pc ^ (r & 7))]);
    }
    __syncwarp();
}
template <bool READ_ONLY>
__device__ __forceinline__ void coalesced_load_rows(uint4 (&v)[8], int lane, const __nv_bfloat16* src, long long ld, int row0, int col0,
                                                    int M, int N) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = i * 4 + (lane >> 3);
        const int pc = lane & 7;
        const int row = row0 + r, col = col0 + pc * 8;
        if (row < M && col < N) {
            if constexpr (READ_ONLY) v[i] = ld_nc_u4(src + (long long)row * ld + col);             // aux: never written by this kernel
            else v[i] = *reinterpret_cast<const uint4*>(src + (long long)row * ld + col);             // the weight tile this CTA alone updates
        } else {
            v[i] = make_uint4(0, 0, 0, 0);
        }
    }
}
__device__ __forceinline__ void stage_to_row_owner(uint4* stage, int lane, const uint4 (&v)[8], uint4 (&rowv)[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = i * 4 + (lane >> 3);
        const int pc = lane & 7;
        stage[r * 8 + (pc ^ (r & 7))] = v[i];
    }
    __syncwarp();
#pragma unroll
    for (int pc = 0; pc < 8; ++pc) rowv[pc] = stage[lane * 8 + (pc ^ (lane & 7))];
    __syncwarp();
}

template <int KIND>
struct LocalEpi {
    using Params = LocalEpiParams;
    static constexpr int SMEM_BYTES = GEMM_SMEM_BYTES_STAGED;
    static constexpr int EXTRA_WARPS = 0;
    static constexpr int LAUNCH_BOUND_PAD = 0;
    static constexpr int PRODUCTS = 1;
    static __device__ __forceinline__ void side_loop(const Params&, const GemmCore&, int, int, uint32_t*) {}
    static __device__ __forceinline__ void before_loads(const Params&, const GemmCore&, int, int) {}
    struct ProducerState {};
    static __device__ __forceinline__ int kblock(const Params&, const GemmCore&, int kb, int, ProducerState&) { return kb; }
    // a second bf16 [M,N] operand is read by these epilogues: aux, or (SGD) the weight tile itself
    static constexpr bool IS_SGD = (KIND == EPI_SGD || KIND == EPI_SGD_SR);
    static constexpr bool READS = (KIND == EPI_RELU_BWD || KIND == EPI_GELU_BWD || KIND == EPI_ADD || IS_SGD);
    struct State {};
    static __device__ __forceinline__ void init(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_skipped_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void tail(const Params&, const GemmCore&, const EpiCtx&, State&) {}

    static __device__ __forceinline__ void process_tile(const Params& p, const GemmCore& g, const EpiCtx& ctx, uint32_t taddr, State&) {
        const int lane = ctx.lane;
        const int row0 = ctx.m0 + ctx.q * 32;                        // first of this warp's 32 rows
        const int colbase = ctx.n0 + ctx.half * (BN / 2);
        uint4* stage = reinterpret_cast<uint4*>(ctx.stage);
        const __nv_bfloat16* rd = IS_SGD ? reinterpret_cast<const __nv_bfloat16*>(p.C) : p.aux;
        const long long rd_ld = IS_SGD ? p.ldc : p.ldaux;

        if constexpr (KIND == EPI_STORE_F32) {                       // fp32 rows are 128 B per 32 columns already
            const int row = row0 + lane;
#pragma unroll 1
            for (int c = 0; c < BN / 2 / 32; ++c) {
                uint32_t acc[32];
                tmem_ld_32x32(taddr + c * 32, acc);
                tmem_ld_wait();
                if (row >= g.M) continue;
                float* out = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + colbase + c * 32;
#pragma unroll
                for (int v = 0; v < 8; ++v) {
                    if (colbase + c * 32 + v * 4 >= g.N) break;
                    *reinterpret_cast<float4*>(out + v * 4) = make_float4(p.alpha * __uint_as_float(acc[v * 4]), p.alpha * __uint_as_float(acc[v * 4 + 1]),
                                                                         p.alpha * __uint_as_float(acc[v * 4 + 2]), p.alpha * __uint_as_float(acc[v * 4 + 3]));
                }
            }
            return;
        }

        // Packed ReLU sign mask, 1 bit per element, in a TILE-BLOCKED layout: the 16 bytes of (row, 128-column chunk cc) live at
        //   ((row / 128) * n_chunks + cc) * 2048 + (row % 128) * 16,   n_chunks = ceil(N / 128)
        // so the 32 rows of a warp are 512 contiguous bytes (whole lines; round 1's row-major [M, N/8] layout scattered 32
        // 16-byte pieces over 32 lines - ncu: tensor pipe 85 % for this epilogue against 97 % for a plain store).
        // m0..m3 are named registers on purpose: an indexed array lands in local memory (the cp loop is not unrolled).
        uint32_t m0 = 0u, m1 = 0u, m2 = 0u, m3 = 0u;
        // colbase < N: with N % 256 == 128 the second column half of the last n-tile lies wholly outside the matrix
        const bool mrow_ok = (row0 + lane) < g.M && colbase < g.N;
        [[maybe_unused]] const long long mask_off =
            ((long long)(row0 >> 7) * ((g.N + 127) >> 7) + (colbase >> 7)) * 2048 + (long long)((row0 & 127) + lane) * 16;
        if constexpr (KIND == EPI_RELU_BWD_MASK) {
            if (mrow_ok) {
                const uint4 mv = ld_nc_u4(reinterpret_cast<const uint8_t*>(p.aux) + mask_off);
                m0 = mv.x; m1 = mv.y; m2 = mv.z; m3 = mv.w;
            }
        }
        uint4 rd_next[8];
        if constexpr (READS) coalesced_load_rows<!IS_SGD>(rd_next, lane, rd, rd_ld, row0, colbase, g.M, g.N);
#pragma unroll 1
        for (int cp = 0; cp < 2; ++cp) {                             // 64 columns = one 128-byte line per row
            const int col0 = colbase + cp * 64;
            uint4 rdrow[8];
            if constexpr (READS) {
                stage_to_row_owner(stage, lane, rd_next, rdrow);
                if (cp == 0) coalesced_load_rows<!IS_SGD>(rd_next, lane, rd, rd_ld, row0, col0 + 64, g.M, g.N);   // prefetch the next 64 columns
            }
            uint32_t a0[32], a1[32];
            tmem_ld_32x32(taddr + cp * 64, a0);
            tmem_ld_32x32(taddr + cp * 64 + 32, a1);
            tmem_ld_wait();
            uint4 outv[8], prev[8];
            [[maybe_unused]] const uint32_t mw0 = cp == 0 ? m0 : m2, mw1 = cp == 0 ? m1 : m3;   // this 64-column step's mask words (BWD)
            [[maybe_unused]] uint32_t nw0 = 0u, nw1 = 0u;                                        // ... and the ones being built (FWD)
#pragma unroll
            for (int pc = 0; pc < 8; ++pc) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(pc < 4 ? a0[pc * 8 + j] : a1[(pc - 4) * 8 + j]);
                if constexpr (KIND == EPI_STORE || KIND == EPI_RELU || KIND == EPI_RELU_MASK || KIND == EPI_GELU || KIND == EPI_ADD) {
                    if (p.bias != nullptr) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int col = col0 + pc * 8 + j;
                            f[j] += col < g.N ? __ldg(p.bias + col) : 0.0f;
                        }
                    }
                }
                if constexpr (KIND == EPI_RELU || KIND == EPI_RELU_MASK) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const bool keep = !(f[j] <= 0.0f);                           // NaN stays NaN, like torch.where(x<=0,0,x)
                        if constexpr (KIND == EPI_RELU_MASK) {
                            if (pc < 4) nw0 |= (keep ? 1u : 0u) << ((pc & 3) * 8 + j);
                            else nw1 |= (keep ? 1u : 0u) << ((pc & 3) * 8 + j);
                        }
                        f[j] = keep ? f[j] : 0.0f;
                    }
                } else if constexpr (KIND == EPI_RELU_BWD_MASK) {
                    const uint32_t w = (pc < 4 ? mw0 : mw1) >> ((pc & 3) * 8);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = ((w >> j) & 1u) ? f[j] : 0.0f;   // masked_fill_(x <= 0, 0)
                } else if constexpr (KIND == EPI_GELU) {
                    prev[pc].x = pack_bf16x2(f[0], f[1]); prev[pc].y = pack_bf16x2(f[2], f[3]);
                    prev[pc].z = pack_bf16x2(f[4], f[5]); prev[pc].w = pack_bf16x2(f[6], f[7]);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
                } else if constexpr (READS) {
                    const uint32_t aw[4] = {rdrow[pc].x, rdrow[pc].y, rdrow[pc].z, rdrow[pc].w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float lo = bf16_lo(aw[j]), hi = bf16_hi(aw[j]);
                        if constexpr (KIND == EPI_RELU_BWD) {
                            f[2 * j] = lo <= 0.0f ? 0.0f : f[2 * j];                // masked_fill_(x <= 0, 0)
                            f[2 * j + 1] = hi <= 0.0f ? 0.0f : f[2 * j + 1];
                        } else if constexpr (KIND == EPI_GELU_BWD) {
                            f[2 * j] *= gelu_grad_f(lo);
                            f[2 * j + 1] *= gelu_grad_f(hi);
                        } else if constexpr (KIND == EPI_ADD) {
                            f[2 * j] += lo;
                            f[2 * j + 1] += hi;
                        } else {                                                    // SGD: W - alpha * dW
                            f[2 * j] = lo - p.alpha * f[2 * j];
                            f[2 * j + 1] = hi - p.alpha * f[2 * j + 1];
                        }
                    }
                }
                if constexpr (KIND == EPI_SGD_SR) {
                    const unsigned long long e0 = (unsigned long long)(row0 + lane) * (unsigned long long)p.ldc + (unsigned long long)(col0 + pc * 8);
                    outv[pc].x = pack_bf16x2_sr(f[0], f[1], p.sr_seed, e0); outv[pc].y = pack_bf16x2_sr(f[2], f[3], p.sr_seed, e0 + 2);
                    outv[pc].z = pack_bf16x2_sr(f[4], f[5], p.sr_seed, e0 + 4); outv[pc].w = pack_bf16x2_sr(f[6], f[7], p.sr_seed, e0 + 6);
                } else {
                    outv[pc].x = pack_bf16x2(f[0], f[1]); outv[pc].y = pack_bf16x2(f[2], f[3]);
                    outv[pc].z = pack_bf16x2(f[4], f[5]); outv[pc].w = pack_bf16x2(f[6], f[7]);
                }
            }
            if constexpr (KIND == EPI_RELU_MASK) {
                if (cp == 0) { m0 = nw0; m1 = nw1; } else { m2 = nw0; m3 = nw1; }
            }
            if constexpr (KIND == EPI_GELU)
                stage_store_rows(stage, lane, prev, p.C2, p.ldc2, row0, col0, g.M, g.N);
            stage_store_rows(stage, lane, outv, reinterpret_cast<__nv_bfloat16*>(p.C), p.ldc, row0, col0, g.M, g.N);
        }
        if constexpr (KIND == EPI_RELU_MASK) {
            if (mrow_ok)
                st_u4(reinterpret_cast<uint8_t*>(p.C2) + mask_off, make_uint4(m0, m1, m2, m3));
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
    int planes = 1;                       // 3: split fp32 operands (three bf16 planes stacked along the storage rows)
    long long a_plane = 0, b_plane = 0;   // storage rows between consecutive planes
};

int validate_operands(const GemmOperands& o);
int make_core_and_maps(const GemmOperands& o, GemmCore* core, CUtensorMap* tmA, CUtensorMap* tmB, bool pair);

bool use_pair_kernel();   // runtime switch (B200_GEMM_PAIR=0 selects the 1-CTA kernel)

template <class Epi>
int launch_gemm(const GemmOperands& o, const typename Epi::Params& ep, int max_ctas, cudaStream_t stream,
                int tile_offset_num = 0, int tile_offset_den = 1, bool overlap_epilogue = false, int group_m = 0) {
    GemmCore core;
    alignas(64) CUtensorMap tmA, tmB;
    // overlap_epilogue: the epilogue is long (peer-memory pushes, cross-rank reduction), so use the kernel whose two
    // 128x256 accumulator stages let it run under the next tile's MMAs, rather than the 256x256-per-CTA pair kernel
    // whose single accumulator set leaves the epilogue exposed.
    const bool pair = use_pair_kernel() && !overlap_epilogue;
    int rc = make_core_and_maps(o, &core, &tmA, &tmB, pair);
    if (rc != 0) return rc;
    if (group_m > 0) core.group_m = group_m;   // rasterisation group (128-row blocks); callers align it with a row shard
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
        gemm_tcgen05_kernel<Epi><<<grid, GEMM_THREADS + 32 * Epi::EXTRA_WARPS, Epi::SMEM_BYTES, stream>>>(tmA, tmB, core, ep);
        return check_cuda(cudaGetLastError(), "gemm launch");
    }
    const int pair_tiles = ((core.tiles_m + 3) / 4) * core.tiles_n;      // 512 x 256 per CTA pair
    core.tile_offset = (int)((long long)pair_tiles * tile_offset_num / tile_offset_den);
    int pairs = cap / 2;
    if (pairs < 1) pairs = 1;
    if (pairs > pair_tiles) pairs = pair_tiles;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(GEMM_THREADS + 32 * Epi::EXTRA_WARPS);
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

// Host side of the emulation: build every rank's tensor maps / core / policy parameters, upload the table, launch ONE
// cooperative grid of world * ctas_per_rank CTAs.  `rots[r] / world` is rank r's tile-walk rotation (launch_gemm's
// tile_offset_num / tile_offset_den).  `table`: device scratch of EMU_TABLE_BYTES.
constexpr int EMU_MAX_WORLD = 8;
constexpr int EMU_TABLE_BYTES = 8 * 1024;
template <class Epi, class P>
int launch_gemm_emu(int world, const GemmOperands* ops, const P* params, const int* rots, int ctas_per_rank, bool pair,
                      int group_m, void* table, cudaStream_t st) {
    static_assert(sizeof(EmuRank<Epi>) * EMU_MAX_WORLD <= EMU_TABLE_BYTES, "emulation table too small");
    EmuRank<Epi> host[EMU_MAX_WORLD];
    for (int r = 0; r < world; ++r) {
        if (int rc = make_core_and_maps(ops[r], &host[r].core, &host[r].tmA, &host[r].tmB, pair)) return rc;
        if (group_m > 0) host[r].core.group_m = group_m;
        const int tiles = pair ? ((host[r].core.tiles_m + 3) / 4) * host[r].core.tiles_n : host[r].core.tiles_m * host[r].core.tiles_n;
        host[r].core.tile_offset = (int)((long long)tiles * rots[r] / world);
        host[r].ep = params[r];
    }
    if (int rc = check_cuda(cudaMemcpyAsync(table, host, sizeof(EmuRank<Epi>) * world, cudaMemcpyHostToDevice, st), "emu: table upload")) return rc;
    if (int rc = check_cuda(cudaStreamSynchronize(st), "emu: table upload sync")) return rc;   // `host` is a stack array
    void* fn = pair ? (void*)gemm_tcgen05_emu_kernel<Epi, true> : (void*)gemm_tcgen05_emu_kernel<Epi, false>;
    if (int rc = check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, Epi::SMEM_BYTES), "emu: smem attribute")) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(world * ctas_per_rank);
    cfg.blockDim = dim3(GEMM_THREADS + 32 * Epi::EXTRA_WARPS);
    cfg.dynamicSmemBytes = Epi::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    at[na].id = cudaLaunchAttributeCooperative;     // all CTAs co-resident, or the launch fails: the slices wait on each other
    at[na].val.cooperative = 1;
    ++na;
    if (pair) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = 2; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    const EmuRank<Epi>* tab = reinterpret_cast<const EmuRank<Epi>*>(table);
    if (pair) return check_cuda(cudaLaunchKernelEx(&cfg, gemm_tcgen05_emu_kernel<Epi, true>, tab, ctas_per_rank), "emu launch (pair)");
    return check_cuda(cudaLaunchKernelEx(&cfg, gemm_tcgen05_emu_kernel<Epi, false>, tab, ctas_per_rank), "emu launch (1-CTA)");
}


}  // namespace b200
