// Fused GEMM + collective: ONE persistent tcgen05 kernel that computes its output tiles and, tile by tile,
// reduces them across ranks over NVLink peer memory.
//
//   push    (epilogue warps) each finished accumulator tile is converted to bf16, transposed through shared memory so
//           that every warp store covers whole 128-byte lines, and STORED INTO THE OWNER RANK'S INBOX (a peer pointer);
//           a release flag per (tile, source) tells the owner it has landed.
//   reduce  (side warps, concurrently with the TMA / MMA / epilogue warps) poll the owned tiles assigned to this CTA,
//           in any order; a tile whose `world` partials have all arrived is summed in fp32 in a fixed source order, the
//           optional optimizer step is applied (W -= alpha * sum) and the result is written to the local output and,
//           in broadcast mode, to every peer's output, followed by a per-tile ready flag.
//   done    the kernel returns when its owned tiles are reduced and, in broadcast mode, every tile of the local output
//           has been delivered.
// Compute-bound calls run on the 256x256-per-CTA pair mainloop; link-bound ones on the double-buffered 128x256
// mainloop whose second accumulator stage absorbs NVLink back-pressure in the epilogue (chosen per call, see below).
//
// One kernel therefore covers the reference's
//   DDP   dW GEMM + all_reduce(SUM) + SGD      (train_ffns.py:66/68 + :165 + :172)   interleaved owners, broadcast, SGD
//   FSDP  dW GEMM + reduce_scatter + shard SGD (:66/68 + :255-256 + :258-259)        row-sharded owners, local, SGD
//   MP    row-parallel GEMM + all_reduce       (:57 + :303, :68 + :309)              interleaved owners, broadcast, SUM
//   MP-SP row-parallel GEMM + reduce_scatter   (sequence-parallel extension)         row-sharded owners, local, SUM
//
// Variants (template parameters of FusedEpiT, all sharing the protocol above): the number of side warps, the side
// warps' polling back-off, and the optimizer applied in the owner's reduction (SGD from FusedParams, or Adam with the
// owner holding the fp32 moments of its tiles - FusedAdamParams).
//
// Inbox slots are double-buffered by call parity so a fast rank may start the next call while a slow one is still
// reducing; flags carry a per-channel sequence number (monotonic, never reset).  With row-sharded owners ranks rotate
// their tile order so that at any moment they push to different owners (interleaved owners spread by construction).
#include "gemm_kernel.cuh"
#include "../include/b200ffn.h"
#include <cstdlib>
#include <type_traits>

namespace b200 {

constexpr int MAX_WORLD_F = 8;
constexpr int TILE_ELEMS = BM * BN;                 // 32768 bf16 = 64 KiB per partial tile
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int MAX_OWNED_PER_CTA = 64;               // the poller tracks its owned tiles in a 64-bit mask

enum FusedOp : int { FOP_SUM = 0, FOP_SGD = 1 };
enum OwnerMode : int { OWN_INTERLEAVED = 0, OWN_ROW_SHARDED = 1 };

struct FusedParams {
    void* inbox[MAX_WORLD_F];        // symmetric inbox base per rank: [2 parities][owned_cap][world][TILE_ELEMS] bf16
    uint32_t* flags[MAX_WORLD_F];    // symmetric flags per rank: arrive[owned_cap][8] then ready[ready_cap]
    void* out[MAX_WORLD_F];          // result base per rank (peers are touched only when broadcast)
    long long ldo;
    int out_row_offset;              // first global row held by `out` (row-sharded outputs)
    int rank, world;
    uint32_t seq;                    // per-channel call counter (>= 1), identical on every rank
    int parity;                      // which inbox half this call uses
    int owner_mode, mpo;             // mpo = m-tiles per owner (row-sharded)
    int op, broadcast;
    float alpha;
    int owned_cap;
    int owned_count;                 // tiles this rank owns in this call
    int ready_base;                  // word offset of the ready flags inside the flag block
    // NVLS form (FusedEpiT<.., NVLS = 1>): partials stay LOCAL, the owner sums them inside the NVSwitch.
    __nv_bfloat16* part_local;       // this rank's partial tiles [2 parities][part_cap][TILE_ELEMS], a multicast-bound block
    const __nv_bfloat16* part_mc;    // the same block through its multicast address (multimem.ld_reduce sums all ranks' copies)
    __nv_bfloat16* out_mc;           // the output through its multicast address (multimem.st stores into every replica)
    int part_cap;
    uint32_t sr_seed;                // optimizer forms: 0 = round the new weight to nearest, else stochastic rounding (common.cuh)
};

// Adam variant (OPT = 1): the owner of a tile also owns its optimizer state - fp32 moments in a compact per-owned-tile
// layout [owned slot][128][256], i.e. 1/world of the state per rank (ZeRO-1 for DDP, the natural layout for FSDP).
// Formula as b200_adam_bf16 (elementwise.cu).  Validated in round 2 (tests/test_emu_gpu.py::test_emu_dw_adam; 2 GPUs: test_fused_adam_matches_unfused_adam).
struct FusedAdamParams : FusedParams {
    float* adam_m;
    float* adam_v;
    float lr, beta1, beta2, eps, bc1, bc2, gscale;
};

__host__ __device__ __forceinline__ int owner_of(const FusedParams& p, const GemmCore& g, int m_blk, int n_blk, int& k_local) {
    if (p.owner_mode == OWN_ROW_SHARDED) {
        k_local = (m_blk % p.mpo) * g.tiles_n + n_blk;
        return m_blk / p.mpo;
    }
    const int lin = m_blk * g.tiles_n + n_blk;
    k_local = lin / p.world;
    return lin % p.world;
}
__host__ __device__ __forceinline__ void owned_tile_coords(const FusedParams& p, const GemmCore& g, int k_local, int& m_blk, int& n_blk) {
    if (p.owner_mode == OWN_ROW_SHARDED) {
        m_blk = p.rank * p.mpo + k_local / g.tiles_n;
        n_blk = k_local % g.tiles_n;
    } else {
        const int lin = k_local * p.world + p.rank;
        m_blk = lin / g.tiles_n;
        n_blk = lin % g.tiles_n;
    }
}
__host__ __device__ __forceinline__ size_t inbox_slot(const FusedParams& p, int k_local, int src) {
    return ((size_t)(p.parity * p.owned_cap + k_local) * p.world + src) * TILE_ELEMS;
}
__host__ __device__ __forceinline__ size_t part_slot(const FusedParams& p, int lin_tile) {
    return ((size_t)p.parity * p.part_cap + lin_tile) * TILE_ELEMS;
}
// sum over every rank's copy of 8 bf16 (fp32 accumulation inside the switch), and a store into every rank's copy
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void epi_group_sync() {       // the 8 epilogue warps only (named barrier 1)
    asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
}
__device__ __forceinline__ uint4 pack_row8(const uint32_t* a) {
    uint4 v;
    v.x = pack_bf16x2(__uint_as_float(a[0]), __uint_as_float(a[1]));
    v.y = pack_bf16x2(__uint_as_float(a[2]), __uint_as_float(a[3]));
    v.z = pack_bf16x2(__uint_as_float(a[4]), __uint_as_float(a[5]));
    v.w = pack_bf16x2(__uint_as_float(a[6]), __uint_as_float(a[7]));
    return v;
}

__device__ __forceinline__ void side_group_sync(int threads) {   // the side warps only (named barrier 2)
    asm volatile("bar.sync 2, %0;" ::"r"(threads) : "memory");
}

// SIDE_WARPS: warps that reduce owned tiles under the MMAs.  PAD: threads added to the launch bound so that a one-warp
// barrier kernel of a concurrent stream (FSDP's copy-engine prefetch) finds registers on every SM.  ptxas budgets
// registers for the bound rounded up to 128 threads: 320 + 64 -> 168 per thread, anything above 384 -> 128.
// OPT: 0 = SUM / SGD from FusedParams, 1 = Adam (FusedAdamParams).
//
// Round-2 changes, from the ncu captures of the round-1 kernel (profiles/ncu/prof_fused_dw_world1.*: tensor pipe 86 % vs 97 %
// for the local kernel, 2.9x the instructions):
//   * ONE side warp polls (in arrival order, exponential back-off 0.25 -> 2 us) while the others are parked on a hardware
//     barrier; round 1 had every side warp re-scan all of its tiles every 200 ns.
//   * In the CTA-pair kernel the arrival flags of a tile's two halves are published together AFTER the second half has
//     been handed back to the MMA warp: the system-scope fence (microseconds over NVLink) no longer sits between the two
//     halves' TMEM drains, i.e. on the path that gates the next tile's first MMAs.
//   * Completion is counted with the barrier instead of shared-memory atomics.
//   * NVLS = 1 (link-bound all-reduce / reduce-scatter of activations, SUM or in-place add): every rank stores its partial
//     tile LOCALLY into a multicast-bound buffer and only sends the arrival flag; the owner's side warps read the tile
//     through the multicast address (multimem.ld_reduce: the NVSwitch pulls the `world` copies and returns their fp32-
//     accumulated sum) and, in broadcast mode, write the result with multimem.st (the switch replicates it).  Per GPU and
//     direction ~1.1x the activation bytes cross NVLink instead of the ~1.75x of push + broadcast.
template <int SIDE_WARPS, int PAD, int OPT = 0, int NVLS = 0>
struct FusedEpiT {
    static_assert(!(NVLS == 1 && OPT == 1), "the NVLS form serves activations (SUM / in-place add), not optimizer updates");
    using Params = std::conditional_t<OPT == 1, FusedAdamParams, FusedParams>;
    static constexpr int SMEM_BYTES = GEMM_SMEM_BYTES_STAGED;
    static constexpr int EXTRA_WARPS = SIDE_WARPS;   // side warps reduce/broadcast owned tiles under the MMAs
    static constexpr int LAUNCH_BOUND_PAD = PAD;
    static constexpr int PRODUCTS = 1;
    struct State {
        int k_local;               // inbox slot of the tile being pushed
        int owner;
        int pend_k, pend_owner;    // first half of a pair tile, pushed but not yet published (owner < 0: none)
    };

    static __device__ __forceinline__ void before_loads(const Params&, const GemmCore&, int, int) {}
    struct ProducerState {};
    static __device__ __forceinline__ int kblock(const Params&, const GemmCore&, int kb, int, ProducerState&) { return kb; }
    static __device__ __forceinline__ void init(const Params&, const GemmCore&, const EpiCtx&, State& st) {
        st.k_local = 0;
        st.owner = 0;
        st.pend_k = 0;
        st.pend_owner = -1;
    }

    // ---- push this warp's 32 rows x 128 columns of the accumulator into the owner's inbox ------------------
    static __device__ __forceinline__ void process_tile(const Params& p, const GemmCore& g, const EpiCtx& ctx, uint32_t taddr, State& st) {
        st.owner = owner_of(p, g, ctx.m_blk, ctx.n_blk, st.k_local);
        __nv_bfloat16* dst_tile;
        if constexpr (NVLS == 1) dst_tile = p.part_local + part_slot(p, ctx.m_blk * g.tiles_n + ctx.n_blk);   // local store; the switch fetches it
        else dst_tile = reinterpret_cast<__nv_bfloat16*>(p.inbox[st.owner]) + inbox_slot(p, st.k_local, p.rank);
        const int lane = ctx.lane;
        uint4* stage = reinterpret_cast<uint4*>(ctx.stage);
#pragma unroll 1
        for (int cp = 0; cp < 2; ++cp) {                             // 64 columns = one 128-byte line per row
            uint32_t a0[32], a1[32];
            tmem_ld_32x32(taddr + cp * 64, a0);
            tmem_ld_32x32(taddr + cp * 64 + 32, a1);
            tmem_ld_wait();
#pragma unroll
            for (int pc = 0; pc < 4; ++pc) {                         // row-owner layout -> swizzled smem
                stage[lane * 8 + (pc ^ (lane & 7))] = pack_row8(&a0[pc * 8]);
                stage[lane * 8 + ((pc + 4) ^ (lane & 7))] = pack_row8(&a1[pc * 8]);
            }
            __syncwarp();
            const int col_in_tile = ctx.half * (BN / 2) + cp * 64;
#pragma unroll
            for (int i = 0; i < 8; ++i) {                            // 8 lanes per row: whole 128-byte lines
                const int r = i * 4 + (lane >> 3);
                const int pc = lane & 7;
                st_u4(dst_tile + (size_t)(ctx.q * 32 + r) * BN + col_in_tile + pc * 8, stage[r * 8 + (pc ^ (r & 7))]);
            }
            __syncwarp();
        }
    }

    // ---- after the accumulator is handed back: publish the push ---------------------------------------------
    // Pair kernel: the first half (ctx.h == 0) only notes what it pushed; the second half publishes both, so that the
    // group barrier and the system-scope fence come after BOTH accumulators are back with the MMA warp.
    static __device__ __forceinline__ void publish(const Params& p, const EpiCtx& ctx, State& st, bool current) {
        epi_group_sync();                                            // all 8 warps' stores (both halves) are issued
        if (ctx.ewarp == 0 && ctx.lane == 0 && (current || st.pend_owner >= 0)) {
            fence_acq_rel_sys();
            if (st.pend_owner >= 0) st_release_sys(p.flags[st.pend_owner] + (size_t)st.pend_k * MAX_WORLD_F + p.rank, p.seq);
            if (current) st_release_sys(p.flags[st.owner] + (size_t)st.k_local * MAX_WORLD_F + p.rank, p.seq);
        }
        st.pend_owner = -1;
    }
    static __device__ __forceinline__ void after_tile(const Params& p, const GemmCore&, const EpiCtx& ctx, State& st) {
        if (ctx.h == 0) {
            st.pend_owner = st.owner;
            st.pend_k = st.k_local;
            return;
        }
        publish(p, ctx, st, true);
    }
    // a pair kernel's CTA whose 128-row half lies beyond M: nothing was pushed for it, but the epilogue warps must stay
    // in step (one group barrier per pair tile)
    static __device__ __forceinline__ void after_skipped_tile(const Params& p, const GemmCore&, const EpiCtx& ctx, State& st) {
        if (ctx.h == 0) {
            st.pend_owner = -1;
            return;
        }
        publish(p, ctx, st, false);
    }

    static __device__ __forceinline__ void tail(const Params& p, const GemmCore& g, const EpiCtx& ctx, State&) {
        if (p.broadcast) {                                           // every tile of MY output has been delivered
            const int num_tiles = g.tiles_m * g.tiles_n;
            for (int t = g.cta + ctx.ewarp * g.ncta; t < num_tiles; t += g.ncta * EPI_WARPS)
                if (ctx.lane == 0) wait_flag_ge(p.flags[p.rank] + p.ready_base + t, p.seq, g.err, 0x700);
        }
    }

    static __device__ __forceinline__ bool arrived(const Params& p, int k_local, int lane) {
        const uint32_t* f = p.flags[p.rank] + (size_t)k_local * MAX_WORLD_F;
        bool ok = true;
        if (lane < p.world) ok = (int32_t)(ld_acquire_sys(f + lane) - p.seq) >= 0;
        return __all_sync(0xffffffffu, ok);
    }

    // Side warps: reduce this CTA's owned tiles (k = cta + j * ncta) as their partials land - in ANY order (ranks may walk
    // their tiles in different orders), concurrently with the TMA/MMA/epilogue warps.  Warp 0 finds the next complete tile
    // and announces it through `scratch[0]`; all side warps then reduce their rows of it.  Returns when all are done.
    static __device__ __forceinline__ void side_loop(const Params& p, const GemmCore& g, int swarp, int lane, uint32_t* scratch) {
        const int n_mine = g.cta < p.owned_count ? (p.owned_count - 1 - g.cta) / g.ncta + 1 : 0;
        volatile uint32_t* slot = scratch;
        unsigned long long done = 0ull;                              // (leader only)
        int remaining = n_mine;
        while (true) {
            if (swarp == 0) {
                int found = -1;
                if (remaining > 0) {
                    long long t0 = clock64();
                    unsigned int backoff = 256, misses = 0;
                    int cand = 0;
                    while (((done >> cand) & 1ull) != 0ull) ++cand;  // lowest tile not yet reduced: the likeliest to be complete
                    while (found < 0) {
                        if (arrived(p, g.cta + cand * g.ncta, lane)) { found = cand; break; }
                        if ((++misses & 3u) == 0u) {                 // every 4th miss: any other tile (row-sharded walks differ per rank)
                            for (int j = cand + 1; j < n_mine; ++j) {
                                if ((done >> j) & 1ull) continue;
                                if (arrived(p, g.cta + j * g.ncta, lane)) { found = j; break; }
                            }
                            if (found >= 0) break;
                        }
                        __nanosleep(backoff);
                        backoff = backoff < 2048 ? backoff * 2 : 2048;
                        if ((clock64() - t0) > B200_FLAG_TIMEOUT_CYCLES) device_fail(g.err, ERR_FLAG_TIMEOUT, 0x680);
                    }
                    done |= 1ull << found;
                    --remaining;
                }
                if (lane == 0) *slot = (uint32_t)found;
            }
            side_group_sync(32 * EXTRA_WARPS);                       // the announcement is visible to every side warp
            const int j = (int)*slot;
            if (j < 0) break;
            const int k = g.cta + j * g.ncta;
            reduce_rows(p, g, k, swarp, lane);
            side_group_sync(32 * EXTRA_WARPS);                       // every side warp's result stores are issued; slot reusable
            if (swarp == 0 && p.broadcast) {                         // publish the finished tile to every rank
                int m_blk, n_blk;
                owned_tile_coords(p, g, k, m_blk, n_blk);
                if (lane < p.world) {
                    fence_acq_rel_sys();                             // ... and ordered system-wide before the flag
                    st_release_sys(p.flags[lane] + p.ready_base + m_blk * g.tiles_n + n_blk, p.seq);
                }
            }
        }
    }

    // one side warp's 64 rows of owned tile k: sum the `world` partials, apply the op, store (and broadcast).
    // Two rows at a time with every source's load issued before the first add, so up to 2*world 16-byte loads are
    // in flight per lane instead of one.
    static __device__ __forceinline__ void reduce_rows(const Params& p, const GemmCore& g, int k_local, int swarp, int lane) {
        constexpr int ROWS = BM / EXTRA_WARPS;
        static_assert(ROWS % 8 == 0, "the NVLS path batches 8 rows");
        int m_blk, n_blk;
        owned_tile_coords(p, g, k_local, m_blk, n_blk);
        const __nv_bfloat16* inb = reinterpret_cast<const __nv_bfloat16*>(p.inbox[p.rank]);
        const int col = n_blk * BN + lane * 8;                       // 32 lanes x 8 columns = one 256-wide row
        if (col >= g.N) return;
        if constexpr (NVLS == 1) {
            // 8 rows (8 in-switch reductions) in flight per lane: a multimem.ld_reduce is a round trip through the switch
            constexpr int NB = 8;
            const __nv_bfloat16* src = p.part_mc + part_slot(p, m_blk * g.tiles_n + n_blk) + lane * 8;
#pragma unroll 1
            for (int r0 = 0; r0 < ROWS; r0 += NB) {
                uint4 v[NB];
#pragma unroll
                for (int rr = 0; rr < NB; ++rr) v[rr] = multimem_ld_reduce_bf16x8(src + (size_t)(swarp * ROWS + r0 + rr) * BN);
#pragma unroll
                for (int rr = 0; rr < NB; ++rr) {
                    const int grow = m_blk * BM + swarp * ROWS + r0 + rr;
                    if (grow >= g.M) continue;
                    const size_t off = (size_t)(grow - p.out_row_offset) * p.ldo + col;
                    uint4 o = v[rr];
                    if (p.op == FOP_SGD) {                           // in-place form: out = out - alpha * sum (residual / bias pre-filled)
                        const uint4 ww = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.out[p.rank]) + off);
                        o.x = pack_bf16x2(bf16_lo(ww.x) - p.alpha * bf16_lo(o.x), bf16_hi(ww.x) - p.alpha * bf16_hi(o.x));
                        o.y = pack_bf16x2(bf16_lo(ww.y) - p.alpha * bf16_lo(o.y), bf16_hi(ww.y) - p.alpha * bf16_hi(o.y));
                        o.z = pack_bf16x2(bf16_lo(ww.z) - p.alpha * bf16_lo(o.z), bf16_hi(ww.z) - p.alpha * bf16_hi(o.z));
                        o.w = pack_bf16x2(bf16_lo(ww.w) - p.alpha * bf16_lo(o.w), bf16_hi(ww.w) - p.alpha * bf16_hi(o.w));
                    }
                    if (p.broadcast) multimem_st_bf16x8(p.out_mc + off, o);
                    else st_u4(reinterpret_cast<__nv_bfloat16*>(p.out[p.rank]) + off, o);
                }
            }
            return;
        }
        constexpr int RB = OPT == 1 ? 1 : 2;                         // Adam keeps 16 more floats live per row
#pragma unroll 1
        for (int r0 = 0; r0 < ROWS; r0 += RB) {
            uint4 v[RB][MAX_WORLD_F];
#pragma unroll
            for (int rr = 0; rr < RB; ++rr) {
                const int row_in_tile = swarp * ROWS + r0 + rr;
#pragma unroll
                for (int s = 0; s < MAX_WORLD_F; ++s)
                    if (s < p.world) v[rr][s] = ld_cv_u4(inb + inbox_slot(p, k_local, s) + (size_t)row_in_tile * BN + lane * 8);
            }
            uint4 w[RB];
#pragma unroll
            for (int rr = 0; rr < RB; ++rr) {
                const int grow = m_blk * BM + swarp * ROWS + r0 + rr;
                if ((OPT == 1 || p.op == FOP_SGD) && grow < g.M)
                    w[rr] = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.out[p.rank]) + (size_t)(grow - p.out_row_offset) * p.ldo + col);
            }
#pragma unroll
            for (int rr = 0; rr < RB; ++rr) {
                const int grow = m_blk * BM + swarp * ROWS + r0 + rr;
                if (grow >= g.M) continue;
                float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
                for (int s = 0; s < MAX_WORLD_F; ++s) {              // fixed order: deterministic, rank-independent
                    if (s < p.world) {
                        const uint4 x = v[rr][s];
                        acc[0] += bf16_lo(x.x); acc[1] += bf16_hi(x.x); acc[2] += bf16_lo(x.y); acc[3] += bf16_hi(x.y);
                        acc[4] += bf16_lo(x.z); acc[5] += bf16_hi(x.z); acc[6] += bf16_lo(x.w); acc[7] += bf16_hi(x.w);
                    }
                }
                const size_t off = (size_t)(grow - p.out_row_offset) * p.ldo + col;
                if constexpr (OPT == 1) {
                    const size_t so = ((size_t)k_local * BM + (size_t)(swarp * ROWS + r0 + rr)) * BN + lane * 8;
                    float4* mp = reinterpret_cast<float4*>(p.adam_m + so);
                    float4* vp = reinterpret_cast<float4*>(p.adam_v + so);
                    const float4 m0 = mp[0], m1 = mp[1], v0 = vp[0], v1 = vp[1];
                    float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
                    float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
                    const uint4 ww = w[rr];
                    const float wf[8] = {bf16_lo(ww.x), bf16_hi(ww.x), bf16_lo(ww.y), bf16_hi(ww.y),
                                         bf16_lo(ww.z), bf16_hi(ww.z), bf16_lo(ww.w), bf16_hi(ww.w)};
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float gr = p.gscale * acc[j];
                        mm[j] = p.beta1 * mm[j] + (1.0f - p.beta1) * gr;
                        vv[j] = p.beta2 * vv[j] + (1.0f - p.beta2) * gr * gr;
                        acc[j] = wf[j] - p.lr * ((mm[j] / p.bc1) / (sqrtf(vv[j] / p.bc2) + p.eps));
                    }
                    mp[0] = make_float4(mm[0], mm[1], mm[2], mm[3]); mp[1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
                    vp[0] = make_float4(vv[0], vv[1], vv[2], vv[3]); vp[1] = make_float4(vv[4], vv[5], vv[6], vv[7]);
                } else if (p.op == FOP_SGD) {
                    const uint4 ww = w[rr];
                    acc[0] = bf16_lo(ww.x) - p.alpha * acc[0]; acc[1] = bf16_hi(ww.x) - p.alpha * acc[1];
                    acc[2] = bf16_lo(ww.y) - p.alpha * acc[2]; acc[3] = bf16_hi(ww.y) - p.alpha * acc[3];
                    acc[4] = bf16_lo(ww.z) - p.alpha * acc[4]; acc[5] = bf16_hi(ww.z) - p.alpha * acc[5];
                    acc[6] = bf16_lo(ww.w) - p.alpha * acc[6]; acc[7] = bf16_hi(ww.w) - p.alpha * acc[7];
                }
                uint4 o;
                if ((OPT == 1 || p.op == FOP_SGD) && p.sr_seed != 0u) {
                    const unsigned long long e0 = (unsigned long long)grow * (unsigned long long)p.ldo + (unsigned long long)col;
                    o.x = pack_bf16x2_sr(acc[0], acc[1], p.sr_seed, e0); o.y = pack_bf16x2_sr(acc[2], acc[3], p.sr_seed, e0 + 2);
                    o.z = pack_bf16x2_sr(acc[4], acc[5], p.sr_seed, e0 + 4); o.w = pack_bf16x2_sr(acc[6], acc[7], p.sr_seed, e0 + 6);
                } else {
                    o.x = pack_bf16x2(acc[0], acc[1]); o.y = pack_bf16x2(acc[2], acc[3]);
                    o.z = pack_bf16x2(acc[4], acc[5]); o.w = pack_bf16x2(acc[6], acc[7]);
                }
                if (p.broadcast) {
                    for (int d = 0; d < p.world; ++d) {
                        const int q = (p.rank + d) % p.world;        // local first, then peers, staggered
                        st_u4(reinterpret_cast<__nv_bfloat16*>(p.out[q]) + off, o);
                    }
                } else {
                    st_u4(reinterpret_cast<__nv_bfloat16*>(p.out[p.rank]) + off, o);
                }
            }
        }
    }
};

}  // namespace b200

using namespace b200;

extern "C" long long b200_fused_inbox_bytes(int owned_cap, int world) {
    return 2ll * owned_cap * world * TILE_ELEMS * 2;
}
extern "C" long long b200_fused_flag_bytes(int owned_cap, int ready_cap) {
    return ((long long)owned_cap * MAX_WORLD_F + ready_cap) * 4;
}
extern "C" int b200_fused_owned_tiles(int M, int N, int world, int owner_mode) {
    const int tm = (M + BM - 1) / BM, tn = (N + BN - 1) / BN;
    if (owner_mode == OWN_ROW_SHARDED) return (tm / world) * tn;
    return (tm * tn + world - 1) / world;
}

// Host mirrors of the scheduling maps, for CPU tests (tests/test_native_maps_cpu.py): the SAME inline functions the
// kernels use, evaluated on the host.
// out[0] = owner rank, out[1] = inbox slot (owned-tile index) on that owner, out[2..3] = the (m_blk, n_blk) that the
// owner maps that slot back to.
extern "C" int b200_debug_fused_map(int M, int N, int world, int owner_mode, int m_blk, int n_blk, int* out) {
    GemmCore g{};
    g.M = M; g.N = N;
    g.tiles_m = (M + BM - 1) / BM; g.tiles_n = (N + BN - 1) / BN;
    FusedParams p{};
    p.world = world; p.owner_mode = owner_mode;
    p.mpo = owner_mode == OWN_ROW_SHARDED ? g.tiles_m / world : 1;
    int k = 0;
    const int owner = owner_of(p, g, m_blk, n_blk, k);
    p.rank = owner;
    int mb = -1, nb = -1;
    owned_tile_coords(p, g, k, mb, nb);
    out[0] = owner; out[1] = k; out[2] = mb; out[3] = nb;
    return 0;
}
// order[i] = m_blk * tiles_n + n_blk of the i-th tile in the 1-CTA kernel's rasterised walk
extern "C" int b200_debug_tile_order(int tiles_m, int tiles_n, int group_m, int* order) {
    GemmCore g{};
    g.tiles_m = tiles_m; g.tiles_n = tiles_n; g.group_m = group_m;
    for (int t = 0; t < tiles_m * tiles_n; ++t) {
        int m = 0, n = 0;
        tile_coords(g, t, m, n);
        order[t] = m * tiles_n + n;
    }
    return 0;
}

// D = A*B^T-style GEMM (same operand conventions as b200_gemm_bf16) whose output tiles are reduced across
// `world` ranks inside the kernel.  See include/b200ffn.h.
struct AdamArgs {
    float* m;
    float* v;
    float lr, beta1, beta2, eps, gscale;
    int step;
};

// Everything one rank's call needs, shared by the real entry points and the single-launch emulation.
struct FusedCall {
    const void* A; long long lda; int a_major;
    const void* B; long long ldb; int b_major;
    int M, N, K;
    void* const* out_peers; long long ldo; int out_row_offset;
    void* const* inbox_peers; void* const* flag_peers;
    int owned_cap, ready_cap, rank, world;
    unsigned int seq;
    int owner_mode, op, broadcast;
    float alpha;
    int max_ctas;
    const AdamArgs* adam;
    void* part_local = nullptr;      // NVLS form: all three set (see FusedParams)
    const void* part_mc = nullptr;
    void* out_mc = nullptr;
};

// Runtime choice of the kernel variant (b200_fused_config): side warps 2 (384 threads, 168 registers) or 4 (448 threads,
// 128 registers); overlap -1 = decide per call, 0 = CTA-pair mainloop, 1 = double-buffered 1-CTA mainloop.
static int g_side_warps = -1, g_overlap = -2;
static void fused_config_defaults() {
    if (g_side_warps < 0) {
        const char* e = getenv("B200_FUSED_SIDE_WARPS");
        g_side_warps = (e != nullptr && e[0] == '2') ? 2 : 4;
    }
    if (g_overlap == -2) {
        const char* e = getenv("B200_FUSED_OVERLAP");
        g_overlap = e == nullptr ? -1 : (e[0] == '1' ? 1 : (e[0] == '0' ? 0 : -1));
    }
}
extern "C" int b200_fused_config(int side_warps, int overlap) {
    fused_config_defaults();
    if (side_warps == 2 || side_warps == 4) g_side_warps = side_warps;
    if (overlap >= -1 && overlap <= 1) g_overlap = overlap;
    return g_side_warps * 10 + (g_overlap + 1);
}

static int fused_build(const FusedCall& c, FusedParams* out, int* grid_out) {
    if (c.world < 1 || c.world > MAX_WORLD_F || c.rank < 0 || c.rank >= c.world) { set_last_error("gemm_allreduce: bad rank/world"); return B200_ERR_ARG; }
    const bool nvls = c.part_mc != nullptr;
    if (c.out_peers == nullptr || (!nvls && c.inbox_peers == nullptr) || c.flag_peers == nullptr) { set_last_error("gemm_allreduce: null peer table"); return B200_ERR_ARG; }
    if (nvls && (c.part_local == nullptr || (c.broadcast && c.out_mc == nullptr) || c.adam != nullptr || c.world < 2 ||
                 (reinterpret_cast<uintptr_t>(c.part_local) & 15) || (reinterpret_cast<uintptr_t>(c.part_mc) & 15) || (reinterpret_cast<uintptr_t>(c.out_mc) & 15))) {
        set_last_error("gemm_allreduce_nvls: needs 16-byte aligned part_local / part_mc (and out_mc when broadcasting), world >= 2, no Adam"); return B200_ERR_ARG;
    }
    if (c.seq == 0) { set_last_error("gemm_allreduce: seq must start at 1"); return B200_ERR_ARG; }
    if (c.ldo % 8 || c.ldo < c.N) { set_last_error("gemm_allreduce: ldo must be >= N and a multiple of 8"); return B200_ERR_ARG; }
    const int tm = (c.M + BM - 1) / BM, tn = (c.N + BN - 1) / BN;
    FusedParams p{};
    p.ldo = c.ldo; p.out_row_offset = c.out_row_offset; p.rank = c.rank; p.world = c.world; p.seq = c.seq; p.parity = (int)(c.seq & 1u);
    p.owner_mode = c.owner_mode; p.op = c.op; p.broadcast = c.broadcast; p.alpha = c.alpha; p.owned_cap = c.owned_cap;
    if (c.owner_mode == OWN_ROW_SHARDED) {
        if (tm % c.world || (c.M % (BM * c.world))) { set_last_error("gemm_allreduce: row-sharded owners need M %% (128*world) == 0 (M=%d world=%d)", c.M, c.world); return B200_ERR_ARG; }
        p.mpo = tm / c.world;
        p.owned_count = p.mpo * tn;
    } else {
        p.mpo = 1;
        const int total = tm * tn;
        p.owned_count = total / c.world + (c.rank < total % c.world ? 1 : 0);
    }
    if (p.owned_count > c.owned_cap || b200_fused_owned_tiles(c.M, c.N, c.world, c.owner_mode) > c.owned_cap) { set_last_error("gemm_allreduce: inbox too small (%d owned tiles, capacity %d)", p.owned_count, c.owned_cap); return B200_ERR_ARG; }
    if (c.broadcast && tm * tn > c.ready_cap) { set_last_error("gemm_allreduce: ready-flag capacity %d < %d tiles", c.ready_cap, tm * tn); return B200_ERR_ARG; }
    p.ready_base = c.owned_cap * MAX_WORLD_F;
    p.part_local = reinterpret_cast<__nv_bfloat16*>(c.part_local);
    p.part_mc = reinterpret_cast<const __nv_bfloat16*>(c.part_mc);
    p.out_mc = reinterpret_cast<__nv_bfloat16*>(c.out_mc);
    p.part_cap = c.ready_cap;
    // weight updates only (alpha > 0); the in-place ADD form of the activations (alpha = -1) always rounds to nearest
    p.sr_seed = (c.adam != nullptr || (c.op == FOP_SGD && c.alpha > 0.0f)) ? update_rounding_seed() : 0u;
    if (nvls && tm * tn > c.ready_cap) { set_last_error("gemm_allreduce_nvls: partial-tile capacity %d < %d tiles", c.ready_cap, tm * tn); return B200_ERR_ARG; }
    for (int i = 0; i < c.world; ++i) {
        p.inbox[i] = nvls ? nullptr : c.inbox_peers[i];
        p.flags[i] = reinterpret_cast<uint32_t*>(c.flag_peers[i]);
        const bool need_out = (c.broadcast && !nvls) || i == c.rank;       // NVLS broadcasts through out_mc; only its own copy is read
        p.out[i] = need_out ? c.out_peers[i] : nullptr;
        if ((!nvls && p.inbox[i] == nullptr) || p.flags[i] == nullptr || (need_out && p.out[i] == nullptr)) { set_last_error("gemm_allreduce: null peer pointer for rank %d", i); return B200_ERR_ARG; }
        if ((reinterpret_cast<uintptr_t>(p.inbox[i]) & 15) || (need_out && (reinterpret_cast<uintptr_t>(p.out[i]) & 15))) { set_last_error("gemm_allreduce: peer pointers must be 16-byte aligned"); return B200_ERR_ARG; }
    }
    // every rank must use the same grid: owned tile k is reduced by CTA k %% grid on its owner
    int grid_cap = c.max_ctas > 0 ? c.max_ctas : num_sms();
    int grid = tm * tn < grid_cap ? tm * tn : grid_cap;
    if ((p.owned_count + grid - 1) / grid > MAX_OWNED_PER_CTA) { set_last_error("gemm_allreduce: too many owned tiles per CTA"); return B200_ERR_ARG; }
    *out = p;
    *grid_out = grid;
    return 0;
}

static int fused_build_adam(const FusedCall& c, const FusedParams& p, FusedAdamParams* ap) {
    const AdamArgs* adam = c.adam;
    if (adam->m == nullptr || adam->v == nullptr || adam->step < 1 || (reinterpret_cast<uintptr_t>(adam->m) & 15) || (reinterpret_cast<uintptr_t>(adam->v) & 15)) {
        set_last_error("gemm_allreduce_adam: moments must be 16-byte aligned device pointers and step >= 1"); return B200_ERR_ARG;
    }
    static_cast<FusedParams&>(*ap) = p;
    ap->op = FOP_SGD;                      // (unused by the Adam instantiation; keeps the weight-tile load path identical)
    ap->adam_m = adam->m; ap->adam_v = adam->v;
    ap->lr = adam->lr; ap->beta1 = adam->beta1; ap->beta2 = adam->beta2; ap->eps = adam->eps; ap->gscale = adam->gscale;
    ap->bc1 = 1.0f - powf(adam->beta1, (float)adam->step);
    ap->bc2 = 1.0f - powf(adam->beta2, (float)adam->step);
    return 0;
}

// Link-bound calls (NVLink time comparable to the MMA time, e.g. the MP all-reduce at 8 ranks) see bursts of store
// back-pressure in the epilogue; the kernel with two accumulator stages absorbs them under the next tile's MMAs,
// while compute-bound calls (the dW GEMMs) are faster on the 256x256-per-CTA pair kernel.
static bool fused_pick_overlap(const FusedCall& c) {
    fused_config_defaults();
    if (g_overlap >= 0) return g_overlap == 1;
    if (c.world <= 1) return false;
    const double bytes = (c.broadcast ? 2.0 : 1.0) * (c.world - 1) / c.world * (double)c.M * c.N * 2.0;
    const double link_ms = bytes / 770e9 * 1e3, mma_ms = 2.0 * c.M * c.N * (double)c.K / 1.4e15 * 1e3;
    return link_ms > 0.6 * mma_ms;
}

static int fused_entry(const FusedCall& c, void* stream) {
    FusedParams p{};
    int grid = 0;
    if (int rc = fused_build(c, &p, &grid)) return rc;
    GemmOperands o{c.A, c.lda, c.a_major, c.B, c.ldb, c.b_major, c.M, c.N, c.K};
    // Interleaved owners already spread consecutive tiles over ranks; row-sharded owners would make every rank push to
    // the same owner at once, so there each rank starts its walk at a different offset.
    const int rot = c.owner_mode == OWN_ROW_SHARDED ? c.rank : 0;
    const bool overlap = fused_pick_overlap(c);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (c.adam != nullptr) {
        FusedAdamParams ap{};
        if (int rc = fused_build_adam(c, p, &ap)) return rc;
        if (g_side_warps == 2) return launch_gemm<FusedEpiT<2, 0, 1>>(o, ap, c.max_ctas, st, rot, c.world, overlap);
        return launch_gemm<FusedEpiT<4, 32, 1>>(o, ap, c.max_ctas, st, rot, c.world, overlap);
    }
    if (c.part_mc != nullptr) return launch_gemm<FusedEpiT<4, 32, 0, 1>>(o, p, c.max_ctas, st, rot, c.world, overlap);
    if (g_side_warps == 2) return launch_gemm<FusedEpiT<2, 0>>(o, p, c.max_ctas, st, rot, c.world, overlap);
    return launch_gemm<FusedEpiT<4, 32>>(o, p, c.max_ctas, st, rot, c.world, overlap);
}

// NVLS form of b200_gemm_allreduce_bf16 (op SUM or the in-place add; no optimizer update): partial tiles stay in this
// rank's `part_local` ([2][ready_cap][128*256] bf16 = b200_fused_partial_bytes(ready_cap)), a block that every rank has
// bound to ONE multicast object whose address is `part_mc`; `out_mc` is the multicast address of the output (needed when
// broadcast != 0; out_peers[rank] is still the local copy).  Needs NVLS (b200_vmm_query).
extern "C" long long b200_fused_partial_bytes(int ready_cap) { return 2ll * ready_cap * TILE_ELEMS * 2; }
extern "C" int b200_gemm_allreduce_nvls_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                                             int M, int N, int K, void* const* out_peers, long long ldo, int out_row_offset,
                                             void* part_local, const void* part_mc, void* out_mc, void* const* flag_peers,
                                             int owned_cap, int ready_cap, int rank, int world, unsigned int seq, int owner_mode,
                                             int op, int broadcast, float alpha, int max_ctas, void* stream) {
    FusedCall c{A, lda, a_major, B, ldb, b_major, M, N, K, out_peers, ldo, out_row_offset, nullptr, flag_peers, owned_cap, ready_cap,
                rank, world, seq, owner_mode, op, broadcast, alpha, max_ctas, nullptr};
    c.part_local = part_local; c.part_mc = part_mc; c.out_mc = out_mc;
    if (part_mc == nullptr) { set_last_error("gemm_allreduce_nvls: part_mc is null"); return B200_ERR_ARG; }
    return fused_entry(c, stream);
}

extern "C" int b200_gemm_allreduce_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                                        int M, int N, int K, void* const* out_peers, long long ldo, int out_row_offset,
                                        void* const* inbox_peers, void* const* flag_peers, int owned_cap, int ready_cap,
                                        int rank, int world, unsigned int seq, int owner_mode, int op, int broadcast,
                                        float alpha, int max_ctas, void* stream) {
    FusedCall c{A, lda, a_major, B, ldb, b_major, M, N, K, out_peers, ldo, out_row_offset, inbox_peers, flag_peers, owned_cap, ready_cap,
                rank, world, seq, owner_mode, op, broadcast, alpha, max_ctas, nullptr};
    return fused_entry(c, stream);
}

// Same kernel with the Adam update in the owner's reduction: out = W - lr * mhat / (sqrt(vhat) + eps) with the gradient
// gscale * sum_ranks(A B^T).  `adam_m` / `adam_v`: LOCAL fp32 moments of this rank's owned tiles, compact layout
// [owned slot][128][256] = b200_fused_adam_state_bytes(owned_cap) bytes each, zero-initialised by the caller; `step` >= 1.
extern "C" long long b200_fused_adam_state_bytes(int owned_cap) { return (long long)owned_cap * TILE_ELEMS * 4; }
extern "C" int b200_gemm_allreduce_adam_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                                             int M, int N, int K, void* const* out_peers, long long ldo, int out_row_offset,
                                             void* const* inbox_peers, void* const* flag_peers, int owned_cap, int ready_cap,
                                             int rank, int world, unsigned int seq, int owner_mode, int broadcast,
                                             float* adam_m, float* adam_v, float lr, float beta1, float beta2, float eps, int step,
                                             float grad_scale, int max_ctas, void* stream) {
    AdamArgs a{adam_m, adam_v, lr, beta1, beta2, eps, grad_scale, step};
    FusedCall c{A, lda, a_major, B, ldb, b_major, M, N, K, out_peers, ldo, out_row_offset, inbox_peers, flag_peers, owned_cap, ready_cap,
                rank, world, seq, owner_mode, FOP_SGD, broadcast, 1.0f, max_ctas, &a};
    return fused_entry(c, stream);
}

// ---- single-launch emulation of `world` ranks on ONE GPU (tests, ncu) ------------------------------------------------
// Every rank's call is built exactly as above (its own operands, owner maps, tile rotation); the `world` kernels then run
// as slices of one cooperative grid (gemm_tcgen05_emu_kernel).  Per-rank inputs are arrays of `world` entries; the peer
// tables are the same for every rank, as in a real run.  `table` = device scratch of b200_emu_table_bytes() bytes.
extern "C" long long b200_emu_table_bytes(void) { return EMU_TABLE_BYTES; }

extern "C" int b200_emu_gemm_allreduce_bf16(int world, const void* const* A, long long lda, int a_major, const void* const* B, long long ldb,
                                            int b_major, int M, int N, int K, void* const* out_peers, long long ldo,
                                            void* const* inbox_peers, void* const* flag_peers, int owned_cap, int ready_cap,
                                            unsigned int seq, int owner_mode, int op, int broadcast, float alpha,
                                            float* const* adam_m, float* const* adam_v, float lr, float beta1, float beta2, float eps,
                                            int step, float grad_scale, int ctas_per_rank, void* table, void* stream) {
    if (world < 1 || world > MAX_WORLD_F || A == nullptr || B == nullptr || table == nullptr) { set_last_error("emu_gemm_allreduce: bad arguments"); return B200_ERR_ARG; }
    if (ctas_per_rank < 2 || (ctas_per_rank & 1) || world * ctas_per_rank > num_sms()) { set_last_error("emu_gemm_allreduce: ctas_per_rank must be even and world * ctas_per_rank <= %d", num_sms()); return B200_ERR_ARG; }
    fused_config_defaults();
    const bool adam = adam_m != nullptr;
    GemmOperands ops[MAX_WORLD_F];
    FusedParams ps[MAX_WORLD_F];
    FusedAdamParams aps[MAX_WORLD_F];
    AdamArgs aa[MAX_WORLD_F];
    int rots[MAX_WORLD_F];
    bool overlap = false;
    for (int r = 0; r < world; ++r) {
        const int row0 = (owner_mode == OWN_ROW_SHARDED && !broadcast) ? r * (M / world) : 0;
        aa[r] = AdamArgs{adam ? adam_m[r] : nullptr, adam ? adam_v[r] : nullptr, lr, beta1, beta2, eps, grad_scale, step};
        FusedCall c{A[r], lda, a_major, B[r], ldb, b_major, M, N, K, out_peers, ldo, row0, inbox_peers, flag_peers, owned_cap, ready_cap,
                    r, world, seq, owner_mode, op, broadcast, alpha, ctas_per_rank, adam ? &aa[r] : nullptr};
        int grid = 0;
        if (int rc = fused_build(c, &ps[r], &grid)) return rc;
        if (!broadcast) {                   // a rank only ever dereferences its own output pointer; the table holds every rank's
            for (int i = 0; i < world; ++i) ps[r].out[i] = i == r ? out_peers[i] : nullptr;
        }
        if (adam) { if (int rc = fused_build_adam(c, ps[r], &aps[r])) return rc; }
        ops[r] = GemmOperands{A[r], lda, a_major, B[r], ldb, b_major, M, N, K};
        rots[r] = owner_mode == OWN_ROW_SHARDED ? r : 0;
        overlap = fused_pick_overlap(c);
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool pair = use_pair_kernel() && !overlap;
    if (adam) {
        if (g_side_warps == 2) return launch_gemm_emu<FusedEpiT<2, 0, 1>>(world, ops, aps, rots, ctas_per_rank, pair, 0, table, st);
        return launch_gemm_emu<FusedEpiT<4, 32, 1>>(world, ops, aps, rots, ctas_per_rank, pair, 0, table, st);
    }
    if (g_side_warps == 2) return launch_gemm_emu<FusedEpiT<2, 0>>(world, ops, ps, rots, ctas_per_rank, pair, 0, table, st);
    return launch_gemm_emu<FusedEpiT<4, 32>>(world, ops, ps, rots, ctas_per_rank, pair, 0, table, st);
}
