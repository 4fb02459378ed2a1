// All-gather -> GEMM in ONE kernel (sequence-parallel MP: the column-parallel GEMM whose A operand is the token-sharded
// activation, SURVEY.md 2.6 C4/C5 "SP form"; no counterpart in the reference, which all-reduces instead,
// train_ffns.py:303,:309).
//
//   A is [M, K] bf16, K-major, contiguous, and lives at the SAME offset of a symmetric block on every rank.  Rank q
//   holds the final values of rows [q*M/world, (q+1)*M/world) only (its token shard).  The kernel computes
//   C = epilogue(A_full * B^T) and, as a side effect, completes the local copy of A_full.
//   (Second form, FSDP backward: A is MN-major, storage [K, M], sharded along K - the contracted dimension.  The side
//   warps work exactly as below on storage rows; the producer then waits per k-block instead of per tile, see kblock.)
//
//   side warps   (2 per CTA, concurrently with the TMA / MMA / epilogue warps) pull the other ranks' shards over NVLink
//                into the local buffer, 16/32 KiB chunks in the order the GEMM will need them (ring order starting at
//                the next rank), and count finished chunks per 128-row block (red.release.gpu).
//   TMA producer before loading a tile's A rows it waits until those 128-row blocks are complete (ld.acquire.gpu +
//                a generic->async proxy fence); tiles are walked starting at the rank's own shard, so the first wave
//                needs no remote data at all and the pull runs ahead of the MMAs from then on.
//   cross-rank   one flag per (reader, owner): "owner's kernel for call `seq` has started", i.e. everything that was
//                stream-ordered before it - the kernel that produced its shard - has completed.  Nothing else is
//                exchanged; the kernel never waits for a peer to finish READING (see the reuse rules in
//                docs/KERNELS.md section 2.5 - callers keep a barrier-ed collective between reuse of a shard buffer).
//
// The same kernel serves FSDP's forward parameter all-gather (train_ffns.py:203-209): there the WEIGHT is the row-sharded
// A operand (activations are kept transposed, [features, tokens], so that every forward GEMM is W * act^T) -
// parallel/fused_fsdp_t.py.
//
// STATUS: validated on B200 in round 2 - single-GPU multi-rank emulation (tests/test_emu_gpu.py, 2/4/8 ranks, both forms) and
// 2 GPUs (tests/test_multigpu.py: mp_spf, FSDP in-kernel gather).  Selected with TrainConfig.mp_mode = "spf" / fsdp_gather = "fused".
#include "gemm_kernel.cuh"
#include "peer_sync.cuh"
#include "../include/b200ffn.h"

namespace b200 {

constexpr int GATHER_SIDE_WARPS = 2;        // 320 + 64 = 384 threads: one more warp would cap registers at 128/thread
constexpr int GATHER_BATCH = 16;            // 16-byte loads in flight per lane (peer reads: ~2-3 us round trip)

struct GatherParams {
    LocalEpiParams base;
    const uint4* src[MAX_WORLD];            // every rank's A buffer
    uint4* dst;                             // the local A buffer (= src[rank])
    uint32_t* started[MAX_WORLD];           // symmetric flag arrays: started[q][r] = last call rank r has entered
    uint32_t* ready;                        // local: finished chunks per 128-row block, zeroed before the launch
    long long shard_vecs;                   // 16-byte vectors per token shard
    int chunk_vecs;                         // vectors per pull chunk
    int chunks_per_block;                   // chunks per 128-row block
    int blocks_per_shard;                   // 128-row blocks per shard
    int rank, world;
    uint32_t seq;
    int k_sharded;                          // 0: A is K-major [M, K], shards = row blocks of M (tokens / weight rows)
                                            // 1: A is MN-major [K, M], shards = row blocks of K (FSDP backward: the contracted dim)
    int kb_rot;                             // k_sharded: first k-block of my own shard (every tile's k-loop starts there)
};

__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void red_release_gpu_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Pull order.  Chunk u of the (world-1) * chunks_per_shard remote chunks: ring distance d = 1 + u / chunks_per_shard
// (the rank after mine first: its rows follow mine in the rotated tile walk), chunk j = u % chunks_per_shard of it.
struct ChunkRef {
    int peer;
    long long first_vec;                    // vector index in the full A buffer
    int block;                              // 128-row block the chunk belongs to
};
__host__ __device__ __forceinline__ ChunkRef chunk_ref(long long u, int rank, int world, long long shard_vecs, int chunk_vecs,
                                                       int chunks_per_block, int blocks_per_shard) {
    const long long cps = shard_vecs / chunk_vecs;
    const int d = 1 + (int)(u / cps);
    const long long j = u % cps;
    ChunkRef r;
    r.peer = (rank + d) % world;
    r.first_vec = (long long)r.peer * shard_vecs + j * chunk_vecs;
    r.block = r.peer * blocks_per_shard + (int)(j / chunks_per_block);
    return r;
}

// K-sharded form: iteration kb of a tile's k-loop loads k-block kbe; `ring` is the position, counted from the start of
// my own shard, of the 128-row storage block that holds it, and `block` that block's index (= the `ready` counter).
__host__ __device__ __forceinline__ void k_walk(int kb, int kb_rot, int num_kb, int& kbe, int& ring, int& block) {
    kbe = kb + kb_rot;
    if (kbe >= num_kb) kbe -= num_kb;
    ring = (kb * BK) >> 7;
    block = (kbe * BK) >> 7;
}

template <int KIND>
struct GatherEpi {
    using Local = LocalEpi<KIND>;
    using Params = GatherParams;
    using State = typename Local::State;
    static constexpr int SMEM_BYTES = GEMM_SMEM_BYTES_STAGED;
    static constexpr int EXTRA_WARPS = GATHER_SIDE_WARPS;
    static constexpr int LAUNCH_BOUND_PAD = 0;       // 384 threads: the epilogue warps keep 168 registers
    static constexpr int PRODUCTS = 1;

    static __device__ __forceinline__ void init(const Params& p, const GemmCore& g, const EpiCtx& c, State& s) { Local::init(p.base, g, c, s); }
    static __device__ __forceinline__ void process_tile(const Params& p, const GemmCore& g, const EpiCtx& c, uint32_t taddr, State& s) {
        Local::process_tile(p.base, g, c, taddr, s);
    }
    static __device__ __forceinline__ void after_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_skipped_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void tail(const Params&, const GemmCore&, const EpiCtx&, State&) {}

    // TMA producer thread: the A rows of blocks [m_blk, m_blk + n) must be complete in the local buffer
    static __device__ __forceinline__ void before_loads(const Params& p, const GemmCore& g, int m_blk, int n) {
        if (p.world == 1 || p.k_sharded) return;
        for (int i = 0; i < n; ++i) {
            const int b = m_blk + i;
            if (b >= g.tiles_m || b / p.blocks_per_shard == p.rank) continue;      // ragged tail / my own shard
            long long t0 = clock64();
            uint32_t spins = 0;
            while (ld_acquire_gpu(p.ready + b) < (uint32_t)p.chunks_per_block) {
                __nanosleep(64);
                if (((++spins) & 0xff) == 0 && (clock64() - t0) > B200_FLAG_TIMEOUT_CYCLES) device_fail(g.err, ERR_FLAG_TIMEOUT, 0xA00 + (b & 0xff));
            }
        }
        fence_proxy_async_global();          // the rows were written with generic stores; TMA reads them via the async proxy
    }

    // K-sharded A (storage rows = k): a tile needs every shard, in any order.  Each tile's k-loop starts at my own
    // shard's k-range and walks the ring, the same order the side warps pull in; a block is waited for once (blocks
    // are visited in ring order, so "complete up to ring position p" is all the producer has to remember).
    struct ProducerState {
        int known_upto = 0;
    };
    static __device__ __forceinline__ int kblock(const Params& p, const GemmCore& g, int kb, int num_kb, ProducerState& ps) {
        if (!p.k_sharded || p.world == 1) return kb;
        int kbe, ring, b;
        k_walk(kb, p.kb_rot, num_kb, kbe, ring, b);
        if (ring >= p.blocks_per_shard && ring >= ps.known_upto) {
            long long t0 = clock64();
            uint32_t spins = 0;
            while (ld_acquire_gpu(p.ready + b) < (uint32_t)p.chunks_per_block) {
                __nanosleep(64);
                if (((++spins) & 0xff) == 0 && (clock64() - t0) > B200_FLAG_TIMEOUT_CYCLES) device_fail(g.err, ERR_FLAG_TIMEOUT, 0xB00 + (b & 0xff));
            }
            fence_proxy_async_global();
            ps.known_upto = ring + 1;
        }
        return kbe;
    }

    static __device__ __forceinline__ void side_loop(const Params& p, const GemmCore& g, int swarp, int lane, uint32_t*) {
        if (p.world == 1) return;
        if (g.cta == 0 && swarp == 0 && lane < p.world && lane != p.rank) {
            fence_acq_rel_sys();
            st_release_sys(p.started[lane] + p.rank, p.seq);     // "my shard is final" (this kernel is stream-ordered after its producer)
        }
        const long long total = (p.shard_vecs / p.chunk_vecs) * (p.world - 1);
        const long long W = (long long)g.ncta * EXTRA_WARPS;
        uint32_t seen = 0;
        for (long long u = (long long)swarp * g.ncta + g.cta; u < total; u += W) {
            const ChunkRef c = chunk_ref(u, p.rank, p.world, p.shard_vecs, p.chunk_vecs, p.chunks_per_block, p.blocks_per_shard);
            if (!((seen >> c.peer) & 1u)) {
                if (lane == 0) wait_flag_ge(p.started[p.rank] + c.peer, p.seq, g.err, 0x900 + c.peer);
                __syncwarp();
                seen |= 1u << c.peer;
            }
            const uint4* s = p.src[c.peer] + c.first_vec;
            uint4* d = p.dst + c.first_vec;
#pragma unroll 1
            for (int i = lane; i < p.chunk_vecs; i += 32 * GATHER_BATCH) {
                uint4 v[GATHER_BATCH];
#pragma unroll
                for (int b = 0; b < GATHER_BATCH; ++b)
                    if (i + b * 32 < p.chunk_vecs) v[b] = ld_cv_u4(s + i + b * 32);
#pragma unroll
                for (int b = 0; b < GATHER_BATCH; ++b)
                    if (i + b * 32 < p.chunk_vecs) st_u4(d + i + b * 32, v[b]);
            }
            fence_proxy_async_global();
            __syncwarp();                                        // every lane's stores are issued ...
            if (lane == 0) {
                fence_acq_rel_gpu();                             // ... and ordered before the count
                red_release_gpu_add(p.ready + c.block, 1u);
            }
        }
    }
};

template <int KIND>
static int launch_kind(const GemmOperands& o, const GatherParams& p, int max_ctas, cudaStream_t st) {
    if (p.k_sharded) return launch_gemm<GatherEpi<KIND>>(o, p, max_ctas, st);      // any tile order: every tile needs every shard
    // walk the tiles starting at my own shard; rasterisation groups no larger than a shard keep the walk shard by shard
    const int group_m = p.blocks_per_shard < 16 ? p.blocks_per_shard : 16;
    return launch_gemm<GatherEpi<KIND>>(o, p, max_ctas, st, p.rank, p.world, false, group_m);
}

}  // namespace b200

using namespace b200;

// Pull-chunk size for a K-wide bf16 A operand: 32 KiB when a 128-row block is a whole number of them, else 16 KiB
// (K % 64 == 0 makes a block a multiple of 16 KiB).
static int gather_chunk_bytes(int row_len) { return (256ll * row_len) % 32768 == 0 ? 32768 : 16384; }

extern "C" long long b200_gather_gemm_ready_bytes(int rows) { return (long long)((rows + BM - 1) / BM) * 4; }

// Host mirror of the pull order for CPU tests: out = {peer, first_vec, block, chunk_vecs, chunks_per_block, total_chunks}
extern "C" int b200_debug_gather_chunk(int M, int K, int rank, int world, long long u, long long* out) {
    if (world < 1 || world > MAX_WORLD || M % (BM * world) || K % BK) return B200_ERR_ARG;
    const int cb = gather_chunk_bytes(K);
    const long long shard_vecs = (long long)(M / world) * K * 2 / 16;
    const int chunk_vecs = cb / 16, cpb = (int)(256ll * K / cb), bps = M / world / BM;
    out[3] = chunk_vecs; out[4] = cpb; out[5] = (shard_vecs / chunk_vecs) * (world - 1);
    if (u < 0 || u >= out[5]) return 0;
    const ChunkRef c = chunk_ref(u, rank, world, shard_vecs, chunk_vecs, cpb, bps);
    out[0] = c.peer; out[1] = c.first_vec; out[2] = c.block;
    return 0;
}

// Host mirror of the row-sharded form's tile walk in the CTA-pair kernel: first_row[i] / n_blk[i] of the i-th pair tile
// (512 rows) that rank `rank` visits, with the rotation and rasterisation group launch_kind() chooses.
extern "C" int b200_debug_gather_tile_walk(int M, int N, int rank, int world, int* first_row, int* n_blk, int capacity) {
    if (world < 1 || world > MAX_WORLD || M % (BM * world) || rank < 0 || rank >= world) return B200_ERR_ARG;
    GemmCore g{};
    g.M = M; g.N = N;
    g.tiles_m = M / BM; g.tiles_n = (N + BN - 1) / BN;
    const int bps = M / world / BM;
    g.group_m = bps < 16 ? bps : 16;
    const int pair_tiles_m = (g.tiles_m + 3) >> 2, tiles = pair_tiles_m * g.tiles_n;
    const int offset = (int)((long long)tiles * rank / world);
    for (int i = 0; i < tiles && i < capacity; ++i) {
        int t = i + offset;
        if (t >= tiles) t -= tiles;
        int pm = 0, nb = 0;
        pair_tile_coords(g, pair_tiles_m, t, pm, nb);
        first_row[i] = pm * 4 * BM;
        n_blk[i] = nb;
    }
    return tiles;
}

// Host mirror of the K-sharded k-loop walk: out = {kbe, ring, block, kb_rot, num_kb, blocks_per_shard}
extern "C" int b200_debug_gather_kwalk(int K, int rank, int world, int kb, int* out) {
    if (world < 1 || world > MAX_WORLD || K % (BM * world) || rank < 0 || rank >= world) return B200_ERR_ARG;
    const int num_kb = K / BK, kb_rot = rank * (num_kb / world);
    out[3] = kb_rot; out[4] = num_kb; out[5] = K / world / BM;
    if (kb < 0 || kb >= num_kb) return 0;
    k_walk(kb, kb_rot, num_kb, out[0], out[1], out[2]);
    return 0;
}

struct GatherCall {
    void* const* a_peers; int a_major;
    const void* B; long long ldb; int b_major;
    void* Cout; long long ldc;
    int M, N, K, epilogue;
    const void* aux; long long ldaux; void* out2; long long ldc2;
    const float* bias; float alpha;
    void* const* started_peers; void* ready;
    int rank, world;
    unsigned int seq;
};

static int gather_build(const GatherCall& c, GatherParams* out, GemmOperands* o) {
    const int world = c.world, rank = c.rank, M = c.M, N = c.N, K = c.K, epilogue = c.epilogue;
    if (world < 1 || world > MAX_WORLD || rank < 0 || rank >= world) { set_last_error("gather_gemm: bad rank/world"); return B200_ERR_ARG; }
    if (c.a_peers == nullptr || c.started_peers == nullptr || c.ready == nullptr || c.Cout == nullptr) { set_last_error("gather_gemm: null argument"); return B200_ERR_ARG; }
    if (c.seq == 0) { set_last_error("gather_gemm: seq must start at 1"); return B200_ERR_ARG; }
    if (c.a_major != MAJOR_K && c.a_major != MAJOR_MN) { set_last_error("gather_gemm: a_major must be 0 (K) or 1 (MN)"); return B200_ERR_ARG; }
    // storage of A: `rows` rows of `row_len` contiguous elements, sharded by rows
    const int rows = c.a_major == MAJOR_K ? M : K, row_len = c.a_major == MAJOR_K ? K : M;
    if (rows % (BM * world)) { set_last_error("gather_gemm: the sharded dimension (%d) must be a multiple of 128 * world (%d)", rows, BM * world); return B200_ERR_ARG; }
    if (row_len % BK) { set_last_error("gather_gemm: the row length of A (%d) must be a multiple of 64", row_len); return B200_ERR_ARG; }
    GatherParams p{};
    for (int i = 0; i < world; ++i) {
        if (c.a_peers[i] == nullptr || c.started_peers[i] == nullptr || (reinterpret_cast<uintptr_t>(c.a_peers[i]) & 15)) {
            set_last_error("gather_gemm: peer %d pointer null or not 16-byte aligned", i); return B200_ERR_ARG;
        }
        p.src[i] = reinterpret_cast<const uint4*>(c.a_peers[i]);
        p.started[i] = reinterpret_cast<uint32_t*>(c.started_peers[i]);
    }
    p.dst = reinterpret_cast<uint4*>(c.a_peers[rank]);
    p.ready = reinterpret_cast<uint32_t*>(c.ready);
    const int cb = gather_chunk_bytes(row_len);
    p.shard_vecs = (long long)(rows / world) * row_len * 2 / 16;
    p.chunk_vecs = cb / 16;
    p.chunks_per_block = (int)(256ll * row_len / cb);
    p.blocks_per_shard = rows / world / BM;
    p.rank = rank; p.world = world; p.seq = c.seq;
    p.k_sharded = c.a_major == MAJOR_MN ? 1 : 0;
    p.kb_rot = p.k_sharded ? rank * (K / BK / world) : 0;
    p.base.C = c.Cout; p.base.ldc = c.ldc; p.base.aux = reinterpret_cast<const __nv_bfloat16*>(c.aux); p.base.ldaux = c.ldaux;
    p.base.C2 = reinterpret_cast<__nv_bfloat16*>(c.out2); p.base.ldc2 = c.ldc2; p.base.bias = c.bias; p.base.alpha = c.alpha;
    const bool needs_aux = epilogue == EPI_RELU_BWD || epilogue == EPI_GELU_BWD || epilogue == EPI_RELU_BWD_MASK || epilogue == EPI_ADD;
    const bool needs_out2 = epilogue == EPI_GELU || epilogue == EPI_RELU_MASK;
    if (needs_aux && (c.aux == nullptr || (reinterpret_cast<uintptr_t>(c.aux) & 15))) { set_last_error("gather_gemm: epilogue %d needs a 16-byte aligned aux", epilogue); return B200_ERR_ARG; }
    if (needs_out2 && (c.out2 == nullptr || (reinterpret_cast<uintptr_t>(c.out2) & 15))) { set_last_error("gather_gemm: epilogue %d needs a 16-byte aligned out2", epilogue); return B200_ERR_ARG; }
    const long long mask_pitch = ((N + 127) / 128) * 16;             // bytes per row of a packed ReLU sign mask
    if ((epilogue == EPI_RELU_BWD || epilogue == EPI_GELU_BWD || epilogue == EPI_ADD) && (c.ldaux % 8 || c.ldaux < N)) { set_last_error("gather_gemm: ldaux must be >= N and a multiple of 8"); return B200_ERR_ARG; }
    if (epilogue == EPI_GELU && (c.ldc2 % 8 || c.ldc2 < N)) { set_last_error("gather_gemm: ldc2 must be >= N and a multiple of 8"); return B200_ERR_ARG; }
    if (epilogue == EPI_RELU_MASK && (c.ldc2 % 16 || c.ldc2 < mask_pitch)) { set_last_error("gather_gemm: RELU_MASK needs out2 = uint8 mask with pitch %% 16 == 0 and >= %lld bytes", mask_pitch); return B200_ERR_ARG; }
    if (epilogue == EPI_RELU_BWD_MASK && (c.ldaux % 16 || c.ldaux < mask_pitch)) { set_last_error("gather_gemm: RELU_BWD_MASK needs aux = uint8 mask with pitch %% 16 == 0 and >= %lld bytes", mask_pitch); return B200_ERR_ARG; }
    if (reinterpret_cast<uintptr_t>(c.Cout) & 15) { set_last_error("gather_gemm: C must be 16-byte aligned"); return B200_ERR_ARG; }
    if (c.ldc % 8 || c.ldc < N) { set_last_error("gather_gemm: ldc must be >= N and a multiple of 8"); return B200_ERR_ARG; }
    *out = p;
    *o = GemmOperands{c.a_peers[rank], (long long)row_len, c.a_major, c.B, c.ldb, c.b_major, M, N, K};
    return 0;
}

extern "C" int b200_gather_gemm_bf16(void* const* a_peers, int a_major, const void* B, long long ldb, int b_major, void* Cout, long long ldc,
                                     int M, int N, int K, int epilogue, const void* aux, long long ldaux, void* out2, long long ldc2,
                                     const float* bias, float alpha, void* const* started_peers, void* ready, int rank, int world,
                                     unsigned int seq, int max_ctas, void* stream) {
    GatherCall c{a_peers, a_major, B, ldb, b_major, Cout, ldc, M, N, K, epilogue, aux, ldaux, out2, ldc2, bias, alpha, started_peers, ready, rank, world, seq};
    GatherParams p{};
    GemmOperands o{};
    if (int rc = gather_build(c, &p, &o)) return rc;
    const int rows = a_major == MAJOR_K ? M : K;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (world > 1) {
        if (int rc = check_cuda(cudaMemsetAsync(ready, 0, (size_t)b200_gather_gemm_ready_bytes(rows), st), "gather_gemm: clear ready counters")) return rc;
    }
    switch (epilogue) {
        case EPI_STORE: return launch_kind<EPI_STORE>(o, p, max_ctas, st);
        case EPI_RELU: return launch_kind<EPI_RELU>(o, p, max_ctas, st);
        case EPI_GELU: return launch_kind<EPI_GELU>(o, p, max_ctas, st);
        case EPI_RELU_BWD: return launch_kind<EPI_RELU_BWD>(o, p, max_ctas, st);
        case EPI_GELU_BWD: return launch_kind<EPI_GELU_BWD>(o, p, max_ctas, st);
        case EPI_ADD: return launch_kind<EPI_ADD>(o, p, max_ctas, st);
        case EPI_RELU_MASK: return launch_kind<EPI_RELU_MASK>(o, p, max_ctas, st);
        case EPI_RELU_BWD_MASK: return launch_kind<EPI_RELU_BWD_MASK>(o, p, max_ctas, st);
        default: set_last_error("gather_gemm: unsupported epilogue %d", epilogue); return B200_ERR_ARG;
    }
}

// ---- single-launch emulation of `world` ranks on ONE GPU (tests, ncu); see gemm_kernel.cuh -----------------------------
// Per-rank arrays of `world` entries: B (each rank's weight shard under MP, the same weight under FSDP), C, aux, out2, ready.
// a_peers / started_peers are the shared peer tables.  Epilogues: STORE, RELU_MASK, RELU_BWD_MASK, GELU (the ones the
// trainers use with this kernel).
template <int KIND>
static int emu_kind(int world, const GemmOperands* ops, const GatherParams* ps, int ctas_per_rank, void* table, cudaStream_t st) {
    int rots[MAX_WORLD];
    for (int r = 0; r < world; ++r) rots[r] = ps[0].k_sharded ? 0 : r;
    const int group_m = ps[0].k_sharded ? 0 : (ps[0].blocks_per_shard < 16 ? ps[0].blocks_per_shard : 16);
    return launch_gemm_emu<GatherEpi<KIND>>(world, ops, ps, rots, ctas_per_rank, use_pair_kernel(), group_m, table, st);
}

extern "C" int b200_emu_gather_gemm_bf16(int world, void* const* a_peers, int a_major, const void* const* B, long long ldb, int b_major,
                                         void* const* C, long long ldc, int M, int N, int K, int epilogue, const void* const* aux,
                                         long long ldaux, void* const* out2, long long ldc2, void* const* started_peers,
                                         void* const* ready, unsigned int seq, int ctas_per_rank, void* table, void* stream) {
    if (world < 1 || world > MAX_WORLD || B == nullptr || C == nullptr || ready == nullptr || table == nullptr) { set_last_error("emu_gather_gemm: bad arguments"); return B200_ERR_ARG; }
    if (ctas_per_rank < 2 || (ctas_per_rank & 1) || world * ctas_per_rank > num_sms()) { set_last_error("emu_gather_gemm: ctas_per_rank must be even and world * ctas_per_rank <= %d", num_sms()); return B200_ERR_ARG; }
    GatherParams ps[MAX_WORLD];
    GemmOperands ops[MAX_WORLD];
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int rows = a_major == MAJOR_K ? M : K;
    for (int r = 0; r < world; ++r) {
        GatherCall c{a_peers, a_major, B[r], ldb, b_major, C[r], ldc, M, N, K, epilogue, aux ? aux[r] : nullptr, ldaux, out2 ? out2[r] : nullptr, ldc2,
                     nullptr, 1.0f, started_peers, ready[r], r, world, seq};
        if (int rc = gather_build(c, &ps[r], &ops[r])) return rc;
        if (world > 1) {
            if (int rc = check_cuda(cudaMemsetAsync(ready[r], 0, (size_t)b200_gather_gemm_ready_bytes(rows), st), "emu_gather_gemm: clear ready counters")) return rc;
        }
    }
    switch (epilogue) {
        case EPI_STORE: return emu_kind<EPI_STORE>(world, ops, ps, ctas_per_rank, table, st);
        case EPI_GELU: return emu_kind<EPI_GELU>(world, ops, ps, ctas_per_rank, table, st);
        case EPI_RELU_MASK: return emu_kind<EPI_RELU_MASK>(world, ops, ps, ctas_per_rank, table, st);
        case EPI_RELU_BWD_MASK: return emu_kind<EPI_RELU_BWD_MASK>(world, ops, ps, ctas_per_rank, table, st);
        default: set_last_error("emu_gather_gemm: unsupported epilogue %d", epilogue); return B200_ERR_ARG;
    }
}
