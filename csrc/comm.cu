// Peer-memory communication runtime: symmetric allocations shared with CUDA IPC (one process per GPU) or
// direct peer access (one process, many GPUs), a device-side cross-rank barrier, and standalone all-reduce /
// all-gather / reduce-scatter kernels that move data with loads/stores on NVLink-mapped peer pointers - no NCCL.
//
// Replaces the transport under the reference's collectives (train_ffns.py:165,203,255,303,309; test_nccl.py:16,
// 25,35) and the CUDA-IPC hand-off it gets from torch.multiprocessing (train_ffns.py:186; test_mp_barrier_gpus.py).
//
// Synchronisation protocol: every rank owns a flag array `flags[block][src_rank]` (uint32) inside a symmetric
// allocation.  Block b of rank r signals peer p by storing a monotonically increasing value to
// flags_p[b][r] with st.release.sys and waits until its own flags[b][q] >= value for every q (ld.acquire.sys).
// Values never reset, so there is no reset race; each kernel consumes three values per call (`3*epoch+k`).
// Index->block assignment is identical in every phase and on every rank, so a per-block barrier is enough to
// order "peer q finished phase 1 of the elements I am about to read".
#include "peer_sync.cuh"
#include "../include/b200ffn.h"

namespace b200 {

constexpr int COLL_THREADS = 256;   // small footprint: must be able to co-reside with a 320-thread GEMM CTA on every SM
constexpr int COLL_MAX_BLOCKS = COLL_FLAG_BLOCKS;   // one 256-thread CTA per SM at most (they co-reside with GEMM CTAs)

__global__ void __launch_bounds__(32) barrier_kernel(PeerTable flags, int rank, int world, uint32_t value, unsigned int* err) {
    block_barrier(flags, rank, world, value, err);
}

__device__ __forceinline__ void add8(float (&acc)[8], const uint4& v) {
    acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x); acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
    acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z); acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
}
// fp32 payloads (the reference's test_nccl.py uses fp32): a 16-byte vector is 4 floats
__device__ __forceinline__ void add4(float (&acc)[8], const uint4& v) {
    acc[0] += __uint_as_float(v.x); acc[1] += __uint_as_float(v.y); acc[2] += __uint_as_float(v.z); acc[3] += __uint_as_float(v.w);
}
__device__ __forceinline__ uint4 pack4f(const float (&f)[8]) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
}
__device__ __forceinline__ uint4 pack8f(const float (&f)[8]) {
    uint4 o;
    o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
    o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
    return o;
}

// Copy vectors [lo, hi) with CU independent 16-byte loads in flight per thread (peer reads have ~2 us latency: one
// load at a time would cap a 32-CTA kernel near 50 GB/s).
constexpr int CU = 8;
__device__ __forceinline__ void copy_batched(uint4* dst, const uint4* src, long long lo, long long hi, long long t0, long long tstride) {
    for (long long i = lo + t0; i < hi; i += tstride * CU) {
        uint4 v[CU];
#pragma unroll
        for (int u = 0; u < CU; ++u) {
            const long long idx = i + u * tstride;
            if (idx < hi) v[u] = ld_cv_u4(src + idx);
        }
#pragma unroll
        for (int u = 0; u < CU; ++u) {
            const long long idx = i + u * tstride;
            if (idx < hi) st_u4(dst + idx, v[u]);
        }
    }
}

// slice s of V 16-byte vectors split over `world` ranks
__device__ __forceinline__ void slice_bounds(long long V, int world, int s, long long& lo, long long& hi) {
    lo = V * s / world;
    hi = V * (s + 1) / world;
}

// MODE 0: all-reduce (two-shot: reduce my slice from all peers, publish, gather the other slices)
// MODE 1: reduce-scatter (first half only; result stays in my slice of my buffer)
// MODE 2: all-gather (second half only; my slice of my buffer is the input)
// MODE 3: one-shot all-reduce for latency-bound sizes (every rank sums all peers; <= 1 vector per thread)
template <int MODE, bool F32>
__global__ void __launch_bounds__(COLL_THREADS) coll_kernel(PeerTable buf, long long offset_bytes, long long V, PeerTable flags,
                                                            int rank, int world, uint32_t epoch, unsigned int* err) {
    const long long tstride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ uint8_t* sbuf[MAX_WORLD];                        // dynamic indexing of kernel params would spill
    if (threadIdx.x < MAX_WORLD) sbuf[threadIdx.x] = reinterpret_cast<uint8_t*>(buf.p[threadIdx.x]);
    __syncthreads();
    uint4* mine = reinterpret_cast<uint4*>(sbuf[rank] + offset_bytes);

    block_barrier(flags, rank, world, 3 * epoch + 1, err);     // every peer's kernel has started: inputs are final

    if (MODE == 3) {
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const bool active = t0 < V;
        if (active) {
            for (int q = 0; q < world; ++q) {
                const uint4* src = reinterpret_cast<const uint4*>(sbuf[q] + offset_bytes);
                if (F32) add4(acc, ld_cv_u4(src + t0)); else add8(acc, ld_cv_u4(src + t0));
            }
        }
        block_barrier(flags, rank, world, 3 * epoch + 2, err); // everyone has read everything: safe to overwrite
        if (active) st_u4(mine + t0, F32 ? pack4f(acc) : pack8f(acc));
        return;
    }

    if (MODE == 0 || MODE == 1) {
        long long lo, hi;
        slice_bounds(V, world, rank, lo, hi);
        constexpr int RU = 2;
        for (long long i = lo + t0; i < hi; i += tstride * RU) {
            uint4 v[RU][MAX_WORLD];
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                const long long idx = i + u * tstride;
#pragma unroll
                for (int q = 0; q < MAX_WORLD; ++q)
                    if (q < world && idx < hi) v[u][q] = ld_cv_u4(reinterpret_cast<const uint4*>(sbuf[q] + offset_bytes) + idx);
            }
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                const long long idx = i + u * tstride;
                if (idx >= hi) continue;
                float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
                for (int q = 0; q < MAX_WORLD; ++q)            // fixed order: deterministic sum
                    if (q < world) { if (F32) add4(acc, v[u][q]); else add8(acc, v[u][q]); }
                st_u4(mine + idx, F32 ? pack4f(acc) : pack8f(acc));
            }
        }
    }
    if (MODE == 0) block_barrier(flags, rank, world, 3 * epoch + 2, err);   // all slices reduced and visible
    if (MODE == 0 || MODE == 2) {
        for (int dq = 1; dq < world; ++dq) {
            const int q = (rank + dq) % world;                 // staggered: no two ranks hammer the same peer
            long long lo, hi;
            slice_bounds(V, world, q, lo, hi);
            const uint4* src = reinterpret_cast<const uint4*>(sbuf[q] + offset_bytes);
            copy_batched(mine, src, lo, hi, t0, tstride);
        }
    }
    block_barrier(flags, rank, world, 3 * epoch + 3, err);     // peers are done reading my buffer
}

// All-gather INTO a local buffer: dst[q*shard .. (q+1)*shard) <- peer q's shard (src_peers[q] + src_offset).
// Used for the FSDP parameter unshard (train_ffns.py:203): the shards stay where the optimizer updates them and
// the gathered copy lands directly in the contiguous buffer the GEMM's TMA descriptor reads (no cat, :209).
__global__ void __launch_bounds__(COLL_THREADS) gather_into_kernel(uint4* dst, PeerTable src, long long src_offset_bytes, long long Vs,
                                                                   PeerTable flags, int rank, int world, uint32_t epoch, unsigned int* err) {
    const long long tstride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ uint8_t* sbuf[MAX_WORLD];
    if (threadIdx.x < MAX_WORLD) sbuf[threadIdx.x] = reinterpret_cast<uint8_t*>(src.p[threadIdx.x]);
    __syncthreads();
    block_barrier(flags, rank, world, 3 * epoch + 1, err);     // every peer's shard is final (its kernel has started)
    for (int dq = 0; dq < world; ++dq) {
        const int q = (rank + dq) % world;
        const uint4* s = reinterpret_cast<const uint4*>(sbuf[q] + src_offset_bytes);
        uint4* d = dst + (long long)q * Vs;
        copy_batched(d, s, 0, Vs, t0, tstride);
    }
    block_barrier(flags, rank, world, 3 * epoch + 3, err);     // peers have finished reading my shard
}

static int make_table(PeerTable* t, void* const* peers, int world, const char* what) {
    if (world < 1 || world > MAX_WORLD) { set_last_error("%s: world must be 1..%d", what, MAX_WORLD); return B200_ERR_ARG; }
    if (peers == nullptr) { set_last_error("%s: null peer table", what); return B200_ERR_ARG; }
    for (int i = 0; i < MAX_WORLD; ++i) t->p[i] = i < world ? peers[i] : nullptr;
    for (int i = 0; i < world; ++i)
        if (t->p[i] == nullptr || (reinterpret_cast<uintptr_t>(t->p[i]) & 15)) { set_last_error("%s: peer %d pointer null or misaligned", what, i); return B200_ERR_ARG; }
    return 0;
}

static int coll_launch(int mode, bool f32, void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                       int world, unsigned int epoch, int blocks, void* stream, const char* what) {
    PeerTable buf, flags;
    if (int rc = make_table(&buf, peers, world, what)) return rc;
    if (int rc = make_table(&flags, flag_peers, world, what)) return rc;
    if (rank < 0 || rank >= world) { set_last_error("%s: bad rank", what); return B200_ERR_ARG; }
    if (bytes <= 0 || (bytes % 16) || (offset_bytes % 16) || offset_bytes < 0) { set_last_error("%s: offset and size must be multiples of 16 bytes", what); return B200_ERR_ARG; }
    const long long V = bytes / 16;
    if (blocks <= 0) {
        const long long want = (V / world + COLL_THREADS * 4 - 1) / (COLL_THREADS * 4);
        blocks = (int)(want < 1 ? 1 : (want > COLL_MAX_BLOCKS ? COLL_MAX_BLOCKS : want));
    }
    if (blocks > COLL_MAX_BLOCKS) blocks = COLL_MAX_BLOCKS;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    unsigned int* err = device_error_slot();
#define B200_COLL(MODE_, GRID_)                                                                                              \
    do {                                                                                                                      \
        if (f32) coll_kernel<MODE_, true><<<GRID_, COLL_THREADS, 0, st>>>(buf, offset_bytes, V, flags, rank, world, epoch, err); \
        else coll_kernel<MODE_, false><<<GRID_, COLL_THREADS, 0, st>>>(buf, offset_bytes, V, flags, rank, world, epoch, err);   \
    } while (0)
    if (mode == 0 && V <= 16 * COLL_THREADS) {                 // latency-bound: one-shot, <= 1 vector per thread
        const int b = (int)((V + COLL_THREADS - 1) / COLL_THREADS);
        B200_COLL(3, b);
    } else if (mode == 0) {
        B200_COLL(0, blocks);
    } else if (mode == 1) {
        B200_COLL(1, blocks);
    } else {
        B200_COLL(2, blocks);
    }
#undef B200_COLL
    return check_cuda(cudaGetLastError(), what);
}

}  // namespace b200

using namespace b200;

// ---- symmetric memory ---------------------------------------------------------------------------------------
extern "C" int b200_symm_alloc(void** ptr, long long bytes) {
    if (ptr == nullptr || bytes <= 0) { set_last_error("symm_alloc: bad arguments"); return B200_ERR_ARG; }
    void* p = nullptr;
    if (int rc = check_cuda(cudaMalloc(&p, (size_t)bytes), "symm_alloc cudaMalloc")) return rc;
    if (int rc = check_cuda(cudaMemset(p, 0, (size_t)bytes), "symm_alloc cudaMemset")) { cudaFree(p); return rc; }
    if (int rc = check_cuda(cudaDeviceSynchronize(), "symm_alloc sync")) { cudaFree(p); return rc; }
    *ptr = p;
    return 0;
}
extern "C" int b200_symm_free(void* ptr) { return ptr ? check_cuda(cudaFree(ptr), "symm_free") : 0; }
extern "C" int b200_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
extern "C" int b200_ipc_get_handle(void* handle_out, void* ptr) {
    if (handle_out == nullptr || ptr == nullptr) { set_last_error("ipc_get_handle: null argument"); return B200_ERR_ARG; }
    return check_cuda(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle_out), ptr), "cudaIpcGetMemHandle");
}
extern "C" int b200_ipc_open_handle(void** peer_ptr, const void* handle) {
    if (peer_ptr == nullptr || handle == nullptr) { set_last_error("ipc_open_handle: null argument"); return B200_ERR_ARG; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    return check_cuda(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
}
extern "C" int b200_ipc_close_handle(void* peer_ptr) { return peer_ptr ? check_cuda(cudaIpcCloseMemHandle(peer_ptr), "cudaIpcCloseMemHandle") : 0; }
extern "C" int b200_enable_peer_access(int peer_device) {
    int cur = 0, can = 0;
    cudaGetDevice(&cur);
    if (peer_device == cur) return 0;
    if (int rc = check_cuda(cudaDeviceCanAccessPeer(&can, cur, peer_device), "cudaDeviceCanAccessPeer")) return rc;
    if (!can) { set_last_error("device %d cannot access peer %d", cur, peer_device); return B200_ERR_STATE; }
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
    return check_cuda(e, "cudaDeviceEnablePeerAccess");
}

// ---- collectives ---------------------------------------------------------------------------------------------
extern "C" int b200_barrier(void* const* flag_peers, int rank, int world, unsigned int epoch, void* stream) {
    PeerTable flags;
    if (int rc = make_table(&flags, flag_peers, world, "barrier")) return rc;
    barrier_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flags, rank, world, 3 * epoch + 1, device_error_slot());
    return check_cuda(cudaGetLastError(), "barrier launch");
}
extern "C" int b200_all_reduce_bf16(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                                    int world, unsigned int epoch, int blocks, void* stream) {
    return coll_launch(0, false, peers, offset_bytes, bytes, flag_peers, rank, world, epoch, blocks, stream, "all_reduce");
}
extern "C" int b200_reduce_scatter_bf16(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers,
                                        int rank, int world, unsigned int epoch, int blocks, void* stream) {
    return coll_launch(1, false, peers, offset_bytes, bytes, flag_peers, rank, world, epoch, blocks, stream, "reduce_scatter");
}
extern "C" int b200_all_reduce_f32(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                                   int world, unsigned int epoch, int blocks, void* stream) {
    return coll_launch(0, true, peers, offset_bytes, bytes, flag_peers, rank, world, epoch, blocks, stream, "all_reduce_f32");
}
extern "C" int b200_reduce_scatter_f32(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers,
                                       int rank, int world, unsigned int epoch, int blocks, void* stream) {
    return coll_launch(1, true, peers, offset_bytes, bytes, flag_peers, rank, world, epoch, blocks, stream, "reduce_scatter_f32");
}
extern "C" int b200_all_gather(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                               int world, unsigned int epoch, int blocks, void* stream) {
    return coll_launch(2, false, peers, offset_bytes, bytes, flag_peers, rank, world, epoch, blocks, stream, "all_gather");
}
extern "C" int b200_all_gather_into(void* dst, void* const* src_peers, long long src_offset_bytes, long long shard_bytes,
                                    void* const* flag_peers, int rank, int world, unsigned int epoch, int blocks, void* stream) {
    PeerTable src, flags;
    if (int rc = make_table(&src, src_peers, world, "all_gather_into")) return rc;
    if (int rc = make_table(&flags, flag_peers, world, "all_gather_into")) return rc;
    if (dst == nullptr || (reinterpret_cast<uintptr_t>(dst) & 15) || shard_bytes <= 0 || (shard_bytes % 16) || (src_offset_bytes % 16)) {
        set_last_error("all_gather_into: dst/offset/size must be 16-byte aligned"); return B200_ERR_ARG;
    }
    if (rank < 0 || rank >= world) { set_last_error("all_gather_into: bad rank"); return B200_ERR_ARG; }
    const long long Vs = shard_bytes / 16;
    if (blocks <= 0) {
        const long long want = (Vs + COLL_THREADS * 4 - 1) / (COLL_THREADS * 4);
        blocks = (int)(want < 1 ? 1 : (want > COLL_MAX_BLOCKS ? COLL_MAX_BLOCKS : want));
    }
    if (blocks > COLL_MAX_BLOCKS) blocks = COLL_MAX_BLOCKS;
    gather_into_kernel<<<blocks, COLL_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(reinterpret_cast<uint4*>(dst), src, src_offset_bytes, Vs,
                                                                                           flags, rank, world, epoch, device_error_slot());
    return check_cuda(cudaGetLastError(), "all_gather_into launch");
}
// Same contract as b200_all_gather_into, but the bytes move on the COPY ENGINES (peer-to-peer DMA over NVLink): no SM
// is used, so the gather overlaps a GEMM that owns every SM's registers and shared memory.  Two one-warp barrier
// kernels order it against the peers: before (every peer's shard is final) and after (every peer has finished
// reading mine).  This is the FSDP parameter prefetch (train_ffns.py:203, :236-249).
extern "C" int b200_all_gather_into_dma(void* dst, void* const* src_peers, long long src_offset_bytes, long long shard_bytes,
                                        void* const* flag_peers, int rank, int world, unsigned int epoch, void* stream) {
    PeerTable src, flags;
    if (int rc = make_table(&src, src_peers, world, "all_gather_into_dma")) return rc;
    if (int rc = make_table(&flags, flag_peers, world, "all_gather_into_dma")) return rc;
    if (dst == nullptr || shard_bytes <= 0 || src_offset_bytes < 0) { set_last_error("all_gather_into_dma: bad arguments"); return B200_ERR_ARG; }
    if (rank < 0 || rank >= world) { set_last_error("all_gather_into_dma: bad rank"); return B200_ERR_ARG; }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    unsigned int* err = device_error_slot();
    barrier_kernel<<<1, 32, 0, st>>>(flags, rank, world, 3 * epoch + 1, err);
    if (int rc = check_cuda(cudaGetLastError(), "all_gather_into_dma barrier")) return rc;
    for (int dq = 0; dq < world; ++dq) {
        const int q = (rank + dq) % world;                 // staggered: ranks pull from different peers at any moment
        const char* s = reinterpret_cast<const char*>(src.p[q]) + src_offset_bytes;
        char* d = reinterpret_cast<char*>(dst) + (long long)q * shard_bytes;
        if (int rc = check_cuda(cudaMemcpyAsync(d, s, (size_t)shard_bytes, cudaMemcpyDeviceToDevice, st), "all_gather_into_dma copy")) return rc;
    }
    barrier_kernel<<<1, 32, 0, st>>>(flags, rank, world, 3 * epoch + 3, err);
    return check_cuda(cudaGetLastError(), "all_gather_into_dma barrier");
}
extern "C" long long b200_coll_flag_bytes(void) { return (long long)COLL_MAX_BLOCKS * MAX_WORLD * 4; }
