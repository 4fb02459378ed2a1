// Peer tables and the cross-rank block barrier shared by the collective kernels (comm.cu) and the NVLS kernels (vmm.cu).
//
// Every rank owns a flag array `flags[block][src_rank]` (uint32) inside a symmetric allocation.  Block b of rank r
// signals peer p by storing a monotonically increasing value to flags_p[b][r] (st.release.sys after an acq_rel system
// fence) and waits until its own flags[b][q] >= value for every q (ld.acquire.sys).  Values never reset.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int MAX_WORLD = 8;
constexpr int COLL_FLAG_BLOCKS = 148;   // barrier slots per flag block = the largest grid a collective kernel may use

struct PeerTable {
    void* p[MAX_WORLD];
};

__device__ __forceinline__ void block_barrier(const PeerTable& flags, int rank, int world, uint32_t value,
                                              unsigned int* err, uint32_t tag_base = 0x500) {
    __syncthreads();                                    // every thread's prior writes are issued
    if (threadIdx.x < world) {
        fence_acq_rel_sys();                            // ... and ordered before the signal
        uint32_t* dst = reinterpret_cast<uint32_t*>(flags.p[threadIdx.x]) + blockIdx.x * MAX_WORLD + rank;
        st_release_sys(dst, value);
        const uint32_t* mine = reinterpret_cast<const uint32_t*>(flags.p[rank]) + blockIdx.x * MAX_WORLD + threadIdx.x;
        wait_flag_ge(mine, value, err, tag_base + threadIdx.x);
    }
    __syncthreads();
}

}  // namespace b200
