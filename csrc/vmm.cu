// EXPERIMENTAL (round 1: compiled and unit-tested on the host side only; the device side has GPU tests queued).
//
// cuMem VMM symmetric blocks + NVLS multicast objects - what the in-switch reduction instructions
// (multimem.ld_reduce / multimem.st) need and what plain CUDA-IPC blocks (comm.cu) cannot give:
//   * a block is created with cuMemCreate as a POSIX-fd-exportable allocation, mapped locally, and its fd is handed
//     to the peers (SCM_RIGHTS over a unix socket, below), who import and map it: same peer-table contract as the
//     IPC path;
//   * rank 0 creates a multicast object for `world` devices, every rank adds its device and binds its block to it,
//     and maps the multicast handle: a store to that VA lands in every replica, a multimem.ld_reduce on it returns
//     the switch-side sum of all replicas.
// SURVEY.md 5.8 lists this as the target runtime; DESIGN.md section 8 lists it as the next step for the link-bound
// MP all-reduce.  Nothing in the default training paths depends on this file.
#include "peer_sync.cuh"
#include "../include/b200ffn.h"
#include <cudaTypedefs.h>
#include <cerrno>
#include <cstring>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

namespace b200 {

// ---------------------------------------------------------------------------------------------
// file-descriptor passing between the ranks of one node (SCM_RIGHTS)
// ---------------------------------------------------------------------------------------------
static int fill_addr(sockaddr_un* addr, const char* path) {
    memset(addr, 0, sizeof(*addr));
    addr->sun_family = AF_UNIX;
    if (path == nullptr || strlen(path) == 0 || strlen(path) >= sizeof(addr->sun_path)) {
        set_last_error("fd passing: socket path empty or longer than %zu bytes", sizeof(addr->sun_path) - 1);
        return B200_ERR_ARG;
    }
    strncpy(addr->sun_path, path, sizeof(addr->sun_path) - 1);
    return 0;
}

}  // namespace b200

using namespace b200;

// Bind a datagram socket at `path` (the receiver side).  Returns the socket fd (>= 0) or a negative error.
extern "C" int b200_fd_listen(const char* path) {
    sockaddr_un addr;
    if (int rc = fill_addr(&addr, path)) return rc;
    int s = socket(AF_UNIX, SOCK_DGRAM, 0);
    if (s < 0) { set_last_error("fd_listen: socket(): %s", strerror(errno)); return B200_ERR_STATE; }
    unlink(path);
    if (bind(s, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) {
        set_last_error("fd_listen: bind(%s): %s", path, strerror(errno));
        close(s);
        return B200_ERR_STATE;
    }
    // a peer that failed before sending must not leave this rank blocked in recvmsg for ever: b200_fd_recv gives up after 60 s
    timeval tv{60, 0};
    setsockopt(s, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    return s;
}

// Send file descriptor `fd` (plus a 4-byte tag) to the socket bound at `path`.
extern "C" int b200_fd_send(const char* path, int fd, int tag) {
    sockaddr_un addr;
    if (int rc = fill_addr(&addr, path)) return rc;
    int s = socket(AF_UNIX, SOCK_DGRAM, 0);
    if (s < 0) { set_last_error("fd_send: socket(): %s", strerror(errno)); return B200_ERR_STATE; }
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    iovec iov{&tag, sizeof(tag)};
    msghdr msg{};
    msg.msg_name = &addr;
    msg.msg_namelen = sizeof(addr);
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
    const ssize_t n = sendmsg(s, &msg, 0);
    const int err = errno;
    close(s);
    if (n != (ssize_t)sizeof(tag)) { set_last_error("fd_send: sendmsg(%s): %s", path, strerror(err)); return B200_ERR_STATE; }
    return 0;
}

// Receive one descriptor on listening socket `sock`; *tag_out gets the sender's tag.  Returns the new fd.
extern "C" int b200_fd_recv(int sock, int* tag_out) {
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    int tag = 0;
    iovec iov{&tag, sizeof(tag)};
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    const ssize_t n = recvmsg(sock, &msg, 0);
    if (n != (ssize_t)sizeof(tag)) { set_last_error("fd_recv: recvmsg(): %s", strerror(errno)); return B200_ERR_STATE; }
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    if (c == nullptr || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) {
        set_last_error("fd_recv: message carried no descriptor");
        return B200_ERR_STATE;
    }
    int fd = -1;
    memcpy(&fd, CMSG_DATA(c), sizeof(int));
    if (tag_out) *tag_out = tag;
    return fd;
}

extern "C" int b200_fd_close(int fd) { return close(fd) == 0 ? 0 : B200_ERR_STATE; }

// ---------------------------------------------------------------------------------------------
// driver entry points (resolved at run time; the library keeps loading on a GPU-less box)
// ---------------------------------------------------------------------------------------------
namespace b200 {

struct VmmApi {
    PFN_cuMemCreate create = nullptr;
    PFN_cuMemRelease release = nullptr;
    PFN_cuMemAddressReserve reserve = nullptr;
    PFN_cuMemAddressFree addr_free = nullptr;
    PFN_cuMemMap map = nullptr;
    PFN_cuMemUnmap unmap = nullptr;
    PFN_cuMemSetAccess set_access = nullptr;
    PFN_cuMemExportToShareableHandle export_h = nullptr;
    PFN_cuMemImportFromShareableHandle import_h = nullptr;
    PFN_cuMemGetAllocationGranularity granularity = nullptr;
    PFN_cuMulticastCreate mc_create = nullptr;
    PFN_cuMulticastAddDevice mc_add = nullptr;
    PFN_cuMulticastBindMem mc_bind = nullptr;
    PFN_cuMulticastUnbind mc_unbind = nullptr;
    PFN_cuMulticastGetGranularity mc_gran = nullptr;
    PFN_cuDeviceGetAttribute dev_attr = nullptr;
    bool ok = false;
};

static VmmApi* vmm_api() {
    static VmmApi api;
    static bool tried = false;
    if (tried) return api.ok ? &api : nullptr;
    tried = true;
#define B200_LOAD(field, name)                                             \
    api.field = reinterpret_cast<decltype(api.field)>(driver_entry(name)); \
    if (api.field == nullptr) return nullptr;
    B200_LOAD(create, "cuMemCreate")
    B200_LOAD(release, "cuMemRelease")
    B200_LOAD(reserve, "cuMemAddressReserve")
    B200_LOAD(addr_free, "cuMemAddressFree")
    B200_LOAD(map, "cuMemMap")
    B200_LOAD(unmap, "cuMemUnmap")
    B200_LOAD(set_access, "cuMemSetAccess")
    B200_LOAD(export_h, "cuMemExportToShareableHandle")
    B200_LOAD(import_h, "cuMemImportFromShareableHandle")
    B200_LOAD(granularity, "cuMemGetAllocationGranularity")
    B200_LOAD(mc_create, "cuMulticastCreate")
    B200_LOAD(mc_add, "cuMulticastAddDevice")
    B200_LOAD(mc_bind, "cuMulticastBindMem")
    B200_LOAD(mc_unbind, "cuMulticastUnbind")
    B200_LOAD(mc_gran, "cuMulticastGetGranularity")
    B200_LOAD(dev_attr, "cuDeviceGetAttribute")
#undef B200_LOAD
    api.ok = true;
    return &api;
}

static int check_cu(CUresult r, const char* what) {
    if (r == CUDA_SUCCESS) return 0;
    set_last_error("%s failed (CUresult %d)", what, (int)r);
    return B200_ERR_DRIVER;
}

static CUmemAllocationProp block_prop(int dev) {
    CUmemAllocationProp p{};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = dev;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

// map `handle` (size bytes) at a fresh VA with read/write access for device `dev`
static int map_handle(VmmApi* a, CUmemGenericAllocationHandle handle, size_t size, size_t align, int dev, void** out) {
    CUdeviceptr va = 0;
    if (int rc = check_cu(a->reserve(&va, size, align, 0, 0), "cuMemAddressReserve")) return rc;
    if (int rc = check_cu(a->map(va, size, 0, handle, 0), "cuMemMap")) { a->addr_free(va, size); return rc; }
    CUmemAccessDesc acc{};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    if (int rc = check_cu(a->set_access(va, size, &acc, 1), "cuMemSetAccess")) { a->unmap(va, size); a->addr_free(va, size); return rc; }
    *out = reinterpret_cast<void*>(va);
    return 0;
}

}  // namespace b200

// out[0] = POSIX-fd export supported, out[1] = multicast (NVLS) supported, for the current device.
extern "C" int b200_vmm_query(int* out) {
    if (out == nullptr) { set_last_error("vmm_query: null out"); return B200_ERR_ARG; }
    out[0] = out[1] = 0;
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    int dev = 0;
    if (int rc = check_cuda(cudaGetDevice(&dev), "vmm_query cudaGetDevice")) return rc;
    if (int rc = check_cu(a->dev_attr(&out[0], CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev), "cuDeviceGetAttribute")) return rc;
    return check_cu(a->dev_attr(&out[1], CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "cuDeviceGetAttribute");
}

// Granularity (bytes) block sizes must be a multiple of; with_multicast != 0 takes the multicast object's
// recommended granularity for `world` devices into account.
extern "C" long long b200_vmm_granularity(int world, int with_multicast) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    int dev = 0;
    cudaGetDevice(&dev);
    CUmemAllocationProp p = block_prop(dev);
    size_t g = 0;
    if (check_cu(a->granularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity")) return B200_ERR_DRIVER;
    if (with_multicast) {
        CUmulticastObjectProp mp{};
        mp.numDevices = (unsigned)world;
        mp.size = g;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        if (check_cu(a->mc_gran(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity")) return B200_ERR_DRIVER;
        if (mg > g) g = mg;
    }
    return (long long)g;
}

// Create + map a shareable block on the current device.  bytes must be a multiple of b200_vmm_granularity().
// *handle_out is the allocation handle (needed for multicast binding and for b200_vmm_free), *fd_out the POSIX fd
// to pass to the peers (close it after sending).
extern "C" int b200_vmm_alloc(long long bytes, void** ptr_out, unsigned long long* handle_out, int* fd_out) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    if (bytes <= 0 || ptr_out == nullptr || handle_out == nullptr || fd_out == nullptr) { set_last_error("vmm_alloc: bad arguments"); return B200_ERR_ARG; }
    int dev = 0;
    if (int rc = check_cuda(cudaGetDevice(&dev), "vmm_alloc cudaGetDevice")) return rc;
    if (int rc = check_cuda(cudaFree(nullptr), "vmm_alloc context init")) return rc;
    CUmemAllocationProp p = block_prop(dev);
    CUmemGenericAllocationHandle h = 0;
    if (int rc = check_cu(a->create(&h, (size_t)bytes, &p, 0), "cuMemCreate")) return rc;
    void* va = nullptr;
    if (int rc = map_handle(a, h, (size_t)bytes, 0, dev, &va)) { a->release(h); return rc; }
    int fd = -1;
    if (int rc = check_cu(a->export_h(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle")) {
        a->unmap(reinterpret_cast<CUdeviceptr>(va), (size_t)bytes);
        a->addr_free(reinterpret_cast<CUdeviceptr>(va), (size_t)bytes);
        a->release(h);
        return rc;
    }
    if (int rc = check_cuda(cudaMemset(va, 0, (size_t)bytes), "vmm_alloc memset")) return rc;
    *ptr_out = va;
    *handle_out = (unsigned long long)h;
    *fd_out = fd;
    return 0;
}

// Import a peer's block from its fd and map it for the current device.
extern "C" int b200_vmm_import(int fd, long long bytes, void** ptr_out, unsigned long long* handle_out) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    if (fd < 0 || bytes <= 0 || ptr_out == nullptr || handle_out == nullptr) { set_last_error("vmm_import: bad arguments"); return B200_ERR_ARG; }
    int dev = 0;
    if (int rc = check_cuda(cudaGetDevice(&dev), "vmm_import cudaGetDevice")) return rc;
    CUmemGenericAllocationHandle h = 0;
    if (int rc = check_cu(a->import_h(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle")) return rc;
    void* va = nullptr;
    if (int rc = map_handle(a, h, (size_t)bytes, 0, dev, &va)) { a->release(h); return rc; }
    *ptr_out = va;
    *handle_out = (unsigned long long)h;
    return 0;
}

extern "C" int b200_vmm_free(void* ptr, long long bytes, unsigned long long handle) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    if (ptr != nullptr) {
        a->unmap(reinterpret_cast<CUdeviceptr>(ptr), (size_t)bytes);
        a->addr_free(reinterpret_cast<CUdeviceptr>(ptr), (size_t)bytes);
    }
    if (handle != 0) a->release((CUmemGenericAllocationHandle)handle);
    return 0;
}

// ---- NVLS multicast ------------------------------------------------------------------------------------------
// Rank 0: create the multicast object for `world` devices and export its fd.
extern "C" int b200_mc_create(int world, long long bytes, unsigned long long* mc_out, int* fd_out) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    if (world < 2 || bytes <= 0 || mc_out == nullptr || fd_out == nullptr) { set_last_error("mc_create: bad arguments"); return B200_ERR_ARG; }
    CUmulticastObjectProp mp{};
    mp.numDevices = (unsigned)world;
    mp.size = (size_t)bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle mc = 0;
    if (int rc = check_cu(a->mc_create(&mc, &mp), "cuMulticastCreate")) return rc;
    int fd = -1;
    if (int rc = check_cu(a->export_h(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle(multicast)")) { a->release(mc); return rc; }
    *mc_out = (unsigned long long)mc;
    *fd_out = fd;
    return 0;
}

// Other ranks: import the multicast object from rank 0's fd.
extern "C" int b200_mc_import(int fd, unsigned long long* mc_out) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    if (fd < 0 || mc_out == nullptr) { set_last_error("mc_import: bad arguments"); return B200_ERR_ARG; }
    CUmemGenericAllocationHandle mc = 0;
    if (int rc = check_cu(a->import_h(&mc, reinterpret_cast<void*>(static_cast<intptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle(multicast)")) return rc;
    *mc_out = (unsigned long long)mc;
    return 0;
}

// Every rank, in two collective steps separated by a host barrier: (1) add my device; (2) bind my block and map the
// multicast VA.  step 1: mem_handle/ptr_out unused.  step 2 returns the multicast pointer in *ptr_out.
extern "C" int b200_mc_attach(unsigned long long mc, int step, unsigned long long mem_handle, long long bytes, void** ptr_out) {
    VmmApi* a = vmm_api();
    if (a == nullptr) return B200_ERR_DRIVER;
    int dev = 0;
    if (int rc = check_cuda(cudaGetDevice(&dev), "mc_attach cudaGetDevice")) return rc;
    if (step == 1) return check_cu(a->mc_add((CUmemGenericAllocationHandle)mc, dev), "cuMulticastAddDevice");
    if (step != 2 || ptr_out == nullptr || bytes <= 0) { set_last_error("mc_attach: bad arguments"); return B200_ERR_ARG; }
    if (int rc = check_cu(a->mc_bind((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem_handle, 0, (size_t)bytes, 0), "cuMulticastBindMem")) return rc;
    return map_handle(a, (CUmemGenericAllocationHandle)mc, (size_t)bytes, 0, dev, ptr_out);
}

// ---- NVLS collectives (in-switch reduction / replication) ----------------------------------------------------------
// On the multicast address of a block every rank has bound:
//   all-reduce, two-shot   rank r asks the switch for the SUM over all replicas of slice r (multimem.ld_reduce, fp32
//                          accumulation: SASS LDGMC...HPADD) and stores it through the multicast address into every
//                          replica (multimem.st).  ~1x the message per GPU and direction (peer-pointer two-shot: ~2x).
//   all-reduce, one-shot   latency-bound sizes: every rank reduces the WHOLE vector through the switch and stores its own
//                          copy; no second phase.
//   reduce-scatter         first half of the two-shot form, result stored locally into slice r.
//   all-gather             rank r multimem.st's its slice: the switch replicates it into every copy.
// Same flag protocol as comm.cu (values 3*epoch+k, one barrier slot per block, identical index->block maps everywhere).
namespace b200 {

constexpr int NVLS_THREADS = 256;
constexpr int NVLS_MAX_BLOCKS = COLL_FLAG_BLOCKS;
constexpr int NVLS_UNROLL = 4;                       // in-switch reductions in flight per thread (each is a switch round trip)

template <bool F32>
__device__ __forceinline__ uint4 mm_ld_reduce(const uint4* p) {
    uint4 v;
    if constexpr (F32)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    else
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(uint4* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// MODE 0 two-shot all-reduce, 1 one-shot all-reduce, 2 reduce-scatter, 3 all-gather.  `mc` / `local`: the operand through
// the multicast address / this rank's own copy.
template <int MODE, bool F32>
__global__ void __launch_bounds__(NVLS_THREADS) nvls_coll_kernel(uint4* mc, uint4* local, long long V, PeerTable flags, int rank, int world,
                                                                 uint32_t epoch, unsigned int* err) {
    const long long tstride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long lo = 0, hi = V;
    if (MODE != 1) { lo = V * rank / world; hi = V * (rank + 1) / world; }
    block_barrier(flags, rank, world, 3 * epoch + 1, err, 0x800);          // every replica holds its input
    if (MODE == 1) {                                                       // <= NVLS_UNROLL vectors per thread (launcher's contract)
        uint4 v[NVLS_UNROLL];
#pragma unroll
        for (int u = 0; u < NVLS_UNROLL; ++u) {
            const long long i = t0 + u * tstride;
            if (i < V) v[u] = mm_ld_reduce<F32>(mc + i);
        }
        block_barrier(flags, rank, world, 3 * epoch + 2, err, 0x800);      // everyone has read everything: safe to overwrite
#pragma unroll
        for (int u = 0; u < NVLS_UNROLL; ++u) {
            const long long i = t0 + u * tstride;
            if (i < V) st_u4(local + i, v[u]);
        }
        return;
    }
    for (long long i = lo + t0; i < hi; i += tstride * NVLS_UNROLL) {
        uint4 v[NVLS_UNROLL];
#pragma unroll
        for (int u = 0; u < NVLS_UNROLL; ++u) {
            const long long idx = i + u * tstride;
            if (idx < hi) v[u] = MODE == 3 ? ld_cv_u4(local + idx) : mm_ld_reduce<F32>(mc + idx);
        }
#pragma unroll
        for (int u = 0; u < NVLS_UNROLL; ++u) {
            const long long idx = i + u * tstride;
            if (idx < hi) {
                if (MODE == 2) st_u4(local + idx, v[u]);
                else mm_st(mc + idx, v[u]);
            }
        }
    }
    block_barrier(flags, rank, world, 3 * epoch + 3, err, 0x800);          // every slice has landed in every replica / been read
}

}  // namespace b200

static int nvls_launch(int mode, bool allow_one_shot, bool f32, void* mc_ptr, void* local_ptr, long long bytes, void* const* flag_peers, int rank, int world,
                       unsigned int epoch, int blocks, void* stream, const char* what) {
    if (mc_ptr == nullptr || local_ptr == nullptr || flag_peers == nullptr || bytes <= 0 || (bytes % 16) || world < 2 || world > 8 || rank < 0 || rank >= world) {
        set_last_error("%s: bad arguments", what);
        return B200_ERR_ARG;
    }
    PeerTable flags{};
    for (int i = 0; i < world; ++i) {
        if (flag_peers[i] == nullptr) { set_last_error("%s: null flag pointer for rank %d", what, i); return B200_ERR_ARG; }
        flags.p[i] = flag_peers[i];
    }
    const long long V = bytes / 16;
    if (mode == 0 && allow_one_shot && V <= (long long)NVLS_UNROLL * NVLS_THREADS * 32) mode = 1;     // <= 512 KiB: one-shot
    if (mode == 1) {
        blocks = (int)((V + NVLS_THREADS * NVLS_UNROLL - 1) / (NVLS_THREADS * NVLS_UNROLL));
    } else if (blocks <= 0) {
        // one vector per thread until every SM has a CTA (mid sizes are latency-bound: spread them), then deeper loops
        const long long want = (V / world + NVLS_THREADS - 1) / NVLS_THREADS;
        blocks = (int)(want < 1 ? 1 : (want > NVLS_MAX_BLOCKS ? NVLS_MAX_BLOCKS : want));
    }
    if (blocks > NVLS_MAX_BLOCKS) blocks = NVLS_MAX_BLOCKS;
    if (blocks < 1) blocks = 1;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    uint4* mc = reinterpret_cast<uint4*>(mc_ptr);
    uint4* lc = reinterpret_cast<uint4*>(local_ptr);
    unsigned int* err = device_error_slot();
#define B200_NVLS(MODE_)                                                                                                      \
    do {                                                                                                                      \
        if (f32) nvls_coll_kernel<MODE_, true><<<blocks, NVLS_THREADS, 0, st>>>(mc, lc, V, flags, rank, world, epoch, err);     \
        else nvls_coll_kernel<MODE_, false><<<blocks, NVLS_THREADS, 0, st>>>(mc, lc, V, flags, rank, world, epoch, err);        \
    } while (0)
    switch (mode) {
        case 0: B200_NVLS(0); break;
        case 1: B200_NVLS(1); break;
        case 2: B200_NVLS(2); break;
        default: B200_NVLS(3); break;
    }
#undef B200_NVLS
    return check_cuda(cudaGetLastError(), what);
}

// In-place collectives on `bytes` at `mc_ptr` (multicast address, from b200_mc_attach step 2) / `local_ptr` (this rank's
// own mapping of the same bytes); `flag_peers`: an ordinary peer table of b200_coll_flag_bytes() flag blocks.
extern "C" int b200_all_reduce_nvls_bf16(void* mc_ptr, long long bytes, void* const* flag_peers, int rank, int world,
                                         unsigned int epoch, int blocks, void* stream) {
    // (round-1 signature: no local pointer - always the two-shot form, which needs none)
    return nvls_launch(0, false, false, mc_ptr, mc_ptr, bytes, flag_peers, rank, world, epoch, blocks, stream, "all_reduce_nvls");
}
extern "C" int b200_nvls_collective(int op, int f32, void* mc_ptr, void* local_ptr, long long bytes, void* const* flag_peers, int rank,
                                    int world, unsigned int epoch, int blocks, void* stream) {
    if (op < 0 || op > 2) { set_last_error("nvls_collective: op must be 0 (all-reduce), 1 (reduce-scatter) or 2 (all-gather)"); return B200_ERR_ARG; }
    return nvls_launch(op == 0 ? 0 : op + 1, true, f32 != 0, mc_ptr, local_ptr, bytes, flag_peers, rank, world, epoch, blocks, stream, "nvls_collective");
}
