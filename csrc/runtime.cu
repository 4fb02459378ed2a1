// Host runtime: error plumbing, driver entry points (no link-time libcuda dependency, so the library also
// loads on a GPU-less box), TMA descriptor encoding, device properties.
#include "common.cuh"
#include "../include/b200ffn.h"
#include <cudaTypedefs.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace b200 {

static thread_local char g_last_error[512] = "";

void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_last_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    return static_cast<int>(e);
}

void* driver_entry(const char* name) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || fn == nullptr) {
        cudaGetLastError();
        set_last_error("driver entry point %s unavailable", name);
        return nullptr;
    }
    return fn;
}

int encode_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer,
                        uint64_t outer_stride_elems, uint32_t box_inner, uint32_t box_outer) {
    static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
    if (encode == nullptr) {
        encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(driver_entry("cuTensorMapEncodeTiled"));
        if (encode == nullptr) return B200_ERR_DRIVER;
    }
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {outer_stride_elems * 2};   // bytes, dims 1..rank-1
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled failed (%d): base=%p inner=%llu outer=%llu stride=%llu box=%ux%u", (int)r,
                       base, (unsigned long long)inner, (unsigned long long)outer,
                       (unsigned long long)outer_stride_elems, box_inner, box_outer);
        return B200_ERR_DRIVER;
    }
    return 0;
}

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    int& c = cached[dev & 63];
    if (c == 0) {
        if (cudaDeviceGetAttribute(&c, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || c <= 0) c = 148;
    }
    return c;
}

static unsigned int* g_err_slot = nullptr;
static std::mutex g_err_mutex;

unsigned int* device_error_slot() {
    if (g_err_slot != nullptr) return g_err_slot;
    std::lock_guard<std::mutex> lock(g_err_mutex);
    if (g_err_slot == nullptr) {
        void* p = nullptr;
        if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
            cudaGetLastError();
            return nullptr;   // kernels tolerate a null slot (they still trap on timeout)
        }
        memset(p, 0, 64);
        g_err_slot = static_cast<unsigned int*>(p);
    }
    return g_err_slot;
}

static unsigned int g_update_rounding_seed = 0;
unsigned int update_rounding_seed() { return g_update_rounding_seed; }

}  // namespace b200

using namespace b200;

// Rounding of every optimizer update launched AFTER this call (process-wide): 0 = round to nearest (default), anything
// else = stochastic rounding of the new bf16 weight with that seed.  Change the seed every step; use the same seed on
// every rank (replicas then round identically).
extern "C" int b200_set_update_rounding(unsigned int seed) {
    g_update_rounding_seed = seed;
    return 0;
}

extern "C" const char* b200_last_error(void) { return g_last_error; }
extern "C" int b200_version(void) { return 100; }
extern "C" int b200_num_sms(void) { return num_sms(); }

extern "C" int b200_device_error(unsigned int* out) {
    unsigned int* s = device_error_slot();
    if (s == nullptr) { if (out) memset(out, 0, 16); return 0; }
    if (out) memcpy(out, s, 16);
    return static_cast<int>(s[0]);
}

extern "C" int b200_clear_device_error(void) {
    unsigned int* s = device_error_slot();
    if (s) memset(s, 0, 16);
    return 0;
}
