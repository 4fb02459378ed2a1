// Memory-bound helpers: optimizer steps, standalone activation kernels (the unfused forms of what the GEMM
// epilogues do), Philox data generation, casts, a deterministic metric reduction.  All are grid-stride,
// 16-byte vectorised, sized to a multiple of the SM count.
#include "common.cuh"
#include "../include/b200ffn.h"
#include <curand_kernel.h>
#include <mutex>

namespace b200 {

constexpr int EW_THREADS = 256;

static inline int ew_grid(long long nvec) {
    long long blocks = (nvec + EW_THREADS - 1) / EW_THREADS;
    long long cap = (long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
    f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 o;
    o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
    o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
    return o;
}
__device__ __forceinline__ void load_grad8(const void* g, long long i8, bool f32, float (&f)[8]) {
    if (f32) {
        const float4* p = reinterpret_cast<const float4*>(g) + 2 * i8;
        float4 a = p[0], b = p[1];
        f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
    } else {
        unpack8(reinterpret_cast<const uint4*>(g)[i8], f);
    }
}

// ------------------------------------------------------------------ SGD
__device__ __forceinline__ uint4 pack8_rounded(const float (&f)[8], uint32_t seed, unsigned long long e0) {
    if (seed == 0u) return pack8(f);
    uint4 o;
    o.x = pack_bf16x2_sr(f[0], f[1], seed, e0); o.y = pack_bf16x2_sr(f[2], f[3], seed, e0 + 2);
    o.z = pack_bf16x2_sr(f[4], f[5], seed, e0 + 4); o.w = pack_bf16x2_sr(f[6], f[7], seed, e0 + 6);
    return o;
}
__device__ __forceinline__ __nv_bfloat16 bf16_rounded(float f, uint32_t seed, unsigned long long e) {
    if (seed == 0u) return __float2bfloat16_rn(f);
    const uint32_t r = hash32(seed ^ hash32((uint32_t)(e >> 1) * 0x9e3779b9u + (uint32_t)(e >> 33)));
    return __ushort_as_bfloat16((unsigned short)bf16_bits_sr(f, (e & 1ull) ? (r >> 16) : (r & 0xffffu)));
}

__global__ void __launch_bounds__(EW_THREADS) sgd_kernel(__nv_bfloat16* p, const void* g, long long n, float lr, int grad_f32, uint32_t sr_seed) {
    const long long n8 = n / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    uint4* pv = reinterpret_cast<uint4*>(p);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        float w[8], gr[8];
        unpack8(pv[i], w);
        load_grad8(g, i, grad_f32 != 0, gr);
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] -= lr * gr[j];
        pv[i] = pack8_rounded(w, sr_seed, (unsigned long long)i * 8ull);
    }
    // scalar tail
    for (long long i = n8 * 8 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float gr = grad_f32 ? reinterpret_cast<const float*>(g)[i] : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g)[i]);
        p[i] = bf16_rounded(__bfloat162float(p[i]) - lr * gr, sr_seed, (unsigned long long)i);
    }
}

// ------------------------------------------------------------------ Adam
// 8 elements per thread: one 16-byte load/store of the bf16 parameter, 16- or 32-byte loads of the gradient, two float4
// each for the fp32 moments (26 bytes of traffic per element, all in full 32-byte sectors).
__device__ __forceinline__ float adam_one(float w, float gr, float& m, float& v, float lr, float b1, float b2, float eps, float bc1, float bc2) {
    m = b1 * m + (1.0f - b1) * gr;
    v = b2 * v + (1.0f - b2) * gr * gr;
    return w - lr * ((m / bc1) / (sqrtf(v / bc2) + eps));
}
__global__ void __launch_bounds__(EW_THREADS) adam_kernel(__nv_bfloat16* p, const void* g, float* m, float* v, long long n, float lr,
                                                          float b1, float b2, float eps, float bc1, float bc2, int grad_f32, uint32_t sr_seed) {
    const long long n8 = n / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    uint4* pv = reinterpret_cast<uint4*>(p);
    float4* mv = reinterpret_cast<float4*>(m);
    float4* vv = reinterpret_cast<float4*>(v);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        float w[8], gr[8];
        unpack8(pv[i], w);
        load_grad8(g, i, grad_f32 != 0, gr);
        const float4 m0 = mv[2 * i], m1 = mv[2 * i + 1], v0 = vv[2 * i], v1 = vv[2 * i + 1];
        float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
        float vs[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = adam_one(w[j], gr[j], mm[j], vs[j], lr, b1, b2, eps, bc1, bc2);
        mv[2 * i] = make_float4(mm[0], mm[1], mm[2], mm[3]); mv[2 * i + 1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
        vv[2 * i] = make_float4(vs[0], vs[1], vs[2], vs[3]); vv[2 * i + 1] = make_float4(vs[4], vs[5], vs[6], vs[7]);
        pv[i] = pack8_rounded(w, sr_seed, (unsigned long long)i * 8ull);
    }
    for (long long i = n8 * 8 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {   // scalar tail
        const float gr = grad_f32 ? reinterpret_cast<const float*>(g)[i] : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g)[i]);
        float mi = m[i], vi = v[i];
        p[i] = bf16_rounded(adam_one(__bfloat162float(p[i]), gr, mi, vi, lr, b1, b2, eps, bc1, bc2), sr_seed, (unsigned long long)i);
        m[i] = mi;
        v[i] = vi;
    }
}

// ------------------------------------------------------------------ fp32 parameters (biases are kept in fp32)
// OPT 0: p -= lr * g     OPT 1: Adam.  Scalar accesses: these vectors are a few thousand elements.
template <int OPT>
__global__ void __launch_bounds__(EW_THREADS) opt_f32_kernel(float* p, const float* g, float* m, float* v, long long n, float lr,
                                                             float b1, float b2, float eps, float bc1, float bc2) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if constexpr (OPT == 0) {
            p[i] -= lr * g[i];
        } else {
            float mi = m[i], vi = v[i];
            p[i] = adam_one(p[i], g[i], mi, vi, lr, b1, b2, eps, bc1, bc2);
            m[i] = mi;
            v[i] = vi;
        }
    }
}

// ------------------------------------------------------------------ column sums (bias gradient: db = sum_t dy[t, :])
// Stage 1: CTA (cb, rc) sums rows [rc * rows_per_chunk, ...) of columns [cb * 256, +256): thread = (8-column group, row
// lane); 16-byte loads, whole 512-byte row segments per warp-pair.  Stage 2: one thread per column adds the chunk
// partials in a fixed order (deterministic).  HBM-bound: reads the matrix once.
constexpr int CS_COLS = 256, CS_ROWLANES = EW_THREADS / (CS_COLS / 8);   // 32 column groups x 8 row lanes
__global__ void __launch_bounds__(EW_THREADS) colsum_partial_kernel(float* partial, const __nv_bfloat16* x, long long ld, int T, int N,
                                                                    int rows_per_chunk) {
    __shared__ float sh[CS_ROWLANES][CS_COLS + 8];
    const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int col = blockIdx.x * CS_COLS + cg * 8;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(T, r0 + rows_per_chunk);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (col < N) {
        for (int r = r0 + rl; r < r1; r += CS_ROWLANES) {
            float f[8];
            unpack8(ld_nc_u4(x + (long long)r * ld + col), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[rl][cg * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;                                        // 256 threads = 256 columns
    if (blockIdx.x * CS_COLS + c < N) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < CS_ROWLANES; ++k) t += sh[k][c];
        partial[(long long)blockIdx.y * N + blockIdx.x * CS_COLS + c] = t;
    }
}
__global__ void __launch_bounds__(EW_THREADS) colsum_final_kernel(float* out, const float* partial, int N, int chunks, int accumulate) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= N) return;
    float t = accumulate ? out[c] : 0.f;
    for (int k = 0; k < chunks; ++k) t += partial[(long long)k * N + c];
    out[c] = t;
}

// ------------------------------------------------------------------ transpose (bf16 [R, C] -> [C, R])
// 64 x 64 tiles through shared memory: 128-byte row segments on both the read and the write side.  Used where a trainer
// keeps activations feature-major (FSDP with the parameter all-gather inside the GEMMs, parallel/fused_fsdp_t.py).
__global__ void __launch_bounds__(256) transpose_kernel(__nv_bfloat16* out, long long ldo, const __nv_bfloat16* in, long long ldi, int R, int C) {
    __shared__ __nv_bfloat16 tile[64][66];
    const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;            // 32 x 8
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = r0 + ty + i * 8, c = c0 + tx * 2;
        __nv_bfloat162 v = __floats2bfloat162_rn(0.f, 0.f);
        if (r < R && c + 1 < C) v = *reinterpret_cast<const __nv_bfloat162*>(in + (long long)r * ldi + c);
        else if (r < R && c < C) v.x = in[(long long)r * ldi + c];
        tile[ty + i * 8][tx * 2] = v.x;
        tile[ty + i * 8][tx * 2 + 1] = v.y;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = c0 + ty + i * 8, r = r0 + tx * 2;                // output row = input column
        if (c >= C) continue;
        __nv_bfloat162 v;
        v.x = tile[tx * 2][ty + i * 8];
        v.y = tile[tx * 2 + 1][ty + i * 8];
        if (r + 1 < R) *reinterpret_cast<__nv_bfloat162*>(out + (long long)c * ldo + r) = v;
        else if (r < R) out[(long long)c * ldo + r] = v.x;
    }
}

// ------------------------------------------------------------------ out[t, :] = bias (+ residual[t, :])
// Pre-fills the output of a fused row-parallel GEMM + all-reduce whose in-place mode then adds the cross-rank sum.
__global__ void __launch_bounds__(EW_THREADS) bias_rows_kernel(__nv_bfloat16* out, long long ldo, const __nv_bfloat16* res, long long ldr,
                                                               const float* bias, int T, int N) {
    const int n8 = N / 8;
    const long long total = (long long)T * n8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int t = (int)(i / n8), c = (int)(i % n8) * 8;
        float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (res != nullptr) unpack8(*reinterpret_cast<const uint4*>(res + (long long)t * ldr + c), f);
        if (bias != nullptr) {
            const float4 b0 = *reinterpret_cast<const float4*>(bias + c), b1 = *reinterpret_cast<const float4*>(bias + c + 4);
            f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w; f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
        }
        *reinterpret_cast<uint4*>(out + (long long)t * ldo + c) = pack8(f);
    }
}

// ------------------------------------------------------------------ activations / add
template <int OP>   // 0 relu fwd (out=relu(x)), 1 relu bwd (out=out*(x>0)), 2 add (out=a+b)
__global__ void __launch_bounds__(EW_THREADS) ew_kernel(__nv_bfloat16* out, const __nv_bfloat16* a, const __nv_bfloat16* b, long long n) {
    const long long n8 = n / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        float x[8], y[8];
        unpack8(reinterpret_cast<const uint4*>(a)[i], x);
        if (OP == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) y[j] = x[j] <= 0.f ? 0.f : x[j];
        } else if (OP == 1) {
            unpack8(reinterpret_cast<const uint4*>(out)[i], y);
#pragma unroll
            for (int j = 0; j < 8; ++j) y[j] = x[j] <= 0.f ? 0.f : y[j];
        } else {
            unpack8(reinterpret_cast<const uint4*>(b)[i], y);
#pragma unroll
            for (int j = 0; j < 8; ++j) y[j] += x[j];
        }
        reinterpret_cast<uint4*>(out)[i] = pack8(y);
    }
    for (long long i = n8 * 8 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float x = __bfloat162float(a[i]);
        float y;
        if (OP == 0) y = x <= 0.f ? 0.f : x;
        else if (OP == 1) y = x <= 0.f ? 0.f : __bfloat162float(out[i]);
        else y = x + __bfloat162float(b[i]);
        out[i] = __float2bfloat16_rn(y);
    }
}

// ------------------------------------------------------------------ Philox normal generator
__global__ void __launch_bounds__(EW_THREADS) randn_kernel(__nv_bfloat16* out, long long n, unsigned long long seed, float scale) {
    const long long n8 = (n + 7) / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        curandStatePhilox4_32_10_t st;
        curand_init(seed, /*subsequence=*/(unsigned long long)i, /*offset=*/0, &st);
        const float4 a = curand_normal4(&st);
        const float4 b = curand_normal4(&st);
        const float f[8] = {a.x * scale, a.y * scale, a.z * scale, a.w * scale, b.x * scale, b.y * scale, b.z * scale, b.w * scale};
        if (i * 8 + 8 <= n) {
            reinterpret_cast<uint4*>(out)[i] = pack8(f);
        } else {
            for (int j = 0; i * 8 + j < n; ++j) out[i * 8 + j] = __float2bfloat16_rn(f[j]);
        }
    }
}

// ------------------------------------------------------------------ casts
__global__ void __launch_bounds__(EW_THREADS) cast_f32_bf16_kernel(__nv_bfloat16* out, const float* in, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = __float2bfloat16_rn(in[i]);
}
__global__ void __launch_bounds__(EW_THREADS) cast_bf16_f32_kernel(float* out, const __nv_bfloat16* in, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = __bfloat162float(in[i]);
}

// ------------------------------------------------------------------ deterministic sum of squares
constexpr int SUMSQ_MAX_BLOCKS = 1184;   // 148 SMs x 8
struct SumsqScratch {
    unsigned int arrived;                // must be 0 on entry; reset to 0 by the finishing CTA
    unsigned int pad[3];
    float partial[SUMSQ_MAX_BLOCKS];
};

__device__ __forceinline__ float block_sum(float v, float* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    if (threadIdx.x < 32) {
        r = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    }
    __syncthreads();
    return r;   // valid in thread 0
}

__device__ __forceinline__ void sumsq_body(float* out, const __nv_bfloat16* x, long long n, SumsqScratch* sc) {
    __shared__ float sh[EW_THREADS / 32];
    __shared__ bool is_last;
    const long long n8 = n / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        float f[8];
        unpack8(ld_nc_u4(reinterpret_cast<const uint4*>(x) + i), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
    }
    for (long long i = n8 * 8 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float f = __bfloat162float(x[i]);
        acc += f * f;
    }
    const float bsum = block_sum(acc, sh);
    if (threadIdx.x == 0) {
        sc->partial[blockIdx.x] = bsum;
        __threadfence();
        const unsigned int prev = atomicAdd(&sc->arrived, 1u);
        is_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float t = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) t += __ldcg(&sc->partial[i]);   // fixed order per thread
        const float total = block_sum(t, sh);
        if (threadIdx.x == 0) {
            out[0] = total;
            sc->arrived = 0;   // ready for the next call on this scratch
        }
    }
}

__global__ void __launch_bounds__(EW_THREADS) sumsq_kernel(float* out, const __nv_bfloat16* x, long long n, SumsqScratch* sc) {
    sumsq_body(out, x, n, sc);
}

// ------------------------------------------------------------------ sum of squares without any allocation
// Scratch for b200_sumsq_bf16_static: one arrival counter + one partial per CTA, for at most 296 CTAs (two per SM on
// a 148-SM part).  It is part of the library image: every GPU that loads libb200ffn.so carries these 1280 bytes
// until the process exits.  All calls on a device share it.
constexpr int SUMSQ_STATIC_MAX_BLOCKS = 296;
constexpr long long SUMSQ_STATIC_MAX_ELEMS = 1ll << 31;
constexpr int SUMSQ_STATIC_BYTES = 1280;                     // 16-byte header + 296 floats, rounded up to 256
static_assert(16 + SUMSQ_STATIC_MAX_BLOCKS * 4 <= SUMSQ_STATIC_BYTES, "static scratch too small");
static __device__ __align__(256) unsigned char g_sumsq_static_scratch[SUMSQ_STATIC_BYTES];

// In host code `g_sumsq_static_scratch` is only a handle for the runtime's symbol table, not a device address, so the
// address each device uses is looked up once, with that device current, and cached here.
static void* g_sumsq_static_addr[64] = {nullptr};
static std::mutex g_sumsq_static_mutex;

static int sumsq_static_address(void** out) {
    int dev = 0;
    if (int rc = check_cuda(cudaGetDevice(&dev), "sumsq_static cudaGetDevice")) return rc;
    if (dev < 0 || dev >= 64) { set_last_error("sumsq_static: device ordinal %d out of range", dev); return B200_ERR_ARG; }
    std::lock_guard<std::mutex> lock(g_sumsq_static_mutex);
    if (g_sumsq_static_addr[dev] == nullptr) {
        void* p = nullptr;
        if (int rc = check_cuda(cudaGetSymbolAddress(&p, g_sumsq_static_scratch), "sumsq_static cudaGetSymbolAddress")) return rc;
        if (p == nullptr) { set_last_error("sumsq_static: cudaGetSymbolAddress returned a null address"); return B200_ERR_STATE; }
        g_sumsq_static_addr[dev] = p;
    }
    *out = g_sumsq_static_addr[dev];
    return 0;
}

__global__ void __launch_bounds__(EW_THREADS) fill_zero_kernel(uint4* buf, long long n16) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) buf[i] = z;
}

static int check_ptr16(const void* p, const char* name) {
    if (p == nullptr) { set_last_error("%s: null pointer", name); return B200_ERR_ARG; }
    if (reinterpret_cast<uintptr_t>(p) & 15) { set_last_error("%s: pointer must be 16-byte aligned", name); return B200_ERR_ARG; }
    return 0;
}

}  // namespace b200

using namespace b200;
#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int b200_sgd_bf16(void* param, const void* grad, long long n, float lr, int grad_f32, void* stream) {
    if (n <= 0) return 0;
    if (int rc = check_ptr16(param, "sgd param")) return rc;
    if (int rc = check_ptr16(grad, "sgd grad")) return rc;
    sgd_kernel<<<ew_grid(n / 8 + 1), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(param), grad, n, lr, grad_f32, update_rounding_seed());
    return check_cuda(cudaGetLastError(), "sgd launch");
}

extern "C" int b200_adam_bf16(void* param, const void* grad, float* m, float* v, long long n, float lr, float beta1, float beta2,
                              float eps, int step, int grad_f32, void* stream) {
    if (n <= 0) return 0;
    if (param == nullptr || grad == nullptr || m == nullptr || v == nullptr || step < 1) { set_last_error("adam: bad arguments"); return B200_ERR_ARG; }
    if (int rc = check_ptr16(param, "adam param")) return rc;
    if (int rc = check_ptr16(grad, "adam grad")) return rc;
    if (int rc = check_ptr16(m, "adam m")) return rc;
    if (int rc = check_ptr16(v, "adam v")) return rc;
    const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
    adam_kernel<<<ew_grid(n / 8 + 1), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(param), grad, m, v, n, lr, beta1, beta2, eps, bc1, bc2, grad_f32, update_rounding_seed());
    return check_cuda(cudaGetLastError(), "adam launch");
}

extern "C" int b200_sgd_f32(float* param, const float* grad, long long n, float lr, void* stream) {
    if (n <= 0) return 0;
    if (param == nullptr || grad == nullptr) { set_last_error("sgd_f32: null pointer"); return B200_ERR_ARG; }
    opt_f32_kernel<0><<<ew_grid(n), EW_THREADS, 0, ST(stream)>>>(param, grad, nullptr, nullptr, n, lr, 0.f, 0.f, 0.f, 1.f, 1.f);
    return check_cuda(cudaGetLastError(), "sgd_f32 launch");
}

extern "C" int b200_adam_f32(float* param, const float* grad, float* m, float* v, long long n, float lr, float beta1, float beta2,
                             float eps, int step, void* stream) {
    if (n <= 0) return 0;
    if (param == nullptr || grad == nullptr || m == nullptr || v == nullptr || step < 1) { set_last_error("adam_f32: bad arguments"); return B200_ERR_ARG; }
    const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
    opt_f32_kernel<1><<<ew_grid(n), EW_THREADS, 0, ST(stream)>>>(param, grad, m, v, n, lr, beta1, beta2, eps, bc1, bc2);
    return check_cuda(cudaGetLastError(), "adam_f32 launch");
}

// chunks of rows such that the first stage launches about 4 CTAs per SM
static int colsum_chunks(int T, int N) {
    const int col_blocks = (N + CS_COLS - 1) / CS_COLS;
    int chunks = (num_sms() * 4 + col_blocks - 1) / col_blocks;
    const int max_chunks = (T + 63) / 64;                             // at least 64 rows per chunk
    if (chunks > max_chunks) chunks = max_chunks;
    return chunks < 1 ? 1 : chunks;
}
extern "C" long long b200_colsum_scratch_bytes(int T, int N) { return (long long)colsum_chunks(T, N) * N * 4; }

extern "C" int b200_colsum_bf16(float* out, const void* x, long long ld, int T, int N, int accumulate, void* scratch,
                                long long scratch_bytes, void* stream) {
    if (T <= 0 || N <= 0) return 0;
    if (out == nullptr || scratch == nullptr) { set_last_error("colsum: null out/scratch"); return B200_ERR_ARG; }
    if (int rc = check_ptr16(x, "colsum x")) return rc;
    if (N % 8 || ld % 8 || ld < N) { set_last_error("colsum: N and ld must be multiples of 8 and ld >= N"); return B200_ERR_ARG; }
    if (scratch_bytes < b200_colsum_scratch_bytes(T, N)) { set_last_error("colsum: scratch must hold %lld bytes", b200_colsum_scratch_bytes(T, N)); return B200_ERR_ARG; }
    const int chunks = colsum_chunks(T, N);
    const int rows_per_chunk = (T + chunks - 1) / chunks;
    dim3 grid((N + CS_COLS - 1) / CS_COLS, chunks);
    colsum_partial_kernel<<<grid, EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<float*>(scratch), reinterpret_cast<const __nv_bfloat16*>(x), ld, T, N, rows_per_chunk);
    if (int rc = check_cuda(cudaGetLastError(), "colsum partial launch")) return rc;
    colsum_final_kernel<<<(N + EW_THREADS - 1) / EW_THREADS, EW_THREADS, 0, ST(stream)>>>(out, reinterpret_cast<const float*>(scratch), N, chunks, accumulate);
    return check_cuda(cudaGetLastError(), "colsum final launch");
}

extern "C" int b200_transpose_bf16(void* out, long long ldo, const void* in, long long ldi, int R, int C, void* stream) {
    if (R <= 0 || C <= 0) return 0;
    if (out == nullptr || in == nullptr || (reinterpret_cast<uintptr_t>(out) & 3) || (reinterpret_cast<uintptr_t>(in) & 3) || (ldo % 2) || (ldi % 2) || ldo < R || ldi < C) {
        set_last_error("transpose: 4-byte aligned pointers, even pitches, ldo >= R and ldi >= C required"); return B200_ERR_ARG;
    }
    dim3 grid((C + 63) / 64, (R + 63) / 64);
    transpose_kernel<<<grid, 256, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), ldo, reinterpret_cast<const __nv_bfloat16*>(in), ldi, R, C);
    return check_cuda(cudaGetLastError(), "transpose launch");
}

extern "C" int b200_bias_rows_bf16(void* out, long long ldo, const void* residual, long long ldr, const float* bias, int T, int N, void* stream) {
    if (T <= 0 || N <= 0) return 0;
    if (int rc = check_ptr16(out, "bias_rows out")) return rc;
    if (residual != nullptr) { if (int rc = check_ptr16(residual, "bias_rows residual")) return rc; }
    if (bias != nullptr) { if (int rc = check_ptr16(bias, "bias_rows bias")) return rc; }
    if (N % 8 || ldo % 8 || (residual != nullptr && ldr % 8)) { set_last_error("bias_rows: N and pitches must be multiples of 8"); return B200_ERR_ARG; }
    bias_rows_kernel<<<ew_grid((long long)T * (N / 8)), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), ldo,
        reinterpret_cast<const __nv_bfloat16*>(residual), ldr, bias, T, N);
    return check_cuda(cudaGetLastError(), "bias_rows launch");
}

extern "C" int b200_relu_bf16(void* out, const void* x, long long n, void* stream) {
    if (n <= 0) return 0;
    if (int rc = check_ptr16(out, "relu out")) return rc;
    if (int rc = check_ptr16(x, "relu x")) return rc;
    ew_kernel<0><<<ew_grid(n / 8 + 1), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), reinterpret_cast<const __nv_bfloat16*>(x), nullptr, n);
    return check_cuda(cudaGetLastError(), "relu launch");
}

extern "C" int b200_relu_bwd_bf16(void* dy, const void* x, long long n, void* stream) {
    if (n <= 0) return 0;
    if (int rc = check_ptr16(dy, "relu_bwd dy")) return rc;
    if (int rc = check_ptr16(x, "relu_bwd x")) return rc;
    ew_kernel<1><<<ew_grid(n / 8 + 1), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(x), nullptr, n);
    return check_cuda(cudaGetLastError(), "relu_bwd launch");
}

extern "C" int b200_add_bf16(void* out, const void* a, const void* b, long long n, void* stream) {
    if (n <= 0) return 0;
    if (int rc = check_ptr16(out, "add out")) return rc;
    if (int rc = check_ptr16(a, "add a")) return rc;
    if (int rc = check_ptr16(b, "add b")) return rc;
    ew_kernel<2><<<ew_grid(n / 8 + 1), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b), n);
    return check_cuda(cudaGetLastError(), "add launch");
}

extern "C" int b200_randn_bf16(void* out, long long n, unsigned long long seed, float scale, void* stream) {
    if (n <= 0) return 0;
    if (int rc = check_ptr16(out, "randn out")) return rc;
    randn_kernel<<<ew_grid((n + 7) / 8), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), n, seed, scale);
    return check_cuda(cudaGetLastError(), "randn launch");
}

extern "C" int b200_cast_f32_to_bf16(void* out, const float* in, long long n, void* stream) {
    if (n <= 0) return 0;
    if (out == nullptr || in == nullptr) { set_last_error("cast: null pointer"); return B200_ERR_ARG; }
    cast_f32_bf16_kernel<<<ew_grid(n), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), in, n);
    return check_cuda(cudaGetLastError(), "cast launch");
}

extern "C" int b200_cast_bf16_to_f32(float* out, const void* in, long long n, void* stream) {
    if (n <= 0) return 0;
    if (out == nullptr || in == nullptr) { set_last_error("cast: null pointer"); return B200_ERR_ARG; }
    cast_bf16_f32_kernel<<<ew_grid(n), EW_THREADS, 0, ST(stream)>>>(out, reinterpret_cast<const __nv_bfloat16*>(in), n);
    return check_cuda(cudaGetLastError(), "cast launch");
}

extern "C" long long b200_sumsq_scratch_bytes(void) { return (long long)sizeof(SumsqScratch); }

extern "C" int b200_sumsq_bf16(float* out, const void* x, long long n, void* scratch, long long scratch_bytes, void* stream) {
    if (out == nullptr || scratch == nullptr) { set_last_error("sumsq: null out/scratch"); return B200_ERR_ARG; }
    if (n < 0 || (n > 0 && x == nullptr)) { set_last_error("sumsq: bad input"); return B200_ERR_ARG; }
    if (n > 0) { if (int rc = check_ptr16(x, "sumsq x")) return rc; }
    if (scratch_bytes < (long long)sizeof(SumsqScratch) || (reinterpret_cast<uintptr_t>(scratch) & 15)) {
        set_last_error("sumsq: scratch must be 16-byte aligned and hold %lld bytes", (long long)sizeof(SumsqScratch)); return B200_ERR_ARG;
    }
    int grid = ew_grid(n / 8 + 1);
    if (grid > SUMSQ_MAX_BLOCKS) grid = SUMSQ_MAX_BLOCKS;
    sumsq_kernel<<<grid, EW_THREADS, 0, ST(stream)>>>(out, reinterpret_cast<const __nv_bfloat16*>(x), n, reinterpret_cast<SumsqScratch*>(scratch));
    return check_cuda(cudaGetLastError(), "sumsq launch");
}

extern "C" long long b200_sumsq_static_max_elems(void) { return SUMSQ_STATIC_MAX_ELEMS; }

// Same result as b200_sumsq_bf16, but the caller brings no scratch and nothing is allocated here or later: the
// partial sums go through the library's own fixed device array.  See include/b200ffn.h for who needs this.
extern "C" int b200_sumsq_bf16_static(float* out, const void* x, long long n, void* stream) {
    if (out == nullptr) { set_last_error("sumsq_static: null out"); return B200_ERR_ARG; }
    if (n < 0 || (n > 0 && x == nullptr)) { set_last_error("sumsq_static: bad input"); return B200_ERR_ARG; }
    if (n > SUMSQ_STATIC_MAX_ELEMS) {
        set_last_error("sumsq_static: n = %lld exceeds the %lld-element limit of the built-in scratch; "
                       "use b200_sumsq_bf16 with caller-provided scratch for larger inputs", n, SUMSQ_STATIC_MAX_ELEMS);
        return B200_ERR_ARG;
    }
    if (n > 0) { if (int rc = check_ptr16(x, "sumsq_static x")) return rc; }
    void* scratch = nullptr;
    if (int rc = sumsq_static_address(&scratch)) return rc;          // first call on this device does the look-up
    cudaStream_t st = ST(stream);
    // the arrival counter has to read zero when the kernel starts, whatever an earlier (possibly aborted) call left
    if (int rc = check_cuda(cudaMemsetAsync(scratch, 0, 16, st), "sumsq_static counter reset")) return rc;
    int grid = ew_grid(n / 8 + 1);
    if (grid > SUMSQ_STATIC_MAX_BLOCKS) grid = SUMSQ_STATIC_MAX_BLOCKS;
    sumsq_kernel<<<grid, EW_THREADS, 0, st>>>(out, reinterpret_cast<const __nv_bfloat16*>(x), n, reinterpret_cast<SumsqScratch*>(scratch));
    return check_cuda(cudaGetLastError(), "sumsq_static launch");
}

extern "C" int b200_fill_zero(void* buf, long long bytes, void* stream) {
    if (bytes <= 0) return 0;
    if (int rc = check_ptr16(buf, "fill_zero buf")) return rc;
    if (bytes % 16) { set_last_error("fill_zero: bytes must be a multiple of 16"); return B200_ERR_ARG; }
    fill_zero_kernel<<<ew_grid(bytes / 16), EW_THREADS, 0, ST(stream)>>>(reinterpret_cast<uint4*>(buf), bytes / 16);
    return check_cuda(cudaGetLastError(), "fill_zero launch");
}
