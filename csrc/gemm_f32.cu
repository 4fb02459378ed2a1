// fp32 GEMMs on the bf16 tensor cores (reference-parity mode: the reference trains in fp32, train_ffns.py:36,:149).
//
// tcgen05 has no fp32-input MMA (kind::tf32 keeps 10 mantissa bits - a precision downgrade against the reference's true
// fp32 cuBLAS GEMMs).  Instead every fp32 operand is split into three bf16 planes, x = x0 + x1 + x2 with
// x0 = bf16(x), x1 = bf16(x - x0), x2 = bf16(x - x0 - x1)   (3 x 8 = 24 mantissa bits: exact up to the last fp32 ulp),
// and the product is accumulated from the six plane products of weight >= 2^-16, smallest first, in the fp32 TMEM
// accumulator:  A.B ~= A1.B1 + A0.B2 + A2.B0 + A0.B1 + A1.B0 + A0.B0.
// bf16 x bf16 products are exact in fp32, so the only roundings are the accumulator's (measured, tools/f32_probe.py: within
// 1.6e-7 of sum |a||b| at K = 256; a three-block forward within 2.6e-6 of its scale against cuBLAS fp32's ~7e-7; TF32 would
// give ~5e-4, bf16 ~4e-3) - at 6x the bf16 MMA work, ~5x the speed of cuBLAS's fp32 GEMM on this part.
// The mainloop is the ordinary one (gemm_kernel.cuh) walked
// six times over plane-shifted TMA coordinates (Epi::PRODUCTS = 6); the epilogues below are the fp32 forms of LocalEpi.
#include "gemm_kernel.cuh"
#include "../include/b200ffn.h"

namespace b200 {

struct F32EpiParams {
    float* C;
    long long ldc;
    const float* aux;
    long long ldaux;
    float* C2;
    long long ldc2;
    const float* bias;
    float alpha;
};

// KIND: the B200_EPI_* numbers (STORE, RELU, GELU, RELU_BWD, GELU_BWD, ADD, SGD); every tensor fp32.
template <int KIND>
struct F32Epi {
    using Params = F32EpiParams;
    static constexpr int SMEM_BYTES = GEMM_SMEM_BYTES_STAGED;
    static constexpr int EXTRA_WARPS = 0;
    static constexpr int LAUNCH_BOUND_PAD = 0;
    static constexpr int PRODUCTS = 6;
    static __device__ __forceinline__ void side_loop(const Params&, const GemmCore&, int, int, uint32_t*) {}
    static __device__ __forceinline__ void before_loads(const Params&, const GemmCore&, int, int) {}
    struct ProducerState {};
    static __device__ __forceinline__ int kblock(const Params&, const GemmCore&, int kb, int, ProducerState&) { return kb; }
    struct State {};
    static __device__ __forceinline__ void init(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void after_skipped_tile(const Params&, const GemmCore&, const EpiCtx&, State&) {}
    static __device__ __forceinline__ void tail(const Params&, const GemmCore&, const EpiCtx&, State&) {}

    static constexpr bool READS = (KIND == EPI_RELU_BWD || KIND == EPI_GELU_BWD || KIND == EPI_ADD || KIND == EPI_SGD);
    static constexpr bool BIAS = (KIND == EPI_STORE || KIND == EPI_RELU || KIND == EPI_GELU || KIND == EPI_ADD);

    // Each lane owns one row; 32 fp32 columns are one 128-byte line, so row-owner accesses are already whole lines.
    static __device__ __forceinline__ void process_tile(const Params& p, const GemmCore& g, const EpiCtx& ctx, uint32_t taddr, State&) {
        const int row = ctx.m0 + ctx.q * 32 + ctx.lane;
        const int colbase = ctx.n0 + ctx.half * (BN / 2);
        const float* rd = KIND == EPI_SGD ? p.C : p.aux;
        const long long rd_ld = KIND == EPI_SGD ? p.ldc : p.ldaux;
#pragma unroll 1
        for (int c = 0; c < BN / 2 / 32; ++c) {
            uint32_t acc[32];
            tmem_ld_32x32(taddr + c * 32, acc);
            tmem_ld_wait();
            if (row >= g.M) continue;
            const int col0 = colbase + c * 32;
#pragma unroll
            for (int v = 0; v < 8; ++v) {
                const int col = col0 + v * 4;
                if (col >= g.N) break;
                float f[4] = {__uint_as_float(acc[v * 4]), __uint_as_float(acc[v * 4 + 1]), __uint_as_float(acc[v * 4 + 2]), __uint_as_float(acc[v * 4 + 3])};
                if constexpr (BIAS) {
                    if (p.bias != nullptr) {
                        const float4 b = *reinterpret_cast<const float4*>(p.bias + col);
                        f[0] += b.x; f[1] += b.y; f[2] += b.z; f[3] += b.w;
                    }
                }
                float r[4] = {0.f, 0.f, 0.f, 0.f};
                if constexpr (READS) {
                    const float4 a = *reinterpret_cast<const float4*>(rd + (long long)row * rd_ld + col);
                    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
                }
                if constexpr (KIND == EPI_GELU)
                    *reinterpret_cast<float4*>(p.C2 + (long long)row * p.ldc2 + col) = make_float4(f[0], f[1], f[2], f[3]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if constexpr (KIND == EPI_STORE) f[j] *= p.alpha;
                    else if constexpr (KIND == EPI_RELU) f[j] = f[j] <= 0.0f ? 0.0f : f[j];          // NaN stays NaN (:47-48)
                    else if constexpr (KIND == EPI_GELU) f[j] = gelu_f(f[j]);
                    else if constexpr (KIND == EPI_RELU_BWD) f[j] = r[j] <= 0.0f ? 0.0f : f[j];      // masked_fill_(x <= 0, 0) (:50-52)
                    else if constexpr (KIND == EPI_GELU_BWD) f[j] *= gelu_grad_f(r[j]);
                    else if constexpr (KIND == EPI_ADD) f[j] += r[j];
                    else if constexpr (KIND == EPI_SGD) f[j] = r[j] - p.alpha * f[j];                // W -= lr * dW (:114, :172)
                }
                *reinterpret_cast<float4*>(p.C + (long long)row * p.ldc + col) = make_float4(f[0], f[1], f[2], f[3]);
            }
        }
    }
};

// fp32 [rows, cols] (pitch ld_in) -> three bf16 planes [3][rows][cols] (contiguous), 8 elements per thread.
__global__ void __launch_bounds__(256) split3_kernel(__nv_bfloat16* out, const float* in, long long rows, int cols, long long ld_in) {
    const int c8 = cols / 8;
    const long long total = rows * c8;
    const long long plane = rows * (long long)cols;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long r = i / c8;
        const int c = (int)(i % c8) * 8;
        const float4 a = *reinterpret_cast<const float4*>(in + r * ld_in + c), b = *reinterpret_cast<const float4*>(in + r * ld_in + c + 4);
        const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        float p0[8], p1[8], p2[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float h0 = __bfloat162float(__float2bfloat16_rn(x[j]));
            const float r1 = x[j] - h0;                                   // exact (Sterbenz-like: h0 carries x's leading bits)
            const float h1 = __bfloat162float(__float2bfloat16_rn(r1));
            p0[j] = h0; p1[j] = h1; p2[j] = r1 - h1;
        }
        uint4 o0, o1, o2;
        o0.x = pack_bf16x2(p0[0], p0[1]); o0.y = pack_bf16x2(p0[2], p0[3]); o0.z = pack_bf16x2(p0[4], p0[5]); o0.w = pack_bf16x2(p0[6], p0[7]);
        o1.x = pack_bf16x2(p1[0], p1[1]); o1.y = pack_bf16x2(p1[2], p1[3]); o1.z = pack_bf16x2(p1[4], p1[5]); o1.w = pack_bf16x2(p1[6], p1[7]);
        o2.x = pack_bf16x2(p2[0], p2[1]); o2.y = pack_bf16x2(p2[2], p2[3]); o2.z = pack_bf16x2(p2[4], p2[5]); o2.w = pack_bf16x2(p2[6], p2[7]);
        const long long o = r * cols + c;
        *reinterpret_cast<uint4*>(out + o) = o0;
        *reinterpret_cast<uint4*>(out + plane + o) = o1;
        *reinterpret_cast<uint4*>(out + 2 * plane + o) = o2;
    }
}

}  // namespace b200

using namespace b200;

// in: fp32 [rows, cols] with row pitch ld_in (elements); out: bf16 [3][rows][cols].  cols % 8 == 0, 16-byte aligned.
extern "C" int b200_split_f32(void* out, const float* in, long long rows, int cols, long long ld_in, void* stream) {
    if (rows <= 0 || cols <= 0) return 0;
    if (out == nullptr || in == nullptr || (reinterpret_cast<uintptr_t>(out) & 15) || (reinterpret_cast<uintptr_t>(in) & 15) || (cols % 8) || (ld_in % 4) || ld_in < cols) {
        set_last_error("split_f32: 16-byte aligned pointers, cols %% 8 == 0 and ld_in %% 4 == 0 required"); return B200_ERR_ARG;
    }
    long long blocks = (rows * (cols / 8) + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    split3_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(reinterpret_cast<__nv_bfloat16*>(out), in, rows, cols, ld_in);
    return check_cuda(cudaGetLastError(), "split_f32 launch");
}

// C[M,N] (fp32) = epilogue(sum_k A(m,k) B(n,k)) with A and B given as the three bf16 planes b200_split_f32 makes of an fp32
// matrix ([3][rows][cols] contiguous; majors and the meaning of rows / cols as in b200_gemm_bf16; lda / ldb = cols).
extern "C" int b200_gemm_f32(const void* A3, long long lda, int a_major, const void* B3, long long ldb, int b_major,
                             float* C, long long ldc, int M, int N, int K, int epilogue, const float* aux, long long ldaux,
                             float* C2, long long ldc2, const float* bias, float alpha, int max_ctas, void* stream) {
    GemmOperands o{A3, lda, a_major, B3, ldb, b_major, M, N, K};
    o.planes = 3;
    o.a_plane = a_major == MAJOR_K ? M : K;
    o.b_plane = b_major == MAJOR_K ? N : K;
    if (C == nullptr || (reinterpret_cast<uintptr_t>(C) & 15) || (ldc % 4) || ldc < N) { set_last_error("gemm_f32: C must be 16-byte aligned with ldc %% 4 == 0 and ldc >= N"); return B200_ERR_ARG; }
    const bool needs_aux = epilogue == EPI_RELU_BWD || epilogue == EPI_GELU_BWD || epilogue == EPI_ADD;
    if (needs_aux && (aux == nullptr || (reinterpret_cast<uintptr_t>(aux) & 15) || (ldaux % 4) || ldaux < N)) { set_last_error("gemm_f32: epilogue %d needs a 16-byte aligned fp32 aux[M,N]", epilogue); return B200_ERR_ARG; }
    if (epilogue == EPI_GELU && (C2 == nullptr || (reinterpret_cast<uintptr_t>(C2) & 15) || (ldc2 % 4) || ldc2 < N)) { set_last_error("gemm_f32: GELU needs a second fp32 output for the pre-activation"); return B200_ERR_ARG; }
    if (bias != nullptr && (reinterpret_cast<uintptr_t>(bias) & 15)) { set_last_error("gemm_f32: bias must be 16-byte aligned"); return B200_ERR_ARG; }
    F32EpiParams ep{C, ldc, aux, ldaux, C2, ldc2, bias, alpha};
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (epilogue) {
        case EPI_STORE:    return launch_gemm<F32Epi<EPI_STORE>>(o, ep, max_ctas, st);
        case EPI_RELU:     return launch_gemm<F32Epi<EPI_RELU>>(o, ep, max_ctas, st);
        case EPI_GELU:     return launch_gemm<F32Epi<EPI_GELU>>(o, ep, max_ctas, st);
        case EPI_RELU_BWD: return launch_gemm<F32Epi<EPI_RELU_BWD>>(o, ep, max_ctas, st);
        case EPI_GELU_BWD: return launch_gemm<F32Epi<EPI_GELU_BWD>>(o, ep, max_ctas, st);
        case EPI_ADD:      return launch_gemm<F32Epi<EPI_ADD>>(o, ep, max_ctas, st);
        case EPI_SGD:      return launch_gemm<F32Epi<EPI_SGD>>(o, ep, max_ctas, st);
        default: set_last_error("gemm_f32: unsupported epilogue %d", epilogue); return B200_ERR_ARG;
    }
}
