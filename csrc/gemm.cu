// C-ABI entry points for the single-GPU tcgen05 GEMM family (see include/b200ffn.h).
#include "gemm_kernel.cuh"
#include "../include/b200ffn.h"
#include <cstdlib>

namespace b200 {

int validate_operands(const GemmOperands& o) {
    if (o.M <= 0 || o.N <= 0 || o.K <= 0) { set_last_error("gemm: M,N,K must be positive (got %d,%d,%d)", o.M, o.N, o.K); return B200_ERR_ARG; }
    if (o.A == nullptr || o.B == nullptr) { set_last_error("gemm: null operand"); return B200_ERR_ARG; }
    if ((o.a_major != MAJOR_K && o.a_major != MAJOR_MN) || (o.b_major != MAJOR_K && o.b_major != MAJOR_MN)) {
        set_last_error("gemm: major must be 0 (K) or 1 (MN)"); return B200_ERR_ARG;
    }
    // TMA: base 16-byte aligned, row pitch a multiple of 16 bytes; vector epilogue: N % 8 == 0.
    if ((reinterpret_cast<uintptr_t>(o.A) & 15) || (reinterpret_cast<uintptr_t>(o.B) & 15)) {
        set_last_error("gemm: operand base pointers must be 16-byte aligned"); return B200_ERR_ARG;
    }
    if ((o.lda % 8) || (o.ldb % 8) || (o.N % 8)) {
        set_last_error("gemm: lda, ldb and N must be multiples of 8 elements (lda=%lld ldb=%lld N=%d)", o.lda, o.ldb, o.N);
        return B200_ERR_ARG;
    }
    const long long a_inner = o.a_major == MAJOR_K ? o.K : o.M;
    const long long b_inner = o.b_major == MAJOR_K ? o.K : o.N;
    if (o.lda < a_inner || o.ldb < b_inner) { set_last_error("gemm: leading dimension smaller than the row length"); return B200_ERR_ARG; }
    return 0;
}

bool use_pair_kernel() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("B200_GEMM_PAIR");
        cached = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return cached == 1;
}

int make_core_and_maps(const GemmOperands& o, GemmCore* core, CUtensorMap* tmA, CUtensorMap* tmB, bool pair) {
    int rc = validate_operands(o);
    if (rc != 0) return rc;
    core->M = o.M; core->N = o.N; core->K = o.K;
    core->a_major = o.a_major; core->b_major = o.b_major;
    core->tiles_m = (o.M + BM - 1) / BM;
    core->tiles_n = (o.N + BN - 1) / BN;
    core->group_m = 16;
    core->tile_offset = 0;
    core->err = device_error_slot();
    core->a_plane = (int)o.a_plane; core->b_plane = (int)o.b_plane;
    const long long a_rows = (o.a_major == MAJOR_K ? o.M : o.K) + (long long)(o.planes - 1) * o.a_plane;   // storage rows incl. the stacked planes
    const long long b_rows = (o.b_major == MAJOR_K ? o.N : o.K) + (long long)(o.planes - 1) * o.b_plane;
    if (o.planes > 1 && ((o.a_major == MAJOR_MN || o.b_major == MAJOR_MN) && (o.K % BK))) {
        set_last_error("gemm (split operands): K must be a multiple of 64 when an operand is MN-major (K=%d)", o.K); return B200_ERR_ARG;
    }
    // K-major boxes: the pair kernel stages 256 A rows and 128 B rows per CTA, the 1-CTA kernel 128 and 256
    if (o.a_major == MAJOR_K) rc = encode_tmap_bf16_2d(tmA, o.A, o.K, a_rows, o.lda, BK, pair ? 256 : BM);
    else                      rc = encode_tmap_bf16_2d(tmA, o.A, o.M, a_rows, o.lda, 64, BK);
    if (rc != 0) return rc;
    // the pair kernel stages 128 B rows per CTA, the 1-CTA kernel all 256
    if (o.b_major == MAJOR_K) rc = encode_tmap_bf16_2d(tmB, o.B, o.K, b_rows, o.ldb, BK, pair ? 128 : BN);
    else                      rc = encode_tmap_bf16_2d(tmB, o.B, o.N, b_rows, o.ldb, 64, BK);
    return rc;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemm_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                              void* C, long long ldc, int M, int N, int K, int epilogue, const void* aux,
                              long long ldaux, void* C2, long long ldc2, const float* bias, float alpha,
                              int max_ctas, void* stream) {
    GemmOperands o{A, lda, a_major, B, ldb, b_major, M, N, K};
    if (C == nullptr) { set_last_error("gemm: null output"); return B200_ERR_ARG; }
    if ((reinterpret_cast<uintptr_t>(C) & 15) || (ldc % 8)) { set_last_error("gemm: C must be 16-byte aligned with ldc %% 8 == 0"); return B200_ERR_ARG; }
    if (ldc < N) { set_last_error("gemm: ldc < N"); return B200_ERR_ARG; }
    const bool needs_aux = epilogue == EPI_RELU_BWD || epilogue == EPI_GELU_BWD || epilogue == EPI_ADD;
    if (needs_aux && (aux == nullptr || (reinterpret_cast<uintptr_t>(aux) & 15) || (ldaux % 8) || ldaux < N)) {
        set_last_error("gemm: epilogue %d needs a 16-byte aligned aux[M,N] with ldaux %% 8 == 0", epilogue); return B200_ERR_ARG;
    }
    if (epilogue == EPI_GELU && (C2 == nullptr || (reinterpret_cast<uintptr_t>(C2) & 15) || (ldc2 % 8) || ldc2 < N)) {
        set_last_error("gemm: GELU epilogue needs a second output C2[M,N] for the pre-activation"); return B200_ERR_ARG;
    }
    if (epilogue == EPI_RELU_MASK && (C2 == nullptr || (reinterpret_cast<uintptr_t>(C2) & 15) || (ldc2 % 16) || ldc2 * 8 < ((N + 127) / 128) * 128)) {
        set_last_error("gemm: RELU_MASK needs C2 = uint8 mask[M, ldc2] with ldc2 %% 16 == 0 and ldc2 >= ceil(N/128)*16 bytes"); return B200_ERR_ARG;
    }
    if (epilogue == EPI_RELU_BWD_MASK && (aux == nullptr || (reinterpret_cast<uintptr_t>(aux) & 15) || (ldaux % 16) || ldaux * 8 < ((N + 127) / 128) * 128)) {
        set_last_error("gemm: RELU_BWD_MASK needs aux = uint8 mask[M, ldaux] with ldaux %% 16 == 0 and ldaux >= ceil(N/128)*16 bytes"); return B200_ERR_ARG;
    }
    LocalEpiParams ep{C, ldc, reinterpret_cast<const __nv_bfloat16*>(aux), ldaux, reinterpret_cast<__nv_bfloat16*>(C2), ldc2, bias, alpha, epilogue == EPI_SGD ? update_rounding_seed() : 0u};
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (epilogue) {
        case EPI_STORE:     return launch_gemm<LocalEpi<EPI_STORE>>(o, ep, max_ctas, st);
        case EPI_RELU:      return launch_gemm<LocalEpi<EPI_RELU>>(o, ep, max_ctas, st);
        case EPI_GELU:      return launch_gemm<LocalEpi<EPI_GELU>>(o, ep, max_ctas, st);
        case EPI_RELU_BWD:  return launch_gemm<LocalEpi<EPI_RELU_BWD>>(o, ep, max_ctas, st);
        case EPI_GELU_BWD:  return launch_gemm<LocalEpi<EPI_GELU_BWD>>(o, ep, max_ctas, st);
        case EPI_ADD:       return launch_gemm<LocalEpi<EPI_ADD>>(o, ep, max_ctas, st);
        case EPI_SGD:       return ep.sr_seed != 0u ? launch_gemm<LocalEpi<EPI_SGD_SR>>(o, ep, max_ctas, st)
                                                    : launch_gemm<LocalEpi<EPI_SGD>>(o, ep, max_ctas, st);
        case EPI_STORE_F32: return launch_gemm<LocalEpi<EPI_STORE_F32>>(o, ep, max_ctas, st);
        case EPI_RELU_MASK: return launch_gemm<LocalEpi<EPI_RELU_MASK>>(o, ep, max_ctas, st);
        case EPI_RELU_BWD_MASK: return launch_gemm<LocalEpi<EPI_RELU_BWD_MASK>>(o, ep, max_ctas, st);
        default: set_last_error("gemm: unknown epilogue %d", epilogue); return B200_ERR_ARG;
    }
}
