/* b200ffn - C ABI of the native sm_100a library (libb200ffn.so).
 *
 * Every function returns 0 on success, a negative B200_ERR_* for argument/setup errors, or a positive
 * cudaError_t.  b200_last_error() returns a human-readable message for the last failure on this thread.
 * All kernels are enqueued on the caller's stream (a cudaStream_t passed as void*; NULL = legacy default
 * stream) and never synchronise the device.  Pointers are device pointers unless stated otherwise.
 *
 * Reference citations are into martin-kukla/distributed-llm-code-samples `train_ffns.py`.
 */
#ifndef B200FFN_H
#define B200FFN_H

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ERR_ARG (-1)      /* bad argument (shape, alignment, null pointer) */
#define B200_ERR_DRIVER (-2)   /* driver entry point missing / tensor-map encode failed */
#define B200_ERR_STATE (-3)    /* runtime not initialised / wrong order of calls */
#define B200_ERR_DEVICE (-4)   /* a kernel recorded a device-side failure (see b200_device_error) */

/* ---- runtime ------------------------------------------------------------------------------------------- */
const char* b200_last_error(void);
int b200_version(void);
/* Copies the 4-word device failure record {code, block, thread, aux} (pinned host memory, readable even
 * after a kernel trapped) into out[4]; returns the code (0 = none). */
int b200_device_error(unsigned int* out);
int b200_clear_device_error(void);
int b200_num_sms(void);

/* ---- GEMM: D[M,N] = sum_k A(m,k) * B(n,k), bf16 in, fp32 accumulate in TMEM -------------------------------
 * a_major / b_major: 0 = K-major (row-major [MN,K]; ld = row pitch in elements)
 *                    1 = MN-major (row-major [K,MN])
 * The reference's call sites map as (x:[T,D], W1:[F,D], W2:[D,F]):
 *   h  = x  @ W1^T  (:42 via :55)  A=x  K-major,  B=W1 K-major,  epilogue RELU/GELU
 *   y  = a  @ W2^T  (:42 via :57)  A=a  K-major,  B=W2 K-major,  epilogue STORE/ADD(residual)
 *   da = dy @ W2    (:45 'bc,cd->bd')  A=dy K-major, B=W2 MN-major, epilogue RELU_BWD/GELU_BWD
 *   dx = dh @ W1    (:45)              A=dh K-major, B=W1 MN-major, epilogue STORE/ADD
 *   dW = dy^T @ x   (:45 'bc,bd->cd')  A=dy MN-major, B=x MN-major, epilogue STORE/STORE_F32/SGD
 * epilogue (B200_EPI_*):
 *   0 STORE      C = acc (+bias)                      4 GELU_BWD  C = acc * gelu'(aux)
 *   1 RELU       C = relu(acc + bias)                 5 ADD       C = acc + aux (+bias)
 *   2 GELU       C = gelu(acc+bias), C2 = acc+bias    6 SGD       C = C - alpha*acc   (in place on the weight)
 *   3 RELU_BWD   C = acc * (aux > 0)                  7 STORE_F32 C(fp32) = alpha*acc
 * Preconditions: 16-byte aligned bases; lda, ldb, ldc, ldaux, ldc2 and N multiples of 8 elements.
 * max_ctas: 0 = one CTA per SM; >0 caps the persistent grid (leaves SMs for a concurrent comm kernel). */
#define B200_EPI_STORE 0
#define B200_EPI_RELU 1
#define B200_EPI_GELU 2
#define B200_EPI_RELU_BWD 3
#define B200_EPI_GELU_BWD 4
#define B200_EPI_ADD 5
#define B200_EPI_SGD 6
#define B200_EPI_STORE_F32 7
int b200_gemm_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                   void* C, long long ldc, int M, int N, int K, int epilogue, const void* aux, long long ldaux,
                   void* C2, long long ldc2, const float* bias, float alpha, int max_ctas, void* stream);

/* ---- elementwise / optimizer / data --------------------------------------------------------------------- */
/* p -= lr * g  (:114, :172, :259, :312).  grad_f32 != 0: g is fp32, else bf16.  n % 8 == 0 not required. */
int b200_sgd_bf16(void* param, const void* grad, long long n, float lr, int grad_f32, void* stream);
/* Adam on bf16 params with fp32 moments (extension; the reference only has SGD). step is 1-based. */
int b200_adam_bf16(void* param, const void* grad, float* m, float* v, long long n, float lr, float beta1,
                   float beta2, float eps, int step, int grad_f32, void* stream);
/* out = relu(x) (:47-48) and dy = dy * (x > 0) in place (:50-52), standalone (unfused) forms. */
int b200_relu_bf16(void* out, const void* x, long long n, void* stream);
int b200_relu_bwd_bf16(void* dy, const void* x, long long n, void* stream);
int b200_add_bf16(void* out, const void* a, const void* b, long long n, void* stream);
/* out[i] = scale * N(0,1), Philox4x32-10 counter-based (device-side stand-in for mock_data :144-151). */
int b200_randn_bf16(void* out, long long n, unsigned long long seed, float scale, void* stream);
int b200_cast_f32_to_bf16(void* out, const float* in, long long n, void* stream);
int b200_cast_bf16_to_f32(float* out, const void* in, long long n, void* stream);
/* Deterministic sum of squares of a bf16 vector -> out[0] (fp32).  Two-stage: per-CTA partials in `scratch`
 * then a fixed-order final pass by the last CTA to finish.  scratch must hold b200_sumsq_scratch_bytes()
 * bytes and its first 4 bytes (the arrival counter) must be zero on entry; the kernel re-zeroes it. */
long long b200_sumsq_scratch_bytes(void);
int b200_sumsq_bf16(float* out, const void* x, long long n, void* scratch, long long scratch_bytes, void* stream);
/* Writes `bytes` of zeros (L2 flush between timed iterations). */
int b200_fill_zero(void* buf, long long bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200FFN_H */
