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

/* Rounding of the new bf16 weight in EVERY optimizer update launched after this call (GEMM SGD epilogue, fused
 * all-reduce / reduce-scatter + SGD / Adam, b200_sgd_bf16, b200_adam_bf16): 0 = round to nearest (default); any other
 * value = stochastic rounding seeded with it (random bits = hash(seed, element index), so replicas agree).  Without an
 * fp32 master copy an update below half a bf16 ulp of the weight rounds away - at the reference's LR = 1e-5
 * (train_ffns.py:29) that is every update; stochastic rounding keeps its expectation.  Use a new seed each step. */
int b200_set_update_rounding(unsigned int seed);

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
 *   8 RELU_MASK      as RELU, and C2 receives one sign bit per element (uint8 buffer, tile-blocked: the 16 bytes of
 *                    (row, 128-column chunk cc) at ((row/128) * ceil(N/128) + cc) * 2048 + (row%128) * 16; size
 *                    ceil(M/128)*128 rows x ceil(N/128)*16 bytes; ldc2 = bytes per row of that shape, % 16 == 0)
 *   9 RELU_BWD_MASK  C = acc where the bit in aux (that mask, same layout, ldaux in BYTES) is set, else 0 - the backward
 *                    of ReLU (:50-52) from 1 bit/element instead of re-reading the bf16 activation
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
#define B200_EPI_RELU_MASK 8
#define B200_EPI_RELU_BWD_MASK 9
int b200_gemm_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                   void* C, long long ldc, int M, int N, int K, int epilogue, const void* aux, long long ldaux,
                   void* C2, long long ldc2, const float* bias, float alpha, int max_ctas, void* stream);

/* ---- fp32 GEMM on the bf16 tensor cores (reference-parity mode; csrc/gemm_f32.cu) ------------------------------------
 * The reference trains in fp32 (train_ffns.py:36, :149).  b200_split_f32 turns an fp32 matrix [rows, cols] (row pitch
 * ld_in) into three bf16 planes [3][rows][cols] whose sum is the matrix to the last fp32 ulp; b200_gemm_f32 multiplies
 * two such operands with six plane products accumulated in the fp32 TMEM accumulator (smallest first): fp32-class
 * accuracy (measured 1.6e-7 of sum |a||b| at K = 256; TF32 would give ~5e-4) at 6x the bf16 MMA work.  Majors / M,N,K as b200_gemm_bf16 (lda / ldb = cols of the split matrix); every other
 * tensor (C, aux, C2, bias) is fp32.  Epilogues: STORE (C = alpha*acc + bias), RELU, GELU (+C2 = pre-activation),
 * RELU_BWD, GELU_BWD, ADD, SGD (C -= alpha*acc).  cols % 8 == 0; MN-major operands need K % 64 == 0. */
int b200_split_f32(void* planes_out, const float* in, long long rows, int cols, long long ld_in, void* stream);
int b200_gemm_f32(const void* A3, long long lda, int a_major, const void* B3, long long ldb, int b_major,
                  float* C, long long ldc, int M, int N, int K, int epilogue, const float* aux, long long ldaux,
                  float* C2, long long ldc2, const float* bias, float alpha, int max_ctas, void* stream);

/* ---- elementwise / optimizer / data --------------------------------------------------------------------- */
/* p -= lr * g  (:114, :172, :259, :312).  grad_f32 != 0: g is fp32, else bf16.  n % 8 == 0 not required. */
int b200_sgd_bf16(void* param, const void* grad, long long n, float lr, int grad_f32, void* stream);
/* Adam on bf16 params with fp32 moments (extension; the reference only has SGD). step is 1-based. */
int b200_adam_bf16(void* param, const void* grad, float* m, float* v, long long n, float lr, float beta1,
                   float beta2, float eps, int step, int grad_f32, void* stream);
/* fp32 parameters (biases stay fp32 whatever the weights' dtype): p -= lr * g, and Adam. */
int b200_sgd_f32(float* param, const float* grad, long long n, float lr, void* stream);
int b200_adam_f32(float* param, const float* grad, float* m, float* v, long long n, float lr, float beta1, float beta2,
                  float eps, int step, void* stream);
/* Bias gradient: out[n] (+)= sum_t x[t, n] over a bf16 [T, N] matrix of row pitch ld (elements).  Deterministic two-stage
 * sum; `scratch`: b200_colsum_scratch_bytes(T, N) bytes of device memory.  accumulate != 0 adds to out.  The reference
 * has no bias (train_ffns.py:35 "no bias as simplification"); this is the north-star extension's db = dy.sum(0). */
long long b200_colsum_scratch_bytes(int T, int N);
int b200_colsum_bf16(float* out, const void* x, long long ld, int T, int N, int accumulate, void* scratch,
                     long long scratch_bytes, void* stream);
/* out[t, :] = bias (+ residual[t, :]): pre-fills the output of a fused row-parallel GEMM + all-reduce / reduce-scatter whose
 * in-place mode (op 1, alpha = -1) then adds the cross-rank sum - how bias and residual reach the fused MP kernels.
 * bias (fp32 [N]) and residual (bf16) may each be NULL. */
int b200_bias_rows_bf16(void* out, long long ldo, const void* residual, long long ldr, const float* bias, int T, int N, void* stream);
/* out[c, r] = in[r, c] for a bf16 [R, C] matrix (pitches in elements, even).  Feature-major activations of the FSDP trainer
 * whose parameter all-gather runs inside the GEMMs (the weight is then the gathered A operand).  No reference counterpart. */
int b200_transpose_bf16(void* out, long long ldo, const void* in, long long ldi, int R, int C, void* stream);
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
/* Allocation-free form of b200_sumsq_bf16 (no counterpart in the reference, which never calls a kernel of its own).
 * For hosts that cannot hand this library any device memory beyond the input and the 4-byte result, and cannot let
 * it allocate either: a plain C program driving the kernels without a tensor framework behind it; a plug-in inside a
 * process whose device memory already belongs in full to another allocator (PyTorch's caching allocator, a serving
 * arena), where a raw cudaMalloc fails or costs a device-wide synchronisation; a step loop that must not take the
 * driver's allocation lock.  The per-CTA partial sums go through a 1280-byte array built into the library image,
 * which every GPU that loads the library holds until the process exits.
 *   - limit: n <= b200_sumsq_static_max_elems() (2^31 elements); larger inputs are refused with B200_ERR_ARG and a
 *     message - use b200_sumsq_bf16 with your own scratch for those.
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, that device must be
 *     current); make that first call before starting any stream capture.
 *   - no cudaMalloc / cudaFree / allocating library call is made, on the first call or later.
 *   - every call on a device uses the same array: the arrival counter is reset on the stream before each launch,
 *     but two calls whose kernels overlap on the GPU (different streams) corrupt each other - serialise them.
 *   - the looked-up address is cached per device ordinal for the life of the process; cudaDeviceReset() on that device
 *     invalidates it (the module is reloaded at a new address) - do not reset a device this entry point has used. */
long long b200_sumsq_static_max_elems(void);
int b200_sumsq_bf16_static(float* out, const void* x, long long n, void* stream);
/* Writes `bytes` of zeros (L2 flush between timed iterations). */
int b200_fill_zero(void* buf, long long bytes, void* stream);

/* ---- peer-memory runtime (one process per GPU; NVLink/NVSwitch) ------------------------------------------------
 * Symmetric memory: every rank allocates the same size with b200_symm_alloc (zero-filled cudaMalloc), exports a
 * CUDA-IPC handle, and opens every peer's handle; the resulting table of `world` base pointers (own pointer at
 * index `rank`) is what the collectives and fused kernels take as `*_peers` (HOST arrays of device pointers).
 * This is the explicit form of the hand-off the reference gets implicitly by pickling CUDA tensors to child
 * processes (train_ffns.py:186; test_mp_barrier_gpus.py:33).  In a single process driving several GPUs use
 * b200_enable_peer_access and the raw pointers instead. */
int b200_symm_alloc(void** ptr, long long bytes);
int b200_symm_free(void* ptr);
int b200_ipc_handle_bytes(void);
int b200_ipc_get_handle(void* handle_out, void* ptr);
int b200_ipc_open_handle(void** peer_ptr, const void* handle);
int b200_ipc_close_handle(void* peer_ptr);
int b200_enable_peer_access(int peer_device);

/* Collectives on a symmetric buffer: the operand lives at `offset_bytes` inside every rank's allocation.
 * `flag_peers` is a symmetric allocation of b200_coll_flag_bytes() bytes reserved for one stream of collectives;
 * `epoch` must increase by one per call on that flag block (same value on every rank).  `blocks` = 0 picks a
 * size-dependent grid (<= 64 CTAs).  Sizes and offsets are multiples of 16 bytes; for all_gather /
 * reduce_scatter `bytes` is the FULL buffer and must be a multiple of 16 * world.
 *   all_reduce      in-place SUM of bf16 (test_nccl.py:25; train_ffns.py:165,:303,:309)
 *   all_gather      rank r's slice r is the input; every slice valid on return (test_nccl.py:16; train_ffns.py:203)
 *   reduce_scatter  every rank's full buffer is an input; slice r of rank r holds the SUM on return
 *                   (test_nccl.py:35; train_ffns.py:255-256)
 *   barrier         device-side cross-rank barrier (the role of mp.Barrier in test_mp_barrier_gpus.py:17) */
long long b200_coll_flag_bytes(void);
int b200_barrier(void* const* flag_peers, int rank, int world, unsigned int epoch, void* stream);
int b200_all_reduce_bf16(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                         int world, unsigned int epoch, int blocks, void* stream);
int b200_all_reduce_f32(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                        int world, unsigned int epoch, int blocks, void* stream);
int b200_reduce_scatter_f32(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers,
                            int rank, int world, unsigned int epoch, int blocks, void* stream);
int b200_all_gather(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers, int rank,
                    int world, unsigned int epoch, int blocks, void* stream);
/* dst (local, world * shard_bytes) <- every rank's shard at src_peers[q] + src_offset_bytes (FSDP unshard, :203). */
int b200_all_gather_into(void* dst, void* const* src_peers, long long src_offset_bytes, long long shard_bytes,
                         void* const* flag_peers, int rank, int world, unsigned int epoch, int blocks, void* stream);
/* Same, moved by the copy engines (P2P DMA) between two one-warp barrier kernels: uses no SM, so it overlaps GEMMs that
 * fill every SM.  The FSDP prefetch path. */
int b200_all_gather_into_dma(void* dst, void* const* src_peers, long long src_offset_bytes, long long shard_bytes,
                             void* const* flag_peers, int rank, int world, unsigned int epoch, void* stream);
int b200_reduce_scatter_bf16(void* const* peers, long long offset_bytes, long long bytes, void* const* flag_peers,
                             int rank, int world, unsigned int epoch, int blocks, void* stream);

/* ---- fused GEMM + collective -------------------------------------------------------------------------------
 * One tcgen05 kernel: computes D = A.B (operand conventions of b200_gemm_bf16), pushes each finished tile to the
 * inbox of the rank that owns it (peer stores over NVLink, overlapped with the next tile's MMAs), and the owner
 * sums the `world` partials in fp32, applies `op` and writes the result.
 *   owner_mode 0  tiles interleaved over ranks (all-reduce forms)      op 0  out  = sum
 *   owner_mode 1  owner = rank holding that block of rows (M/world)     op 1  out -= alpha * sum   (SGD on the weight)
 *   broadcast  1  result written to every rank's `out` and the kernel returns once the local `out` is complete
 *              0  result written to the local `out` only (reduce-scatter forms; `out` holds rows
 *                 [out_row_offset, out_row_offset + M/world))
 * `inbox_peers` / `flag_peers`: symmetric allocations of b200_fused_inbox_bytes(owned_cap, world) and
 * b200_fused_flag_bytes(owned_cap, ready_cap) bytes, reserved for one stream of fused calls ("channel");
 * owned_cap >= b200_fused_owned_tiles(M, N, world, owner_mode), ready_cap >= number of 128x256 tiles.
 * `seq` starts at 1 and increases by one per call on the channel (same on every rank).  Every rank must pass the
 * same shapes and max_ctas.  Reference sites: train_ffns.py:66/68+165+172 (DDP), :255-259 (FSDP), :57+303 and
 * :68+309 (MP). */
long long b200_fused_inbox_bytes(int owned_cap, int world);
long long b200_fused_flag_bytes(int owned_cap, int ready_cap);
int b200_fused_owned_tiles(int M, int N, int world, int owner_mode);
int b200_gemm_allreduce_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                             int M, int N, int K, void* const* out_peers, long long ldo, int out_row_offset,
                             void* const* inbox_peers, void* const* flag_peers, int owned_cap, int ready_cap,
                             int rank, int world, unsigned int seq, int owner_mode, int op, int broadcast,
                             float alpha, int max_ctas, void* stream);

/* Same kernel with an ADAM update in the owner's reduction instead of SGD (the north star's "sharded SGD/Adam update"):
 * out = W - lr * mhat / (sqrt(vhat) + eps), gradient = grad_scale * sum over ranks of A B^T, formula as b200_adam_bf16.
 * `adam_m` / `adam_v`: LOCAL fp32 moments of the tiles this rank owns, compact layout [owned slot][128][256]
 * (b200_fused_adam_state_bytes(owned_cap) bytes each, zeroed by the caller before the first step): every rank holds
 * 1/world of the optimizer state, for DDP as well as FSDP.  `step` >= 1 (bias correction).  No counterpart in the
 * reference (SGD only, train_ffns.py:172).  */
long long b200_fused_adam_state_bytes(int owned_cap);
/* Kernel variant for the fused calls of this process: side_warps 2 (384 threads, 168 registers) or 4 (448 threads, 128
 * registers; default); overlap -1 = per call, 0 = CTA-pair mainloop, 1 = double-buffered 1-CTA mainloop.  Out-of-range
 * values leave a setting unchanged.  Returns side_warps * 10 + (overlap + 1).  Every rank must use the same settings. */
int b200_fused_config(int side_warps, int overlap);
int b200_gemm_allreduce_adam_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                                  int M, int N, int K, void* const* out_peers, long long ldo, int out_row_offset,
                                  void* const* inbox_peers, void* const* flag_peers, int owned_cap, int ready_cap,
                                  int rank, int world, unsigned int seq, int owner_mode, int broadcast,
                                  float* adam_m, float* adam_v, float lr, float beta1, float beta2, float eps, int step,
                                  float grad_scale, int max_ctas, void* stream);

/* ---- all-gather -> GEMM in one kernel (csrc/gather_gemm.cu) -------------------------------------------------------
 * C[M,N] = epilogue(A_full[M,K] * B^T) where A is token-sharded: a_peers[q] is rank q's [M,K] bf16 buffer (K-major,
 * contiguous, 16-byte aligned; 128 for whole-line chunks) in which only rows [q*M/world, (q+1)*M/world) are
 * final when q's call starts (a_peers[q] for q != rank is only ever read inside those rows, so it may be a shard's
 * address minus q shards: that is how FSDP's row-sharded weights are gathered, parallel/fused_fsdp_t.py).  The kernel pulls the other shards into a_peers[rank] (complete when it returns) while
 * the MMAs start on the local rows.  M % (128*world) == 0, K % 64 == 0.  B / epilogue / aux / out2 / bias / alpha as
 * b200_gemm_bf16 (epilogues STORE, RELU, GELU, RELU_BWD, GELU_BWD, ADD, RELU_MASK, RELU_BWD_MASK).
 * a_major = 1 (MN-major) selects the second form: A is stored [K, M] and sharded along K (FSDP's backward pass, where
 * the sharded dimension of the weight is the contracted one); K % (128*world) == 0, M % 64 == 0; every tile's k-loop then
 * starts at the rank's own k-range and waits per k-block.  `ready` holds one counter per 128 storage rows in both forms
 * (b200_gather_gemm_ready_bytes(M) or (K)).
 * `started_peers[q]`: symmetric array of >= `world` uint32 reserved for this stream of calls, zero-initialised;
 * `seq` starts at 1 and increases by one per call (same on every rank).  `ready`: LOCAL scratch of
 * b200_gather_gemm_ready_bytes(M) bytes, cleared by the call itself (stream-ordered).
 * The kernel never waits for peers to finish reading a shard: the caller must not overwrite its shard before a later
 * collective that every rank enters after this call (docs/KERNELS.md 2.5).
 * No counterpart in the reference (it all-reduces, train_ffns.py:303,:309); this is the sequence-parallel form
 * SURVEY.md 2.6 plans for C4/C5.  */
long long b200_gather_gemm_ready_bytes(int M);
int b200_gather_gemm_bf16(void* const* a_peers, int a_major, const void* B, long long ldb, int b_major, void* C, long long ldc,
                          int M, int N, int K, int epilogue, const void* aux, long long ldaux, void* out2, long long ldc2,
                          const float* bias, float alpha, void* const* started_peers, void* ready, int rank, int world,
                          unsigned int seq, int max_ctas, void* stream);
/* host mirror of the pull order (CPU tests): out = {peer, first 16-byte vector, 128-row block, vectors per chunk,
 * chunks per block, total chunks} for pull chunk `u` */
int b200_debug_gather_chunk(int M, int K, int rank, int world, long long u, long long* out);
/* host mirror of the row-sharded form's rotated pair-tile walk (CPU tests); returns the number of pair tiles */
int b200_debug_gather_tile_walk(int M, int N, int rank, int world, int* first_row, int* n_blk, int capacity);
/* host mirror of the K-sharded form's k-loop walk: out = {k-block loaded in iteration kb, ring position of its 128-row
 * storage block, that block's index, kb_rot, num_kb, blocks per shard} */
int b200_debug_gather_kwalk(int K, int rank, int world, int kb, int* out);

/* ---- several ranks of the fused kernels as slices of ONE cooperative launch on one GPU (tests, ncu) ---------------------
 * Kernels that spin on one another's flags cannot be separate launches on one device (nothing makes them co-resident),
 * so CTAs [r * ctas_per_rank, (r + 1) * ctas_per_rank) of a single grid run rank r's kernel body - its own operands, owner
 * maps, tile walk, flags - against peer tables whose entries are all local memory.  Per-rank inputs are HOST arrays of
 * `world` device pointers; peer tables as in the real entry points; `table`: b200_emu_table_bytes() bytes of device
 * scratch; ctas_per_rank even, world * ctas_per_rank <= SM count.  adam_m == NULL selects `op`, else Adam.  No counterpart
 * in the reference (its multi-rank paths need as many GPUs as ranks). */
long long b200_emu_table_bytes(void);
int b200_emu_gemm_allreduce_bf16(int world, const void* const* A, long long lda, int a_major, const void* const* B, long long ldb,
                                 int b_major, int M, int N, int K, void* const* out_peers, long long ldo,
                                 void* const* inbox_peers, void* const* flag_peers, int owned_cap, int ready_cap,
                                 unsigned int seq, int owner_mode, int op, int broadcast, float alpha,
                                 float* const* adam_m, float* const* adam_v, float lr, float beta1, float beta2, float eps,
                                 int step, float grad_scale, int ctas_per_rank, void* table, void* stream);
int b200_emu_gather_gemm_bf16(int world, void* const* a_peers, int a_major, const void* const* B, long long ldb, int b_major,
                              void* const* C, long long ldc, int M, int N, int K, int epilogue, const void* const* aux,
                              long long ldaux, void* const* out2, long long ldc2, void* const* started_peers,
                              void* const* ready, unsigned int seq, int ctas_per_rank, void* table, void* stream);

/* ---- cuMem VMM blocks + NVLS multicast (csrc/vmm.cu) -----------------------------------------------------------
 * The runtime the in-switch reductions need: symmetric blocks the fused MP trainer allocates with
 * SymmWorld.alloc(multicast=True) where the box has NVLS.  Host-side fd passing is unit-tested on CPU; the device side
 * runs in tests/test_multigpu.py::test_vmm_* (B200, 2 GPUs: profiles/tests/).
 * Descriptor passing between the ranks of one node (SCM_RIGHTS over a unix datagram socket): */
int b200_fd_listen(const char* path);              /* returns the listening socket fd */
int b200_fd_send(const char* path, int fd, int tag);
int b200_fd_recv(int sock, int* tag_out);          /* returns the received fd */
int b200_fd_close(int fd);
/* Shareable blocks: out[0] = POSIX-fd export supported, out[1] = multicast supported (current device). */
int b200_vmm_query(int* out);
long long b200_vmm_granularity(int world, int with_multicast);
int b200_vmm_alloc(long long bytes, void** ptr_out, unsigned long long* handle_out, int* fd_out);
int b200_vmm_import(int fd, long long bytes, void** ptr_out, unsigned long long* handle_out);
int b200_vmm_free(void* ptr, long long bytes, unsigned long long handle);
/* Multicast object: rank 0 creates + exports, the others import; then every rank attaches in two collective steps
 * (1: add device; host barrier; 2: bind its block, map the multicast VA -> *ptr_out). */
int b200_mc_create(int world, long long bytes, unsigned long long* mc_out, int* fd_out);
int b200_mc_import(int fd, unsigned long long* mc_out);
int b200_mc_attach(unsigned long long mc, int step, unsigned long long mem_handle, long long bytes, void** ptr_out);
/* In-place bf16 all-reduce through the switch: multimem.ld_reduce (fp32 accumulate) of my slice, multimem.st of the
 * result to every replica.  Counterpart of train_ffns.py:165/:303 on NVLS. */
int b200_all_reduce_nvls_bf16(void* mc_ptr, long long bytes, void* const* flag_peers, int rank, int world,
                              unsigned int epoch, int blocks, void* stream);
/* In-place NVLS collectives on `bytes` of a multicast-bound block: op 0 all-reduce (one-shot below 512 KiB, else
 * two-shot), 1 reduce-scatter (slice `rank` of the local copy holds the sum), 2 all-gather (slice `rank` is the input).
 * f32 != 0: fp32 payload, else bf16 with fp32 accumulation in the switch.  `mc_ptr` / `local_ptr`: the operand through
 * the multicast address / this rank's own mapping.  Counterparts of test_nccl.py:16,25,35 on NVLS. */
int b200_nvls_collective(int op, int f32, void* mc_ptr, void* local_ptr, long long bytes, void* const* flag_peers, int rank,
                         int world, unsigned int epoch, int blocks, void* stream);
/* NVLS form of b200_gemm_allreduce_bf16 (op 0 SUM or op 1 with alpha = -1: in-place add; no optimizer): partial tiles
 * stay in this rank's `part_local` (b200_fused_partial_bytes(ready_cap) bytes of a block every rank bound to one
 * multicast object, address `part_mc`); the owner sums them in the switch (multimem.ld_reduce, fp32 accumulate) and, with
 * broadcast != 0, stores the result through `out_mc`, the output's multicast address.  Reference sites: :303, :309. */
long long b200_fused_partial_bytes(int ready_cap);
int b200_gemm_allreduce_nvls_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                                  int M, int N, int K, void* const* out_peers, long long ldo, int out_row_offset,
                                  void* part_local, const void* part_mc, void* out_mc, void* const* flag_peers,
                                  int owned_cap, int ready_cap, int rank, int world, unsigned int seq, int owner_mode,
                                  int op, int broadcast, float alpha, int max_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200FFN_H */
