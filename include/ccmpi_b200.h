/*
 * ccmpi_b200 -- C ABI of the native runtime and the sm_100a kernels.
 *
 * One shared library (libccmpi_b200.so) holds the device runtime (symmetric heap, peer mapping,
 * NVLS multicast, signal pads, device-resident communicators) and every CUDA kernel; a second
 * one (libccmpi_host.so, see ccmpi_host.h) is the CPU shared-memory transport. Python binds
 * both with ctypes, so every entry point takes plain pointers / integers / a cudaStream_t.
 *
 * Reference parity notes cite files of anaykulkarni/collective-communication-mpi.
 * All functions return 0 on success and a negative ccmpi_status on failure;
 * ccmpi_last_error() returns a human readable message for the calling thread.
 */
#ifndef CCMPI_B200_H_
#define CCMPI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCMPI_MAX_RANKS 8   /* one NVSwitch domain */
#define CCMPI_MAX_VRANKS 4  /* emulated ranks per single-GPU launch (tests) */
#define CCMPI_MAX_BLOCKS 160
/* OR into a collective's `max_blocks`: even grid, launched as 2-CTA clusters (whole SM pairs), for a
 * collective that runs next to a CTA-pair GEMM whose CTAs wait on peers */
#define CCMPI_BLOCKS_PAIRED 0x10000

/* signal pad layout, in uint32 words, one pad per (communicator, rank) */
#define CCMPI_PAD_WORDS 16384                     /* 64 KiB */
#define CCMPI_PAD_BARRIER 0                       /* [block][src rank] epochs */
#define CCMPI_PAD_EPOCH (CCMPI_MAX_BLOCKS * CCMPI_MAX_RANKS) /* [block] local epoch */
#define CCMPI_PAD_USER 2048                       /* fused kernels: chunk / tile flags */
#define CCMPI_PAD_USER_WORDS (CCMPI_PAD_WORDS - CCMPI_PAD_USER)

typedef enum {
    CCMPI_OK = 0,
    CCMPI_ERR_CUDA = -1,
    CCMPI_ERR_ARG = -2,
    CCMPI_ERR_UNSUPPORTED = -3,
    CCMPI_ERR_SYS = -4,
    CCMPI_ERR_LIMIT = -5
} ccmpi_status;

typedef enum {
    CCMPI_F32 = 0,
    CCMPI_BF16 = 1,
    CCMPI_F16 = 2,
    CCMPI_F64 = 3,
    CCMPI_I32 = 4,
    CCMPI_I64 = 5,
    CCMPI_U8 = 6
} ccmpi_dtype;

/* MPI.SUM / MPI.MIN / MPI.MAX of the reference (mpi_wrapper/comm.py:88-93) */
typedef enum { CCMPI_SUM = 0, CCMPI_MIN = 1, CCMPI_MAX = 2 } ccmpi_op;

typedef enum {
    CCMPI_AR_AUTO = 0,
    CCMPI_AR_ONESHOT = 1, /* every rank pulls every peer, rank-ordered reduce (deterministic) */
    CCMPI_AR_TWOSHOT = 2, /* reduce own slice from peers, then gather slices */
    CCMPI_AR_NVLS = 3,    /* multimem.ld_reduce + multimem.st through the NVSwitch */
    CCMPI_AR_TWOSHOT_GENERIC = 4 /* two-shot with the kernel sized for any group (TWOSHOT uses the group-size-
                                    specialised kernel when dtype/alignment allow: +3..30 % measured, same bits) */
} ccmpi_ar_algo;

/*
 * Device-resident communicator: what a kernel needs to reach its peers. `heap[i]` is member i's
 * symmetric segment as mapped into THIS process, so `heap[i] + off` names the same logical
 * buffer on every member. Replaces MPI.Comm handles (comm.py:5-7) and the result of
 * MPI_Comm_split (comm.py:38-39) on the device side.
 */
typedef struct ccmpi_dev_comm {
    int32_t rank;
    int32_t size;
    unsigned char* heap[CCMPI_MAX_RANKS];
    unsigned char* mc_heap;          /* multicast alias of the member heaps, or NULL */
    uint32_t* pad[CCMPI_MAX_RANKS];  /* member i's signal pad for this communicator */
} ccmpi_dev_comm;

const char* ccmpi_last_error(void);

/* Every device-side wait (flag spins across GPUs, mbarrier waits) gives up with a printed message and a
 * trap after this long; default 120 s. Applies to the current device, to kernels launched afterwards.
 * (The reference's MPI calls wait forever, mpi_wrapper/comm.py:15-16; a GPU kernel that does so cannot be
 * stopped, hence a generous, adjustable limit instead.) */
int ccmpi_set_spin_timeout_ns(unsigned long long ns);
int ccmpi_version(void);

/* ------------------------------------------------------------------------------------------
 * Runtime (csrc/runtime.cpp): symmetric heap over cuMemCreate / POSIX-fd export / cuMemMap,
 * fd passing over abstract unix sockets, NVLS multicast objects.
 * ---------------------------------------------------------------------------------------- */

/* Driver/device capability probe; fills a small int array:
 * [0] device count, [1] cc major, [2] cc minor, [3] SM count, [4] multicast supported,
 * [5] VMM supported, [6] posix-fd handle supported, [7] heap granularity (bytes, low 31 bits) */
int ccmpi_probe(int device, int* out8);

/* Create this rank's segment of a symmetric heap of `bytes` (rounded up to the granularity) on
 * `device`, and map it. Returns an opaque heap handle (>0). */
int64_t ccmpi_heap_create(int device, size_t bytes);
size_t ccmpi_heap_bytes(int64_t heap);
void* ccmpi_heap_local_ptr(int64_t heap);

/* Exchange segment handles with the other `nranks-1` processes of the job and map their
 * segments. `job` names the rendezvous (abstract unix sockets "\0ccmpi-<job>-<rank>"), the
 * python layer guarantees every rank is listening before anyone connects. */
int ccmpi_heap_listen(int64_t heap, const char* job, int rank);
int ccmpi_heap_connect_all(int64_t heap, const char* job, int rank, int nranks);
void* ccmpi_heap_peer_ptr(int64_t heap, int peer);

/* NVLS multicast over a subset of ranks (`members`, world ranks, sorted): the group's first
 * member creates the object and ships its fd; all members add their device, bind their whole
 * segment and map the multicast VA. Phases are separate because cuMulticastBindMem must wait
 * until every member has added its device (python barriers in between).
 * Returns a group id >= 0, or <0 (CCMPI_ERR_UNSUPPORTED when the platform has no multicast). */
int ccmpi_mc_create(int64_t heap, const char* job, int rank, const int* members, int nmembers,
                    int group_tag);
int ccmpi_mc_bind(int64_t heap, int group);
void* ccmpi_mc_ptr(int64_t heap, int group);
int ccmpi_heap_destroy(int64_t heap);

/* Emulated "heap" for single-GPU tests: nranks plain device buffers on one device. */
int64_t ccmpi_heap_create_emulated(int device, size_t bytes, int nranks);

/* ------------------------------------------------------------------------------------------
 * Collective kernels (csrc/collectives.cu). `comms` points to `nvr` device communicators in
 * HOST memory (nvr = 1 in production; nvr > 1 runs that many emulated ranks inside one
 * cooperative launch on one GPU). Buffers are heap offsets (bytes) unless named *_ptr.
 * ---------------------------------------------------------------------------------------- */
int ccmpi_barrier(const ccmpi_dev_comm* comms, int nvr, void* stream);

/* MPI_Allreduce (comm.py:22) and the reduce-then-broadcast myAllreduce (comm.py:63-107),
 * re-done as one-shot / two-shot / NVLS kernels. dst is a heap offset too. `scale` is applied
 * to float results (1.0f = none) and `out_dtype` may differ from `dtype` for float types
 * (DP gradient path: sum, scale by 1/dp, cast). tmp_off: count*sizeof(dtype) bytes of symmetric
 * scratch for the two-shot variant. */
int ccmpi_allreduce(const ccmpi_dev_comm* comms, int nvr, int algo, int dtype, int op,
                    size_t src_off, size_t dst_off, size_t tmp_off, size_t count, float scale,
                    int out_dtype, int max_blocks, void* stream);

/* MPI_Allgather (comm.py:29). Generalised to a 2-D strided form so the last-axis gathers of
 * func_impl.py:89-90,107-108 need no transposes: each rank contributes `rows` rows of
 * `row_bytes`; in dst, row i of rank r lands at dst + i*dst_row_stride + r*row_bytes.
 * (flat MPI semantics: rows = 1, row_bytes = block bytes.) */
int ccmpi_allgather(const ccmpi_dev_comm* comms, int nvr, size_t src_off, size_t dst_off,
                    size_t rows, size_t row_bytes, size_t src_row_stride, size_t dst_row_stride,
                    int use_mc, int max_blocks, void* stream);

/* MPI_Reduce_scatter_block (comm.py:36) with the same 2-D form for func_impl.py:182-186:
 * src has `rows` rows of size*row_bytes; rank r gets column block r of every row, reduced. */
int ccmpi_reduce_scatter(const ccmpi_dev_comm* comms, int nvr, int dtype, int op, size_t src_off,
                         size_t dst_off, size_t rows, size_t row_elems, size_t src_row_stride,
                         size_t dst_row_stride, int use_mc, int max_blocks, void* stream);

/* MPI_Alltoall (comm.py:61) / myAlltoall (comm.py:110-159): block j of src goes to rank j's
 * dst block `rank`; every peer push is in flight at once. staggered != 0 uses the pairwise
 * (rank+i)%n visiting order of myAlltoall2 (comm.py:162-199). */
int ccmpi_alltoall(const ccmpi_dev_comm* comms, int nvr, size_t src_off, size_t dst_off,
                   size_t seg_bytes, int staggered, int max_blocks, void* stream);

/* put-with-signal mailbox behind Send/Recv/Isend/Irecv/Sendrecv (comm.py:86-193) */
int ccmpi_put_signal(const ccmpi_dev_comm* comms, int nvr, int peer, size_t src_off,
                     size_t dst_off, size_t bytes, int flag_index, void* stream);
int ccmpi_wait_signal(const ccmpi_dev_comm* comms, int nvr, int peer, int flag_index,
                      void* stream);

/* ------------------------------------------------------------------------------------------
 * tcgen05 GEMM and fused GEMM + collective kernels (csrc/gemm_tcgen05.cu)
 * C[M,N] = A[M,K] * B[K,N], bf16 inputs, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A stored row-major [M,K]     1: stored row-major [K,M]
 *   b_mn_major = 0: B stored row-major [N,K]     1: stored row-major [K,N]
 *   (the FC GEMMs absent from the reference: y = x W, dx = dy W^T, dW = x^T dy)
 * ---------------------------------------------------------------------------------------- */
int ccmpi_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                    int ldc, int a_mn_major, int b_mn_major, int out_dtype, int accumulate,
                    int num_sms, void* stream);
/* The same GEMM with elementwise work fused into its epilogue, so the activation / residual passes between
 * the FC GEMMs cost no extra trip through HBM (bf16 C, N % 32 == 0, CTA-pair kernel):
 *   epi_op 1: C = A*B and epi_out2 = gelu(C)            (tanh-approximated; C is kept for backward)
 *          2: C = (A*B) * gelu'(epi_in)                 (backward through the activation)
 *          3: C = A*B + epi_in                          (residual / gradient accumulation)
 * epi_in / epi_out2: [M][epi_ld] bf16. `trace`: NULL, or 16 uint64 per CTA that the kernel fills with where
 * each warp role spent its cycles (utils/ktrace.py decodes them). */
int ccmpi_gemm_bf16_ex(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                       int a_mn_major, int b_mn_major, int out_dtype, int accumulate, int epi_op, const void* epi_in,
                       void* epi_out2, long long epi_ld, void* trace, int num_sms, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused GEMM + collective kernels (csrc/fused_gemm.cu, csrc/fused_policies.cuh). One kernel
 * computes tcgen05 tiles AND moves them / their operands over NVLink (peer stores or NVLS
 * multimem) so the transfer overlaps the math tile by tile. Per-rank operand pointers come as
 * arrays of `nvr` entries; buffers that peers touch are symmetric-heap offsets. `flag_base`
 * selects a region of the signal pad's user words, `epoch` must grow by one per launch on a
 * given (communicator, flag_base) and be the same on every rank; buffers that peers write into
 * (x_full, slots) should alternate between two copies from launch to launch.
 * ---------------------------------------------------------------------------------------- */

/* all-gather -> GEMM on the gathered K axis (func_impl.py:76-91 + the consuming FC GEMM):
 * C[M][N] = [x_0 | ... | x_{n-1}] * B with x_r = rank r's [M][K/n] shard at x_off (row stride
 * ldx elements); xfull_off: [M][K] staging the peers push into.
 *
 * Ready flags count the pushing CTAs of each source rank. `flag_target` is the value the flags must
 * reach for this launch: the caller's running sum, over all launches on these flags, of
 * ccmpi_fused_ag_ctas() -- so shapes and SM budgets may change between launches. With flag_target = 0
 * the kernel waits for epoch * (this launch's CTA count), which is only right when every launch on
 * these flags has that count. */
int ccmpi_fused_ag_gemm(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                        const void* const* B, void* const* C, int M, int N, int K, int ldb, int ldc,
                        int b_kn, int out_dtype, int flag_base, unsigned epoch, unsigned flag_target,
                        int num_comm_ctas, int num_sms, int use_mc, void* stream);
/* kind 0: K-gather kernels (ccmpi_fused_ag_gemm, ccmpi_fused_ag_gemm_ag), 1: M-gather (ccmpi_fused_ag_gemm_m) */
int ccmpi_fused_ag_ctas(int kind, int M, int N, int nvr, int num_sms);
/* ccmpi_fused_ag_gemm with a fused elementwise epilogue (as ccmpi_gemm_bf16_ex; one epi_in / epi_out2 pointer
 * per rank, or NULL) and an optional device trace. */
int ccmpi_fused_ag_gemm_ex(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                           const void* const* B, void* const* C, int M, int N, int K, int ldb, int ldc, int b_kn,
                           int out_dtype, int flag_base, unsigned epoch, unsigned flag_target, int epi_op,
                           const void* const* epi_in, void* const* epi_out2, long long epi_ld, void* trace,
                           int num_comm_ctas, int num_sms, int use_mc, void* stream);
/* ccmpi_fused_gemm_rs whose reducer adds `residual` ([M][N/n] bf16 per rank, row stride res_ld; NULL: none) to
 * the reduced result before the single rounding, and an optional device trace. */
int ccmpi_fused_gemm_rs_ex(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                           size_t slots_off, void* const* C, int M, int N, int K, int lda, int ldb, int ldc, int a_km,
                           int b_kn, int out_dtype, int flag_base, unsigned epoch, const void* const* residual,
                           long long res_ld, void* trace, int num_sms, void* stream);

/* GEMM -> reduce-scatter on N (the dx GEMM + func_impl.py:150-187): C[M][N/n] = rank-ordered sum
 * over ranks of (A*B)[:, own column block]; slots_off: [n][M][N/n] staging (out dtype). */
int ccmpi_fused_gemm_rs(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                        size_t slots_off, void* const* C, int M, int N, int K, int lda, int ldb, int ldc,
                        int a_km, int b_kn, int out_dtype, int flag_base, unsigned epoch, int num_sms,
                        void* stream);

/* GEMM -> all-reduce (row-parallel fc_o / fc2 forward): out_off: symmetric [M][N] result. */
int ccmpi_fused_gemm_ar(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                        size_t slots_off, size_t out_off, int M, int N, int K, int lda, int ldb, int a_km,
                        int b_kn, int out_dtype, int flag_base, unsigned epoch, int num_sms, int use_mc,
                        void* stream);

/* GEMM whose epilogue all-gathers the output on N (the FC GEMM + func_impl.py:94-109):
 * yfull_off: symmetric [M][N*n]; this rank's A*B lands in column block `rank` on every member. */
int ccmpi_fused_gemm_ag(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                        size_t yfull_off, int M, int N, int K, int lda, int ldb, int a_km, int b_kn,
                        int out_dtype, int num_sms, int use_mc, void* stream);

/* Run on B200s in round 2 (tests/test_gpu_seq.py): the token-axis ("sequence
 * parallel") forms of the two kernels above (csrc/fused_gemm_m.cu, csrc/fused_policies_m.cuh).
 * The reference shards only the feature axis; these cover layouts that shard the activation rows.
 * M/n must be a multiple of 256. CTA-pair kernels only.
 *   ag_gemm_m: C[M][N] = [x_0 ; ... ; x_{n-1}] * B, x_r = rank r's [M/n][K] rows at x_off (row
 *              stride ldx); xfull_off: symmetric [M][K] that the members' rows are pushed into.
 *   gemm_rs_m: C[M/n][N] = rank-ordered sum over ranks of (A*B)[own rows]; slots_off: symmetric
 *              [n][M/n][N] staging of the output dtype. */
int ccmpi_fused_ag_gemm_m(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                          const void* const* B, void* const* C, int M, int N, int K, int ldb, int ldc,
                          int b_kn, int out_dtype, int flag_base, unsigned epoch, unsigned flag_target, int num_sms,
                          void* stream);
int ccmpi_fused_gemm_rs_m(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                          size_t slots_off, void* const* C, int M, int N, int K, int lda, int ldb, int ldc,
                          int a_km, int b_kn, int out_dtype, int flag_base, unsigned epoch, int num_sms,
                          void* stream);

/* Run on B200s in round 2 (tests/test_gpu_seq.py): the reference's whole forward step
 * -- gather the input, multiply, gather the output (func_impl.py:76-109) -- in one kernel:
 * ccmpi_fused_ag_gemm's K-gather plus ccmpi_fused_gemm_ag's epilogue. yfull_off: symmetric
 * [M][N*n]; x_off / xfull_off as in ccmpi_fused_ag_gemm. */
int ccmpi_fused_ag_gemm_ag(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                           const void* const* B, size_t yfull_off, int M, int N, int K, int ldb, int b_kn,
                           int out_dtype, int flag_base, unsigned epoch, unsigned flag_target, int num_sms,
                           void* stream);

/* Run on B200s in round 2 (tests/test_gpu_elementwise.py; default in the models): fused elementwise kernels of the
 * transformer-FC block (csrc/elementwise.cu), bf16 in / bf16 out, fp32 math, one pass over HBM each.
 * The reference has no compute; these sit between the GEMMs of models/fc_stack.py.
 *   gate:  a[T][H] = v * sigmoid(q * k) with qkv[T][3H] = q | k | v; backward writes dq | dk | dv.
 *   gelu:  tanh approximation, forward and backward, n contiguous elements.
 *   mse:   loss[0] = mean((y - t)^2) and dy = 2 (y - t) / n in one pass; `workspace` holds
 *          ccmpi_mse_workspace_bytes() bytes, zeroed once by the caller before the first call. */
int ccmpi_gate_fwd(const void* qkv, long long ld_qkv, void* a, long long ld_a, int T, int H, void* stream);
int ccmpi_gate_bwd(const void* da, long long ld_da, const void* qkv, long long ld_qkv, void* dqkv,
                   long long ld_d, int T, int H, void* stream);
int ccmpi_gelu_fwd(const void* u, void* y, long long n, void* stream);
int ccmpi_gelu_bwd(const void* dy, const void* u, void* du, long long n, void* stream);
size_t ccmpi_mse_workspace_bytes(void);
int ccmpi_mse_fwd_bwd(const void* y, const void* t, void* dy, long long n, void* workspace, void* loss,
                      void* stream);

/* Run and measured in round 2 (tests/test_gpu_tight.py; the default): the pull-style reduce-scatter and the two-shot
 * all-reduce with their in-flight buffers sized by the communicator (2 / 4 / 8 members) instead of for 8
 * (csrc/collectives_tight.cu). fp32 / bf16, SUM / MIN / MAX, 16-byte aligned operands; anything else returns
 * CCMPI_ERR_UNSUPPORTED. Results are bit-identical to ccmpi_reduce_scatter / ccmpi_allreduce(CCMPI_AR_TWOSHOT). */
int ccmpi_reduce_scatter_tight(const ccmpi_dev_comm* comms, int nvr, int dtype, int op, size_t src_off, size_t dst_off,
                               size_t rows, size_t row_elems, size_t src_row_stride, size_t dst_row_stride,
                               int max_blocks, void* stream);
int ccmpi_allreduce_twoshot_tight(const ccmpi_dev_comm* comms, int nvr, int dtype, int op, size_t src_off, size_t dst_off,
                                  size_t tmp_off, size_t count, float scale, int max_blocks, void* stream);

/* Weight-gradient GEMM with the SGD update in its epilogue, W[M][N] -= lr * (A * B), bf16 W updated in
 * place from the fp32 accumulator (csrc/gemm_sgd.cu). For layers whose gradient needs no data-parallel
 * reduction. Operand layout flags as in ccmpi_gemm_bf16. The _sr form rounds the new weights
 * stochastically (step >= 0; random bits from (element, step, seed)) so that updates below one bf16 ulp
 * are not lost; step < 0 = round to nearest, which is what the plain form does. The reference has no
 * optimizer; it is part of the training step this framework adds around its communicators. */
int ccmpi_gemm_sgd_bf16(const void* A, const void* B, void* W, int M, int N, int K, int lda, int ldb, int ldw,
                        int a_mn_major, int b_mn_major, float lr, int num_sms, void* stream);
int ccmpi_gemm_sgd_bf16_sr(const void* A, const void* B, void* W, int M, int N, int K, int lda, int ldb, int ldw,
                           int a_mn_major, int b_mn_major, float lr, int step, unsigned seed, int num_sms, void* stream);
/* w[i] = SR(w[i] - lr * g[i]) on a flat run of n bf16 weights (n % 8 == 0): the optimizer pass for gradients
 * that came out of a collective. base_index = position of w[0] in the model's flat parameter numbering. */
int ccmpi_sgd_sr_bf16(void* w, const void* g, long long n, float lr, unsigned long long base_index, unsigned step,
                      unsigned seed, void* stream);
/* Data-parallel gradient all-reduce fused with that update (NVLS, bf16): every member sums its slice of all
 * replicas' gradients in the switch, updates its slice of the weights and multicasts the new weights to every
 * replica. grad_off / w_off: symmetric offsets of the bucket's gradients / weights; lr_eff = lr / dp.
 * (The data-parallel all-reduce the reference implies with dp_comm, func_impl.py:60-62.) */
int ccmpi_allreduce_sgd_nvls(const ccmpi_dev_comm* comms, int nvr, size_t grad_off, size_t w_off, size_t count,
                             float lr_eff, unsigned long long base_index, unsigned step, unsigned seed, int max_blocks,
                             void* stream);

/* Block-scaled fp8 (OCP MXFP8: e4m3 values, one UE8M0 scale per 1 x 32 block of K) on the generic CTA-pair
 * kernel, plain and fused with the collectives (csrc/fused_gemm_mx.cu). Operands K-major: A [M][K], B [N][K];
 * scale arrays in the chunked layout the tensor cores read (512-byte chunks of 128 rows x 4 scales, K-fastest;
 * ccmpi_quantize_mxfp8 writes it). BASELINE config 3's fp8 forward products of func_impl.py:76-91 / :150-187.
 *   ag_gemm_mx: xq_off / sfl_off = this rank's [M][K/n] shard and its scales, xfull_off / sffull_off = gathered
 *               staging; the kernel pushes payload AND scales (half the bytes of the bf16 gather).
 *   gemm_rs_mx: fp8 operands, partial tiles and result in out_dtype; optional bf16 residual as in gemm_rs_ex. */
int ccmpi_gemm_mx(const void* A, const void* sfa, const void* B, const void* sfb, void* C, int M, int N, int K, int lda,
                  int ldb, int ldc, int out_dtype, int num_sms, void* stream);
int ccmpi_fused_ag_gemm_mx(const ccmpi_dev_comm* comms, int nvr, size_t xq_off, int ldx, size_t sfl_off, size_t xfull_off,
                           size_t sffull_off, const void* const* B, const void* const* sfb, void* const* C, int M, int N,
                           int K, int ldb, int ldc, int out_dtype, int flag_base, unsigned epoch, unsigned flag_target,
                           void* trace, int num_sms, void* stream);
int ccmpi_fused_gemm_rs_mx(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* sfa,
                           const void* const* B, const void* const* sfb, size_t slots_off, void* const* C, int M, int N,
                           int K, int lda, int ldb, int ldc, int out_dtype, int flag_base, unsigned epoch,
                           const void* const* residual, long long res_ld, void* trace, int num_sms, void* stream);
int ccmpi_quantize_mxfp8(const void* x, long long ldx, void* q, long long ldq, void* sf, int R, int K, void* stream);

/* Run on a B200 in round 2 (tests/test_gpu_fp8.py): fp8 e4m3 GEMM on the CTA-pair
 * kernel, C[M][N] = scale * A[M][K] * B[N][K]^T, both operands K-major (tcgen05.mma.kind::f8f6f4).
 * First step towards the block-scaled fp8 FC layers of BASELINE.md config 3; the reference has no
 * GEMM at all. */
int ccmpi_gemm_fp8_e4m3(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                        int ldc, float scale, int out_dtype, int num_sms, void* stream);

/* Run on a B200 in round 2 (tests/test_gpu_fp8.py, tests/test_gpu_mx_fused.py): block-scaled fp8 GEMM (OCP MXFP8:
 * e4m3 values, one UE8M0 scale per 1 x 32 block), C[M][N] = (A * 2^sfa) (B * 2^sfb)^T with both operands
 * K-major, on tcgen05.mma.kind::mxf8f6f4.block_scale (csrc/gemm_mxfp8.cu). sfa / sfb are in the tensor
 * cores' chunked layout: 512-byte chunks of 128 rows x 4 scales, K-fastest, padded to whole chunks
 * (collective_communication_mpi_b200/ops/gemm_fp8.py::swizzle_scales). K must be a multiple of 128. */
int ccmpi_gemm_mxfp8(const void* A, const void* sfa, const void* B, const void* sfb, void* C, int M, int N,
                     int K, int lda, int ldb, int ldc, int out_dtype, int num_sms, void* stream);

/* No counterpart in the reference: it never allocates device memory because it has no device
 * code. This is the GEMM -> reduce-scatter kernel of ccmpi_fused_gemm_rs for a ONE-GPU
 * communicator (C[M][N] = A * B), for callers that must not allocate device memory when they
 * call -- servers that gave the GPU to their own arena, plug-ins under another framework's
 * memory planner, loops that cannot take cudaMalloc's device-wide synchronisation. Its staging
 * buffer and flag words live in a static device array inside the library (64 KiB + 16 MiB, held
 * by every GPU that loads the library until the process exits), so the caller passes only A, B
 * and C; nothing is allocated, by this call or by anything it uses.
 *   - M * N * sizeof(output element) must be <= ccmpi_gemm_rs_local_noalloc_max_bytes() (16 MiB);
 *     larger problems return CCMPI_ERR_LIMIT. N must be a multiple of 128.
 *   - All calls on a device share the array and it keeps the previous call's contents; each call
 *     zeroes the flag words on `stream` first. Calls that overlap on the GPU race on it.
 *   - The first call on each device resolves the array's device address; make that call before
 *     any stream capture on that device.
 * Operand layout flags as in ccmpi_gemm_bf16. */
int ccmpi_gemm_rs_local_noalloc(const void* A, const void* B, void* C, int M, int N, int K, int lda,
                                int ldb, int ldc, int a_km, int b_kn, int out_dtype, void* stream);
size_t ccmpi_gemm_rs_local_noalloc_max_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* CCMPI_B200_H_ */
