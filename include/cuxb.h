/*
 * cuxb.h -- C ABI of libcuxb.so: B200 (sm_100a) kernels for the cux operator set.
 *
 * This is the drop-in boundary.  Everything above it (the cux-compatible C++
 * host layer in hpc_b200/cux/, the Python ctypes mirror in hpc_b200/capi.py, or
 * the reference's own plugins -- see INTEGRATION.md) reaches CUDA only through
 * these functions: plain pointers and sizes, no CUDA, torch or C++ types.
 *
 * Conventions
 *   - every function returns 0 on success, else a non-zero code (a cudaError_t
 *     value, or one of CUXB_ERR_*); cuxb_last_error_string() describes the most
 *     recent failure on the calling thread.  Nothing throws.  The reference's
 *     behaviour for the same failures is log + exit/abort
 *     (0-frameworks/cux/include/util/util.h:69-95); the C++ layer maps non-zero
 *     codes onto that.
 *   - device pointers are opaque `void*` / typed pointers to DEVICE memory
 *     unless the name ends in `_host`.
 *   - `stream` is a cudaStream_t passed as void*; NULL is the legacy default
 *     stream, which is what cux's GpuTimer records on
 *     (0-frameworks/cux/include/util/timer.h:26-27).
 *   - matrices are row-major with explicit leading dimensions, as cux passes
 *     them (0-frameworks/cux/src/operator/gemm.cpp:458-463: lda=K, ldb=N, ldc=N).
 *   - kernels never fall back to the CPU.  A missing device, a non-sm_100
 *     device or an unsupported shape is an error.
 *   - concurrency: a context owns one set of workspaces (block partials and
 *     tickets, scan tile states, split-K scratch, broadcast flags), so the
 *     kernel entry points of ONE context must be issued in stream order -- one
 *     stream at a time, or streams the caller has ordered with events.  This
 *     is cux's own model (one OpAssistor, default stream,
 *     0-frameworks/cux/include/executor.h:74-103).  Work that is meant to
 *     overlap on several streams uses one context per stream; contexts are
 *     cheap (a few KiB of device memory until a workspace is first needed).
 *     Calls may come from any host thread, but not from two at once for the
 *     same context, except the *_host entry points, which serialise
 *     themselves.  An entry point that takes a context makes that context's
 *     device the calling thread's current device and leaves it so.
 *   - CUDA graphs: the single-GPU kernel entry points (reductions, scans
 *     including cuxb_scan_i32_noalloc, GEMMs, linears, convert, fill) may be
 *     recorded by stream capture and replayed: no protocol state is baked
 *     into a launch's arguments (the scan counts its epoch on the device, the
 *     reduction and split-K tickets reset themselves).  Call once outside the
 *     capture first, so that workspaces exist and nothing allocates while
 *     capturing.  A recorded launch holds the ADDRESS of the context's workspace.  When a later call on the
 *     same context needs a larger one (a longer scan, a split-K GEMM with more tiles) the old workspace is
 *     kept allocated until cuxb_ctx_destroy, so graphs recorded before stay replayable
 *     (tests/test_gpu_graphs.py); a replay and a direct call must still not overlap on the GPU when the
 *     graph was recorded with the CURRENT workspace.  The multi-GPU reductions and scan (cuxb_mg_dot_f32 ...
 *     cuxb_mg_scan_i32) can be captured too: their exchange epoch is counted
 *     in the exchange buffer.  NOT capturable: the *_host entry points (they
 *     synchronise) and the fused-broadcast GEMMs (cuxb_mg_*gemm*,
 *     cuxb_mg_linear_*: their flag targets are counted on the host, per call).
 */
#ifndef CUXB_H_
#define CUXB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUXB_ERR_INVALID 100001   /* bad argument (null pointer, negative size, misalignment) */
#define CUXB_ERR_UNSUPPORTED 100002 /* shape / device not supported by the sm_100a kernels */
#define CUXB_ERR_TOO_LARGE 100003 /* size beyond a documented maximum */

/* Element type tags, numerically equal to cux::TypeFlag
 * (0-frameworks/cux/include/util/util.h:22-28) plus BFLOAT16. */
#define CUXB_FLOAT32 0
#define CUXB_INT32 1
#define CUXB_FLOAT16 2
#define CUXB_INT8 3
#define CUXB_BFLOAT16 4

typedef struct cuxb_ctx cuxb_ctx;

/* ---- context: per-device workspace (block partials, scan tile states, TMA
 * descriptors, staging buffers).  Replaces what cux::OpAssistor owns
 * (0-frameworks/cux/include/operator/op_assistor.h:100-143: device, cuBLAS handle). */
int cuxb_ctx_create(int device, cuxb_ctx **out);
int cuxb_ctx_destroy(cuxb_ctx *ctx);
int cuxb_ctx_device(const cuxb_ctx *ctx);
int cuxb_ctx_sm_count(const cuxb_ctx *ctx);
const char *cuxb_last_error_string(void);
const char *cuxb_version(void);
/* Number of kernel launches this library has issued in this process (all contexts). */
long long cuxb_launch_count(void);

/* ---- memory, replacing the cudaMalloc/cudaMemcpy/cudaMemset/cudaFree calls inside
 * cux::Array4D (0-frameworks/cux/include/array.h:84-114, src/array.cpp:29-44). */
int cuxb_malloc(void **dptr, size_t bytes);
int cuxb_free(void *dptr);
int cuxb_malloc_host(void **hptr, size_t bytes); /* pinned */
int cuxb_free_host(void *hptr);
int cuxb_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream);
int cuxb_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream);
int cuxb_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream);
int cuxb_memset(void *dst, int value, size_t bytes, void *stream);
int cuxb_set_device(int device);
int cuxb_device_count(int *count);
int cuxb_device_sync(void);
int cuxb_device_name(int device, char *buf, size_t buflen);
/* The cudaDeviceProp fields cux reads (0-frameworks/cux/include/executor.h:45-66,
 * util/launch_config.h:56-58).  Any out pointer may be NULL. */
int cuxb_device_info(int device, int *sm_count, int *cc_major, int *cc_minor, int *max_threads_per_sm,
                     int *max_threads_per_block, int *warp_size, size_t *global_mem_bytes,
                     size_t *shared_mem_per_block);

/* ---- streams and events, replacing cux::GpuTimer
 * (0-frameworks/cux/include/util/timer.h:16-38). */
int cuxb_stream_create(void **stream);
int cuxb_stream_destroy(void *stream);
int cuxb_stream_sync(void *stream);
int cuxb_event_create(void **event);
int cuxb_event_destroy(void *event);
int cuxb_event_record(void *event, void *stream);
int cuxb_event_sync(void *event);
int cuxb_event_elapsed_ms(void *start, void *stop, float *ms);

/* ---- reductions (HBM-bound).  Results OVERWRITE *res (device pointer); the
 * summation order is fixed, so results are run-to-run identical.
 *
 * cuxb_dot_f32 replaces the body of DotGpuKernelIF::func
 *   (0-frameworks/cux/include/operator/kernel_interface.h:30-35; kernels
 *   DotDeviceV0/V1/V2 + cublasSdot, src/operator/dot_product.cu:9-112,203-204).
 * cuxb_sum_f32 replaces ReduceKernelFp32V1/V2 (cuda/reduce_fp32.cu:33-116).
 * cuxb_nrm2_f32 replaces Nrm2GpuKernelIF::func (kernel_interface.h:43-48;
 *   Nrm2DeviceV0 + Sqrt + cublasSnrm2, src/operator/nrm2.cu:7-23,58-59); n<=0 -> 0,
 *   n==1 -> |x0| as Nrm2HostV0 does (src/operator/nrm2.cpp:8-9). */
int cuxb_dot_f32(cuxb_ctx *ctx, void *stream, long long n, const float *a, const float *b, float *res);
int cuxb_sum_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res);
int cuxb_nrm2_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res);
/* int32 sum with two's-complement wrap-around (phase 1 of the multi-GPU scan). */
int cuxb_sum_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *x, int32_t *res);

/* ---- inclusive prefix scan, out[i] = seed + in[0] + ... + in[i], int32 with
 * wrap-around.  Replaces VectorScanNormal/SIMDv1/SIMDv2
 * (simd/x86/vector_scan.cpp:21-79) in the int element type of
 * tbb/base_parallel_scan.cpp:62-65.  in == out (in place) is allowed.  The kernel is launched
 * cooperatively (its CTAs wait on each other's tile totals), so scans issued on different streams of
 * one GPU run one after the other. */
int cuxb_scan_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *in, int32_t *out, int32_t seed);
/* fp32 variant (same recurrence; the order of additions differs from the serial
 * reference). */
int cuxb_scan_f32(cuxb_ctx *ctx, void *stream, long long n, const float *in, float *out, float seed);
/* The same int32 scan for callers that must not allocate device memory (a plug-in inside an
 * application that owns all of the GPU's memory, a fixed-budget service, a loop that cannot take the
 * device-wide stall of cudaMalloc).  It has no counterpart in the reference.  It takes no cuxb_ctx --
 * creating one allocates -- and runs on the current device, keeping its tile-state scratch in a static
 * array inside the library (about 128 KiB per GPU, held from library load to process exit).
 * Limit: n <= 67,108,864 (2^26) elements; larger n returns CUXB_ERR_TOO_LARGE.  The first call on each
 * device resolves the array's device address; make it before any stream capture starts -- and make it
 * EARLY: CUDA loads a library's device code lazily by default, so it is this first call (not the loading
 * of libcuxb.so) that brings the kernel and the array onto the GPU, and that one-off step does need a few
 * hundred KiB of device memory.  A host application that is about to hand all remaining memory to its own
 * pool should therefore make one small warm-up call per device first (or run with
 * CUDA_MODULE_LOADING=EAGER); every later call leaves the amount of free device memory unchanged
 * (tests/test_zz_no_alloc.py checks exactly that).  It runs on the CURRENT device: the other entry points
 * switch the calling thread's current device to their context's and do not switch back, so after calls on
 * several GPUs set the device (cuxb_set_device) before calling this one.  Calls on the same GPU share the
 * scratch, so two calls whose kernels may run at the same time race. */
int cuxb_scan_i32_noalloc(void *stream, long long n, const int32_t *in, int32_t *out, int32_t seed);

/* ---- multi-GPU (one process or thread per GPU, all on one NVLink/NVSwitch box).  The scalar
 * exchange is FUSED into the reduction kernel: every rank's finishing CTA stores its partial into every
 * rank's exchange buffer through peer-mapped pointers and adds the world's partials in rank order, so
 * no collective is launched and every rank gets the identical result.  Replaces the
 * MPI_Allreduce / MPI_Exscan of the reference's MPI tutorials
 * (/root/reference/mpi/base_reduce_alltoall_scan.cpp:62,101) on GPU partials.
 *   peer_bufs[r] = rank r's exchange buffer as seen from THIS process (cudaIpc / symmetric memory /
 *   plain pointers with peer access in a single process), CUXB_MG_EXCHANGE_BYTES each, zero-filled
 *   once before the first call.  Every rank must make the same sequence of cuxb_mg_* calls.
 *   world <= 8.  world == 1 degenerates to the single-GPU kernel.  The kernels of different ranks wait
 *   for each other, so they must be able to run at the same time: one GPU per rank, and -- when one
 *   process drives several GPUs -- every kernel launched once per device beforehand (its first launch
 *   loads the module, which can wait for the peer device to go idle).  A rank that never shows up makes
 *   the others fail with a launch error after 20 s rather than hang. */
#define CUXB_MG_EXCHANGE_BYTES 256
int cuxb_mg_dot_f32(cuxb_ctx *ctx, void *stream, long long n, const float *a, const float *b, float *res,
                    int world, int rank, void *const *peer_bufs);
int cuxb_mg_sum_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res, int world, int rank,
                    void *const *peer_bufs);
int cuxb_mg_nrm2_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res, int world, int rank,
                     void *const *peer_bufs);
/* *seed_out (device) = sum of the int32 totals of the lower ranks' slices (wrap-around). */
int cuxb_mg_exscan_total_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *x, int32_t *seed_out,
                             int world, int rank, void *const *peer_bufs);
/* Sharded inclusive scan: rank r holds the r-th contiguous slice; out continues the lower ranks' sum. */
int cuxb_mg_scan_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *in, int32_t *out, int world,
                     int rank, void *const *peer_bufs);
/* Row-sharded fp16 GEMM with the broadcast of B fused into the kernel (the MPI GEMM of
 * /root/reference/mpi/alg_matrix_multiply.cpp:52-74 -- B to every worker, row blocks of A and C -- without a
 * separate send): the root rank (is_root != 0) multiplies with b_root in place; any other rank passes the
 * root's B as mapped into its address space (b_root) and a local buffer of the same shape (b_local) that
 * the kernel fills over NVLink, column panel by column panel, while it is already multiplying the panels
 * that have arrived.  Same requirements as cuxb_hgemm_f16_f32acc.  The root must leave B untouched until
 * every rank's call has finished. */
int cuxb_mg_hgemm_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const void *A,
                             int lda, const void *b_root, void *b_local, int ldb, float beta, void *C, int ldc,
                             int c_dtype, int is_root);
/* The same for fp32 SGEMM (cfg 3): on a non-root rank the first CTAs of the launch to start (16..64 of
 * them: what the last wave of tiles would leave idle) copy B from the root into b_local, 128 rows of B per
 * readiness flag, and the other CTAs multiply, each waiting only for the k-panel it is about to load.  Needs cuxb_sgemm_f32's vectorised shape class (M, N multiples of 128,
 * K of 16, leading dimensions multiples of 4, 16-byte aligned pointers), else CUXB_ERR_UNSUPPORTED. */
int cuxb_mg_sgemm_f32(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const float *A, int lda,
                      const float *b_root, float *b_local, int ldb, float beta, float *C, int ldc, int is_root);
/* And for the int8 linear (cfg 5; ncclBroadcast of Bt in SURVEY 8e): Bt[N,K] is pulled n-block panel
 * (256 rows of Bt) by n-block panel.  Same requirements as cuxb_linear_s8_s32. */
int cuxb_mg_linear_s8_s32(cuxb_ctx *ctx, void *stream, int M, int N, int K, const int8_t *A,
                          const int8_t *bt_root, int8_t *bt_local, int32_t *C, int is_root);
/* The 16-bit linears (LinearNormalBf16, linear.cpp:137-151, and its fp16 twin) the same way. */
int cuxb_mg_linear_bf16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A, const void *bt_root,
                               void *bt_local, void *C, int c_dtype, int is_root);
int cuxb_mg_linear_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A, const void *bt_root,
                              void *bt_local, void *C, int c_dtype, int is_root);
/* Chained broadcast.  By default every non-root rank pulls B straight from the root, whose NVLink egress
 * then carries B (world - 1) times -- fine at 2 GPUs, the bottleneck at 8.  After cuxb_mg_set_chain the
 * b_root / bt_root argument of the calls above means "the copy of B on the rank I pull from" (rank r pulls
 * from rank r - 1's b_local, rank 1 from the root), so that every GPU sends B once.  The flag buffers are
 * peer-mapped device memory of CUXB_MG_CHAIN_FLAGS_BYTES, zeroed once before the first call:
 * flags_local = this rank's; flags_up = that of the rank pulled from (NULL when that rank is the root: its
 * B is complete); flags_down = that of the rank that pulls from this one (NULL for the last rank).  Every
 * rank WAITS only on its own flags and WRITES its neighbours': when a panel is complete here the kernel
 * stores the call's number into the panel's word of flags_down, and when it has read all of a call's B it
 * says so in flags_up -- so a wait is a local poll and never an NVLink round trip.  A panel is pulled once
 * this rank's flags say it has landed upstream, and a call does not overwrite b_local before the downstream
 * rank has read all of the previous call's.  All non-root ranks must make the same sequence of chained
 * calls.  Host-side state only: takes effect for later calls; NULL, NULL, NULL switches back to pulling
 * from the root. */
#define CUXB_MG_CHAIN_FLAGS_BYTES 16640
int cuxb_mg_set_chain(cuxb_ctx *ctx, void *flags_local, void *flags_up, void *flags_down);
/* NVSwitch multicast all-gather (3..8 ranks on a box whose GPUs sit behind an NVSwitch).  The chain above
 * needs world - 1 hops before the last rank sees the first panel, and every hop moves all of B.  Here the
 * root copies nothing and every other rank owns every (world-1)-th 16 KiB box of B (SGEMM: every
 * (world-1)-th row): its kernel pulls THOSE boxes from the root -- so the root's NVLink egress carries B
 * once, spread over all receivers -- and stores them to a MULTICAST address, which the switch replicates
 * into every rank's copy of B.  Every panel is filled by all receivers at once at the full ingress rate of
 * a GPU; each rank announces its share of a panel group with one multicast store into the flags, and a
 * rank starts loading a panel once all world - 1 announcements for it are in its own instance of the flags.
 * After this call, for the cuxb_mg_*gemm / cuxb_mg_linear_* calls of this context (non-root ranks; the root
 * needs no call -- it multiplies with its own B as before):
 *   b_root / bt_root   = the root's B as mapped into this process (what this rank's share is pulled from)
 *   b_local / bt_local = this rank's copy (what its GEMM reads)
 *   b_mc               = the multicast address of the ranks' copies (a multicast object bound to one
 *                        allocation per GPU: cuMulticastCreate / cuMulticastBindMem, torch symmetric memory's
 *                        multicast_ptr, or cuxb_mg_create's buffers); the copies must have b_local's layout
 *   flags_local        = this rank's instance of the flags (CUXB_MG_MCAST_FLAGS_BYTES, zeroed once before
 *                        the first call); flags_mc = the multicast address of the flags
 * A rank does not overwrite the copies before every receiver has reported (through the flags) that it has
 * read all of the previous call's B.  Every non-root rank must make the same sequence of calls.  The root
 * must leave B untouched until every rank's call has finished.  flags_local == NULL switches the mode off.
 * Host-side state only; a chain set with cuxb_mg_set_chain is ignored while this mode is on.  Has no
 * counterpart in the reference; the transfer it replaces is the MPI_Send of B to every worker,
 * mpi/alg_matrix_multiply.cpp:57-62. */
#define CUXB_MG_MCAST_FLAGS_BYTES 16640
int cuxb_mg_set_multicast(cuxb_ctx *ctx, int world, int rank, void *flags_local, void *flags_mc, void *b_mc);
int cuxb_enable_peer_access(int device, int peer_device);

/* ---- the multi-GPU driver: ONE host thread driving g GPUs of one NVLink box.  cux::Executor::Initialize(dev)
 * (0-frameworks/cux/src/executor.cpp:10-22) generalised to g devices, for the master / worker flow of
 * mpi/alg_matrix_multiply.cpp:52-80 (row blocks of A and C per worker, one B, results gathered and checked)
 * without MPI.  cuxb_mg_create(ndev, devices, &mg) -- devices == NULL means 0..ndev-1 -- creates one cuxb_ctx and one stream
 * per device, enables peer access between all pairs, allocates the exchange buffers, the broadcast flags
 * and (on demand) every rank's copy of B, loads the kernels on every device before any of them can wait on
 * a peer, and routes the broadcast (2 GPUs: pull from the root; more: rank r pulls from rank r-1).
 * The cuxb_mg_all_* calls take ARRAYS indexed by rank -- sizes and pointers to memory on that rank's GPU --
 * and issue one asynchronous launch per device on that device's stream; cuxb_mg_sync waits for all.  Rank r
 * holds the r-th contiguous slice of a vector, or M[r] rows of A and C; B (Bt) lives on rank 0 only and is
 * delivered inside the GEMM kernels.  Reductions leave the same scalar at every res[r].  Not thread-safe:
 * one host thread per cuxb_mg.  The caller's current device is changed by these calls. */
typedef struct cuxb_mg cuxb_mg;
int cuxb_mg_create(int ndev, const int *devices, cuxb_mg **out);
int cuxb_mg_init(int ndev, cuxb_mg **out);      /* = cuxb_mg_create(ndev, NULL, out) */
int cuxb_mg_destroy(cuxb_mg *mg);
int cuxb_mg_world(const cuxb_mg *mg);
int cuxb_mg_device(const cuxb_mg *mg, int rank);
cuxb_ctx *cuxb_mg_ctx(cuxb_mg *mg, int rank);   /* owned by mg */
void *cuxb_mg_stream(cuxb_mg *mg, int rank);    /* the stream the cuxb_mg_all_* launches of this rank go to */
int cuxb_mg_sync(cuxb_mg *mg);
int cuxb_mg_all_dot_f32(cuxb_mg *mg, const long long *n, const float *const *a, const float *const *b, float *const *res);
int cuxb_mg_all_sum_f32(cuxb_mg *mg, const long long *n, const float *const *x, float *const *res);
int cuxb_mg_all_nrm2_f32(cuxb_mg *mg, const long long *n, const float *const *x, float *const *res);
int cuxb_mg_all_scan_i32(cuxb_mg *mg, const long long *n, const int32_t *const *in, int32_t *const *out);
int cuxb_mg_all_sgemm_f32(cuxb_mg *mg, const int *M, int N, int K, float alpha, const float *const *A, int lda,
                          const float *B_root, int ldb, float beta, float *const *C, int ldc);
int cuxb_mg_all_hgemm_f16_f32acc(cuxb_mg *mg, const int *M, int N, int K, float alpha, const void *const *A, int lda,
                                 const void *B_root, int ldb, float beta, void *const *C, int ldc, int c_dtype);
int cuxb_mg_all_linear_s8_s32(cuxb_mg *mg, const int *M, int N, int K, const int8_t *const *A, const int8_t *Bt_root,
                              int32_t *const *C);

/* ---- one PROCESS per GPU (torchrun, MPI): the peer pointers the cuxb_mg_* kernel entry points take are made
 * by exporting a cuxb_malloc'ed buffer as a 64-byte handle, passing the handle to the other processes over
 * whatever channel the launcher offers (an all-gather of bytes), and importing it there.  Thin wrappers of
 * cudaIpcGetMemHandle / cudaIpcOpenMemHandle / cudaIpcCloseMemHandle; the buffer must be a whole cuxb_malloc
 * allocation.  The counterpart of the MPI_Send / MPI_Recv plumbing of mpi/alg_matrix_multiply.cpp:57-72. */
#define CUXB_IPC_HANDLE_BYTES 64
int cuxb_ipc_export(void *dptr, void *handle_out);
int cuxb_ipc_import(const void *handle, void **dptr_out);
int cuxb_ipc_close(void *dptr);

/* ---- GEMM: C = alpha * A[M,K] * B[K,N] + beta * C, row-major.
 *
 * cuxb_sgemm_f32 replaces GemmGpuKernelIF::func for TypeFlag::FLOAT32
 *   (kernel_interface.h:61-72; GemmDeviceV0/V1/V2 + cublasSgemm,
 *   src/operator/gemm.cu:17-130,237-238; cuda/gemm_fp32.cu:102-528).  CUDA cores (FFMA).
 * cuxb_hgemm_f16_f32acc replaces the FLOAT16 entry (cublasSgemmEx,
 *   src/operator/gemm.cu:266-269) and cuda/gemm_fp16_wmma.cu:123-700: fp16 A,B,
 *   fp32 accumulate on tcgen05/TMEM, alpha/beta in fp32.  c_dtype is CUXB_FLOAT16
 *   (cux convention) or CUXB_FLOAT32.  Requires K % 8 == 0, N % 8 == 0, lda/ldb/ldc
 *   multiples of 8 elements and 16-byte aligned base pointers (TMA). */
int cuxb_sgemm_f32(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const float *A,
                   int lda, const float *B, int ldb, float beta, float *C, int ldc);
int cuxb_hgemm_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha,
                          const void *A, int lda, const void *B, int ldb, float beta, void *C,
                          int ldc, int c_dtype);
/* The same with bf16 operands (c_dtype CUXB_BFLOAT16 or CUXB_FLOAT32); numerically the contraction of
 * LinearNormalBf16 (simd/x86/linear.cpp:137-151) in cux's B[K,N] layout. */
int cuxb_gemm_bf16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha,
                          const void *A, int lda, const void *B, int ldb, float beta, void *C,
                          int ldc, int c_dtype);

/* ---- linear: C[M,N] = A[M,K] * Bt[N,K]^T (simd/x86/linear.cpp:2).
 * cuxb_linear_s8_s32 replaces LinearNormalInt8 / LinearAvx512Int8
 *   (simd/x86/linear.cpp:199-237): signed int8 x signed int8 -> int32, bit-exact.
 * cuxb_linear_bf16_f32acc replaces LinearNormalBf16 / LinearAvx512Bf16
 *   (simd/x86/linear.cpp:137-182); c_dtype CUXB_BFLOAT16 or CUXB_FLOAT32.
 * cuxb_linear_f16_f32acc: the same contraction for fp16 inputs
 *   (LinearAvx512AhBfCh, linear.cpp:186-196); c_dtype CUXB_FLOAT16 or CUXB_FLOAT32.
 * Requires K % 16 == 0 bytes-aligned rows (K % 16 for int8, K % 8 for 16-bit). */
int cuxb_linear_s8_s32(cuxb_ctx *ctx, void *stream, int M, int N, int K, const int8_t *A,
                       const int8_t *Bt, int32_t *C);
int cuxb_linear_bf16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A,
                            const void *Bt, void *C, int c_dtype);
int cuxb_linear_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A,
                           const void *Bt, void *C, int c_dtype);

/* ---- element conversion on the device, replacing the scalar host loop
 * Array4D::PrecsCpuCvt (0-frameworks/cux/include/array.h:124-144) for fp32 -> fp16/bf16
 * (round to nearest even) and back. */
int cuxb_convert(cuxb_ctx *ctx, void *stream, long long n, const void *src, int src_dtype, void *dst,
                 int dst_dtype);
/* Random fill on the device with the distribution of DataFiller::RandomFill
 * (0-frameworks/cux/include/util/data_filler.h:41-60): integers in [min_value, max_value], or, with
 * decimal_pose > 0, [min_value, max_value) plus that many random decimal digits.  Element i is a pure
 * function of (seed, i).  Backs Array4D::Fill(..., ON_DEVICE), which the reference does not implement
 * (src/array.cpp:47-49). */
int cuxb_fill_random(cuxb_ctx *ctx, void *stream, long long n, void *dst, int dtype, int min_value, int max_value,
                     int decimal_pose, unsigned long long seed);

/* ---- host-buffer entry points: what a cux plugin that is handed HOST arrays
 * calls.  Inputs are copied host->device, the kernel runs, the result is
 * copied back and the call returns after the context's streams have drained.
 * The GEMMs and the scan, which send back as much as they take in, run as a
 * pipeline once more than 32 MiB is moved: row (element) chunks, chunk i+1
 * uploading while chunk i is computed and chunk i-1 downloads, on three streams
 * (16384^3 fp16: 33.8 -> 22.5 ms, the upload of A and B alone is 19.5 ms).  The
 * result is bit-identical to one kernel over the whole problem.  Pinned host
 * memory (cuxb_malloc_host) makes the copies asynchronous DMA; pageable memory
 * works but is slower.  Device staging buffers are owned by ctx and reused. */
int cuxb_dot_f32_host(cuxb_ctx *ctx, long long n, const float *a, const float *b, float *res);
int cuxb_sum_f32_host(cuxb_ctx *ctx, long long n, const float *x, float *res);
int cuxb_nrm2_f32_host(cuxb_ctx *ctx, long long n, const float *x, float *res);
int cuxb_scan_i32_host(cuxb_ctx *ctx, long long n, const int32_t *in, int32_t *out, int32_t seed);
int cuxb_sgemm_f32_host(cuxb_ctx *ctx, int M, int N, int K, float alpha, const float *A, int lda,
                        const float *B, int ldb, float beta, float *C, int ldc);
int cuxb_hgemm_f16_f32acc_host(cuxb_ctx *ctx, int M, int N, int K, float alpha, const void *A,
                               int lda, const void *B, int ldb, float beta, void *C, int ldc,
                               int c_dtype);
int cuxb_linear_s8_s32_host(cuxb_ctx *ctx, int M, int N, int K, const int8_t *A, const int8_t *Bt,
                            int32_t *C);

/* ---- occupancy hook, replacing Operator::QueryPotentialOccupancy on
 * KernelIF::config_kernel (0-frameworks/cux/src/operator/operator.cpp:11-23).
 * kernel: "dot_f32", "sum_f32", "nrm2_f32", "scan_i32", "sgemm_f32", "hgemm_f16". */
int cuxb_kernel_occupancy(cuxb_ctx *ctx, const char *kernel, int *active_blocks_per_sm,
                          double *occupancy, int *threads_per_block, int *dynamic_smem_bytes);

/* Scratch memory (device memory beyond the inputs and outputs) the context would hold after a call of this
 * size, for callers that plan a memory budget: what cux leaves to cudaMalloc inside Array4D / the cuBLAS handle
 * (0-frameworks/cux/include/operator/op_assistor.h:100-143).  kernel: "dot_f32", "sum_f32", "nrm2_f32", "sum_i32"
 * (m_or_n = n; 0: their 16 KiB of partials belong to every context), "scan_i32", "scan_f32" (m_or_n = n),
 * "sgemm_f32", "hgemm_f16", "gemm_bf16", "linear_s8", "linear_bf16", "linear_f16" (m_or_n = M, N, K).  The
 * *_host entry points additionally stage their operands on the device.  Pure host arithmetic; ctx may be NULL. */
int cuxb_workspace_bytes(const cuxb_ctx *ctx, const char *kernel, long long m_or_n, long long N, long long K, size_t *bytes);

/* ---- measured ceilings of this GPU, for a harness that prints roofline fractions: the counterpart of the
 * reference's evaluation helpers (pocket-ai/eval/cuda/bandwidth_eval.hpp:257-300, timed device-to-device
 * copies; pocket-ai/eval/cuda/max_flops.cuh:11-94, timed fma chains).  Both calls allocate their own
 * buffers, run on the context's private stream, block until done and return the best of `reps` launches.
 * cuxb_probe_hbm: which = 0 read-only, 1 write-only, 2 copy (a copy of `bytes` counts as 2 x bytes moved);
 * at least 64 MiB, use several times the 126 MB L2.  cuxb_probe_fp32: packed = 0 scalar FFMA, 1 packed FFMA2. */
int cuxb_probe_hbm(cuxb_ctx *ctx, int which, size_t bytes, int reps, double *gb_per_s);
int cuxb_probe_fp32(cuxb_ctx *ctx, int packed, int reps, double *tflops);
/* Tensor-pipe ceiling (the tcgen05 counterpart of max_flops.cuh:11-94): CTA pairs issue tcgen05.mma
 * 256 x 256 x 16 (kind CUXB_FLOAT16 / CUXB_BFLOAT16, TFLOP/s) or 256 x 256 x 32 (CUXB_INT8, TOP/s) back to
 * back on operands resident in shared memory.  data = 0: all-zero operands, 1: pseudo-random operands (the
 * data sets the power draw and, on a power-capped board, the clock).  *best_tflops = fastest single launch
 * (burst); *mean_tflops = `reps` launches of about 4 ms each timed back to back (sustained when reps covers a
 * second or more). */
int cuxb_probe_tensor(cuxb_ctx *ctx, int kind, int data, int reps, double *best_tflops, double *mean_tflops);
/* One entry of the NVLink pair matrix (pocket-ai/eval/cuda/device_query.hpp:134-186 lists peer access and
 * topology; this measures it): GB/s between `peer_device`'s memory and this context's GPU.  mode 0 = a
 * kernel on this GPU loads from the peer and stores locally; 1 = it loads locally and stores to the peer;
 * 2 = cudaMemcpyPeerAsync peer -> here (copy engines); 3 = one thread per SM pulls with bulk async copies
 * peer -> shared memory -> here through `depth` (1..12) 16 KiB buffers, the way the tensor GEMM's broadcast
 * warp does.  Single process with both devices visible; `bytes` (multiple of 16 KiB, >= 16 MiB) is
 * allocated on both GPUs for the duration of the call and peer access is enabled both ways. */
int cuxb_probe_nvlink(cuxb_ctx *ctx, int peer_device, int mode, int depth, size_t bytes, int reps, double *gb_per_s);

#ifdef __cplusplus
}
#endif
#endif /* CUXB_H_ */
