// host_api.cu -- host-buffer entry points and the occupancy hook.
//
// The *_host functions are what a cux plugin that is handed HOST arrays calls
// (cux itself stages through Array4D::GetGpuData / GetCpuData with synchronous
// cudaMemcpy on pageable memory, 0-frameworks/cux/include/array.h:84-114).  The
// reductions copy their inputs host->device on the context's private stream, launch
// the kernel there, copy the scalar back and wait.  The GEMMs and the scan move as
// much data out as in, so they are pipelines: the problem is cut into row (element)
// chunks, chunk i+1 is uploaded on one stream while chunk i is computed on a second
// and chunk i-1 is downloaded on a third -- PCIe carries both directions at once and
// the kernels hide behind the copies.  Staging buffers live in the cuxb_ctx and are
// reused between calls; every call returns with all three streams idle.
#include <string.h>

#include "common.cuh"

namespace cuxb {
int scan_i32_device_seed(cuxb_ctx *ctx, cudaStream_t st, long long n, const int32_t *in, int32_t *out, int32_t seed,
                         const int32_t *seed_ptr);
const void *reduce_kernel_symbol(const char *name);
const void *scan_kernel_symbol();
int scan_smem_bytes();
const void *sgemm_kernel_symbol();
const void *umma_kernel_symbol();
int umma_smem_bytes();
}  // namespace cuxb

using namespace cuxb;

#define CUXB_TRY(expr)      \
  do {                      \
    int rc__ = (expr);      \
    if (rc__) return rc__;  \
  } while (0)

static int h2d(cuxb_ctx *ctx, int slot, const void *h, size_t bytes, void **d) {
  CUXB_TRY(stage_reserve(ctx, slot, bytes ? bytes : 16, d));
  if (bytes) CUXB_CUDA(cudaMemcpyAsync(*d, h, bytes, cudaMemcpyHostToDevice, ctx->stage_stream));
  return 0;
}
static int d2h_sync(cuxb_ctx *ctx, void *h, const void *d, size_t bytes) {
  if (bytes) CUXB_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stage_stream));
  CUXB_CUDA(cudaStreamSynchronize(ctx->stage_stream));
  return 0;
}

// ---- pipelines
static int pipe_init(cuxb_ctx *ctx) {
  if (ctx->host_pipe_ready) return 0;
  CUXB_CUDA(cudaStreamCreateWithFlags(&ctx->up_stream, cudaStreamNonBlocking));
  CUXB_CUDA(cudaStreamCreateWithFlags(&ctx->down_stream, cudaStreamNonBlocking));
  for (int i = 0; i <= kHostMaxChunks; i++) CUXB_CUDA(cudaEventCreateWithFlags(&ctx->ev_up[i], cudaEventDisableTiming));
  for (int i = 0; i < kHostMaxChunks; i++) CUXB_CUDA(cudaEventCreateWithFlags(&ctx->ev_k[i], cudaEventDisableTiming));
  ctx->host_pipe_ready = true;
  return 0;
}
// all three streams idle again, whatever happened (returns rc unchanged when it is already an error)
static int pipe_drain(cuxb_ctx *ctx, int rc) {
  cudaError_t e1 = cudaStreamSynchronize(ctx->up_stream);
  cudaError_t e2 = cudaStreamSynchronize(ctx->stage_stream);
  cudaError_t e3 = cudaStreamSynchronize(ctx->down_stream);
  if (rc) return rc;
  cudaError_t e = e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3);
  if (e != cudaSuccess) {
    set_error("host pipeline: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}
// Rows per chunk, a multiple of `gran` (the kernel's tile height).  One chunk when the whole transfer is
// small enough that latency, not bandwidth, is what it costs.  Otherwise between 2 and kHostMaxChunks
// chunks, picked by a small cost model of the three-stream pipeline:
//   upload of B and of the first chunk of A  +  max(kernels, remaining uploads, downloads but the last)
//   +  download of the last chunk  +  a fixed cost per chunk,
// where a chunk of q tile-rows has q * tiles_per_row tiles and its kernel takes ceil(that / slots) waves
// of the GPU's `slots` CTA (cluster) slots -- cutting a kernel-bound problem into equal eighths can cost
// more in partial waves than the overlap buys.  8192^3 fp32 (64 tile-rows of 64 tiles, 296 slots, kernel
// 18 ms vs 15 ms of copies): chunks of 23, 23, 18 tile-rows, 14 waves like the uncut problem.  16384^3
// fp16 and 8192^3 int8 are copy-bound and get the finest cut.
struct PipeCost {
  double up_b_bytes, up_row_bytes, down_row_bytes;  // B once; A (+C) and C per row
  double wave_seconds;                              // one full wave of tiles at the kernel's usual rate
};
static int pipe_chunk_rows(long long rows, int gran, int tiles_per_row, int slots, const PipeCost &c) {
  const double kPcie = 50e9, kPerChunk = 30e-6;
  const long long blocks = (rows + gran - 1) / gran;
  const double moved = c.up_b_bytes + (c.up_row_bytes + c.down_row_bytes) * (double)rows;
  if (moved < (double)(32ll << 20) || blocks < 4) return (int)rows;
  long long best_q = blocks;
  double best_t = -1;
  const long long q_min = (blocks + kHostMaxChunks - 1) / kHostMaxChunks;
  for (long long q = q_min; q <= blocks; q++) {
    const long long chunks = (blocks + q - 1) / q, last = blocks - (chunks - 1) * q;
    const long long waves = (chunks - 1) * ((q * tiles_per_row + slots - 1) / slots) + (last * tiles_per_row + slots - 1) / slots;
    const double first_rows = (double)(chunks > 1 ? q : last) * gran, last_rows = (double)last * gran;
    const double kernels = (double)waves * c.wave_seconds;
    const double ups = c.up_row_bytes * ((double)rows - first_rows) / kPcie;
    const double downs = c.down_row_bytes * ((double)rows - last_rows) / kPcie;
    double mid = kernels > ups ? kernels : ups;
    if (downs > mid) mid = downs;
    const double t = (c.up_b_bytes + c.up_row_bytes * first_rows) / kPcie + mid + c.down_row_bytes * last_rows / kPcie +
                     kPerChunk * (double)chunks;
    if (best_t < 0 || t < best_t) best_t = t, best_q = q;
    if (q == blocks) break;
  }
  return (int)(best_q * gran < rows ? best_q * gran : rows);
}
static int up_2d(cuxb_ctx *ctx, void *d, const void *h, int rows, int cols, int ld, int ebytes) {
  if (rows > 0)
    CUXB_CUDA(cudaMemcpy2DAsync(d, (size_t)cols * ebytes, h, (size_t)ld * ebytes, (size_t)cols * ebytes, rows,
                                cudaMemcpyHostToDevice, ctx->up_stream));
  return 0;
}
static int down_2d(cuxb_ctx *ctx, void *h, const void *d, int rows, int cols, int ld, int ebytes) {
  if (rows > 0)
    CUXB_CUDA(cudaMemcpy2DAsync(h, (size_t)ld * ebytes, d, (size_t)cols * ebytes, (size_t)cols * ebytes, rows,
                                cudaMemcpyDeviceToHost, ctx->down_stream));
  return 0;
}

// C[M,N] = f(A[M,K], B) by row chunks.  `b_rows x b_cols` is the shape of B as stored (B[K,N] or Bt[N,K]);
// `launch(r, rows, dA, dB, dC)` issues the kernel for rows [r, r+rows) on ctx->stage_stream.
template <typename Launch>
static int gemm_host_pipeline(cuxb_ctx *ctx, int M, int N, int K, const void *A, int lda, int abytes, const void *B,
                              int b_rows, int b_cols, int ldb, void *C, int ldc, int cbytes, bool load_c, int gran,
                              int tile_n, int slots, double ops_per_second, Launch launch) {
  CUXB_TRY(pipe_init(ctx));
  char *da, *db, *dc;
  CUXB_TRY(stage_reserve(ctx, 0, (size_t)M * K * abytes + 16, (void **)&da));
  CUXB_TRY(stage_reserve(ctx, 1, (size_t)b_rows * b_cols * abytes + 16, (void **)&db));
  CUXB_TRY(stage_reserve(ctx, 2, (size_t)M * N * cbytes + 16, (void **)&dc));
  PipeCost cost;
  cost.up_b_bytes = (double)b_rows * b_cols * abytes;
  cost.up_row_bytes = (double)K * abytes + (load_c ? (double)N * cbytes : 0.0);
  cost.down_row_bytes = (double)N * cbytes;
  cost.wave_seconds = 2.0 * gran * tile_n * (double)K * slots / ops_per_second;
  const int per = pipe_chunk_rows(M, gran, (N + tile_n - 1) / tile_n, slots, cost);
  int rc = 0;
  do {
    if ((rc = up_2d(ctx, db, B, b_rows, b_cols, ldb, abytes))) break;
    int ci = 0;
    for (int r = 0; r < M && !rc; r += per, ci++) {
      const int rows = M - r < per ? M - r : per;
      if ((rc = up_2d(ctx, da + (size_t)r * K * abytes, (const char *)A + (size_t)r * lda * abytes, rows, K, lda, abytes))) break;
      if (load_c && (rc = up_2d(ctx, dc + (size_t)r * N * cbytes, (const char *)C + (size_t)r * ldc * cbytes, rows, N, ldc, cbytes))) break;
      cudaError_t e = cudaEventRecord(ctx->ev_up[ci], ctx->up_stream);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stage_stream, ctx->ev_up[ci], 0);
      if (e != cudaSuccess) { set_error("host pipeline: %s", cudaGetErrorString(e)); rc = (int)e; break; }
      if ((rc = launch(r, rows, da + (size_t)r * K * abytes, db, dc + (size_t)r * N * cbytes))) break;
      e = cudaEventRecord(ctx->ev_k[ci], ctx->stage_stream);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->down_stream, ctx->ev_k[ci], 0);
      if (e != cudaSuccess) { set_error("host pipeline: %s", cudaGetErrorString(e)); rc = (int)e; break; }
      rc = down_2d(ctx, (char *)C + (size_t)r * ldc * cbytes, dc + (size_t)r * N * cbytes, rows, N, ldc, cbytes);
    }
  } while (0);
  return pipe_drain(ctx, rc);
}

CUXB_API int cuxb_dot_f32_host(cuxb_ctx *ctx, long long n, const float *a, const float *b, float *res) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(n >= 0 && res && (n == 0 || (a && b)), CUXB_ERR_INVALID, "cuxb_dot_f32_host: bad argument");
  std::lock_guard<std::mutex> g(*ctx->mu);
  void *da, *db, *dr;
  CUXB_TRY(h2d(ctx, 0, a, (size_t)n * 4, &da));
  CUXB_TRY(h2d(ctx, 1, b, (size_t)n * 4, &db));
  CUXB_TRY(stage_reserve(ctx, 2, 16, &dr));
  CUXB_TRY(cuxb_dot_f32(ctx, ctx->stage_stream, n, (const float *)da, (const float *)db, (float *)dr));
  return d2h_sync(ctx, res, dr, 4);
}

static int red1_host(cuxb_ctx *ctx, int which, long long n, const float *x, float *res) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(n >= 0 && res && (n == 0 || x), CUXB_ERR_INVALID, "cuxb reduction (host): bad argument");
  std::lock_guard<std::mutex> g(*ctx->mu);
  void *dx, *dr;
  CUXB_TRY(h2d(ctx, 0, x, (size_t)n * 4, &dx));
  CUXB_TRY(stage_reserve(ctx, 2, 16, &dr));
  if (which == 0)
    CUXB_TRY(cuxb_sum_f32(ctx, ctx->stage_stream, n, (const float *)dx, (float *)dr));
  else
    CUXB_TRY(cuxb_nrm2_f32(ctx, ctx->stage_stream, n, (const float *)dx, (float *)dr));
  return d2h_sync(ctx, res, dr, 4);
}
CUXB_API int cuxb_sum_f32_host(cuxb_ctx *ctx, long long n, const float *x, float *res) { return red1_host(ctx, 0, n, x, res); }
CUXB_API int cuxb_nrm2_f32_host(cuxb_ctx *ctx, long long n, const float *x, float *res) { return red1_host(ctx, 1, n, x, res); }

CUXB_API int cuxb_scan_i32_host(cuxb_ctx *ctx, long long n, const int32_t *in, int32_t *out, int32_t seed) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(n >= 0 && (n == 0 || (in && out)), CUXB_ERR_INVALID, "cuxb_scan_i32_host: bad argument");
  if (n == 0) return 0;
  std::lock_guard<std::mutex> g(*ctx->mu);
  CUXB_TRY(pipe_init(ctx));
  int32_t *d;
  CUXB_TRY(stage_reserve(ctx, 0, (size_t)n * 4, (void **)&d));
  // element chunks, scanned in place on the device; chunk i continues from the last output of chunk
  // i-1, which the kernel reads from device memory (same stream, so it is there)
  long long per = n;
  if (n * 8 >= (32ll << 20)) {
    per = (n + 7) / 8;
    per = (per + kScanTileElems - 1) / kScanTileElems * kScanTileElems;
  }
  int rc = 0, ci = 0;
  for (long long off = 0; off < n && !rc; off += per, ci++) {
    const long long len = n - off < per ? n - off : per;
    cudaError_t e = cudaMemcpyAsync(d + off, in + off, (size_t)len * 4, cudaMemcpyHostToDevice, ctx->up_stream);
    if (e == cudaSuccess) e = cudaEventRecord(ctx->ev_up[ci], ctx->up_stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stage_stream, ctx->ev_up[ci], 0);
    if (e != cudaSuccess) { set_error("cuxb_scan_i32_host: %s", cudaGetErrorString(e)); rc = (int)e; break; }
    if ((rc = scan_i32_device_seed(ctx, ctx->stage_stream, len, d + off, d + off, off == 0 ? seed : 0,
                                   off == 0 ? nullptr : d + off - 1)))
      break;
    e = cudaEventRecord(ctx->ev_k[ci], ctx->stage_stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->down_stream, ctx->ev_k[ci], 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out + off, d + off, (size_t)len * 4, cudaMemcpyDeviceToHost, ctx->down_stream);
    if (e != cudaSuccess) { set_error("cuxb_scan_i32_host: %s", cudaGetErrorString(e)); rc = (int)e; break; }
  }
  return pipe_drain(ctx, rc);
}

CUXB_API int cuxb_sgemm_f32_host(cuxb_ctx *ctx, int M, int N, int K, float alpha, const float *A, int lda,
                                 const float *B, int ldb, float beta, float *C, int ldc) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(M > 0 && N > 0 && K > 0 && A && B && C, CUXB_ERR_INVALID, "cuxb_sgemm_f32_host: bad argument");
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID, "cuxb_sgemm_f32_host: leading dimension too small");
  std::lock_guard<std::mutex> g(*ctx->mu);
  return gemm_host_pipeline(ctx, M, N, K, A, lda, 4, B, K, N, ldb, C, ldc, 4, beta != 0.f, 128, 128, 2 * ctx->sm_count, 60e12,
                            [&](int, int rows, const void *da, const void *db, void *dc) {
                              return cuxb_sgemm_f32(ctx, ctx->stage_stream, rows, N, K, alpha, (const float *)da, K,
                                                    (const float *)db, N, beta, (float *)dc, N);
                            });
}

CUXB_API int cuxb_hgemm_f16_f32acc_host(cuxb_ctx *ctx, int M, int N, int K, float alpha, const void *A, int lda,
                                        const void *B, int ldb, float beta, void *C, int ldc, int c_dtype) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(M > 0 && N > 0 && K > 0 && A && B && C, CUXB_ERR_INVALID, "cuxb_hgemm_f16_f32acc_host: bad argument");
  CUXB_REQUIRE(c_dtype == CUXB_FLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID, "bad c_dtype");
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID, "cuxb_hgemm_f16_f32acc_host: leading dimension too small");
  CUXB_REQUIRE(K % 8 == 0 && N % 8 == 0, CUXB_ERR_UNSUPPORTED, "cuxb_hgemm_f16_f32acc_host: K and N must be multiples of 8");
  const int cb = c_dtype == CUXB_FLOAT16 ? 2 : 4;
  std::lock_guard<std::mutex> g(*ctx->mu);
  return gemm_host_pipeline(ctx, M, N, K, A, lda, 2, B, K, N, ldb, C, ldc, cb, beta != 0.f, 256, 256, ctx->sm_count / 2, 1.6e15,
                            [&](int, int rows, const void *da, const void *db, void *dc) {
                              return cuxb_hgemm_f16_f32acc(ctx, ctx->stage_stream, rows, N, K, alpha, da, K, db, N, beta,
                                                           dc, N, c_dtype);
                            });
}

CUXB_API int cuxb_linear_s8_s32_host(cuxb_ctx *ctx, int M, int N, int K, const int8_t *A, const int8_t *Bt,
                                     int32_t *C) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(M > 0 && N > 0 && K > 0 && A && Bt && C, CUXB_ERR_INVALID, "cuxb_linear_s8_s32_host: bad argument");
  std::lock_guard<std::mutex> g(*ctx->mu);
  return gemm_host_pipeline(ctx, M, N, K, A, K, 1, Bt, N, K, K, C, N, 4, false, 256, 256, ctx->sm_count / 2, 3.2e15,
                            [&](int, int rows, const void *da, const void *db, void *dc) {
                              return cuxb_linear_s8_s32(ctx, ctx->stage_stream, rows, N, K, (const int8_t *)da,
                                                        (const int8_t *)db, (int32_t *)dc);
                            });
}

CUXB_API int cuxb_kernel_occupancy(cuxb_ctx *ctx, const char *kernel, int *active_blocks, double *occupancy,
                                   int *threads_per_block, int *dynamic_smem_bytes) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(kernel && active_blocks && occupancy, CUXB_ERR_INVALID, "cuxb_kernel_occupancy: null argument");
  const void *fn = nullptr;
  int threads = 0, smem = 0;
  if ((fn = reduce_kernel_symbol(kernel)) != nullptr) {
    threads = 512;
  } else if (!strcmp(kernel, "scan_i32")) {
    fn = scan_kernel_symbol(), threads = 512 + 32, smem = scan_smem_bytes();
    CUXB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  } else if (!strcmp(kernel, "sgemm_f32")) {
    fn = sgemm_kernel_symbol(), threads = 256;
  } else if (!strcmp(kernel, "hgemm_f16")) {
    fn = umma_kernel_symbol(), threads = 256, smem = umma_smem_bytes();
    CUXB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  CUXB_REQUIRE(fn != nullptr, CUXB_ERR_INVALID, "cuxb_kernel_occupancy: unknown kernel '%s'", kernel);
  // same arithmetic as LaunchConfig::QueryPotentialOccupancy
  // (0-frameworks/cux/include/util/launch_config.h:50-61)
  CUXB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(active_blocks, fn, threads, smem));
  cudaDeviceProp prop;
  CUXB_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
  const int active_warps = *active_blocks * threads / prop.warpSize;
  const int max_warps = prop.maxThreadsPerMultiProcessor / prop.warpSize;
  *occupancy = (double)active_warps / max_warps;
  if (threads_per_block) *threads_per_block = threads;
  if (dynamic_smem_bytes) *dynamic_smem_bytes = smem;
  return 0;
}

// Scratch a call would need from the context (beyond its inputs and outputs), so that a caller with a memory
// budget can plan: see cuxb.h.  Pure host arithmetic; ctx may be NULL (a 148-SM B200 is assumed).
namespace cuxb { int sgemm_pick_splits(int tiles, int K, int room); }
CUXB_API int cuxb_workspace_bytes(const cuxb_ctx *ctx, const char *kernel, long long m_or_n, long long N, long long K,
                                  size_t *bytes) {
  CUXB_REQUIRE(kernel != nullptr && bytes != nullptr, CUXB_ERR_INVALID, "cuxb_workspace_bytes: null argument");
  CUXB_REQUIRE(m_or_n >= 0 && N >= 0 && K >= 0, CUXB_ERR_INVALID, "cuxb_workspace_bytes: negative size");
  const int sms = ctx ? ctx->sm_count : 148;
  *bytes = 0;
  if (!strcmp(kernel, "dot_f32") || !strcmp(kernel, "sum_f32") || !strcmp(kernel, "nrm2_f32") || !strcmp(kernel, "sum_i32")) {
    *bytes = 0;   // block partials and the ticket are part of every context (kReduceMaxBlocks * 8 + 256 bytes)
  } else if (!strcmp(kernel, "scan_i32") || !strcmp(kernel, "scan_f32")) {
    // one 64-bit state word per tile of 8192 elements, capacity doubling from 4096 tiles, + 32 control words
    long long tiles = div_up(m_or_n > 0 ? m_or_n : 1, kScanTileElems), cap = 4096;
    CUXB_REQUIRE(tiles <= kScanMaxTiles, CUXB_ERR_TOO_LARGE, "cuxb_workspace_bytes: n beyond the scan's maximum");
    while (cap < tiles) cap *= 2;
    *bytes = (size_t)(cap + 32) * 8;
  } else if (!strcmp(kernel, "sgemm_f32")) {
    // split-K only: per-tile tickets + the parked accumulators of every split
    const long long tiles = div_up(m_or_n, 128) * div_up(N, 128);
    const int splits = tiles > 0 && tiles <= 1024 ? sgemm_pick_splits((int)tiles, (int)K, 2 * sms) : 1;
    *bytes = splits > 1 ? 4096 + (size_t)tiles * splits * 16 * 256 * 16 : 0;
  } else if (!strcmp(kernel, "hgemm_f16") || !strcmp(kernel, "gemm_bf16") || !strcmp(kernel, "linear_s8") ||
             !strcmp(kernel, "linear_bf16") || !strcmp(kernel, "linear_f16")) {
    // the soft wave rendezvous of long problems: one 256-byte counter block
    const long long kb = div_up(K, !strcmp(kernel, "linear_s8") ? 128 : 64);
    const long long tiles = div_up(m_or_n, 256) * div_up(N, 256);
    *bytes = (tiles >= 16 * (sms / 2) && kb >= 128) ? 256 : 0;
  } else {
    CUXB_REQUIRE(false, CUXB_ERR_INVALID, "cuxb_workspace_bytes: unknown kernel '%s'", kernel);
  }
  return 0;
}

// test hook (not in cuxb.h): rows per chunk the host pipelines would use; kind 0 = fp32 sgemm, 1 = fp16
// gemm (fp16 out), 2 = int8 linear.  Pure host arithmetic, callable without a GPU.
CUXB_API int cuxb_debug_host_chunk_rows(int kind, int M, int N, int K, int sm_count) {
  PipeCost c;
  const int ab = kind == 0 ? 4 : (kind == 1 ? 2 : 1), cb = kind == 1 ? 2 : 4;
  const int gran = kind == 0 ? 128 : 256, slots = kind == 0 ? 2 * sm_count : sm_count / 2;
  const double rate = kind == 0 ? 60e12 : (kind == 1 ? 1.6e15 : 3.2e15);
  c.up_b_bytes = (double)K * N * ab;
  c.up_row_bytes = (double)K * ab;
  c.down_row_bytes = (double)N * cb;
  c.wave_seconds = 2.0 * gran * gran * (double)K * slots / rate;
  return pipe_chunk_rows(M, gran, (N + gran - 1) / gran, slots, c);
}
