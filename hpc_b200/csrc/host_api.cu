// host_api.cu -- host-buffer entry points and the occupancy hook.
//
// The *_host functions are what a cux plugin that is handed HOST arrays calls
// (cux itself stages through Array4D::GetGpuData / GetCpuData with synchronous
// cudaMemcpy on pageable memory, 0-frameworks/cux/include/array.h:84-114).  Each call
// copies its inputs host->device on the context's private stream, launches the kernel
// on the same stream, copies the result back and waits for the stream.  Staging
// buffers live in the cuxb_ctx and are reused between calls.
#include <string.h>

#include "common.cuh"

namespace cuxb {
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
  void *d;
  CUXB_TRY(h2d(ctx, 0, in, (size_t)n * 4, &d));
  CUXB_TRY(cuxb_scan_i32(ctx, ctx->stage_stream, n, (const int32_t *)d, (int32_t *)d, seed));  // in place
  return d2h_sync(ctx, out, d, (size_t)n * 4);
}

// copy a row-major [rows, cols] host matrix with pitch ld into a DENSE device matrix
static int h2d_2d(cuxb_ctx *ctx, int slot, const void *h, int rows, int cols, int ld, int ebytes, void **d) {
  CUXB_TRY(stage_reserve(ctx, slot, (size_t)rows * cols * ebytes + 16, d));
  CUXB_CUDA(cudaMemcpy2DAsync(*d, (size_t)cols * ebytes, h, (size_t)ld * ebytes, (size_t)cols * ebytes, rows,
                              cudaMemcpyHostToDevice, ctx->stage_stream));
  return 0;
}
static int d2h_2d_sync(cuxb_ctx *ctx, void *h, const void *d, int rows, int cols, int ld, int ebytes) {
  CUXB_CUDA(cudaMemcpy2DAsync(h, (size_t)ld * ebytes, d, (size_t)cols * ebytes, (size_t)cols * ebytes, rows,
                              cudaMemcpyDeviceToHost, ctx->stage_stream));
  CUXB_CUDA(cudaStreamSynchronize(ctx->stage_stream));
  return 0;
}

CUXB_API int cuxb_sgemm_f32_host(cuxb_ctx *ctx, int M, int N, int K, float alpha, const float *A, int lda,
                                 const float *B, int ldb, float beta, float *C, int ldc) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(M > 0 && N > 0 && K > 0 && A && B && C, CUXB_ERR_INVALID, "cuxb_sgemm_f32_host: bad argument");
  std::lock_guard<std::mutex> g(*ctx->mu);
  void *da, *db, *dc;
  CUXB_TRY(h2d_2d(ctx, 0, A, M, K, lda, 4, &da));
  CUXB_TRY(h2d_2d(ctx, 1, B, K, N, ldb, 4, &db));
  if (beta != 0.f)
    CUXB_TRY(h2d_2d(ctx, 2, C, M, N, ldc, 4, &dc));
  else
    CUXB_TRY(stage_reserve(ctx, 2, (size_t)M * N * 4, &dc));
  CUXB_TRY(cuxb_sgemm_f32(ctx, ctx->stage_stream, M, N, K, alpha, (const float *)da, K, (const float *)db, N, beta,
                          (float *)dc, N));
  return d2h_2d_sync(ctx, C, dc, M, N, ldc, 4);
}

CUXB_API int cuxb_hgemm_f16_f32acc_host(cuxb_ctx *ctx, int M, int N, int K, float alpha, const void *A, int lda,
                                        const void *B, int ldb, float beta, void *C, int ldc, int c_dtype) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(M > 0 && N > 0 && K > 0 && A && B && C, CUXB_ERR_INVALID, "cuxb_hgemm_f16_f32acc_host: bad argument");
  CUXB_REQUIRE(c_dtype == CUXB_FLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID, "bad c_dtype");
  CUXB_REQUIRE(K % 8 == 0 && N % 8 == 0, CUXB_ERR_UNSUPPORTED, "cuxb_hgemm_f16_f32acc_host: K and N must be multiples of 8");
  const int cb = c_dtype == CUXB_FLOAT16 ? 2 : 4;
  std::lock_guard<std::mutex> g(*ctx->mu);
  void *da, *db, *dc;
  CUXB_TRY(h2d_2d(ctx, 0, A, M, K, lda, 2, &da));
  CUXB_TRY(h2d_2d(ctx, 1, B, K, N, ldb, 2, &db));
  if (beta != 0.f)
    CUXB_TRY(h2d_2d(ctx, 2, C, M, N, ldc, cb, &dc));
  else
    CUXB_TRY(stage_reserve(ctx, 2, (size_t)M * N * cb, &dc));
  CUXB_TRY(cuxb_hgemm_f16_f32acc(ctx, ctx->stage_stream, M, N, K, alpha, da, K, db, N, beta, dc, N, c_dtype));
  return d2h_2d_sync(ctx, C, dc, M, N, ldc, cb);
}

CUXB_API int cuxb_linear_s8_s32_host(cuxb_ctx *ctx, int M, int N, int K, const int8_t *A, const int8_t *Bt,
                                     int32_t *C) {
  CUXB_TRY(ensure_device(ctx));
  CUXB_REQUIRE(M > 0 && N > 0 && K > 0 && A && Bt && C, CUXB_ERR_INVALID, "cuxb_linear_s8_s32_host: bad argument");
  std::lock_guard<std::mutex> g(*ctx->mu);
  void *da, *db, *dc;
  CUXB_TRY(h2d(ctx, 0, A, (size_t)M * K, &da));
  CUXB_TRY(h2d(ctx, 1, Bt, (size_t)N * K, &db));
  CUXB_TRY(stage_reserve(ctx, 2, (size_t)M * N * 4, &dc));
  CUXB_TRY(cuxb_linear_s8_s32(ctx, ctx->stage_stream, M, N, K, (const int8_t *)da, (const int8_t *)db, (int32_t *)dc));
  return d2h_sync(ctx, C, dc, (size_t)M * N * 4);
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
