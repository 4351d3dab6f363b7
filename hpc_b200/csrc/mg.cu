// mg.cu -- the multi-GPU driver of the C ABI: ONE host process (thread) driving g GPUs of one NVLink box.
//
// The reference's only multi-worker GEMM is the MPI tutorial (mpi/alg_matrix_multiply.cpp:52-80): the
// master sends row blocks of A and all of B to the workers, every worker multiplies, the master gathers
// the rows of C.  Its single-device harness is cux::Executor::Initialize(dev)
// (0-frameworks/cux/src/executor.cpp:10-22).  cuxb_mg is that Initialize generalised to g devices: it owns
// one cuxb_ctx and one stream per device, enables peer access between all pairs, allocates the exchange
// buffers of the fused reductions / scan and the flag buffers and copies of B of the fused broadcast, loads
// every kernel on every device up front (kernels of different devices wait for each other, and a first
// launch that still has to load its module can wait for the peer device to go idle -- which a kernel
// spinning on a peer never does), and routes the broadcast (2 GPUs: the receiver pulls from the root;
// more: rank r pulls from rank r-1's copy, so every GPU sends B once).  The cuxb_mg_all_* calls issue one
// asynchronous launch per device, in rank order, and return; cuxb_mg_sync waits for all of them.
//
// Everything below is host code over the kernels' own entry points (cuxb_mg_dot_f32 ... cuxb_mg_linear_s8_s32):
// no kernel lives here.
#include <string.h>

#include "common.cuh"

struct cuxb_mg {
  int world;
  int device[8];
  cuxb_ctx *ctx[8];
  cudaStream_t stream[8];
  void *exchange[8];        // CUXB_MG_EXCHANGE_BYTES each, zeroed once
  void *chain_flags[8];     // CUXB_MG_CHAIN_FLAGS_BYTES each, zeroed once (world > 2)
  void *b_copy[8];          // rank r's copy of B (r >= 1), grown on demand
  size_t b_copy_bytes;
  bool chain_on;
  bool warmed[3];           // GEMM kinds whose kernels have been loaded on every device
};

namespace cuxb {

static int mg_check(const char *name, const cuxb_mg *mg) {
  CUXB_REQUIRE(mg != nullptr && mg->world >= 1 && mg->world <= 8, CUXB_ERR_INVALID, "%s: null or corrupt cuxb_mg", name);
  return 0;
}

// One tiny call of every kernel family on `ctx`'s device, so that its code is resident before any launch
// on this device has to wait for another GPU.  `scratch` is >= 64 KiB of device memory on that device.
static int mg_warm_up(cuxb_ctx *ctx, cudaStream_t st, char *scratch) {
  float *f = (float *)scratch;
  int32_t *i32 = (int32_t *)(scratch + 16384);
  int rc;
  CUXB_CUDA(cudaMemsetAsync(scratch, 0, 65536, st));
  if ((rc = cuxb_dot_f32(ctx, st, 1024, f, f, f + 2048))) return rc;
  if ((rc = cuxb_sum_f32(ctx, st, 1024, f, f + 2048))) return rc;
  if ((rc = cuxb_nrm2_f32(ctx, st, 1024, f, f + 2048))) return rc;
  if ((rc = cuxb_sum_i32(ctx, st, 1024, i32, i32 + 2048))) return rc;
  if ((rc = cuxb_scan_i32(ctx, st, 1024, i32, i32 + 2048, 0))) return rc;
  return 0;
}

// The GEMM kernels are much larger; they are loaded by a real (small) problem the first time a GEMM of that
// kind is asked for: see mg_warm_gemm below.
static int mg_grow_b(cuxb_mg *mg, size_t bytes) {
  if (mg->b_copy_bytes >= bytes) return 0;
  for (int r = 1; r < mg->world; r++) {
    CUXB_CUDA(cudaSetDevice(mg->device[r]));
    CUXB_CUDA(cudaDeviceSynchronize());
    if (mg->b_copy[r]) CUXB_CUDA(cudaFree(mg->b_copy[r]));
    mg->b_copy[r] = nullptr;
    CUXB_CUDA(cudaMalloc(&mg->b_copy[r], bytes));
  }
  mg->b_copy_bytes = bytes;
  return 0;
}

// Chain routing for rank r (see cuxb_mg_set_chain): flags of this rank, of the rank it pulls from (none
// when that is the root) and of the rank that pulls from it (none for the last).
static int mg_route(cuxb_mg *mg) {
  if (mg->chain_on || mg->world <= 2) return 0;
  for (int r = 1; r < mg->world; r++) {
    int rc = cuxb_mg_set_chain(mg->ctx[r], mg->chain_flags[r], r >= 2 ? mg->chain_flags[r - 1] : nullptr,
                               r + 1 < mg->world ? mg->chain_flags[r + 1] : nullptr);
    if (rc) return rc;
  }
  mg->chain_on = true;
  return 0;
}

// the B a rank's kernel reads from: the root's own B for rank 1 (and for everybody with two GPUs), else
// the copy of the rank before it
static const void *mg_upstream(const cuxb_mg *mg, int r, const void *b_root) {
  return (r <= 1 || mg->world <= 2) ? b_root : (const void *)mg->b_copy[r - 1];
}

}  // namespace cuxb

using namespace cuxb;

CUXB_API int cuxb_mg_create(int ndev, const int *devices, cuxb_mg **out) {
  CUXB_REQUIRE(out != nullptr, CUXB_ERR_INVALID, "cuxb_mg_create: out is null");
  *out = nullptr;
  CUXB_REQUIRE(ndev >= 1 && ndev <= 8, CUXB_ERR_INVALID, "cuxb_mg_create: %d devices (1..8 supported)", ndev);
  int count = 0;
  CUXB_CUDA(cudaGetDeviceCount(&count));
  cuxb_mg *mg = new cuxb_mg();
  memset(mg, 0, sizeof(*mg));
  mg->world = ndev;
  for (int r = 0; r < ndev; r++) {
    mg->device[r] = devices ? devices[r] : r;
    if (mg->device[r] < 0 || mg->device[r] >= count) {
      set_error("cuxb_mg_create: device %d out of range (%d devices)", mg->device[r], count);
      delete mg;
      return CUXB_ERR_INVALID;
    }
    for (int q = 0; q < r; q++)
      if (mg->device[q] == mg->device[r]) {
        set_error("cuxb_mg_create: device %d listed twice (one rank per GPU)", mg->device[r]);
        delete mg;
        return CUXB_ERR_INVALID;
      }
  }
  int rc = 0;
  for (int a = 0; a < ndev && !rc; a++)
    for (int b = 0; b < ndev && !rc; b++)
      if (a != b) rc = cuxb_enable_peer_access(mg->device[a], mg->device[b]);
  for (int r = 0; r < ndev && !rc; r++) {
    rc = cuxb_ctx_create(mg->device[r], &mg->ctx[r]);
    if (rc) break;
    cudaError_t e = cudaStreamCreateWithFlags(&mg->stream[r], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&mg->exchange[r], CUXB_MG_EXCHANGE_BYTES);
    if (e == cudaSuccess) e = cudaMemset(mg->exchange[r], 0, CUXB_MG_EXCHANGE_BYTES);
    if (e == cudaSuccess) e = cudaMalloc(&mg->chain_flags[r], CUXB_MG_CHAIN_FLAGS_BYTES);
    if (e == cudaSuccess) e = cudaMemset(mg->chain_flags[r], 0, CUXB_MG_CHAIN_FLAGS_BYTES);
    void *scratch = nullptr;
    if (e == cudaSuccess) e = cudaMalloc(&scratch, 65536);
    if (e != cudaSuccess) {
      set_error("cuxb_mg_create: %s (device %d)", cudaGetErrorString(e), mg->device[r]);
      rc = (int)e;
      break;
    }
    rc = mg_warm_up(mg->ctx[r], mg->stream[r], (char *)scratch);
    cudaStreamSynchronize(mg->stream[r]);
    cudaFree(scratch);
  }
  if (rc) {
    cuxb_mg_destroy(mg);
    return rc;
  }
  *out = mg;
  return 0;
}

// SURVEY 8(b) names the constructor cuxb_mg_init(ndev, &mg): devices 0 .. ndev-1
CUXB_API int cuxb_mg_init(int ndev, cuxb_mg **out) { return cuxb_mg_create(ndev, nullptr, out); }

CUXB_API int cuxb_mg_destroy(cuxb_mg *mg) {
  if (!mg) return 0;
  for (int r = 0; r < mg->world; r++) {
    cudaSetDevice(mg->device[r]);
    cudaDeviceSynchronize();
    if (mg->ctx[r]) cuxb_ctx_destroy(mg->ctx[r]);
    if (mg->stream[r]) cudaStreamDestroy(mg->stream[r]);
    if (mg->exchange[r]) cudaFree(mg->exchange[r]);
    if (mg->chain_flags[r]) cudaFree(mg->chain_flags[r]);
    if (mg->b_copy[r]) cudaFree(mg->b_copy[r]);
  }
  delete mg;
  return 0;
}

CUXB_API int cuxb_mg_world(const cuxb_mg *mg) { return mg ? mg->world : 0; }
CUXB_API int cuxb_mg_device(const cuxb_mg *mg, int rank) {
  return (mg && rank >= 0 && rank < mg->world) ? mg->device[rank] : -1;
}
CUXB_API cuxb_ctx *cuxb_mg_ctx(cuxb_mg *mg, int rank) {
  return (mg && rank >= 0 && rank < mg->world) ? mg->ctx[rank] : nullptr;
}
CUXB_API void *cuxb_mg_stream(cuxb_mg *mg, int rank) {
  return (mg && rank >= 0 && rank < mg->world) ? (void *)mg->stream[rank] : nullptr;
}
CUXB_API int cuxb_mg_sync(cuxb_mg *mg) {
  int rc = mg_check("cuxb_mg_sync", mg);
  if (rc) return rc;
  for (int r = 0; r < mg->world; r++) {
    CUXB_CUDA(cudaSetDevice(mg->device[r]));
    CUXB_CUDA(cudaStreamSynchronize(mg->stream[r]));
  }
  return 0;
}

// ---- sharded reductions and scan: rank r holds the r-th contiguous slice; n[r], the input pointers and
// the result pointers are per rank (device memory of that rank's GPU).  Every rank ends with the same
// scalar (reductions) / with its slice of the global scan.
namespace cuxb {
static void *const *mg_peers(const cuxb_mg *mg) { return (void *const *)mg->exchange; }
}  // namespace cuxb

CUXB_API int cuxb_mg_all_dot_f32(cuxb_mg *mg, const long long *n, const float *const *a, const float *const *b,
                                 float *const *res) {
  int rc = mg_check("cuxb_mg_all_dot_f32", mg);
  if (rc) return rc;
  CUXB_REQUIRE(n && a && b && res, CUXB_ERR_INVALID, "cuxb_mg_all_dot_f32: null argument array");
  for (int r = 0; r < mg->world; r++)
    if ((rc = cuxb_mg_dot_f32(mg->ctx[r], mg->stream[r], n[r], a[r], b[r], res[r], mg->world, r, mg_peers(mg)))) return rc;
  return 0;
}
CUXB_API int cuxb_mg_all_sum_f32(cuxb_mg *mg, const long long *n, const float *const *x, float *const *res) {
  int rc = mg_check("cuxb_mg_all_sum_f32", mg);
  if (rc) return rc;
  CUXB_REQUIRE(n && x && res, CUXB_ERR_INVALID, "cuxb_mg_all_sum_f32: null argument array");
  for (int r = 0; r < mg->world; r++)
    if ((rc = cuxb_mg_sum_f32(mg->ctx[r], mg->stream[r], n[r], x[r], res[r], mg->world, r, mg_peers(mg)))) return rc;
  return 0;
}
CUXB_API int cuxb_mg_all_nrm2_f32(cuxb_mg *mg, const long long *n, const float *const *x, float *const *res) {
  int rc = mg_check("cuxb_mg_all_nrm2_f32", mg);
  if (rc) return rc;
  CUXB_REQUIRE(n && x && res, CUXB_ERR_INVALID, "cuxb_mg_all_nrm2_f32: null argument array");
  for (int r = 0; r < mg->world; r++)
    if ((rc = cuxb_mg_nrm2_f32(mg->ctx[r], mg->stream[r], n[r], x[r], res[r], mg->world, r, mg_peers(mg)))) return rc;
  return 0;
}
CUXB_API int cuxb_mg_all_scan_i32(cuxb_mg *mg, const long long *n, const int32_t *const *in, int32_t *const *out) {
  int rc = mg_check("cuxb_mg_all_scan_i32", mg);
  if (rc) return rc;
  CUXB_REQUIRE(n && in && out, CUXB_ERR_INVALID, "cuxb_mg_all_scan_i32: null argument array");
  for (int r = 0; r < mg->world; r++)
    if ((rc = cuxb_mg_scan_i32(mg->ctx[r], mg->stream[r], n[r], in[r], out[r], mg->world, r, mg_peers(mg)))) return rc;
  return 0;
}

// ---- row-sharded GEMMs: rank r multiplies its M[r] rows of A (and C) with the one B that lives on rank 0;
// the broadcast of B runs inside the kernels (cuxb_mg_*gemm*), into copies this object owns.
namespace cuxb {

enum { MG_SGEMM = 0, MG_HGEMM = 1, MG_LIN8 = 2 };

// First use of a GEMM kind: run it once per device on a tiny problem of its own, synchronously, so that the
// kernel (for the SGEMM also its copying instantiation) is loaded before launches start waiting on peers.
static int mg_warm_gemm(cuxb_mg *mg, int kind) {
  static_assert(sizeof(float) == 4, "");
  for (int r = 0; r < mg->world; r++) {
    CUXB_CUDA(cudaSetDevice(mg->device[r]));
    char *s = nullptr;
    const int M = 128, N = 256, K = 128;
    CUXB_CUDA(cudaMalloc((void **)&s, 4 * (size_t)(M * K + 2 * K * N + M * N)));
    CUXB_CUDA(cudaMemset(s, 0, 4 * (size_t)(M * K + 2 * K * N + M * N)));
    float *A = (float *)s, *B = A + M * K, *B2 = B + K * N, *C = B2 + K * N;
    int rc = 0;
    cuxb_ctx *c = mg->ctx[r];
    cudaStream_t st = mg->stream[r];
    if (kind == MG_SGEMM) {
      rc = cuxb_sgemm_f32(c, st, M, N, K, 1.f, A, K, B, N, 0.f, C, N);
      if (!rc && r > 0) rc = cuxb_mg_sgemm_f32(c, st, M, N, K, 1.f, A, K, B, B2, N, 0.f, C, N, 0);   // pulls from itself
    } else if (kind == MG_HGEMM) {
      rc = cuxb_hgemm_f16_f32acc(c, st, M, N, K, 1.f, A, K, B, N, 0.f, C, N, CUXB_FLOAT16);
      if (!rc) rc = cuxb_hgemm_f16_f32acc(c, st, M, N, K, 1.f, A, K, B, N, 0.f, C, N, CUXB_FLOAT32);
    } else {
      rc = cuxb_linear_s8_s32(c, st, M, N, K, (const int8_t *)A, (const int8_t *)B, (int32_t *)C);
    }
    cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(s);
    if (rc) return rc;
    CUXB_CUDA(e);
  }
  return 0;
}

static int mg_prepare_gemm(const char *name, cuxb_mg *mg, int kind, const int *M, size_t b_bytes) {
  int rc = mg_check(name, mg);
  if (rc) return rc;
  CUXB_REQUIRE(M != nullptr, CUXB_ERR_INVALID, "%s: null row-count array", name);
  if (!mg->warmed[kind]) {
    // the self-pull of the SGEMM warm-up must not run through a chain
    for (int r = 1; r < mg->world && mg->chain_on; r++) cuxb_mg_set_chain(mg->ctx[r], nullptr, nullptr, nullptr);
    mg->chain_on = false;
    if ((rc = mg_warm_gemm(mg, kind))) return rc;
    mg->warmed[kind] = true;
  }
  if ((rc = mg_grow_b(mg, b_bytes))) return rc;
  return mg_route(mg);
}

}  // namespace cuxb

CUXB_API int cuxb_mg_all_sgemm_f32(cuxb_mg *mg, const int *M, int N, int K, float alpha, const float *const *A, int lda,
                                   const float *B_root, int ldb, float beta, float *const *C, int ldc) {
  int rc = mg_prepare_gemm("cuxb_mg_all_sgemm_f32", mg, MG_SGEMM, M, (size_t)K * ldb * sizeof(float));
  if (rc) return rc;
  CUXB_REQUIRE(A && C && B_root, CUXB_ERR_INVALID, "cuxb_mg_all_sgemm_f32: null pointer");
  for (int r = 0; r < mg->world; r++) {
    rc = cuxb_mg_sgemm_f32(mg->ctx[r], mg->stream[r], M[r], N, K, alpha, A[r], lda,
                           (const float *)mg_upstream(mg, r, B_root), (float *)mg->b_copy[r], ldb, beta, C[r], ldc, r == 0);
    if (rc) return rc;
  }
  return 0;
}

CUXB_API int cuxb_mg_all_hgemm_f16_f32acc(cuxb_mg *mg, const int *M, int N, int K, float alpha, const void *const *A,
                                          int lda, const void *B_root, int ldb, float beta, void *const *C, int ldc,
                                          int c_dtype) {
  int rc = mg_prepare_gemm("cuxb_mg_all_hgemm_f16_f32acc", mg, MG_HGEMM, M, (size_t)K * ldb * 2);
  if (rc) return rc;
  CUXB_REQUIRE(A && C && B_root, CUXB_ERR_INVALID, "cuxb_mg_all_hgemm_f16_f32acc: null pointer");
  for (int r = 0; r < mg->world; r++) {
    rc = cuxb_mg_hgemm_f16_f32acc(mg->ctx[r], mg->stream[r], M[r], N, K, alpha, A[r], lda, mg_upstream(mg, r, B_root),
                                  mg->b_copy[r], ldb, beta, C[r], ldc, c_dtype, r == 0);
    if (rc) return rc;
  }
  return 0;
}

CUXB_API int cuxb_mg_all_linear_s8_s32(cuxb_mg *mg, const int *M, int N, int K, const int8_t *const *A,
                                       const int8_t *Bt_root, int32_t *const *C) {
  int rc = mg_prepare_gemm("cuxb_mg_all_linear_s8_s32", mg, MG_LIN8, M, (size_t)N * K);
  if (rc) return rc;
  CUXB_REQUIRE(A && C && Bt_root, CUXB_ERR_INVALID, "cuxb_mg_all_linear_s8_s32: null pointer");
  for (int r = 0; r < mg->world; r++) {
    rc = cuxb_mg_linear_s8_s32(mg->ctx[r], mg->stream[r], M[r], N, K, A[r], (const int8_t *)mg_upstream(mg, r, Bt_root),
                               (int8_t *)mg->b_copy[r], C[r], r == 0);
    if (rc) return rc;
  }
  return 0;
}
