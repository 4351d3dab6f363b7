// ref_shim.cpp -- C entry points over the REFERENCE's own host functions.
//
// TEST INFRASTRUCTURE ONLY (see oracle/oracle.c header).  Built by
// oracle/build.py into oracle/_ref/libcuxref.so when /root/reference exists.
// No reference source is stored in this repository: the standalone samples
// are compiled where they lie (their main() renamed on the command line) and
// the cux host kernels -- whose translation units only build with MSVC
//  -- are pulled in by line range at build time into
// oracle/_ref/*.inc (git-ignored) and #included below.
//
//   gen_cux_gemm.inc   <- 0-frameworks/cux/src/operator/gemm.cpp:11-275
//   gen_cux_dot.inc    <- 0-frameworks/cux/src/operator/dot_product.cpp:5-13
//   gen_cux_nrm2.inc   <- 0-frameworks/cux/src/operator/nrm2.cpp:5-16
//   gen_hgemm_host.inc <- cuda/gemm_fp16_wmma.cu:75-90
//   gen_reduce_host.inc<- cuda/reduce_fp32.cu:26-31
#include <immintrin.h>
#include <omp.h>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cuda_fp16.h>  // host-side __half2float / __float2half (RNE)

namespace cuxref {
#include "gen_cux_gemm.inc"
#include "gen_cux_dot.inc"
#include "gen_cux_nrm2.inc"
}  // namespace cuxref

namespace cudaref {
#include "gen_hgemm_host.inc"
#include "gen_reduce_host.inc"
}  // namespace cudaref

// simd/x86/linear.cpp (compiled in place, main -> ref_linear_main)
typedef uint16_t bfloat16_t_;
void LinearNormalFp32(const int M, const int N, const int K, const float *A, const float *B, float *C);
void LinearAvx512Fp32(const int M, const int N, const int K, const float *A, const float *B, float *C);
void LinearNormalBf16(const int M, const int N, const int K, const uint16_t *A, const uint16_t *B, uint16_t *C);
void LinearAvx512Bf16(const int M, const int N, const int K, const uint16_t *A, const uint16_t *B, uint16_t *C);
void LinearNormalInt8(const int M, const int N, const int K, const int8_t *A, const int8_t *B, int32_t *C);
void LinearAvx512Int8(const int M, const int N, const int K, const int8_t *A, const int8_t *B, int32_t *C);
// simd/x86/vector_scan.cpp (main -> ref_scan_main)
void VectorScanNormal(const float *vec_in, const int len, float *vec_out);
void VectorScanSIMDv1(const float *vec_in, const int len, float *vec_out);
void VectorScanSIMDv2(const float *vec_in, const int len, float *vec_out);
// openmp/alg_matrix_multiply.cpp (main -> ref_ompmm_main)
void MatrixMulSerial(const int M, const int N, const int K, const float ALPHA, const float *A,
                     const int lda, const float *B, const int ldb, float *C, const int ldc);
void MatrixMulMP(const int M, const int N, const int K, const float ALPHA, const float *A,
                 const int lda, const float *B, const int ldb, float *C, const int ldc);

// std/alg_vector_dot_product.cpp (main -> ref_stddot_main): the reference's all-cores dot product
void DotProductMTv1(float *src0, float *src1, int len, int num_threads, float *dst);

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API int ref_cpu_has_avx512(void) {
  return __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
         __builtin_cpu_supports("avx512vnni");
}
REF_API int ref_cpu_has_avx512bf16(void) { return __builtin_cpu_supports("avx512bf16"); }
REF_API int ref_num_threads(void) { return omp_get_max_threads(); }

// version: 0..5 = GemmHostV0..V5
REF_API int ref_cux_gemm(int version, int M, int N, int K, float alpha, const float *A, int lda,
                         const float *B, int ldb, float beta, float *C, int ldc) {
  switch (version) {
    case 0: cuxref::GemmHostV0(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc); return 0;
    case 1: cuxref::GemmHostV1(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc); return 0;
    case 2: cuxref::GemmHostV2(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc); return 0;
    case 3: cuxref::GemmHostV3(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc); return 0;
    case 4: cuxref::GemmHostV4(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc); return 0;
    case 5: cuxref::GemmHostV5(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc); return 0;
  }
  return -1;
}
// DotProductMTv1 (std/alg_vector_dot_product.cpp:19-56): one std::thread per slice, partials added in
// slice order.  The sample clears num_threads floats at dst (:28), so dst gets that much room here.
REF_API void ref_std_dot_mt(int len, const float *a, const float *b, int num_threads, float *res) {
  if (num_threads < 1) num_threads = 1;
  float *dst = (float *)calloc((size_t)num_threads + 1, sizeof(float));
  DotProductMTv1(const_cast<float *>(a), const_cast<float *>(b), len, num_threads, dst);
  *res = dst[0];
  free(dst);
}
REF_API void ref_cux_dot_v0(int len, const float *a, const float *b, float *res) {
  cuxref::DotHostV0(len, a, b, res);
}
REF_API void ref_cux_nrm2_v0(int n, const float *x, float *res) {
  cuxref::Nrm2HostV0(n, const_cast<float *>(x), res);
}
REF_API void ref_reduce_host(const float *x, float *y, int n) { cudaref::ReduceHost(x, y, n); }
REF_API void ref_hgemm_host(int M, int N, int K, const uint16_t *A, int lda, const uint16_t *B,
                            int ldb, uint16_t *C, int ldc) {
  cudaref::GemmHostV1(M, N, K, reinterpret_cast<const half *>(A), lda,
                      reinterpret_cast<const half *>(B), ldb, reinterpret_cast<half *>(C), ldc);
}
REF_API void ref_omp_gemm(int parallel, int M, int N, int K, float alpha, const float *A, int lda,
                          const float *B, int ldb, float *C, int ldc) {
  if (parallel)
    MatrixMulMP(M, N, K, alpha, A, lda, B, ldb, C, ldc);
  else
    MatrixMulSerial(M, N, K, alpha, A, lda, B, ldb, C, ldc);
}
// variant: 0 = LinearNormalInt8, 1 = LinearAvx512Int8 (u8 x s8, valid for A >= 0 only)
REF_API int ref_linear_s8(int variant, int M, int N, int K, const int8_t *A, const int8_t *B,
                          int32_t *C) {
  if (variant == 0) { LinearNormalInt8(M, N, K, A, B, C); return 0; }
  if (!ref_cpu_has_avx512()) return -2;
  LinearAvx512Int8(M, N, K, A, B, C);
  return 0;
}
// variant: 0 = LinearNormalBf16, 1 = LinearAvx512Bf16
REF_API int ref_linear_bf16(int variant, int M, int N, int K, const uint16_t *A,
                            const uint16_t *B, uint16_t *C) {
  if (variant == 0) { LinearNormalBf16(M, N, K, A, B, C); return 0; }
  if (!ref_cpu_has_avx512bf16()) return -2;
  LinearAvx512Bf16(M, N, K, A, B, C);
  return 0;
}
REF_API void ref_linear_f32(int M, int N, int K, const float *A, const float *B, float *C) {
  LinearNormalFp32(M, N, K, A, B, C);
}
// variant: 0 = VectorScanNormal, 1 = SIMDv1 (len%4==0, out 16B aligned), 2 = SIMDv2 (len%8==0)
REF_API void ref_scan_f32(int variant, const float *in, int len, float *out) {
  if (variant == 0) VectorScanNormal(in, len, out);
  else if (variant == 1) VectorScanSIMDv1(in, len, out);
  else VectorScanSIMDv2(in, len, out);
}
