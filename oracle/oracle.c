/*
 * oracle.c -- CPU restatement of the cjmcv/hpc "cux operator" hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, the smoke()
 * entry of __graft_entry__.py and the cpu_baseline / --impl reference legs of
 * bench.py may load it.  The product path (libcuxb.so, hpc_b200/) never links,
 * imports or calls anything in this directory.
 *
 * Parity status: PINNED.  Every function below is checked bit-for-bit against
 * the reference's own host code compiled from /root/reference (oracle/_ref,
 * see oracle/build.py and tests/test_oracle_vs_ref.py) and against the golden
 * fixtures under tests/golden/ that were generated from that build
 * (tests/golden/make_golden.py).
 *
 * Each function cites the reference lines it restates.  Loop order and
 * rounding points follow the reference exactly, because for floating point the
 * order of the additions IS the specification.  Everything is scalar on
 * purpose; the file is compiled with -ffp-contract=off so that no fused
 * multiply-add is introduced where the reference (compiled by its own
 * Makefile without -mfma for these scalar loops) rounds twice.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------ */
/* fp16 / bf16 <-> fp32, round-to-nearest-even.                              */
/* Follows pocket-ai/util/type.hpp:134-241 (half) and :268-288 (bf16); the   */
/* same RNE result as cux/include/util/half.h:156-203 and CUDA's             */
/* __float2half / __half2float used by cuda/gemm_fp16_wmma.cu:84-86.         */
/* ------------------------------------------------------------------------ */
ORC_API uint16_t orc_f32_to_f16(float f) {
  uint32_t s;
  memcpy(&s, &f, 4);
  uint16_t sign = (uint16_t)((s >> 16) & 0x8000u);
  int32_t exp = (int32_t)((s >> 23) & 0xff) - 127;
  uint32_t man = s & 0x7fffffu;
  if ((s & 0x7fffffffu) == 0) return sign;
  if (exp > 15) {
    if (exp == 128 && man) return 0x7fff;
    return (uint16_t)(sign | 0x7c00u);
  }
  uint16_t u = 0;
  int sticky = 0;
  if (exp >= -14) {
    u = (uint16_t)((((uint32_t)(exp + 15)) & 0x1f) << 10);
    u = (uint16_t)(u | (man >> 13));
  } else {
    int rshift = -14 - exp;
    if (rshift < 32) {
      man |= (1u << 23);
      sticky = (man & ((1u << rshift) - 1)) != 0;
      man >>= rshift;
      u = (uint16_t)((man >> 13) & 0x3ff);
    } else {
      man = 0;
      u = 0;
    }
  }
  int round_bit = (man >> 12) & 1;
  sticky |= (man & ((1u << 12) - 1)) != 0;
  if ((round_bit && sticky) || (round_bit && (u & 1))) u = (uint16_t)(u + 1);
  return (uint16_t)(u | sign);
}

ORC_API float orc_f16_to_f32(uint16_t h) {
  uint32_t sign = (h >> 15) & 1, exp = (h >> 10) & 0x1f, man = h & 0x3ff, f = 0;
  if (exp > 0 && exp < 31) {
    f = (sign << 31) | ((exp + 112) << 23) | (man << 13);
  } else if (exp == 0) {
    if (man) {
      exp = 113;
      while ((man & (1u << 10)) == 0) {
        man <<= 1;
        exp--;
      }
      man &= 0x3ff;
      f = (sign << 31) | (exp << 23) | (man << 13);
    } else {
      f = sign << 31;
    }
  } else {
    f = man ? 0x7fffffffu : ((0xffu << 23) | (sign << 31));
  }
  float out;
  memcpy(&out, &f, 4);
  return out;
}

ORC_API uint16_t orc_f32_to_bf16(float f) {
  uint32_t bits;
  memcpy(&bits, &f, 4);
  if ((bits & 0x7f800000u) != 0x7f800000u) {
    int mant = (bits & (1u << 16)) != 0, rnd = (bits & (1u << 15)) != 0;
    int sticky = (bits & ((1u << 15) - 1)) != 0;
    if ((rnd && sticky) || (rnd && mant)) bits += (1u << 16);
  } else if (bits & ~0xff800000u) {
    bits = 0x7fffffffu;
  }
  return (uint16_t)(bits >> 16);
}

ORC_API float orc_bf16_to_f32(uint16_t b) {
  uint32_t bits = ((uint32_t)b) << 16;
  float out;
  memcpy(&out, &bits, 4);
  return out;
}

ORC_API void orc_cvt_f32_to_f16(long long n, const float *src, uint16_t *dst) {
  for (long long i = 0; i < n; i++) dst[i] = orc_f32_to_f16(src[i]);
}
ORC_API void orc_cvt_f16_to_f32(long long n, const uint16_t *src, float *dst) {
  for (long long i = 0; i < n; i++) dst[i] = orc_f16_to_f32(src[i]);
}
ORC_API void orc_cvt_f32_to_bf16(long long n, const float *src, uint16_t *dst) {
  for (long long i = 0; i < n; i++) dst[i] = orc_f32_to_bf16(src[i]);
}
ORC_API void orc_cvt_bf16_to_f32(long long n, const uint16_t *src, float *dst) {
  for (long long i = 0; i < n; i++) dst[i] = orc_bf16_to_f32(src[i]);
}

/* ------------------------------------------------------------------------ */
/* fp32 GEMM.                                                                */
/* ------------------------------------------------------------------------ */

/* G0: cux GemmHostV0, 0-frameworks/cux/src/operator/gemm.cpp:13-34.
 * C <- beta*C; then per (i,j): t = C; for k: t += (alpha*A[i,k])*B[k,j].
 * `alpha * a * b` parses left to right, so alpha*a is rounded first. */
ORC_API void orc_gemm_f32_v0(int M, int N, int K, float alpha, const float *A, int lda,
                             const float *B, int ldb, float beta, float *C, int ldc) {
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) C[(size_t)i * ldc + j] *= beta;
  for (int i = 0; i < M; ++i) {
    for (int j = 0; j < N; ++j) {
      float t = C[(size_t)i * ldc + j];
      for (int k = 0; k < K; ++k) {
        float aa = alpha * A[(size_t)i * lda + k];
        float p = aa * B[(size_t)k * ldb + j];
        t += p;
      }
      C[(size_t)i * ldc + j] = t;
    }
  }
}

/* G1: cux GemmHostV1, gemm.cpp:38-58 (ikj order; identical per-element
 * summation order to G0, accumulating in memory instead of a register).
 * Rows are independent, so the optional OpenMP split over i (as in
 * openmp/alg_matrix_multiply.cpp:33-51, MatrixMulMP) does not change a bit. */
ORC_API void orc_gemm_f32_ikj(int M, int N, int K, float alpha, const float *A, int lda,
                              const float *B, int ldb, float beta, float *C, int ldc,
                              int threads) {
  (void)threads;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
  for (int i = 0; i < M; ++i) {
    float *c = C + (size_t)i * ldc;
    for (int j = 0; j < N; ++j) c[j] *= beta;
    for (int k = 0; k < K; ++k) {
      float aa = alpha * A[(size_t)i * lda + k];
      const float *b = B + (size_t)k * ldb;
      for (int j = 0; j < N; ++j) {
        float p = aa * b[j];
        c[j] += p;
      }
    }
  }
}

/* G7: cuda/gemm_fp32.cu:48-61 GemmHostV1 -- C = 0; C[i,j] += A[i,k]*B[k,j]. */
ORC_API void orc_gemm_f32_plain(int M, int N, int K, const float *A, int lda, const float *B,
                                int ldb, float *C, int ldc) {
  for (int i = 0; i < M; ++i) {
    float *c = C + (size_t)i * ldc;
    for (int j = 0; j < N; ++j) {
      float t = 0.f;
      for (int k = 0; k < K; ++k) {
        float p = A[(size_t)i * lda + k] * B[(size_t)k * ldb + j];
        t += p;
      }
      c[j] = t;
    }
  }
}

/* ------------------------------------------------------------------------ */
/* fp16 GEMM, fp32 accumulate.                                               */
/* H0: cuda/gemm_fp16_wmma.cu:75-90 GemmHostV1:                              */
/*   acc(float) += h2f(A[i,k]) * h2f(B[k,j]);  C[i,j] = f2h(acc)             */
/* Generalised with cux's alpha/beta (gemm.cu:266-269 cublasSgemmEx          */
/* semantics: fp32 compute, fp16 storage): C = f2h(alpha*acc + beta*h2f(C)). */
/* With alpha=1, beta=0 this is exactly H0.  c_is_f32 != 0 stores the fp32   */
/* value instead of rounding to half.                                        */
/* ------------------------------------------------------------------------ */
ORC_API void orc_hgemm_f16(int M, int N, int K, float alpha, const uint16_t *A, int lda,
                           const uint16_t *B, int ldb, float beta, void *C, int ldc,
                           int c_is_f32, int threads) {
  (void)threads;
  /* Same arithmetic as the reference loop -- for every (i, j): acc = 0; acc += a[i,k] * b[k,j] for k = 0..K-1 in
   * increasing order, each product and each sum rounded to fp32 -- but walked column BLOCK by column block with
   * k outside j, so that B is read along its rows.  The sequence of operations that produces any one C[i,j] is
   * unchanged, so the results are bit-identical to the i-j-k walk (tests/test_oracle_vs_ref.py pins that against
   * the reference's own function); it is only ~20x faster, which lets the GPU tests check more rows.
   * The fp16 -> fp32 conversions are exact, so doing them once per block changes nothing either. */
  enum { JB = 256 };
  const int nblk = (N + JB - 1) / JB;
  float *Af = (float *)malloc((size_t)(M > 0 ? M : 1) * (K > 0 ? K : 1) * sizeof(float));
  for (int i = 0; i < M; ++i)
    for (int k = 0; k < K; ++k) Af[(size_t)i * K + k] = orc_f16_to_f32(A[(size_t)i * lda + k]);
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 1)
  for (int jb = 0; jb < nblk; ++jb) {
    const int j0 = jb * JB, jn = (N - j0 < JB) ? N - j0 : JB;
    float *Bf = (float *)malloc((size_t)(K > 0 ? K : 1) * JB * sizeof(float));
    for (int k = 0; k < K; ++k)
      for (int j = 0; j < jn; ++j) Bf[(size_t)k * JB + j] = orc_f16_to_f32(B[(size_t)k * ldb + j0 + j]);
    for (int i = 0; i < M; ++i) {
      float acc[JB];
      for (int j = 0; j < jn; ++j) acc[j] = 0.f;
      for (int k = 0; k < K; ++k) {
        const float a = Af[(size_t)i * K + k];
        const float *brow = Bf + (size_t)k * JB;
        for (int j = 0; j < jn; ++j) {
          float p = a * brow[j];
          acc[j] += p;
        }
      }
      for (int j = 0; j < jn; ++j) {
        float r;
        if (beta == 0.0f) {
          r = alpha * acc[j];
        } else {
          float c0 = c_is_f32 ? ((float *)C)[(size_t)i * ldc + j0 + j]
                              : orc_f16_to_f32(((uint16_t *)C)[(size_t)i * ldc + j0 + j]);
          float t0 = alpha * acc[j], t1 = beta * c0;
          r = t0 + t1;
        }
        if (c_is_f32)
          ((float *)C)[(size_t)i * ldc + j0 + j] = r;
        else
          ((uint16_t *)C)[(size_t)i * ldc + j0 + j] = orc_f32_to_f16(r);
      }
    }
    free(Bf);
  }
  free(Af);
}

/* ------------------------------------------------------------------------ */
/* linear: C = A * B^T with B stored [N,K] (simd/x86/linear.cpp:2,55-66).    */
/* ------------------------------------------------------------------------ */

/* L0: LinearNormalInt8, simd/x86/linear.cpp:199-211.  int8 x int8 -> int32. */
ORC_API void orc_linear_s8(int M, int N, int K, const int8_t *A, const int8_t *B, int32_t *C,
                           int threads) {
  (void)threads;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
  for (int i = 0; i < M; ++i) {
    for (int j = 0; j < N; ++j) {
      int sum = 0;
      const int8_t *a = A + (size_t)i * K, *b = B + (size_t)j * K;
      for (int k = 0; k < K; ++k) sum += (int)a[k] * (int)b[k];
      C[(size_t)i * N + j] = sum;
    }
  }
}

/* H3: LinearNormalBf16, simd/x86/linear.cpp:137-151.
 * C[i,j] = bf16( sum_k f32(A[i,k]) * f32(B[j,k]) ), serial fp32 sum. */
ORC_API void orc_linear_bf16(int M, int N, int K, const uint16_t *A, const uint16_t *B,
                             void *C, int c_is_f32, int threads) {
  (void)threads;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
  for (int i = 0; i < M; ++i) {
    for (int j = 0; j < N; ++j) {
      float sum = 0.0f;
      for (int k = 0; k < K; ++k) {
        float a = orc_bf16_to_f32(A[(size_t)i * K + k]);
        float b = orc_bf16_to_f32(B[(size_t)j * K + k]);
        float p = a * b;
        sum += p;
      }
      if (c_is_f32)
        ((float *)C)[(size_t)i * N + j] = sum;
      else
        ((uint16_t *)C)[(size_t)i * N + j] = orc_f32_to_bf16(sum);
    }
  }
}

/* L2: LinearNormalFp32, simd/x86/linear.cpp:55-66. */
ORC_API void orc_linear_f32(int M, int N, int K, const float *A, const float *B, float *C) {
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      float sum = 0;
      for (int k = 0; k < K; ++k) {
        float p = A[(size_t)i * K + k] * B[(size_t)j * K + k];
        sum += p;
      }
      C[(size_t)i * N + j] = sum;
    }
}

/* ------------------------------------------------------------------------ */
/* Reductions.                                                               */
/* ------------------------------------------------------------------------ */

/* R0: ReduceHost, cuda/reduce_fp32.cu:26-31 -- serial fp32 sum. */
ORC_API void orc_sum_f32(long long n, const float *x, float *y) {
  float t = 0;
  for (long long i = 0; i < n; ++i) t += x[i];
  *y = t;
}

/* R2: DotHostV0, cux/src/operator/dot_product.cpp:7-13 (identical recurrence:
 * DotProductSerial, std/alg_vector_dot_product.cpp:11-16). */
ORC_API void orc_dot_f32(long long n, const float *a, const float *b, float *res) {
  float t = 0;
  for (long long i = 0; i < n; ++i) {
    float p = a[i] * b[i];
    t += p;
  }
  *res = t;
}

/* R3: DotHostV1, dot_product.cpp:17-40 -- four SSE lanes over 8-element
 * steps, lanes summed 0+1+2+3, then a scalar tail. */
ORC_API void orc_dot_f32_v1(long long n, const float *a, const float *b, float *res) {
  float result = 0;
  if (n > 8) {
    float lane[4] = {0, 0, 0, 0};
    for (long long i = 0; i < n - 7; i += 8)
      for (int l = 0; l < 4; ++l) {
        float p0 = a[i + l] * b[i + l], p1 = a[i + 4 + l] * b[i + 4 + l];
        float s = p0 + p1;
        lane[l] += s;
      }
    result = lane[0] + lane[1] + lane[2] + lane[3];
  }
  for (long long i = n / 8 * 8; i < n; ++i) {
    float p = a[i] * b[i];
    result += p;
  }
  *res = result;
}

/* R5: Nrm2HostV0, cux/src/operator/nrm2.cpp:7-16. */
ORC_API void orc_nrm2_f32(long long n, const float *x, float *res) {
  if (n <= 0) {
    *res = 0.0f;
    return;
  }
  if (n == 1) {
    *res = fabsf(x[0]);
    return;
  }
  float s = 0.0f;
  for (long long i = 0; i < n; ++i) {
    float p = x[i] * x[i];
    s += p;
  }
  *res = sqrtf(s);
}

/* fp64 "truth" for float-fill tolerance tests (no reference counterpart; used
 * only to bound the error of BOTH the fp32-serial reference and the GPU). */
ORC_API void orc_dot_f64(long long n, const float *a, const float *b, double *res,
                         double *abs_sum) {
  double t = 0, s = 0;
  for (long long i = 0; i < n; ++i) {
    double p = (double)a[i] * (double)(b ? b[i] : 1.0f);
    t += p;
    s += fabs(p);
  }
  *res = t;
  if (abs_sum) *abs_sum = s;
}

/* ------------------------------------------------------------------------ */
/* Prefix scan.                                                              */
/* ------------------------------------------------------------------------ */

/* S0: VectorScanNormal, simd/x86/vector_scan.cpp:21-26 (fp32 reference). */
ORC_API void orc_scan_f32(long long n, const float *in, float *out) {
  if (n <= 0) return;
  out[0] = in[0];
  for (long long i = 1; i < n; ++i) out[i] = in[i] + out[i - 1];
}

/* S0/S2 in int32: the same recurrence with the int element type of
 * tbb/base_parallel_scan.cpp:62-65.  Adds are done in uint32_t so that
 * wrap-around is defined (two's complement) and identical to the GPU. */
ORC_API void orc_scan_i32(long long n, const int32_t *in, int32_t *out, int32_t seed) {
  uint32_t acc = (uint32_t)seed;
  for (long long i = 0; i < n; ++i) {
    acc += (uint32_t)in[i];
    out[i] = (int32_t)acc;
  }
}

ORC_API void orc_sum_i32(long long n, const int32_t *in, int32_t *out) {
  uint32_t acc = 0;
  for (long long i = 0; i < n; ++i) acc += (uint32_t)in[i];
  *out = (int32_t)acc;
}

/* ------------------------------------------------------------------------ */
/* T0: ResultChecker::CheckArray, cux/include/operator/op_assistor.h:38-46.  */
/* diff = (sum_i |ref_i - x_i|) * scale, accumulated in float; pass iff      */
/* diff < 1e-3 (FLT_PRECISION_ERR_THRESH, util.h:137).                       */
/* ------------------------------------------------------------------------ */
ORC_API float orc_cux_check_f32(const float *ref, const float *x, long long len, float scale) {
  float diff = 0.0f;
  for (long long i = 0; i < len; ++i) {
    float t = ref[i] - x[i];
    diff += (t >= 0 ? t : -t);
  }
  return diff * scale;
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
