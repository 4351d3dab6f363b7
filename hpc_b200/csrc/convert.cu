// convert.cu -- element type conversion on the device (fp32 <-> fp16 / bf16, int32 <-> int8).
// Replaces the scalar host loop Array4D::PrecsCpuCvt (0-frameworks/cux/include/array.h:124-144),
// which cux::Executor::BindAndFill runs for every array and every extra dtype
// (src/executor.cpp:62-78).  fp32 -> 16-bit rounds to nearest even, the same result as
// cux::half::float2half (include/util/half.h:156-203) and pocket-ai's ConvertFp32ToHalf /
// ConvertFp32ToBf16 (pocket-ai/util/type.hpp:134-190,268-283).  int32 -> int8 truncates like the
// C++ conversion `dst[i] = src[i]`.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace cuxb {

template <typename S, typename D> __device__ __forceinline__ D cvt(S v);
template <> __device__ __forceinline__ __half cvt<float, __half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt<float, __nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ float cvt<__half, float>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float cvt<__nv_bfloat16, float>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ int8_t cvt<int32_t, int8_t>(int32_t v) { return (int8_t)v; }
template <> __device__ __forceinline__ int32_t cvt<int8_t, int32_t>(int8_t v) { return (int32_t)v; }
template <> __device__ __forceinline__ int32_t cvt<float, int32_t>(float v) { return (int32_t)v; }
template <> __device__ __forceinline__ float cvt<int32_t, float>(int32_t v) { return (float)v; }
template <> __device__ __forceinline__ int8_t cvt<float, int8_t>(float v) { return (int8_t)v; }
template <> __device__ __forceinline__ float cvt<int8_t, float>(int8_t v) { return (float)v; }

template <typename S, typename D>
__global__ void __launch_bounds__(256) convert_kernel(const S *__restrict__ src, D *__restrict__ dst, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = cvt<S, D>(src[i]);
}

template <typename S, typename D>
static int run_convert(cuxb_ctx *ctx, cudaStream_t st, long long n, const void *src, void *dst) {
  long long blocks = div_up(n, 256 * 8);
  const long long cap = (long long)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  convert_kernel<S, D><<<(unsigned)blocks, 256, 0, st>>>((const S *)src, (D *)dst, n);
  CUXB_LAUNCHED();
  return 0;
}

// ---- device-side random fill.  Same distribution as cux's DataFiller::RandomFill
// (0-frameworks/cux/include/util/data_filler.h:41-60: an integer in [min, max], plus decimal_pose random
// decimal digits), but counter-based: element i depends only on (seed, i), so the host can reproduce any
// element and the result does not depend on the launch shape.  Array4D::Fill(..., ON_DEVICE) -- which the
// reference rejects ("does not supported for now", src/array.cpp:47-49) -- ends up here.
__host__ __device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
template <typename D> __device__ __forceinline__ D from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_float<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ int32_t from_float<int32_t>(float v) { return (int32_t)v; }
template <> __device__ __forceinline__ int8_t from_float<int8_t>(float v) { return (int8_t)v; }

template <typename D>
__global__ void __launch_bounds__(256) fill_kernel(D *__restrict__ dst, long long n, int lo, unsigned int mod, unsigned int dec,
                                                   unsigned long long seed) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned long long h = splitmix64(seed + (unsigned long long)i * 0xD1342543DE82EF95ull);
    float v = (float)(lo + (int)((unsigned int)(h >> 32) % mod));
    if (dec > 1) v += (float)((unsigned int)h % dec) / (float)dec;
    dst[i] = from_float<D>(v);
  }
}

template <typename D>
static int run_fill(cuxb_ctx *ctx, cudaStream_t st, long long n, void *dst, int lo, unsigned int mod, unsigned int dec,
                    unsigned long long seed) {
  long long blocks = div_up(n, 256 * 8);
  const long long cap = (long long)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  fill_kernel<D><<<(unsigned)blocks, 256, 0, st>>>((D *)dst, n, lo, mod, dec, seed);
  CUXB_LAUNCHED();
  return 0;
}

}  // namespace cuxb

using namespace cuxb;

CUXB_API int cuxb_fill_random(cuxb_ctx *ctx, void *stream, long long n, void *dst, int dtype, int min_value,
                              int max_value, int decimal_pose, unsigned long long seed) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(n >= 0 && (n == 0 || dst != nullptr), CUXB_ERR_INVALID, "cuxb_fill_random: bad argument");
  CUXB_REQUIRE(decimal_pose >= 0 && decimal_pose <= 6, CUXB_ERR_INVALID, "cuxb_fill_random: decimal_pose must be 0..6");
  if (n == 0) return 0;
  if (decimal_pose > 0) max_value--;  // as DataFiller::RandomFill: [min, max) once decimals are added
  CUXB_REQUIRE(max_value >= min_value, CUXB_ERR_INVALID, "cuxb_fill_random: empty range");
  const unsigned int mod = (unsigned int)(max_value - min_value + 1);
  unsigned int dec = 1;
  for (int i = 0; i < decimal_pose; i++) dec *= 10;
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case CUXB_FLOAT32: return run_fill<float>(ctx, st, n, dst, min_value, mod, dec, seed);
    case CUXB_FLOAT16: return run_fill<__half>(ctx, st, n, dst, min_value, mod, dec, seed);
    case CUXB_BFLOAT16: return run_fill<__nv_bfloat16>(ctx, st, n, dst, min_value, mod, dec, seed);
    case CUXB_INT32: return run_fill<int32_t>(ctx, st, n, dst, min_value, mod, dec, seed);
    case CUXB_INT8: return run_fill<int8_t>(ctx, st, n, dst, min_value, mod, dec, seed);
  }
  set_error("cuxb_fill_random: unknown dtype %d", dtype);
  return CUXB_ERR_INVALID;
}

CUXB_API int cuxb_convert(cuxb_ctx *ctx, void *stream, long long n, const void *src, int sd, void *dst, int dd) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(n >= 0, CUXB_ERR_INVALID, "cuxb_convert: negative length");
  if (n == 0) return 0;
  CUXB_REQUIRE(src && dst, CUXB_ERR_INVALID, "cuxb_convert: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
#define CV(SD, DD, S, D) \
  if (sd == SD && dd == DD) return run_convert<S, D>(ctx, st, n, src, dst);
  CV(CUXB_FLOAT32, CUXB_FLOAT16, float, __half)
  CV(CUXB_FLOAT32, CUXB_BFLOAT16, float, __nv_bfloat16)
  CV(CUXB_FLOAT16, CUXB_FLOAT32, __half, float)
  CV(CUXB_BFLOAT16, CUXB_FLOAT32, __nv_bfloat16, float)
  CV(CUXB_INT32, CUXB_INT8, int32_t, int8_t)
  CV(CUXB_INT8, CUXB_INT32, int8_t, int32_t)
  CV(CUXB_FLOAT32, CUXB_INT32, float, int32_t)
  CV(CUXB_INT32, CUXB_FLOAT32, int32_t, float)
  CV(CUXB_FLOAT32, CUXB_INT8, float, int8_t)
  CV(CUXB_INT8, CUXB_FLOAT32, int8_t, float)
#undef CV
  set_error("cuxb_convert: unsupported conversion %d -> %d", sd, dd);
  return CUXB_ERR_UNSUPPORTED;
}
