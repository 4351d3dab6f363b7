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

}  // namespace cuxb

using namespace cuxb;

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
