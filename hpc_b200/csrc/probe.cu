// probe.cu -- measured ceilings of the GPU the library is running on: HBM read / write / copy streams and
// the FP32 rate of the CUDA cores.  The B200 counterpart of the reference's evaluation helpers
// (pocket-ai/eval/cuda/bandwidth_eval.hpp:257-300: device-to-device bandwidth from timed copies;
// pocket-ai/eval/cuda/max_flops.cuh:11-94: independent fma chains timed on the device), exported so that a
// harness can print its roofline denominators from the same process that ran the kernels.  The standalone
// versions with more variants live in tools/hbm_peak.cu and tools/fp32_peak.cu.
#include "common.cuh"

namespace cuxb {

__global__ void __launch_bounds__(512) probe_read(const float4 *__restrict__ a, size_t n4, float *out) {
  float s = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + 3 * stride < n4; i += 4 * stride) {
    float4 x[4];
#pragma unroll
    for (int u = 0; u < 4; u++)
      asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(x[u].x), "=f"(x[u].y), "=f"(x[u].z), "=f"(x[u].w) : "l"(a + i + u * stride));
    s += x[0].x + x[1].y + x[2].z + x[3].w;
  }
  if (s == 123.456f) out[0] = s;  // never true for the zero-filled buffer: keeps the loads alive
}
__global__ void __launch_bounds__(512) probe_write(float4 *a, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) __stcs(a + i, make_float4(1, 2, 3, 4));
}
__global__ void __launch_bounds__(512) probe_copy(const float4 *__restrict__ a, float4 *b, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + 3 * stride < n4; i += 4 * stride) {
    const float4 x0 = __ldcs(a + i), x1 = __ldcs(a + i + stride), x2 = __ldcs(a + i + 2 * stride), x3 = __ldcs(a + i + 3 * stride);
    __stcs(b + i, x0), __stcs(b + i + stride, x1), __stcs(b + i + 2 * stride, x2), __stcs(b + i + 3 * stride, x3);
  }
}

constexpr int kProbeIters = 4096, kProbeAcc = 8;
template <bool PACKED>
__global__ void __launch_bounds__(256) probe_fma(float *out, float a, float b) {
  float s = 0;
  if (PACKED) {
    float2 acc[kProbeAcc];
#pragma unroll
    for (int i = 0; i < kProbeAcc; i++) acc[i] = make_float2(threadIdx.x * 1e-6f + i, i);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    for (int it = 0; it < kProbeIters; it++)
#pragma unroll
      for (int i = 0; i < kProbeAcc; i++) acc[i] = __ffma2_rn(acc[i], a2, b2);
#pragma unroll
    for (int i = 0; i < kProbeAcc; i++) s += acc[i].x + acc[i].y;
  } else {
    float acc[kProbeAcc];
#pragma unroll
    for (int i = 0; i < kProbeAcc; i++) acc[i] = threadIdx.x * 1e-6f + i;
    for (int it = 0; it < kProbeIters; it++)
#pragma unroll
      for (int i = 0; i < kProbeAcc; i++) acc[i] = fmaf(acc[i], a, b);
#pragma unroll
    for (int i = 0; i < kProbeAcc; i++) s += acc[i];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

struct ProbeBuffers {
  void *a = nullptr, *b = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  ~ProbeBuffers() {
    if (a) cudaFree(a);
    if (b) cudaFree(b);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
  }
};

}  // namespace cuxb

using namespace cuxb;

// which: 0 = read-only stream, 1 = write-only stream, 2 = copy (bytes are counted once per direction, so a
// copy of `bytes` moves 2 x bytes).  Best of `reps` launches after two warm-up launches.
CUXB_API int cuxb_probe_hbm(cuxb_ctx *ctx, int which, size_t bytes, int reps, double *gb_per_s) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(gb_per_s != nullptr && which >= 0 && which <= 2 && reps >= 1 && reps <= 1000, CUXB_ERR_INVALID,
               "cuxb_probe_hbm: bad argument");
  CUXB_REQUIRE(bytes >= (1u << 20) && bytes % 16 == 0, CUXB_ERR_INVALID,
               "cuxb_probe_hbm: bytes must be a multiple of 16 and at least 1 MiB (use several times the L2 size)");
  ProbeBuffers pb;
  CUXB_CUDA(cudaMalloc(&pb.a, bytes));
  CUXB_CUDA(cudaMemset(pb.a, 0, bytes));
  if (which == 2) {
    CUXB_CUDA(cudaMalloc(&pb.b, bytes));
    CUXB_CUDA(cudaMemset(pb.b, 0, bytes));
  }
  CUXB_CUDA(cudaEventCreate(&pb.e0));
  CUXB_CUDA(cudaEventCreate(&pb.e1));
  const size_t n4 = bytes / 16;
  const int grid = ctx->sm_count * 4;
  float best = 1e30f;
  for (int r = 0; r < reps + 2; r++) {
    CUXB_CUDA(cudaEventRecord(pb.e0, ctx->stage_stream));
    if (which == 0) probe_read<<<grid, 512, 0, ctx->stage_stream>>>((const float4 *)pb.a, n4, (float *)ctx->red_partials);
    else if (which == 1) probe_write<<<grid, 512, 0, ctx->stage_stream>>>((float4 *)pb.a, n4);
    else probe_copy<<<grid, 512, 0, ctx->stage_stream>>>((const float4 *)pb.a, (float4 *)pb.b, n4);
    CUXB_LAUNCHED();
    CUXB_CUDA(cudaEventRecord(pb.e1, ctx->stage_stream));
    CUXB_CUDA(cudaEventSynchronize(pb.e1));
    float ms = 0;
    CUXB_CUDA(cudaEventElapsedTime(&ms, pb.e0, pb.e1));
    if (r >= 2 && ms < best) best = ms;
  }
  *gb_per_s = (which == 2 ? 2.0 : 1.0) * (double)bytes / ((double)best * 1e-3) / 1e9;
  return 0;
}

// packed == 0: scalar FFMA chains, packed != 0: FFMA2 (fma.rn.f32x2).  8 CTAs of 256 threads per SM, eight
// independent accumulators per thread.
CUXB_API int cuxb_probe_fp32(cuxb_ctx *ctx, int packed, int reps, double *tflops) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(tflops != nullptr && reps >= 1 && reps <= 1000, CUXB_ERR_INVALID, "cuxb_probe_fp32: bad argument");
  ProbeBuffers pb;
  const int grid = ctx->sm_count * 8;
  CUXB_CUDA(cudaMalloc(&pb.a, (size_t)grid * 256 * sizeof(float)));
  CUXB_CUDA(cudaEventCreate(&pb.e0));
  CUXB_CUDA(cudaEventCreate(&pb.e1));
  float best = 1e30f;
  for (int r = 0; r < reps + 2; r++) {
    CUXB_CUDA(cudaEventRecord(pb.e0, ctx->stage_stream));
    if (packed) probe_fma<true><<<grid, 256, 0, ctx->stage_stream>>>((float *)pb.a, 1.0001f, 1e-7f);
    else probe_fma<false><<<grid, 256, 0, ctx->stage_stream>>>((float *)pb.a, 1.0001f, 1e-7f);
    CUXB_LAUNCHED();
    CUXB_CUDA(cudaEventRecord(pb.e1, ctx->stage_stream));
    CUXB_CUDA(cudaEventSynchronize(pb.e1));
    float ms = 0;
    CUXB_CUDA(cudaEventElapsedTime(&ms, pb.e0, pb.e1));
    if (r >= 2 && ms < best) best = ms;
  }
  const double flop = (packed ? 4.0 : 2.0) * kProbeIters * kProbeAcc * (double)grid * 256;
  *tflops = flop / ((double)best * 1e-3) / 1e12;
  return 0;
}
