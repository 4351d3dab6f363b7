// fp32_peak.cu -- measured FP32 peak of the CUDA cores on this GPU, scalar FFMA vs packed FFMA2
// (fma.rn.f32x2, new on sm_100).  Same idea as the reference's MaxFlopsFp32CudaCore
// (pocket-ai/eval/cuda/max_flops.cuh:11-47: independent fma chains timed on the device), sized for a
// B200: every SM gets 8 CTAs of 256 threads, 8 independent accumulators per thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu && ./tools/fp32_peak
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kAcc = 8;

__global__ void __launch_bounds__(256) ffma_chain(float *out, float a, float b) {
  float acc[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; i++) acc[i] = threadIdx.x * 1e-6f + i;
  for (int it = 0; it < kIters; it++)
#pragma unroll
    for (int i = 0; i < kAcc; i++) acc[i] = fmaf(acc[i], a, b);
  float s = 0;
#pragma unroll
  for (int i = 0; i < kAcc; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) ffma2_chain(float *out, float a, float b) {
  float2 acc[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; i++) acc[i] = make_float2(threadIdx.x * 1e-6f + i, i);
  const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
  for (int it = 0; it < kIters; it++)
#pragma unroll
    for (int i = 0; i < kAcc; i++) acc[i] = __ffma2_rn(acc[i], a2, b2);
  float s = 0;
#pragma unroll
  for (int i = 0; i < kAcc; i++) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FFMA2 interleaved with shared-memory loads at the SGEMM ratio (6 LDS.128 per 32 FFMA2)
__global__ void __launch_bounds__(256) ffma2_lds_mix(float *out, float a) {
  __shared__ float4 buf[512];
  for (int i = threadIdx.x; i < 512; i += 256) buf[i] = make_float4(i * 1e-7f, 1e-7f, 2e-7f, 3e-7f);
  __syncthreads();
  float2 acc[32];
#pragma unroll
  for (int i = 0; i < 32; i++) acc[i] = make_float2(i, threadIdx.x);
  const float2 a2 = make_float2(a, a);
  unsigned base = (unsigned)__cvta_generic_to_shared(&buf[(threadIdx.x >> 1) & 7]);
  for (int it = 0; it < kIters / 4; it++) {
    float4 v[6];
#pragma unroll
    for (int j = 0; j < 6; j++)
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v[j].x), "=f"(v[j].y), "=f"(v[j].z), "=f"(v[j].w) : "r"(base + j * 256 + (it & 7) * 16));
#pragma unroll
    for (int i = 0; i < 32; i++) {
      const float4 t = v[i % 6];
      acc[i] = __ffma2_rn(a2, make_float2(t.x, t.y + t.z), acc[i]);
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 32; i++) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int grid = p.multiProcessorCount * 8;
  float *d;
  cudaMalloc(&d, (size_t)grid * 256 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int which = 0; which < 3; which++) {
    float best = 1e9f;
    for (int rep = 0; rep < 8; rep++) {
      cudaEventRecord(e0);
      if (which == 0) ffma_chain<<<grid, 256>>>(d, 1.0001f, 1e-7f);
      else if (which == 1) ffma2_chain<<<grid, 256>>>(d, 1.0001f, 1e-7f);
      else ffma2_lds_mix<<<grid, 256>>>(d, 1.0001f);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep >= 2 && ms < best) best = ms;
    }
    double flop = which == 0 ? 2.0 * kIters * kAcc * grid * 256
                 : which == 1 ? 4.0 * kIters * kAcc * grid * 256
                              : 4.0 * (kIters / 4) * 32 * grid * 256;
    printf("%s: %.3f ms  %.2f TFLOP/s\n", which == 0 ? "FFMA  (scalar)" : which == 1 ? "FFMA2 (packed)" : "FFMA2 + LDS.128 mix", best,
           flop / best * 1e-9);
  }
  printf("%s SMs=%d clock=%d kHz err=%s\n", p.name, p.multiProcessorCount, p.clockRate, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
