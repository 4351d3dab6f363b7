// hbm_peak.cu -- measured HBM ceilings on this GPU: read-only, write-only and copy streams, the
// B200 counterpart of the reference's device-to-device bandwidth test
// (pocket-ai/eval/cuda/bandwidth_eval.hpp:257-300, which times cudaMemcpy D2D and counts bytes twice).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/hbm_peak tools/hbm_peak.cu && ./tools/hbm_peak
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(512) k_read(const float4 *__restrict__ a, size_t n4, float *out) {
  float s = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + 3 * stride < n4; i += 4 * stride) {
    float4 x0, x1, x2, x3;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x0.x), "=f"(x0.y), "=f"(x0.z), "=f"(x0.w) : "l"(a + i));
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x1.x), "=f"(x1.y), "=f"(x1.z), "=f"(x1.w) : "l"(a + i + stride));
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x2.x), "=f"(x2.y), "=f"(x2.z), "=f"(x2.w) : "l"(a + i + 2 * stride));
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x3.x), "=f"(x3.y), "=f"(x3.z), "=f"(x3.w) : "l"(a + i + 3 * stride));
    s += x0.x + x1.y + x2.z + x3.w;
  }
  if (s == 123.456f) out[0] = s;
}
__global__ void __launch_bounds__(512) k_write(float4 *a, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) __stcs(a + i, make_float4(1, 2, 3, 4));
}
__global__ void __launch_bounds__(512) k_copy(const float4 *__restrict__ a, float4 *b, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + 3 * stride < n4; i += 4 * stride) {
    const float4 x0 = __ldcs(a + i), x1 = __ldcs(a + i + stride), x2 = __ldcs(a + i + 2 * stride), x3 = __ldcs(a + i + 3 * stride);
    __stcs(b + i, x0), __stcs(b + i + stride, x1), __stcs(b + i + 2 * stride, x2), __stcs(b + i + 3 * stride, x3);
  }
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const size_t bytes = 2ull << 30, n4 = bytes / 16;  // 2 GiB per buffer, far above the 126 MB L2
  float4 *a, *b;
  float *out;
  cudaMalloc(&a, bytes), cudaMalloc(&b, bytes), cudaMalloc(&out, 4);
  cudaMemset(a, 0, bytes), cudaMemset(b, 0, bytes);
  const int grid = p.multiProcessorCount * 4;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  for (int which = 0; which < 4; which++) {
    float best = 1e9f;
    for (int rep = 0; rep < 8; rep++) {
      cudaEventRecord(e0);
      if (which == 0) k_read<<<grid, 512>>>(a, n4, out);
      else if (which == 1) k_write<<<grid, 512>>>(b, n4);
      else if (which == 2) k_copy<<<grid, 512>>>(a, b, n4);
      else cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep >= 2 && ms < best) best = ms;
    }
    const double moved = (which >= 2 ? 2.0 : 1.0) * bytes;
    printf("%-22s %.3f ms  %.0f GB/s\n", which == 0 ? "read  (2 GiB)" : which == 1 ? "write (2 GiB)" : which == 2 ? "copy  (2+2 GiB, kernel)" : "copy  (2+2 GiB, cudaMemcpy)", best, moved / best * 1e-6);
  }
  printf("%s err=%s\n", p.name, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
