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

// copy through shared memory with 1-D bulk copies (TMA): one CTA per SM, a ring of kSlots x kChunk bytes,
// one thread issues cp.async.bulk global->shared (mbarrier) and shared->global (bulk group).
template <int kChunk, int kSlots>
__global__ void __launch_bounds__(32) k_copy_bulk(const char *__restrict__ a, char *b, size_t bytes) {
  extern __shared__ __align__(128) char ring[];
  __shared__ __align__(8) unsigned long long bar[kSlots];
  const size_t nchunks = bytes / kChunk;
  if (threadIdx.x != 0) return;
  for (int s = 0; s < kSlots; s++)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&bar[s])));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);
  auto load = [&](size_t c, int s) {
    const unsigned bs = (unsigned)__cvta_generic_to_shared(&bar[s]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bs), "r"(kChunk) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(ring_s + s * kChunk), "l"(a + c * kChunk), "r"(kChunk), "r"(bs) : "memory");
  };
  size_t c = blockIdx.x, issued = 0, done = 0;
  // prologue: fill the ring except one slot (a slot is reloaded only after its store has been read out)
  size_t cl = c;
  for (; issued < kSlots - 1 && cl < nchunks; issued++, cl += gridDim.x) load(cl, (int)(issued % kSlots));
  for (; c < nchunks; c += gridDim.x, done++) {
    const int s = (int)(done % kSlots);
    const unsigned bs = (unsigned)__cvta_generic_to_shared(&bar[s]);
    const unsigned parity = (unsigned)((done / kSlots) & 1);
    unsigned ok = 0;
    while (!ok)
      asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}"
                   : "=r"(ok) : "r"(bs), "r"(parity) : "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(b + c * kChunk), "r"(ring_s + s * kChunk), "r"(kChunk) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (cl < nchunks) {
      // the slot about to be refilled was stored (kSlots - 1) iterations ago: wait until that store has read it
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kSlots - 2) : "memory");
      load(cl, (int)(issued % kSlots));
      issued++, cl += gridDim.x;
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// mixed variants, 512 threads + ring of 6 x 32 KiB, one CTA per SM (the scan kernel's shape):
//   MODE 0: bulk load (TMA) -> LDS.128 -> STG.128        MODE 1: cp.async 16 B loads -> bulk store (TMA)
template <int MODE>
__global__ void __launch_bounds__(512, 1) k_copy_mixed(const char *__restrict__ a, char *b, size_t bytes) {
  constexpr int kChunk = 32768, kSlots = 6, kAhead = 3;
  extern __shared__ __align__(128) char ring[];
  __shared__ __align__(8) unsigned long long bar[kSlots];
  const size_t nchunks = bytes / kChunk;
  const int tid = threadIdx.x;
  const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);
  if (tid == 0) {
    for (int s = 0; s < kSlots; s++)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&bar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto load = [&](size_t c, int s) {
    if (MODE == 0) {
      if (tid == 0) {
        const unsigned bs = (unsigned)__cvta_generic_to_shared(&bar[s]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bs), "r"(kChunk) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(ring_s + s * kChunk), "l"(a + c * kChunk), "r"(kChunk), "r"(bs) : "memory");
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; j++)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_s + s * kChunk + (j * 512 + tid) * 16),
                     "l"(a + c * kChunk + (size_t)(j * 512 + tid) * 16) : "memory");
    }
  };
  size_t cl = blockIdx.x;
  int issued = 0;
  for (; issued < kAhead; issued++, cl += gridDim.x) {
    if (cl < nchunks) load(cl, issued % kSlots);
    if (MODE == 1) asm volatile("cp.async.commit_group;" ::: "memory");
  }
  int done = 0;
  for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x, done++) {
    const int s = done % kSlots;
    // refill: the slot being loaded now was consumed kSlots - kAhead iterations ago
    if (MODE == 1 && tid == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kSlots - kAhead - 1) : "memory");
    if (MODE == 1) __syncthreads();
    if (cl < nchunks) load(cl, issued % kSlots);
    if (MODE == 1) asm volatile("cp.async.commit_group;" ::: "memory");
    issued++, cl += gridDim.x;
    if (MODE == 0) {
      const unsigned bs = (unsigned)__cvta_generic_to_shared(&bar[s]);
      const unsigned parity = (unsigned)((done / kSlots) & 1);
      unsigned ok = 0;
      while (!ok)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}"
                     : "=r"(ok) : "r"(bs), "r"(parity) : "memory");
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const float4 v = *reinterpret_cast<const float4 *>(ring + s * kChunk + (j * 512 + tid) * 16);
        __stcs(reinterpret_cast<float4 *>(b + c * kChunk + (size_t)(j * 512 + tid) * 16), v);
      }
      __syncthreads();  // everybody has read the slot before thread 0 may refill it (kSlots - kAhead passes later)
    } else {
      asm volatile("cp.async.wait_group %0;" ::"n"(kAhead) : "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (tid == 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(b + c * kChunk), "r"(ring_s + s * kChunk), "r"(kChunk) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }
  if (MODE == 1 && tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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
  cudaFuncSetAttribute(k_copy_bulk<32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6);
  cudaFuncSetAttribute(k_copy_bulk<16384, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 12);
  cudaFuncSetAttribute(k_copy_mixed<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6);
  cudaFuncSetAttribute(k_copy_mixed<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6);
  for (int which = 0; which < 8; which++) {
    float best = 1e9f;
    for (int rep = 0; rep < 8; rep++) {
      cudaEventRecord(e0);
      if (which == 0) k_read<<<grid, 512>>>(a, n4, out);
      else if (which == 1) k_write<<<grid, 512>>>(b, n4);
      else if (which == 2) k_copy<<<grid, 512>>>(a, b, n4);
      else if (which == 3) cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice);
      else if (which == 4) k_copy_bulk<32768, 6><<<p.multiProcessorCount, 32, 32768 * 6>>>((const char *)a, (char *)b, bytes);
      else if (which == 5) k_copy_bulk<16384, 12><<<p.multiProcessorCount, 32, 16384 * 12>>>((const char *)a, (char *)b, bytes);
      else if (which == 6) k_copy_mixed<0><<<p.multiProcessorCount, 512, 32768 * 6>>>((const char *)a, (char *)b, bytes);
      else k_copy_mixed<1><<<p.multiProcessorCount, 512, 32768 * 6>>>((const char *)a, (char *)b, bytes);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep >= 2 && ms < best) best = ms;
    }
    const double moved = (which >= 2 ? 2.0 : 1.0) * bytes;
    printf("%-22s %.3f ms  %.0f GB/s\n", which == 0 ? "read  (2 GiB)" : which == 1 ? "write (2 GiB)" : which == 2 ? "copy  (2+2 GiB, kernel)" : which == 3 ? "copy  (2+2 GiB, cudaMemcpy)" : which == 4 ? "copy  (bulk 32K x 6)" : which == 5 ? "copy  (bulk 16K x 12)" : which == 6 ? "copy  (bulk load, STG)" : "copy  (cp.async, bulk store)", best, moved / best * 1e-6);
  }
  printf("%s err=%s\n", p.name, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
