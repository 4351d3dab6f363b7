// lds_patterns.cu -- micro-benchmark: how many shared-memory wavefronts does one warp-wide LDS.128
// cost for a given lane -> 16-byte-chunk mapping?  (Used to pick the thread layout of sgemm.cu.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/lds_patterns tools/lds_patterns.cu
//   ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum ./tools/lds_patterns
#include <cstdio>
#include <cuda_runtime.h>

template <int P> __device__ __forceinline__ int chunk_of(int lane) {
  switch (P) {
    case 0: return lane >> 2;                               // 8 distinct, quads of lanes share
    case 1: return lane & 3;                                // 4 distinct, period 4
    case 2: return lane >> 3;                               // 4 distinct, quarter-warp uniform
    case 3: return lane & 7;                                // 8 distinct, period 8
    case 4: return (lane >> 1) & 7;                         // 8 distinct, pairs share, period 16
    case 5: return ((lane & 16) >> 3) | (lane & 1);         // 4 distinct
    case 6: return lane;                                    // 32 distinct
    case 7: return 0;                                       // all same
    case 8: return lane >> 4;                               // 2 distinct, half-warp uniform
    case 9: return (lane >> 2) & 3;                         // 4 distinct, quads share, period 16
    case 10: return (lane >> 1) & 3;                        // 4 distinct, pairs share, period 8
    case 11: return lane & 1;                               // 2 distinct alternating
    case 12: return lane & 15;                              // 16 distinct, period 16
    case 13: return lane >> 1;                              // 16 distinct, pairs share
    case 14: return ((lane >> 3) << 1) | (lane & 1);        // 8 distinct: quarter id, lane bit0
    case 15: return (lane & 3) * 2;                         // 4 distinct, strided by 32 B, period 4
    case 16: return (lane >> 2) * 2;                        // 8 distinct strided by 32 B (256 B span)
    default: return 0;
  }
}

template <int P> __global__ void lds_pat(float *out, int iters) {
  __shared__ float4 buf[128];
  for (int i = threadIdx.x; i < 128; i += 32) buf[i] = make_float4(i, 1, 2, 3);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  unsigned base = (unsigned)__cvta_generic_to_shared(&buf[chunk_of<P>(lane)]);
  float4 acc = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; it++) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(base + (it & 1) * 1024));
    acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
  }
  out[threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
}

template <int P> void run(float *d) {
  lds_pat<P><<<1, 32>>>(d, 1000);
  cudaDeviceSynchronize();
}

int main() {
  float *d;
  cudaMalloc(&d, 4096);
  run<0>(d); run<1>(d); run<2>(d); run<3>(d); run<4>(d); run<5>(d); run<6>(d); run<7>(d); run<8>(d);
  run<9>(d); run<10>(d); run<11>(d); run<12>(d); run<13>(d); run<14>(d); run<15>(d); run<16>(d);
  printf("done %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
