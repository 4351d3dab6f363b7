// sgemm_inner.cu -- how fast can the SGEMM micro-kernel go when nothing else is in the way?
// The inner product of hpc_b200/csrc/sgemm.cu (8x8 micro-tile as 4x8 FFMA2, 4 LDS.128 per k, fragments
// double-buffered) run from shared memory that is never refilled, in three variants:
//   0: no barrier            -- the LDS/FFMA2 schedule alone
//   1: + __syncthreads per k-tile of 16
//   2: + the 8 STS of the A transpose and a cp.async wait per k-tile (the shared-memory side of the refill)
//   3: variant 0 with the A fragment fetched as 4 LDS.64 instead of 2 LDS.128 (is the cost per
//      instruction or per register written?)
//   5: 16x8 micro-tile (CTA tile 256x128, 1 CTA per SM, up to 255 registers): 6 LDS.128 per 64 FFMA2
//      instead of 8; barrier per k-tile
//   4: A kept row-major in shared memory ([m][k], what cp.async could fill without a transpose), one
//      LDS.128 = 4 k of one row, accumulator pairs along n; barrier per k-tile, no STS
// Same launch shape as the real kernel: 256 threads, 2 CTAs per SM, <= 128 registers.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/sgemm_inner tools/sgemm_inner.cu && ./tools/sgemm_inner
#include <cstdio>
#include <cuda_runtime.h>

constexpr int BM = 128, BN = 128, BK = 16, LDA = BM + 4;

// variant 4
__global__ void __launch_bounds__(256, 2) inner_rowmajor_a(float *out, int ktiles) {
  constexpr int LDK = BK + 4;  // 80-byte rows: 8 consecutive rows land on 8 different 16-byte bank groups
  __shared__ __align__(16) float As[2][BM][LDK];
  __shared__ __align__(16) float Bs[2][BK][BN];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 2 * BM * LDK; i += 256) (&As[0][0][0])[i] = 1e-3f * (i & 7);
  for (int i = tid; i < 2 * BK * BN; i += 256) (&Bs[0][0][0])[i] = 1e-3f * (i & 3);
  __syncthreads();
  const int warp_m = warp >> 2, warp_n = warp & 3;
  const int lane_m = (lane >> 1) & 7, lane_n = ((lane & 16) >> 3) | (lane & 1);
  const int m0 = warp_m * 64 + lane_m;  // rows m0 + 8 r + 32 h
  const int tn0 = warp_n * 32 + lane_n * 4;
  float2 acc2[8][4];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc2[i][j] = make_float2(0.f, 0.f);
  for (int kt = 0; kt < ktiles; kt++) {
    const int buf = kt & 1;
#pragma unroll
    for (int g = 0; g < BK / 4; g++) {
      float4 a[8];
#pragma unroll
      for (int i = 0; i < 8; i++) a[i] = *reinterpret_cast<const float4 *>(&As[buf][m0 + 8 * (i & 3) + 32 * (i >> 2)][g * 4]);
#pragma unroll
      for (int kk = 0; kk < 4; kk++) {
        const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][g * 4 + kk][tn0]);
        const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][g * 4 + kk][tn0 + 16]);
        const float2 bp[4] = {make_float2(b0.x, b0.y), make_float2(b0.z, b0.w), make_float2(b1.x, b1.y), make_float2(b1.z, b1.w)};
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const float av = kk == 0 ? a[i].x : kk == 1 ? a[i].y : kk == 2 ? a[i].z : a[i].w;
          const float2 aa = make_float2(av, av);
#pragma unroll
          for (int j = 0; j < 4; j++) acc2[i][j] = __ffma2_rn(aa, bp[j], acc2[i][j]);
        }
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) s += acc2[i][j].x + acc2[i][j].y;
  out[blockIdx.x * 256 + tid] = s;
}

// variant 5
__global__ void __launch_bounds__(256, 1) inner_16x8(float *out, int ktiles) {
  constexpr int BM2 = 256, LDA2 = BM2;  // (no pad: static shared memory stops at 48 KB)
  __shared__ __align__(16) float As[2][BK][LDA2];
  __shared__ __align__(16) float Bs[2][BK][BN];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 2 * BK * LDA2; i += 256) (&As[0][0][0])[i] = 1e-3f * (i & 7);
  for (int i = tid; i < 2 * BK * BN; i += 256) (&Bs[0][0][0])[i] = 1e-3f * (i & 3);
  __syncthreads();
  const int warp_m = warp >> 2, warp_n = warp & 3;   // warp tile 128 x 32
  const int lane_m = (lane >> 1) & 7, lane_n = ((lane & 16) >> 3) | (lane & 1);
  const int tm0 = warp_m * 128 + lane_m * 4, tn0 = warp_n * 32 + lane_n * 4;  // rows tm0 + 32 q + {0..3}, q < 4
  float2 acc2[8][8];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc2[i][j] = make_float2(0.f, 0.f);
  float2 af[2][8];
  float bf[2][8];
  auto lfrag = [&](int buf, int k, int s) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const float4 a = *reinterpret_cast<const float4 *>(&As[buf][k][tm0 + 32 * q]);
      af[s][2 * q] = make_float2(a.x, a.y), af[s][2 * q + 1] = make_float2(a.z, a.w);
    }
    const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tn0]);
    const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tn0 + 16]);
    bf[s][0] = b0.x, bf[s][1] = b0.y, bf[s][2] = b0.z, bf[s][3] = b0.w;
    bf[s][4] = b1.x, bf[s][5] = b1.y, bf[s][6] = b1.z, bf[s][7] = b1.w;
  };
  for (int kt = 0; kt < ktiles; kt++) {
    const int buf = kt & 1;
    lfrag(buf, 0, 0);
#pragma unroll
    for (int k = 0; k < BK; k++) {
      if (k + 1 < BK) lfrag(buf, k + 1, (k + 1) & 1);
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const float2 bb = make_float2(bf[k & 1][j], bf[k & 1][j]);
#pragma unroll
        for (int ip = 0; ip < 8; ip++) acc2[ip][j] = __ffma2_rn(af[k & 1][ip], bb, acc2[ip][j]);
      }
    }
    __syncthreads();
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) s += acc2[i][j].x + acc2[i][j].y;
  out[blockIdx.x * 256 + tid] = s;
}

template <int VARIANT>
__global__ void __launch_bounds__(256, 2) inner(float *out, int ktiles) {
  __shared__ __align__(16) float As[2][BK][LDA];
  __shared__ __align__(16) float Bs[2][BK][BN];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 2 * BK * LDA; i += 256) (&As[0][0][0])[i] = 1e-3f * (i & 7);
  for (int i = tid; i < 2 * BK * BN; i += 256) (&Bs[0][0][0])[i] = 1e-3f * (i & 3);
  __syncthreads();
  const int warp_m = warp >> 2, warp_n = warp & 3;
  const int lane_m = (lane >> 1) & 7, lane_n = ((lane & 16) >> 3) | (lane & 1);
  const int tm0 = warp_m * 64 + lane_m * 4, tn0 = warp_n * 32 + lane_n * 4;
  const int a_row = tid >> 2, a_k4 = (tid & 3) * 4;

  float2 acc2[4][8];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc2[i][j] = make_float2(0.f, 0.f);
  float2 af[2][4];
  float bf[2][8];
  auto lfrag = [&](int buf, int k, int s) {
    float4 a0, a1;
    if (VARIANT == 3) {
      const float2 p0 = *reinterpret_cast<const float2 *>(&As[buf][k][tm0]);
      const float2 p1 = *reinterpret_cast<const float2 *>(&As[buf][k][tm0 + 2]);
      const float2 p2 = *reinterpret_cast<const float2 *>(&As[buf][k][tm0 + 32]);
      const float2 p3 = *reinterpret_cast<const float2 *>(&As[buf][k][tm0 + 34]);
      a0 = make_float4(p0.x, p0.y, p1.x, p1.y), a1 = make_float4(p2.x, p2.y, p3.x, p3.y);
    } else {
      a0 = *reinterpret_cast<const float4 *>(&As[buf][k][tm0]);
      a1 = *reinterpret_cast<const float4 *>(&As[buf][k][tm0 + 32]);
    }
    const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tn0]);
    const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tn0 + 16]);
    af[s][0] = make_float2(a0.x, a0.y), af[s][1] = make_float2(a0.z, a0.w);
    af[s][2] = make_float2(a1.x, a1.y), af[s][3] = make_float2(a1.z, a1.w);
    bf[s][0] = b0.x, bf[s][1] = b0.y, bf[s][2] = b0.z, bf[s][3] = b0.w;
    bf[s][4] = b1.x, bf[s][5] = b1.y, bf[s][6] = b1.z, bf[s][7] = b1.w;
  };
  float4 ra[2] = {make_float4(1e-3f, 2e-3f, 3e-3f, 4e-3f), make_float4(1e-3f, 2e-3f, 3e-3f, 4e-3f)};
  for (int kt = 0; kt < ktiles; kt++) {
    const int buf = kt & 1;
    lfrag(buf, 0, 0);
#pragma unroll
    for (int k = 0; k < BK; k++) {
      if (k + 1 < BK) lfrag(buf, k + 1, (k + 1) & 1);
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const float2 bb = make_float2(bf[k & 1][j], bf[k & 1][j]);
#pragma unroll
        for (int ip = 0; ip < 4; ip++) acc2[ip][j] = __ffma2_rn(af[k & 1][ip], bb, acc2[ip][j]);
      }
    }
    if (VARIANT >= 2) {
#pragma unroll
      for (int i = 0; i < 2; i++) {
        const int m = a_row + 64 * i;
        As[buf ^ 1][a_k4 + 0][m] = ra[i].x;
        As[buf ^ 1][a_k4 + 1][m] = ra[i].y;
        As[buf ^ 1][a_k4 + 2][m] = ra[i].z;
        As[buf ^ 1][a_k4 + 3][m] = ra[i].w;
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    if (VARIANT >= 1) __syncthreads();
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) s += acc2[i][j].x + acc2[i][j].y;
  out[blockIdx.x * 256 + tid] = s;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int grid = p.multiProcessorCount * 2, ktiles = 4096;
  float *d;
  cudaMalloc(&d, (size_t)grid * 256 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int v = 0; v < 6; v++) {
    float best = 1e9f;
    for (int rep = 0; rep < 6; rep++) {
      cudaEventRecord(e0);
      if (v == 0) inner<0><<<grid, 256>>>(d, ktiles);
      else if (v == 1) inner<1><<<grid, 256>>>(d, ktiles);
      else if (v == 2) inner<2><<<grid, 256>>>(d, ktiles);
      else if (v == 3) inner<3><<<grid, 256>>>(d, ktiles);
      else if (v == 4) inner_rowmajor_a<<<grid, 256>>>(d, ktiles);
      else inner_16x8<<<grid / 2, 256>>>(d, ktiles);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep >= 2 && ms < best) best = ms;
    }
    const double flop = 2.0 * BM * BN * BK * (double)ktiles * grid;  // variant 5: half the CTAs, twice the tile
    printf("variant %d: %.3f ms  %.2f TFLOP/s\n", v, best, flop / best * 1e-9);
  }
  printf("%s SMs=%d err=%s\n", p.name, p.multiProcessorCount, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
