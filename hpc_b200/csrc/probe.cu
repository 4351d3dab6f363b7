// probe.cu -- measured ceilings of the GPU the library is running on: HBM read / write / copy streams and
// the FP32 rate of the CUDA cores.  The B200 counterpart of the reference's evaluation helpers
// (pocket-ai/eval/cuda/bandwidth_eval.hpp:257-300: device-to-device bandwidth from timed copies;
// pocket-ai/eval/cuda/max_flops.cuh:11-94: independent fma chains timed on the device), exported so that a
// harness can print its roofline denominators from the same process that ran the kernels.  The standalone
// versions with more variants live in tools/hbm_peak.cu and tools/fp32_peak.cu.
#include "common.cuh"
#include "umma_ptx.cuh"

namespace cuxb {

__global__ void __launch_bounds__(512) probe_read(const float4 *__restrict__ a, size_t n4, float *out) {
  float s = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {
    float4 x[4];
#pragma unroll
    for (int u = 0; u < 4; u++)
      asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(x[u].x), "=f"(x[u].y), "=f"(x[u].z), "=f"(x[u].w) : "l"(a + i + u * stride));
    s += x[0].x + x[1].y + x[2].z + x[3].w;
  }
  // the last (up to three) vectors per thread, so that every byte that is counted is also read
  for (; i < n4; i += stride) {
    float4 x;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "l"(a + i));
    s += x.x + x.w;
  }
  if (s == 123.456f) out[0] = s;  // never true for the zero-filled buffer: keeps the loads alive
}
__global__ void __launch_bounds__(512) probe_write(float4 *a, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) __stcs(a + i, make_float4(1, 2, 3, 4));
}
__global__ void __launch_bounds__(512) probe_copy(const float4 *__restrict__ a, float4 *b, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {
    const float4 x0 = __ldcs(a + i), x1 = __ldcs(a + i + stride), x2 = __ldcs(a + i + 2 * stride), x3 = __ldcs(a + i + 3 * stride);
    __stcs(b + i, x0), __stcs(b + i + stride, x1), __stcs(b + i + 2 * stride, x2), __stcs(b + i + 3 * stride, x3);
  }
  for (; i < n4; i += stride)
    __stcs(b + i, __ldcs(a + i));
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


// ---- tensor-pipe ceiling: tcgen05.mma 256 x 256 x (16 | 32) issued back to back by CTA pairs on operands
// that never leave shared memory -- the GEMM kernel's mainloop with its TMA feed taken away.  What this
// measures is the rate the tensor pipe sustains when nothing else can be the limit, on the same
// instruction shape the product kernel uses (cta_group::2, 128-byte swizzled K-major tiles).
constexpr int kTpMmaPerChunk = 256;   // one chunk = a k-loop of 4096 (f16) / 8192 (int8) elements
constexpr int kTpThreads = 128;
constexpr int kTpTileBytes = 128 * 128;  // 128 rows of 128 bytes: this CTA's half of A, and of B

template <int KIND>
__global__ void __launch_bounds__(kTpThreads, 1) probe_tensor_kernel(int chunks, unsigned int fill_seed) {
  extern __shared__ uint8_t tp_raw[];
  const uint32_t base = (smem_u32(tp_raw) + 1023u) & ~1023u;
  uint8_t *gen = tp_raw + (base - smem_u32(tp_raw));
  const uint32_t sa = base, sb = base + kTpTileBytes;
  const uint32_t bar = base + 2 * kTpTileBytes, tmem_slot = bar + 16;
  const int warp = threadIdx.x >> 5;
  const bool is_leader = cluster_ctarank() == 0;
  // operand fill: zeros (seed 0) or hashed values -- fp16 / bf16 in (-1, 1), int8 over the whole range.  The
  // data decides the power the tensor pipe draws, and on a power-capped part the clock.
  for (int i = threadIdx.x; i < 2 * kTpTileBytes / 4; i += kTpThreads) {
    uint32_t w = 0;
    if (fill_seed) {
      uint32_t h = (uint32_t)i * 2654435761u ^ fill_seed ^ (blockIdx.x * 40503u);
      h ^= h >> 15, h *= 2246822519u, h ^= h >> 13;
      if (KIND == KIND_S8) w = h;
      else if (KIND == KIND_F16) w = (h & 0x83ff83ffu) | 0x38003800u;   // sign, mantissa, exponent of [0.5, 1)
      else w = (h & 0x807f807fu) | 0x3f003f00u;
    }
    reinterpret_cast<uint32_t *>(gen)[i] = w;
  }
  fence_proxy_async();
  if (warp == 0 && elect_one()) {
    mbar_init(bar, 1), mbar_init(bar + 8, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<2>(tmem_slot, 256);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(gen + (tmem_slot - base));
  if (warp == 1 && is_leader && elect_one()) {
    constexpr uint32_t idesc = make_idesc<KIND, false>(256, 256);
    for (int c = 0; c < chunks; c++) {
      if (c >= 2) mbar_wait(bar + 8 * (c & 1), (uint32_t)(((c - 2) >> 1) & 1));   // chunk c-2 has retired
      for (int m = 0; m < kTpMmaPerChunk; m++) {
        const int k = m & 3;
        umma<2, KIND>(tmem_base, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                      m != 0);
      }
      umma_commit<2>(bar + 8 * (c & 1));
    }
    for (int c = (chunks >= 2 ? chunks - 2 : 0); c < chunks; c++) mbar_wait(bar + 8 * (c & 1), (uint32_t)((c >> 1) & 1));
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc<2>(tmem_base, 256);
}

// ---- NVLink ceiling between two GPUs, with the data moved the ways the multi-GPU kernels move it.
// mode 0: kernel pull  (ld from the peer, st local; 16 x 16-byte loads in flight per thread -- the SGEMM copy CTAs)
// mode 1: kernel push  (ld local, st to the peer)
// mode 3: bulk-copy pull, one thread per SM: cp.async.bulk peer -> shared -> local through a ring of `depth`
//         16 KiB buffers (depth 2 is what the tensor GEMM's broadcast warp keeps in flight)
__global__ void __launch_bounds__(256) probe_link_copy(const float4 *__restrict__ src, float4 *__restrict__ dst, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 15 * stride < n4; i += 16 * stride) {
    float4 v[16];
#pragma unroll
    for (int u = 0; u < 16; u++)
      asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(src + i + u * stride) : "memory");
#pragma unroll
    for (int u = 0; u < 16; u++) dst[i + u * stride] = v[u];
  }
  for (; i < n4; i += stride) dst[i] = src[i];
}

constexpr int kLinkBox = 16384;
__global__ void __launch_bounds__(32) probe_link_bulk(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst,
                                                      size_t boxes, int depth) {
  extern __shared__ __align__(128) uint8_t lk_raw[];
  const uint32_t base = (smem_u32(lk_raw) + 127u) & ~127u;
  const uint32_t bars = base + (uint32_t)depth * kLinkBox;
  if (threadIdx.x != 0) return;
  for (int d = 0; d < depth; d++) mbar_init(bars + 8 * d, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const size_t G = gridDim.x, me = blockIdx.x;
  const size_t mine = boxes > me ? (boxes - me + G - 1) / G : 0;
  auto issue = [&](size_t i) {
    const int d = (int)(i % (size_t)depth);
    mbar_expect_tx(bars + 8 * d, kLinkBox);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(base + (uint32_t)d * kLinkBox), "l"(src + (me + i * G) * kLinkBox), "r"(kLinkBox), "r"(bars + 8 * d) : "memory");
  };
  for (size_t i = 0; i < mine && i < (size_t)depth; i++) issue(i);
  for (size_t i = 0; i < mine; i++) {
    const int d = (int)(i % (size_t)depth);
    mbar_wait(bars + 8 * d, (uint32_t)((i / (size_t)depth) & 1));
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(dst + (me + i * G) * kLinkBox), "r"(base + (uint32_t)d * kLinkBox), "r"(kLinkBox) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (i + depth < mine) {
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the buffer has been read out: refill it
      issue(i + depth);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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
  CUXB_REQUIRE(bytes >= (64u << 20) && bytes % 16 == 0, CUXB_ERR_INVALID,
               "cuxb_probe_hbm: bytes must be a multiple of 16 and at least 64 MiB (use several times the 126 MB L2)");
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

// kind: CUXB_FLOAT16, CUXB_BFLOAT16 or CUXB_INT8 operands (fp32 / int32 accumulate).  data: 0 = all-zero
// operands (the least power the pipe can draw), 1 = pseudo-random operands.  *best = the fastest single
// launch (burst), *mean = all `reps` launches timed back to back as one region (with enough reps -- each
// launch is about 4 ms -- that is the sustained rate under the board's power limit).
CUXB_API int cuxb_probe_tensor(cuxb_ctx *ctx, int kind, int data, int reps, double *best_tflops, double *mean_tflops) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(best_tflops != nullptr && mean_tflops != nullptr && reps >= 1 && reps <= 100000, CUXB_ERR_INVALID,
               "cuxb_probe_tensor: bad argument");
  CUXB_REQUIRE(kind == CUXB_FLOAT16 || kind == CUXB_BFLOAT16 || kind == CUXB_INT8, CUXB_ERR_INVALID,
               "cuxb_probe_tensor: kind must be CUXB_FLOAT16, CUXB_BFLOAT16 or CUXB_INT8");
  ProbeBuffers pb;
  CUXB_CUDA(cudaEventCreate(&pb.e0));
  CUXB_CUDA(cudaEventCreate(&pb.e1));
  const int clusters = ctx->sm_count / 2, chunks = 200;
  const int smem = 2 * kTpTileBytes + 1024 + 256;
  void (*kern)(int, unsigned int) = kind == CUXB_INT8 ? probe_tensor_kernel<KIND_S8>
                                    : kind == CUXB_FLOAT16 ? probe_tensor_kernel<KIND_F16> : probe_tensor_kernel<KIND_BF16>;
  CUXB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * 2, 1, 1);
  cfg.blockDim = dim3(kTpThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = ctx->stage_stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  const unsigned int seed = data ? 0x9e3779b9u : 0u;
  const double ops = 2.0 * 256 * 256 * (kind == CUXB_INT8 ? 32 : 16) * (double)kTpMmaPerChunk * chunks * clusters;
  float best = 1e30f;
  for (int r = 0; r < 3; r++) {   // warm-up, then single launches for the burst figure
    CUXB_CUDA(cudaEventRecord(pb.e0, ctx->stage_stream));
    CUXB_CUDA(cudaLaunchKernelEx(&cfg, kern, chunks, seed));
    g_launches.fetch_add(1);
    CUXB_CUDA(cudaEventRecord(pb.e1, ctx->stage_stream));
    CUXB_CUDA(cudaEventSynchronize(pb.e1));
    float ms = 0;
    CUXB_CUDA(cudaEventElapsedTime(&ms, pb.e0, pb.e1));
    if (r >= 1 && ms < best) best = ms;
  }
  CUXB_CUDA(cudaEventRecord(pb.e0, ctx->stage_stream));
  for (int r = 0; r < reps; r++) {
    CUXB_CUDA(cudaLaunchKernelEx(&cfg, kern, chunks, seed));
    g_launches.fetch_add(1);
  }
  CUXB_CUDA(cudaEventRecord(pb.e1, ctx->stage_stream));
  CUXB_CUDA(cudaEventSynchronize(pb.e1));
  float ms_all = 0;
  CUXB_CUDA(cudaEventElapsedTime(&ms_all, pb.e0, pb.e1));
  *best_tflops = ops / ((double)best * 1e-3) / 1e12;
  *mean_tflops = ops * reps / ((double)ms_all * 1e-3) / 1e12;
  return 0;
}

// Bandwidth from `peer_device`'s memory into this context's GPU (modes 0, 2, 3) or the other way (mode 1):
// one entry of the pair matrix a harness prints next to the 900 GB/s per direction of NVLink 5 -- the B200
// restatement of the reference's peer-access / topology queries (pocket-ai/eval/cuda/device_query.hpp:134-186),
// which list who can reach whom but measure nothing.  mode 0 / 1 / 3: see the kernels above; mode 2:
// cudaMemcpyPeerAsync (the copy engines).  depth: buffers in flight per SM for mode 3 (1..12), else ignored.
// Allocates `bytes` on both GPUs for the duration of the call and enables peer access both ways.
CUXB_API int cuxb_probe_nvlink(cuxb_ctx *ctx, int peer_device, int mode, int depth, size_t bytes, int reps,
                               double *gb_per_s) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(gb_per_s != nullptr && reps >= 1 && reps <= 1000 && mode >= 0 && mode <= 3, CUXB_ERR_INVALID,
               "cuxb_probe_nvlink: bad argument");
  CUXB_REQUIRE(peer_device != ctx->device, CUXB_ERR_INVALID, "cuxb_probe_nvlink: peer_device is this context's own GPU");
  CUXB_REQUIRE(bytes >= (size_t)kLinkBox * 1024 && bytes % kLinkBox == 0, CUXB_ERR_INVALID,
               "cuxb_probe_nvlink: bytes must be a multiple of 16 KiB, at least 16 MiB");
  CUXB_REQUIRE(mode != 3 || (depth >= 1 && depth <= 12), CUXB_ERR_INVALID, "cuxb_probe_nvlink: depth must be 1..12");
  rc = cuxb_enable_peer_access(ctx->device, peer_device);
  if (rc) return rc;
  rc = cuxb_enable_peer_access(peer_device, ctx->device);
  if (rc) return rc;
  ProbeBuffers pb;   // a = local, b = on the peer
  CUXB_CUDA(cudaMalloc(&pb.a, bytes));
  CUXB_CUDA(cudaMemset(pb.a, 0, bytes));
  CUXB_CUDA(cudaSetDevice(peer_device));
  cudaError_t e = cudaMalloc(&pb.b, bytes);
  if (e == cudaSuccess) e = cudaMemset(pb.b, 1, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaSetDevice(ctx->device);
  CUXB_CUDA(e);
  CUXB_CUDA(cudaEventCreate(&pb.e0));
  CUXB_CUDA(cudaEventCreate(&pb.e1));
  const int smem = depth * kLinkBox + 128 + 8 * 16;
  if (mode == 3) CUXB_CUDA(cudaFuncSetAttribute(probe_link_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  float best = 1e30f;
  for (int r = 0; r < reps + 2; r++) {
    CUXB_CUDA(cudaEventRecord(pb.e0, ctx->stage_stream));
    if (mode == 0) probe_link_copy<<<ctx->sm_count * 2, 256, 0, ctx->stage_stream>>>((const float4 *)pb.b, (float4 *)pb.a, bytes / 16);
    else if (mode == 1) probe_link_copy<<<ctx->sm_count * 2, 256, 0, ctx->stage_stream>>>((const float4 *)pb.a, (float4 *)pb.b, bytes / 16);
    else if (mode == 2) CUXB_CUDA(cudaMemcpyPeerAsync(pb.a, ctx->device, pb.b, peer_device, bytes, ctx->stage_stream));
    else probe_link_bulk<<<ctx->sm_count, 32, smem, ctx->stage_stream>>>((const uint8_t *)pb.b, (uint8_t *)pb.a, bytes / kLinkBox, depth);
    if (mode != 2) CUXB_LAUNCHED();
    CUXB_CUDA(cudaEventRecord(pb.e1, ctx->stage_stream));
    CUXB_CUDA(cudaEventSynchronize(pb.e1));
    float ms = 0;
    CUXB_CUDA(cudaEventElapsedTime(&ms, pb.e0, pb.e1));
    if (r >= 2 && ms < best) best = ms;
  }
  *gb_per_s = (double)bytes / ((double)best * 1e-3) / 1e9;
  return 0;
}
