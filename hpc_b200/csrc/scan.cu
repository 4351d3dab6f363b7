// scan.cu -- inclusive prefix scan, single pass, HBM-bound (8 bytes per element).
//
// Replaces VectorScanNormal / VectorScanSIMDv1 / VectorScanSIMDv2
// (simd/x86/vector_scan.cpp:21-79): out[0]=in[0]; out[i]=in[i]+out[i-1], in the
// int element type of tbb/base_parallel_scan.cpp:62-65 (adds are done in
// uint32, i.e. two's-complement wrap-around, identical to the CPU oracle) and,
// as a second instantiation, in fp32.
//
// One pass over the data ("decoupled look-back"): the input is cut into tiles
// of 4096 elements; a CTA takes the next tile number from a global ticket (so
// a CTA only ever waits on tiles that are already running), scans its tile in
// registers (128-bit loads, shuffle scans), publishes the tile aggregate in a
// 64-bit state word, then walks back over its predecessors' words -- 32 at a
// time, one per lane -- until it meets one that already carries an inclusive
// prefix.  value(32) | epoch(30) | status(2) share one word, so a single 64-bit
// load observes a consistent (status, value) pair and the words never need
// clearing between launches: a word from an older launch has a stale epoch and
// reads as "not ready".
//
// Scratch memory: 8 bytes per tile (2 KiB per Mi elements) + two counters,
// owned by the cuxb_ctx.  Launches that share a ctx must not overlap.
#include "common.cuh"

namespace cuxb {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 16;  // per thread: 4 x 128-bit
static_assert(kScanThreads * kScanItems == kScanTileElems, "tile size");

constexpr unsigned long long ST_AGG = 1ull, ST_PREFIX = 2ull;

__device__ __forceinline__ unsigned long long pack_state(unsigned int value, unsigned int epoch,
                                                         unsigned long long status) {
  return ((unsigned long long)value << 32) | ((unsigned long long)(epoch & 0x3fffffffu) << 2) | status;
}
__device__ __forceinline__ unsigned long long ld_state(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_state(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <typename T> __device__ __forceinline__ unsigned int to_bits(T v);
template <> __device__ __forceinline__ unsigned int to_bits<unsigned int>(unsigned int v) { return v; }
template <> __device__ __forceinline__ unsigned int to_bits<float>(float v) { return __float_as_uint(v); }
template <typename T> __device__ __forceinline__ T from_bits(unsigned int v);
template <> __device__ __forceinline__ unsigned int from_bits<unsigned int>(unsigned int v) { return v; }
template <> __device__ __forceinline__ float from_bits<float>(unsigned int v) { return __uint_as_float(v); }

template <typename T> struct V4;
template <> struct V4<unsigned int> { typedef uint4 type; };
template <> struct V4<float> { typedef float4 type; };

// Look-back executed by warp 0.  Returns the exclusive prefix of `tile` (seed included
// through tile 0's published prefix).
template <typename T>
__device__ __forceinline__ T look_back(const unsigned long long *state, unsigned int tile,
                                       unsigned int epoch) {
  const int lane = threadIdx.x & 31;
  const unsigned long long want_epoch = (unsigned long long)(epoch & 0x3fffffffu);
  T running = 0;
  long long pos = (long long)tile - 1;
  while (true) {
    const long long idx = pos - lane;
    const bool valid = idx >= 0;  // lanes that fall before tile 0 contribute nothing
    unsigned long long w = 0;
    bool ready;
    unsigned int spin = 0;
    do {
      if (valid) w = ld_state(state + idx);
      ready = !valid || (((w & 3ull) != 0) && (((w >> 2) & 0x3fffffffull) == want_epoch));
      if (++spin > 200000000u) __trap();  // a protocol bug must fail the launch, never hang the GPU
    } while (!__all_sync(0xffffffffu, ready));
    const bool is_prefix = valid && ((w & 3ull) == ST_PREFIX);
    const unsigned int pm = __ballot_sync(0xffffffffu, is_prefix);
    const T val = (idx >= 0) ? from_bits<T>((unsigned int)(w >> 32)) : (T)0;
    // nearest predecessor with a full prefix: take it and every aggregate in front of it
    const int stop = pm ? (__ffs(pm) - 1) : 31;
    T contrib = (lane <= stop) ? val : (T)0;
    // fixed-order fold across lanes (lane 0 = nearest tile)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
    running += contrib;
    if (pm) break;
    pos -= 32;
  }
  return running;
}

template <typename T>
__global__ void __launch_bounds__(kScanThreads, 6)
scan_kernel(const T *in, T *out, long long n, T seed, unsigned long long *__restrict__ state,
            unsigned int *__restrict__ counter, unsigned int epoch, unsigned int num_tiles) {
  typedef typename V4<T>::type V;
  __shared__ unsigned int s_tile;
  __shared__ T s_warp[kScanThreads / 32];
  __shared__ T s_prefix;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) s_tile = atomicAdd(&counter[0], 1u);
  __syncthreads();
  const unsigned int tile = s_tile;
  const long long base = (long long)tile * kScanTileElems + warp * 512 + lane * 4;
  const bool full = ((long long)(tile + 1) * kScanTileElems <= n) && ((((uintptr_t)in | (uintptr_t)out) & 15) == 0);

  T v[4][4];
  if (full) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      V t = __ldcs(reinterpret_cast<const V *>(in + base + j * 128));
      v[j][0] = t.x, v[j][1] = t.y, v[j][2] = t.z, v[j][3] = t.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const long long i = base + j * 128 + c;
        v[j][c] = (i < n) ? in[i] : (T)0;
      }
  }

  // thread-local inclusive scan of each 4-vector, then a shuffle scan per row j,
  // rows chained through `carry`
  T carry = 0;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    v[j][1] += v[j][0];
    v[j][2] += v[j][1];
    v[j][3] += v[j][2];
    T incl = v[j][3];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      T t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    T excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = 0;
    const T off = carry + excl;
#pragma unroll
    for (int c = 0; c < 4; c++) v[j][c] += off;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 31) s_warp[warp] = carry;
  __syncthreads();

  T warp_off = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; w++) {
    const T t = s_warp[w];
    if (w < warp) warp_off += t;
    block_total += t;
  }

  if (warp == 0) {
    T prefix;
    if (tile == 0) {
      prefix = seed;
      if (lane == 0) st_state(state, pack_state(to_bits<T>(seed + block_total), epoch, ST_PREFIX));
    } else {
      if (lane == 0) st_state(state + tile, pack_state(to_bits<T>(block_total), epoch, ST_AGG));
      prefix = look_back<T>(state, tile, epoch);
      if (lane == 0) st_state(state + tile, pack_state(to_bits<T>(prefix + block_total), epoch, ST_PREFIX));
    }
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  const T add = s_prefix + warp_off;

  if (full) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      V t;
      t.x = v[j][0] + add, t.y = v[j][1] + add, t.z = v[j][2] + add, t.w = v[j][3] + add;
      __stcs(reinterpret_cast<V *>(out + base + j * 128), t);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const long long i = base + j * 128 + c;
        if (i < n) out[i] = v[j][c] + add;
      }
  }

  // the last tile to finish puts both counters back to zero for the next launch
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(&counter[1], 1u);
    if (done == num_tiles - 1) {
      counter[0] = 0;
      counter[1] = 0;
    }
  }
}

template <typename T>
static int scan_entry(const char *name, cuxb_ctx *ctx, void *stream, long long n, const T *in, T *out,
                      T seed) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(n >= 0, CUXB_ERR_INVALID, "%s: negative length %lld", name, n);
  if (n == 0) return 0;
  CUXB_REQUIRE(in && out, CUXB_ERR_INVALID, "%s: null pointer", name);
  const long long tiles = div_up(n, kScanTileElems);
  CUXB_REQUIRE(tiles <= kScanMaxTiles, CUXB_ERR_TOO_LARGE,
               "%s: n=%lld exceeds the maximum of %lld elements", name, n,
               kScanMaxTiles * kScanTileElems);
  cudaStream_t st = (cudaStream_t)stream;
  if (ctx->scan_state_tiles < tiles) {
    // grow the tile-state scratch (zeroed: epoch 0 is never used by a launch)
    long long cap = ctx->scan_state_tiles ? ctx->scan_state_tiles : 4096;
    while (cap < tiles) cap *= 2;
    if (ctx->scan_state) {
      CUXB_CUDA(cudaDeviceSynchronize());
      CUXB_CUDA(cudaFree(ctx->scan_state));
      ctx->scan_state = nullptr;
      ctx->scan_state_tiles = 0;
    }
    CUXB_CUDA(cudaMalloc((void **)&ctx->scan_state, cap * sizeof(unsigned long long)));
    CUXB_CUDA(cudaMemset(ctx->scan_state, 0, cap * sizeof(unsigned long long)));
    ctx->scan_state_tiles = cap;
  }
  ctx->scan_epoch = (ctx->scan_epoch + 1) & 0x3fffffffu;
  if (ctx->scan_epoch == 0) {
    // 2^30 launches later the epoch wraps: clear the words once and skip 0
    CUXB_CUDA(cudaMemsetAsync(ctx->scan_state, 0, ctx->scan_state_tiles * sizeof(unsigned long long), st));
    ctx->scan_epoch = 1;
  }
  scan_kernel<T><<<(unsigned int)tiles, kScanThreads, 0, st>>>(in, out, n, seed, ctx->scan_state,
                                                                ctx->scan_counter, ctx->scan_epoch,
                                                                (unsigned int)tiles);
  CUXB_LAUNCHED();
  return 0;
}

const void *scan_kernel_symbol() { return (const void *)scan_kernel<unsigned int>; }

// ---------------------------------------------------------------------------------------------
// Allocation-free variant (cuxb_scan_i32_noalloc).
//
// The scan needs scratch memory besides its input and output: one 64-bit state word per tile and
// two counters.  The entry points above keep that scratch in the cuxb_ctx, which is created with
// cudaMalloc and may grow it.  Some callers cannot do that: a plug-in that runs inside a host
// application which has already handed all device memory to its own pool, a service with a fixed
// memory budget where a late cudaMalloc may fail, or a latency-sensitive loop that cannot afford
// the device-wide synchronisation an allocation implies.  For them the scratch lives in the
// file-scope array below.  It is part of the library image, so every GPU that loads libcuxb.so
// holds these ~128 KiB until the process exits, whether or not the entry point is ever called.
//
// Size limit: kNoAllocMaxElems elements (one state word per 4096-element tile).
constexpr long long kNoAllocMaxElems = 1ll << 26;  // 67,108,864 int32 values = 256 MiB of input
constexpr int kNoAllocMaxTiles = (int)(kNoAllocMaxElems / kScanTileElems);  // 16384 words = 128 KiB
constexpr int kNoAllocCounterWords = 32;  // 256 bytes after the state words; [0],[1] are used
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocMaxTiles + kNoAllocCounterWords];

// Host-side bookkeeping, one slot per device ordinal.  `scratch` is the DEVICE address of
// g_noalloc_scratch on that GPU: in host code the array's name is only a handle for the runtime,
// not something a kernel can dereference, so the address is asked for once per device with
// cudaGetSymbolAddress while that device is current.
constexpr int kNoAllocMaxDevices = 64;
struct NoAllocDevice {
  unsigned long long *scratch;
  unsigned int epoch;
  bool ready;
};
static NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
static std::mutex g_noalloc_mu;

}  // namespace cuxb

using namespace cuxb;

CUXB_API int cuxb_scan_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *in, int32_t *out,
                           int32_t seed) {
  return scan_entry<unsigned int>("cuxb_scan_i32", ctx, stream, n, (const unsigned int *)in,
                                  (unsigned int *)out, (unsigned int)seed);
}
CUXB_API int cuxb_scan_f32(cuxb_ctx *ctx, void *stream, long long n, const float *in, float *out,
                           float seed) {
  return scan_entry<float>("cuxb_scan_f32", ctx, stream, n, in, out, seed);
}

// See include/cuxb.h.  Works on the CURRENT device; takes no cuxb_ctx because creating one allocates.
// Calls no cudaMalloc / cudaFree / allocating library routine.  The first call on each device looks
// up the device address of the static scratch array (cudaGetSymbolAddress); make that first call
// outside any stream capture.  All calls on one GPU share the same scratch: two calls whose kernels
// overlap on that GPU (different streams, or different host threads) race on it and produce garbage.
CUXB_API int cuxb_scan_i32_noalloc(void *stream, long long n, const int32_t *in, int32_t *out, int32_t seed) {
  CUXB_REQUIRE(n >= 0, CUXB_ERR_INVALID, "cuxb_scan_i32_noalloc: negative length %lld", n);
  CUXB_REQUIRE(n <= kNoAllocMaxElems, CUXB_ERR_TOO_LARGE,
               "cuxb_scan_i32_noalloc: n=%lld is larger than the %lld elements the built-in scratch array covers; "
               "use cuxb_scan_i32 (which allocates) or split the input and chain the pieces through `seed`",
               n, kNoAllocMaxElems);
  if (n == 0) return 0;
  CUXB_REQUIRE(in && out, CUXB_ERR_INVALID, "cuxb_scan_i32_noalloc: null pointer");
  int dev = -1;
  CUXB_CUDA(cudaGetDevice(&dev));
  CUXB_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, CUXB_ERR_UNSUPPORTED,
               "cuxb_scan_i32_noalloc: device ordinal %d is outside the %d supported", dev, kNoAllocMaxDevices);
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  NoAllocDevice &d = g_noalloc_dev[dev];
  if (!d.ready) {
    // one-time set-up for this GPU: resolve the array's device address (no allocation happens here)
    unsigned long long *ptr = nullptr;
    CUXB_CUDA(cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch));
    CUXB_REQUIRE(ptr != nullptr, CUXB_ERR_UNSUPPORTED, "cuxb_scan_i32_noalloc: static scratch has no device address");
    d.scratch = ptr;
    d.epoch = 0;   // the array starts zero-filled, and no launch ever uses epoch 0
    d.ready = true;
  }
  // The array still holds whatever the previous call left.  State words are told apart by the
  // launch epoch, but the two tile counters must read zero when the kernel starts; a kernel that
  // was aborted half-way would have left them set, so they are cleared on the stream every time.
  unsigned int *counters = reinterpret_cast<unsigned int *>(d.scratch + kNoAllocMaxTiles);
  CUXB_CUDA(cudaMemsetAsync(counters, 0, 2 * sizeof(unsigned int), st));
  d.epoch = (d.epoch + 1) & 0x3fffffffu;
  if (d.epoch == 0) {  // epoch wrapped after 2^30 calls: wipe the state words once and skip 0
    CUXB_CUDA(cudaMemsetAsync(d.scratch, 0, (size_t)kNoAllocMaxTiles * sizeof(unsigned long long), st));
    d.epoch = 1;
  }
  const unsigned int tiles = (unsigned int)div_up(n, kScanTileElems);
  scan_kernel<unsigned int><<<tiles, kScanThreads, 0, st>>>((const unsigned int *)in, (unsigned int *)out, n,
                                                            (unsigned int)seed, d.scratch, counters, d.epoch, tiles);
  CUXB_LAUNCHED();
  return 0;
}
