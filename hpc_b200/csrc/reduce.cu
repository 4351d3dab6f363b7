// reduce.cu -- fp32 sum / dot / nrm2 and int32 sum, HBM-bound, one launch each.
//
// Replaces DotDeviceV0/V1/V2 + cublasSdot (0-frameworks/cux/src/operator/
// dot_product.cu:9-112,203-204), Nrm2DeviceV0 + Sqrt + cublasSnrm2
// (src/operator/nrm2.cu:7-23,58-59) and ReduceKernelFp32V1/V2
// (cuda/reduce_fp32.cu:33-116).  Those kernels read one element per thread and
// finish with one fp32 atomicAdd per block, so their result depends on block
// scheduling.  Here:
//   * a persistent grid (2 CTAs of 512 threads per SM) streams the input with
//     128-bit ld.global.nc.L1::no_allocate loads, four independent loads in
//     flight per thread and per operand;
//   * each thread keeps four accumulators, combined in a fixed order; warps
//     combine with a shuffle butterfly, the CTA through shared memory;
//   * each CTA writes ONE partial to scratch memory and takes a ticket; the
//     CTA that draws the last ticket adds the partials in index order and
//     OVERWRITES *res (sqrt fused for nrm2).  No floating-point atomics, so
//     the result is a pure function of (input, n, grid size).
//
// Algorithmic bytes: sum/nrm2 4n, dot 8n (read once, nothing written but 4 B).
#include <string.h>

#include "common.cuh"

namespace cuxb {

enum { OP_SUM = 0, OP_DOT = 1, OP_SUMSQ = 2 };
constexpr int kRedThreads = 512;
constexpr int kRedBlocksPerSM = 2;

__device__ __forceinline__ float4 ldg_stream(const float4 *p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ---- multi-GPU: the scalar exchange is part of the reduction kernel.  Every rank's finishing CTA
// writes its partial (value | epoch in one 64-bit word) straight into EVERY rank's exchange buffer
// over NVLink (peer-mapped pointers), then waits until its own buffer holds this call's word from
// every rank and adds them up in rank order -- so all ranks get the same, order-fixed result without
// a separate collective launch.  Slots alternate with the epoch's parity: a rank that runs one call
// ahead writes the other half, and it cannot run two ahead because it needs everybody's word first.
// The epoch is a call counter kept in the rank's OWN exchange buffer (word 16; only this warp ever
// touches it), so every rank counts the same sequence without the host -- and a call recorded into
// a CUDA graph runs under a fresh epoch on every replay.
struct MgExchange {
  int world, rank;              // world <= 1: single GPU, no exchange
  int mode;                     // 0: sum over all ranks   1: exclusive prefix (sum of the lower ranks)
  unsigned long long *peer[8];  // peer[r] = rank r's exchange buffer (>= CUXB_MG_EXCHANGE_BYTES)
};
__device__ __forceinline__ void st_sys_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
template <typename T> __device__ __forceinline__ unsigned int bits_of(T v);
template <> __device__ __forceinline__ unsigned int bits_of<float>(float v) { return __float_as_uint(v); }
template <> __device__ __forceinline__ unsigned int bits_of<unsigned int>(unsigned int v) { return v; }
template <typename T> __device__ __forceinline__ T of_bits(unsigned int v);
template <> __device__ __forceinline__ float of_bits<float>(unsigned int v) { return __uint_as_float(v); }
template <> __device__ __forceinline__ unsigned int of_bits<unsigned int>(unsigned int v) { return v; }

// Executed by warp 0 of the finishing CTA.  Returns the combined value to every lane.
template <typename T>
__device__ __forceinline__ T mg_exchange(const MgExchange &mg, T mine) {
  const int lane = threadIdx.x & 31;
  unsigned long long *dst = nullptr, *own = nullptr;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    if (r == lane) dst = mg.peer[r];
    if (r == mg.rank) own = mg.peer[r];
  }
  unsigned long long calls = 0;                       // exchanges completed on this buffer
  if (lane == 0) calls = ld_sys_u64(own + 16);
  calls = __shfl_sync(0xffffffffu, calls, 0);
  const unsigned int epoch = (unsigned int)(calls % 0xffffffffull) + 1u;   // never 0 (zeroed buffers)
  const int slot = (int)(calls & 1) * 8;   // strict alternation (the epoch's own parity repeats once at its wrap)
  const unsigned long long word = ((unsigned long long)bits_of<T>(mine) << 32) | epoch;
  if (lane < mg.world) st_sys_u64(dst + slot + mg.rank, word);
  unsigned long long w = 0;
  if (lane < mg.world) {
    const unsigned long long t0 = global_timer_ns();
    while (true) {
      w = ld_sys_u64(own + slot + lane);
      if ((unsigned int)w == epoch) break;
      if (global_timer_ns() - t0 > 20000000000ull) __trap();  // a rank never arrived: fail, do not hang
    }
  }
  __syncwarp();
  if (lane == 0) st_sys_u64(own + 16, calls + 1);   // everybody's word for this call has been read
  const T val = (lane < mg.world) ? of_bits<T>((unsigned int)(w >> 32)) : (T)0;
  T total = 0;
  for (int r = 0; r < mg.world; r++) {  // rank order: every rank adds in the same order
    const T vr = __shfl_sync(0xffffffffu, val, r);
    if (mg.mode == 0 || r < mg.rank) total += vr;
  }
  return total;
}

template <typename T> struct Vec4;
template <> struct Vec4<float> { typedef float4 type; };
template <> struct Vec4<unsigned int> { typedef uint4 type; };

// fixed association: ((x*x' + y*y') + (z*z' + w*w'))
template <int OP, typename T, typename V>
__device__ __forceinline__ T fold4(const V &a, const V &b) {
  if (OP == OP_SUM) return (a.x + a.y) + (a.z + a.w);
  if (OP == OP_DOT) return (a.x * b.x + a.y * b.y) + (a.z * b.z + a.w * b.w);
  return (a.x * a.x + a.y * a.y) + (a.z * a.z + a.w * a.w);
}
template <int OP, typename T>
__device__ __forceinline__ T fold1(T a, T b) {
  if (OP == OP_SUM) return a;
  if (OP == OP_DOT) return a * b;
  return a * a;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// All threads of the CTA call this; the CTA total is returned to every thread of warp 0.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T *smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  T t = 0;
  if (warp == 0) {
    t = lane < (kRedThreads / 32) ? smem[lane] : (T)0;
    t = warp_sum(t);
  }
  __syncthreads();
  return t;
}

// VEC: 4 = 128-bit loads (pointers 16-byte aligned), 1 = scalar path for misaligned input.
template <int OP, typename T, int VEC>
__global__ void __launch_bounds__(kRedThreads, kRedBlocksPerSM)
reduce_kernel(const T *__restrict__ a, const T *__restrict__ b, long long n,
              T *__restrict__ partials, unsigned int *__restrict__ ticket, T *__restrict__ res,
              int finish_sqrt, const MgExchange mg) {
  __shared__ T smem[kRedThreads / 32];
  __shared__ bool is_last;
  typedef typename Vec4<T>::type V;
  const long long tid = (long long)blockIdx.x * kRedThreads + threadIdx.x;
  const long long stride = (long long)gridDim.x * kRedThreads;
  T acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;

  if (VEC == 4) {
    const long long nvec = n >> 2;
    const V *a4 = reinterpret_cast<const V *>(a);
    const V *b4 = reinterpret_cast<const V *>(OP == OP_DOT ? b : a);
    long long i = tid;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
      V x0 = ldg_stream(a4 + i), x1 = ldg_stream(a4 + i + stride);
      V x2 = ldg_stream(a4 + i + 2 * stride), x3 = ldg_stream(a4 + i + 3 * stride);
      V y0 = x0, y1 = x1, y2 = x2, y3 = x3;
      if (OP == OP_DOT) {
        y0 = ldg_stream(b4 + i), y1 = ldg_stream(b4 + i + stride);
        y2 = ldg_stream(b4 + i + 2 * stride), y3 = ldg_stream(b4 + i + 3 * stride);
      }
      acc0 += fold4<OP, T>(x0, y0);
      acc1 += fold4<OP, T>(x1, y1);
      acc2 += fold4<OP, T>(x2, y2);
      acc3 += fold4<OP, T>(x3, y3);
    }
    // at most three vectors per thread are left: issue their loads together, then fold (a loop
    // would wait for each load in turn -- most of a short vector's run time)
    {
      const bool p0 = i < nvec, p1 = i + stride < nvec, p2 = i + 2 * stride < nvec;
      V x0 = {}, x1 = {}, x2 = {};
      if (p0) x0 = ldg_stream(a4 + i);
      if (p1) x1 = ldg_stream(a4 + i + stride);
      if (p2) x2 = ldg_stream(a4 + i + 2 * stride);
      V y0 = x0, y1 = x1, y2 = x2;
      if (OP == OP_DOT) {
        if (p0) y0 = ldg_stream(b4 + i);
        if (p1) y1 = ldg_stream(b4 + i + stride);
        if (p2) y2 = ldg_stream(b4 + i + 2 * stride);
      }
      if (p0) acc0 += fold4<OP, T>(x0, y0);
      if (p1) acc0 += fold4<OP, T>(x1, y1);
      if (p2) acc0 += fold4<OP, T>(x2, y2);
    }
    // n % 4 trailing scalars: one per thread of CTA 0
    const long long rem = (nvec << 2) + tid;
    if (tid < 4 && rem < n) acc1 += fold1<OP, T>(a[rem], OP == OP_DOT ? b[rem] : a[rem]);
  } else {
    for (long long i = tid; i < n; i += stride)
      acc0 += fold1<OP, T>(a[i], OP == OP_DOT ? b[i] : a[i]);
  }
  T v = (acc0 + acc1) + (acc2 + acc3);
  v = block_sum(v, smem);

  // one partial per CTA, then a ticket; the last CTA folds the partials in index order
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = v;
    __threadfence();
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  T p = 0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kRedThreads) p += __ldcg(partials + i);
  p = block_sum(p, smem);
  if (mg.world > 1 && threadIdx.x < 32) p = mg_exchange<T>(mg, p);  // warp 0 holds the CTA total
  if (threadIdx.x == 0) {
    if (finish_sqrt) {
      // Nrm2HostV0 (nrm2.cpp:8-9): n == 1 -> |x0| exactly (single GPU)
      float f = (n == 1 && mg.world <= 1) ? fabsf((float)a[0]) : sqrtf((float)p);
      res[0] = (T)f;
    } else {
      res[0] = p;
    }
    *ticket = 0;  // at rest for the next launch on this stream
  }
}

static int reduce_grid(const cuxb_ctx *ctx, long long n) {
  long long want = div_up(n > 0 ? n : 1, (long long)kRedThreads * 4 * 4);
  long long cap = (long long)ctx->sm_count * kRedBlocksPerSM;
  if (cap > kReduceMaxBlocks) cap = kReduceMaxBlocks;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

template <int OP, typename T>
static int launch_reduce(cuxb_ctx *ctx, cudaStream_t st, long long n, const T *a, const T *b, T *res,
                         int finish_sqrt, void *partials, unsigned int *ticket, int grid_cap,
                         const MgExchange &mg) {
  const int grid = grid_cap;
  const bool aligned = (((uintptr_t)a | (uintptr_t)(OP == OP_DOT ? b : a)) & 15) == 0;
  if (aligned)
    reduce_kernel<OP, T, 4><<<grid, kRedThreads, 0, st>>>(a, b, n, (T *)partials, ticket, res, finish_sqrt, mg);
  else
    reduce_kernel<OP, T, 1><<<grid, kRedThreads, 0, st>>>(a, b, n, (T *)partials, ticket, res, finish_sqrt, mg);
  CUXB_LAUNCHED();
  return 0;
}

template <int OP, typename T>
static int reduce_entry(const char *name, cuxb_ctx *ctx, void *stream, long long n, const T *a,
                        const T *b, T *res, int finish_sqrt, const MgExchange *mgp = nullptr) {
  MgExchange mg;
  memset(&mg, 0, sizeof(mg));
  if (mgp) mg = *mgp;
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(res != nullptr, CUXB_ERR_INVALID, "%s: result pointer is null", name);
  CUXB_REQUIRE(n >= 0, CUXB_ERR_INVALID, "%s: negative length %lld", name, n);
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0 && mg.world <= 1) {
    CUXB_CUDA(cudaMemsetAsync(res, 0, sizeof(T), st));
    return 0;
  }
  CUXB_REQUIRE(n == 0 || (a != nullptr && (OP != OP_DOT || b != nullptr)), CUXB_ERR_INVALID,
               "%s: input pointer is null", name);
  return launch_reduce<OP, T>(ctx, st, n, a, b, res, finish_sqrt, ctx->red_partials,
                              ctx->red_ticket, reduce_grid(ctx, n), mg);
}

const void *reduce_kernel_symbol(const char *name) {
  if (!strcmp(name, "dot_f32")) return (const void *)reduce_kernel<OP_DOT, float, 4>;
  if (!strcmp(name, "sum_f32")) return (const void *)reduce_kernel<OP_SUM, float, 4>;
  if (!strcmp(name, "nrm2_f32")) return (const void *)reduce_kernel<OP_SUMSQ, float, 4>;
  return nullptr;
}

}  // namespace cuxb

using namespace cuxb;

CUXB_API int cuxb_dot_f32(cuxb_ctx *ctx, void *stream, long long n, const float *a, const float *b,
                          float *res) {
  return reduce_entry<OP_DOT, float>("cuxb_dot_f32", ctx, stream, n, a, b, res, 0);
}
CUXB_API int cuxb_sum_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res) {
  return reduce_entry<OP_SUM, float>("cuxb_sum_f32", ctx, stream, n, x, nullptr, res, 0);
}
CUXB_API int cuxb_nrm2_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res) {
  return reduce_entry<OP_SUMSQ, float>("cuxb_nrm2_f32", ctx, stream, n, x, nullptr, res, 1);
}
CUXB_API int cuxb_sum_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *x, int32_t *res) {
  return reduce_entry<OP_SUM, unsigned int>("cuxb_sum_i32", ctx, stream, n, (const unsigned int *)x,
                                            nullptr, (unsigned int *)res, 0);
}

// ---- multi-GPU entry points: the same kernels with the scalar exchange fused into the finishing CTA
namespace cuxb {
int mg_prepare(const char *name, cuxb_ctx *ctx, int world, int rank, void *const *peer_bufs, int mode, MgExchange *mg) {
  CUXB_REQUIRE(ctx != nullptr, CUXB_ERR_INVALID, "%s: null ctx", name);
  CUXB_REQUIRE(world >= 1 && world <= 8 && rank >= 0 && rank < world, CUXB_ERR_INVALID,
               "%s: world=%d rank=%d (1..8 ranks supported)", name, world, rank);
  memset(mg, 0, sizeof(*mg));
  mg->world = world, mg->rank = rank, mg->mode = mode;
  if (world == 1) return 0;
  CUXB_REQUIRE(peer_bufs != nullptr, CUXB_ERR_INVALID, "%s: peer_bufs is null", name);
  for (int r = 0; r < world; r++) {
    CUXB_REQUIRE(peer_bufs[r] != nullptr && ((uintptr_t)peer_bufs[r] & 7) == 0, CUXB_ERR_INVALID,
                 "%s: peer buffer %d is null or not 8-byte aligned", name, r);
    mg->peer[r] = (unsigned long long *)peer_bufs[r];
  }
  return 0;
}
}  // namespace cuxb

CUXB_API int cuxb_mg_dot_f32(cuxb_ctx *ctx, void *stream, long long n, const float *a, const float *b, float *res,
                             int world, int rank, void *const *peer_bufs) {
  MgExchange mg;
  int rc = mg_prepare("cuxb_mg_dot_f32", ctx, world, rank, peer_bufs, 0, &mg);
  if (rc) return rc;
  return reduce_entry<OP_DOT, float>("cuxb_mg_dot_f32", ctx, stream, n, a, b, res, 0, &mg);
}
CUXB_API int cuxb_mg_sum_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res, int world,
                             int rank, void *const *peer_bufs) {
  MgExchange mg;
  int rc = mg_prepare("cuxb_mg_sum_f32", ctx, world, rank, peer_bufs, 0, &mg);
  if (rc) return rc;
  return reduce_entry<OP_SUM, float>("cuxb_mg_sum_f32", ctx, stream, n, x, nullptr, res, 0, &mg);
}
CUXB_API int cuxb_mg_nrm2_f32(cuxb_ctx *ctx, void *stream, long long n, const float *x, float *res, int world,
                              int rank, void *const *peer_bufs) {
  MgExchange mg;
  int rc = mg_prepare("cuxb_mg_nrm2_f32", ctx, world, rank, peer_bufs, 0, &mg);
  if (rc) return rc;
  return reduce_entry<OP_SUMSQ, float>("cuxb_mg_nrm2_f32", ctx, stream, n, x, nullptr, res, 1, &mg);
}
// exclusive prefix over ranks of the per-GPU int32 totals -> *seed_out (device): phase 1 of the sharded scan
CUXB_API int cuxb_mg_exscan_total_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *x, int32_t *seed_out,
                                      int world, int rank, void *const *peer_bufs) {
  MgExchange mg;
  int rc = mg_prepare("cuxb_mg_exscan_total_i32", ctx, world, rank, peer_bufs, 1, &mg);
  if (rc) return rc;
  if (world == 1) {  // nothing in front of the only rank
    CUXB_REQUIRE(seed_out != nullptr, CUXB_ERR_INVALID, "cuxb_mg_exscan_total_i32: null seed_out");
    CUXB_CUDA(cudaMemsetAsync(seed_out, 0, 4, (cudaStream_t)stream));
    return 0;
  }
  return reduce_entry<OP_SUM, unsigned int>("cuxb_mg_exscan_total_i32", ctx, stream, n, (const unsigned int *)x, nullptr,
                                            (unsigned int *)seed_out, 0, &mg);
}
