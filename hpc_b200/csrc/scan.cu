// scan.cu -- inclusive prefix scan, single pass, HBM-bound (8 bytes per element).
//
// Replaces VectorScanNormal / VectorScanSIMDv1 / VectorScanSIMDv2
// (simd/x86/vector_scan.cpp:21-79): out[0]=in[0]; out[i]=in[i]+out[i-1], in the
// int element type of tbb/base_parallel_scan.cpp:62-65 (adds are done in
// uint32, i.e. two's-complement wrap-around, identical to the CPU oracle) and,
// as a second instantiation, in fp32.
//
// One pass over the data: the input is cut into tiles of 8192 elements; a persistent
// grid of one resident CTA per SM walks the tiles round-robin (so a CTA only ever waits
// on tiles owned by CTAs that are running).  cp.async streams tiles into a 192 KiB
// shared-memory ring two passes ahead; a tile is scanned in registers (shuffle
// scans), its aggregate is published in a 64-bit state word, and its exclusive prefix
// is the CTA's own running prefix plus the aggregates of the tiles in between (see
// "Prefix gathering" below) -- fetched one pass late so the L2 round trip is hidden.
// value(32) | epoch(30) | status(2) share one word, so a single 64-bit
// load observes a consistent (status, value) pair and the words never need
// clearing between launches: a word from an older launch has a stale epoch and
// reads as "not ready".  The epoch itself lives in DEVICE memory next to the state
// words: every CTA reads it when it starts and the last CTA to finish (a ticket)
// advances it, so a launch recorded into a CUDA graph gets a fresh epoch on every
// replay -- nothing about a launch's protocol state is baked into its arguments.
//
// Scratch memory: 8 bytes per tile (1 KiB per Mi elements), owned by the
// cuxb_ctx.  Launches that share a ctx must not overlap.
#include <string.h>

#include "common.cuh"

namespace cuxb {

constexpr int kScanThreads = kScanTileElems / 16;
constexpr int kScanItems = 16;  // per thread: 4 x 128-bit
constexpr int kScanCtasPerSM = 1;  // 128 KiB of shared memory per CTA
static_assert(kScanThreads * kScanItems == kScanTileElems, "tile size");

constexpr unsigned long long ST_AGG = 1ull;  // status field: 0 = not published (or stale epoch)

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
// launch number -> epoch in [1, 2^30 - 1]: never 0, which is what zero-filled scratch holds
__device__ __forceinline__ unsigned int scan_epoch_of(unsigned long long launches_done) {
  return (unsigned int)(launches_done % 0x3fffffffull) + 1u;
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

// ---------------------------------------------------------------------------------------------
// Prefix gathering (warp 0).  With a persistent grid of G CTAs walking the tiles round-robin, the
// tile G places back is this CTA's own previous tile, whose inclusive prefix it still holds in a
// register.  A tile's exclusive prefix is that value plus the AGGREGATES of the (at most G-1)
// tiles in between -- and every CTA publishes its aggregate before it looks at anybody else's, so
// nothing ever waits on another CTA's look-back.  (The classic decoupled look-back waits for a
// predecessor's full prefix; on a grid that runs in lock step that chains several L2 round trips
// per pass and capped this kernel at ~2.8 TB/s.)  The aggregates for tile k are fetched at the
// top of pass k+1 -- by then every CTA has published them -- and consumed at the end of that pass
// together with the store of tile k, so the L2 round trip overlaps the scan of tile k+1.
constexpr int kGatherWords = 5;  // state words per lane: covers G-1 <= 160 predecessors in one round
static_assert(kGatherWords * 32 >= 148 * kScanCtasPerSM - 1, "one gather round must cover a B200's worth of CTAs");

template <typename T>
struct Gather {
  unsigned long long w[kGatherWords];
  unsigned int tile, count;
};

template <typename T>
__device__ __forceinline__ void gather_issue(Gather<T> &g, const unsigned long long *state, unsigned int tile,
                                             unsigned int count) {
  const int lane = threadIdx.x & 31;
  g.tile = tile, g.count = count;
#pragma unroll
  for (int j = 0; j < kGatherWords; j++)
    g.w[j] = ((unsigned)(j * 32 + lane) < count) ? ld_state(state + (tile - 1 - (j * 32 + lane))) : 0ull;
}

// Sum of the aggregates of tiles g.tile-1 ... g.tile-g.count, added in a fixed order.
template <typename T>
__device__ __forceinline__ T gather_finish(Gather<T> &g, const unsigned long long *state, unsigned int epoch) {
  const int lane = threadIdx.x & 31;
  const unsigned long long want_epoch = (unsigned long long)(epoch & 0x3fffffffu);
  T sum = 0;
  for (unsigned int base = 0; base < g.count; base += 32 * kGatherWords) {  // one trip when G-1 <= 160
    if (base) {
#pragma unroll
      for (int j = 0; j < kGatherWords; j++)
        g.w[j] = (base + j * 32 + lane < g.count) ? ld_state(state + (g.tile - 1 - (base + j * 32 + lane))) : 0ull;
    }
    unsigned int spin = 0;
    while (true) {
      bool waiting = false;
#pragma unroll
      for (int j = 0; j < kGatherWords; j++) {
        const bool mine = base + j * 32 + lane < g.count;
        const bool ready = ((g.w[j] & 3ull) != 0) && (((g.w[j] >> 2) & 0x3fffffffull) == want_epoch);
        if (mine && !ready) {
          g.w[j] = ld_state(state + (g.tile - 1 - (base + j * 32 + lane)));  // poll again
          waiting = true;
        }
      }
      if (!__any_sync(0xffffffffu, waiting)) break;
      if (++spin > 100000000u) __trap();  // a protocol bug must fail the launch, never hang the GPU
    }
#pragma unroll
    for (int j = 0; j < kGatherWords; j++)
      if (base + j * 32 + lane < g.count) sum += from_bits<T>((unsigned int)(g.w[j] >> 32));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);  // fixed order
  return sum;
}

// One tile = kScanTileElems consecutive elements; thread (warp w, lane l) owns, for j = 0..3, the
// 4 elements at  w*512 + j*128 + l*4  -- every warp load/store instruction covers 512 contiguous bytes.
template <typename T>
struct TileRegs {
  T v[4][4];
};

__device__ __forceinline__ long long tile_elem_base(unsigned int tile) {
  return (long long)tile * kScanTileElems + (threadIdx.x >> 5) * 512 + (threadIdx.x & 31) * 4;  // warp w: 512 elements
}

// Shared-memory ring (kScanRing tile slots, see the pipeline description further down): every scan
// thread owns the same four 16-byte chunks of each slot -- filled by its own cp.async, read, rewritten
// and read again by itself only.  Layout: see slot_chunk().
// uint4 index of this thread's chunk j of ring slot `slot`: the slot holds the tile in GLOBAL order (warp w's
// 512 elements are 2 KiB contiguous, as in memory), so a warp's share of a finished tile can leave with one
// bulk copy.  A warp instruction touches 32 consecutive 16-byte chunks: conflict-free.
__device__ __forceinline__ int slot_chunk(int slot, int j) {
  return slot * (kScanTileElems / 4) + (threadIdx.x >> 5) * 128 + j * 32 + (threadIdx.x & 31);
}

template <typename T>
__device__ __forceinline__ void prefetch_tile(uint4 *ring, int slot, const T *in, unsigned int tile) {
  const long long base = tile_elem_base(tile);
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(ring + slot_chunk(slot, j));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(in + base + j * 128) : "memory");
  }
}

template <typename T>
__device__ __forceinline__ void store_tile(const TileRegs<T> &r, T add, T *out, long long n, unsigned int tile, bool vec_ok) {
  typedef typename V4<T>::type V;
  const long long base = tile_elem_base(tile);
  if (vec_ok && (long long)(tile + 1) * kScanTileElems <= n) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      V t;
      t.x = r.v[j][0] + add, t.y = r.v[j][1] + add, t.z = r.v[j][2] + add, t.w = r.v[j][3] + add;
      __stcs(reinterpret_cast<V *>(out + base + j * 128), t);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const long long i = base + j * 128 + c;
        if (i < n) out[i] = r.v[j][c] + add;
      }
  }
}

// Named barriers (barrier 0 is left to __syncthreads).
__device__ __forceinline__ void named_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }
constexpr int kScanCtaThreads = kScanThreads + 32;  // 16 scan warps + 1 helper warp
constexpr int BAR_TOTALS = 1;                        // warp totals of a tile are in shared memory
constexpr int BAR_PREFIX = 2;                        // (+ pass % 4) the tile's exclusive prefix is in shared memory

// Software pipeline, in passes (one tile per CTA per pass):
//   pass p - kPrefetch : cp.async of tile p into its ring slot starts
//   pass p             : scan warps scan tile p in registers, park the scanned values in the slot,
//                        publish the warp totals; the helper warp publishes the tile AGGREGATE
//   pass p + kGatherLag: the helper fetches the other CTAs' aggregates for tile p (published two
//                        passes ago, so the loads come back ready) -> exclusive prefix of tile p
//   pass p + kStoreLag : scan warps add the prefix and store tile p; its slot is free again
#ifndef CUXB_SCAN_PREFETCH
#define CUXB_SCAN_PREFETCH 2
#endif
#ifndef CUXB_SCAN_GATHER_LAG
#define CUXB_SCAN_GATHER_LAG 2
#endif
#ifndef CUXB_SCAN_STORE_LAG
#define CUXB_SCAN_STORE_LAG 3
#endif
#ifndef CUXB_SCAN_BULK_STORE
#define CUXB_SCAN_BULK_STORE 1
#endif
constexpr bool kBulkStore = CUXB_SCAN_BULK_STORE != 0;   // finished tiles leave shared memory by cp.async.bulk
constexpr int kPrefetch = CUXB_SCAN_PREFETCH, kGatherLag = CUXB_SCAN_GATHER_LAG, kStoreLag = CUXB_SCAN_STORE_LAG;
constexpr int kScanRing = kPrefetch + kStoreLag + 1;
static_assert(kStoreLag > kGatherLag, "a tile's prefix must be settled before the pass that stores it");
static_assert(kScanRing >= kPrefetch + kStoreLag + 1, "a slot must be stored before it is refilled");
constexpr int kScanSmemBytes = kScanRing * kScanTileElems * 4;  // 192 KiB

// Persistent CTAs, one per SM: CTA b scans tiles b, b+G, b+2G, ... (G = gridDim.x; every CTA is
// resident, which is what makes waiting on another CTA's aggregate safe).  Roles:
//   * 16 scan warps: each thread owns the same four 16-byte chunks of every ring slot (filled by
//     its own cp.async, so cp.async.wait_group is all the synchronisation the data needs).
//   * 1 helper warp: aggregate publication and prefix gathering; its L2 round trips never sit on
//     the scan warps' critical path.
template <typename T>
__global__ void __launch_bounds__(kScanCtaThreads, 1)
scan_kernel(const T *in, T *out, long long n, T seed, const T *__restrict__ seed_ptr,
            unsigned long long *__restrict__ state, unsigned long long *ctl, unsigned int num_tiles) {
  // ctl[0]: launches completed on this scratch, ctl[1]: ticket of CTAs that have finished this launch
  const unsigned long long launches_done = ld_state(ctl);
  const unsigned int epoch = scan_epoch_of(launches_done);
  if (seed_ptr != nullptr) seed += *seed_ptr;  // sharded scan: the exclusive offset computed on the device
  typedef typename V4<T>::type V;
  extern __shared__ __align__(16) unsigned char scan_smem[];
  uint4 *ring = reinterpret_cast<uint4 *>(scan_smem);
  __shared__ T s_warp[2][kScanThreads / 32];
  __shared__ T s_add[4][kScanThreads / 32];  // per tile (mod 4) and scan warp: prefix + warp offset
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool vec_ok = ((((uintptr_t)in | (uintptr_t)out) & 15) == 0);
  const unsigned int G = gridDim.x;
  const unsigned int first_tile = blockIdx.x;
  if (first_tile >= num_tiles) return;
  const int passes = (int)((num_tiles - first_tile + G - 1) / G);  // tiles this CTA owns

  if (warp == kScanThreads / 32) {
    // ===================================== helper warp =====================================
    // Lane w < 16 keeps, for the last few passes, the exclusive offset of scan warp w inside its
    // tile; all lanes keep the tile totals.  Settling tile q leaves  prefix(q) + offset(w)  for
    // every scan warp in s_add[q & 3][w]: the scan warps need one shared-memory read per tile.
    T running = 0;  // inclusive prefix of the last tile whose prefix was settled
    T tot[kGatherLag + 1], off[kGatherLag + 1];
#pragma unroll
    for (int i = 0; i <= kGatherLag; i++) tot[i] = 0, off[i] = 0;
    auto settle = [&](int q, T total_q, T off_q) {
      const unsigned int tile = first_tile + (unsigned int)q * G;
      Gather<T> g;
      const bool first = tile < G;
      gather_issue<T>(g, state, tile, first ? tile : G - 1);
      const T prefix = (first ? seed : running) + gather_finish<T>(g, state, epoch);
      running = prefix + total_q;
      if (lane < kScanThreads / 32) s_add[q & 3][lane] = prefix + off_q;
      __threadfence_block();
      named_bar_arrive(BAR_PREFIX + (q & 3), kScanCtaThreads);
    };
    for (int p = 0; p < passes; p++) {
      named_bar_sync(BAR_TOTALS, kScanCtaThreads);
#pragma unroll
      for (int i = kGatherLag; i > 0; i--) tot[i] = tot[i - 1], off[i] = off[i - 1];
      const T x = (lane < kScanThreads / 32) ? s_warp[p & 1][lane] : (T)0;
      T incl = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      T excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 0;
      tot[0] = __shfl_sync(0xffffffffu, incl, 31);
      off[0] = excl;
      if (lane == 0) st_state(state + first_tile + (unsigned int)p * G, pack_state(to_bits<T>(tot[0]), epoch, ST_AGG));
      if (p >= kGatherLag) settle(p - kGatherLag, tot[kGatherLag], off[kGatherLag]);
    }
    // drain: the last kGatherLag tiles
#pragma unroll
    for (int d = kGatherLag; d > 0; d--) {
      const int q = passes - d;
      if (q >= 0) settle(q, tot[d - 1], off[d - 1]);
    }
    return;
  }

  // ======================================= scan warps =======================================
  auto tile_of = [&](int q) { return first_tile + (unsigned int)q * G; };
  auto is_async = [&](int q) {
    return vec_ok && q < passes && (long long)(tile_of(q) + 1) * kScanTileElems <= n;
  };
  auto slot_ptr = [&](int slot, int j) { return ring + slot_chunk(slot, j); };
  // add the settled prefix to a parked tile and write it out
  auto finish_tile = [&](int q, int slot) {
    named_bar_sync(BAR_PREFIX + (q & 3), kScanCtaThreads);
    const T add = s_add[q & 3][warp];
    TileRegs<T> r;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const V t = *reinterpret_cast<const V *>(slot_ptr(slot, j));
      r.v[j][0] = t.x, r.v[j][1] = t.y, r.v[j][2] = t.z, r.v[j][3] = t.w;
    }
    const unsigned int tile = tile_of(q);
    if (kBulkStore && vec_ok && (long long)(tile + 1) * kScanTileElems <= n) {
      // add the prefix in place, then ONE bulk copy per warp moves its 2 KiB of the tile to global memory:
      // the copy engine streams it out while the warp is already scanning the next tile (128 STG.128 per
      // warp and tile became one instruction).  Only this warp ever touches this part of the slot, so a
      // warp-level barrier is all the ordering the generic-proxy writes need before the async-proxy read.
#pragma unroll
      for (int j = 0; j < 4; j++) {
        V t;
        t.x = r.v[j][0] + add, t.y = r.v[j][1] + add, t.z = r.v[j][2] + add, t.w = r.v[j][3] + add;
        *reinterpret_cast<V *>(slot_ptr(slot, j)) = t;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        const unsigned src = (unsigned)__cvta_generic_to_shared(ring + slot * (kScanTileElems / 4) + warp * 128);
        T *dst = out + (long long)tile * kScanTileElems + warp * 512;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 2048;" ::"l"(dst), "r"(src) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    } else {
      store_tile(r, add, out, n, tile, vec_ok);
    }
  };
  // before a slot is refilled, the bulk copy that emptied it (issued a whole pass earlier) must have read it
  auto slot_reusable = [&]() {
    if (kBulkStore) {
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      __syncwarp();
    }
  };
  auto next_slot = [](int s) { return s + 1 == kScanRing ? 0 : s + 1; };

  // prologue: start the first kPrefetch tiles (one commit group per pass, even when empty, so
  // that wait_group counts passes)
#pragma unroll
  for (int d = 0; d < kPrefetch; d++) {
    if (is_async(d)) prefetch_tile(ring, d, in, tile_of(d));
    asm volatile("cp.async.commit_group;" ::: "memory");
  }

  // ring positions: tile p sits in slot p % kScanRing (kept as counters -- no division in the loop)
  int slot_cur = 0, slot_pf = kPrefetch % kScanRing, slot_st = (kScanRing - kStoreLag) % kScanRing;
  for (int p = 0; p < passes; p++) {
    if (is_async(p + kPrefetch)) {
      slot_reusable();
      prefetch_tile(ring, slot_pf, in, tile_of(p + kPrefetch));
    }
    asm volatile("cp.async.commit_group;" ::: "memory");

    // ---- fetch this tile into registers
    TileRegs<T> cur;
    if (is_async(p)) {
      asm volatile("cp.async.wait_group %0;" ::"n"(kPrefetch) : "memory");
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const V t = *reinterpret_cast<const V *>(slot_ptr(slot_cur, j));
        cur.v[j][0] = t.x, cur.v[j][1] = t.y, cur.v[j][2] = t.z, cur.v[j][3] = t.w;
      }
    } else {  // ragged last tile or unaligned pointers: guarded scalar loads
      const long long base = tile_elem_base(tile_of(p));
#pragma unroll
      for (int j = 0; j < 4; j++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const long long i = base + j * 128 + c;
          cur.v[j][c] = (i < n) ? in[i] : (T)0;
        }
    }

    // ---- scan the tile: thread-local, then a shuffle scan per row j, rows chained through `carry`
    T carry = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      cur.v[j][1] += cur.v[j][0];
      cur.v[j][2] += cur.v[j][1];
      cur.v[j][3] += cur.v[j][2];
      T incl = cur.v[j][3];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      T excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 0;
      const T off = carry + excl;
      V t;
      t.x = cur.v[j][0] + off, t.y = cur.v[j][1] + off, t.z = cur.v[j][2] + off, t.w = cur.v[j][3] + off;
      *reinterpret_cast<V *>(slot_ptr(slot_cur, j)) = t;  // park the warp-local scan in this thread's chunk
      carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 31) s_warp[p & 1][warp] = carry;
    named_bar_sync(BAR_TOTALS, kScanCtaThreads);  // warp totals handed to the helper warp

    if (p >= kStoreLag) finish_tile(p - kStoreLag, slot_st);
    slot_cur = next_slot(slot_cur), slot_pf = next_slot(slot_pf), slot_st = next_slot(slot_st);
  }
  // drain: the last kStoreLag tiles
#pragma unroll
  for (int d = kStoreLag; d > 0; d--) {
    const int q = passes - d;
    if (q >= 0) finish_tile(q, q % kScanRing);
  }
  if (kBulkStore && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // shared memory stays valid until read
  // This CTA's helper warp settled its last tile before the store above could run, so the CTA no
  // longer reads anybody's state words.  The last CTA through here opens the next epoch.
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(ctl + 1, 1ull) == (unsigned long long)gridDim.x - 1) {
      st_state(ctl + 1, 0ull);
      __threadfence();
      st_state(ctl, launches_done + 1);
    }
  }
}

// The CTAs of one scan wait on each other's aggregates, so they must all be resident at once.  A
// cooperative launch makes the driver guarantee that (or refuse): two scans issued on different
// streams are then run one after the other instead of interleaving into a dead wait.
template <typename T>
static int launch_scan(int device, cudaStream_t st, unsigned int grid, const T *in, T *out, long long n, T seed,
                       const T *seed_ptr, unsigned long long *state, unsigned long long *ctl, unsigned int tiles) {
  {
    // once per instantiation and device (the attribute call costs about a microsecond of host time)
    static std::atomic<unsigned long long> smem_set{0};
    const unsigned long long bit = 1ull << (device & 63);
    if (!(smem_set.load(std::memory_order_acquire) & bit)) {
      CUXB_CUDA(cudaFuncSetAttribute(scan_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kScanSmemBytes));
      smem_set.fetch_or(bit, std::memory_order_release);
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(kScanCtaThreads, 1, 1);
  cfg.dynamicSmemBytes = kScanSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CUXB_CUDA(cudaLaunchKernelEx(&cfg, scan_kernel<T>, in, out, n, seed, seed_ptr, state, ctl, tiles));
  g_launches.fetch_add(1);
  return 0;
}

template <typename T>
static int scan_entry(const char *name, cuxb_ctx *ctx, void *stream, long long n, const T *in, T *out,
                      T seed, const T *seed_ptr = nullptr) {
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
      // not freed: a CUDA graph recorded with the smaller workspace must stay replayable
      if ((rc = retire_workspace(ctx, ctx->scan_state))) return rc;
      ctx->scan_state = nullptr;
      ctx->scan_state_tiles = 0;
    }
    // state words + 32 control words (launch counter, finish ticket), all zero: launch 0 uses epoch 1
    CUXB_CUDA(cudaMalloc((void **)&ctx->scan_state, (cap + 32) * sizeof(unsigned long long)));
    CUXB_CUDA(cudaMemset(ctx->scan_state, 0, (cap + 32) * sizeof(unsigned long long)));
    ctx->scan_state_tiles = cap;
    ctx->scan_epoch = 0;
  }
  // The device-side epoch repeats after 2^30 - 1 launches.  A word that old can only survive in a
  // tile slot no launch has touched since; wipe the words every 2^29 calls so it cannot be mistaken
  // for a fresh one (graph replays do not pass through here: see cuxb.h).
  ctx->scan_epoch = (ctx->scan_epoch + 1) & 0x1fffffffu;
  if (ctx->scan_epoch == 0)
    CUXB_CUDA(cudaMemsetAsync(ctx->scan_state, 0, ctx->scan_state_tiles * sizeof(unsigned long long), st));
  long long grid = (long long)ctx->sm_count * kScanCtasPerSM;  // all CTAs resident (see scan_kernel)
  if (grid > tiles) grid = tiles;
  return launch_scan<T>(ctx->device, st, (unsigned int)grid, in, out, n, seed, seed_ptr, ctx->scan_state,
                        ctx->scan_state + ctx->scan_state_tiles, (unsigned int)tiles);
}

// for host_api.cu: a scan whose seed is `seed` plus (when seed_ptr != null) an int32 read on the device
int scan_i32_device_seed(cuxb_ctx *ctx, cudaStream_t st, long long n, const int32_t *in, int32_t *out, int32_t seed,
                         const int32_t *seed_ptr) {
  return scan_entry<unsigned int>("cuxb_scan_i32_host", ctx, (void *)st, n, (const unsigned int *)in,
                                  (unsigned int *)out, (unsigned int)seed, (const unsigned int *)seed_ptr);
}

const void *scan_kernel_symbol() { return (const void *)scan_kernel<unsigned int>; }
int scan_smem_bytes() { return kScanSmemBytes; }

// ---------------------------------------------------------------------------------------------
// Allocation-free variant (cuxb_scan_i32_noalloc).
//
// The scan needs scratch memory besides its input and output: one 64-bit state word per tile.  The entry points above keep that scratch in the cuxb_ctx, which is created with
// cudaMalloc and may grow it.  Some callers cannot do that: a plug-in that runs inside a host
// application which has already handed all device memory to its own pool, a service with a fixed
// memory budget where a late cudaMalloc may fail, or a latency-sensitive loop that cannot afford
// the device-wide synchronisation an allocation implies.  For them the scratch lives in the
// file-scope array below.  It is part of the library image: every GPU on which this library's device
// code gets loaded holds these ~128 KiB until the process exits.  (CUDA loads device code lazily by
// default, so that happens at the first call on the device, not when libcuxb.so is opened -- which is
// why that first call should be made early, while memory is still available; see cuxb.h.)
//
// Size limit: kNoAllocMaxElems elements.  The array holds one state word per 4096 elements of that
// limit; the kernel's tiles are at least that large, so a call uses a prefix of the array.
constexpr long long kNoAllocMaxElems = 1ll << 26;  // 67,108,864 int32 values = 256 MiB of input
constexpr int kNoAllocMaxTiles = (int)(kNoAllocMaxElems / 4096);  // 16384 words = 128 KiB
static_assert(kScanTileElems >= 4096, "the static scratch is sized for tiles of at least 4096 elements");
constexpr int kNoAllocCounterWords = 32;  // 256 bytes after the state words: [0] launch counter, [1] finish ticket
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocMaxTiles + kNoAllocCounterWords];

// Host-side bookkeeping, one slot per device ordinal.  `scratch` is the DEVICE address of
// g_noalloc_scratch on that GPU: in host code the array's name is only a handle for the runtime,
// not something a kernel can dereference, so the address is asked for once per device with
// cudaGetSymbolAddress while that device is current.
constexpr int kNoAllocMaxDevices = 64;
struct NoAllocDevice {
  unsigned long long *scratch;
  unsigned int epoch;
  int sm_count;
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
    CUXB_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
    d.scratch = ptr;
    d.epoch = 0;   // host-side call counter (only for the periodic wipe below)
    d.ready = true;
  }
  // The array still holds whatever the previous call left.  Nothing has to be cleared for that:
  // every state word carries the epoch of the launch that wrote it, every launch runs under a fresh
  // epoch (counted on the device, in the word after the state words), and a word with any other
  // epoch reads as "not published yet".  Only every 2^29 calls are the words wiped (on the caller's
  // stream, ahead of the kernel), long before an epoch value could come round again.
  d.epoch = (d.epoch + 1) & 0x1fffffffu;
  if (d.epoch == 0)
    CUXB_CUDA(cudaMemsetAsync(d.scratch, 0, (size_t)kNoAllocMaxTiles * sizeof(unsigned long long), st));
  const unsigned int tiles = (unsigned int)div_up(n, kScanTileElems);
  unsigned int grid = (unsigned int)d.sm_count * kScanCtasPerSM;  // every CTA resident at once
  if (grid > tiles) grid = tiles;
  return launch_scan<unsigned int>(dev, st, grid, (const unsigned int *)in, (unsigned int *)out, n, (unsigned int)seed,
                                   (const unsigned int *)nullptr, d.scratch, d.scratch + kNoAllocMaxTiles, tiles);
}

// Sharded scan across GPUs (rank r holds the r-th contiguous slice): one reduction kernel whose
// finishing CTA exchanges the per-GPU totals over NVLink and leaves this rank's exclusive offset in
// device memory, then the scan kernel, which reads the offset from there.  Two launches, no host
// synchronisation, no NCCL call.  Replaces MPI_Exscan + local scan
// (mpi/base_reduce_alltoall_scan.cpp:101 in the reference's MPI tutorial).
extern "C" int cuxb_mg_exscan_total_i32(cuxb_ctx *, void *, long long, const int32_t *, int32_t *, int, int,
                                        void *const *);
CUXB_API int cuxb_mg_scan_i32(cuxb_ctx *ctx, void *stream, long long n, const int32_t *in, int32_t *out, int world,
                              int rank, void *const *peer_bufs) {
  CUXB_REQUIRE(ctx != nullptr, CUXB_ERR_INVALID, "cuxb_mg_scan_i32: null ctx");
  int rc = cuxb_mg_exscan_total_i32(ctx, stream, n, in, ctx->mg_seed, world, rank, peer_bufs);
  if (rc) return rc;
  return scan_entry<unsigned int>("cuxb_mg_scan_i32", ctx, stream, n, (const unsigned int *)in, (unsigned int *)out, 0u,
                                  (const unsigned int *)ctx->mg_seed);
}
