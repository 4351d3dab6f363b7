// umma_gemm.cu -- tensor-core GEMM for sm_100a: TMA -> shared memory -> tcgen05.mma
// (accumulators in TMEM) -> tcgen05.ld epilogue.  One kernel template serves
//
//   cuxb_hgemm_f16_f32acc   fp16  A[M,K] x B[K,N]    (cux "gemm", FLOAT16 kernels:
//                           0-frameworks/cux/src/operator/gemm.cu:266-269 cublasSgemmEx;
//                           cuda/gemm_fp16_wmma.cu:123-700 wmma study, fp32 accumulate
//                           as its CPU oracle GemmHostV1 :75-90 does)
//   cuxb_linear_f16/bf16    16-bit A[M,K] x Bt[N,K]^T (simd/x86/linear.cpp:137-196)
//   cuxb_linear_s8_s32      int8  A[M,K] x Bt[N,K]^T -> int32 (linear.cpp:199-237)
//
// Design (see DESIGN.md "tensor-core GEMM"):
//   * persistent grid, one CTA per SM; CTAs are paired into 2-CTA clusters and the
//     pair issues ONE tcgen05.mma.cta_group::2 of 256 x 256 x 16 per step: each CTA
//     holds 128 rows of A and 128 columns of B, so neither operand is loaded twice
//     inside the pair.  (kCtaGroup = 1 is the single-CTA 128 x 256 variant.)
//   * warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane, leader CTA only),
//     warp 2 = TMEM allocator, warps 4..7 = epilogue (one TMEM lane quarter each).
//   * kStages-deep shared-memory ring (128B-swizzled tiles, 64 k-elements = 128 bytes
//     per row) guarded by full/empty mbarriers; the full barrier lives in the leader
//     CTA and receives the byte counts of BOTH CTAs' TMA loads.
//   * two 256-column fp32/int32 accumulators in TMEM (all 512 columns): the epilogue of
//     tile t (tcgen05.ld -> alpha/beta -> convert -> 16-byte global stores) overlaps the
//     mainloop of tile t+1.
//   * B given as [K,N] (cux layout) is consumed as an "MN-major" UMMA operand: TMA boxes
//     of 64 n x 64 k land as 128-byte rows and the descriptor walks them with
//     LBO = 8 KiB (next 64 columns) / SBO = 1 KiB (next 8 k-rows); no transpose pass.
//   * ragged M/N/K: TMA zero-fills out-of-bounds reads, stores are guarded.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <string.h>

#include <type_traits>

#include "common.cuh"
#include "umma_ptx.cuh"

namespace cuxb {

constexpr int UG_BLOCK_M = 128;   // rows of A per CTA
constexpr int UG_BLOCK_N = 256;   // columns per MMA (per CTA pair / per CTA when kCtaGroup==1)
constexpr int UG_ROW_BYTES = 128; // bytes of K per shared-memory row (one swizzle span)
constexpr int UG_THREADS = 256;
constexpr int UG_GROUP_M = 8;

struct UmmaArgs {
  int M, N, K;         // problem
  int ldc;             // elements
  void *C;
  float alpha, beta;
  int tiles_m, tiles_n;  // in units of (UG_BLOCK_M * kCtaGroup) x UG_BLOCK_N
  int k_blocks;
  int group_m;           // tile rows per rasterisation group (L2 reuse)
  int stages, copy_bufs; // shared-memory split of this launch: operand ring depth, broadcast staging buffers
  int tma_epilogue;      // 1: C leaves through shared memory and TMA stores (beta == 0, plain launches); 0: direct stores
  int n_halves;          // 256-column halves per tile: 1 = 256 x 256 tiles, two accumulators (epilogue of tile t overlaps the
                         // mainloop of t+1); 2 = 256 x 512 tiles, one accumulator: A is loaded once per 512 columns
  unsigned int *wave_sync;  // arrival counter of the soft per-tile rendezvous of the TMA producers (null: off)
  // fused B broadcast (multi-GPU, non-root ranks): column panel p of B (= n-block p) may be loaded
  // once b_ready[p] >= ready_target; the CTA's otherwise idle warp 3 pulls B from the root GPU.
  unsigned int *b_ready;
  unsigned int ready_target;
  // chained broadcast (cuxb_mg_set_chain): B is pulled from the previous rank's copy instead of the
  // root's, so every GPU's NVLink egress carries B once however many ranks there are.  Flags live in
  // peer-mapped memory and every rank WAITS only on its own: word p of a rank's flags is written by the
  // rank it pulls from (the call epoch of that rank's last completed fill of panel p), word kChainDone by
  // the rank that pulls from it (the last call that rank has read completely), word kChainEpoch by itself.
  unsigned int *chain_local;        // ours (null: no chain)
  unsigned int *chain_up;           // the rank we pull from (null: its B is complete -- the root)
  unsigned int *chain_down;         // the rank that pulls from us (null: nobody)
  unsigned int *done_ctr;           // local: panels published, grows by tiles_n per call
  unsigned int done_target;
  // NVSwitch multicast all-gather (cuxb_mg_set_multicast), world >= 3.  The root copies nothing.  Each of
  // the other ranks owns every (world-1)-th 16 KiB box of B: its broadcast warps pull THOSE boxes from
  // the root (so the root's NVLink egress carries B once, spread over all receivers) and store them to
  // the multicast address of the ranks' copies -- the switch replicates each store into every rank's
  // copy, the sender's own included.  Every panel is therefore filled by all receivers at once, at the
  // full ingress rate of a GPU, and nobody waits for a neighbour to finish first.  Rank r announces
  // "my share of panel group g of call e has landed everywhere" by a multicast store of e into flag word
  // (r-1)*kMcMaxGroups + g; a TMA producer waits until the world-1 words of its panel's group carry e.
  // mc_mode: 0 off, 1 root (multiplies with its own B, takes no part), 2 receiver.
  int mc_mode;
  unsigned int mc_epoch;            // this call's number on these flags (host-counted, same on every rank)
  unsigned int *mc_flags_local;     // this rank's instance of the flags
  unsigned int *mc_flags_mc;        // the multicast address of the flags
  int mc_world, mc_rank;
  int mc_group;                     // panels per announcement (flag word g covers panels [g*mc_group, (g+1)*mc_group))
  unsigned int *mc_done_ctr;        // local: CTAs that have finished, grows by the grid size per call
  unsigned int mc_done_target;
};
constexpr int kMcMaxGroups = 512;   // flag words per sending rank
constexpr int kMcDoneBase = 4100;    // flag word kMcDoneBase + r: last call whose B rank r has read completely
constexpr int kChainDone = 4096;    // word written by the rank that pulls from us: the last call it has read completely
constexpr int kChainEpoch = 4097;   // our own count of completed calls

template <int CG>
struct UmmaCfg {
  static constexpr int kBRows = UG_BLOCK_N / CG;                    // rows (K-major) / columns (MN-major) of B per CTA
  static constexpr int kABytes = UG_BLOCK_M * UG_ROW_BYTES;         // 16 KiB
  static constexpr int kBBytes = kBRows * UG_ROW_BYTES;             // 16 / 32 KiB
  static constexpr int kStageBytes = kABytes + kBBytes;
  // Shared memory is split at launch time between the operand ring and the staging buffers of the fused
  // broadcast: a plain GEMM runs kStages stages and no staging; a launch that also copies B runs one stage
  // less and kCopyBufs 16 KiB staging buffers -- an NVLink pull is a round trip of several microseconds
  // when the remote GPU is busy, and with two boxes in flight per SM (4.7 MB per GPU) the copy was
  // latency-bound at 635 GB/s on a link that carries 770.
  static constexpr int kStages = (CG == 2) ? 6 : 4;
  static constexpr int kStagesCopying = (CG == 2) ? 5 : 3;
  static constexpr int kCopyBufs = (CG == 2) ? 4 : 2;
  static constexpr int kMaxStages = 8;         // barrier slots reserved per kind
  static constexpr int kCopyBufBytes = 16384;  // one 256 x 32 box of 16-bit B or 256 x 64 of int8 (broadcast staging)
  static constexpr int smem_bytes(int stages, int bufs) { return stages * kStageBytes + 1024 /*align*/ + 1024 /*barriers*/ + bufs * kCopyBufBytes; }
  // wide tiles (two 256-column halves of B per stage): kStagesWide stages of kABytes + 2 * kBBytes
  static constexpr int kStagesWide = (CG == 2) ? 4 : 2;
  static constexpr int smem_bytes_wide() { return kStagesWide * (kABytes + 2 * kBBytes) + 1024 + 1024; }
  static constexpr int kEpiBytes = 4 * 8192;   // TMA-store epilogue: two 4 KiB units per epilogue warp (plain launches only)
  static constexpr int kSmemBytes = smem_bytes(kStages, 0) + kEpiBytes > smem_bytes(kStagesCopying, kCopyBufs)
                                        ? smem_bytes(kStages, 0) + kEpiBytes : smem_bytes(kStagesCopying, kCopyBufs);
  static_assert(kSmemBytes <= 232448, "more shared memory than an sm_100 CTA can have");
};

// out-type helpers ---------------------------------------------------------------------------
template <int CT> struct OutT;
template <> struct OutT<CUXB_FLOAT16> { typedef __half type; };
template <> struct OutT<CUXB_BFLOAT16> { typedef __nv_bfloat16 type; };
template <> struct OutT<CUXB_FLOAT32> { typedef float type; };
template <> struct OutT<CUXB_INT32> { typedef int32_t type; };

__device__ __forceinline__ float out_to_f32(__half v) { return __half2float(v); }
__device__ __forceinline__ float out_to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float out_to_f32(float v) { return v; }
__device__ __forceinline__ void f32_to_out(float f, __half &o) { o = __float2half_rn(f); }
__device__ __forceinline__ void f32_to_out(float f, __nv_bfloat16 &o) { o = __float2bfloat16_rn(f); }
__device__ __forceinline__ void f32_to_out(float f, float &o) { o = f; }

// 8 consecutive outputs of one row: v[0..7] are raw accumulator words
template <int KIND, typename OT>
__device__ __forceinline__ void store8(OT *cp, const uint32_t *v, float alpha, float beta, bool full_cols,
                                       int cols_left) {
  if constexpr (KIND == KIND_S8) {
    int32_t *ip = reinterpret_cast<int32_t *>(cp);
    if (full_cols) {
      *reinterpret_cast<uint4 *>(ip) = make_uint4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<uint4 *>(ip + 4) = make_uint4(v[4], v[5], v[6], v[7]);
    } else {
      for (int c = 0; c < 8 && c < cols_left; c++) ip[c] = (int32_t)v[c];
    }
  } else {
  OT o[8];
  if (beta != 0.f) {
    OT c0[8];
    if (full_cols) {
      if (sizeof(OT) == 2) {
        *reinterpret_cast<uint4 *>(c0) = *reinterpret_cast<const uint4 *>(cp);
      } else {
        *reinterpret_cast<uint4 *>(c0) = *reinterpret_cast<const uint4 *>(cp);
        *reinterpret_cast<uint4 *>(c0 + 4) = *reinterpret_cast<const uint4 *>(cp + 4);
      }
    } else {
      for (int c = 0; c < 8; c++) c0[c] = (c < cols_left) ? cp[c] : OT();
    }
#pragma unroll
    for (int c = 0; c < 8; c++) {
      // two separately rounded products and one add, as the CPU oracle does (no fma contraction)
      const float t0 = __fmul_rn(alpha, __uint_as_float(v[c]));
      const float t1 = __fmul_rn(beta, out_to_f32(c0[c]));
      f32_to_out(__fadd_rn(t0, t1), o[c]);
    }
  } else {
#pragma unroll
    for (int c = 0; c < 8; c++) f32_to_out(__fmul_rn(alpha, __uint_as_float(v[c])), o[c]);
  }
  if (full_cols) {
    if (sizeof(OT) == 2) {
      *reinterpret_cast<uint4 *>(cp) = *reinterpret_cast<const uint4 *>(o);
    } else {
      *reinterpret_cast<uint4 *>(cp) = *reinterpret_cast<const uint4 *>(o);
      *reinterpret_cast<uint4 *>(cp + 4) = *reinterpret_cast<const uint4 *>(o + 4);
    }
  } else {
    for (int c = 0; c < 8 && c < cols_left; c++) cp[c] = o[c];
  }
  }
}

// --------------------------------------------------------------------------------- the kernel
template <int KIND, bool B_MN, int CT, int CG>
__global__ void __launch_bounds__(UG_THREADS, 1)
umma_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_bsrc, const __grid_constant__ CUtensorMap map_bdst,
                 const __grid_constant__ CUtensorMap map_c, const UmmaArgs p) {
  typedef UmmaCfg<CG> Cfg;
  typedef typename OutT<CT>::type OT;
  const int kStages = p.stages;          // operand ring depth of this launch (see UmmaCfg)
  const int kBufs = p.copy_bufs;         // staging buffers of the broadcast warp
  constexpr int kElemBytes = (KIND == KIND_S8) ? 1 : 2;
  constexpr int kBlockK = UG_ROW_BYTES / kElemBytes;  // 64 or 128 elements
  constexpr int kMmaK = 32 / kElemBytes;              // 16 or 32 elements = 32 bytes
  constexpr int kMmaPerStage = kBlockK / kMmaK;       // 4
  constexpr uint32_t kIdesc = make_idesc<KIND, B_MN>(UG_BLOCK_M * CG, UG_BLOCK_N);
  constexpr uint32_t kTmemCols = 512;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int NH = p.n_halves;                                    // 256-column halves per tile
  const uint32_t stage_bytes = Cfg::kABytes + (uint32_t)NH * Cfg::kBBytes;
  const int tile_n = UG_BLOCK_N * NH;
  const int num_acc = NH == 2 ? 1 : 2;                          // TMEM accumulators in rotation
  const uint32_t bar_base = smem_base + kStages * stage_bytes;
  // barrier layout (8 bytes each): full[kStages], empty[kStages], tmem_full[2], tmem_empty[2], tmem ptr
  const uint32_t full_bar = bar_base, empty_bar = bar_base + 8 * Cfg::kMaxStages;
  const uint32_t tfull_bar = bar_base + 16 * Cfg::kMaxStages, tempty_bar = tfull_bar + 16;
  const uint32_t tmem_slot = tempty_bar + 16;
  const uint32_t copy_bar = bar_base + 512;           // one mbarrier per broadcast staging buffer
  const uint32_t copy_buf = bar_base + 1024;          // kBufs x kCopyBufBytes, 1024-byte aligned
  uint8_t *smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t *tmem_slot_ptr = reinterpret_cast<volatile uint32_t *>(smem_gen + (tmem_slot - smem_base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;

  if (warp == 0 && elect_one()) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < kStages; s++) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    for (int a = 0; a < 2; a++) {
      mbar_init(tfull_bar + 8 * a, 1);
      mbar_init(tempty_bar + 8 * a, 4 * CG);  // one arrival per epilogue warp of every CTA in the pair
    }
    for (int a = 0; a < kBufs; a++) mbar_init(copy_bar + 8 * a, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<CG>(tmem_slot, kTmemCols);
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int num_clusters = gridDim.x / CG;
  const int cluster_id = blockIdx.x / CG;
  const int total_tiles = p.tiles_m * p.tiles_n;
  const int per_group = p.group_m * p.tiles_n;

  auto tile_coord = [&](int t, int &tm, int &tn) {
    const int group = t / per_group;
    const int first_m = group * p.group_m;
    const int gsz = min(p.tiles_m - first_m, p.group_m);
    tm = first_m + (t % per_group) % gsz;
    tn = (t % per_group) / gsz;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs of a pair; one elected lane) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int wave = 0;
      for (int t = cluster_id; t < total_tiles; t += num_clusters, wave++) {
        int tm, tn;
        tile_coord(t, tm, tn);
        const int m0 = (tm * CG + (int)cta_rank) * UG_BLOCK_M;
        const int n0 = tn * tile_n + (int)cta_rank * Cfg::kBRows;   // (+ 256 per half)
        if (p.wave_sync != nullptr && wave > 0 && is_leader) {
          // Soft rendezvous: the clusters of a wave share operand panels through the L2, which only
          // works while they read the same k-slices at about the same time.  Wait (a few microseconds
          // at most -- this is a hint, never a dependency) until every cluster has reached this wave.
          const long long full = (long long)num_clusters * wave;
          const long long target = ((long long)(wave + 1) * num_clusters <= total_tiles)
                                       ? full : full - num_clusters + (total_tiles - (long long)wave * num_clusters);
          atomicAdd(p.wave_sync, 1u);
          const unsigned long long t0 = global_timer_ns_();
          while ((long long)ld_acquire_gpu_u32(p.wave_sync) < target && global_timer_ns_() - t0 < 4000ull) {}
        }
        if (p.mc_mode == 2) {
          // multicast all-gather: this panel is complete here once every receiving rank (ourselves
          // included) has announced its share of the panel's group for this call
          for (int o = 0; o < p.mc_world - 1; o++)
            wait_sys_reached(p.mc_flags_local + o * kMcMaxGroups + tn / p.mc_group, p.mc_epoch);
          asm volatile("fence.proxy.async;" ::: "memory");
        } else if (p.b_ready != nullptr && p.mc_mode == 0) {
          // this column panel of B is being pulled from the root GPU: wait until every copy warp has
          // delivered its share, then order the TMA (async-proxy) reads behind that observation
          const unsigned long long t0 = global_timer_ns_();
          while (ld_acquire_gpu_u32(p.b_ready + tn) < p.ready_target)
            if (global_timer_ns_() - t0 > 20000000000ull) __trap();
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        for (int kb = 0; kb < p.k_blocks; kb++) {
          mbar_wait(empty_bar + 8 * stage, phase ^ 1);
          const uint32_t sa = smem_base + stage * stage_bytes;
          const uint32_t fb = full_bar + 8 * stage;
          if (is_leader) mbar_expect_tx(fb, stage_bytes * CG);
          const int k0 = kb * kBlockK;
          tma_load_2d<CG>(sa, &map_a, fb, k0, m0);
          auto load_b = [&](auto width) {
            constexpr int kHalves = decltype(width)::value;
#pragma unroll
            for (int h = 0; h < kHalves; h++) {
              const uint32_t sb = sa + Cfg::kABytes + h * Cfg::kBBytes;
              const int nh0 = n0 + h * UG_BLOCK_N;
              if (B_MN) {
                // B[K,N]: boxes of 64 n x kBlockK k, one per 64 columns
#pragma unroll
                for (int c = 0; c < Cfg::kBRows / 64; c++)
                  tma_load_2d<CG>(sb + c * (kBlockK * UG_ROW_BYTES), &map_b, fb, nh0 + c * 64, k0);
              } else {
                tma_load_2d<CG>(sb, &map_b, fb, k0, nh0);
              }
            }
          };
          if (NH == 1) load_b(std::integral_constant<int, 1>{});
          else load_b(std::integral_constant<int, 2>{});
          if (++stage == kStages) stage = 0, phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA, one elected lane) =====================
    if (is_leader && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = cluster_id; t < total_tiles; t += num_clusters) {
        mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * UG_BLOCK_N;
        for (int kb = 0; kb < p.k_blocks; kb++) {
          mbar_wait(full_bar + 8 * stage, phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * stage_bytes;
          // (the tile width is a launch parameter, but this loop is the issue rate of the tensor pipe: it is
          // written out for both widths with compile-time bounds -- a run-time inner loop cost narrow tiles 8 %)
          auto issue_block = [&](auto width) {
            constexpr int kHalves = decltype(width)::value;
#pragma unroll
            for (int k = 0; k < kMmaPerStage; k++) {
              // K-major: 32 bytes further along the 128-byte swizzled row;
              // MN-major: kMmaK k-rows of 128 bytes further down.
              const uint64_t adesc = make_smem_desc(sa + k * 32, 16, 1024);
#pragma unroll
              for (int h = 0; h < kHalves; h++) {   // wide tiles: the same A slice against both halves of B
                const uint32_t sb = sa + Cfg::kABytes + h * Cfg::kBBytes;
                const uint64_t bdesc = B_MN ? make_smem_desc(sb + k * (kMmaK * UG_ROW_BYTES), kBlockK * UG_ROW_BYTES, 1024)
                                            : make_smem_desc(sb + k * 32, 16, 1024);
                umma<CG, KIND>(tmem_d + h * UG_BLOCK_N, adesc, bdesc, kIdesc, (kb | k) != 0);
              }
            }
          };
          if (NH == 1) issue_block(std::integral_constant<int, 1>{});
          else issue_block(std::integral_constant<int, 2>{});
          umma_commit<CG>(empty_bar + 8 * stage);  // frees the smem slot (in both CTAs) when the MMAs retire
          if (++stage == kStages) stage = 0, phase ^= 1;
        }
        umma_commit<CG>(tfull_bar + 8 * acc);      // accumulator complete -> epilogue (both CTAs)
        if (++acc == num_acc) acc = 0, acc_phase ^= 1;
      }
    }
    __syncwarp();
  } else if (warp == 3) {
    // ===================== broadcast warp (multi-GPU, non-root ranks only) =====================
    // Pulls B from the root GPU's memory over NVLink into the local copy, panel by panel in the
    // order the tiles consume them (panel = the 256 output columns of one n-block): TMA load (peer
    // global -> shared), TMA store (shared -> local global), two 16 KiB boxes in flight.  B[K,N] (B_MN): a
    // panel is 256 columns x K rows, boxes of 256 columns x 32 rows.  Bt[N,K] (K-major): a panel is
    // 256 rows x K columns, boxes of 256 k x kPullRows rows.  Box c of a panel belongs to CTA
    // (c mod grid).  When this CTA's share of a panel has landed it bumps the panel's counter (release).
    if (p.b_ready != nullptr && elect_one()) {
      constexpr int kPullRows = (KIND == KIND_S8) ? 64 : 32;
      constexpr uint32_t kPullBytes = 256u * kPullRows * (KIND == KIND_S8 ? 1 : 2);
      static_assert(kPullBytes <= (uint32_t)Cfg::kCopyBufBytes, "broadcast staging buffer too small");
      constexpr int kRowBoxes = UG_BLOCK_N / kPullRows;          // K-major only: row chunks per panel
      const int boxes_per_panel = B_MN ? (p.K + kPullRows - 1) / kPullRows : ((p.K + 255) / 256) * kRowBoxes;
      const int G = gridDim.x, me = blockIdx.x;
      // Which boxes of a panel are ours?  Pull / chain: every box, so box c belongs to CTA (c mod grid).
      // Multicast all-gather: this RANK owns boxes c = ro + nr * j (ro = rank - 1, nr = world - 1), and of
      // those, j = me, me + G, ... belong to this CTA.
      const int nr = p.mc_mode == 2 ? p.mc_world - 1 : 1, ro = p.mc_mode == 2 ? p.mc_rank - 1 : 0;
      const int rank_boxes = (boxes_per_panel - ro + nr - 1) / nr;
      const int mine_per_panel = (rank_boxes - me + G - 1) / G;
      const long long total = (long long)mine_per_panel * p.tiles_n;
      auto coords = [&](long long i, int &panel, int &x, int &y) {
        panel = (int)(i / mine_per_panel);
        const int c = ro + nr * (me + (int)(i % mine_per_panel) * G);
        if (B_MN) x = panel * UG_BLOCK_N, y = c * kPullRows;
        else x = (c / kRowBoxes) * 256, y = panel * UG_BLOCK_N + (c % kRowBoxes) * kPullRows;
      };
      // chain: our call epoch is one more than the last call we completed (our own counter word).  Nothing
      // of the previous call may be overwritten before the rank that pulls from us has read all of it:
      // it says so by writing the call's epoch into word kChainDone of OUR flags.
      unsigned int epoch = 0;
      if (p.chain_local != nullptr) {
        epoch = ld_acquire_sys_u32(p.chain_local + kChainEpoch) + 1;
        if (p.chain_down != nullptr) wait_sys_reached(p.chain_local + kChainDone, epoch - 1);
      }
      if (p.mc_mode == 2) {
        // our stores land in EVERY rank's copy: nothing of the previous call may be overwritten before
        // every receiver has read all of it (we ourselves have: calls of one context run in stream order)
        for (int r = 1; r < p.mc_world; r++)
          if (r != p.mc_rank) wait_sys_reached(p.mc_flags_local + kMcDoneBase + r, p.mc_epoch - 1);
      }
      int up_known = 0;  // panels of the upstream copy known to be complete
      // Every flag a rank waits on lives in its OWN memory: the rank we pull from writes "panel p of call e
      // is complete here" into word p of our flags (a store over NVLink that nobody waits for), so a look at
      // it is a local load, not an NVLink round trip -- cheap enough to look before every load is issued and
      // to go on with other work (the stores of boxes already in flight) when the panel is not there yet.
      const bool has_up = p.chain_local != nullptr && p.chain_up != nullptr;
      auto up_ready = [&](long long i) -> bool {
        const int panel = (int)(i / mine_per_panel);
        if (!has_up || panel < up_known) return true;
        if ((int)(ld_acquire_sys_u32(p.chain_local + panel) - epoch) < 0) return false;
        asm volatile("fence.proxy.async;" ::: "memory");
        up_known = panel + 1;
        return true;
      };
      auto issue = [&](long long i) {
        int panel, x, y;
        coords(i, panel, x, y);
        const uint32_t b = (uint32_t)(i % kBufs);
        const uint32_t bar = copy_bar + 8 * b, buf = copy_buf + Cfg::kCopyBufBytes * b;
        mbar_expect_tx(bar, kPullBytes);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(buf), "l"(&map_bsrc), "r"(bar), "r"(x), "r"(y) : "memory");
      };
      // this CTA's share of `panel` is in place: count it; whoever completes the panel tells the next rank
      // (gpu-scope release per arrival; only the one that completes a panel pays for the sys-scope
      // store -- its acquire of the counter carries the other CTAs' writes along.  The atomic's result
      // is looked at one box later (settle), so its round trip to L2 is not on the copy loop's path.)
      unsigned int pend_old = 0;
      int pend_panel = -1;
      auto settle = [&]() {
        if (pend_panel < 0) return;
        if (pend_old + 1 == p.ready_target) {
          // the panel is complete here: tell the rank that pulls from us (a store into ITS flags)
          if (p.chain_down != nullptr) st_release_sys_u32(p.chain_down + pend_panel, epoch);
          if (atom_add_acq_rel_gpu(p.done_ctr, 1u) + 1 == p.done_target) {
            // every panel is here, so we have read all we need from upstream: count the call, and tell
            // the rank we pulled from that it may overwrite its copy
            st_release_sys_u32(p.chain_local + kChainEpoch, epoch);
            if (p.chain_up != nullptr) st_release_sys_u32(p.chain_up + kChainDone, epoch);
          }
        }
        pend_panel = -1;
      };
      auto arrive = [&](int panel) {
        if (p.mc_mode == 2) {
          // our stores of this group are complete (wait_group 0 above); order them, system-wide, before
          // the count -- whoever completes the count announces this rank's share to every rank at once
          const int grp = panel / p.mc_group;   // `panel` is the last panel of its group
          asm volatile("fence.acq_rel.sys;" ::: "memory");
          if (atom_add_acq_rel_gpu(p.b_ready + grp, 1u) + 1 == p.ready_target) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            asm volatile("multimem.st.release.sys.global.u32 [%0], %1;"
                         ::"l"(p.mc_flags_mc + (p.mc_rank - 1) * kMcMaxGroups + grp), "r"(p.mc_epoch) : "memory");
          }
          return;
        }
        if (p.chain_local == nullptr) {
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.b_ready + panel) : "memory");
          return;
        }
        settle();
        pend_old = atom_add_acq_rel_gpu(p.b_ready + panel, 1u);
        pend_panel = panel;
      };
      int pend_arrive = -1;   // panel whose last store is committed but not yet confirmed
      // Up to kBufs boxes are in flight.  Box j uses buffer j mod kBufs, free again once the store of box
      // j - kBufs has read it -- checked one store late (wait_group.read 1), so that this thread never sits
      // waiting for the store it has just issued.  A load is issued only when its panel is complete upstream.
      long long ni = 0;   // next box to issue
      auto pump = [&](long long upto) {
        while (ni < total && ni < upto && up_ready(ni)) issue(ni), ni++;
      };
      pump(kBufs);
      for (long long i = 0; i < total; i++) {
        int panel, x, y;
        coords(i, panel, x, y);
        const uint32_t b = (uint32_t)(i % kBufs);
        const uint32_t bar = copy_bar + 8 * b, buf = copy_buf + Cfg::kCopyBufBytes * b;
        if (ni <= i && (pend_arrive >= 0 || pend_panel >= 0)) {
          // Announcements are normally made one box late (see below).  Here nothing of ours is in flight and
          // we are about to wait for the upstream rank: a downstream rank may be waiting for exactly the
          // panel whose announcement is pending, so make it first.
          if (pend_arrive >= 0) {
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async;" ::: "memory");
            __threadfence();
            arrive(pend_arrive);
            pend_arrive = -1;
          }
          settle();
        }
        if (ni <= i) {
          // nothing of ours is in flight and the next panel is not complete upstream: now we do have to wait
          const unsigned long long t0 = global_timer_ns_();
          while (!up_ready(i))
            if (global_timer_ns_() - t0 > 20000000000ull) __trap();
          issue(i), ni = i + 1;
        }
        mbar_wait(bar, (uint32_t)((i / kBufs) & 1));
        settle();
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                     ::"l"(&map_bdst), "r"(buf), "r"(x), "r"(y) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        // The store that completed a panel one box ago has had a whole box time to land: confirm it now
        // (all but the newest bulk group complete) and announce the panel.  Waiting for it right after its
        // commit instead would leave this thread idle for a full store round trip once per panel -- with
        // 3-4 boxes per panel and CTA that cost a quarter of the copy rate.
        if (pend_arrive >= 0) {
          asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");
          asm volatile("fence.proxy.async;" ::: "memory");
          __threadfence();
          arrive(pend_arrive);
          pend_arrive = -1;
        }
        // When does this CTA tell the others that its share has landed?  Pull: after every panel.  Multicast
        // all-gather: after every mc_group panels (each announcement is a multicast round trip).
        bool last_of_panel = ((i + 1) % mine_per_panel) == 0;
        if (p.mc_mode == 2 && last_of_panel)
          last_of_panel = ((panel + 1) % p.mc_group) == 0 || panel + 1 == p.tiles_n;
        if (last_of_panel) pend_arrive = panel;
        if (ni < total) {
          // every store but the one just committed has read its buffer: buffers up to box i - 1's are free
          asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          pump(i + kBufs);
        }
      }
      if (pend_arrive >= 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // every store is complete
        asm volatile("fence.proxy.async;" ::: "memory");
        __threadfence();
        arrive(pend_arrive);
      }
      // CTAs that own no box of a panel still have to be counted
      if (mine_per_panel == 0)
        for (int panel = 0; panel < p.tiles_n; panel++)
          if (p.mc_mode != 2 || (panel + 1) % p.mc_group == 0 || panel + 1 == p.tiles_n) arrive(panel);
      settle();
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> global =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    int acc = 0;
    uint32_t acc_phase = 0;
    OT *C = reinterpret_cast<OT *>(p.C);
    const bool use_tma_epi = p.tma_epilogue != 0;
    const uint32_t epi_buf = bar_base + 1024 + (uint32_t)(warp - 4) * 8192u;   // two 4 KiB units per epilogue warp
    int unit_buf = 0;
    for (int t = cluster_id; t < total_tiles; t += num_clusters) {
      int tm, tn;
      tile_coord(t, tm, tn);
      const int row = (tm * CG + (int)cta_rank) * UG_BLOCK_M + q * 32 + lane;
      const int n0 = tn * tile_n;
      mbar_wait(tfull_bar + 8 * acc, acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * UG_BLOCK_N;
#pragma unroll 1
      for (int cb = 0; cb < tile_n / 32; cb++) {
        uint32_t v[32];
        tmem_ld32(taddr + cb * 32, v);
        tmem_ld_wait();
        if (cb == tile_n / 32 - 1) {
          // every word of this accumulator is now in registers: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CG == 2) mbar_arrive_leader(tempty_bar + 8 * acc);
            else mbar_arrive_local(tempty_bar + 8 * acc);
          }
        }
        const int col = n0 + cb * 32;
        if (use_tma_epi) {
          // C leaves through shared memory: every lane writes its row's 32 values of this chunk into a
          // 32-row x 128-byte unit (128-byte swizzle: 16-byte piece j of row r sits at piece j ^ (r & 7), so the
          // 32 lanes' STS.128 are conflict-free), and when a unit is full one lane hands it to the TMA, which
          // writes 32 full 128-byte rows -- instead of every lane scattering 16-byte stores over 32 rows.
          constexpr int kChunksPerUnit = (sizeof(OT) == 2) ? 2 : 1;   // 64 or 32 columns per unit
          const int cu = cb % kChunksPerUnit;
          const uint32_t ubuf = epi_buf + (uint32_t)(unit_buf * 4096);
          if (cu == 0) {
            // the store that last read this buffer (two units ago) must have finished reading it
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
          }
          const uint32_t rowaddr = ubuf + (uint32_t)lane * 128u;
          if constexpr (sizeof(OT) == 2) {
            alignas(16) OT o[32];
#pragma unroll
            for (int c = 0; c < 32; c++) f32_to_out(__fmul_rn(p.alpha, __uint_as_float(v[c])), o[c]);
            const uint4 *o4 = reinterpret_cast<const uint4 *>(o);
#pragma unroll
            for (int j = 0; j < 4; j++) {
              const uint32_t piece = (uint32_t)((cu * 4 + j) ^ (lane & 7));
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + piece * 16u), "r"(o4[j].x), "r"(o4[j].y), "r"(o4[j].z), "r"(o4[j].w) : "memory");
            }
          } else {
            uint32_t o[32];
#pragma unroll
            for (int c = 0; c < 32; c++) o[c] = (KIND == KIND_S8) ? v[c] : __float_as_uint(__fmul_rn(p.alpha, __uint_as_float(v[c])));
#pragma unroll
            for (int j = 0; j < 8; j++) {
              const uint32_t piece = (uint32_t)(j ^ (lane & 7));
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + piece * 16u), "r"(o[4 * j]), "r"(o[4 * j + 1]), "r"(o[4 * j + 2]), "r"(o[4 * j + 3]) : "memory");
            }
          }
          if (cu == kChunksPerUnit - 1) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            const int ucol = col - cu * 32, urow = row - lane;
            if (lane == 0 && urow < p.M && ucol < p.N) {   // (the TMA clips a unit that sticks out of C)
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                           ::"l"(&map_c), "r"(ubuf), "r"(ucol), "r"(urow) : "memory");
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            unit_buf ^= 1;
          }
        } else if (row < p.M && col < p.N) {
          OT *cp = C + (size_t)row * p.ldc + col;
#pragma unroll
          for (int g = 0; g < 4; g++) {
            const int left = p.N - (col + g * 8);
            if (left > 0) store8<KIND, OT>(cp + g * 8, v + g * 8, p.alpha, p.beta, left >= 8, left);
          }
        }
      }
      if (++acc == num_acc) acc = 0, acc_phase ^= 1;
    }
    if (use_tma_epi && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // C is written before the CTA leaves
  }

  // teardown
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<CG>(tmem_base, kTmemCols);
  if (p.mc_mode == 2 && threadIdx.x == 0) {
    // every tile of this CTA has been multiplied, so all of its loads of B are done.  The last CTA of
    // the launch tells the other receivers that this rank is finished with the call's B (they may then overwrite it).
    if (atom_add_acq_rel_gpu(p.mc_done_ctr, 1u) + 1 == p.mc_done_target)
      asm volatile("multimem.st.release.sys.global.u32 [%0], %1;" ::"l"(p.mc_flags_mc + kMcDoneBase + p.mc_rank), "r"(p.mc_epoch) : "memory");
  }
}

// ------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = (EncodeTiledFn)p;
  return fn;
}

// 2-D row-major tensor [rows, cols] of `ebytes`-byte elements, row pitch ld elements;
// box = box_cols x box_rows, 128-byte swizzle.  Descriptors are cached per context: a cux-style
// harness calls the same GEMM on the same buffers over and over (warm-up + timed run, gemm.cpp:534-541).
static int make_map(cuxb_ctx *ctx, CUtensorMap *map, const void *base, int ebytes, CUtensorMapDataType dt,
                    uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols, uint32_t box_rows,
                    bool swizzle128 = true, bool must_be_local = false) {
  static_assert(sizeof(CUtensorMap) == 128, "descriptor size");
  {
    std::lock_guard<std::mutex> g(*ctx->tma_mu);
    for (int i = 0; i < ctx->tma_cache_n; i++) {
      const TmaCacheEntry &e = ctx->tma_cache[i];
      if (e.base == base && e.rows == rows && e.cols == cols && e.ld == ld && e.box_cols == box_cols &&
          e.box_rows == box_rows && e.dtype == (int)dt * 2 + (swizzle128 ? 1 : 0)) {
        memcpy(map, e.map, sizeof(CUtensorMap));
        return 0;
      }
    }
  }
  if (must_be_local) {
    // The operand loads of the mainloop are issued by CTA pairs (cp.async.bulk.tensor ... cta_group::2, and
    // any future cluster-multicast variant): TMA multicast loads here read only this GPU's memory, since on
    // B200 a TMA multicast load from another GPU's memory hung the GPU.  Operands that live on a peer are
    // brought over first (the fused broadcast does that with plain, per-CTA TMA loads) -- so a matrix given
    // to the mainloop must be this device's own.  Checked once per descriptor (cache miss), not per launch.
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, base) == cudaSuccess && at.type == cudaMemoryTypeDevice)
      CUXB_REQUIRE(at.device == ctx->device, CUXB_ERR_UNSUPPORTED,
                   "GEMM operand %p lives on GPU %d, not on this context's GPU %d: operands of the tensor-core mainloop "
                   "must be local (use the cuxb_mg_* entry points to bring B over)", base, at.device, ctx->device);
    else
      cudaGetLastError();
  }
  EncodeTiledFn enc = get_encode();
  CUXB_REQUIRE(enc != nullptr, CUXB_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * (uint64_t)ebytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, dt, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUXB_REQUIRE(r == CUDA_SUCCESS, CUXB_ERR_INVALID,
               "cuTensorMapEncodeTiled failed (%d): base=%p rows=%llu cols=%llu ld=%llu box=%ux%u", (int)r, base,
               (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_cols, box_rows);
  std::lock_guard<std::mutex> g(*ctx->tma_mu);
  TmaCacheEntry &e = ctx->tma_cache[ctx->tma_cache_next];
  e.base = base, e.rows = rows, e.cols = cols, e.ld = ld, e.box_cols = box_cols, e.box_rows = box_rows, e.dtype = (int)dt * 2 + (swizzle128 ? 1 : 0);
  memcpy(e.map, map, sizeof(CUtensorMap));
  ctx->tma_cache_next = (ctx->tma_cache_next + 1) % 16;
  if (ctx->tma_cache_n < 16) ctx->tma_cache_n++;
  return 0;
}


template <int KIND, bool B_MN, int CT, int CG>
static int launch_umma(cuxb_ctx *ctx, cudaStream_t st, int M, int N, int K, float alpha, const void *A, int lda,
                       const void *B, int ldb, float beta, void *C, int ldc, const void *b_pull_from = nullptr,
                       int mc_role = 0) {
  typedef UmmaCfg<CG> Cfg;
  constexpr int ebytes = (KIND == KIND_S8) ? 1 : 2;
  constexpr int kBlockK = UG_ROW_BYTES / ebytes;
  const CUtensorMapDataType dt = (KIND == KIND_S8)    ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                 : (KIND == KIND_F16) ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                      : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMap ma, mb;
  int rc = make_map(ctx, &ma, A, ebytes, dt, (uint64_t)M, (uint64_t)K, (uint64_t)lda, kBlockK, UG_BLOCK_M, true, true);
  if (rc) return rc;
  if (B_MN)
    rc = make_map(ctx, &mb, B, ebytes, dt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, 64, kBlockK, true, true);
  else
    rc = make_map(ctx, &mb, B, ebytes, dt, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, kBlockK, Cfg::kBRows, true, true);
  if (rc) return rc;

  UmmaArgs p;
  p.stages = Cfg::kStages, p.copy_bufs = 0;
  p.M = M, p.N = N, p.K = K, p.ldc = ldc, p.C = C, p.alpha = alpha, p.beta = beta;
  p.tiles_m = (int)div_up(M, UG_BLOCK_M * CG);
  // 256 x 512 tiles (see UmmaArgs::n_halves): CTA pairs only, and not in launches that also copy B
  // A wide tile reads A once per 512 columns (a quarter less shared-memory fill per flop), but it owns all of
  // TMEM, so its epilogue is not hidden behind the next tile's mainloop, and it halves the number of tiles the
  // waves are cut from.  Measured against the 256 x 256 tiles with both issue loops fully unrolled
  // (tools/epilogue_sweep.py, 10 shapes, fp16 and int8): never faster by more than 2 % (16384^3), 5-25 % slower
  // below that -- so it is a tuning hook (cuxb_debug_ctx_set what = 5), not a default.
  int nh = 1;
  if (CG == 2 && b_pull_from == nullptr && mc_role == 0 && ctx->dbg_wide == 2) nh = 2;
  p.n_halves = nh;
  p.tiles_n = (int)div_up(N, UG_BLOCK_N * nh);
  p.k_blocks = (int)div_up(K, kBlockK);
  p.group_m = ctx->dbg_group_m > 0 ? ctx->dbg_group_m : UG_GROUP_M;
  p.wave_sync = nullptr;
  p.b_ready = nullptr, p.ready_target = 0;
  p.chain_local = nullptr, p.chain_up = nullptr, p.chain_down = nullptr, p.done_ctr = nullptr, p.done_target = 0;
  const int total = p.tiles_m * p.tiles_n;
  int clusters = ctx->sm_count / CG;
  if (clusters > total) clusters = total;

  p.mc_mode = 0, p.mc_epoch = 0, p.mc_flags_local = nullptr, p.mc_flags_mc = nullptr;
  p.mc_world = 0, p.mc_rank = 0, p.mc_group = 1, p.mc_done_ctr = nullptr, p.mc_done_target = 0;
  CUtensorMap msrc = mb, mdst = mb;  // unused unless B is pulled from another GPU
  if (mc_role == 2) {
    // NVSwitch multicast all-gather, receiving rank: B = this rank's copy (what the GEMM reads),
    // b_pull_from = the root's B as mapped here (what our share is pulled from), ctx->mc_b = the multicast
    // address of the ranks' copies (where our share is stored to).
    CUXB_REQUIRE(p.tiles_n <= 4096, CUXB_ERR_TOO_LARGE, "fused B broadcast: N too large");
    CUXB_REQUIRE(b_pull_from != nullptr && ctx->mc_b != nullptr, CUXB_ERR_INVALID,
                 "multicast broadcast: the root's B or the multicast address of the copies is missing");
    if (ctx->bcast_ready == nullptr) {
      CUXB_CUDA(cudaMalloc((void **)&ctx->bcast_ready, (4096 + 32) * sizeof(unsigned int)));
      CUXB_CUDA(cudaMemset(ctx->bcast_ready, 0, (4096 + 32) * sizeof(unsigned int)));
      ctx->bcast_done_total = 0, ctx->bcast_calls = 0, ctx->bcast_grid = 0, ctx->bcast_panels = 0;
    }
    p.mc_mode = 2;
    p.mc_epoch = ++ctx->mc_tab[ctx->mc_cur].calls;
    p.mc_flags_local = ctx->mc_flags_local, p.mc_flags_mc = ctx->mc_flags_mc;
    p.mc_world = ctx->mc_world, p.mc_rank = ctx->mc_rank;
    // about eight announcements per call (the first wave of tiles needs ~ten panels before it can start
    // anyway; every announcement drains the announcing thread's stores through the switch)
    p.mc_group = p.tiles_n / 8 < 1 ? 1 : (p.tiles_n / 8 > 8 ? 8 : p.tiles_n / 8);
    if (ctx->dbg_mc_group > 0) p.mc_group = ctx->dbg_mc_group;
    while ((p.tiles_n + p.mc_group - 1) / p.mc_group > kMcMaxGroups) p.mc_group *= 2;
    const uint64_t rows = B_MN ? (uint64_t)K : (uint64_t)N, cols = B_MN ? (uint64_t)N : (uint64_t)K;
    const uint32_t box_rows = (KIND == KIND_S8) ? 64 : 32;
    rc = make_map(ctx, &msrc, b_pull_from, ebytes, dt, rows, cols, (uint64_t)ldb, 256, box_rows, false);
    if (rc) return rc;
    rc = make_map(ctx, &mdst, ctx->mc_b, ebytes, dt, rows, cols, (uint64_t)ldb, 256, box_rows, false);
    if (rc) return rc;
    // (a different announcement group changes which counters a call touches: encoded into the panel count)
    if (ctx->bcast_grid != clusters * CG || ctx->bcast_panels != p.tiles_n + 65536 * p.mc_group) {
      CUXB_CUDA(cudaMemsetAsync(ctx->bcast_ready, 0, 4096 * sizeof(unsigned int), st));
      ctx->bcast_calls = 0, ctx->bcast_grid = clusters * CG, ctx->bcast_panels = p.tiles_n + 65536 * p.mc_group;
    }
    ctx->bcast_calls++;
    p.b_ready = ctx->bcast_ready;
    p.ready_target = ctx->bcast_calls * (unsigned int)(clusters * CG);
    p.mc_done_ctr = ctx->bcast_ready + 4096 + 8;  // beyond the panel counters, never reset
    ctx->mc_done_total += (unsigned int)(clusters * CG);
    p.mc_done_target = ctx->mc_done_total;
  } else if (b_pull_from != nullptr) {
    CUXB_REQUIRE(p.tiles_n <= 4096, CUXB_ERR_TOO_LARGE, "fused B broadcast: N too large");
    // un-swizzled copy boxes (see the broadcast warp): 256 elements wide, 32 rows (64 for int8) = 16 KiB
    const uint64_t rows = B_MN ? (uint64_t)K : (uint64_t)N, cols = B_MN ? (uint64_t)N : (uint64_t)K;
    const uint32_t box_rows = (KIND == KIND_S8) ? 64 : 32;
    rc = make_map(ctx, &msrc, b_pull_from, ebytes, dt, rows, cols, (uint64_t)ldb, 256, box_rows, false);
    if (rc) return rc;
    rc = make_map(ctx, &mdst, B, ebytes, dt, rows, cols, (uint64_t)ldb, 256, box_rows, false);
    if (rc) return rc;
    if (ctx->bcast_ready == nullptr) {
      CUXB_CUDA(cudaMalloc((void **)&ctx->bcast_ready, (4096 + 32) * sizeof(unsigned int)));
      CUXB_CUDA(cudaMemset(ctx->bcast_ready, 0, (4096 + 32) * sizeof(unsigned int)));
      ctx->bcast_done_total = 0;
      ctx->bcast_calls = 0;
      ctx->bcast_grid = 0;
      ctx->bcast_panels = 0;
    }
    // counters only grow: every copy warp adds 1 to each of the call's panels.  A different grid size
    // or panel count would break the arithmetic, so start over from zero when either changes.
    if (ctx->bcast_grid != clusters * CG || ctx->bcast_panels != p.tiles_n) {
      CUXB_CUDA(cudaMemsetAsync(ctx->bcast_ready, 0, 4096 * sizeof(unsigned int), st));
      ctx->bcast_calls = 0;
      ctx->bcast_grid = clusters * CG;
      ctx->bcast_panels = p.tiles_n;
    }
    ctx->bcast_calls++;
    p.b_ready = ctx->bcast_ready;
    p.ready_target = ctx->bcast_calls * (unsigned int)(clusters * CG);
    if (ctx->chain_local != nullptr) {
      p.chain_local = ctx->chain_local, p.chain_up = const_cast<unsigned int *>(ctx->chain_up), p.chain_down = const_cast<unsigned int *>(ctx->chain_down);
      p.done_ctr = ctx->bcast_ready + 4096;  // beyond the panel counters, never reset
      ctx->bcast_done_total += (unsigned int)p.tiles_n;
      p.done_target = ctx->bcast_done_total;
    }
  }

  // Long problems only (>= 16 waves of tiles with >= 128 k-blocks each): keep the clusters of a wave
  // reading the same k-slices at about the same time (see the producer).  16384^3 fp16: DRAM reads
  // 14-21 GB -> 8.8 GB, -4 % time at fixed clocks, +1 % under the power cap.  Short kernels lose more to
  // the counter's memset and the waits than the L2 hits return (8192^3 int8: 0.341 -> 0.375 ms), so
  // they run without it.
  if (!ctx->dbg_no_wave_sync && total >= 16 * clusters && p.k_blocks >= 128) {
    if (ctx->wave_sync == nullptr) CUXB_CUDA(cudaMalloc((void **)&ctx->wave_sync, 256));
    CUXB_CUDA(cudaMemsetAsync(ctx->wave_sync, 0, 4, st));
    p.wave_sync = ctx->wave_sync;
  }
  auto kern = umma_gemm_kernel<KIND, B_MN, CT, CG>;
  {
    // once per instantiation and device (the attribute call costs about a microsecond of host time)
    static std::atomic<unsigned long long> smem_set{0};
    const unsigned long long bit = 1ull << (ctx->device & 63);
    if (!(smem_set.load(std::memory_order_acquire) & bit)) {
      CUXB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
      smem_set.fetch_or(bit, std::memory_order_release);
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * CG, 1, 1);
  cfg.blockDim = dim3(UG_THREADS, 1, 1);
  // a launch that also copies B trades one operand stage for deeper staging (see UmmaCfg)
  const bool copying = p.b_ready != nullptr;
  p.stages = copying ? Cfg::kStagesCopying : Cfg::kStages;
  p.copy_bufs = copying ? ((ctx->dbg_copy_bufs >= 2 && ctx->dbg_copy_bufs <= Cfg::kCopyBufs) ? ctx->dbg_copy_bufs : Cfg::kCopyBufs) : 0;
  cfg.dynamicSmemBytes = Cfg::smem_bytes(p.stages, p.copy_bufs);
  if (nh == 2) {
    p.stages = Cfg::kStagesWide, p.copy_bufs = 0;
    cfg.dynamicSmemBytes = Cfg::smem_bytes_wide();
  }
  // C through shared memory and TMA stores: when nothing has to be read from C and the launch does not copy B
  // (a copying launch needs that shared memory for its staging buffers)
  CUtensorMap mc_out = mb;
  p.tma_epilogue = 0;
  // ... and when it pays: the staged epilogue is much cheaper per byte of C (no scattered 16-byte stores), but
  // its shared-memory traffic competes with the operand feed of the tensor pipe.  4-byte outputs (int32, fp32):
  // always (+4 ... +70 % across tools/epilogue_sweep.py's shapes); 2-byte outputs: up to K = 4096 (+1 ... +30 %);
  // beyond that the mainloop dominates and the direct stores are as good or up to a few per cent better.
  constexpr int kOutBytes = (CT == CUXB_FLOAT16 || CT == CUXB_BFLOAT16) ? 2 : 4;
  const bool epi_pays = kOutBytes == 4 || p.k_blocks <= 64 || ctx->dbg_epilogue == 2;
  if (!copying && p.b_ready == nullptr && beta == 0.f && ctx->dbg_epilogue != 1 && epi_pays) {
    constexpr int cbytes = (CT == CUXB_FLOAT16 || CT == CUXB_BFLOAT16) ? 2 : 4;
    const CUtensorMapDataType cdt = CT == CUXB_FLOAT16    ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                    : CT == CUXB_BFLOAT16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                    : CT == CUXB_FLOAT32  ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                                          : CU_TENSOR_MAP_DATA_TYPE_INT32;
    rc = make_map(ctx, &mc_out, C, cbytes, cdt, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 128 / cbytes, 32, true, false);
    if (rc) return rc;
    p.tma_epilogue = 1;
    cfg.dynamicSmemBytes += Cfg::kEpiBytes;
  }
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CUXB_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, msrc, mdst, mc_out, p));
  g_launches.fetch_add(1);
  return 0;
}

template <int KIND, bool B_MN, int CT>
static int dispatch_cg(cuxb_ctx *ctx, cudaStream_t st, int M, int N, int K, float alpha, const void *A, int lda,
                       const void *B, int ldb, float beta, void *C, int ldc) {
  const int cg = ctx->dbg_cta_group ? ctx->dbg_cta_group : 2;
  if (cg == 1) return launch_umma<KIND, B_MN, CT, 1>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
  return launch_umma<KIND, B_MN, CT, 2>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
}

static int check_common(const char *name, cuxb_ctx *ctx, int M, int N, int K, const void *A, const void *B,
                        const void *C, int ebytes, int lda, int ldb, int ldc, int cbytes) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(M >= 0 && N >= 0 && K >= 0, CUXB_ERR_INVALID, "%s: negative dimension", name);
  if (M == 0 || N == 0) return -1;  // nothing to do
  CUXB_REQUIRE(K > 0, CUXB_ERR_UNSUPPORTED, "%s: K must be positive", name);
  CUXB_REQUIRE(A && B && C, CUXB_ERR_INVALID, "%s: null pointer", name);
  CUXB_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0 && ((uintptr_t)C & 15) == 0,
               CUXB_ERR_UNSUPPORTED, "%s: A, B and C must be 16-byte aligned (TMA / vector stores)", name);
  CUXB_REQUIRE((lda * ebytes) % 16 == 0 && (ldb * ebytes) % 16 == 0 && (ldc * cbytes) % 16 == 0,
               CUXB_ERR_UNSUPPORTED, "%s: row pitches must be multiples of 16 bytes (lda=%d ldb=%d ldc=%d)", name,
               lda, ldb, ldc);
  return 0;
}

// multicast all-gather role of a cuxb_mg_* call on this context: 0 = off (or the root, which takes no
// part: it multiplies with its own B), 2 = receiving rank
static int mc_role_of(const cuxb_ctx *ctx, int is_root) {
  return (ctx != nullptr && ctx->mc_flags_local != nullptr && !is_root) ? 2 : 0;
}

const void *umma_kernel_symbol() { return (const void *)umma_gemm_kernel<KIND_F16, true, CUXB_FLOAT16, 2>; }
int umma_smem_bytes() { return UmmaCfg<2>::kSmemBytes; }

}  // namespace cuxb

using namespace cuxb;

// Tuning / test hooks (not in cuxb.h).  They act on ONE context -- nothing process-wide is touched, so a
// caller that flips one cannot change the kernels another context launches.
//   what 0: MMA variant, value 0 = default (CTA pair), 1 = single CTA, 2 = CTA pair
//   what 1: rasterisation group height (tile rows per L2 group), 0 = default, 1..64
//   what 2: soft per-wave rendezvous of the TMA producers, value 1 = on (default), 0 = off
//   what 3: multicast broadcast, panels per announcement, 0 = default (about eight announcements per call)
//   what 4: staging buffers of the broadcast warp (boxes in flight per SM), 0 = default, 2..4
//   what 5: tile width, 0 = default, 1 = 256 x 256 tiles, 2 = 256 x 512 tiles
//   what 6: epilogue, 0 = default (through shared memory and TMA stores where it pays), 1 = direct stores, 2 = staged wherever possible
CUXB_API int cuxb_debug_ctx_set(cuxb_ctx *ctx, int what, int value) {
  if (ctx == nullptr) return CUXB_ERR_INVALID;
  if (what == 0 && value >= 0 && value <= 2) return ctx->dbg_cta_group = value, 0;
  if (what == 1 && value >= 0 && value <= 64) return ctx->dbg_group_m = value, 0;
  if (what == 2) return ctx->dbg_no_wave_sync = value ? 0 : 1, 0;
  if (what == 3 && value >= 0 && value <= 64) return ctx->dbg_mc_group = value, 0;
  if (what == 4 && value >= 0 && value <= 4) return ctx->dbg_copy_bufs = value, 0;
  if (what == 5 && value >= 0 && value <= 2) return ctx->dbg_wide = value, 0;
  if (what == 6 && value >= 0 && value <= 2) return ctx->dbg_epilogue = value, 0;
  return CUXB_ERR_INVALID;
}

CUXB_API int cuxb_hgemm_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const void *A,
                                   int lda, const void *B, int ldb, float beta, void *C, int ldc, int c_dtype) {
  CUXB_REQUIRE(c_dtype == CUXB_FLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_hgemm_f16_f32acc: c_dtype must be CUXB_FLOAT16 or CUXB_FLOAT32");
  int rc = check_common("cuxb_hgemm_f16_f32acc", ctx, M, N, K, A, B, C, 2, lda, ldb, ldc,
                        c_dtype == CUXB_FLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID, "cuxb_hgemm_f16_f32acc: leading dimension too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (c_dtype == CUXB_FLOAT16)
    return dispatch_cg<KIND_F16, true, CUXB_FLOAT16>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
  return dispatch_cg<KIND_F16, true, CUXB_FLOAT32>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
}

// bf16 operands in cux's gemm layout (B[K,N]); the reference's TypeFlag has no bf16 (util.h:22-28) --
// this is the bf16 twin of the FLOAT16 gemm kernel, same tcgen05 pipeline with bf16 instruction format.
CUXB_API int cuxb_gemm_bf16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const void *A,
                                   int lda, const void *B, int ldb, float beta, void *C, int ldc, int c_dtype) {
  CUXB_REQUIRE(c_dtype == CUXB_BFLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_gemm_bf16_f32acc: c_dtype must be CUXB_BFLOAT16 or CUXB_FLOAT32");
  int rc = check_common("cuxb_gemm_bf16_f32acc", ctx, M, N, K, A, B, C, 2, lda, ldb, ldc,
                        c_dtype == CUXB_BFLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID, "cuxb_gemm_bf16_f32acc: leading dimension too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (c_dtype == CUXB_BFLOAT16)
    return dispatch_cg<KIND_BF16, true, CUXB_BFLOAT16>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
  return dispatch_cg<KIND_BF16, true, CUXB_FLOAT32>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
}

CUXB_API int cuxb_linear_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A, const void *Bt,
                                    void *C, int c_dtype) {
  CUXB_REQUIRE(c_dtype == CUXB_FLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_linear_f16_f32acc: c_dtype must be CUXB_FLOAT16 or CUXB_FLOAT32");
  int rc = check_common("cuxb_linear_f16_f32acc", ctx, M, N, K, A, Bt, C, 2, K, K, N, c_dtype == CUXB_FLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (c_dtype == CUXB_FLOAT16)
    return dispatch_cg<KIND_F16, false, CUXB_FLOAT16>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N);
  return dispatch_cg<KIND_F16, false, CUXB_FLOAT32>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N);
}

CUXB_API int cuxb_linear_bf16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A,
                                     const void *Bt, void *C, int c_dtype) {
  CUXB_REQUIRE(c_dtype == CUXB_BFLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_linear_bf16_f32acc: c_dtype must be CUXB_BFLOAT16 or CUXB_FLOAT32");
  int rc = check_common("cuxb_linear_bf16_f32acc", ctx, M, N, K, A, Bt, C, 2, K, K, N, c_dtype == CUXB_BFLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (c_dtype == CUXB_BFLOAT16)
    return dispatch_cg<KIND_BF16, false, CUXB_BFLOAT16>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N);
  return dispatch_cg<KIND_BF16, false, CUXB_FLOAT32>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N);
}

CUXB_API int cuxb_linear_s8_s32(cuxb_ctx *ctx, void *stream, int M, int N, int K, const int8_t *A, const int8_t *Bt,
                                int32_t *C) {
  int rc = check_common("cuxb_linear_s8_s32", ctx, M, N, K, A, Bt, C, 1, K, K, N, 4);
  if (rc) return rc < 0 ? 0 : rc;
  return dispatch_cg<KIND_S8, false, CUXB_INT32>(ctx, (cudaStream_t)stream, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N);
}

// Row-sharded fp16 GEMM across GPUs with the B broadcast fused into the kernel: rank r computes
// C_r = alpha * A_r * B + beta * C_r for its own rows.  The root rank reads its B in place; every other
// rank's kernel pulls B from the root's memory over NVLink (b_root is the root's B as mapped into this
// process) into b_local while its mainloop is already working on the panels that have arrived.
// This is the decomposition of the reference's MPI GEMM (mpi/alg_matrix_multiply.cpp:52-74: B sent
// to every worker, row blocks of A and C) with the MPI_Send of B folded into the compute kernel.
// The root must not modify B until every rank's call has completed.
CUXB_API int cuxb_mg_hgemm_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const void *A,
                                      int lda, const void *b_root, void *b_local, int ldb, float beta, void *C,
                                      int ldc, int c_dtype, int is_root) {
  CUXB_REQUIRE(c_dtype == CUXB_FLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_mg_hgemm_f16_f32acc: c_dtype must be CUXB_FLOAT16 or CUXB_FLOAT32");
  const void *B = is_root ? b_root : b_local;
  int rc = check_common("cuxb_mg_hgemm_f16_f32acc", ctx, M, N, K, A, B, C, 2, lda, ldb, ldc,
                        c_dtype == CUXB_FLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  CUXB_REQUIRE(b_root != nullptr && ((uintptr_t)b_root & 15) == 0, CUXB_ERR_INVALID,
               "cuxb_mg_hgemm_f16_f32acc: b_root is null or not 16-byte aligned");
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID, "cuxb_mg_hgemm_f16_f32acc: leading dimension too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int mc = mc_role_of(ctx, is_root);
  const void *pull = is_root ? nullptr : b_root;
  if (c_dtype == CUXB_FLOAT16)
    return launch_umma<KIND_F16, true, CUXB_FLOAT16, 2>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, pull, mc);
  return launch_umma<KIND_F16, true, CUXB_FLOAT32, 2>(ctx, st, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, pull, mc);
}

// Row-sharded int8 linear with the broadcast of Bt fused into the kernel (the same decomposition, for
// simd/x86/linear.cpp:199-211's C = A * Bt^T): Bt[N,K] lives on the root; every other rank's kernel
// pulls it over NVLink, n-block panel by n-block panel, while multiplying the panels that arrived.
CUXB_API int cuxb_mg_linear_s8_s32(cuxb_ctx *ctx, void *stream, int M, int N, int K, const int8_t *A,
                                   const int8_t *bt_root, int8_t *bt_local, int32_t *C, int is_root) {
  const void *Bt = is_root ? (const void *)bt_root : (const void *)bt_local;
  int rc = check_common("cuxb_mg_linear_s8_s32", ctx, M, N, K, A, Bt, C, 1, K, K, N, 4);
  if (rc) return rc < 0 ? 0 : rc;
  CUXB_REQUIRE(bt_root != nullptr && ((uintptr_t)bt_root & 15) == 0, CUXB_ERR_INVALID,
               "cuxb_mg_linear_s8_s32: bt_root is null or not 16-byte aligned");
  const int mc = mc_role_of(ctx, is_root);
  const void *pull = is_root ? nullptr : (const void *)bt_root;
  return launch_umma<KIND_S8, false, CUXB_INT32, 2>(ctx, (cudaStream_t)stream, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N,
                                                    pull, mc);
}

// The same for the 16-bit linears (linear.cpp:137-182's layout): Bt[N,K] pulled in 256-k x 32-row boxes.
CUXB_API int cuxb_mg_linear_bf16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A,
                                        const void *bt_root, void *bt_local, void *C, int c_dtype, int is_root) {
  CUXB_REQUIRE(c_dtype == CUXB_BFLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_mg_linear_bf16_f32acc: c_dtype must be CUXB_BFLOAT16 or CUXB_FLOAT32");
  const void *Bt = is_root ? bt_root : (const void *)bt_local;
  int rc = check_common("cuxb_mg_linear_bf16_f32acc", ctx, M, N, K, A, Bt, C, 2, K, K, N, c_dtype == CUXB_BFLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  CUXB_REQUIRE(bt_root != nullptr && ((uintptr_t)bt_root & 15) == 0, CUXB_ERR_INVALID,
               "cuxb_mg_linear_bf16_f32acc: bt_root is null or not 16-byte aligned");
  const int mc = mc_role_of(ctx, is_root);
  const void *pull = is_root ? nullptr : bt_root;
  cudaStream_t st = (cudaStream_t)stream;
  if (c_dtype == CUXB_BFLOAT16)
    return launch_umma<KIND_BF16, false, CUXB_BFLOAT16, 2>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N, pull, mc);
  return launch_umma<KIND_BF16, false, CUXB_FLOAT32, 2>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N, pull, mc);
}
CUXB_API int cuxb_mg_linear_f16_f32acc(cuxb_ctx *ctx, void *stream, int M, int N, int K, const void *A,
                                       const void *bt_root, void *bt_local, void *C, int c_dtype, int is_root) {
  CUXB_REQUIRE(c_dtype == CUXB_FLOAT16 || c_dtype == CUXB_FLOAT32, CUXB_ERR_INVALID,
               "cuxb_mg_linear_f16_f32acc: c_dtype must be CUXB_FLOAT16 or CUXB_FLOAT32");
  const void *Bt = is_root ? bt_root : (const void *)bt_local;
  int rc = check_common("cuxb_mg_linear_f16_f32acc", ctx, M, N, K, A, Bt, C, 2, K, K, N, c_dtype == CUXB_FLOAT16 ? 2 : 4);
  if (rc) return rc < 0 ? 0 : rc;
  CUXB_REQUIRE(bt_root != nullptr && ((uintptr_t)bt_root & 15) == 0, CUXB_ERR_INVALID,
               "cuxb_mg_linear_f16_f32acc: bt_root is null or not 16-byte aligned");
  const int mc = mc_role_of(ctx, is_root);
  const void *pull = is_root ? nullptr : bt_root;
  cudaStream_t st = (cudaStream_t)stream;
  if (c_dtype == CUXB_FLOAT16)
    return launch_umma<KIND_F16, false, CUXB_FLOAT16, 2>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N, pull, mc);
  return launch_umma<KIND_F16, false, CUXB_FLOAT32, 2>(ctx, st, M, N, K, 1.f, A, K, Bt, K, 0.f, C, N, pull, mc);
}
