// sgemm.cu -- fp32 GEMM on CUDA cores (FFMA), C = alpha*A*B + beta*C, row-major.
//
// Replaces GemmDeviceV0/V1/V2 + cublasSgemm (0-frameworks/cux/src/operator/
// gemm.cu:17-130,237-238) and the GemmKernelv1..v5 study (cuda/gemm_fp32.cu:
// 102-528).  The reference's best kernel is a 32x32 CTA tile with a 4x4 register
// micro-tile and no software pipelining; this one is built for the B200 SM:
//
//   CTA tile 128x128x16, 256 threads (8 warps as 2x4), 8x8 register micro-tile
//   per thread split 4+4 in both directions so that every shared-memory read is
//   a conflict-free LDS.128 at the hardware minimum of 2 wavefronts -- 4 LDS.128
//   (8 wavefronts) feed 64 FFMA.
//   A is stored k-major (transposed) in shared memory, B as it lies.
//   Tile t+1 is fetched while tile t is multiplied: B with cp.async straight into
//   the other shared-memory stage, A with LDG.128 into registers (issued before
//   the FFMA block, transposed into shared memory after it); two stages, one
//   __syncthreads per k-tile.  Register fragments are double-buffered inside the
//   tile.  CTAs are rasterised in groups of 8 tile-rows so that the B panel a
//   wave needs stays in the 126 MB L2.
//   2 CTAs per SM (<=128 registers, 33 KB shared each).
//
// Numerics: acc = fma(a, b, acc) over k in increasing order, then
// C = alpha*acc + beta*C0 (beta == 0: C0 is not read).  With the reference's
// integer-valued fills every partial sum is exact, so the result is bit-identical
// to GemmHostV0 (gemm.cpp:13-34); for general floats see tests/ for the tolerance.
#include "common.cuh"

namespace cuxb {

constexpr int SG_BM = 128, SG_BN = 128, SG_BK = 16;
constexpr int SG_THREADS = 256;
constexpr int SG_LDA = SG_BM + 4;  // padded k-major A rows
constexpr int SG_GROUP_M = 8;

struct SgemmArgs {
  int M, N, K;
  float alpha, beta;
  const float *A;
  int lda;
  const float *B;
  int ldb;
  float *C;
  int ldc;
  int tiles_m, tiles_n;
  // split-K (problems with fewer tiles than the GPU has room for): CTA = (tile, split); split s
  // multiplies k in [s*k_per_split, (s+1)*k_per_split), parks its accumulators in `scratch`, and the
  // last CTA to finish a tile adds the splits in index order and runs the epilogue.
  int splits, k_per_split;
  float4 *scratch;        // [tile][split][16][256] float4
  unsigned int *tickets;  // [tile], zero at rest
  // fused B broadcast (PULL instantiation, multi-GPU non-root ranks): the first n_copy CTAs to start
  // copy B from the root GPU (Bsrc, peer memory) into B, k-panel by k-panel, and exit; the others
  // multiply and wait for pull_ready[panel] to reach ready_target before loading rows of that panel.
  const float *Bsrc;
  unsigned int *pull_ready;   // [panel], grows by n_copy per call
  unsigned int *pull_ticket;  // arrival ticket, grows by the grid size per call
  unsigned int ticket_base, ready_target;
  int panels, n_copy;
  // chained broadcast (cuxb_mg_set_chain; the protocol is described in umma_gemm.cu): peer-mapped flags;
  // a rank waits only on its own: word p written by the rank it pulls from, word 4096 by the rank that
  // pulls from it, word 4097 (its count of completed calls) by itself
  unsigned int *chain_local;
  unsigned int *chain_up, *chain_down;
  unsigned int *done_ctr;
  unsigned int done_target;
  // NVSwitch multicast all-gather (cuxb_mg_set_multicast; the protocol is described in umma_gemm.cu):
  // mc_mode 2 = receiving rank, whose copy CTAs pull the rows of B this rank owns (every (world-1)-th row)
  // from the root and store them to the multicast address Bdst -- every rank's copy at once -- and whose
  // other CTAs wait, k-panel by k-panel, for the world-1 announcements in their own instance of the flags.
  int mc_mode;
  unsigned int mc_epoch;
  float *Bdst;
  unsigned int *mc_flags_local, *mc_flags_mc;
  int mc_world, mc_rank;
  unsigned int *mc_done_ctr;
  unsigned int mc_done_target;
};
constexpr int kSgMcMaxGroups = 512;   // same flag layout as umma_gemm.cu (kMcMaxGroups, kMcDoneBase)
constexpr int kSgMcDoneBase = 4100;   // same flag layout as umma_gemm.cu's kMcDoneBase

constexpr int kSgPullInFlight = 16;  // 16-byte loads in flight per thread: 64 KiB per copy CTA over NVLink
constexpr int kSgPullPanelK = 128;   // rows of B per readiness flag

__device__ __forceinline__ unsigned int sg_ld_acquire(const unsigned int *p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long sg_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned int sg_ld_acquire_sys(const unsigned int *p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void sg_wait_sys_reached(const unsigned int *p, unsigned int want) {
  const unsigned long long t0 = sg_timer_ns();
  while ((int)(sg_ld_acquire_sys(p) - want) < 0)
    if (sg_timer_ns() - t0 > 20000000000ull) __trap();
}

// One copy CTA's share of the broadcast: float4 e of a panel (row-major over rows x N/4) belongs to
// CTA (e / 256) mod n_copy; sixteen loads in flight per thread.
__device__ __noinline__ void sgemm_pull_b(const SgemmArgs &p, int c) {
  const int tid = threadIdx.x;
  const int n4 = p.N >> 2;
  const int stride = p.n_copy * SG_THREADS;
  const float *src = p.Bsrc;
  float *dst = p.mc_mode == 2 ? p.Bdst : const_cast<float *>(p.B);
  const int nr = p.mc_mode == 2 ? p.mc_world - 1 : 1, ro = p.mc_mode == 2 ? p.mc_rank - 1 : 0;   // rows we own: ro + nr * j
  unsigned int epoch = 0;
  if (p.mc_mode == 2) {
    // our stores land in every rank's copy: nothing of the previous call may be overwritten before every
    // receiver has read all of it
    if (tid == 0)
      for (int r = 1; r < p.mc_world; r++)
        if (r != p.mc_rank) sg_wait_sys_reached(p.mc_flags_local + kSgMcDoneBase + r, p.mc_epoch - 1);
    __syncthreads();
  }
  if (p.chain_local != nullptr) {
    // our call epoch; nothing of the previous call may be overwritten before the next rank has read it
    if (tid == 0) {
      epoch = sg_ld_acquire_sys(p.chain_local + 4097) + 1;     // our own count of completed calls
      if (p.chain_down != nullptr) sg_wait_sys_reached(p.chain_local + 4096, epoch - 1);   // written by the rank that pulls from us
    }
    __syncthreads();
  }
  for (int pnl = 0; pnl < p.panels; pnl++) {
    if (p.chain_up != nullptr) {
      // the upstream copy of this panel is complete: the upstream rank wrote that into OUR flags (a local poll)
      if (tid == 0) sg_wait_sys_reached(p.chain_local + pnl, epoch);
      __syncthreads();
    }
    const int r0 = pnl * kSgPullPanelK;
    const int panel_rows = min(kSgPullPanelK, p.K - r0);
    const int rows = (panel_rows - ro + nr - 1) / nr;   // the rows of this panel that are ours
    const long long total = (long long)rows * n4;  // < 2^31 (checked on the host)
    for (long long e0 = c * SG_THREADS + tid; e0 < total; e0 += (long long)kSgPullInFlight * stride) {
      float4 v[kSgPullInFlight];
#pragma unroll
      for (int u = 0; u < kSgPullInFlight; u++) {
        const long long e = e0 + (long long)u * stride;
        if (e < total) {
          const int j = (int)e / n4, col4 = (int)e - j * n4, row = ro + nr * j;
          const float *sp = src + (size_t)(r0 + row) * p.ldb + (size_t)col4 * 4;
          // A coherent load: in a chain the upstream GPU is still writing other panels of this buffer
          // while we read, so the non-coherent (.nc) path, which assumes memory that is read-only for
          // the whole kernel, is not allowed here.  The acquire on the panel's flag above orders it.
          asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(sp) : "memory");
        }
      }
#pragma unroll
      for (int u = 0; u < kSgPullInFlight; u++) {
        const long long e = e0 + (long long)u * stride;
        if (e < total) {
          const int j = (int)e / n4, col4 = (int)e - j * n4, row = ro + nr * j;
          float *dp = dst + (size_t)(r0 + row) * p.ldb + (size_t)col4 * 4;
          if (p.mc_mode == 2)   // a multicast address is written with multimem.st: the switch replicates the store
            asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dp), "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
          else
            *reinterpret_cast<float4 *>(dp) = v[u];
        }
      }
    }
    if (p.mc_mode == 2) __threadfence_system(); else __threadfence();
    __syncthreads();
    if (tid == 0) {
      unsigned int old;
      asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(p.pull_ready + pnl) : "memory");
      if (p.mc_mode == 2 && old + 1 == p.ready_target) {
        // last copier of the panel: one multicast store tells every rank that this rank's share has landed
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        asm volatile("multimem.st.release.sys.global.u32 [%0], %1;"
                     ::"l"(p.mc_flags_mc + (p.mc_rank - 1) * kSgMcMaxGroups + pnl), "r"(p.mc_epoch) : "memory");
      }
      if (p.chain_local != nullptr && old + 1 == p.ready_target) {
        // last copier of the panel: tell the rank that pulls from us (a store into ITS flags), and count the panel
        if (p.chain_down != nullptr)
          asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.chain_down + pnl), "r"(epoch) : "memory");
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(p.done_ctr) : "memory");
        if (old + 1 == p.done_target) {
          // every panel is here: count the call, and tell the rank we pulled from that we have read all of its copy
          asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.chain_local + 4097), "r"(epoch) : "memory");
          if (p.chain_up != nullptr)
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.chain_up + 4096), "r"(epoch) : "memory");
        }
      }
    }
  }
}

template <bool FAST, bool SPLITK, bool PULL = false>
__global__ void __launch_bounds__(SG_THREADS, 2) sgemm_kernel(const SgemmArgs p) {
  __shared__ __align__(16) float As[2][SG_BK][SG_LDA];
  __shared__ __align__(16) float Bs[2][SG_BK][SG_BN];

  const int tid = threadIdx.x;
  int block_id = blockIdx.x;
  if constexpr (PULL) {
    // Roles by order of arrival: the first n_copy CTAs to start copy B and exit, the others multiply
    // tile (arrival - n_copy).  A CTA that waits for a panel therefore only ever waits for CTAs that
    // are already running and that never wait themselves -- no deadlock however few CTAs fit at once.
    __shared__ unsigned int s_slot;
    if (tid == 0) s_slot = atomicAdd(p.pull_ticket, 1u) - p.ticket_base;
    __syncthreads();
    const int slot = (int)s_slot;
    if (slot < p.n_copy) {
      sgemm_pull_b(p, slot);
      return;
    }
    block_id = slot - p.n_copy;
  }
  const int lane = tid & 31, warp = tid >> 5;
  const int warp_m = warp >> 2, warp_n = warp & 3;   // 2 x 4 warps
  // 8 x 4 lanes.  An LDS.128 costs max(2, distinct 16-byte chunks per 4-lane group) wavefronts
  // (measured on B200, tools/lds_patterns.cu), so within every group of four consecutive lanes the
  // A index may take two values (lane bit 1) and the B index two values (lane bit 0): both
  // fragment loads then cost the minimum of 2 wavefronts.
  const int lane_m = (lane >> 1) & 7, lane_n = ((lane & 16) >> 3) | (lane & 1);
  // rows owned: tm0 + {0..3} and tm0 + 32 + {0..3};  cols: tn0 + {0..3} and tn0 + 16 + {0..3}
  const int tm0 = warp_m * 64 + lane_m * 4;
  const int tn0 = warp_n * 32 + lane_n * 4;

  // grouped rasterisation
  const int pid = SPLITK ? block_id / p.splits : block_id;
  const int split = SPLITK ? block_id - pid * p.splits : 0;
  const int k_begin = SPLITK ? split * p.k_per_split : 0;
  const int k_end = SPLITK ? min(p.K, k_begin + p.k_per_split) : p.K;
  const int per_group = SG_GROUP_M * p.tiles_n;
  const int group = pid / per_group;
  const int first_m = group * SG_GROUP_M;
  const int gsz = min(p.tiles_m - first_m, SG_GROUP_M);
  const int tile_m = first_m + (pid % per_group) % gsz;
  const int tile_n = (pid % per_group) / gsz;
  const int m_base = tile_m * SG_BM, n_base = tile_n * SG_BN;

  // global->shared assignment
  const int a_row = tid >> 2, a_k4 = (tid & 3) * 4;  // + 64*i rows
  const int b_k = tid >> 5, b_n4 = (tid & 31) * 4;   // + 8*i k-rows

  float4 ra[2], rb[2];
  // ---- global -> registers / shared.  FAST: B goes straight to shared memory with cp.async (no
  // staging registers), A through registers because it is transposed on the way.  The loads are
  // volatile asm with a memory clobber so they are ISSUED here, ahead of the FFMA block of the
  // current tile (left to itself the compiler sinks them next to the stores and exposes the full
  // global-memory latency once per k-tile).
  auto gload = [&](int k0, int nbuf) {
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const int m = m_base + a_row + 64 * i;
      const int k = k0 + a_k4;
      const int kb = k0 + b_k + 8 * i;
      const int n = n_base + b_n4;
      if (FAST) {
        const float *ap = p.A + (size_t)m * p.lda + k;
        asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(ra[i].x), "=f"(ra[i].y), "=f"(ra[i].z), "=f"(ra[i].w) : "l"(ap) : "memory");
        const float *bp = p.B + (size_t)kb * p.ldb + n;
        const unsigned sdst = (unsigned)__cvta_generic_to_shared(&Bs[nbuf][b_k + 8 * i][b_n4]);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst), "l"(bp) : "memory");
      } else {
        float t[4];
#pragma unroll
        for (int c = 0; c < 4; c++) t[c] = (m < p.M && k + c < k_end) ? p.A[(size_t)m * p.lda + k + c] : 0.f;
        ra[i] = make_float4(t[0], t[1], t[2], t[3]);
#pragma unroll
        for (int c = 0; c < 4; c++) t[c] = (kb < k_end && n + c < p.N) ? p.B[(size_t)kb * p.ldb + n + c] : 0.f;
        rb[i] = make_float4(t[0], t[1], t[2], t[3]);
      }
    }
    if (FAST) asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const int m = a_row + 64 * i;
      As[buf][a_k4 + 0][m] = ra[i].x;
      As[buf][a_k4 + 1][m] = ra[i].y;
      As[buf][a_k4 + 2][m] = ra[i].z;
      As[buf][a_k4 + 3][m] = ra[i].w;
      if (!FAST) *reinterpret_cast<float4 *>(&Bs[buf][b_k + 8 * i][b_n4]) = rb[i];
    }
    if (FAST) asm volatile("cp.async.wait_group 0;" ::: "memory");
  };

  // Accumulators are kept as row PAIRS so the inner product runs on packed FFMA2 (fma.rn.f32x2,
  // sm_100): acc2[ip][j] = (C[2ip][j], C[2ip+1][j]).  FFMA2 has the same FLOP rate as FFMA but
  // needs half the issue slots (measured, tools/fp32_peak.cu), which is what a 4-LDS-per-64-FMA
  // loop is short of.  Each lane is an ordinary IEEE fp32 fma, so results are unchanged.
  float2 acc2[4][8];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc2[i][j] = make_float2(0.f, 0.f);

  // PULL: rows [q*128, q*128+128) of B may be read once every copy CTA has delivered its share of panel q
  int panels_known = 0;
  auto need_panel = [&](int k0) {
    const int q = k0 / kSgPullPanelK;
    if (q < panels_known) return;
    if (p.mc_mode == 2) {   // every receiving rank's share of the panel has landed in our copy
      for (int o = 0; o < p.mc_world - 1; o++) sg_wait_sys_reached(p.mc_flags_local + o * kSgMcMaxGroups + q, p.mc_epoch);
      panels_known = q + 1;
      return;
    }
    const unsigned long long t0 = sg_timer_ns();
    while ((int)(sg_ld_acquire(p.pull_ready + q) - p.ready_target) < 0)
      if (sg_timer_ns() - t0 > 20000000000ull) __trap();
    panels_known = q + 1;
  };

  const int KT = (max(k_end - k_begin, 0) + SG_BK - 1) / SG_BK;
  if constexpr (PULL) need_panel(k_begin);
  gload(k_begin, 0);
  sstore(0);
  __syncthreads();

  float2 af[2][4];  // (a0,a1) (a2,a3) (a4,a5) (a6,a7)
  float bf[2][8];
  auto lfrag = [&](int buf, int k, int s) {
    const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][k][tm0]);
    const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][k][tm0 + 32]);
    const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tn0]);
    const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tn0 + 16]);
    af[s][0] = make_float2(a0.x, a0.y), af[s][1] = make_float2(a0.z, a0.w);
    af[s][2] = make_float2(a1.x, a1.y), af[s][3] = make_float2(a1.z, a1.w);
    bf[s][0] = b0.x, bf[s][1] = b0.y, bf[s][2] = b0.z, bf[s][3] = b0.w;
    bf[s][4] = b1.x, bf[s][5] = b1.y, bf[s][6] = b1.z, bf[s][7] = b1.w;
  };

  for (int kt = 0; kt < KT; kt++) {
    const int buf = kt & 1;
    const bool more = kt + 1 < KT;
    if constexpr (PULL) {
      if (more) need_panel(k_begin + (kt + 1) * SG_BK);
    }
    if (more) gload(k_begin + (kt + 1) * SG_BK, buf ^ 1);
    lfrag(buf, 0, 0);
#pragma unroll
    for (int k = 0; k < SG_BK; k++) {
      if (k + 1 < SG_BK) lfrag(buf, k + 1, (k + 1) & 1);
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const float2 bb = make_float2(bf[k & 1][j], bf[k & 1][j]);
#pragma unroll
        for (int ip = 0; ip < 4; ip++) acc2[ip][j] = __ffma2_rn(af[k & 1][ip], bb, acc2[ip][j]);
      }
    }
    if (more) sstore(buf ^ 1);
    __syncthreads();
  }

  float acc[8][8];
#pragma unroll
  for (int ip = 0; ip < 4; ip++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[2 * ip][j] = acc2[ip][j].x, acc[2 * ip + 1][j] = acc2[ip][j].y;

  if (SPLITK) {
    // park this split's accumulators (coalesced: float4 f of every thread is contiguous across the CTA)
    float4 *mine = p.scratch + ((size_t)pid * p.splits + split) * (16 * SG_THREADS);
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int h = 0; h < 2; h++)
        __stcg(mine + (i * 2 + h) * SG_THREADS + tid,
               make_float4(acc[i][h * 4 + 0], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]));
    __shared__ bool last_split;
    __threadfence();
    __syncthreads();
    if (tid == 0) last_split = (atomicAdd(p.tickets + pid, 1u) == (unsigned int)p.splits - 1);
    __syncthreads();
    if (!last_split) return;
    __threadfence();
    // the last CTA of the tile adds all splits in index order (its own included): a fixed order
    const float4 *all = p.scratch + (size_t)pid * p.splits * (16 * SG_THREADS);
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
    for (int sp = 0; sp < p.splits; sp++) {
      const float4 *src = all + (size_t)sp * (16 * SG_THREADS);
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const float4 v = __ldcg(src + (i * 2 + h) * SG_THREADS + tid);
          acc[i][h * 4 + 0] += v.x, acc[i][h * 4 + 1] += v.y, acc[i][h * 4 + 2] += v.z, acc[i][h * 4 + 3] += v.w;
        }
    }
    if (tid == 0) p.tickets[pid] = 0;  // at rest for the next launch
  }

  // epilogue
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int m = m_base + tm0 + (i & 3) + (i >> 2) * 32;
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int n = n_base + tn0 + h * 16;
      float *cp = p.C + (size_t)m * p.ldc + n;
      if (FAST) {
        float4 o;
        // two separately rounded products and one add -- never contracted into an fma -- exactly as the
        // tensor-core epilogue (umma_gemm.cu store8) and the CPU oracle's alpha*acc + beta*c do
        if (p.beta != 0.f) {
          const float4 c0 = *reinterpret_cast<const float4 *>(cp);
          o.x = __fadd_rn(__fmul_rn(p.alpha, acc[i][h * 4 + 0]), __fmul_rn(p.beta, c0.x));
          o.y = __fadd_rn(__fmul_rn(p.alpha, acc[i][h * 4 + 1]), __fmul_rn(p.beta, c0.y));
          o.z = __fadd_rn(__fmul_rn(p.alpha, acc[i][h * 4 + 2]), __fmul_rn(p.beta, c0.z));
          o.w = __fadd_rn(__fmul_rn(p.alpha, acc[i][h * 4 + 3]), __fmul_rn(p.beta, c0.w));
        } else {
          o.x = __fmul_rn(p.alpha, acc[i][h * 4 + 0]);
          o.y = __fmul_rn(p.alpha, acc[i][h * 4 + 1]);
          o.z = __fmul_rn(p.alpha, acc[i][h * 4 + 2]);
          o.w = __fmul_rn(p.alpha, acc[i][h * 4 + 3]);
        }
        *reinterpret_cast<float4 *>(cp) = o;
      } else if (m < p.M) {
#pragma unroll
        for (int c = 0; c < 4; c++)
          if (n + c < p.N) {
            float o = __fmul_rn(p.alpha, acc[i][h * 4 + c]);
            if (p.beta != 0.f) o = __fadd_rn(o, __fmul_rn(p.beta, cp[c]));
            cp[c] = o;
          }
      }
    }
  }
  if constexpr (PULL) {
    if (p.mc_mode == 2) {
      // this CTA has read all it needs of B; the last CTA of the launch tells the other receivers so
      __syncthreads();
      if (tid == 0) {
        unsigned int old;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(p.mc_done_ctr) : "memory");
        if (old + 1 == p.mc_done_target)
          asm volatile("multimem.st.release.sys.global.u32 [%0], %1;" ::"l"(p.mc_flags_mc + kSgMcDoneBase + p.mc_rank), "r"(p.mc_epoch) : "memory");
      }
    }
  }
}

// Split count for a problem of `tiles` output tiles and depth K on a GPU with room for `room` CTAs.
// Pure host arithmetic (tests/test_abi.py exercises it without a GPU through cuxb_debug_sgemm_splits).
// Pick the split count whose last wave is fullest (efficiency = work / (waves * room)); splitting costs
// one round trip of the accumulators through L2, so it has to buy at least 8 %.  At least 64 k per
// split, at most 32 splits and 1024 CTAs; then rounded so that every split is a whole number of k-tiles.
int sgemm_pick_splits(int tiles, int K, int room) {
  int splits = 1;
  if (tiles <= 1024 && K >= 256) {
    auto eff = [&](int sp) {
      const int ctas = tiles * sp;
      const int waves = (ctas + room - 1) / room;
      return (double)ctas / ((double)waves * room);
    };
    double best = eff(1) * 1.08;
    const int max_sp = K / 64 < 32 ? K / 64 : 32;
    for (int sp = 2; sp <= max_sp && tiles * sp <= 1024; sp++)
      if (eff(sp) > best) best = eff(sp), splits = sp;
  }
  if (splits > 1) {
    int kps = (K + splits - 1) / splits;
    kps = (kps + SG_BK - 1) / SG_BK * SG_BK;
    splits = (K + kps - 1) / kps;
  }
  return splits;
}

// Copy CTAs of the fused-broadcast launch: the slots the last wave of tiles leaves idle (16..63), else 64.
int sgemm_pick_copy_ctas(int tiles, int room) {
  const int spare = (tiles % room) ? room - tiles % room : 0;
  return (spare >= 16 && spare < 64) ? spare : 64;
}

const void *sgemm_kernel_symbol() { return (const void *)sgemm_kernel<true, false>; }

}  // namespace cuxb

using namespace cuxb;

CUXB_API int cuxb_sgemm_f32(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha,
                            const float *A, int lda, const float *B, int ldb, float beta, float *C,
                            int ldc) {
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(M >= 0 && N >= 0 && K >= 0, CUXB_ERR_INVALID, "cuxb_sgemm_f32: negative dimension");
  if (M == 0 || N == 0) return 0;
  CUXB_REQUIRE(A && B && C, CUXB_ERR_INVALID, "cuxb_sgemm_f32: null pointer");
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID,
               "cuxb_sgemm_f32: leading dimension smaller than the row length");
  SgemmArgs p;
  p.M = M, p.N = N, p.K = K, p.alpha = alpha, p.beta = beta;
  p.A = A, p.lda = lda, p.B = B, p.ldb = ldb, p.C = C, p.ldc = ldc;
  p.tiles_m = (M + SG_BM - 1) / SG_BM;
  p.tiles_n = (N + SG_BN - 1) / SG_BN;
  const bool fast = (M % SG_BM == 0) && (N % SG_BN == 0) && (K % SG_BK == 0) && K > 0 &&
                    (lda % 4 == 0) && (ldb % 4 == 0) && (ldc % 4 == 0) &&
                    ((((uintptr_t)A | (uintptr_t)B | (uintptr_t)C) & 15) == 0);
  const int tiles = p.tiles_m * p.tiles_n;
  cudaStream_t st = (cudaStream_t)stream;
  // split-K when the tiles alone would leave most of the GPU idle (room: 2 CTAs per SM) -- the
  // 64 x 4096 x 64 case of cuda/gemm_fp32.cu:162-258,714-718 and cux's demo shape
  p.splits = 1, p.k_per_split = K > 0 ? K : 1, p.scratch = nullptr, p.tickets = nullptr;
  p.Bsrc = nullptr, p.pull_ready = nullptr, p.pull_ticket = nullptr, p.ticket_base = 0, p.ready_target = 0, p.panels = 0, p.n_copy = 0;
  p.chain_local = nullptr, p.chain_up = nullptr, p.chain_down = nullptr, p.done_ctr = nullptr, p.done_target = 0;
  const int room = 2 * ctx->sm_count;  // CTAs resident at once
  const int splits = sgemm_pick_splits(tiles, K, room);
  if (splits > 1) {
    // layout: [0, 4096) per-tile tickets (tiles <= 1024; zero at rest), then the parked accumulators
    const size_t need = 4096 + (size_t)tiles * splits * 16 * SG_THREADS * sizeof(float4);
    if (ctx->sg_scratch_bytes < need) {
      if (ctx->sg_scratch) {
        // not freed: a CUDA graph recorded with the smaller scratch must stay replayable
        if ((rc = retire_workspace(ctx, ctx->sg_scratch))) return rc;
        ctx->sg_scratch = nullptr, ctx->sg_scratch_bytes = 0;
      }
      CUXB_CUDA(cudaMalloc(&ctx->sg_scratch, need));
      CUXB_CUDA(cudaMemset(ctx->sg_scratch, 0, 4096));
      ctx->sg_scratch_bytes = need;
    }
    p.splits = splits, p.k_per_split = (K + splits - 1) / splits;
    p.k_per_split = (p.k_per_split + SG_BK - 1) / SG_BK * SG_BK;
    p.tickets = (unsigned int *)ctx->sg_scratch;
    p.scratch = (float4 *)((char *)ctx->sg_scratch + 4096);
  }
  const unsigned int grid = (unsigned int)(tiles * p.splits);
  if (p.splits > 1) {
    if (fast) sgemm_kernel<true, true><<<grid, SG_THREADS, 0, st>>>(p);
    else sgemm_kernel<false, true><<<grid, SG_THREADS, 0, st>>>(p);
  } else {
    if (fast) sgemm_kernel<true, false><<<grid, SG_THREADS, 0, st>>>(p);
    else sgemm_kernel<false, false><<<grid, SG_THREADS, 0, st>>>(p);
  }
  CUXB_LAUNCHED();
  return 0;
}

// Row-sharded SGEMM with the broadcast of B fused into the kernel (mpi/alg_matrix_multiply.cpp:52-74:
// B to every worker, row blocks of A and C).  The root multiplies with b_root in place.  On any other
// rank the first n_copy CTAs to start copy B from the root's memory (b_root as mapped into this
// process) into b_local, 128 rows of B per readiness flag, and exit; every other CTA multiplies one
// tile, waiting only for the k-panel it is about to load.  n_copy = the slots the last wave of tiles
// would leave idle anyway, so that the copy does not push tiles into an extra wave (8192^3: 4096 tiles
// on 296 slots leave 48), at most 64; 64 when there is no such slack.  Sixteen 16-byte loads in flight
// per copy thread keep NVLink busy with that few CTAs.  Needs the vectorised path: M, N multiples of
// 128, K of 16, lda/ldb/ldc multiples of 4, 16-byte aligned pointers.
CUXB_API int cuxb_mg_sgemm_f32(cuxb_ctx *ctx, void *stream, int M, int N, int K, float alpha, const float *A,
                               int lda, const float *b_root, float *b_local, int ldb, float beta, float *C,
                               int ldc, int is_root) {
  if (is_root) return cuxb_sgemm_f32(ctx, stream, M, N, K, alpha, A, lda, b_root, ldb, beta, C, ldc);
  const int mc = (ctx != nullptr && ctx->mc_flags_local != nullptr) ? 2 : 0;   // multicast all-gather, receiving rank
  int rc = ensure_device(ctx);
  if (rc) return rc;
  CUXB_REQUIRE(M >= 0 && N >= 0 && K >= 0, CUXB_ERR_INVALID, "cuxb_mg_sgemm_f32: negative dimension");
  if (M == 0 || N == 0) return 0;
  CUXB_REQUIRE(A && b_root && b_local && C, CUXB_ERR_INVALID, "cuxb_mg_sgemm_f32: null pointer");
  CUXB_REQUIRE(lda >= K && ldb >= N && ldc >= N, CUXB_ERR_INVALID,
               "cuxb_mg_sgemm_f32: leading dimension smaller than the row length");
  const bool fast = (M % SG_BM == 0) && (N % SG_BN == 0) && (K % SG_BK == 0) && K > 0 && (lda % 4 == 0) &&
                    (ldb % 4 == 0) && (ldc % 4 == 0) &&
                    ((((uintptr_t)A | (uintptr_t)b_root | (uintptr_t)b_local | (uintptr_t)C) & 15) == 0);
  CUXB_REQUIRE(fast, CUXB_ERR_UNSUPPORTED,
               "cuxb_mg_sgemm_f32: needs M,N %% 128 == 0, K %% 16 == 0, lda/ldb/ldc %% 4 == 0 and 16-byte aligned pointers");
  SgemmArgs p;
  p.M = M, p.N = N, p.K = K, p.alpha = alpha, p.beta = beta;
  p.A = A, p.lda = lda, p.B = b_local, p.ldb = ldb, p.C = C, p.ldc = ldc;
  p.tiles_m = M / SG_BM, p.tiles_n = N / SG_BN;
  p.splits = 1, p.k_per_split = K, p.scratch = nullptr, p.tickets = nullptr;
  p.panels = (K + kSgPullPanelK - 1) / kSgPullPanelK;
  CUXB_REQUIRE(p.panels <= 4096 && (long long)kSgPullPanelK * (N / 4) < (1ll << 31), CUXB_ERR_TOO_LARGE,
               "cuxb_mg_sgemm_f32: K or N too large for the fused broadcast");
  cudaStream_t st = (cudaStream_t)stream;
  if (ctx->sg_pull == nullptr) {
    CUXB_CUDA(cudaMalloc((void **)&ctx->sg_pull, (4096 + 32) * sizeof(unsigned int)));
    CUXB_CUDA(cudaMemset(ctx->sg_pull, 0, (4096 + 32) * sizeof(unsigned int)));
    ctx->sg_pull_calls = 0, ctx->sg_pull_ticket_base = 0, ctx->sg_pull_panels = 0, ctx->sg_pull_ncopy = 0;
    ctx->sg_pull_done_total = 0;
  }
  const int tiles = p.tiles_m * p.tiles_n;
  const int room = 2 * ctx->sm_count;
  p.n_copy = sgemm_pick_copy_ctas(tiles, room);
  // the flags only grow, by n_copy per call for each of the call's panels: a different panel count or
  // n_copy would leave them uneven, so start over from zero when either changes (ordered on the stream)
  if (ctx->sg_pull_panels != p.panels || ctx->sg_pull_ncopy != p.n_copy) {
    CUXB_CUDA(cudaMemsetAsync(ctx->sg_pull, 0, 4096 * sizeof(unsigned int), st));
    ctx->sg_pull_calls = 0;
    ctx->sg_pull_panels = p.panels;
    ctx->sg_pull_ncopy = p.n_copy;
  }
  const unsigned int grid = (unsigned int)(tiles + p.n_copy);
  p.Bsrc = b_root;
  p.pull_ready = ctx->sg_pull;
  p.pull_ticket = ctx->sg_pull + 4096;
  p.ticket_base = ctx->sg_pull_ticket_base;
  ctx->sg_pull_ticket_base += grid;                      // unsigned wrap-around is fine
  ctx->sg_pull_calls++;
  p.ready_target = ctx->sg_pull_calls * (unsigned int)p.n_copy;
  p.chain_local = ctx->chain_local, p.chain_up = const_cast<unsigned int *>(ctx->chain_up), p.chain_down = const_cast<unsigned int *>(ctx->chain_down);
  p.done_ctr = ctx->sg_pull + 4097, p.done_target = 0;
  if (ctx->chain_local != nullptr) {
    ctx->sg_pull_done_total += (unsigned int)p.panels;
    p.done_target = ctx->sg_pull_done_total;
  }
  p.mc_mode = mc, p.mc_epoch = 0, p.Bdst = nullptr, p.mc_flags_local = nullptr, p.mc_flags_mc = nullptr;
  p.mc_world = 0, p.mc_rank = 0, p.mc_done_ctr = nullptr, p.mc_done_target = 0;
  if (mc != 0) {
    CUXB_REQUIRE(ctx->mc_b != nullptr && p.panels <= kSgMcMaxGroups, CUXB_ERR_INVALID,
                 "cuxb_mg_sgemm_f32: multicast mode needs the multicast address of the copies of B, and K <= 65536");
    p.chain_local = nullptr, p.chain_up = nullptr, p.chain_down = nullptr;   // the chain is ignored in multicast mode
    p.mc_epoch = ++ctx->mc_tab[ctx->mc_cur].calls;
    p.mc_flags_local = ctx->mc_flags_local, p.mc_flags_mc = ctx->mc_flags_mc;
    p.mc_world = ctx->mc_world, p.mc_rank = ctx->mc_rank;
    p.Bdst = (float *)ctx->mc_b;            // the multicast address of every rank's copy
    p.mc_done_ctr = ctx->sg_pull + 4100;    // local: CTAs finished (never reset)
    ctx->sg_mc_done_total += grid - (unsigned int)p.n_copy;   // the multiplying CTAs report
    p.mc_done_target = ctx->sg_mc_done_total;
  }
  sgemm_kernel<true, false, true><<<grid, SG_THREADS, 0, st>>>(p);
  CUXB_LAUNCHED();
  return 0;
}

// test hooks (not in cuxb.h): the host-side planning arithmetic, callable without a GPU
CUXB_API int cuxb_debug_sgemm_splits(int M, int N, int K, int sm_count) {
  const int tiles = ((M + SG_BM - 1) / SG_BM) * ((N + SG_BN - 1) / SG_BN);
  return sgemm_pick_splits(tiles, K, 2 * sm_count);
}
CUXB_API int cuxb_debug_sgemm_copy_ctas(int M, int N, int sm_count) {
  return sgemm_pick_copy_ctas((M / SG_BM) * (N / SG_BN), 2 * sm_count);
}
