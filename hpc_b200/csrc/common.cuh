// common.cuh -- shared internals of libcuxb.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <mutex>

#include "../../include/cuxb.h"

#define CUXB_API extern "C" __attribute__((visibility("default")))

namespace cuxb {

void set_error(const char *fmt, ...);
extern std::atomic<long long> g_launches;

// Evaluate a CUDA runtime call; on failure record the message and return the code.
#define CUXB_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t e__ = (expr);                                                             \
    if (e__ != cudaSuccess) {                                                             \
      ::cuxb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                        \
      return (int)e__;                                                                    \
    }                                                                                     \
  } while (0)

#define CUXB_REQUIRE(cond, code, ...)  \
  do {                                 \
    if (!(cond)) {                     \
      ::cuxb::set_error(__VA_ARGS__);  \
      return (code);                   \
    }                                  \
  } while (0)

// Called after every <<< >>> / cudaLaunchKernelEx.
#define CUXB_LAUNCHED()                       \
  do {                                        \
    ::cuxb::g_launches.fetch_add(1);          \
    CUXB_CUDA(cudaGetLastError());            \
  } while (0)

static inline long long div_up(long long a, long long b) { return (a + b - 1) / b; }

// Scan tile-state capacity: one 64-bit word per tile.
constexpr int kScanTileElems = 8192;                 // 512 threads x 16 ints
constexpr long long kScanMaxTiles = (1ll << 31) / kScanTileElems;  // n < 2^31 (cux uses int lengths)
constexpr int kReduceMaxBlocks = 2048;
constexpr int kHostMaxChunks = 16;  // *_host entry points: at most this many row / element chunks in flight

// A cuTensorMapEncodeTiled result, keyed by everything that went into it.
struct TmaCacheEntry {
  const void *base;
  uint64_t rows, cols, ld;
  uint32_t box_cols, box_rows;
  int dtype;
  alignas(64) unsigned char map[128];
};

}  // namespace cuxb

struct cuxb_ctx {
  int device;
  int sm_count;
  int cc_major, cc_minor;
  // reductions: per-block partials (float / int32 share the storage) + arrival ticket
  void *red_partials;        // kReduceMaxBlocks * 8 bytes
  unsigned int *red_ticket;  // zero at rest
  // scan: tile states (64-bit), dynamic tile counter, launch epoch
  unsigned long long *scan_state;
  long long scan_state_tiles;
  unsigned int scan_epoch;
  unsigned int mg_epoch;       // unused (the exchange epoch lives in the exchange buffer, on the device)
  int *mg_seed;                // device int32: the sharded scan's exclusive offset
  unsigned int *bcast_ready;   // fused B broadcast: per-panel arrival counters (grow by grid size per call)
  unsigned int bcast_calls;
  int bcast_grid, bcast_panels;
  unsigned int *wave_sync;         // tcgen05 GEMM: per-launch arrival counter of the soft wave rendezvous
  unsigned int bcast_done_total;   // chained broadcast: panels published so far (all calls)
  unsigned int *chain_local;       // chained broadcast flags (cuxb_mg_set_chain), peer-mapped
  const unsigned int *chain_up, *chain_down;
  // NVSwitch multicast broadcast (cuxb_mg_set_multicast): peer- / multicast-mapped flag buffers
  int mc_world, mc_rank;
  unsigned int *mc_flags_local, *mc_flags_mc;
  void *mc_b;                      // multicast address of the ranks' copies of B
  // call counters (the call epoch; same on every rank), one per flags buffer this context has been given:
  // a caller may keep several broadcasts (several B's, each with its own flags) alive and switch between them
  struct { const void *flags; unsigned int calls; } mc_tab[8];
  int mc_cur;                      // index into mc_tab of the active flags
  unsigned int mc_done_total;      // receiver: tensor-GEMM CTAs launched so far in multicast mode (all calls)
  unsigned int sg_mc_done_total;   // the same for the SGEMM (its device-side counter is separate)
  // sgemm split-K: parked accumulators + per-tile tickets
  void *sg_scratch;
  size_t sg_scratch_bytes;
  // sgemm fused B broadcast: [0,4096) per-panel arrival counters, [4096] role ticket
  unsigned int *sg_pull;
  unsigned int sg_pull_calls, sg_pull_ticket_base;
  int sg_pull_panels, sg_pull_ncopy;
  unsigned int sg_pull_done_total;
  // Workspaces that a larger call has superseded.  They stay allocated until cuxb_ctx_destroy: a CUDA graph
  // recorded earlier holds their address and must stay replayable (bounded: sizes double, so a handful).
  void *retired[48];
  int retired_n;
  // host-API staging
  void *stage[4];
  size_t stage_bytes[4];
  cudaStream_t stage_stream;   // the *_host entry points launch their kernels here
  // *_host pipelines: uploads and downloads run on their own streams so that PCIe carries both
  // directions while the kernels of earlier chunks run; events order chunk i's upload -> kernel -> download
  cudaStream_t up_stream, down_stream;
  cudaEvent_t ev_up[cuxb::kHostMaxChunks + 1], ev_k[cuxb::kHostMaxChunks];
  bool host_pipe_ready;
  // TMA descriptor cache
  cuxb::TmaCacheEntry tma_cache[16];  // most recent TMA descriptors (encoding one costs microseconds)
  int tma_cache_n;
  int tma_cache_next;
  // tuning / test hooks (cuxb_debug_ctx_set): per context, zero = library default
  int dbg_cta_group, dbg_group_m, dbg_no_wave_sync, dbg_mc_group, dbg_copy_bufs, dbg_wide, dbg_epilogue;
  std::mutex *mu;      // serialises the *_host entry points (they share the staging buffers)
  std::mutex *tma_mu;  // guards the descriptor cache (taken inside kernels' entry points, never with mu's order reversed)
};

namespace cuxb {
int ensure_device(cuxb_ctx *ctx);
int stage_reserve(cuxb_ctx *ctx, int slot, size_t bytes, void **out);
// keep a superseded workspace alive until the context is destroyed (frees at once only if the list is full)
int retire_workspace(cuxb_ctx *ctx, void *p);
}  // namespace cuxb
