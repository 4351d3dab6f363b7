// runtime.cu -- context, memory, stream/event entry points of the cuxb C ABI.
// These replace the CUDA-runtime calls cux makes from Array4D, GpuTimer and
// OpAssistor (0-frameworks/cux/include/array.h:84-114, util/timer.h:16-38,
// operator/op_assistor.h:100-143) so that host code needs no CUDA header.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace cuxb {

static thread_local char t_err[512] = "no error";
std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

int ensure_device(cuxb_ctx *ctx) {
  CUXB_REQUIRE(ctx != nullptr, CUXB_ERR_INVALID, "null cuxb_ctx");
  int cur = -1;
  CUXB_CUDA(cudaGetDevice(&cur));
  if (cur != ctx->device) CUXB_CUDA(cudaSetDevice(ctx->device));
  return 0;
}

int retire_workspace(cuxb_ctx *ctx, void *p) {
  if (p == nullptr) return 0;
  if (ctx->retired_n < (int)(sizeof(ctx->retired) / sizeof(ctx->retired[0]))) {
    ctx->retired[ctx->retired_n++] = p;
    return 0;
  }
  CUXB_CUDA(cudaDeviceSynchronize());   // list full (never in practice): fall back to freeing now
  CUXB_CUDA(cudaFree(p));
  return 0;
}

int stage_reserve(cuxb_ctx *ctx, int slot, size_t bytes, void **out) {
  if (ctx->stage_bytes[slot] < bytes) {
    if (ctx->stage[slot]) CUXB_CUDA(cudaFree(ctx->stage[slot]));
    ctx->stage[slot] = nullptr;
    ctx->stage_bytes[slot] = 0;
    CUXB_CUDA(cudaMalloc(&ctx->stage[slot], bytes));
    ctx->stage_bytes[slot] = bytes;
  }
  *out = ctx->stage[slot];
  return 0;
}

}  // namespace cuxb

using namespace cuxb;

CUXB_API const char *cuxb_last_error_string(void) { return t_err; }
CUXB_API const char *cuxb_version(void) { return "cuxb 0.1 (sm_100a)"; }
CUXB_API long long cuxb_launch_count(void) { return g_launches.load(); }

CUXB_API int cuxb_ctx_create(int device, cuxb_ctx **out) {
  CUXB_REQUIRE(out != nullptr, CUXB_ERR_INVALID, "cuxb_ctx_create: out is null");
  *out = nullptr;
  int count = 0;
  CUXB_CUDA(cudaGetDeviceCount(&count));
  CUXB_REQUIRE(device >= 0 && device < count, CUXB_ERR_INVALID,
               "cuxb_ctx_create: device %d out of range (%d devices)", device, count);
  CUXB_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUXB_CUDA(cudaGetDeviceProperties(&prop, device));
  // No fallback path exists: the kernels are sm_100a-only (tcgen05/TMEM/TMA).
  CUXB_REQUIRE(prop.major == 10, CUXB_ERR_UNSUPPORTED,
               "cuxb needs a Blackwell sm_100 device, found sm_%d%d (%s)", prop.major, prop.minor,
               prop.name);
  cuxb_ctx *c = new cuxb_ctx();
  memset(c, 0, sizeof(*c));
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->cc_major = prop.major;
  c->cc_minor = prop.minor;
  c->mu = new std::mutex();
  c->tma_mu = new std::mutex();
  cudaError_t e;
  e = cudaMalloc(&c->red_partials, kReduceMaxBlocks * 8);
  if (e == cudaSuccess) e = cudaMalloc((void **)&c->red_ticket, 256);
  if (e == cudaSuccess) e = cudaMemset(c->red_ticket, 0, 256);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->stage_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("cuxb_ctx_create: %s", cudaGetErrorString(e));
    cuxb_ctx_destroy(c);
    return (int)e;
  }
  c->scan_epoch = 0;
  c->mg_epoch = 0;
  c->mg_seed = (int *)((char *)c->red_ticket + 128);  // inside the zeroed 256-byte block
  *out = c;
  return 0;
}

CUXB_API int cuxb_ctx_destroy(cuxb_ctx *c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  if (c->red_partials) cudaFree(c->red_partials);
  if (c->red_ticket) cudaFree(c->red_ticket);
  for (int i = 0; i < c->retired_n; i++) cudaFree(c->retired[i]);
  if (c->scan_state) cudaFree(c->scan_state);
  if (c->bcast_ready) cudaFree(c->bcast_ready);
  if (c->wave_sync) cudaFree(c->wave_sync);
  if (c->sg_scratch) cudaFree(c->sg_scratch);
  if (c->sg_pull) cudaFree(c->sg_pull);
  for (int i = 0; i < 4; i++)
    if (c->stage[i]) cudaFree(c->stage[i]);
  if (c->stage_stream) cudaStreamDestroy(c->stage_stream);
  if (c->host_pipe_ready) {
    cudaStreamDestroy(c->up_stream);
    cudaStreamDestroy(c->down_stream);
    for (int i = 0; i <= kHostMaxChunks; i++) cudaEventDestroy(c->ev_up[i]);
    for (int i = 0; i < kHostMaxChunks; i++) cudaEventDestroy(c->ev_k[i]);
  }
  delete c->mu;
  delete c->tma_mu;
  delete c;
  return 0;
}

CUXB_API int cuxb_ctx_device(const cuxb_ctx *c) { return c ? c->device : -1; }
CUXB_API int cuxb_ctx_sm_count(const cuxb_ctx *c) { return c ? c->sm_count : 0; }

CUXB_API int cuxb_malloc(void **p, size_t bytes) {
  CUXB_REQUIRE(p != nullptr, CUXB_ERR_INVALID, "cuxb_malloc: null out pointer");
  CUXB_CUDA(cudaMalloc(p, bytes ? bytes : 1));
  return 0;
}
CUXB_API int cuxb_free(void *p) {
  if (p) CUXB_CUDA(cudaFree(p));
  return 0;
}
CUXB_API int cuxb_malloc_host(void **p, size_t bytes) {
  CUXB_REQUIRE(p != nullptr, CUXB_ERR_INVALID, "cuxb_malloc_host: null out pointer");
  CUXB_CUDA(cudaMallocHost(p, bytes ? bytes : 1));
  return 0;
}
CUXB_API int cuxb_free_host(void *p) {
  if (p) CUXB_CUDA(cudaFreeHost(p));
  return 0;
}
static int copy(void *dst, const void *src, size_t bytes, void *stream, cudaMemcpyKind kind) {
  if (bytes == 0) return 0;
  CUXB_REQUIRE(dst && src, CUXB_ERR_INVALID, "cuxb_memcpy: null pointer");
  // cux semantics are synchronous cudaMemcpy (array.h:95,110); with an explicit
  // stream the copy is asynchronous on that stream.
  if (stream)
    CUXB_CUDA(cudaMemcpyAsync(dst, src, bytes, kind, (cudaStream_t)stream));
  else
    CUXB_CUDA(cudaMemcpy(dst, src, bytes, kind));
  return 0;
}
CUXB_API int cuxb_memcpy_h2d(void *d, const void *s, size_t n, void *st) { return copy(d, s, n, st, cudaMemcpyHostToDevice); }
CUXB_API int cuxb_memcpy_d2h(void *d, const void *s, size_t n, void *st) { return copy(d, s, n, st, cudaMemcpyDeviceToHost); }
CUXB_API int cuxb_memcpy_d2d(void *d, const void *s, size_t n, void *st) { return copy(d, s, n, st, cudaMemcpyDeviceToDevice); }
CUXB_API int cuxb_memset(void *d, int v, size_t n, void *st) {
  if (n == 0) return 0;
  CUXB_REQUIRE(d, CUXB_ERR_INVALID, "cuxb_memset: null pointer");
  if (st)
    CUXB_CUDA(cudaMemsetAsync(d, v, n, (cudaStream_t)st));
  else
    CUXB_CUDA(cudaMemset(d, v, n));
  return 0;
}
CUXB_API int cuxb_set_device(int device) {
  CUXB_CUDA(cudaSetDevice(device));
  return 0;
}
CUXB_API int cuxb_device_count(int *count) {
  CUXB_REQUIRE(count, CUXB_ERR_INVALID, "null");
  CUXB_CUDA(cudaGetDeviceCount(count));
  return 0;
}
CUXB_API int cuxb_device_sync(void) {
  CUXB_CUDA(cudaDeviceSynchronize());
  return 0;
}
CUXB_API int cuxb_device_name(int device, char *buf, size_t buflen) {
  CUXB_REQUIRE(buf && buflen, CUXB_ERR_INVALID, "null");
  cudaDeviceProp prop;
  CUXB_CUDA(cudaGetDeviceProperties(&prop, device));
  snprintf(buf, buflen, "%s", prop.name);
  return 0;
}

CUXB_API int cuxb_device_info(int device, int *sm_count, int *cc_major, int *cc_minor, int *max_threads_per_sm,
                               int *max_threads_per_block, int *warp_size, size_t *global_mem_bytes,
                               size_t *shared_mem_per_block) {
  cudaDeviceProp prop;
  CUXB_CUDA(cudaGetDeviceProperties(&prop, device));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (max_threads_per_sm) *max_threads_per_sm = prop.maxThreadsPerMultiProcessor;
  if (max_threads_per_block) *max_threads_per_block = prop.maxThreadsPerBlock;
  if (warp_size) *warp_size = prop.warpSize;
  if (global_mem_bytes) *global_mem_bytes = prop.totalGlobalMem;
  if (shared_mem_per_block) *shared_mem_per_block = prop.sharedMemPerBlock;
  return 0;
}

CUXB_API int cuxb_enable_peer_access(int device, int peer_device) {
  int can = 0;
  CUXB_CUDA(cudaDeviceCanAccessPeer(&can, device, peer_device));
  CUXB_REQUIRE(can, CUXB_ERR_UNSUPPORTED, "device %d cannot access device %d directly", device, peer_device);
  int cur = -1;
  CUXB_CUDA(cudaGetDevice(&cur));
  CUXB_CUDA(cudaSetDevice(device));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
  cudaSetDevice(cur);
  CUXB_CUDA(e);
  return 0;
}

// One process per GPU: export / import of device buffers between processes (see cuxb.h).
CUXB_API int cuxb_ipc_export(void *dptr, void *handle_out) {
  static_assert(sizeof(cudaIpcMemHandle_t) == CUXB_IPC_HANDLE_BYTES, "handle size");
  CUXB_REQUIRE(dptr != nullptr && handle_out != nullptr, CUXB_ERR_INVALID, "cuxb_ipc_export: null pointer");
  cudaIpcMemHandle_t h;
  CUXB_CUDA(cudaIpcGetMemHandle(&h, dptr));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}
CUXB_API int cuxb_ipc_import(const void *handle, void **dptr_out) {
  CUXB_REQUIRE(handle != nullptr && dptr_out != nullptr, CUXB_ERR_INVALID, "cuxb_ipc_import: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  *dptr_out = nullptr;
  CUXB_CUDA(cudaIpcOpenMemHandle(dptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
CUXB_API int cuxb_ipc_close(void *dptr) {
  if (dptr) CUXB_CUDA(cudaIpcCloseMemHandle(dptr));
  return 0;
}

// Chained broadcast for the cuxb_mg_*gemm / cuxb_mg_linear_* calls of this context (see cuxb.h).
CUXB_API int cuxb_mg_set_chain(cuxb_ctx *ctx, void *flags_local, void *flags_up, void *flags_down) {
  CUXB_REQUIRE(ctx != nullptr, CUXB_ERR_INVALID, "cuxb_mg_set_chain: null context");
  CUXB_REQUIRE(flags_local != nullptr || (flags_up == nullptr && flags_down == nullptr), CUXB_ERR_INVALID,
               "cuxb_mg_set_chain: flags_up / flags_down given without flags_local");
  CUXB_REQUIRE((((uintptr_t)flags_local | (uintptr_t)flags_up | (uintptr_t)flags_down) & 3) == 0, CUXB_ERR_INVALID,
               "cuxb_mg_set_chain: flag buffers must be 4-byte aligned");
  ctx->chain_local = (unsigned int *)flags_local;
  ctx->chain_up = (const unsigned int *)flags_up;
  ctx->chain_down = (const unsigned int *)flags_down;
  return 0;
}

// NVSwitch multicast all-gather for the cuxb_mg_*gemm / cuxb_mg_linear_* calls of this context (see cuxb.h).
CUXB_API int cuxb_mg_set_multicast(cuxb_ctx *ctx, int world, int rank, void *flags_local, void *flags_mc, void *b_mc) {
  CUXB_REQUIRE(ctx != nullptr, CUXB_ERR_INVALID, "cuxb_mg_set_multicast: null context");
  if (flags_local == nullptr) {   // off
    ctx->mc_flags_local = ctx->mc_flags_mc = nullptr;
    ctx->mc_b = nullptr;
    ctx->mc_world = 0, ctx->mc_rank = 0;
    return 0;
  }
  CUXB_REQUIRE(world >= 3 && world <= 8 && rank >= 0 && rank < world, CUXB_ERR_INVALID,
               "cuxb_mg_set_multicast: world=%d rank=%d (3..8 ranks; with two, the one receiver simply pulls from the root)",
               world, rank);
  CUXB_REQUIRE(rank == 0 || (flags_mc != nullptr && b_mc != nullptr), CUXB_ERR_INVALID,
               "cuxb_mg_set_multicast: a receiving rank needs the multicast addresses of the flags and of the copies of B");
  CUXB_REQUIRE((((uintptr_t)flags_local | (uintptr_t)flags_mc) & 3) == 0 && ((uintptr_t)b_mc & 15) == 0, CUXB_ERR_INVALID,
               "cuxb_mg_set_multicast: flag buffers must be 4-byte aligned, the copies of B 16-byte aligned");
  // the call number belongs to the flags: it continues when the same flags are selected again, and new
  // (zeroed) flags start at 1.  Eight flag buffers are remembered per context; beyond that the oldest
  // entry is reused (its flags must then not be selected again).
  int slot = -1;
  for (int i = 0; i < 8; i++)
    if (ctx->mc_tab[i].flags == flags_local) slot = i;
  if (slot < 0) {
    for (int i = 0; i < 8 && slot < 0; i++)
      if (ctx->mc_tab[i].flags == nullptr) slot = i;
    if (slot < 0) slot = (ctx->mc_cur + 1) % 8;
    ctx->mc_tab[slot].flags = flags_local, ctx->mc_tab[slot].calls = 0;
  }
  ctx->mc_cur = slot;
  ctx->mc_world = world, ctx->mc_rank = rank;
  ctx->mc_flags_local = (unsigned int *)flags_local;
  ctx->mc_flags_mc = (unsigned int *)flags_mc;
  ctx->mc_b = b_mc;
  return 0;
}

CUXB_API int cuxb_stream_create(void **s) {
  CUXB_REQUIRE(s, CUXB_ERR_INVALID, "null");
  cudaStream_t st;
  CUXB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  *s = st;
  return 0;
}
CUXB_API int cuxb_stream_destroy(void *s) {
  if (s) CUXB_CUDA(cudaStreamDestroy((cudaStream_t)s));
  return 0;
}
CUXB_API int cuxb_stream_sync(void *s) {
  CUXB_CUDA(cudaStreamSynchronize((cudaStream_t)s));
  return 0;
}
CUXB_API int cuxb_event_create(void **e) {
  CUXB_REQUIRE(e, CUXB_ERR_INVALID, "null");
  cudaEvent_t ev;
  CUXB_CUDA(cudaEventCreate(&ev));
  *e = ev;
  return 0;
}
CUXB_API int cuxb_event_destroy(void *e) {
  if (e) CUXB_CUDA(cudaEventDestroy((cudaEvent_t)e));
  return 0;
}
CUXB_API int cuxb_event_record(void *e, void *s) {
  CUXB_CUDA(cudaEventRecord((cudaEvent_t)e, (cudaStream_t)s));
  return 0;
}
CUXB_API int cuxb_event_sync(void *e) {
  CUXB_CUDA(cudaEventSynchronize((cudaEvent_t)e));
  return 0;
}
CUXB_API int cuxb_event_elapsed_ms(void *a, void *b, float *ms) {
  CUXB_REQUIRE(ms, CUXB_ERR_INVALID, "null");
  CUXB_CUDA(cudaEventSynchronize((cudaEvent_t)b));
  CUXB_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)a, (cudaEvent_t)b));
  return 0;
}
