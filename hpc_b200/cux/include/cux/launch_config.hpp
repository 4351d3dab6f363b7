// launch_config.hpp -- Config1D/2D/3D and LaunchConfig (0-frameworks/cux/include/util/launch_config.h).
// The structs keep the reference's fields (launch_config.h:17-42) so plugin code that fills them
// compiles unchanged.  The occupancy query goes through cuxb_kernel_occupancy instead of
// cudaOccupancyMaxActiveBlocksPerMultiprocessor (launch_config.h:50-61): the `kernel` argument is a
// cuxb kernel tag obtained from cux::KernelTag(), not a device function pointer.
#pragma once
#include <algorithm>

#include "cux/core.hpp"

namespace cux {

inline int DivUp(int a, int b) { return (a + b - 1) / b; }

struct Config1D {
  int virtual_thread_count = -1;
  int threads_per_block = -1;
  int blocks_per_grid = -1;
  int shared_memory_size = 0;
};
struct Config2D {
  dim3 virtual_thread_count = dim3(0, 0, 0);
  dim3 threads_per_block = dim3(0, 0, 0);
  dim3 blocks_per_grid = dim3(0, 0, 0);
  int shared_memory_size = 0;
};
struct Config3D {
  dim3 virtual_thread_count = dim3(0, 0, 0);
  dim3 threads_per_block = dim3(0, 0, 0);
  dim3 blocks_per_grid = dim3(0, 0, 0);
  int shared_memory_size = 0;
};

// A tag that stands for one of the library's kernels where cux expects `void *config_kernel`.
// Returns nullptr for an unknown name.  Valid names: dot_f32, sum_f32, nrm2_f32, scan_i32,
// sgemm_f32, hgemm_f16.
const void *KernelTag(const char *name);
// nullptr if `p` is not a tag returned by KernelTag (e.g. a foreign __global__ function pointer).
const char *KernelTagName(const void *p);

class LaunchConfig {
public:
  LaunchConfig(Device *device, cuxb_ctx *ctx = nullptr) : device_(device), ctx_(ctx) {}
  void set_ctx(cuxb_ctx *ctx) { ctx_ = ctx; }

  // occupancy = active_warps / max_warps, as launch_config.h:50-61
  void QueryPotentialOccupancy(const void *kernel, int threads_per_block, size_t dynamic_shared_mem,
                               int &active_blocks, double &occupancy);

  // Heuristic configs for simple memory-bound kernels (launch_config.h:116-160).
  Config1D GetKernelConfig1D(int element_count) const {
    Config1D c;
    const int physical = std::min(device_->prop.multiProcessorCount * device_->prop.maxThreadsPerMultiProcessor, element_count);
    c.virtual_thread_count = element_count;
    c.threads_per_block = std::min(1024, device_->prop.maxThreadsPerBlock);
    c.blocks_per_grid = std::min(DivUp(physical, c.threads_per_block), device_->prop.multiProcessorCount);
    return c;
  }
  Config2D GetKernelConfig2D(int xdim, int ydim) const {
    Config2D c;
    const int kThreads = 256;
    const int cols = std::min(xdim, kThreads), rows = std::max(kThreads / cols, 1);
    const int physical = device_->prop.multiProcessorCount * device_->prop.maxThreadsPerMultiProcessor;
    const int max_blocks = std::max(physical / kThreads, 1);
    c.virtual_thread_count = dim3(xdim, ydim, 1);
    c.threads_per_block = dim3(cols, rows, 1);
    const int gx = std::min(DivUp(xdim, cols), max_blocks);
    const int gy = std::min(max_blocks / gx, std::max(ydim / rows, 1));
    c.blocks_per_grid = dim3(gx, gy, 1);
    return c;
  }

private:
  Device *device_;
  cuxb_ctx *ctx_;
};

}  // namespace cux
