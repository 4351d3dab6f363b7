// array4d.hpp -- Array4D, the data container cux operators exchange
// (0-frameworks/cux/include/array.h:73-156, src/array.cpp:9-87): one lazily allocated host buffer and
// one device buffer per element type, PUSH / PUSH_IF_EMPTY / NO_PUSH copy semantics
// (util.h:42-46), Save/Restore snapshots so every kernel starts from the same C.
// Device memory comes from cuxb_malloc and moves with cuxb_memcpy_*; no CUDA header is involved.
#pragma once
#include <vector>

#include "cux/core.hpp"
#include "cux/data_filler.hpp"

namespace cux {

class Array4D {
public:
  explicit Array4D(int num, int channels, int height, int width);
  ~Array4D();
  Array4D(const Array4D &) = delete;
  Array4D &operator=(const Array4D &) = delete;

  inline const std::vector<int> shape() const { return shape_; }
  inline int num_element() { return num_element_; }
  inline void *cpu_data(int type_flag) { return cpu_[type_flag]; }
  inline void *gpu_data(int type_flag) { return gpu_[type_flag]; }

  template <typename DType> DType *GetCpuData(DataFetchMode mode = NO_PUSH) {
    return static_cast<DType *>(FetchCpu(DataType<DType>::kFlag, mode));
  }
  template <typename DType> DType *GetGpuData(DataFetchMode mode = NO_PUSH) {
    return static_cast<DType *>(FetchGpu(DataType<DType>::kFlag, mode));
  }

  // decimal_pose: 0 for integers, 1 for x.x, 2 for x.xx ...  ON_HOST is the reference's path
  // (array.cpp:46-60).  ON_DEVICE -- rejected by the reference -- fills the device copy with
  // cuxb_fill_random (needs SetDeviceContext); the host copy is pulled lazily by GetCpuData(PUSH...).
  void Fill(int min_value, int max_value, int decimal_pose, int type_flag, OpRunMode mode = OpRunMode::ON_HOST);
  // Context used by the device-side extensions (Fill ON_DEVICE).  Borrowed, not owned.
  static void SetDeviceContext(cuxb_ctx *ctx) { DeviceContext() = ctx; }
  // Allocate host copies as pinned memory (cuxb_malloc_host): H2D / D2H then run as DMA at PCIe speed.
  // Off by default so that host-only use needs no GPU.
  static void UsePinnedHostMemory(bool on) { Pinned() = on; }
  void Save(int type_flag, OpRunMode mode, bool is_save_if_empty = true);
  void Restore(int type_flag, OpRunMode mode);

  // Host-side element conversion (array.h:124-144).
  template <typename SrcType, typename DstType> void PrecsCpuCvt() {
    const int s = DataType<SrcType>::kFlag, d = DataType<DstType>::kFlag;
    if (s == d) CUXLOG_ERR("PrecsCpuCvt -> src_data_type == dst_data_type.");
    if (cpu_[s] == nullptr) CUXLOG_ERR("PrecsCpuCvt -> cpu_data_[src_data_type] == nullptr.");
    const SrcType *src = static_cast<const SrcType *>(cpu_[s]);
    DstType *dst = static_cast<DstType *>(FetchCpu(d, NO_PUSH));
    for (int i = 0; i < num_element_; i++) dst[i] = (DstType)src[i];
  }
  // Extension: the same conversion on the device (cuxb_convert), for config-sized arrays where the
  // scalar host loop takes seconds.  Requires the source type to be present on the device.
  void PrecsGpuCvt(cuxb_ctx *ctx, int src_type_flag, int dst_type_flag);

private:
  static cuxb_ctx *&DeviceContext() { static cuxb_ctx *c = nullptr; return c; }
  static bool &Pinned() { static bool p = false; return p; }
  void *HostAlloc(size_t bytes);
  void HostFree(void *p);
  void *FetchCpu(int type_flag, DataFetchMode mode);
  void *FetchGpu(int type_flag, DataFetchMode mode);

  std::vector<int> shape_;
  int num_element_;
  std::vector<void *> cpu_, gpu_;          // indexed by TypeFlag
  std::vector<bool> cpu_pinned_;
  std::vector<void *> cpu_backup_, gpu_backup_;
};

}  // namespace cux
