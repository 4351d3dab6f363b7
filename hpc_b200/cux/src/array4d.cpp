// Array4D over the C ABI (see include/cux/array4d.hpp; semantics of
// 0-frameworks/cux/include/array.h:73-156 and src/array.cpp:9-87).
#include "cux/array4d.hpp"

namespace cux {

Array4D::Array4D(int num, int channels, int height, int width)
    : shape_{num, channels, height, width},
      num_element_(num * channels * height * width),
      cpu_(TYPES_NUM, nullptr), gpu_(TYPES_NUM, nullptr), cpu_pinned_(TYPES_NUM, false),
      cpu_backup_(TYPES_NUM, nullptr), gpu_backup_(TYPES_NUM, nullptr) {}

Array4D::~Array4D() {
  for (int t = 0; t < TYPES_NUM; t++) {
    if (cpu_pinned_[t]) cuxb_free_host(cpu_[t]);
    else std::free(cpu_[t]);
    std::free(cpu_backup_[t]);
    if (gpu_[t]) cuxb_free(gpu_[t]);
    if (gpu_backup_[t]) cuxb_free(gpu_backup_[t]);
  }
}

void *Array4D::FetchCpu(int t, DataFetchMode mode) {
  const size_t bytes = (size_t)num_element_ * TypeSize(t);
  bool created = false;
  if (cpu_[t] == nullptr) {
    if (Pinned()) {
      CUXB_CHECK(cuxb_malloc_host(&cpu_[t], bytes));
      cpu_pinned_[t] = true;
    } else {
      cpu_[t] = std::malloc(bytes ? bytes : 1);
    }
    if (cpu_[t] == nullptr) CUXLOG_ERR("Array4D: out of host memory (%zu bytes).", bytes);
    created = true;
  }
  // PUSH: always refresh from the device copy; PUSH_IF_EMPTY: only when just created (array.h:92-96)
  if ((mode == PUSH || (mode == PUSH_IF_EMPTY && created)) && gpu_[t] != nullptr)
    CUXB_CHECK(cuxb_memcpy_d2h(cpu_[t], gpu_[t], bytes, nullptr));
  return cpu_[t];
}

void *Array4D::FetchGpu(int t, DataFetchMode mode) {
  const size_t bytes = (size_t)num_element_ * TypeSize(t);
  bool created = false;
  if (gpu_[t] == nullptr) {
    CUXB_CHECK(cuxb_malloc(&gpu_[t], bytes));
    created = true;
  }
  if ((mode == PUSH || (mode == PUSH_IF_EMPTY && created)) && cpu_[t] != nullptr)
    CUXB_CHECK(cuxb_memcpy_h2d(gpu_[t], cpu_[t], bytes, nullptr));
  return gpu_[t];
}

void Array4D::Fill(int min_value, int max_value, int decimal_pose, int type_flag, OpRunMode mode) {
  if (mode == OpRunMode::ON_DEVICE) {
    if (DeviceContext() == nullptr) CUXLOG_ERR("Fill Array ON_DEVICE needs Array4D::SetDeviceContext first.");
    void *d = FetchGpu(type_flag, NO_PUSH);
    static unsigned long long call = 0;  // a different stream of values for every array, reproducible run to run
    const unsigned long long seed = DataFiller::Seed() + 0x51ED270B27B4F3CFull * (++call);
    if (min_value == max_value)  // constant fill: the random fill of a one-value range
      CUXB_CHECK(cuxb_fill_random(DeviceContext(), nullptr, num_element_, d, type_flag, min_value, max_value, 0, seed));
    else
      CUXB_CHECK(cuxb_fill_random(DeviceContext(), nullptr, num_element_, d, type_flag, min_value, max_value, decimal_pose, seed));
    if (cpu_[type_flag] != nullptr)  // keep an existing host copy coherent
      CUXB_CHECK(cuxb_memcpy_d2h(cpu_[type_flag], d, (size_t)num_element_ * TypeSize(type_flag), nullptr));
    return;
  }
  TYPE_SWITCH(type_flag, T, {
    T *p = GetCpuData<T>();
    if (min_value == max_value)
      DataFiller::ConstantFill((float)max_value, num_element_, p);
    else
      DataFiller::RandomFill(min_value, max_value, decimal_pose, num_element_, p);
  });
}

void Array4D::Save(int t, OpRunMode mode, bool is_save_if_empty) {
  const size_t bytes = (size_t)num_element_ * TypeSize(t);
  if (mode == ON_HOST) {
    if (is_save_if_empty && cpu_backup_[t] != nullptr) return;
    if (cpu_[t] == nullptr) CUXLOG_ERR("Save -> no host data of type %d.", t);
    if (cpu_backup_[t] == nullptr) cpu_backup_[t] = std::malloc(bytes ? bytes : 1);
    std::memcpy(cpu_backup_[t], cpu_[t], bytes);
  } else {
    if (is_save_if_empty && gpu_backup_[t] != nullptr) return;
    if (gpu_[t] == nullptr) CUXLOG_ERR("Save -> no device data of type %d.", t);
    if (gpu_backup_[t] == nullptr) CUXB_CHECK(cuxb_malloc(&gpu_backup_[t], bytes));
    CUXB_CHECK(cuxb_memcpy_d2d(gpu_backup_[t], gpu_[t], bytes, nullptr));
  }
}

void Array4D::Restore(int t, OpRunMode mode) {
  const size_t bytes = (size_t)num_element_ * TypeSize(t);
  if (mode == ON_HOST) {
    if (cpu_backup_[t] == nullptr) CUXLOG_ERR("Restore -> backup_[%d] does not exist.", t);
    std::memcpy(cpu_[t], cpu_backup_[t], bytes);
  } else {
    if (gpu_backup_[t] == nullptr) CUXLOG_ERR("Restore -> backup_[%d] does not exist.", t);
    CUXB_CHECK(cuxb_memcpy_d2d(gpu_[t], gpu_backup_[t], bytes, nullptr));
  }
}

void Array4D::PrecsGpuCvt(cuxb_ctx *ctx, int s, int d) {
  if (s == d) CUXLOG_ERR("PrecsGpuCvt -> src_data_type == dst_data_type.");
  if (gpu_[s] == nullptr) CUXLOG_ERR("PrecsGpuCvt -> no device data of type %d.", s);
  void *dst = FetchGpu(d, NO_PUSH);
  CUXB_CHECK(cuxb_convert(ctx, nullptr, num_element_, gpu_[s], s, dst, d));
}

}  // namespace cux
