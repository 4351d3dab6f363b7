// Operator base, OpAssistor, LaunchConfig and kernel tags
// (0-frameworks/cux/src/operator/operator.cpp:11-52, include/operator/op_assistor.h:100-143,
// include/util/launch_config.h:50-61).
#include "cux/operator.hpp"

namespace cux {

// ---- kernel tags: stand-ins for `void *config_kernel`
static const char *const kTags[] = {"dot_f32", "sum_f32", "nrm2_f32", "scan_i32", "sgemm_f32", "hgemm_f16"};
const void *KernelTag(const char *name) {
  for (const char *t : kTags)
    if (std::strcmp(t, name) == 0) return t;
  return nullptr;
}
const char *KernelTagName(const void *p) {
  for (const char *t : kTags)
    if (p == (const void *)t) return t;
  return nullptr;
}

void LaunchConfig::QueryPotentialOccupancy(const void *kernel, int threads_per_block, size_t dynamic_shared_mem,
                                           int &active_blocks, double &occupancy) {
  (void)threads_per_block;
  (void)dynamic_shared_mem;
  const char *name = KernelTagName(kernel);
  if (name == nullptr || ctx_ == nullptr) {
    // a foreign device-function pointer cannot be queried through the C ABI
    active_blocks = 0;
    occupancy = 0.0;
    return;
  }
  CUXB_CHECK(cuxb_kernel_occupancy(ctx_, name, &active_blocks, &occupancy, nullptr, nullptr));
}

OpAssistor::OpAssistor(Device *device) : device_(device), ctx_(nullptr), launch_config_(nullptr), checker_(nullptr) {
  // cux creates its cuBLAS handle here and aborts if it cannot (op_assistor.h:104-106)
  if (cuxb_ctx_create(device->id, &ctx_) != 0)
    CUXLOG_ERR("Cannot create the cuxb context: %s", cuxb_last_error_string());
  launch_config_ = new LaunchConfig(device, ctx_);
  checker_ = new ResultChecker();
}
OpAssistor::~OpAssistor() {
  delete checker_;
  delete launch_config_;
  if (ctx_ != nullptr) cuxb_ctx_destroy(ctx_);
}

void Operator::QueryPotentialOccupancy(const void *config_kernel, int kernel_id, int threads_per_block,
                                       int shared_memory_size) {
  if (kernel_id < 0 || kernel_id >= (int)gpu_kernel_occupancys_.size()) return;
  if (config_kernel == nullptr) {  // operator.cpp:13-18: nothing to query -> zeros
    gpu_kernel_active_blocks_[kernel_id] = 0;
    gpu_kernel_occupancys_[kernel_id] = 0.0;
    return;
  }
  assistor_->launch_config()->QueryPotentialOccupancy(config_kernel, threads_per_block, shared_memory_size,
                                                      gpu_kernel_active_blocks_[kernel_id],
                                                      gpu_kernel_occupancys_[kernel_id]);
}

void Operator::PrintRecordedInfo(const OpRunMode &mode, int kernel_id, const KernelInterface *info) {
  if (mode == OpRunMode::ON_HOST) {
    const KernelTimeRecord &r = cpu_timer_record_[kernel_id];
    CUXLOG_COUT("V%d ) I: %f, R: %f -> [ %s ]", kernel_id, r.input, r.run, info->describe_info.c_str());
  } else {
    const KernelTimeRecord &r = gpu_timer_record_[kernel_id];
    CUXLOG_COUT("V%d ) I: %f, W: %f, R: %f, O: %f -> %f [ %s ]", kernel_id, r.input, r.warnup, r.run, r.output,
                gpu_kernel_occupancys_[kernel_id], info->describe_info.c_str());
  }
}

void Operator::ResetByKernelNum(int cpu_kernel_num, int gpu_kernel_num) {
  cpu_timer_record_.resize(cpu_kernel_num);
  gpu_timer_record_.resize(gpu_kernel_num);
  gpu_kernel_occupancys_.resize(gpu_kernel_num);
  gpu_kernel_active_blocks_.resize(gpu_kernel_num);
}

void Operator::MarkTypes(std::vector<int> &type_flags, const std::vector<int> &used) const {
  type_flags.assign(TYPES_NUM, 0);
  for (int t : used) type_flags[t] = 1;
}

}  // namespace cux
