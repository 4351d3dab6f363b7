// Executor and environment functions (0-frameworks/cux/include/executor.h:18-72,
// src/executor.cpp:10-89), over the C ABI.
#include "cux/executor.hpp"

namespace cux {

void InitEnvironment() {
  CUXLOG_INFO("Initialize Environment.");
  static bool registered = false;
  if (!registered) {
    OpFactory &f = OpFactory::GetInstance();
    f.RegisterOpClass("dot", Dot::Creator);
    f.RegisterOpClass("nrm2", Nrm2::Creator);
    f.RegisterOpClass("gemm", Gemm::Creator);
    f.RegisterOpClass("reduce_sum", ReduceSum::Creator);
    f.RegisterOpClass("linear_int8", LinearInt8::Creator);
    f.RegisterOpClass("linear_bf16", LinearBf16::Creator);
    f.RegisterOpClass("scan_i32", ScanI32::Creator);
    registered = true;
  }
  CUXLOG_COUT("* Registered Op: %s.", OpFactory::GetInstance().PrintList().c_str());
  std::fprintf(stdout, "> \n");
}

void CleanUpEnvironment() {
  CUXLOG_INFO("Cleanup Environment.");
  CUXB_CHECK(cuxb_device_sync());
}

static void FillDeviceProp(int id, DeviceProp &p) {
  CUXB_CHECK(cuxb_device_name(id, p.name, sizeof(p.name)));
  CUXB_CHECK(cuxb_device_info(id, &p.multiProcessorCount, &p.major, &p.minor, &p.maxThreadsPerMultiProcessor,
                              &p.maxThreadsPerBlock, &p.warpSize, &p.totalGlobalMem, &p.sharedMemPerBlock));
}

void QueryDevices() {
  int count = 0;
  CUXB_CHECK(cuxb_device_count(&count));
  CUXLOG_COUT("-------------< Device Information >--------------");
  CUXLOG_COUT("-- The number of compute-capable devices: %d.", count);
  for (int id = 0; id < count; id++) {
    DeviceProp p;
    FillDeviceProp(id, p);
    CUXLOG_COUT("* GPU Device %d: \"%s\". ", id, p.name);
    CUXLOG_COUT("* (device) Compute capability: %d.%d. ", p.major, p.minor);
    CUXLOG_COUT("* (device) Number of multiprocessors on device: %d. ", p.multiProcessorCount);
    CUXLOG_COUT("* (device) Maximum resident threads per multiprocessor: %d. ", p.maxThreadsPerMultiProcessor);
    CUXLOG_COUT("* (device) Global memory available on device in bytes: %zu. ", p.totalGlobalMem);
    CUXLOG_COUT("* (block) Shared memory available per block in bytes: %zu. ", p.sharedMemPerBlock);
    CUXLOG_COUT("* (block) Maximum number of threads per block: %d. ", p.maxThreadsPerBlock);
    CUXLOG_COUT("* (thread) Warp size in threads: %d. ", p.warpSize);
  }
  CUXLOG_COUT("-------------------------------------------------");
}

void Executor::Initialize(const int dev_id) {
  device_.id = dev_id;
  CUXB_CHECK(cuxb_set_device(dev_id));
  FillDeviceProp(dev_id, device_.prop);
  op_assistor_ = new OpAssistor(&device_);
}

void Executor::Clear() {
  delete op_;
  op_ = nullptr;
  delete op_assistor_;
  op_assistor_ = nullptr;
}

void Executor::SelectOp(std::string op_name, std::string params) {
  delete op_;
  op_ = OpFactory::GetInstance().CreateOpByType(op_name, op_assistor_, params);
}

void Executor::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (op_ == nullptr) CUXLOG_ERR("The operator has not been selected yet. Please select an operator first.");
  op_->AddPlugin(kernel_if, mode);
}

void Executor::BindAndFill(const std::vector<Array4D *> &inputs, const std::vector<Array4D *> &outputs, int min_value,
                           int max_value, int decimal_pose) {
  if (op_ == nullptr) CUXLOG_ERR("The operator has not been selected yet. Please select an operator first.");
  inputs_ = inputs;
  outputs_ = outputs;
  // fill as fp32 on the host, then mirror into every other element type some kernel of the op uses
  // (executor.cpp:56-78)
  for (Array4D *a : inputs_) a->Fill(min_value, max_value, decimal_pose, TypeFlag::FLOAT32, OpRunMode::ON_HOST);
  for (Array4D *a : outputs_) a->Fill(min_value, max_value, decimal_pose, TypeFlag::FLOAT32, OpRunMode::ON_HOST);
  std::vector<int> used;
  op_->ExtractDataTypes(used);
  for (int t = 0; t < (int)used.size(); t++) {
    if (t == TypeFlag::FLOAT32 || !used[t]) continue;
    TYPE_SWITCH(t, Dst, {
      for (Array4D *a : inputs_) a->PrecsCpuCvt<float, Dst>();
      for (Array4D *a : outputs_) a->PrecsCpuCvt<float, Dst>();
    });
  }
}

void Executor::BindAndFillOnDevice(const std::vector<Array4D *> &inputs, const std::vector<Array4D *> &outputs,
                                   int min_value, int max_value, int decimal_pose) {
  if (op_ == nullptr) CUXLOG_ERR("The operator has not been selected yet. Please select an operator first.");
  inputs_ = inputs;
  outputs_ = outputs;
  Array4D::SetDeviceContext(op_assistor_->ctx());
  std::vector<int> used;
  op_->ExtractDataTypes(used);
  auto prepare = [&](Array4D *a) {
    a->Fill(min_value, max_value, decimal_pose, TypeFlag::FLOAT32, OpRunMode::ON_DEVICE);
    for (int t = 0; t < (int)used.size(); t++)
      if (t != TypeFlag::FLOAT32 && used[t]) a->PrecsGpuCvt(op_assistor_->ctx(), TypeFlag::FLOAT32, t);
  };
  for (Array4D *a : inputs_) prepare(a);
  for (Array4D *a : outputs_) prepare(a);
  CUXB_CHECK(cuxb_device_sync());
}

void Executor::Run(const OpRunMode mode) {
  if (op_ == nullptr) CUXLOG_ERR("The operator has not been selected yet. Please select an operator first.");
  if (mode == OpRunMode::ON_HOST) op_->RunOnHost(inputs_, outputs_);
  else op_->RunOnDevice(inputs_, outputs_);
}

bool Executor::AllChecksPassed() const { return op_assistor_ != nullptr && op_assistor_->checker()->all_passed(); }

}  // namespace cux
