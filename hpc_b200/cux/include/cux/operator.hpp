// operator.hpp -- Operator base class and OpFactory
// (0-frameworks/cux/include/operator/operator.h:18-67, op_factory.h:17-70).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "cux/array4d.hpp"
#include "cux/kernel_interface.hpp"
#include "cux/op_assistor.hpp"
#include "cux/timer.hpp"

namespace cux {

struct KernelTimeRecord {  // operator.h:18-24 ("warnup" is the reference's spelling)
  float run = 0.0;
  float warnup = 0.0;
  float input = 0.0;
  float output = 0.0;
};

class Operator {
public:
  explicit Operator(OpAssistor *assistor) : assistor_(assistor) {}
  virtual ~Operator() {}

  void QueryPotentialOccupancy(const void *config_kernel, int kernel_id, int threads_per_block, int shared_memory_size);
  void PrintRecordedInfo(const OpRunMode &mode, int kernel_id, const KernelInterface *kernel_info);
  void ResetByKernelNum(int cpu_kernel_num, int gpu_kernel_num);

  virtual void Help() const = 0;
  virtual void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) = 0;
  virtual void ExtractDataTypes(std::vector<int> &type_flags) = 0;
  virtual void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) = 0;
  virtual void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) = 0;

private:
  virtual void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) = 0;

public:
  OpAssistor *assistor_;
  // GpuTimer is created lazily: constructing an op must not need a device (host-only use).
  GpuTimer &gpu_timer() { if (!gpu_timer_) gpu_timer_ = new GpuTimer(); return *gpu_timer_; }
  CpuTimer cpu_timer_;
  std::vector<KernelTimeRecord> cpu_timer_record_;
  std::vector<KernelTimeRecord> gpu_timer_record_;
  std::vector<double> gpu_kernel_occupancys_;
  std::vector<int> gpu_kernel_active_blocks_;

protected:
  void MarkTypes(std::vector<int> &type_flags, const std::vector<int> &used) const;
  GpuTimer *gpu_timer_ = nullptr;
};

class OpFactory {
  typedef Operator *(*OpCreator)(OpAssistor *assistor, std::string &params_str);
public:
  Operator *CreateOpByType(std::string type, OpAssistor *assistor, std::string param_str) {
    auto it = creators_.find(type);
    if (it == creators_.end()) {
      CUXLOG_ERR("Can not find Op: %s.\n Registered Op: < %s>", type.c_str(), PrintList().c_str());
      return nullptr;
    }
    return it->second ? it->second(assistor, param_str) : nullptr;
  }
  void RegisterOpClass(std::string type, OpCreator creator) {
    if (creators_.count(type)) {
      CUXLOG_WARN("Op type: %s has already been registered.", type.c_str());
      return;
    }
    creators_[type] = creator;
  }
  std::string PrintList() {
    std::string out;
    for (auto &kv : creators_) out += kv.first + " ";
    return out;
  }
  static OpFactory &GetInstance() {
    static OpFactory factory;
    return factory;
  }
private:
  OpFactory() {}
  std::map<std::string, OpCreator> creators_;
};

}  // namespace cux
