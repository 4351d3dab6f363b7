// executor.hpp -- Executor and the environment functions
// (0-frameworks/cux/include/executor.h:18-103, src/executor.cpp:10-89).
#pragma once
#include "cux/multi_executor.hpp"
#include "cux/ops.hpp"

namespace cux {

// Registers the operators with the OpFactory (executor.h:18-27) -- the three cux ops and the three
// added ones.  Safe to call more than once.
void InitEnvironment();
void CleanUpEnvironment();  // executor.h:29-38 (cudaDeviceReset there; here: device sync)
void QueryDevices();        // executor.h:40-72

class Executor {
public:
  Executor() : op_assistor_(nullptr), op_(nullptr) {}
  void Initialize(const int dev_id);
  void Clear();
  void SelectOp(std::string op_name, std::string params);
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode);
  void BindAndFill(const std::vector<Array4D *> &inputs, const std::vector<Array4D *> &outputs, int min_value,
                   int max_value, int decimal_pose);
  // Extension: the same, but filled and converted on the GPU (cuxb_fill_random + cuxb_convert) -- at the
  // BASELINE sizes (16384^2 matrices) the reference's host fill and scalar conversion loops take seconds
  // per array.  Host copies appear lazily when a host kernel asks for them (GetCpuData(PUSH_IF_EMPTY)).
  void BindAndFillOnDevice(const std::vector<Array4D *> &inputs, const std::vector<Array4D *> &outputs, int min_value,
                           int max_value, int decimal_pose);
  void Run(const OpRunMode mode);

  // Additions (the reference only prints its verdicts, op_assistor.h:59-69):
  bool AllChecksPassed() const;
  OpAssistor *assistor() { return op_assistor_; }
  Operator *op() { return op_; }

private:
  Device device_;
  OpAssistor *op_assistor_;
  Operator *op_;
  std::vector<Array4D *> inputs_;
  std::vector<Array4D *> outputs_;
};

}  // namespace cux
