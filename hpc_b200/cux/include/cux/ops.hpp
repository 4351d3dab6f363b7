// ops.hpp -- the operators: cux's dot / nrm2 / gemm (0-frameworks/cux/include/operator/
// dot_product.h, nrm2.h, gemm.h) and the three added through the same OpFactory:
// reduce_sum, linear_int8, scan_i32.  Each op owns its kernel lists (deleted in the destructor,
// gemm.h:34-41), registers one host kernel "Normal" -- the verification baseline cux's checker
// compares against (op_assistor.h:31-37) -- and the sm_100a device kernels from libcuxb.
#pragma once
#include "cux/operator.hpp"

namespace cux {

// Shared by all ops: kernel lists, plugin registration, type extraction.
template <typename CpuIF, typename GpuIF>
class KernelLists {
public:
  ~KernelLists() {
    for (auto *k : cpu_kernels_) delete k;
    for (auto *k : gpu_kernels_) delete k;
  }
  std::vector<CpuIF *> cpu_kernels_;
  std::vector<GpuIF *> gpu_kernels_;
};

class Dot : public Operator, protected KernelLists<DotCpuKernelIF, DotGpuKernelIF> {
public:
  explicit Dot(OpAssistor *assistor);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *in_a_ = nullptr, *in_b_ = nullptr, *out_ = nullptr;
};

class Nrm2 : public Operator, protected KernelLists<Nrm2CpuKernelIF, Nrm2GpuKernelIF> {
public:
  explicit Nrm2(OpAssistor *assistor);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *in_ = nullptr, *out_ = nullptr;
};

class ReduceSum : public Operator, protected KernelLists<ReduceCpuKernelIF, ReduceGpuKernelIF> {
public:
  explicit ReduceSum(OpAssistor *assistor);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *in_ = nullptr, *out_ = nullptr;
};

struct GemmKernelParam {  // gemm.h:15-24
  float alpha = 1.0;
  float beta = 0.0;
};

class Gemm : public Operator, protected KernelLists<GemmCpuKernelIF, GemmGpuKernelIF> {
public:
  Gemm(OpAssistor *assistor, GemmKernelParam &params);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *A_ = nullptr, *B_ = nullptr, *C_ = nullptr;
  GemmKernelParam params_;
};

class LinearInt8 : public Operator, protected KernelLists<LinearI8CpuKernelIF, LinearI8GpuKernelIF> {
public:
  explicit LinearInt8(OpAssistor *assistor);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *A_ = nullptr, *Bt_ = nullptr, *C_ = nullptr;
};

// bf16 linear (LinearNormalBf16 / LinearAvx512Bf16, simd/x86/linear.cpp:137-182): C[M,N] bf16 = bf16 of the
// fp32 sum over k of A[M,K] * Bt[N,K], all three matrices stored as bf16.  Same kernel interface shape as
// linear_int8 (untyped pointers), so plugins are LinearI8*KernelIF structs with type_flag BFLOAT16.
class LinearBf16 : public Operator, protected KernelLists<LinearI8CpuKernelIF, LinearI8GpuKernelIF> {
public:
  explicit LinearBf16(OpAssistor *assistor);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *A_ = nullptr, *Bt_ = nullptr, *C_ = nullptr;
};

class ScanI32 : public Operator, protected KernelLists<ScanCpuKernelIF, ScanGpuKernelIF> {
public:
  explicit ScanI32(OpAssistor *assistor);
  static Operator *Creator(OpAssistor *assistor, std::string &params_str);
  void Help() const override;
  void AddPlugin(KernelInterface *kernel_if, OpRunMode mode) override;
  void ExtractDataTypes(std::vector<int> &type_flags) override;
  void RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  void RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
private:
  void IoCheckAndSet(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) override;
  Array4D *in_ = nullptr, *out_ = nullptr;
};

}  // namespace cux
