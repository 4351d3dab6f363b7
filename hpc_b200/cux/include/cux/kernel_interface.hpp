// kernel_interface.hpp -- the plugin structs.  Dot / Nrm2 / Gemm entries are field-for-field the
// ones in 0-frameworks/cux/include/operator/kernel_interface.h:16-72, so a plugin written for cux
// (example/dot_product_plugin.cu:32-58, example/gemm_plugin.cu:31-60) fills them unchanged.  The
// entries for the three new operators (reduce_sum, linear_int8, scan_i32) follow the
// same pattern: get_config + func + config_kernel on the device side, func alone on the host side.
#pragma once
#include <functional>
#include <string>

#include "cux/core.hpp"
#include "cux/launch_config.hpp"

namespace cux {

struct KernelInterface {
  TypeFlag type_flag;
  std::string describe_info;
};

// ---- dot
struct DotCpuKernelIF : KernelInterface {
  std::function<void(int len, const void *vec_a, const void *vec_b, void *res)> func;
};
struct DotGpuKernelIF : KernelInterface {
  std::function<Config1D(int len)> get_config;
  std::function<void(Config1D, int len, const void *vec_a, const void *vec_b, void *res)> func;
  void *config_kernel;
};
// ---- nrm2
struct Nrm2CpuKernelIF : KernelInterface {
  std::function<void(int n, const void *x, void *result)> func;
};
struct Nrm2GpuKernelIF : KernelInterface {
  std::function<Config1D(int len)> get_config;
  std::function<void(Config1D, int n, const void *x, void *result)> func;
  void *config_kernel;
};
// ---- gemm
struct GemmCpuKernelIF : KernelInterface {
  std::function<void(const int M, const int N, const int K, const float alpha, const void *A, const int lda,
                     const void *B, const int ldb, const float beta, void *C, const int ldc)> func;
};
struct GemmGpuKernelIF : KernelInterface {
  std::function<Config2D(int M, int N)> get_config;
  std::function<void(Config2D, const int M, const int N, const int K, const float alpha, const void *A,
                     const int lda, const void *B, const int ldb, const float beta, void *C, const int ldc)> func;
  void *config_kernel;
};
// ---- reduce_sum (cuda/reduce_fp32.cu): y = sum(x)
struct ReduceCpuKernelIF : KernelInterface {
  std::function<void(int n, const void *x, void *y)> func;
};
struct ReduceGpuKernelIF : KernelInterface {
  std::function<Config1D(int n)> get_config;
  std::function<void(Config1D, int n, const void *x, void *y)> func;
  void *config_kernel;
};
// ---- linear_int8 (simd/x86/linear.cpp:199-237): C[M,N] int32 = A[M,K] int8 * Bt[N,K]^T int8
struct LinearI8CpuKernelIF : KernelInterface {
  std::function<void(int M, int N, int K, const void *A, const void *Bt, void *C)> func;
};
struct LinearI8GpuKernelIF : KernelInterface {
  std::function<Config2D(int M, int N)> get_config;
  std::function<void(Config2D, int M, int N, int K, const void *A, const void *Bt, void *C)> func;
  void *config_kernel;
};
// ---- scan_i32 (simd/x86/vector_scan.cpp:21-26 in int32): out[i] = in[0] + ... + in[i]
struct ScanCpuKernelIF : KernelInterface {
  std::function<void(int n, const void *in, void *out)> func;
};
struct ScanGpuKernelIF : KernelInterface {
  std::function<Config1D(int n)> get_config;
  std::function<void(Config1D, int n, const void *in, void *out)> func;
  void *config_kernel;
};

}  // namespace cux
