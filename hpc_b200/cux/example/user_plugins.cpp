// user_plugins.cpp -- what a cux user writes to add a kernel to an existing op: build a
// *GpuKernelIF on the heap, fill type_flag / describe_info / get_config / func / config_kernel and
// hand it to Executor::AddPlugin (the op deletes it).  Same role as the reference's
// example/dot_product_plugin.cu:32-58 and example/gemm_plugin.cu:31-60, but the bodies are plain
// C++: they call the C ABI instead of launching a __global__ function, so no nvcc is needed.
#include "executor.h"
#include "operator/kernel_interface.h"

static cuxb_ctx *PluginCtx() {
  static cuxb_ctx *ctx = nullptr;
  if (ctx == nullptr) CUXB_CHECK(cuxb_ctx_create(0, &ctx));
  return ctx;
}

cux::KernelInterface *DotProductGPUPlugin() {
  using namespace cux;
  DotGpuKernelIF *kernel = new DotGpuKernelIF();
  kernel->type_flag = TypeFlag::FLOAT32;
  kernel->describe_info = "Plugin example: dot through cuxb_dot_f32 with its own context";
  kernel->get_config = [](int len) {
    Config1D config;
    config.threads_per_block = 512;
    config.blocks_per_grid = 2 * 148;
    config.virtual_thread_count = len;
    return config;
  };
  kernel->func = [](Config1D, int len, const void *vec_a, const void *vec_b, void *res) {
    CUXB_CHECK(cuxb_dot_f32(PluginCtx(), nullptr, len, (const float *)vec_a, (const float *)vec_b, (float *)res));
  };
  kernel->config_kernel = nullptr;
  return kernel;
}

cux::KernelInterface *GemmGPUPlugin() {
  using namespace cux;
  GemmGpuKernelIF *kernel = new GemmGpuKernelIF();
  kernel->type_flag = TypeFlag::FLOAT32;
  kernel->describe_info = "Plugin example: sgemm through cuxb_sgemm_f32 with its own context";
  kernel->get_config = [](int M, int N) {
    Config2D config;
    config.threads_per_block = dim3(256, 1, 1);
    config.blocks_per_grid = dim3((N + 127) / 128, (M + 127) / 128, 1);
    return config;
  };
  kernel->func = [](Config2D, const int M, const int N, const int K, const float alpha, const void *A, const int lda,
                    const void *B, const int ldb, const float beta, void *C, const int ldc) {
    CUXB_CHECK(cuxb_sgemm_f32(PluginCtx(), nullptr, M, N, K, alpha, (const float *)A, lda, (const float *)B, ldb, beta,
                              (float *)C, ldc));
  };
  kernel->config_kernel = nullptr;
  return kernel;
}
