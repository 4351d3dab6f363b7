// ops.cpp -- dot, nrm2, reduce_sum, gemm, linear_int8, scan_i32 behind cux's Operator interface.
//
// Run protocol (kept from the reference so its printed I/W/R/O columns mean the same thing):
//   RunOnHost   (dot_product.cpp:130-163, gemm.cpp:453-495): per kernel -- fetch host data
//               (PUSH_IF_EMPTY), [gemm: save C once, restore C], timed run, check against kernel 0.
//   RunOnDevice (dot_product.cpp:167-214, gemm.cpp:499-556): per kernel -- config, occupancy,
//               fetch device data (cudaMalloc + H2D on first touch), [gemm: save/restore C on the
//               device], one warm-up launch, timed launch between events on the default stream,
//               D2H, check against device kernel 0 with scale 1/len (op_assistor.h:38-46).
// The device kernels are the sm_100a ones in libcuxb; each std::function body is one C-ABI call.
// The single host kernel per op ("Normal") is the plain loop the checker needs as its benchmark
// in host mode; RunOnDevice never touches it -- there is no CPU fallback.
#include <cmath>

#include "cux/ops.hpp"

namespace cux {

namespace {

template <typename IF> void PushKernel(std::vector<IF *> &list, IF *k) { list.push_back(k); }

Config1D StreamConfig(OpAssistor *a, int threads, int ctas_per_sm) {
  Config1D c;
  c.threads_per_block = threads;
  c.blocks_per_grid = a->device()->prop.multiProcessorCount * ctas_per_sm;
  c.virtual_thread_count = c.threads_per_block * c.blocks_per_grid;
  return c;
}

// Every library call made from a plugin body: non-zero status -> the reference's CUDA_CHECK behaviour.
#define DEVICE_CALL(expr) CUXB_CHECK(expr)

}  // namespace

// =====================================================================================  dot
Dot::Dot(OpAssistor *assistor) : Operator(assistor) {
  {  // host V0: DotHostV0's recurrence (dot_product.cpp:7-13) -- t += a[i]*b[i], serial
    auto *k = new DotCpuKernelIF();
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "Normal";
    k->func = [](int len, const void *va, const void *vb, void *res) {
      const float *a = static_cast<const float *>(va), *b = static_cast<const float *>(vb);
      float acc = 0.f;
      for (int i = 0; i < len; i++) acc += a[i] * b[i];
      *static_cast<float *>(res) = acc;
    };
    cpu_kernels_.push_back(k);
  }
  {  // device V0
    auto *k = new DotGpuKernelIF();
    OpAssistor *as = assistor_;
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "cuxb sm_100a: 128-bit streaming loads, fixed-order two-stage reduction, one launch";
    k->get_config = [as](int) { return StreamConfig(as, 512, 2); };
    k->func = [as](Config1D, int len, const void *a, const void *b, void *res) {
      DEVICE_CALL(cuxb_dot_f32(as->ctx(), nullptr, len, (const float *)a, (const float *)b, (float *)res));
    };
    k->config_kernel = const_cast<void *>(KernelTag("dot_f32"));
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *Dot::Creator(OpAssistor *assistor, std::string &) { return new Dot(assistor); }
void Dot::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: Dot Product.");
  CUXLOG_COUT("* Function: sum += a[i] * b[i]");
  CUXLOG_COUT("* Inputs:  [Two] Array4D with one vector each. ");
  CUXLOG_COUT("* Outputs: [One] Array4D with one element.");
  CUXLOG_COUT("* Params:  [None].");
  CUXLOG_COUT("**************************************************");
}
void Dot::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 2 || out.size() != 1) { Help(); CUXLOG_ERR("Error: The dimensions of the input parameters do not match."); }
  if (in[0]->num_element() != in[1]->num_element() || out[0]->num_element() != 1) { Help(); CUXLOG_ERR("Error: The dimensions of the input parameters do not match."); }
  in_a_ = in[0], in_b_ = in[1], out_ = out[0];
}
void Dot::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<DotCpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<DotGpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void Dot::ExtractDataTypes(std::vector<int> &type_flags) {
  std::vector<int> used;
  for (auto *k : cpu_kernels_) used.push_back(k->type_flag);
  for (auto *k : gpu_kernels_) used.push_back(k->type_flag);
  MarkTypes(type_flags, used);
}
void Dot::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("Dot -> CPU: ");
  IoCheckAndSet(input, output);
  const int len = in_a_->num_element();
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) {
    DotCpuKernelIF *kernel = cpu_kernels_[ki];
    const void *a = nullptr, *b = nullptr;
    void *res = nullptr;
    cpu_timer_record_[ki].input = GET_TIME_DIFF(cpu_timer_, TYPE_SWITCH(kernel->type_flag, T, {
      a = in_a_->GetCpuData<T>(PUSH_IF_EMPTY);
      b = in_b_->GetCpuData<T>(PUSH_IF_EMPTY);
      res = out_->GetCpuData<T>(NO_PUSH);
    }););
    cpu_timer_record_[ki].run = GET_TIME_DIFF(cpu_timer_, kernel->func(len, a, b, res););
    TYPE_SWITCH(kernel->type_flag, T,
                assistor_->checker()->CheckArray(out_->GetCpuData<T>(NO_PUSH), out_->num_element(), 1.0f / len, (int)ki););
  }
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_HOST, (int)ki, cpu_kernels_[ki]);
}
void Dot::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("Dot -> GPU:");
  IoCheckAndSet(input, output);
  const int len = in_a_->num_element();
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) {
    DotGpuKernelIF *kernel = gpu_kernels_[ki];
    Config1D config = kernel->get_config(len);
    QueryPotentialOccupancy(kernel->config_kernel, (int)ki, config.threads_per_block, config.shared_memory_size);
    const void *a = nullptr, *b = nullptr;
    void *res = nullptr;
    gpu_timer_record_[ki].input = GET_TIME_DIFF(gpu_timer(), TYPE_SWITCH(kernel->type_flag, T, {
      a = in_a_->GetGpuData<T>(PUSH_IF_EMPTY);
      b = in_b_->GetGpuData<T>(PUSH_IF_EMPTY);
      res = out_->GetGpuData<T>(NO_PUSH);
    }););
    gpu_timer_record_[ki].warnup = GET_TIME_DIFF(gpu_timer(), kernel->func(config, len, a, b, res););
    // the reference clears the result before the timed run (dot_product.cpp:196): kernels that
    // accumulate (atomicAdd) need it, ours overwrite
    CUXB_CHECK(cuxb_memset(res, 0, TypeSize(kernel->type_flag), nullptr));
    gpu_timer_record_[ki].run = GET_TIME_DIFF(gpu_timer(), kernel->func(config, len, a, b, res););
    gpu_timer_record_[ki].output = GET_TIME_DIFF(gpu_timer(), TYPE_SWITCH(kernel->type_flag, T, { out_->GetCpuData<T>(PUSH); }););
    TYPE_SWITCH(kernel->type_flag, T,
                assistor_->checker()->CheckArray(out_->GetCpuData<T>(NO_PUSH), out_->num_element(), 1.0f / len, (int)ki););
  }
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_DEVICE, (int)ki, gpu_kernels_[ki]);
}

// =====================================================================================  nrm2 / reduce_sum
// Both are "one vector in, one scalar out"; they share the run loops.
namespace {
template <typename Op, typename CpuIF, typename GpuIF>
void RunUnaryHost(Op *op, const char *title, std::vector<CpuIF *> &kernels, Array4D *in, Array4D *out) {
  CUXLOG_COUT("%s -> CPU: ", title);
  const int n = in->num_element();
  for (size_t ki = 0; ki < kernels.size(); ki++) {
    CpuIF *kernel = kernels[ki];
    const void *x = nullptr;
    void *res = nullptr;
    op->cpu_timer_record_[ki].input = GET_TIME_DIFF(op->cpu_timer_, TYPE_SWITCH(kernel->type_flag, T, {
      x = in->template GetCpuData<T>(PUSH_IF_EMPTY);
      res = out->template GetCpuData<T>(NO_PUSH);
    }););
    op->cpu_timer_record_[ki].run = GET_TIME_DIFF(op->cpu_timer_, kernel->func(n, x, res););
    TYPE_SWITCH(kernel->type_flag, T,
                op->assistor_->checker()->CheckArray(out->template GetCpuData<T>(NO_PUSH), out->num_element(), 1.0f / n, (int)ki););
  }
  for (size_t ki = 0; ki < kernels.size(); ki++) op->PrintRecordedInfo(OpRunMode::ON_HOST, (int)ki, kernels[ki]);
}
template <typename Op, typename GpuIF>
void RunUnaryDevice(Op *op, const char *title, std::vector<GpuIF *> &kernels, Array4D *in, Array4D *out) {
  CUXLOG_COUT("%s -> GPU: ", title);
  const int n = in->num_element();
  for (size_t ki = 0; ki < kernels.size(); ki++) {
    GpuIF *kernel = kernels[ki];
    Config1D config = kernel->get_config(n);
    op->QueryPotentialOccupancy(kernel->config_kernel, (int)ki, config.threads_per_block, config.shared_memory_size);
    const void *x = nullptr;
    void *res = nullptr;
    op->gpu_timer_record_[ki].input = GET_TIME_DIFF(op->gpu_timer(), TYPE_SWITCH(kernel->type_flag, T, {
      x = in->template GetGpuData<T>(PUSH_IF_EMPTY);
      res = out->template GetGpuData<T>(NO_PUSH);
    }););
    op->gpu_timer_record_[ki].warnup = GET_TIME_DIFF(op->gpu_timer(), kernel->func(config, n, x, res););
    CUXB_CHECK(cuxb_memset(res, 0, TypeSize(kernel->type_flag), nullptr));  // nrm2.cpp:190
    op->gpu_timer_record_[ki].run = GET_TIME_DIFF(op->gpu_timer(), kernel->func(config, n, x, res););
    op->gpu_timer_record_[ki].output = GET_TIME_DIFF(op->gpu_timer(), TYPE_SWITCH(kernel->type_flag, T, { out->template GetCpuData<T>(PUSH); }););
    TYPE_SWITCH(kernel->type_flag, T,
                op->assistor_->checker()->CheckArray(out->template GetCpuData<T>(NO_PUSH), out->num_element(), 1.0f / n, (int)ki););
  }
  for (size_t ki = 0; ki < kernels.size(); ki++) op->PrintRecordedInfo(OpRunMode::ON_DEVICE, (int)ki, kernels[ki]);
}
}  // namespace

Nrm2::Nrm2(OpAssistor *assistor) : Operator(assistor) {
  {  // host V0: Nrm2HostV0 (nrm2.cpp:7-16): n<=0 -> 0, n==1 -> |x0|, else sqrt(sum x^2)
    auto *k = new Nrm2CpuKernelIF();
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "Normal";
    k->func = [](int n, const void *vx, void *result) {
      const float *x = static_cast<const float *>(vx);
      float *r = static_cast<float *>(result);
      if (n <= 0) { *r = 0.f; return; }
      if (n == 1) { *r = std::fabs(x[0]); return; }
      float s = 0.f;
      for (int i = 0; i < n; i++) s += x[i] * x[i];
      *r = std::sqrt(s);
    };
    cpu_kernels_.push_back(k);
  }
  {
    auto *k = new Nrm2GpuKernelIF();
    OpAssistor *as = assistor_;
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "cuxb sm_100a: streaming sum of squares with the sqrt fused into the finishing CTA (one launch)";
    k->get_config = [as](int) { return StreamConfig(as, 512, 2); };
    k->func = [as](Config1D, int n, const void *x, void *result) {
      DEVICE_CALL(cuxb_nrm2_f32(as->ctx(), nullptr, n, (const float *)x, (float *)result));
    };
    k->config_kernel = const_cast<void *>(KernelTag("nrm2_f32"));
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *Nrm2::Creator(OpAssistor *assistor, std::string &) { return new Nrm2(assistor); }
void Nrm2::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: Nrm2.");
  CUXLOG_COUT("* Function: the Euclidean norm of a vector, sqrt(sum x[i]^2)");
  CUXLOG_COUT("* Inputs:  [One] Array4D with one vector. ");
  CUXLOG_COUT("* Outputs: [One] Array4D with one element.");
  CUXLOG_COUT("* Params:  [None].");
  CUXLOG_COUT("**************************************************");
}
void Nrm2::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 1 || out.size() != 1 || out[0]->num_element() != 1) { Help(); CUXLOG_ERR("Error: The dimensions of the input parameters do not match."); }
  in_ = in[0], out_ = out[0];
}
void Nrm2::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<Nrm2CpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<Nrm2GpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void Nrm2::ExtractDataTypes(std::vector<int> &type_flags) {
  std::vector<int> used;
  for (auto *k : cpu_kernels_) used.push_back(k->type_flag);
  for (auto *k : gpu_kernels_) used.push_back(k->type_flag);
  MarkTypes(type_flags, used);
}
void Nrm2::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  IoCheckAndSet(input, output);
  RunUnaryHost<Nrm2, Nrm2CpuKernelIF, Nrm2GpuKernelIF>(this, "Nrm2", cpu_kernels_, in_, out_);
}
void Nrm2::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  IoCheckAndSet(input, output);
  RunUnaryDevice<Nrm2, Nrm2GpuKernelIF>(this, "Nrm2", gpu_kernels_, in_, out_);
}

ReduceSum::ReduceSum(OpAssistor *assistor) : Operator(assistor) {
  {  // host V0: ReduceHost (cuda/reduce_fp32.cu:26-31)
    auto *k = new ReduceCpuKernelIF();
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "Normal";
    k->func = [](int n, const void *vx, void *vy) {
      const float *x = static_cast<const float *>(vx);
      float acc = 0.f;
      for (int i = 0; i < n; i++) acc += x[i];
      *static_cast<float *>(vy) = acc;
    };
    cpu_kernels_.push_back(k);
  }
  {
    auto *k = new ReduceGpuKernelIF();
    OpAssistor *as = assistor_;
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "cuxb sm_100a: 128-bit streaming loads, fixed-order two-stage reduction, one launch";
    k->get_config = [as](int) { return StreamConfig(as, 512, 2); };
    k->func = [as](Config1D, int n, const void *x, void *y) {
      DEVICE_CALL(cuxb_sum_f32(as->ctx(), nullptr, n, (const float *)x, (float *)y));
    };
    k->config_kernel = const_cast<void *>(KernelTag("sum_f32"));
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *ReduceSum::Creator(OpAssistor *assistor, std::string &) { return new ReduceSum(assistor); }
void ReduceSum::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: ReduceSum.");
  CUXLOG_COUT("* Function: y = sum x[i]");
  CUXLOG_COUT("* Inputs:  [One] Array4D with one vector. ");
  CUXLOG_COUT("* Outputs: [One] Array4D with one element.");
  CUXLOG_COUT("**************************************************");
}
void ReduceSum::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 1 || out.size() != 1 || out[0]->num_element() != 1) { Help(); CUXLOG_ERR("Error: The dimensions of the input parameters do not match."); }
  in_ = in[0], out_ = out[0];
}
void ReduceSum::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<ReduceCpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<ReduceGpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void ReduceSum::ExtractDataTypes(std::vector<int> &type_flags) {
  std::vector<int> used;
  for (auto *k : cpu_kernels_) used.push_back(k->type_flag);
  for (auto *k : gpu_kernels_) used.push_back(k->type_flag);
  MarkTypes(type_flags, used);
}
void ReduceSum::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  IoCheckAndSet(input, output);
  RunUnaryHost<ReduceSum, ReduceCpuKernelIF, ReduceGpuKernelIF>(this, "ReduceSum", cpu_kernels_, in_, out_);
}
void ReduceSum::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  IoCheckAndSet(input, output);
  RunUnaryDevice<ReduceSum, ReduceGpuKernelIF>(this, "ReduceSum", gpu_kernels_, in_, out_);
}

// =====================================================================================  gemm
Gemm::Gemm(OpAssistor *assistor, GemmKernelParam &params) : Operator(assistor), params_(params) {
  {  // host V0: the naive triple loop of GemmHostV0 (gemm.cpp:13-34): C *= beta, then per element
     // t = C; t += (alpha*a)*b over k
    auto *k = new GemmCpuKernelIF();
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "Normal";
    k->func = [](const int M, const int N, const int K, const float alpha, const void *vA, const int lda,
                 const void *vB, const int ldb, const float beta, void *vC, const int ldc) {
      const float *A = static_cast<const float *>(vA), *B = static_cast<const float *>(vB);
      float *C = static_cast<float *>(vC);
      for (int r = 0; r < M; r++) {
        float *crow = C + (size_t)r * ldc;
        const float *arow = A + (size_t)r * lda;
        for (int c = 0; c < N; c++) {
          float t = crow[c] * beta;
          for (int k2 = 0; k2 < K; k2++) t += alpha * arow[k2] * B[(size_t)k2 * ldb + c];
          crow[c] = t;
        }
      }
    };
    cpu_kernels_.push_back(k);
  }
  OpAssistor *as = assistor_;
  auto tile_config = [](int M, int N, int bm, int bn, int threads) {
    Config2D c;
    c.virtual_thread_count = dim3(N, M, 1);
    c.threads_per_block = dim3(threads, 1, 1);
    c.blocks_per_grid = dim3(DivUp(N, bn), DivUp(M, bm), 1);
    return c;
  };
  {  // device V0: fp32 on CUDA cores
    auto *k = new GemmGpuKernelIF();
    k->type_flag = TypeFlag::FLOAT32;
    k->describe_info = "cuxb sm_100a SGEMM: 128x128x16 CTA tile, 8x8 micro-tile on packed FFMA2, double-buffered shared memory, deterministic split-K for small problems";
    k->get_config = [tile_config](int M, int N) { return tile_config(M, N, 128, 128, 256); };
    k->func = [as](Config2D, const int M, const int N, const int K, const float alpha, const void *A, const int lda,
                   const void *B, const int ldb, const float beta, void *C, const int ldc) {
      DEVICE_CALL(cuxb_sgemm_f32(as->ctx(), nullptr, M, N, K, alpha, (const float *)A, lda, (const float *)B, ldb, beta,
                                 (float *)C, ldc));
    };
    k->config_kernel = const_cast<void *>(KernelTag("sgemm_f32"));
    gpu_kernels_.push_back(k);
  }
  {  // device V1: fp16 storage, fp32 accumulate on tcgen05 (the reference's slot: cublasSgemmEx, gemm.cu:266-269)
    auto *k = new GemmGpuKernelIF();
    k->type_flag = TypeFlag::FLOAT16;
    k->describe_info = "cuxb sm_100a HGEMM: TMA -> tcgen05.mma (2-CTA 256x256x16) -> TMEM, fp32 accumulate, fp16 C";
    k->get_config = [tile_config](int M, int N) { return tile_config(M, N, 256, 256, 256); };
    k->func = [as](Config2D, const int M, const int N, const int K, const float alpha, const void *A, const int lda,
                   const void *B, const int ldb, const float beta, void *C, const int ldc) {
      DEVICE_CALL(cuxb_hgemm_f16_f32acc(as->ctx(), nullptr, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, CUXB_FLOAT16));
    };
    k->config_kernel = const_cast<void *>(KernelTag("hgemm_f16"));
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *Gemm::Creator(OpAssistor *assistor, std::string &params_str) {
  GemmKernelParam params;  // "alpha: 1.0, beta: 3.0" (demo.cpp:77); strtof instead of the reference's atoi
  if (params_str.find("alpha:") != std::string::npos)
    params.alpha = std::strtof(StrProcessor::FetchSubStr(params_str, "alpha:", ",").c_str(), nullptr);
  if (params_str.find("beta:") != std::string::npos)
    params.beta = std::strtof(StrProcessor::FetchSubStr(params_str, "beta:", ",").c_str(), nullptr);
  return new Gemm(assistor, params);
}
void Gemm::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: Gemm.");
  CUXLOG_COUT("* Function: C(M, N) = alpha * A(M, K) * B(K, N) + beta * C -> (height, width)");
  CUXLOG_COUT("* Inputs:  [Two] Array4D with one matrix each. ");
  CUXLOG_COUT("* Outputs: [One] Array4D with one matrix.");
  CUXLOG_COUT("* Params:  [Two] alpha / beta -> alpha: 1.0, beta: 0.0");
  CUXLOG_COUT("**************************************************");
}
void Gemm::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 2 || out.size() != 1) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  const std::vector<int> a = in[0]->shape(), b = in[1]->shape(), c = out[0]->shape();
  if (a[WIDTH] != b[HEIGHT] || a[HEIGHT] != c[HEIGHT] || b[WIDTH] != c[WIDTH]) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  A_ = in[0], B_ = in[1], C_ = out[0];
}
void Gemm::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<GemmCpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<GemmGpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void Gemm::ExtractDataTypes(std::vector<int> &type_flags) {
  std::vector<int> used;
  for (auto *k : cpu_kernels_) used.push_back(k->type_flag);
  for (auto *k : gpu_kernels_) used.push_back(k->type_flag);
  MarkTypes(type_flags, used);
}
void Gemm::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("Gemm -> CPU: ");
  IoCheckAndSet(input, output);
  const int M = A_->shape()[HEIGHT], N = B_->shape()[WIDTH], K = B_->shape()[HEIGHT];
  const int lda = K, ldb = N, ldc = N;  // gemm.cpp:458-463
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) {
    GemmCpuKernelIF *kernel = cpu_kernels_[ki];
    const void *A = nullptr, *B = nullptr;
    void *C = nullptr;
    cpu_timer_record_[ki].input = GET_TIME_DIFF(cpu_timer_, TYPE_SWITCH(kernel->type_flag, T, {
      A = A_->GetCpuData<T>(PUSH_IF_EMPTY);
      B = B_->GetCpuData<T>(PUSH_IF_EMPTY);
      C = C_->GetCpuData<T>(PUSH_IF_EMPTY);
    }););
    C_->Save(kernel->type_flag, ON_HOST, /*is_save_if_empty=*/true);  // every kernel starts from the same C (beta)
    C_->Restore(kernel->type_flag, ON_HOST);
    cpu_timer_record_[ki].run = GET_TIME_DIFF(cpu_timer_, kernel->func(M, N, K, params_.alpha, A, lda, B, ldb, params_.beta, C, ldc););
    TYPE_SWITCH(kernel->type_flag, T,
                assistor_->checker()->CheckArray(C_->GetCpuData<T>(NO_PUSH), C_->num_element(), 1.0f / C_->num_element(), (int)ki););
  }
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_HOST, (int)ki, cpu_kernels_[ki]);
}
void Gemm::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("Gemm -> GPU: ");
  IoCheckAndSet(input, output);
  const int M = A_->shape()[HEIGHT], N = B_->shape()[WIDTH], K = B_->shape()[HEIGHT];
  const int lda = K, ldb = N, ldc = N;
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) {
    GemmGpuKernelIF *kernel = gpu_kernels_[ki];
    Config2D config = kernel->get_config(M, N);
    QueryPotentialOccupancy(kernel->config_kernel, (int)ki, config.threads_per_block.x * config.threads_per_block.y,
                            config.shared_memory_size);
    const void *A = nullptr, *B = nullptr;
    void *C = nullptr;
    gpu_timer_record_[ki].input = GET_TIME_DIFF(gpu_timer(), TYPE_SWITCH(kernel->type_flag, T, {
      A = A_->GetGpuData<T>(PUSH_IF_EMPTY);
      B = B_->GetGpuData<T>(PUSH_IF_EMPTY);
      C = C_->GetGpuData<T>(PUSH_IF_EMPTY);
    }););
    C_->Save(kernel->type_flag, ON_DEVICE, true);  // gemm.cpp:531-532
    gpu_timer_record_[ki].warnup = GET_TIME_DIFF(gpu_timer(), kernel->func(config, M, N, K, params_.alpha, A, lda, B, ldb, params_.beta, C, ldc););
    C_->Restore(kernel->type_flag, ON_DEVICE);
    gpu_timer_record_[ki].run = GET_TIME_DIFF(gpu_timer(), kernel->func(config, M, N, K, params_.alpha, A, lda, B, ldb, params_.beta, C, ldc););
    gpu_timer_record_[ki].output = GET_TIME_DIFF(gpu_timer(), TYPE_SWITCH(kernel->type_flag, T, { C_->GetCpuData<T>(PUSH); }););
    TYPE_SWITCH(kernel->type_flag, T,
                assistor_->checker()->CheckArray(C_->GetCpuData<T>(NO_PUSH), C_->num_element(), 1.0f / C_->num_element(), (int)ki););
  }
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_DEVICE, (int)ki, gpu_kernels_[ki]);
}

// =====================================================================================  linear_int8
LinearInt8::LinearInt8(OpAssistor *assistor) : Operator(assistor) {
  {  // host V0: LinearNormalInt8 (simd/x86/linear.cpp:199-211)
    auto *k = new LinearI8CpuKernelIF();
    k->type_flag = TypeFlag::INT8;
    k->describe_info = "Normal";
    k->func = [](int M, int N, int K, const void *vA, const void *vB, void *vC) {
      const int8_t *A = static_cast<const int8_t *>(vA), *B = static_cast<const int8_t *>(vB);
      int32_t *C = static_cast<int32_t *>(vC);
      for (int r = 0; r < M; r++)
        for (int c = 0; c < N; c++) {
          int32_t acc = 0;
          const int8_t *a = A + (size_t)r * K, *b = B + (size_t)c * K;
          for (int k2 = 0; k2 < K; k2++) acc += (int32_t)a[k2] * (int32_t)b[k2];
          C[(size_t)r * N + c] = acc;
        }
    };
    cpu_kernels_.push_back(k);
  }
  {
    auto *k = new LinearI8GpuKernelIF();
    OpAssistor *as = assistor_;
    k->type_flag = TypeFlag::INT8;
    k->describe_info = "cuxb sm_100a: TMA -> tcgen05.mma kind::i8 (2-CTA 256x256x32) -> TMEM int32";
    k->get_config = [](int M, int N) {
      Config2D c;
      c.virtual_thread_count = dim3(N, M, 1);
      c.threads_per_block = dim3(256, 1, 1);
      c.blocks_per_grid = dim3(DivUp(N, 256), DivUp(M, 256), 1);
      return c;
    };
    k->func = [as](Config2D, int M, int N, int K, const void *A, const void *Bt, void *C) {
      DEVICE_CALL(cuxb_linear_s8_s32(as->ctx(), nullptr, M, N, K, (const int8_t *)A, (const int8_t *)Bt, (int32_t *)C));
    };
    k->config_kernel = nullptr;
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *LinearInt8::Creator(OpAssistor *assistor, std::string &) { return new LinearInt8(assistor); }
void LinearInt8::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: LinearInt8.");
  CUXLOG_COUT("* Function: C(M, N) int32 = A(M, K) int8 * Bt(N, K)^T int8");
  CUXLOG_COUT("* Inputs:  [Two] Array4D: A with (height, width) = (M, K), Bt with (N, K). ");
  CUXLOG_COUT("* Outputs: [One] Array4D with (M, N).");
  CUXLOG_COUT("**************************************************");
}
void LinearInt8::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 2 || out.size() != 1) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  const std::vector<int> a = in[0]->shape(), b = in[1]->shape(), c = out[0]->shape();
  if (a[WIDTH] != b[WIDTH] || a[HEIGHT] != c[HEIGHT] || b[HEIGHT] != c[WIDTH]) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  A_ = in[0], Bt_ = in[1], C_ = out[0];
}
void LinearInt8::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<LinearI8CpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<LinearI8GpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void LinearInt8::ExtractDataTypes(std::vector<int> &type_flags) { MarkTypes(type_flags, {INT8}); }
void LinearInt8::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("LinearInt8 -> CPU: ");
  IoCheckAndSet(input, output);
  const int M = A_->shape()[HEIGHT], K = A_->shape()[WIDTH], N = Bt_->shape()[HEIGHT];
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) {
    LinearI8CpuKernelIF *kernel = cpu_kernels_[ki];
    const void *A = nullptr, *B = nullptr;
    int32_t *C = nullptr;
    cpu_timer_record_[ki].input = GET_TIME_DIFF(cpu_timer_, {
      A = A_->GetCpuData<int8_t>(PUSH_IF_EMPTY);
      B = Bt_->GetCpuData<int8_t>(PUSH_IF_EMPTY);
      C = C_->GetCpuData<int32_t>(NO_PUSH);
    });
    cpu_timer_record_[ki].run = GET_TIME_DIFF(cpu_timer_, kernel->func(M, N, K, A, B, C););
    assistor_->checker()->CheckArrayExact(C, C_->num_element(), (int)ki);
  }
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_HOST, (int)ki, cpu_kernels_[ki]);
}
void LinearInt8::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("LinearInt8 -> GPU: ");
  IoCheckAndSet(input, output);
  const int M = A_->shape()[HEIGHT], K = A_->shape()[WIDTH], N = Bt_->shape()[HEIGHT];
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) {
    LinearI8GpuKernelIF *kernel = gpu_kernels_[ki];
    Config2D config = kernel->get_config(M, N);
    QueryPotentialOccupancy(kernel->config_kernel, (int)ki, config.threads_per_block.x, config.shared_memory_size);
    const void *A = nullptr, *B = nullptr;
    void *C = nullptr;
    gpu_timer_record_[ki].input = GET_TIME_DIFF(gpu_timer(), {
      A = A_->GetGpuData<int8_t>(PUSH_IF_EMPTY);
      B = Bt_->GetGpuData<int8_t>(PUSH_IF_EMPTY);
      C = C_->GetGpuData<int32_t>(NO_PUSH);
    });
    gpu_timer_record_[ki].warnup = GET_TIME_DIFF(gpu_timer(), kernel->func(config, M, N, K, A, B, C););
    CUXB_CHECK(cuxb_memset(C, 0xff, (size_t)C_->num_element() * 4, nullptr));  // the timed run must rewrite every element
    gpu_timer_record_[ki].run = GET_TIME_DIFF(gpu_timer(), kernel->func(config, M, N, K, A, B, C););
    gpu_timer_record_[ki].output = GET_TIME_DIFF(gpu_timer(), { C_->GetCpuData<int32_t>(PUSH); });
    assistor_->checker()->CheckArrayExact(C_->GetCpuData<int32_t>(NO_PUSH), C_->num_element(), (int)ki);
  }
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_DEVICE, (int)ki, gpu_kernels_[ki]);
}

// =====================================================================================  linear_bf16
LinearBf16::LinearBf16(OpAssistor *assistor) : Operator(assistor) {
  {  // host V0: LinearNormalBf16 (simd/x86/linear.cpp:137-151): fp32 accumulate over k in order, one rounding to bf16
    auto *k = new LinearI8CpuKernelIF();
    k->type_flag = TypeFlag::BFLOAT16;
    k->describe_info = "Normal";
    k->func = [](int M, int N, int K, const void *vA, const void *vB, void *vC) {
      const bfloat16 *A = static_cast<const bfloat16 *>(vA), *B = static_cast<const bfloat16 *>(vB);
      bfloat16 *C = static_cast<bfloat16 *>(vC);
      for (int r = 0; r < M; r++)
        for (int c = 0; c < N; c++) {
          float acc = 0.f;
          const bfloat16 *a = A + (size_t)r * K, *b = B + (size_t)c * K;
          for (int k2 = 0; k2 < K; k2++) acc += (float)a[k2] * (float)b[k2];
          C[(size_t)r * N + c] = bfloat16(acc);
        }
    };
    cpu_kernels_.push_back(k);
  }
  {
    auto *k = new LinearI8GpuKernelIF();
    OpAssistor *as = assistor_;
    k->type_flag = TypeFlag::BFLOAT16;
    k->describe_info = "cuxb sm_100a: TMA -> tcgen05.mma kind::f16 bf16 operands (2-CTA 256x256x16) -> TMEM fp32 -> bf16 C";
    k->get_config = [](int M, int N) {
      Config2D c;
      c.virtual_thread_count = dim3(N, M, 1);
      c.threads_per_block = dim3(256, 1, 1);
      c.blocks_per_grid = dim3(DivUp(N, 256), DivUp(M, 256), 1);
      return c;
    };
    k->func = [as](Config2D, int M, int N, int K, const void *A, const void *Bt, void *C) {
      DEVICE_CALL(cuxb_linear_bf16_f32acc(as->ctx(), nullptr, M, N, K, A, Bt, C, CUXB_BFLOAT16));
    };
    k->config_kernel = nullptr;
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *LinearBf16::Creator(OpAssistor *assistor, std::string &) { return new LinearBf16(assistor); }
void LinearBf16::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: LinearBf16.");
  CUXLOG_COUT("* Function: C(M, N) bf16 = A(M, K) bf16 * Bt(N, K)^T bf16, fp32 accumulate");
  CUXLOG_COUT("* Inputs:  [Two] Array4D: A with (height, width) = (M, K), Bt with (N, K). ");
  CUXLOG_COUT("* Outputs: [One] Array4D with (M, N).");
  CUXLOG_COUT("**************************************************");
}
void LinearBf16::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 2 || out.size() != 1) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  const std::vector<int> a = in[0]->shape(), b = in[1]->shape(), c = out[0]->shape();
  if (a[WIDTH] != b[WIDTH] || a[HEIGHT] != c[HEIGHT] || b[HEIGHT] != c[WIDTH]) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  A_ = in[0], Bt_ = in[1], C_ = out[0];
}
void LinearBf16::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<LinearI8CpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<LinearI8GpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void LinearBf16::ExtractDataTypes(std::vector<int> &type_flags) { MarkTypes(type_flags, {BFLOAT16}); }
void LinearBf16::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("LinearBf16 -> CPU: ");
  IoCheckAndSet(input, output);
  const int M = A_->shape()[HEIGHT], K = A_->shape()[WIDTH], N = Bt_->shape()[HEIGHT];
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) {
    LinearI8CpuKernelIF *kernel = cpu_kernels_[ki];
    const void *A = nullptr, *B = nullptr;
    bfloat16 *C = nullptr;
    cpu_timer_record_[ki].input = GET_TIME_DIFF(cpu_timer_, {
      A = A_->GetCpuData<bfloat16>(PUSH_IF_EMPTY);
      B = Bt_->GetCpuData<bfloat16>(PUSH_IF_EMPTY);
      C = C_->GetCpuData<bfloat16>(NO_PUSH);
    });
    cpu_timer_record_[ki].run = GET_TIME_DIFF(cpu_timer_, kernel->func(M, N, K, A, B, C););
    assistor_->checker()->CheckArray(C, C_->num_element(), 1.0f / C_->num_element(), (int)ki);
  }
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_HOST, (int)ki, cpu_kernels_[ki]);
}
void LinearBf16::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("LinearBf16 -> GPU: ");
  IoCheckAndSet(input, output);
  const int M = A_->shape()[HEIGHT], K = A_->shape()[WIDTH], N = Bt_->shape()[HEIGHT];
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) {
    LinearI8GpuKernelIF *kernel = gpu_kernels_[ki];
    Config2D config = kernel->get_config(M, N);
    QueryPotentialOccupancy(kernel->config_kernel, (int)ki, config.threads_per_block.x, config.shared_memory_size);
    const void *A = nullptr, *B = nullptr;
    void *C = nullptr;
    gpu_timer_record_[ki].input = GET_TIME_DIFF(gpu_timer(), {
      A = A_->GetGpuData<bfloat16>(PUSH_IF_EMPTY);
      B = Bt_->GetGpuData<bfloat16>(PUSH_IF_EMPTY);
      C = C_->GetGpuData<bfloat16>(NO_PUSH);
    });
    gpu_timer_record_[ki].warnup = GET_TIME_DIFF(gpu_timer(), kernel->func(config, M, N, K, A, B, C););
    CUXB_CHECK(cuxb_memset(C, 0xff, (size_t)C_->num_element() * 2, nullptr));  // the timed run must rewrite every element
    gpu_timer_record_[ki].run = GET_TIME_DIFF(gpu_timer(), kernel->func(config, M, N, K, A, B, C););
    gpu_timer_record_[ki].output = GET_TIME_DIFF(gpu_timer(), { C_->GetCpuData<bfloat16>(PUSH); });
    assistor_->checker()->CheckArray(C_->GetCpuData<bfloat16>(NO_PUSH), C_->num_element(), 1.0f / C_->num_element(), (int)ki);
  }
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_DEVICE, (int)ki, gpu_kernels_[ki]);
}

// =====================================================================================  scan_i32
ScanI32::ScanI32(OpAssistor *assistor) : Operator(assistor) {
  {  // host V0: VectorScanNormal's recurrence (simd/x86/vector_scan.cpp:21-26) on int32, wrap-around adds
    auto *k = new ScanCpuKernelIF();
    k->type_flag = TypeFlag::INT32;
    k->describe_info = "Normal";
    k->func = [](int n, const void *vin, void *vout) {
      const int32_t *in = static_cast<const int32_t *>(vin);
      int32_t *out = static_cast<int32_t *>(vout);
      uint32_t run = 0;
      for (int i = 0; i < n; i++) {
        run += (uint32_t)in[i];
        out[i] = (int32_t)run;
      }
    };
    cpu_kernels_.push_back(k);
  }
  {
    auto *k = new ScanGpuKernelIF();
    OpAssistor *as = assistor_;
    k->type_flag = TypeFlag::INT32;
    k->describe_info = "cuxb sm_100a: single-pass scan: one persistent CTA per SM, cp.async ring, 8192-element tiles, aggregate gathering";
    k->get_config = [](int n) {
      Config1D c;
      c.virtual_thread_count = n;
      c.threads_per_block = 512;
      c.blocks_per_grid = std::min(DivUp(n, 8192), 148);
      c.shared_memory_size = 128 * 1024;
      return c;
    };
    k->func = [as](Config1D, int n, const void *in, void *out) {
      DEVICE_CALL(cuxb_scan_i32(as->ctx(), nullptr, n, (const int32_t *)in, (int32_t *)out, 0));
    };
    k->config_kernel = const_cast<void *>(KernelTag("scan_i32"));
    gpu_kernels_.push_back(k);
  }
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
Operator *ScanI32::Creator(OpAssistor *assistor, std::string &) { return new ScanI32(assistor); }
void ScanI32::Help() const {
  CUXLOG_COUT("***************** Op Helper ********************");
  CUXLOG_COUT("* Name: ScanI32.");
  CUXLOG_COUT("* Function: inclusive prefix sum, out[i] = in[0] + ... + in[i] (int32, wrap-around)");
  CUXLOG_COUT("* Inputs:  [One] Array4D with one vector. ");
  CUXLOG_COUT("* Outputs: [One] Array4D with one vector of the same length.");
  CUXLOG_COUT("**************************************************");
}
void ScanI32::IoCheckAndSet(const std::vector<Array4D *> &in, const std::vector<Array4D *> &out) {
  if (in.size() != 1 || out.size() != 1 || in[0]->num_element() != out[0]->num_element()) { Help(); CUXLOG_ERR("The dimensions of the data do not match."); }
  in_ = in[0], out_ = out[0];
}
void ScanI32::AddPlugin(KernelInterface *kernel_if, OpRunMode mode) {
  if (mode == OpRunMode::ON_HOST) cpu_kernels_.push_back(static_cast<ScanCpuKernelIF *>(kernel_if));
  else gpu_kernels_.push_back(static_cast<ScanGpuKernelIF *>(kernel_if));
  ResetByKernelNum(cpu_kernels_.size(), gpu_kernels_.size());
}
void ScanI32::ExtractDataTypes(std::vector<int> &type_flags) { MarkTypes(type_flags, {INT32}); }
void ScanI32::RunOnHost(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("ScanI32 -> CPU: ");
  IoCheckAndSet(input, output);
  const int n = in_->num_element();
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) {
    ScanCpuKernelIF *kernel = cpu_kernels_[ki];
    const void *x = nullptr;
    int32_t *y = nullptr;
    cpu_timer_record_[ki].input = GET_TIME_DIFF(cpu_timer_, {
      x = in_->GetCpuData<int32_t>(PUSH_IF_EMPTY);
      y = out_->GetCpuData<int32_t>(NO_PUSH);
    });
    cpu_timer_record_[ki].run = GET_TIME_DIFF(cpu_timer_, kernel->func(n, x, y););
    assistor_->checker()->CheckArrayExact(y, n, (int)ki);
  }
  for (size_t ki = 0; ki < cpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_HOST, (int)ki, cpu_kernels_[ki]);
}
void ScanI32::RunOnDevice(const std::vector<Array4D *> &input, const std::vector<Array4D *> &output) {
  CUXLOG_COUT("ScanI32 -> GPU: ");
  IoCheckAndSet(input, output);
  const int n = in_->num_element();
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) {
    ScanGpuKernelIF *kernel = gpu_kernels_[ki];
    Config1D config = kernel->get_config(n);
    QueryPotentialOccupancy(kernel->config_kernel, (int)ki, config.threads_per_block, config.shared_memory_size);
    const void *x = nullptr;
    void *y = nullptr;
    gpu_timer_record_[ki].input = GET_TIME_DIFF(gpu_timer(), {
      x = in_->GetGpuData<int32_t>(PUSH_IF_EMPTY);
      y = out_->GetGpuData<int32_t>(NO_PUSH);
    });
    gpu_timer_record_[ki].warnup = GET_TIME_DIFF(gpu_timer(), kernel->func(config, n, x, y););
    CUXB_CHECK(cuxb_memset(y, 0xff, (size_t)n * 4, nullptr));
    gpu_timer_record_[ki].run = GET_TIME_DIFF(gpu_timer(), kernel->func(config, n, x, y););
    gpu_timer_record_[ki].output = GET_TIME_DIFF(gpu_timer(), { out_->GetCpuData<int32_t>(PUSH); });
    assistor_->checker()->CheckArrayExact(out_->GetCpuData<int32_t>(NO_PUSH), n, (int)ki);
  }
  for (size_t ki = 0; ki < gpu_kernels_.size(); ki++) PrintRecordedInfo(OpRunMode::ON_DEVICE, (int)ki, gpu_kernels_[ki]);
}

}  // namespace cux
