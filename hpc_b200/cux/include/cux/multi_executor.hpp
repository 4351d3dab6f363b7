// multi_executor.hpp -- Executor::Initialize(dev) (0-frameworks/cux/include/executor.h:74-103,
// src/executor.cpp:10-22) generalised to g devices of one box, for the master / worker flow of the
// reference's MPI GEMM (mpi/alg_matrix_multiply.cpp:52-80): the master cuts A and C into row blocks, every
// worker multiplies its block with the one B, the master gathers C and checks it.  Here the "workers" are
// GPUs driven from this one host thread through the C ABI's cuxb_mg driver (include/cuxb.h); B never
// leaves rank 0 on the host side -- the GEMM kernels fetch it over NVLink while they multiply.  The gathered
// result is judged by cux's own ResultChecker (op_assistor.h:31-71): id 0 = the expected values the caller
// got from the host kernel, id 1 = what the GPUs produced.
#pragma once
#include <vector>

#include "cux/op_assistor.hpp"

namespace cux {

class MultiExecutor {
public:
  MultiExecutor() : mg_(nullptr) {}
  ~MultiExecutor() { Clear(); }
  void InitializeMulti(const std::vector<int> &device_ids);
  void Clear();
  int world() const;

  // dot over `len` elements cut into contiguous slices; `expect` = DotHostV0's result
  bool RunDot(const float *a, const float *b, int len, float expect);
  // C = alpha * A[M,K] * B[K,N] + beta * C0, rows of A / C cut into blocks of whole 128-row tiles
  bool RunGemm(int M, int N, int K, float alpha, const float *A, const float *B, float beta, const float *C0,
               const float *expect);
  // inclusive int32 scan over contiguous slices, bit-exact
  bool RunScanI32(const int32_t *in, int len, const int32_t *expect);
  bool AllChecksPassed() const { return checker_.all_passed(); }
  float last_ms() const { return last_ms_; }   // host wall clock around launch-on-all-devices .. all synchronised

private:
  cuxb_mg *mg_;
  ResultChecker checker_;
  float last_ms_ = 0.f;
};

}  // namespace cux
