// op_assistor.hpp -- ResultChecker and OpAssistor (0-frameworks/cux/include/operator/op_assistor.h).
// ResultChecker keeps cux's rule: the first kernel of a Run is the benchmark, every later one passes
// iff  (sum_i |ref_i - out_i|) * scale < 1e-3  (op_assistor.h:31-46, util.h:137).  Additions: the
// verdicts are recorded so callers can read them (the reference only prints), and integer outputs
// can be compared exactly (the reference keeps the benchmark as float, which is inexact above 2^24).
// OpAssistor owns what operators share; where cux keeps a cuBLAS handle (op_assistor.h:104,117-119)
// this one keeps the cuxb context (device workspace) -- created here, destroyed in the destructor.
#pragma once
#include <vector>

#include "cux/core.hpp"
#include "cux/launch_config.hpp"

namespace cux {

class ResultChecker {
public:
  template <typename DType> bool CheckArray(DType *in, int len, float scale, int id) {
    if (id == 0) {  // benchmark
      ref_.resize((size_t)len);
      for (int i = 0; i < len; i++) ref_[i] = (float)in[i];
      return Record(true);
    }
    float diff = 0.f;
    const int n = std::min<int>(len, (int)ref_.size());
    for (int i = 0; i < n; i++) {
      const float t = ref_[i] - (float)in[i];
      diff += (t >= 0 ? t : -t);
    }
    diff *= scale;
    const float thresh = (float)FLT_PRECISION_ERR_THRESH;
    const bool ok = diff < thresh && n == len;
    CUXLOG_COUT("%s: V0 vs V%d -> (diff_each: %f, thresh: %f, out[0]: %f, %f)", ok ? "Pass" : "Fail", id, diff,
                thresh, ref_.empty() ? 0.f : ref_[0], len > 0 ? (float)in[0] : 0.f);
    return Record(ok);
  }
  // Exact comparison for integer results (bit-exact parity for int8 linear and int32 scan).
  bool CheckArrayExact(const int32_t *in, int len, int id) {
    if (id == 0) {
      iref_.assign(in, in + len);
      return Record(true);
    }
    long long bad = 0;
    const int n = std::min<int>(len, (int)iref_.size());
    for (int i = 0; i < n; i++) bad += (iref_[i] != in[i]);
    const bool ok = bad == 0 && n == len;
    CUXLOG_COUT("%s: V0 vs V%d -> (mismatching elements: %lld of %d, out[0]: %d, %d)", ok ? "Pass" : "Fail", id, bad,
                len, iref_.empty() ? 0 : iref_[0], len > 0 ? in[0] : 0);
    return Record(ok);
  }
  const std::vector<bool> &verdicts() const { return verdicts_; }
  bool all_passed() const {
    for (bool v : verdicts_) if (!v) return false;
    return true;
  }
  void clear_verdicts() { verdicts_.clear(); }

private:
  bool Record(bool ok) { verdicts_.push_back(ok); return ok; }
  std::vector<float> ref_;
  std::vector<int32_t> iref_;
  std::vector<bool> verdicts_;
};

class OpAssistor {
public:
  explicit OpAssistor(Device *device);
  ~OpAssistor();
  OpAssistor(const OpAssistor &) = delete;
  OpAssistor &operator=(const OpAssistor &) = delete;
  inline Device *device() { return device_; }
  inline cuxb_ctx *ctx() { return ctx_; }          // in place of cublas_handle()
  inline LaunchConfig *launch_config() { return launch_config_; }
  inline ResultChecker *checker() { return checker_; }

private:
  Device *device_;
  cuxb_ctx *ctx_;
  LaunchConfig *launch_config_;
  ResultChecker *checker_;
};

}  // namespace cux
