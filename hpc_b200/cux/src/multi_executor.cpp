// MultiExecutor: cux's Executor over g GPUs (see cux/multi_executor.hpp), on the cuxb_mg driver of the C ABI.
#include "cux/multi_executor.hpp"

#include <chrono>

#include "cux/core.hpp"

namespace cux {

namespace {
// contiguous slice [lo, hi) of n items for rank r of w; `grain` keeps slice sizes multiples of a tile
void Shard(long long n, int w, int r, long long grain, long long *lo, long long *hi) {
  const long long units = n / grain, base = units / w, rem = units % w;
  *lo = (r * base + (r < rem ? r : rem)) * grain;
  *hi = *lo + (base + (r < rem ? 1 : 0)) * grain;
  if (r == w - 1) *hi = n;   // the last rank also takes what does not fill a unit
}
struct DeviceBuf {   // device memory on rank r's GPU, freed on scope exit
  void *p = nullptr;
  DeviceBuf() {}
  DeviceBuf(const DeviceBuf &) = delete;
  ~DeviceBuf() { if (p) cuxb_free(p); }
  void Alloc(int device, size_t bytes) {
    CUXB_CHECK(cuxb_set_device(device));
    CUXB_CHECK(cuxb_malloc(&p, bytes ? bytes : 4));
  }
};
double NowMs() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

void MultiExecutor::InitializeMulti(const std::vector<int> &device_ids) {
  Clear();
  CUXB_CHECK(cuxb_mg_create((int)device_ids.size(), device_ids.data(), &mg_));
  CUXLOG_COUT("* MultiExecutor: %d GPU(s), one rank per GPU, driven from one host thread.", cuxb_mg_world(mg_));
}

void MultiExecutor::Clear() {
  if (mg_) CUXB_CHECK(cuxb_mg_destroy(mg_));
  mg_ = nullptr;
}

int MultiExecutor::world() const { return mg_ ? cuxb_mg_world(mg_) : 0; }

bool MultiExecutor::RunDot(const float *a, const float *b, int len, float expect) {
  if (!mg_) CUXLOG_ERR("MultiExecutor::RunDot before InitializeMulti.");
  const int w = world();
  std::vector<DeviceBuf> da(w), db(w), dr(w);
  std::vector<long long> n(w);
  std::vector<const float *> pa(w), pb(w);
  std::vector<float *> pr(w);
  for (int r = 0; r < w; r++) {
    long long lo, hi;
    Shard(len, w, r, 1, &lo, &hi);
    n[r] = hi - lo;
    const int dev = cuxb_mg_device(mg_, r);
    da[r].Alloc(dev, n[r] * 4), db[r].Alloc(dev, n[r] * 4), dr[r].Alloc(dev, 4);
    CUXB_CHECK(cuxb_memcpy_h2d(da[r].p, a + lo, n[r] * 4, nullptr));
    CUXB_CHECK(cuxb_memcpy_h2d(db[r].p, b + lo, n[r] * 4, nullptr));
    pa[r] = (const float *)da[r].p, pb[r] = (const float *)db[r].p, pr[r] = (float *)dr[r].p;
  }
  const double t0 = NowMs();
  CUXB_CHECK(cuxb_mg_all_dot_f32(mg_, n.data(), pa.data(), pb.data(), pr.data()));
  CUXB_CHECK(cuxb_mg_sync(mg_));
  last_ms_ = (float)(NowMs() - t0);
  checker_.CheckArray(&expect, 1, 1.0f, 0);
  bool ok = true;
  for (int r = 0; r < w; r++) {   // every rank holds the same combined scalar
    float got = 0.f;
    CUXB_CHECK(cuxb_set_device(cuxb_mg_device(mg_, r)));
    CUXB_CHECK(cuxb_memcpy_d2h(&got, dr[r].p, 4, nullptr));
    ok = checker_.CheckArray(&got, 1, 1.0f, 1 + r) && ok;
  }
  return ok;
}

bool MultiExecutor::RunGemm(int M, int N, int K, float alpha, const float *A, const float *B, float beta, const float *C0,
                            const float *expect) {
  if (!mg_) CUXLOG_ERR("MultiExecutor::RunGemm before InitializeMulti.");
  const int w = world();
  std::vector<DeviceBuf> dA(w), dC(w);
  DeviceBuf dB;
  std::vector<int> rows(w);
  std::vector<long long> first(w);
  std::vector<const float *> pA(w);
  std::vector<float *> pC(w);
  for (int r = 0; r < w; r++) {
    long long lo, hi;
    Shard(M, w, r, 128, &lo, &hi);   // whole 128-row tiles per rank (the fused-broadcast SGEMM's shape class)
    rows[r] = (int)(hi - lo), first[r] = lo;
    const int dev = cuxb_mg_device(mg_, r);
    dA[r].Alloc(dev, (size_t)rows[r] * K * 4), dC[r].Alloc(dev, (size_t)rows[r] * N * 4);
    CUXB_CHECK(cuxb_memcpy_h2d(dA[r].p, A + lo * K, (size_t)rows[r] * K * 4, nullptr));
    CUXB_CHECK(cuxb_memcpy_h2d(dC[r].p, C0 + lo * N, (size_t)rows[r] * N * 4, nullptr));
    pA[r] = (const float *)dA[r].p, pC[r] = (float *)dC[r].p;
  }
  dB.Alloc(cuxb_mg_device(mg_, 0), (size_t)K * N * 4);   // B exists on rank 0 only
  CUXB_CHECK(cuxb_memcpy_h2d(dB.p, B, (size_t)K * N * 4, nullptr));
  const double t0 = NowMs();
  CUXB_CHECK(cuxb_mg_all_sgemm_f32(mg_, rows.data(), N, K, alpha, pA.data(), K, (const float *)dB.p, N, beta, pC.data(), N));
  CUXB_CHECK(cuxb_mg_sync(mg_));
  last_ms_ = (float)(NowMs() - t0);
  std::vector<float> got((size_t)M * N);
  for (int r = 0; r < w; r++) {   // gather the row blocks of C (the master's MPI_Recv loop, alg_matrix_multiply.cpp:64-72)
    CUXB_CHECK(cuxb_set_device(cuxb_mg_device(mg_, r)));
    CUXB_CHECK(cuxb_memcpy_d2h(got.data() + first[r] * N, dC[r].p, (size_t)rows[r] * N * 4, nullptr));
  }
  checker_.CheckArray(const_cast<float *>(expect), M * N, 1.0f / (M * N), 0);
  return checker_.CheckArray(got.data(), M * N, 1.0f / (M * N), 1);
}

bool MultiExecutor::RunScanI32(const int32_t *in, int len, const int32_t *expect) {
  if (!mg_) CUXLOG_ERR("MultiExecutor::RunScanI32 before InitializeMulti.");
  const int w = world();
  std::vector<DeviceBuf> di(w), dout(w);
  std::vector<long long> n(w), first(w);
  std::vector<const int32_t *> pi(w);
  std::vector<int32_t *> po(w);
  for (int r = 0; r < w; r++) {
    long long lo, hi;
    Shard(len, w, r, 1, &lo, &hi);
    n[r] = hi - lo, first[r] = lo;
    const int dev = cuxb_mg_device(mg_, r);
    di[r].Alloc(dev, n[r] * 4), dout[r].Alloc(dev, n[r] * 4);
    CUXB_CHECK(cuxb_memcpy_h2d(di[r].p, in + lo, n[r] * 4, nullptr));
    pi[r] = (const int32_t *)di[r].p, po[r] = (int32_t *)dout[r].p;
  }
  const double t0 = NowMs();
  CUXB_CHECK(cuxb_mg_all_scan_i32(mg_, n.data(), pi.data(), po.data()));
  CUXB_CHECK(cuxb_mg_sync(mg_));
  last_ms_ = (float)(NowMs() - t0);
  std::vector<int32_t> got((size_t)len);
  for (int r = 0; r < w; r++) {
    CUXB_CHECK(cuxb_set_device(cuxb_mg_device(mg_, r)));
    CUXB_CHECK(cuxb_memcpy_d2h(got.data() + first[r], dout[r].p, n[r] * 4, nullptr));
  }
  checker_.CheckArrayExact(expect, len, 0);
  return checker_.CheckArrayExact(got.data(), len, 1);
}

}  // namespace cux
