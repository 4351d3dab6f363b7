// device_test.cpp -- the cases of the reference's DataTest suite (0-frameworks/cux/tests/data_test.cpp:
// 104-143: push to the GPU and back for int32 / float / half / int8, element-type conversion, Save /
// Restore) run against this Array4D on a real device, plus the device-side extensions the reference does
// not have (Fill ON_DEVICE, PrecsGpuCvt, Save / Restore ON_DEVICE, pinned host copies).
// Needs a B200; exit code 0 = all passed.   Usage: cux_device_test
#include <cmath>
#include <cstring>

#include "executor.h"

static int g_fail = 0;
#define EXPECT(cond)                                                      \
  do {                                                                    \
    if (!(cond)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); g_fail++; } \
  } while (0)

// benchmark[i] = i * scale + beta, written as DType, compared as BType (data_test.cpp:8-38)
template <typename DType, typename BType>
static void PushAndBack(int n, int c, int h, int w, float scale, float beta) {
  cux::Array4D in(n, c, h, w);
  const int len = in.num_element();
  EXPECT(len == n * c * h * w);
  std::vector<BType> bench(len);
  for (int i = 0; i < len; i++) bench[i] = (BType)(i * scale + beta);
  DType *cpu = in.GetCpuData<DType>();
  for (int i = 0; i < len; i++) cpu[i] = (DType)bench[i];
  in.GetGpuData<DType>(cux::PUSH);                 // host -> device
  for (int i = 0; i < len; i++) cpu[i] = (DType)0;
  cpu = in.GetCpuData<DType>(cux::PUSH);           // device -> host
  for (int i = 0; i < len; i++) EXPECT((BType)cpu[i] == bench[i]);
  // PUSH_IF_EMPTY must not overwrite an existing copy; NO_PUSH never copies
  cpu[0] = (DType)1;
  in.GetGpuData<DType>(cux::PUSH_IF_EMPTY);
  in.GetCpuData<DType>(cux::PUSH_IF_EMPTY);
  EXPECT((BType)cpu[0] == (BType)1);
  cpu = in.GetCpuData<DType>(cux::PUSH);
  EXPECT((BType)cpu[0] == bench[0]);
}

// host conversion (data_test.cpp:40-66) and the same conversion done on the device
template <typename SrcType, typename DstType, typename BType>
static void Convert(cuxb_ctx *ctx, int n, int c, int h, int w, float scale, float beta) {
  cux::Array4D in(n, c, h, w), dev(n, c, h, w);
  const int len = in.num_element();
  std::vector<BType> bench(len);
  for (int i = 0; i < len; i++) bench[i] = (BType)(i * scale + beta);
  SrcType *src = in.GetCpuData<SrcType>(), *src2 = dev.GetCpuData<SrcType>();
  for (int i = 0; i < len; i++) src[i] = src2[i] = (SrcType)bench[i];
  in.PrecsCpuCvt<SrcType, DstType>();
  DstType *dst = in.GetCpuData<DstType>(cux::PUSH_IF_EMPTY);
  for (int i = 0; i < len; i++) EXPECT((BType)dst[i] == bench[i]);
  dev.GetGpuData<SrcType>(cux::PUSH);
  dev.PrecsGpuCvt(ctx, cux::DataType<SrcType>::kFlag, cux::DataType<DstType>::kFlag);
  DstType *dst2 = dev.GetCpuData<DstType>(cux::PUSH);
  for (int i = 0; i < len; i++) EXPECT((BType)dst2[i] == bench[i]);
}

// Save / Restore on the host (data_test.cpp:68-100) and on the device
template <typename DType, typename BType>
static void Backup(int n, int c, int h, int w, float scale, float beta) {
  cux::Array4D in(n, c, h, w);
  const int len = in.num_element(), type = cux::DataType<DType>::kFlag;
  std::vector<BType> bench(len);
  for (int i = 0; i < len; i++) bench[i] = (BType)(i * scale + beta);
  DType *cpu = in.GetCpuData<DType>();
  for (int i = 0; i < len; i++) cpu[i] = (DType)bench[i];
  in.Save(type, cux::OpRunMode::ON_HOST);
  for (int i = 0; i < len; i++) cpu[i] = (DType)0;
  in.Restore(type, cux::OpRunMode::ON_HOST);
  for (int i = 0; i < len; i++) EXPECT((BType)cpu[i] == bench[i]);
  // device snapshot: what gemm's RunOnDevice relies on so that every kernel starts from the same C
  in.GetGpuData<DType>(cux::PUSH);
  in.Save(type, cux::OpRunMode::ON_DEVICE);
  for (int i = 0; i < len; i++) cpu[i] = (DType)0;
  in.GetGpuData<DType>(cux::PUSH);                 // clobber the device copy with zeros
  in.Restore(type, cux::OpRunMode::ON_DEVICE);
  cpu = in.GetCpuData<DType>(cux::PUSH);
  for (int i = 0; i < len; i++) EXPECT((BType)cpu[i] == bench[i]);
}

int main() {
  cux::InitEnvironment();
  cux::Executor ex;
  ex.Initialize(0);
  cuxb_ctx *ctx = ex.assistor()->ctx();
  cux::Array4D::SetDeviceContext(ctx);
  for (int pinned = 0; pinned < 2; pinned++) {
    cux::Array4D::UsePinnedHostMemory(pinned != 0);
    PushAndBack<int32_t, int32_t>(1, 2, 1, 3, 1.2f, -2);     // data_test.cpp:104-115
    PushAndBack<float, float>(2, 1, 1, 2, 3.5f, 5);
    PushAndBack<cux::half, float>(2, 1, 1, 2, 3.5f, 5);
    PushAndBack<int8_t, int8_t>(1, 3, 1, 3, 123, 32);
    PushAndBack<float, float>(3, 5, 64, 129, 0.25f, -7);       // not a multiple of anything
    Convert<float, cux::half, float>(ctx, 1, 2, 1, 3, 1.5f, -2.0f);   // data_test.cpp:118-129
    Convert<cux::half, float, float>(ctx, 1, 2, 1, 2, 3.5f, -3.0f);
    Convert<int32_t, int8_t, float>(ctx, 1, 5, 1, 1, 6, -2);
    Convert<int8_t, int32_t, float>(ctx, 1, 1, 3, 3, 5, -3);
    Backup<float, float>(1, 3, 1, 3, 12.3f, 32);                // data_test.cpp:132-143
    Backup<cux::half, float>(1, 2, 2, 3, 2.5f, 32);
    Backup<cux::half, float>(1, 2, 2, 3, 0.5f, 32);
    Backup<cux::half, float>(1, 2, 2, 3, 1.0f, 32);
  }
  cux::Array4D::UsePinnedHostMemory(false);
  // Fill ON_DEVICE (rejected by the reference, array.cpp:47-49): same distribution as the host filler
  {
    cux::Array4D a(1, 1, 256, 257);
    a.Fill(-2, 5, 0, cux::FLOAT32, cux::OpRunMode::ON_DEVICE);
    const float *f = a.GetCpuData<float>(cux::PUSH);
    bool seen[8] = {false};
    for (int i = 0; i < a.num_element(); i++) {
      EXPECT(f[i] >= -2 && f[i] <= 5 && f[i] == std::floor(f[i]));
      if (f[i] >= -2 && f[i] <= 5) seen[(int)f[i] + 2] = true;
    }
    for (bool s : seen) EXPECT(s);
  }
  ex.Clear();
  cux::CleanUpEnvironment();
  std::printf(g_fail ? "cux_device_test: %d FAILED\n" : "cux_device_test: all passed\n", g_fail);
  return g_fail ? 1 : 0;
}
