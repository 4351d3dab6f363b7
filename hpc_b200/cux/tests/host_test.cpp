// host_test.cpp -- checks of the cux-compatible host layer that need no GPU.  Modelled on what the
// reference's gtest suite pins (0-frameworks/cux/tests/util/half_test.cpp:8-72,
// util/data_filler_test.cpp:52-71, data_test.cpp:118-143) plus the factory / checker / parser logic.
// Exit code 0 = all passed.   Usage: cux_host_test [missing-op]   ("missing-op" must abort).
#include <cmath>
#include <cstring>

#include "executor.h"

static int g_fail = 0;
#define EXPECT(cond)                                                      \
  do {                                                                    \
    if (!(cond)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); g_fail++; } \
  } while (0)

static void TestHalf() {
  using cux::half;
  // half_test.cpp: construction and + - * / within 0.01
  half a(9.125f), b(1.234f);
  EXPECT(std::fabs((float)a - 9.125f) < 1e-6f);          // exactly representable
  EXPECT(std::fabs((float)b - 1.234f) < 0.001f);
  EXPECT(std::fabs((float)(a + b) - 10.359f) < 0.01f);
  EXPECT(std::fabs((float)(a - b) - 7.891f) < 0.01f);
  EXPECT(std::fabs((float)(a * b) - 11.26f) < 0.01f);
  EXPECT(std::fabs((float)(a / b) - 7.3947f) < 0.01f);
  EXPECT((float)half(1234.f) == 1234.f && (float)half(200) == 200.f);
  // bit patterns: RNE, subnormals, overflow, specials
  EXPECT(half(1.0f).bits() == 0x3c00 && half(-2.0f).bits() == 0xc000 && half(65504.f).bits() == 0x7bff);
  EXPECT(half(65520.f).bits() == 0x7c00 && half(65519.99f).bits() == 0x7bff);
  EXPECT(half(5.9604645e-8f).bits() == 0x0001 && half(2.9802322e-8f).bits() == 0x0000);  // 2^-24, 2^-25 (tie -> even)
  EXPECT(half(1.0009766f).bits() == 0x3c01);                                           // 1 + 2^-10
  EXPECT(half(1.00048828125f).bits() == 0x3c00 && half(1.00146484375f).bits() == 0x3c02);  // ties to even
  EXPECT(half(INFINITY).bits() == 0x7c00 && (half(NAN).bits() & 0x7c00) == 0x7c00 && (half(NAN).bits() & 0x3ff));
  for (uint32_t h = 0; h < 0x10000; h++) {  // every finite half survives half -> float -> half
    if (((h >> 10) & 0x1f) == 0x1f) continue;
    const float f = (float)half::from_bits((uint16_t)h);
    if (half(f).bits() != h) { EXPECT(half(f).bits() == h); break; }
  }
  EXPECT(cux::bfloat16(1.0f).bits() == 0x3f80 && cux::bfloat16(1.00390625f).bits() == 0x3f80 &&
         cux::bfloat16(1.01171875f).bits() == 0x3f82);
}

static void TestDataFiller() {
  float buf[4096];
  cux::DataFiller::ZeroFill(4096, buf);
  for (float v : buf) EXPECT(v == 0.f);
  cux::DataFiller::ConstantFill(2.5f, 4096, buf);
  for (float v : buf) EXPECT(v == 2.5f);
  cux::DataFiller::RandomFill(-2, 5, 0, 4096, buf);
  bool seen[8] = {false};
  for (float v : buf) {
    EXPECT(v >= -2 && v <= 5 && v == std::floor(v));
    if (v >= -2 && v <= 5) seen[(int)v + 2] = true;
  }
  for (bool s : seen) EXPECT(s);  // every value of the range occurs
  cux::DataFiller::RandomFill(0, 3, 2, 4096, buf);  // two decimals: [0,3), multiples of 0.01
  for (float v : buf) EXPECT(v >= 0.f && v < 3.f);
  int32_t ibuf[64];
  cux::DataFiller::StepFill(7, 64, ibuf);
  EXPECT(ibuf[0] == 0 && ibuf[6] == 6 && ibuf[7] == 0);
}

static void TestArrayHostSide() {
  cux::Array4D a(1, 2, 3, 4);
  EXPECT(a.num_element() == 24 && a.shape()[cux::HEIGHT] == 3 && a.shape()[cux::WIDTH] == 4);
  EXPECT(a.cpu_data(cux::FLOAT32) == nullptr);
  a.Fill(-2, 5, 0, cux::FLOAT32);
  float *f = a.GetCpuData<float>();
  EXPECT(f == a.cpu_data(cux::FLOAT32));
  a.PrecsCpuCvt<float, cux::half>();
  a.PrecsCpuCvt<float, int32_t>();
  a.PrecsCpuCvt<float, int8_t>();
  a.PrecsCpuCvt<float, cux::bfloat16>();
  for (int i = 0; i < 24; i++) {
    EXPECT((float)a.GetCpuData<cux::half>()[i] == f[i]);
    EXPECT(a.GetCpuData<int32_t>()[i] == (int32_t)f[i] && a.GetCpuData<int8_t>()[i] == (int8_t)f[i]);
    EXPECT((float)a.GetCpuData<cux::bfloat16>()[i] == f[i]);
  }
  a.Save(cux::FLOAT32, cux::ON_HOST);       // data_test.cpp:132-143
  const float first = f[0];
  f[0] = 99.f;
  a.Save(cux::FLOAT32, cux::ON_HOST, true);  // save-if-empty: must NOT overwrite the snapshot
  a.Restore(cux::FLOAT32, cux::ON_HOST);
  EXPECT(f[0] == first);
  f[0] = 77.f;
  a.Save(cux::FLOAT32, cux::ON_HOST, false);
  f[0] = 1.f;
  a.Restore(cux::FLOAT32, cux::ON_HOST);
  EXPECT(f[0] == 77.f);
}

static void TestCheckerAndParser() {
  cux::ResultChecker chk;
  float ref[4] = {1, 2, 3, 4}, same[4] = {1, 2, 3, 4}, off[4] = {1, 2, 3, 4.1f};
  EXPECT(chk.CheckArray(ref, 4, 0.25f, 0));
  EXPECT(chk.CheckArray(same, 4, 0.25f, 1));
  EXPECT(!chk.CheckArray(off, 4, 0.25f, 2));      // 0.1 * 0.25 = 0.025 >= 1e-3
  off[3] = 4.003f;
  EXPECT(chk.CheckArray(off, 4, 0.25f, 3));       // 0.003 * 0.25 < 1e-3
  EXPECT(!chk.all_passed() && chk.verdicts().size() == 4);
  int32_t ir[3] = {1 << 30, -5, 7}, ib[3] = {(1 << 30) + 1, -5, 7};
  cux::ResultChecker ichk;
  EXPECT(ichk.CheckArrayExact(ir, 3, 0) && ichk.CheckArrayExact(ir, 3, 1) && !ichk.CheckArrayExact(ib, 3, 2));
  std::string p = "alpha: 1.0, beta: 3.0";
  EXPECT(std::strtof(cux::StrProcessor::FetchSubStr(p, "alpha:", ",").c_str(), nullptr) == 1.0f);
  EXPECT(std::strtof(cux::StrProcessor::FetchSubStr(p, "beta:", ",").c_str(), nullptr) == 3.0f);
  EXPECT(cux::DivUp(10, 4) == 3 && cux::TypeSize(cux::FLOAT16) == 2 && cux::TypeSize(cux::INT8) == 1);
  EXPECT(cux::DataType<cux::half>::kFlag == cux::FLOAT16 && cux::DataType<uint8_t>::kFlag == cux::INT8);
  EXPECT(cux::KernelTag("dot_f32") != nullptr && cux::KernelTag("nope") == nullptr);
  EXPECT(cux::KernelTagName(cux::KernelTag("sgemm_f32")) != nullptr && cux::KernelTagName(&p) == nullptr);
}

static void TestFactoryAndHeuristics() {
  cux::InitEnvironment();
  cux::InitEnvironment();  // idempotent
  const std::string list = cux::OpFactory::GetInstance().PrintList();
  for (const char *name : {"dot", "nrm2", "gemm", "reduce_sum", "linear_int8", "scan_i32"})
    EXPECT(list.find(std::string(name) + " ") != std::string::npos);
  cux::Device dev;
  dev.prop.multiProcessorCount = 148, dev.prop.maxThreadsPerMultiProcessor = 2048, dev.prop.maxThreadsPerBlock = 1024;
  cux::LaunchConfig lc(&dev);
  cux::Config1D c1 = lc.GetKernelConfig1D(4096000);
  EXPECT(c1.threads_per_block == 1024 && c1.blocks_per_grid == 148 && c1.virtual_thread_count == 4096000);
  cux::Config2D c2 = lc.GetKernelConfig2D(576, 512);
  EXPECT(c2.threads_per_block.x == 256 && c2.threads_per_block.y == 1 && c2.blocks_per_grid.x == 3);
  int ab = -1;
  double occ = -1;
  lc.QueryPotentialOccupancy(nullptr, 256, 0, ab, occ);  // no context / foreign pointer -> zeros
  EXPECT(ab == 0 && occ == 0.0);
}

int main(int argc, char **argv) {
  if (argc > 1 && !std::strcmp(argv[1], "missing-op")) {
    cux::InitEnvironment();
    std::string none;
    cux::OpFactory::GetInstance().CreateOpByType("no_such_op", nullptr, none);  // CUXLOG_ERR aborts
    return 0;
  }
  TestHalf();
  TestDataFiller();
  TestArrayHostSide();
  TestCheckerAndParser();
  TestFactoryAndHeuristics();
  std::printf("cux_host_test: %s (%d failures)\n", g_fail ? "FAILED" : "ok", g_fail);
  return g_fail ? 1 : 0;
}
