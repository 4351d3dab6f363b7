// data_filler.hpp -- DataFiller with the interface of 0-frameworks/cux/include/util/data_filler.h.
// Same distributions (`rand() % mod + min`, optional decimals, data_filler.h:41-60); the reference
// reseeds with time(0) on every call, which makes runs irreproducible and arrays filled within one
// second identical.  Here the seed is a fixed constant advanced per call.
#pragma once
#include <cstdint>
#include <cstring>

namespace cux {

class DataFiller {
public:
  static uint64_t &Seed() { static uint64_t s = 20240607ull; return s; }
  static uint32_t Next(uint64_t &state) {  // 64-bit LCG (Knuth MMIX constants), top bits out
    state = state * 6364136223846793005ull + 1442695040888963407ull;
    return (uint32_t)(state >> 33);
  }
  template <typename DType> static void ZeroFill(int len, DType *arr) { std::memset((void *)arr, 0, sizeof(DType) * (size_t)len); }
  template <typename DType> static void ConstantFill(float value, int len, DType *arr) {
    if (value == 0) { ZeroFill(len, arr); return; }
    for (int i = 0; i < len; i++) arr[i] = value;
  }
  template <typename DType> static void StepFill(int step, int len, DType *arr) {
    for (int i = 0; i < len; i++) arr[i] = i % step;
  }
  // decimal_pose: 0 -> integers in [min, max]; d > 0 -> [min, max) with d decimal digits
  template <typename DType> static void RandomFill(int min, int max, int decimal_pose, int len, DType *arr) {
    uint64_t st = (Seed() += 0x9e3779b97f4a7c15ull);
    if (decimal_pose > 0) max--;
    const int mod = max - min + 1;
    for (int i = 0; i < len; i++) arr[i] = (int)(Next(st) % (uint32_t)mod) + min;
    if (decimal_pose > 0) {
      int s = 1;
      while (decimal_pose--) s *= 10;
      for (int i = 0; i < len; i++) arr[i] = (DType)((float)arr[i] + (float)(Next(st) % (uint32_t)s) / s);
    }
  }
};

}  // namespace cux
