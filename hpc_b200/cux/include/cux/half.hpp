// half.hpp -- 16-bit storage types for the host side: cux::half (IEEE binary16) and cux::bfloat16.
// Same role as 0-frameworks/cux/include/util/half.h (a host half with float conversion and
// arithmetic through float); conversions round to nearest even like half.h:156-203 and
// pocket-ai/util/type.hpp:134-190,268-283.  Written with the "add the rounding bias, then shift"
// formulation rather than the reference's table of magic constants.
#pragma once
#include <cstdint>
#include <cstring>

namespace cux {

namespace detail {
inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

inline uint16_t float_to_half_bits(float f) {
  const uint32_t x = f2u(f);
  const uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
  const uint32_t mag = x & 0x7fffffffu;
  if (mag >= 0x7f800000u) return (uint16_t)(sign | (mag > 0x7f800000u ? 0x7e00u : 0x7c00u));  // nan / inf
  if (mag >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);  // rounds to >= 65520 -> inf
  if (mag >= 0x38800000u) {                                   // normal half
    const uint32_t lsb = (mag >> 13) & 1u;
    const uint32_t r = mag + 0xfffu + lsb;                    // RNE at bit 13
    return (uint16_t)(sign | ((r - 0x38000000u) >> 13));
  }
  if (mag < 0x33000000u) return sign;                         // below half of the smallest subnormal
  // subnormal half: value = m * 2^-24, m in [1,1023]
  const int e = (int)(mag >> 23);                             // biased fp32 exponent, 102..112
  const uint32_t man = (mag & 0x7fffffu) | 0x800000u;         // 24-bit significand
  const int shift = 126 - e;                                  // 14..24
  const uint32_t q = man >> shift, rem = man & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
  uint32_t m = q;
  if (rem > halfway || (rem == halfway && (q & 1u))) m++;
  return (uint16_t)(sign | m);
}
inline float half_bits_to_float(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  const uint32_t e = (h >> 10) & 0x1fu, m = h & 0x3ffu;
  if (e == 0x1f) return u2f(sign | 0x7f800000u | (m << 13));
  if (e != 0) return u2f(sign | ((e + 112u) << 23) | (m << 13));
  if (m == 0) return u2f(sign);
  const float v = (float)m * (1.0f / 16777216.0f);  // m * 2^-24, exact
  return sign ? -v : v;
}
inline uint16_t float_to_bf16_bits(float f) {
  uint32_t x = f2u(f);
  if ((x & 0x7f800000u) == 0x7f800000u) return (uint16_t)((x & 0x7fffffu) ? 0x7fffu : (x >> 16));
  x += 0x7fffu + ((x >> 16) & 1u);
  return (uint16_t)(x >> 16);
}
inline float bf16_bits_to_float(uint16_t b) { return u2f((uint32_t)b << 16); }
}  // namespace detail

class half {
public:
  half() : bits_(0) {}
  half(float f) : bits_(detail::float_to_half_bits(f)) {}
  half(double d) : bits_(detail::float_to_half_bits((float)d)) {}
  half(int i) : bits_(detail::float_to_half_bits((float)i)) {}
  operator float() const { return detail::half_bits_to_float(bits_); }
  uint16_t bits() const { return bits_; }
  static half from_bits(uint16_t b) { half h; h.bits_ = b; return h; }
  half &operator+=(const half &o) { return *this = half((float)*this + (float)o); }
  half &operator-=(const half &o) { return *this = half((float)*this - (float)o); }
  half &operator*=(const half &o) { return *this = half((float)*this * (float)o); }
  half &operator/=(const half &o) { return *this = half((float)*this / (float)o); }
private:
  uint16_t bits_;
};
inline half operator+(half a, half b) { return half((float)a + (float)b); }
inline half operator-(half a, half b) { return half((float)a - (float)b); }
inline half operator*(half a, half b) { return half((float)a * (float)b); }
inline half operator/(half a, half b) { return half((float)a / (float)b); }

class bfloat16 {
public:
  bfloat16() : bits_(0) {}
  bfloat16(float f) : bits_(detail::float_to_bf16_bits(f)) {}
  bfloat16(int i) : bits_(detail::float_to_bf16_bits((float)i)) {}
  operator float() const { return detail::bf16_bits_to_float(bits_); }
  uint16_t bits() const { return bits_; }
private:
  uint16_t bits_;
};

static_assert(sizeof(half) == 2 && sizeof(bfloat16) == 2, "16-bit storage");

}  // namespace cux
