// Host-side conversions between fp32 and the two 16-bit storage formats (round to nearest even).
#pragma once
#include <stdint.h>
#include <string.h>

namespace b2i {

inline uint16_t host_f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(u >> 16);
}
inline float host_bf16_to_f32(uint16_t h) {
  uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// IEEE binary16: overflow -> inf, subnormals kept, NaN -> quiet NaN.
inline uint16_t host_f32_to_f16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  const uint16_t sign = static_cast<uint16_t>((u >> 16) & 0x8000u);
  const uint32_t abs = u & 0x7fffffffu;
  if (abs > 0x7f800000u) return sign | 0x7e00u;                 // NaN
  if (abs >= 0x47800000u) return sign | 0x7c00u;                // >= 65536 (or inf) -> inf; 65520..65535.99 handled by rounding below
  if (abs < 0x33000001u) return sign;                           // below half of the smallest subnormal -> 0 (tie to even = 0)
  const int e = static_cast<int>(abs >> 23) - 127;              // unbiased exponent
  uint32_t mant = (abs & 0x7fffffu) | 0x800000u;                // 24-bit significand
  int shift;                                                    // bits to drop
  uint32_t base;                                                // exponent field << 10
  if (e >= -14) { shift = 13; base = static_cast<uint32_t>(e + 15) << 10; mant &= 0x7fffffu; }
  else { shift = 13 + (-14 - e); base = 0; }                    // subnormal: keep the implicit bit, shift further
  const uint32_t kept = mant >> shift;
  const uint32_t rem = mant & ((1u << shift) - 1);
  const uint32_t half = 1u << (shift - 1);
  uint32_t r = base + kept;                                     // rounding up may carry into the exponent: that is correct
  if (rem > half || (rem == half && (kept & 1u))) ++r;
  return sign | static_cast<uint16_t>(r);
}
inline float host_f16_to_f32(uint16_t h) {
  const uint32_t sign = static_cast<uint32_t>(h & 0x8000u) << 16;
  const uint32_t e = (h >> 10) & 0x1fu;
  uint32_t m = h & 0x3ffu;
  uint32_t u;
  if (e == 0) {
    if (m == 0) { u = sign; }
    else {                                                      // subnormal: normalise
      int sh = 0;
      while (!(m & 0x400u)) { m <<= 1; ++sh; }
      u = sign | (static_cast<uint32_t>(127 - 15 + 1 - sh) << 23) | ((m & 0x3ffu) << 13);
    }
  } else if (e == 31) {
    u = sign | 0x7f800000u | (m << 13);
  } else {
    u = sign | ((e + 127 - 15) << 23) | (m << 13);
  }
  float f;
  memcpy(&f, &u, 4);
  return f;
}

inline uint16_t host_f32_to_16(float f, bool f16) { return f16 ? host_f32_to_f16(f) : host_f32_to_bf16(f); }
inline float host_16_to_f32(uint16_t h, bool f16) { return f16 ? host_f16_to_f32(h) : host_bf16_to_f32(h); }

}  // namespace b2i
