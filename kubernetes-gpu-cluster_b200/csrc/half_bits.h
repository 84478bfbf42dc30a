// IEEE half <-> float bit conversions for HOST code (loader, RoPE cache): round to nearest even, as torch's .to(float16).
#pragma once
#include <stdint.h>
#include <string.h>

namespace kge {

static inline uint16_t f32_to_f16_bits(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  x &= 0x7fffffffu;
  if (x >= 0x7f800000u) return (uint16_t)(sign | 0x7c00u | (x > 0x7f800000u ? 0x200u : 0u));   // inf / NaN
  if (x >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);                                        // rounds to inf
  if (x < 0x33000001u) return (uint16_t)sign;                                                     // rounds to zero
  if (x < 0x38800000u) {                                                                          // subnormal half
    const int shift = 126 - (int)(x >> 23);                 // 14..24
    const uint32_t man = (x & 0x7fffffu) | 0x800000u;
    uint32_t r = man >> shift;
    const uint32_t rem = man & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (r & 1u))) ++r;
    return (uint16_t)(sign | r);
  }
  uint32_t r = ((x - 0x38000000u) >> 13);
  const uint32_t rem = x & 0x1fffu;
  if (rem > 0x1000u || (rem == 0x1000u && (r & 1u))) ++r;
  return (uint16_t)(sign | r);
}

}  // namespace kge
