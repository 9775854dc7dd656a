// nxs_host.h — host-only pieces shared by the translation units of the library.
#pragma once
#include <cstdint>

#include "nxs_classify.h"

namespace nxs {
// Fills the two decision tables the kernel stages in shared memory:
//   lut[min(kind,3)*32 + (reason < 32 ? reason : 0)] = switch arm 1..7 of onEvent, 0 for `default`;
//   code[arm] = action | new_stage << 3 | cause_tpl << 7 | msg_tpl << 9, the low 12 bits of the
//   verdict of an applied record of that arm (code[0] = 0).
void nxs_build_lut(uint8_t lut[128], uint16_t code[8]);
inline uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
} // namespace nxs
