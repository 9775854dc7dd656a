// nxs_host.h — host-only pieces shared by the translation units of the library.
#pragma once
#include <cstdint>

#include "nxs_classify.h"

namespace nxs {
// Fills the 128-entry (kind, reason) -> code table the kernel stages in shared
// memory: index = min(kind,3)*32 + (reason < 32 ? reason : 0),
// code = action | cause_tpl << 3 | msg_tpl << 5, 0 for every `default` arm.
void nxs_build_lut(uint8_t lut[128]);
inline uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
} // namespace nxs
