// nxs_host.h — host-only pieces shared by the translation units of the library.
#pragma once
#include <cstdint>
#include <thread>

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

// fn(0) .. fn(T-1), fn(0) on the calling thread and the others on threads of their own, all finished on
// return.  A thread that cannot be started (resource limit) costs parallelism, not the call: its share
// runs on the calling thread, and nothing is thrown — these loops run below a C ABI.
template <class Fn> inline void run_parallel(unsigned T, Fn &&fn) {
  constexpr unsigned kMax = 256;
  std::thread th[kMax];
  unsigned started = 0;
  for (unsigned t = 1; t < T; ++t) {
    bool ok = false;
    if (started < kMax) {
      try {
        th[started] = std::thread([&fn, t] { fn(t); });
        ok = true;
        ++started;
      } catch (...) {
      }
    }
    if (!ok) fn(t);
  }
  if (T > 0) fn(0u);
  for (unsigned i = 0; i < started; ++i) th[i].join();
}
} // namespace nxs
