// nxs_foldsim.cpp — CPU walk-through of the kernel's tile algorithm (test only).
//
// Runs, on the host and one tile after another, the same steps the CUDA kernel
// performs per tile — per-record verdict, 32-record bit masks, tile aggregate,
// look-back over predecessor descriptors, job-row emission by the lanes holding a
// job's last record — using the product's own fold algebra from nxs_device.cuh
// (summ_piece / summ_combine / summ_resolve) and decision table (nxs_build_lut).
// It lets `pytest -m "not gpu"` check the algebra and the cross-tile carries
// against the oracle without a GPU.  The look-back is exercised in its general
// form: each tile pretends that a pseudo-random number of its predecessors have
// only published their aggregate so far.
#include <cstdint>
#include <cstring>
#include <vector>

#include "nxs_device.cuh"
#include "nxs_host.h"

using namespace nxs;

namespace {
struct Masks { uint32_t head, tail, he, term, fail, err; };
struct Desc { uint32_t agg_count, inc_count; Summ agg, inc; };
} // namespace

extern "C" int nxs_foldsim(const uint8_t *kind, const uint16_t *reason, const uint8_t *flags, const uint8_t *stage,
                           const uint64_t *key, uint64_t n, uint16_t finished_mask, uint32_t tile_records,
                           uint64_t lookback_seed, uint32_t *verdict, nxs_job_row *rows, uint64_t *n_jobs) {
  if (tile_records == 0 || tile_records % 32) return -1;
  uint8_t lut[128];
  uint16_t code_of_arm[8];
  nxs_build_lut(lut, code_of_arm);
  uint32_t term_mask = 0;
  for (uint32_t a = 1; a <= 4; ++a)
    if ((finished_mask >> new_stage_of(a)) & 1u) term_mask |= 1u << a;
  const uint32_t TILE = tile_records, CHUNKS = TILE / 32;
  const uint64_t n_tiles = (n + TILE - 1) / TILE;
  std::vector<Desc> desc(n_tiles);
  std::vector<uint16_t> code(TILE);
  std::vector<Masks> masks(CHUNKS);
  std::vector<uint32_t> chunk_off(CHUNKS);
  uint64_t rng = lookback_seed;
  *n_jobs = 0;

  for (uint64_t t = 0; t < n_tiles; ++t) {
    const uint64_t start = t * TILE;
    const uint32_t cnt = (uint32_t)((n - start) < TILE ? (n - start) : TILE);
    // ---- phase 1
    for (uint32_t c = 0; c < CHUNKS; ++c) {
      Masks m{0, 0, 0, 0, 0, 0};
      for (uint32_t lane = 0; lane < 32; ++lane) {
        const uint32_t i = c * 32 + lane;
        if (i >= cnt) continue;
        const uint64_t g = start + i;
        const bool head = g == 0 || key[g] != key[g - 1];
        const bool tail = g + 1 == n || key[g] != key[g + 1];
        const uint32_t fl = flags[g], st = stage[g] & 15u;
        const uint32_t k = kind[g] < 3 ? kind[g] : 3u;
        const uint32_t idx = k * 32u + (reason[g] < 32u ? reason[g] : 0u);
        const uint32_t cd = ((fl & 3u) == 3u) ? code_of_arm[lut[idx]] : 0u;
        const uint32_t a = cd & 7u;
        const bool exists = (fl & NXS_FLAG_CHECKPOINT) != 0;
        const bool finished = (finished_mask >> st) & 1u;
        const uint32_t base = exists ? (finished ? (a | 1u << 14) : (a ? applied_verdict(a) : 0u))
                                     : (a | 1u << 12 | NXS_ERR_NO_CHECKPOINT << 15);
        verdict[g] = a ? (base | (cd >> 9) << 9) : 0u;
        code[i] = (uint16_t)cd;
        const bool he = a != 0 && exists;
        m.head |= (uint32_t)head << lane;
        m.tail |= (uint32_t)tail << lane;
        m.he |= (uint32_t)he << lane;
        m.term |= (uint32_t)(he && ((term_mask >> a) & 1u)) << lane;
        m.fail |= (uint32_t)(he && a != NXS_ACTION_RUNNING) << lane;
        m.err |= (uint32_t)(a != 0 && !exists) << lane;
      }
      masks[c] = m;
    }
    // ---- aggregate
    uint32_t total = 0;
    int last_head = -1;
    for (uint32_t c = 0; c < CHUNKS; ++c) {
      chunk_off[c] = total;
      total += (uint32_t)__builtin_popcount(masks[c].tail);
      if (masks[c].head) last_head = (int)(c * 32 + hi_bit(masks[c].head));
    }
    const uint32_t last = cnt - 1;
    const bool closed = (masks[last >> 5].tail >> (last & 31)) & 1u;
    Summ T{0u, F_ANCHORED};
    if (!closed) {
      const uint32_t from = last_head < 0 ? 0u : (uint32_t)last_head;
      Summ S = summ_identity();
      for (uint32_t c = from >> 5; c < CHUNKS; ++c) {
        uint32_t mk = ~0u;
        if (c == (from >> 5)) mk <<= (from & 31);
        const Masks &w = masks[c];
        S = summ_combine(S, summ_piece(w.he & mk, w.term & mk, w.fail & mk, w.err & mk, code.data(), c * 32,
                                       (uint32_t)start));
      }
      T = last_head < 0 ? S : summ_combine(summ_anchor((finished_mask >> (stage[start + from] & 15u)) & 1u), S);
    }
    desc[t].agg_count = total;
    desc[t].agg = T;
    // ---- look-back
    uint32_t excl = 0;
    Summ carry = summ_identity();
    if (t > 0) {
      bool need = !(masks[0].head & 1u);
      rng = splitmix64(rng);
      uint64_t agg_only = rng % 5; // predecessors seen in "aggregate" state before an inclusive one
      int64_t j = (int64_t)t - 1;
      while (true) {
        const bool inclusive = j == 0 || agg_only == 0;
        excl += inclusive ? desc[j].inc_count : desc[j].agg_count;
        if (need) {
          const Summ sj = inclusive ? desc[j].inc : desc[j].agg;
          carry = summ_combine(sj, carry);
          if (sj.f & F_ANCHORED) need = false;
        }
        if (inclusive) break;
        --agg_only;
        --j;
      }
    }
    desc[t].inc_count = excl + total;
    desc[t].inc = (T.f & F_ANCHORED) ? T : summ_combine(carry, T);
    // ---- phase 2
    for (uint32_t c = 0; c < CHUNKS; ++c) {
      const Masks &m = masks[c];
      for (uint32_t lane = 0; lane < 32; ++lane) {
        if (!((m.tail >> lane) & 1u)) continue;
        const uint32_t le = (2u << lane) - 1u;
        const uint32_t pos = excl + chunk_off[c] + (uint32_t)__builtin_popcount(m.tail & (le >> 1));
        uint32_t cc = c, hm = m.head & le;
        while (hm == 0 && cc > 0) {
          --cc;
          hm = masks[cc].head;
        }
        uint32_t hbit = 0;
        Summ S = carry;
        if (hm) {
          hbit = hi_bit(hm);
          S = summ_anchor((finished_mask >> (stage[start + cc * 32 + hbit] & 15u)) & 1u);
        }
        for (uint32_t k = cc; k <= c; ++k) {
          uint32_t mk = ~0u;
          if (k == cc) mk <<= hbit;
          if (k == c) mk &= le;
          const Masks &w = masks[k];
          S = summ_combine(S, summ_piece(w.he & mk, w.term & mk, w.fail & mk, w.err & mk, code.data(), k * 32,
                                         (uint32_t)start));
        }
        uint32_t v, win;
        summ_resolve(S, v, win);
        rows[pos] = nxs_job_row{key[start + c * 32 + lane], v, win};
      }
    }
    *n_jobs = desc[t].inc_count;
  }
  return 0;
}
