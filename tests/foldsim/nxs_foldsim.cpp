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
#include <algorithm>
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

// ---------------------------------------------------------------------------
// Walk-through of the kernel's worker pass as it is since the job pass has one lane per job: every
// worker range of `range_records` records is processed the way classify_range / job_round do it —
// record pass: ballot words per 32-record chunk (MaskRow) and the list of job ends; job pass: every
// run of records [ends[k-1], ends[k]) is cut out of the mask rows (funnel shift for runs of up to 32
// records, word by word for longer ones) and summarised with the product's summ_run_flags /
// summ_run_finish; the run that began before the range is left open (`unres`), the one still open at
// its end is the `suffix`.  Ranges are stitched sequentially with summ_combine (the kernel does that
// in the summary warp and through the look-back).
namespace {
struct Row4 { uint32_t he, tm, fl, er; };
inline uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {
  return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
}
inline uint32_t shr_clamp(uint32_t v, uint32_t n) { return n >= 32u ? 0u : v >> n; }
} // namespace

extern "C" int nxs_foldsim_ranges(const uint8_t *kind, const uint16_t *reason, const uint8_t *flags,
                                  const uint8_t *stage, const uint64_t *key, uint64_t n, uint16_t finished_mask,
                                  uint32_t range_records, uint32_t *verdict, nxs_job_row *rows, uint64_t *n_jobs) {
  if (range_records == 0 || range_records % 32) return -1;
  uint8_t lut[128];
  uint16_t code_of_arm[8];
  nxs_build_lut(lut, code_of_arm);
  uint32_t term_mask = 0;
  for (uint32_t a = 1; a <= 4; ++a)
    if ((finished_mask >> new_stage_of(a)) & 1u) term_mask |= 1u << a;
  const uint32_t RANGE = range_records, CPW = RANGE / 32;
  std::vector<Row4> mrow(CPW + 2);
  std::vector<uint16_t> ends(RANGE + 1);
  std::vector<uint8_t> arm(RANGE);
  Summ open = summ_identity(); // the job open at the end of the previous range, from its first record on
  uint64_t n_rows = 0;

  for (uint64_t start = 0; start < n; start += RANGE) {
    const uint32_t cnt_r = (uint32_t)((n - start) < RANGE ? (n - start) : RANGE);
    const bool first_is_head = start == 0 || key[start] != key[start - 1];
    // ---- record pass
    std::fill(mrow.begin(), mrow.end(), Row4{0, 0, 0, 0});
    uint32_t njobs = 0;
    for (uint32_t i = 0; i < cnt_r; ++i) {
      const uint64_t g = start + i;
      const uint32_t fl = flags[g], st = stage[g] & 15u;
      const uint32_t k = kind[g] < 3 ? kind[g] : 3u;
      const uint32_t am = lut[k * 32u + (reason[g] < 32u ? reason[g] : 0u)];
      const uint32_t cd = ((fl & 3u) == 3u) ? code_of_arm[am] : 0u;
      const uint32_t a = cd & 7u;
      const bool exists = (fl & NXS_FLAG_CHECKPOINT) != 0;
      const bool finished = (finished_mask >> st) & 1u;
      const uint32_t base = exists ? (finished ? (a | 1u << 14) : (a ? applied_verdict(a) : 0u))
                                   : (a | 1u << 12 | NXS_ERR_NO_CHECKPOINT << 15);
      verdict[g] = a ? (base | (cd >> 9) << 9) : 0u;
      arm[i] = (uint8_t)am;
      const bool he = a != 0 && exists;
      Row4 &m = mrow[i >> 5];
      const uint32_t bit = 1u << (i & 31);
      if (he) m.he |= bit;
      if (he && ((term_mask >> a) & 1u)) m.tm |= bit;
      if (he && a != NXS_ACTION_RUNNING) m.fl |= bit;
      if (a != 0 && !exists) m.er |= bit;
      if (g + 1 == n || key[g] != key[g + 1]) ends[njobs++] = (uint16_t)(i + 1);
    }
    const bool closed = njobs > 0 && ends[njobs - 1] == cnt_r;
    if (!closed) ends[njobs] = (uint16_t)cnt_r;
    const uint32_t nruns = njobs + (closed ? 0u : 1u);
    // ---- job pass
    Summ suffix{0u, F_ANCHORED};
    for (uint32_t k = 0; k < nruns; ++k) {
      const uint32_t s = k ? ends[k - 1] : 0u, e1 = ends[k], len = e1 - s;
      const bool anchored = k != 0 || first_is_head;
      const bool stop0 = anchored && ((finished_mask >> (stage[start + s] & 15u)) & 1u);
      const uint32_t gbase = (uint32_t)start;
      Summ S;
      if (len <= 32u) {
        const uint32_t w = s >> 5, sh = s & 31u, m = shr_clamp(0xFFFFFFFFu, 32u - len);
        const Row4 lo = mrow[w], hi = mrow[w + 1];
        uint32_t wm;
        bool liw;
        const uint32_t f = summ_run_flags(funnel_r(lo.he, hi.he, sh) & m, funnel_r(lo.tm, hi.tm, sh) & m,
                                          funnel_r(lo.fl, hi.fl, sh), funnel_r(lo.er, hi.er, sh) & m, stop0, wm, liw);
        const uint32_t wpos = wm ? s + hi_bit(wm) : s;
        const uint32_t codex = wm ? (code_of_arm[arm[wpos]] | F_HAS_WIN | F_HAS_APP) : 0u;
        S = Summ{gbase + wpos, summ_run_finish(f, codex, liw)};
      } else {
        S = summ_identity();
        for (uint32_t pos = s; pos < e1;) {
          const uint32_t ww = pos >> 5, ss = pos & 31u;
          const uint32_t take = std::min(32u - ss, e1 - pos), m = shr_clamp(0xFFFFFFFFu, 32u - take);
          const Row4 r = mrow[ww];
          uint32_t wm;
          bool liw;
          const uint32_t f = summ_run_flags((r.he >> ss) & m, (r.tm >> ss) & m, r.fl >> ss, (r.er >> ss) & m,
                                            pos == s && stop0, wm, liw);
          uint32_t codex = 0, wpos = pos;
          if (wm) {
            wpos = pos + hi_bit(wm);
            codex = code_of_arm[arm[wpos]] | F_HAS_WIN | F_HAS_APP;
          }
          S = summ_combine(S, Summ{gbase + wpos, summ_run_finish(f, codex, liw)});
          pos += take;
        }
      }
      if (anchored) S.f |= F_ANCHORED;
      if (k == njobs) {
        suffix = S;
      } else {
        if (!anchored) S = summ_combine(open, S); // the run began in an earlier range
        uint32_t v, win;
        summ_resolve(S, v, win);
        rows[n_rows++] = nxs_job_row{key[start + e1 - 1], v, win};
      }
    }
    open = (suffix.f & F_ANCHORED) ? suffix : summ_combine(open, suffix);
  }
  *n_jobs = n_rows;
  return 0;
}

// ------------------------------------------------------------------------------------------------
// run_parallel (nxs_host.h), the helper every host loop of the library runs through: every share runs
// exactly once whatever the number of shares — more than the helper has thread slots for included, which
// takes the same path as a thread that cannot be started (the share runs on the calling thread).
#include <atomic>
extern "C" int nxs_foldsim_run_parallel(unsigned shares, unsigned long long *sum_of_squares) {
  std::atomic<unsigned> ran{0};
  std::atomic<unsigned long long> sum{0};
  std::vector<std::atomic<unsigned char>> seen(shares);
  for (auto &x : seen) x = 0;
  nxs::run_parallel(shares, [&](unsigned t) {
    ran += 1;
    sum += (unsigned long long)t * t;
    seen[t] += 1;
  });
  for (auto &x : seen)
    if (x != 1) return -1;
  *sum_of_squares = sum;
  return (int)ran.load();
}

// ------------------------------------------------------------------------------------------------
// The per-record decision as the host-pointer path reconstructs it: the one-byte code of a record
// (4 * arm | 32 * gate | 128 * finished, built here from the product's tables exactly as the kernel's
// set-up code does) expanded by verdict_of_code() — the function nxs_create fills its 256-entry table from.
#include "nxs_kernel.cuh"
extern "C" void nxs_foldsim_verdict_via_code(const uint8_t *kind, const uint16_t *reason, const uint8_t *flags,
                                             const uint8_t *stage, uint64_t n, uint32_t finished_mask, uint8_t *code_out,
                                             uint32_t *verdict_out) {
  uint8_t lut[128];
  uint16_t code[8];
  nxs::nxs_build_lut(lut, code);
  uint32_t term_mask = 0;
  for (uint32_t a = 1; a <= 4; ++a)
    if ((finished_mask >> nxs::new_stage_of(a)) & 1u) term_mask |= 1u << a;
  for (uint64_t i = 0; i < n; ++i) {
    const uint32_t k = kind[i] < 3 ? kind[i] : 3, r = reason[i] < 32 ? reason[i] : 32;
    const uint32_t arm = r < 32 ? lut[k * 32 + r] : 0; // (the kernel's staged table has a zero column 32)
    const uint32_t fl = flags[i] & 7u, st = stage[i] & 15u, fin = (finished_mask >> st) & 1u;
    const uint32_t gate = (fl & 3u) != 3u ? 0u : !(fl & NXS_FLAG_CHECKPOINT) ? 1u : fin ? 2u : 3u;
    const uint32_t byte = 4u * arm | 32u * gate | 128u * fin;
    code_out[i] = (uint8_t)byte;
    verdict_out[i] = nxs::verdict_of_code(byte, code, term_mask);
  }
}
