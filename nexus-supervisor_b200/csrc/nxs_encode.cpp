// nxs_encode.cpp — batched checkpoint-row encoder: the step right after the classification.
//
// superviseAction mutates three columns of the run's row in nexus.checkpoints
// (test-resources/checkpoints.cql:1-23) and writes it back (services/supervisor.go:297-301, :324-328,
// :349-353, :362-364): lifecycle_stage, algorithm_failure_cause, algorithm_failure_details.  For a
// batch of job rows this file materialises those strings — the stage name, the cause format applied to
// the message text (a template pair, so one of 3 x 6 fixed strings) and the winner event's Message —
// back to back in ONE byte arena with an offsets array, ready to be bound to UPDATE statements.  Host
// code: the only variable-length input, the events' messages, lives in host memory and never goes to
// the GPU, so this is a gather of host bytes (two passes: sizes and an exclusive scan, then the copy,
// both split over threads).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "nxs_host.h"

namespace {

// A fixed string of the vocabulary with zero bytes behind it up to a multiple of 32 (and one block more): the
// encoder copies it in whole 32-byte blocks.
struct Tpl {
  std::string buf;
  size_t len = 0;
  void set(const std::string &x) {
    len = x.size();
    buf = x;
    buf.resize(((len + 31) & ~size_t(31)) + 32, '\0');
  }
  size_t size() const { return len; }
  const char *data() const { return buf.data(); }
};
struct Tables {
  Tpl stage[16];
  Tpl cause[4][8];
};

Tables build_tables(const char *const *stage_names) {
  Tables t;
  for (uint32_t s = 0; s < 16; ++s) t.stage[s].set(stage_names ? (stage_names[s] ? stage_names[s] : "") : nxs_stage_name(s));
  for (uint32_t c = 1; c < 4; ++c)
    for (uint32_t m = 0; m < 8; ++m) {
      const char *fmt = nxs_cause_template(c), *msg = nxs_msg_template(m);
      std::string out(strlen(fmt) + strlen(msg) + 1, '\0');
      out.resize((size_t)snprintf(out.data(), out.size(), fmt, msg)); // :298, :325, :350
      t.cause[c][m].set(out);
    }
  for (Tpl &x : t.cause[0]) x.set("");
  return t;
}

inline uint64_t details_len(const nxs_job_row &r, const uint64_t *msg_offsets, uint64_t n_records) {
  if (!msg_offsets || r.winner_idx == NXS_NO_WINNER || r.winner_idx >= n_records) return 0;
  return msg_offsets[r.winner_idx + 1] - msg_offsets[r.winner_idx];
}

} // namespace

static int encode_impl(const nxs_job_row *rows, uint64_t n_rows, const uint8_t *msg_heap, const uint64_t *msg_offsets,
                       uint64_t n_records, const char *const *stage_names, uint8_t *arena, uint64_t arena_capacity,
                       uint64_t *offsets, uint32_t *update_row, uint8_t *update_flags, uint64_t *n_updates,
                       uint64_t *arena_bytes) {
  if ((!rows && n_rows) || !n_updates || !arena_bytes) return NXS_ERR_ARG;
  if (msg_offsets && !msg_heap) return NXS_ERR_ARG;
  const Tables t = build_tables(stage_names);
  const unsigned hw = std::thread::hardware_concurrency();
  const unsigned T = (unsigned)std::min<uint64_t>(hw == 0 ? 4 : hw > 32 ? 32 : hw, (n_rows + 65535) / 65536 + (n_rows == 0));
  std::vector<uint64_t> cnt(T + 1, 0), bytes(T + 1, 0);
  auto range = [&](unsigned k, uint64_t &lo, uint64_t &hi) { lo = n_rows * k / T, hi = n_rows * (k + 1) / T; };
  auto run = [&](auto &&fn) { nxs::run_parallel(T, fn); };
  // pass 1: how many updates and bytes per slice of the rows
  run([&](unsigned k) {
    uint64_t lo, hi, c = 0, b = 0;
    range(k, lo, hi);
    for (uint64_t r = lo; r < hi; ++r) {
      const uint32_t v = rows[r].verdict;
      if (!NXS_V_UPSERT(v)) continue; // delete-only, skipped and failed rows write nothing (:265-279)
      ++c;
      b += t.stage[NXS_V_NEW_STAGE(v)].size();
      if (NXS_V_CAUSE_TPL(v)) b += t.cause[NXS_V_CAUSE_TPL(v)][NXS_V_MSG_TPL(v)].size() + details_len(rows[r], msg_offsets, n_records);
    }
    cnt[k + 1] = c, bytes[k + 1] = b;
  });
  for (unsigned k = 0; k < T; ++k) cnt[k + 1] += cnt[k], bytes[k + 1] += bytes[k];
  *n_updates = cnt[T];
  *arena_bytes = bytes[T];
  if (!arena && !offsets) return NXS_OK; // size query
  if (!arena || !offsets || !update_row || arena_capacity < bytes[T]) return NXS_ERR_CAPACITY;
  // pass 2: the strings, back to back: update u owns offsets[3u .. 3u+3]
  run([&](unsigned k) {
    uint64_t lo, hi;
    range(k, lo, hi);
    uint64_t u = cnt[k], pos = bytes[k];
    const uint64_t region_end = bytes[k + 1];
    const uint64_t heap_bytes = msg_offsets ? msg_offsets[n_records] : 0;
    auto put = [&](const void *src, uint64_t len) { // bytes of unknown origin: exact copy
      memcpy(arena + pos, src, len);
      pos += len;
    };
    // A template string (Tpl: zero-padded to whole 32-byte blocks) is copied
    // in 32-byte blocks — constant-size moves instead of a library call per field — as long as the blocks
    // stay inside this thread's part of the arena; what a block writes behind the field is overwritten by
    // the next field.
    auto put_tpl = [&](const Tpl &str) {
      const uint64_t len = str.size(), padded = (len + 31) & ~uint64_t(31);
      if (pos + padded <= region_end) {
        const char *src = str.data();
        for (uint64_t o = 0; o < padded; o += 32) memcpy(arena + pos + o, src + o, 32);
      } else {
        memcpy(arena + pos, str.data(), len);
      }
      pos += len;
    };
    for (uint64_t r = lo; r < hi; ++r) {
      const uint32_t v = rows[r].verdict;
      if (!NXS_V_UPSERT(v)) continue;
      const uint32_t c = NXS_V_CAUSE_TPL(v);
      const Tpl &stage = t.stage[NXS_V_NEW_STAGE(v)];
      offsets[3 * u] = pos;
      put_tpl(stage); // :297, :324, :349, :362
      offsets[3 * u + 1] = pos;
      if (c) put_tpl(t.cause[c][NXS_V_MSG_TPL(v)]);
      offsets[3 * u + 2] = pos;
      if (c) { // :299, :326, :351
        const uint64_t len = details_len(rows[r], msg_offsets, n_records);
        if (len) {
          const uint64_t at = msg_offsets[rows[r].winner_idx];
          if (len <= 32 && pos + 32 <= region_end && at + 32 <= heap_bytes) { // a short message: one fixed-size block
            memcpy(arena + pos, msg_heap + at, 32);
            pos += len;
          } else {
            put(msg_heap + at, len);
          }
        }
      }
      update_row[u] = (uint32_t)r;
      if (update_flags) update_flags[u] = (uint8_t)((c ? 1u : 0u) | (NXS_V_DELETE_JOB(v) ? 2u : 0u));
      ++u;
    }
  });
  offsets[3 * cnt[T]] = bytes[T];
  return NXS_OK;
}

extern "C" int nxs_encode_checkpoint_rows(const nxs_job_row *rows, uint64_t n_rows, const uint8_t *msg_heap,
                                          const uint64_t *msg_offsets, uint64_t n_records,
                                          const char *const *stage_names, uint8_t *arena, uint64_t arena_capacity,
                                          uint64_t *offsets, uint32_t *update_row, uint8_t *update_flags,
                                          uint64_t *n_updates, uint64_t *arena_bytes) {
  try {
    return encode_impl(rows, n_rows, msg_heap, msg_offsets, n_records, stage_names, arena, arena_capacity, offsets,
                       update_row, update_flags, n_updates, arena_bytes);
  } catch (...) { // (std::bad_alloc from the string tables: nothing is thrown across the C ABI)
    return NXS_ERR_NOMEM;
  }
}
