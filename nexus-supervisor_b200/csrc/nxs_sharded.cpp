// nxs_sharded.cpp — one host batch classified on several GPUs (BASELINE config 3).
//
// The unit of dependence of the path is one job = one checkpoint row: superviseAction reads and
// writes only the row of its (algorithm, id) (services/supervisor.go:264,:301; primary key in
// test-resources/checkpoints.cql:22), so jobs never interact.  A batch is therefore partitioned on
// the host by shard = splitmix64(job_key) % G without ever splitting a job, every shard is
// classified by its own ctx (its own GPU when there are G of them, round-robin otherwise) through
// the ordinary host-pointer entry point, and the results are put back in stream order on the host —
// verdicts at the position of their record, job rows in the order in which the jobs end in the
// stream.  No collective, no peer access: G independent H2D / kernel / D2H pipelines.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <type_traits>
#include <vector>

#include "nxs_host.h"

namespace {

struct ShardBuf {
  nxs_ctx *ctx = nullptr;
  int device = 0;
  uint64_t cap = 0; // records the pinned buffers hold
  uint64_t *key = nullptr;
  uint16_t *reason = nullptr;
  uint8_t *kind = nullptr, *flags = nullptr, *stage = nullptr;
  uint32_t *verdict = nullptr;
  nxs_job_row *rows = nullptr;
  uint64_t n = 0, n_jobs = 0;
  int rc = NXS_OK;
  bool pinned = true; // false only under the CPU test hook (no CUDA runtime to page-lock with)
};

// How a shard gets classified: nxs_classify on the shard's ctx in the product; the CPU tests of the
// partition / gather logic plug in their own function (nxs_internal_sharded_run below).
typedef int (*shard_classify_fn)(void *user, uint32_t shard, const nxs_batch *in, const nxs_out *out);

void free_one(const ShardBuf &b, void *p) {
  if (b.pinned) nxs_free_host(p);
  else free(p);
}

void free_bufs(ShardBuf &b) {
  free_one(b, b.key);
  free_one(b, b.reason);
  free_one(b, b.kind);
  free_one(b, b.flags);
  free_one(b, b.stage);
  free_one(b, b.verdict);
  free_one(b, b.rows);
  b.key = nullptr, b.reason = nullptr, b.kind = b.flags = b.stage = nullptr;
  b.verdict = nullptr, b.rows = nullptr, b.cap = 0;
}

int grow_bufs(ShardBuf &b, uint64_t want) {
  if (want <= b.cap) return NXS_OK;
  free_bufs(b);
  const uint64_t cap = (want + want / 16 + 4095) & ~4095ull;
  void *p[7] = {nullptr};
  const size_t bytes[7] = {cap * 8, cap * 2, cap, cap, cap, cap * 4, cap * sizeof(nxs_job_row)};
  for (int i = 0; i < 7; ++i) {
    if (b.pinned) {
      if (nxs_alloc_host(bytes[i], &p[i]) != NXS_OK) p[i] = nullptr;
    } else {
      p[i] = aligned_alloc(64, (bytes[i] + 63) & ~size_t(63));
    }
    if (!p[i]) {
      for (int j = 0; j < i; ++j) free_one(b, p[j]);
      return NXS_ERR_NOMEM;
    }
  }
  b.key = static_cast<uint64_t *>(p[0]), b.reason = static_cast<uint16_t *>(p[1]);
  b.kind = static_cast<uint8_t *>(p[2]), b.flags = static_cast<uint8_t *>(p[3]), b.stage = static_cast<uint8_t *>(p[4]);
  b.verdict = static_cast<uint32_t *>(p[5]);
  b.rows = static_cast<nxs_job_row *>(p[6]);
  b.cap = cap;
  return NXS_OK;
}

// nxs_shard_of(key, G) = splitmix64(key) % G for every record of a pass, without a division: G a power
// of two is a mask; otherwise the 64-bit hash is folded as (hi * (2^32 mod G) + lo) mod G with the exact
// 32-bit remainder-by-multiplication (a mod d = ((M * a mod 2^64) * d) >> 64 with M = floor((2^64 - 1) / d) + 1,
// exact for every 32-bit a and d; D. Lemire, O. Kaser, N. Kurz, "Faster remainder by direct computation", 2019).
struct ShardMap {
  uint64_t M = 0;
  uint32_t d = 1, c = 0, mask = 0;
  bool pow2 = true;
  explicit ShardMap(uint32_t G) : d(G ? G : 1) {
    pow2 = (d & (d - 1)) == 0;
    mask = d - 1;
    M = ~0ull / d + 1;
    c = (uint32_t)((1ull << 32) % d);
  }
  inline uint32_t mod32(uint32_t a) const { return (uint32_t)(((unsigned __int128)(M * a) * d) >> 64); }
  inline uint32_t operator()(uint64_t key) const {
    const uint64_t h = nxs::splitmix64(key);
    if (pow2) return (uint32_t)h & mask;
    return mod32(mod32((uint32_t)(h >> 32)) * c + mod32((uint32_t)h));
  }
};

} // namespace

struct nxs_sharded {
  uint32_t n_shards = 0;
  uint64_t max_records = 0;
  std::vector<ShardBuf> shard;
  unsigned host_threads = 1;
  char err[256] = {0};
};

extern "C" {

int nxs_sharded_create(uint32_t n_shards, const int *devices, uint32_t n_devices, uint64_t max_records,
                       uint16_t finished_mask, nxs_sharded **out) {
  if (!out) return NXS_ERR_ARG;
  *out = nullptr;
  if (n_shards == 0 || n_shards > 256 || !devices || n_devices == 0) return NXS_ERR_ARG;
  if (max_records == 0 || max_records >= 0xFFFFFFFFull) return NXS_ERR_CAPACITY;
  nxs_sharded *h = new (std::nothrow) nxs_sharded();
  if (!h) return NXS_ERR_NOMEM;
  h->n_shards = n_shards;
  h->max_records = max_records;
  try {
    h->shard.resize(n_shards);
  } catch (...) {
    delete h;
    return NXS_ERR_NOMEM;
  }
  const unsigned hw = std::thread::hardware_concurrency();
  h->host_threads = hw == 0 ? 4u : hw > 32u ? 32u : hw;
  for (uint32_t s = 0; s < n_shards; ++s) {
    h->shard[s].device = devices[s % n_devices];
    // Any shard may receive (almost) the whole batch: the hash decides, not the size.
    const int rc = nxs_create(h->shard[s].device, max_records, finished_mask, &h->shard[s].ctx);
    if (rc != NXS_OK) {
      nxs_sharded_destroy(h);
      return rc;
    }
  }
  *out = h;
  return NXS_OK;
}

void nxs_sharded_destroy(nxs_sharded *h) {
  if (!h) return;
  for (ShardBuf &b : h->shard) {
    nxs_destroy(b.ctx);
    free_bufs(b);
  }
  delete h;
}

const char *nxs_sharded_last_error(const nxs_sharded *h) { return h ? h->err : ""; }

static int classify_sharded_impl(nxs_sharded *h, const nxs_batch *in, const nxs_out *out, shard_classify_fn classify,
                                 void *user);

static int classify_on_ctx(void *user, uint32_t shard, const nxs_batch *in, const nxs_out *out) {
  return nxs_classify(static_cast<nxs_sharded *>(user)->shard[shard].ctx, in, out);
}

int nxs_classify_sharded(nxs_sharded *h, const nxs_batch *in, const nxs_out *out) {
  try {
    return classify_sharded_impl(h, in, out, classify_on_ctx, h);
  } catch (...) { // (std::bad_alloc from the bookkeeping vectors: nothing is thrown across the C ABI)
    if (h) snprintf(h->err, sizeof h->err, "out of host memory");
    return NXS_ERR_NOMEM;
  }
}

// Test hook (not in the public header): the division-free shard map of the host passes, to be compared
// with nxs_shard_of.
void nxs_internal_shard_map(const uint64_t *keys, uint64_t n, uint32_t n_shards, uint32_t *out) {
  const ShardMap map(n_shards);
  for (uint64_t i = 0; i < n; ++i) out[i] = map(keys[i]);
}

// Test hook (not in the public header): the partition and gather of nxs_classify_sharded around a
// caller-supplied per-shard classifier, without any GPU — `pytest -m "not gpu"` runs the product's
// host logic with the oracle standing in for the kernel.
int nxs_internal_sharded_run(uint32_t n_shards, const nxs_batch *in, const nxs_out *out, shard_classify_fn classify,
                             void *user) {
  if (n_shards == 0 || n_shards > 256 || !classify) return NXS_ERR_ARG;
  nxs_sharded h;
  h.n_shards = n_shards;
  h.max_records = 0xFFFFFFFEull;
  h.shard.resize(n_shards);
  for (ShardBuf &b : h.shard) b.pinned = false;
  h.host_threads = 3; // more chunks than one, fewer than the shards of some tests
  if (const char *e = getenv("NXS_HOOK_THREADS")) h.host_threads = (unsigned)std::max(1, atoi(e)); // (host-pass timing)
  int rc;
  try {
    rc = classify_sharded_impl(&h, in, out, classify, user);
  } catch (...) {
    rc = NXS_ERR_NOMEM;
  }
  for (ShardBuf &b : h.shard) free_bufs(b);
  return rc;
}

static int classify_sharded_impl(nxs_sharded *h, const nxs_batch *in, const nxs_out *out, shard_classify_fn classify,
                                 void *user) {
  if (!h || !in || !out || !out->n_jobs) return NXS_ERR_ARG;
  const uint64_t n = in->n;
  if (n > h->max_records) return NXS_ERR_CAPACITY;
  if (n == 0) {
    *out->n_jobs = 0;
    return NXS_OK;
  }
  if (!in->kind || !in->reason_id || !in->flags || !in->cur_stage || !in->job_key || !out->job_rows) {
    snprintf(h->err, sizeof h->err, "null column pointer");
    return NXS_ERR_ARG;
  }
  const uint32_t G = h->n_shards;
  const uint64_t *key = in->job_key;
  const bool verbose = getenv("NXS_VERBOSE") != nullptr;
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  double t_mark = now();
  auto lap = [&](const char *what) {
    const double t = now();
    if (verbose) fprintf(stderr, "nxs sharded: %-28s %8.1f ms\n", what, (t - t_mark) * 1e3);
    t_mark = t;
  };

  // ---- host chunks, cut on job boundaries, for the parallel partition and gather
  const uint64_t chunk_min = h->shard[0].pinned ? (1u << 16) : 64u; // (tiny chunks only under the CPU test hook)
  const unsigned T = (unsigned)std::min<uint64_t>(h->host_threads, (n + chunk_min - 1) / chunk_min);
  std::vector<uint64_t> cut(T + 1);
  cut[0] = 0;
  cut[T] = n;
  for (unsigned t = 1; t < T; ++t) {
    uint64_t i = std::max(cut[t - 1], n * t / T);
    while (i < n && i > 0 && key[i] == key[i - 1]) ++i;
    cut[t] = i;
  }
  // per chunk and shard: records and jobs
  std::vector<uint64_t> rec_cnt((size_t)T * G, 0), job_cnt((size_t)T * G, 0);
  auto parallel = [&](auto &&fn) { nxs::run_parallel(T, fn); };
  // The three host passes below visit RECORDS, not jobs: a loop over the few records of a job ends on a
  // branch nobody can predict, and at a few records per job that branch was most of the time.  The shard
  // is recomputed for every record (a hash and a mask or three multiplications), cursors advance by 0 or 1.
  const ShardMap shard_of(G);
  parallel([&](unsigned t) {
    uint64_t rc[256] = {0}, jc[256] = {0}; // thread-local: neighbours' counters would share cache lines
    // (prev differs from the chunk's first key: the first record begins a job; an empty chunk reads nothing)
    uint64_t prev = cut[t] < cut[t + 1] ? ~key[cut[t]] : 0;
    for (uint64_t i = cut[t]; i < cut[t + 1]; ++i) {
      const uint64_t k = key[i];
      const uint32_t s = shard_of(k);
      rc[s] += 1;
      jc[s] += k != prev;
      prev = k;
    }
    for (uint32_t s = 0; s < G; ++s) rec_cnt[(size_t)t * G + s] = rc[s], job_cnt[(size_t)t * G + s] = jc[s];
  });
  lap("count (records, jobs / shard)");
  // exclusive prefix over the chunks, per shard
  std::vector<uint64_t> rec_off((size_t)T * G), job_off((size_t)T * G), chunk_job0(T + 1, 0);
  for (uint32_t s = 0; s < G; ++s) {
    uint64_t r = 0, j = 0;
    for (unsigned t = 0; t < T; ++t) {
      rec_off[(size_t)t * G + s] = r;
      job_off[(size_t)t * G + s] = j;
      r += rec_cnt[(size_t)t * G + s];
      j += job_cnt[(size_t)t * G + s];
    }
    h->shard[s].n = r;
    h->shard[s].n_jobs = j;
    const int rc = grow_bufs(h->shard[s], r);
    if (rc != NXS_OK) {
      snprintf(h->err, sizeof h->err, "cannot allocate pinned staging for shard %u (%llu records)", s, (unsigned long long)r);
      return rc;
    }
  }
  for (unsigned t = 0; t < T; ++t) {
    uint64_t j = 0;
    for (uint32_t s = 0; s < G; ++s) j += job_cnt[(size_t)t * G + s];
    chunk_job0[t + 1] = chunk_job0[t] + j;
  }
  const uint64_t total_jobs = chunk_job0[T];

  lap("prefix + staging buffers");
  // ---- partition: every shard's records in stream order
  struct Cols { // a shard's staging columns (local copies: the byte stores below may alias anything else)
    uint64_t *key;
    uint16_t *reason;
    uint8_t *kind, *flags, *stage;
  };
  parallel([&](unsigned t) {
    uint64_t w[256];
    Cols d[256];
    for (uint32_t s = 0; s < G; ++s) {
      const ShardBuf &b = h->shard[s];
      w[s] = rec_off[(size_t)t * G + s];
      d[s] = Cols{b.key, b.reason, b.kind, b.flags, b.stage};
    }
    const uint16_t *reason = in->reason_id;
    const uint8_t *kind = in->kind, *flags = in->flags, *stage = in->cur_stage;
    for (uint64_t i = cut[t]; i < cut[t + 1]; ++i) {
      const uint64_t k = key[i];
      const uint32_t s = shard_of(k);
      const uint64_t pos = w[s]++;
      const Cols &c = d[s];
      c.key[pos] = k;
      c.reason[pos] = reason[i];
      c.kind[pos] = kind[i];
      c.flags[pos] = flags[i];
      c.stage[pos] = stage[i];
    }
  });

  lap("partition (scatter)");
  // ---- classify: one thread per shard, each on its own ctx (and GPU)
  {
    auto run = [&](uint32_t s) {
      ShardBuf &b = h->shard[s];
      b.rc = NXS_OK;
      if (b.n == 0) return;
      nxs_batch sin{b.kind, b.reason, b.flags, b.stage, b.key, b.n};
      uint64_t jobs = 0;
      nxs_out sout{out->verdict ? b.verdict : nullptr, b.rows, &jobs}; // (rows only when the caller wants no per-record verdicts)
      b.rc = classify(user, s, &sin, &sout);
      if (b.rc != NXS_OK) return;
      if (jobs != b.n_jobs) b.rc = NXS_ERR_CUDA;
    };
    nxs::run_parallel(G, run);
  }
  lap("classify (all shards)");
  for (uint32_t s = 0; s < G; ++s)
    if (h->shard[s].rc != NXS_OK) {
      snprintf(h->err, sizeof h->err, "shard %u on device %d: %s: %s", s, h->shard[s].device, nxs_strerror(h->shard[s].rc),
               h->shard[s].ctx ? nxs_last_error(h->shard[s].ctx) : "");
      return h->shard[s].rc;
    }

  // ---- gather: the stream is walked once more.  A shard's records and job rows are in stream order among
  // themselves, so every record finds its verdict, and the last record of a job the job's row, at the
  // shard's cursors; a winner index is shard-local and moves with the job's first record.  The row is
  // written for every record of a job and the output cursor advances only behind the last one (a chunk
  // ends on a job boundary, so every such early write is overwritten from inside the chunk).
  auto gather = [&](unsigned t, auto want_verdict) {
    constexpr bool kVerdict = decltype(want_verdict)::value;
    uint64_t rpos[256], r[256]; // this chunk's cursors in every shard's buffers: records and job rows
    const uint32_t *sv[256];
    const nxs_job_row *sr[256];
    for (uint32_t s = 0; s < G; ++s) {
      rpos[s] = rec_off[(size_t)t * G + s];
      r[s] = job_off[(size_t)t * G + s];
      sv[s] = h->shard[s].verdict;
      sr[s] = h->shard[s].rows;
    }
    uint32_t *v = out->verdict;
    nxs_job_row *dst = out->job_rows + chunk_job0[t];
    const uint64_t lo = cut[t], hi = cut[t + 1];
    if (lo == hi) return;
    uint64_t prev = ~key[lo], first_i = lo, p0 = 0;
    auto one = [&](uint64_t i, uint64_t k, bool tail) {
      const uint32_t s = shard_of(k);
      const uint64_t p = rpos[s]++;
      if constexpr (kVerdict) v[i] = sv[s][p];
      const bool head = k != prev;
      first_i = head ? i : first_i; // where the job begins: in the stream, and in its shard
      p0 = head ? p : p0;
      prev = k;
      nxs_job_row row = sr[s][r[s]];
      r[s] += tail;
      const uint32_t moved = (uint32_t)(first_i + (row.winner_idx - p0)); // shard -> stream index
      row.winner_idx = row.winner_idx != NXS_NO_WINNER ? moved : row.winner_idx;
      *dst = row;
      dst += tail;
    };
    for (uint64_t i = lo; i + 1 < hi; ++i) one(i, key[i], key[i + 1] != key[i]);
    one(hi - 1, key[hi - 1], true);
  };
  if (out->verdict) parallel([&](unsigned t) { gather(t, std::true_type{}); });
  else parallel([&](unsigned t) { gather(t, std::false_type{}); });
  lap("gather (verdicts, job rows)");
  *out->n_jobs = total_jobs;
  return NXS_OK;
}

} // extern "C"
