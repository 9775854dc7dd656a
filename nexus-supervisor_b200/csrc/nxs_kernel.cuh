// nxs_kernel.cuh — the classification kernel (sm_100a).
//
// One persistent grid makes a single pass over HBM: it classifies every record
// (onEvent + superviseAction decision, services/supervisor.go:137-374) and folds
// the records of each job into one row.  Each CTA is a warp-specialised pipeline
// over a ring of shared-memory stages, driven by mbarriers only:
//
//   producer warp   claims tile ids from a global counter ONE TILE AHEAD of the ring (the id of the
//                   tile that will use the next stage to come free is already known) and streams the
//                   five input columns of a tile into its stage with 1-D bulk async copies (TMA
//                   engine, SASS UBLKCP)
//   marker warps    as soon as a tile id is claimed — a tile time before the tile's stage is even
//                   free — they read the tile's job keys straight from global memory (they reach
//                   shared memory a little later through L2), compare adjacent keys and leave the
//                   "last record of a job" bit masks and the jobs-per-range counts in the tile's
//                   slot of a small ring of tile descriptions (TileMeta)
//   look-back warps (they take the tiles in turn) add up the ranges' counts, publish the tile's JOB
//                   COUNT, poll predecessor tiles (decoupled look-back on counts only) for the
//                   index of the tile's first job row.  Marking and look-back run a tile ahead of
//                   the ring, so by the time a tile has landed its row positions are known: no
//                   stage ever waits for them
//   worker warps    each warp owns a contiguous range of the tile and makes two passes over it.
//                   Record pass, 32 records at a time, lane per record: table lookups -> packed
//                   verdict -> global; four warp ballots -> one row of bit masks per 32 records, kept
//                   in the stage in place of the columns just consumed, as are the switch arm of
//                   every record and the list of the positions where jobs end.
//                   Job pass, lane per JOB: a lane cuts its job's bits out of the mask rows with
//                   funnel shifts, folds them with a dozen bit operations and stores the 16-byte
//                   job row at its final position — every lane does useful work, whatever the share
//                   of records that end a job
//   summary warp    after the workers: frees the stage, publishes the summary of
//                   the job left open at the tile's end, and writes the few rows
//                   the workers could not close — jobs that began before a
//                   warp's range or before the tile (the latter after a second,
//                   off-the-critical-path look-back on summaries)
//
// Splitting the look-back in two keeps the stage's residency short: the count depends on the keys
// alone, so it is published — and the row positions looked up — while the tile is still on its way,
// and no stage waits for another CTA's classification.
// Scratch: TileDesc[n_tiles] (16 B each) + 3 counters.
#pragma once
#include "nxs_device.cuh"

namespace nxs {

#ifndef NXS_WORKERS
#define NXS_WORKERS 16
#endif
constexpr int kWorkerWarps = NXS_WORKERS;
constexpr int kWorkers = kWorkerWarps * 32;
constexpr int kProducerWarp = kWorkerWarps;
constexpr int kSummaryWarp = kWorkerWarps + 1;
constexpr int kCounterWarp0 = kWorkerWarps + 2; // first of the kCounterWarps look-back warps
#ifndef NXS_COUNTER_WARPS
#define NXS_COUNTER_WARPS 2
#endif
constexpr int kCounterWarps = NXS_COUNTER_WARPS; // look-back warps: tile k of a CTA is looked up by warp k % kCounterWarps
#ifndef NXS_MARKER_WARPS
#define NXS_MARKER_WARPS 8
#endif
constexpr int kMarkerWarps = NXS_MARKER_WARPS; // each marks the job boundaries of its share of the worker ranges of every tile
constexpr int kMarkerWarp0 = kCounterWarp0 + kCounterWarps;
constexpr int kThreads = kWorkers + 64 + 32 * (kCounterWarps + kMarkerWarps);
constexpr int kPendingSlots = 4;                // tiles whose carried-in summary is still being looked up
constexpr uint64_t kWatchdogNs = 8000000000ull; // a stuck wait traps instead of hanging the GPU

// The three tile sizes built into the library (NXS_TILE selects one at run time); each must
// be a multiple of 32 * NXS_WORKERS.
#ifndef NXS_TILE_S
#define NXS_TILE_S 2048
#endif
#ifndef NXS_TILE_M
#define NXS_TILE_M 4096
#endif
#ifndef NXS_TILE_L
#define NXS_TILE_L 8192
#endif
#ifndef NXS_TILE_DEFAULT
#define NXS_TILE_DEFAULT NXS_TILE_L
#endif
// Ring depth: as many stages as fit in the shared-memory budget of a CTA, 2..NXS_MAX_STAGES.
#ifndef NXS_MAX_STAGES
#define NXS_MAX_STAGES 8
#endif
#ifndef NXS_SMEM_BUDGET_KB
#define NXS_SMEM_BUDGET_KB 222 /* one CTA per SM */
#endif
constexpr int kStageBudgetBytes = NXS_SMEM_BUDGET_KB * 1024;
template <int TILE> struct StagesFor {
  static constexpr int per_stage = TILE * 13 + 256 + 40 * NXS_WORKERS; // ~ sizeof(StageBuf<TILE>)
  static constexpr int per_meta = TILE / 4 + 8 * NXS_WORKERS + 64;     // ~ sizeof(TileMeta<TILE>); one slot more than stages
  static constexpr int fit = (kStageBudgetBytes - per_meta) / (per_stage + per_meta);
  static constexpr int value = fit < 2 ? 2 : fit > NXS_MAX_STAGES ? NXS_MAX_STAGES : fit;
  static constexpr int metas = value + 1; // the tile claimed ahead of the ring has its description too
};

// What a worker warp leaves behind for its range of the tile.
struct __align__(8) RangeInfo {
  uint64_t suffix;    // summary of the job still open at the end of the range (restricted to the range)
  uint64_t unres;     // summary of the range's first job when it began before the range
  uint64_t unres_key; // ... and its job key
  uint32_t flags;
  uint32_t has_unres; // set by the lane that left `unres`
};
enum : uint32_t { RANGE_CLOSED = 1u, RANGE_FIRST_IS_HEAD = 4u };

template <int TILE> struct __align__(128) StageBuf {
  // keyx[1] = key of the record before the tile, keyx[2..TILE+1] = the tile's
  // keys (16-byte aligned for the bulk copy), keyx[TILE+2] = key after the tile
  uint64_t keyx[TILE + 4];
  // The four small columns are consumed by the record pass of a worker range and then reused IN PLACE,
  // range by range, for what the job pass reads (a range = RANGE records of one worker warp):
  //   reason[r0 + q]            <- u16: position + 1 (range-local) of the q-th record that ends a job
  //   kind[r0 + i]              <- u8: 4 * switch arm | 32 * gate class | 128 * IsFinished(cur_stage) of record i
  //   flags[r0 + 16 c .. +15]   <- MaskRow of 32-record chunk c of the range (he, tm, fl, er ballots)
  // (every write lands on bytes of records the pass has already read: q <= i, 16 c <= 32 c; stage stays)
  uint16_t reason[TILE];
  uint8_t kind[TILE];
  alignas(16) uint8_t flags[TILE];
  alignas(16) uint8_t stage[TILE];
  RangeInfo range[kWorkerWarps];
};

// What is known about a tile before (and independently of) its records: written by the producer (the
// id), the marker warps and the look-back warp while the tile is still on its way.  A ring of
// STAGES + 1 slots: the tile claimed ahead of the ring has one too.
template <int TILE> struct __align__(16) TileMeta {
  uint2 tailinfo[TILE / 32]; // .x bit b: record 32*chunk + b is the last of its job; .y: shared-memory address of the
                             //   list entry (see StageBuf::reason) of the first job that ends in the chunk
  uint32_t range_cnt[kWorkerWarps]; // jobs ending in each worker range                (marker warps)
  uint32_t range_row[kWorkerWarps]; // index of the first job row of each range        (look-back warp)
  uint32_t first_is_head;           // the tile's first record begins a job            (marker warp 0)
  uint32_t total;                   // jobs ending in the tile                         (look-back warp)
  int32_t tile;                     // tile id, -1 = end of work                       (producer)
  uint32_t pad;
};

// Verdict table entry = packed verdict of the record plus the fold bits above it.
constexpr uint32_t VT_HE = 1u << 17, VT_TERM = 1u << 18, VT_FAIL = 1u << 19, VT_ERR = 1u << 20;
constexpr uint32_t kVerdictMask = 0x1FFFFu;

// A tile whose first job began in an earlier tile, parked by the summary warp until
// the summary look-back has delivered what came before.
struct PendingTile {
  uint64_t unres[kWorkerWarps]; // per range: summary of the rows still to be closed
  uint64_t key[kWorkerWarps];
  uint32_t rowpos[kWorkerWarps];
  uint32_t mask; // ranges with a row to close
  uint32_t tile;
  uint64_t local_summ; // open-job summary of the tile alone
  uint32_t need_inc;   // that summary is not anchored: the inclusive one needs the carry
  uint32_t pad;
};

template <int STAGES> struct Ctl {
  // per stage of the ring
  uint64_t full[STAGES];   // the tile's keys landed        (producer -> workers, job pass)
  uint64_t cols[STAGES];   // its other columns landed      (producer -> workers, record pass)
  uint64_t p1done[STAGES]; // rows written                  (workers -> summary warp, producer)
  uint64_t freeb[STAGES];  // stage reusable                (summary warp -> producer)
  // per slot of the ring of tile descriptions (STAGES + 1 slots: they run a tile ahead of the stages)
  uint64_t claimed[STAGES + 1];    // tile id known           (producer -> everybody)
  uint64_t marked[STAGES + 1];     // job boundaries marked   (marker warps -> look-back warps, workers)
  uint64_t base_ready[STAGES + 1]; // first row index known   (look-back warp -> workers)
  PendingTile pending[kPendingSlots];
  // Tables of the classification pass.  Entries are pre-scaled to byte offsets so that an address is
  // one instruction: vt entry of a record = vt + lut[..] + lut2[..].
  alignas(16) uint32_t vt[64]; // [finished(stage) 0..1][gate class 0..3][arm 0..7] -> verdict | fold bits (both halves equal)
  uint32_t codex[8];           // arm -> action | new_stage << 3 | cause_tpl << 7 | msg_tpl << 9 | F_HAS_WIN | F_HAS_APP (0 for arm 0)
  uint8_t lut[136];            // [min(kind,3)][min(reason,32)] (row stride 33) -> 4 * switch arm (1..7, 0 = `default`)
  alignas(8) uint8_t lut2[2048]; // stage * 8 + (flags & 7) -> 32 * gate class (0 dropped, 1 no checkpoint, 2 finished,
                                 //                            3 apply) + 128 * IsFinished(stage & 15)
};

// Tile descriptor: two self-validating 64-bit words, each published with one
// relaxed store, so no fence is needed on either side.
//   c: job count[0,32) | status[32,34) | generation[34,41)      (look-back warp)
//   s: win_idx[0,32) | Summ.f[32,55) | status[55,57) | generation[57,64)   (summary warp)
struct __align__(16) TileDesc {
  uint64_t c, s;
};
enum : uint32_t { DESC_INVALID = 0, DESC_AGG = 1, DESC_INC = 2 };
constexpr uint32_t kGenMax = 127;
constexpr uint32_t kSummFMask = (1u << F_BITS) - 1u;
__host__ __device__ __forceinline__ uint64_t desc_c(uint32_t count, uint32_t status, uint32_t gen) {
  return (uint64_t)count | (uint64_t)status << 32 | (uint64_t)gen << 34;
}
__host__ __device__ __forceinline__ uint64_t desc_s(Summ sm, uint32_t status, uint32_t gen) {
  return (uint64_t)sm.win_idx | (uint64_t)(sm.f & kSummFMask) << 32 | (uint64_t)status << 55 | (uint64_t)gen << 57;
}
__host__ __device__ __forceinline__ uint32_t desc_c_status(uint64_t c, uint32_t gen) {
  return ((uint32_t)(c >> 34) & 127u) == gen ? (uint32_t)(c >> 32) & 3u : DESC_INVALID;
}
__host__ __device__ __forceinline__ uint32_t desc_s_status(uint64_t s, uint32_t gen) {
  return (uint32_t)(s >> 57) == gen ? (uint32_t)(s >> 55) & 3u : DESC_INVALID;
}
__host__ __device__ __forceinline__ Summ desc_s_summ(uint64_t s) {
  return Summ{(uint32_t)s, (uint32_t)(s >> 32) & kSummFMask};
}

// Entry e = arm | 8 * gate class | 32 * finished of the verdict table: the packed verdict of a record that
// takes switch arm `arm` behind gate `gate` (0 dropped, 1 no checkpoint, 2 finished, 3 applied) plus the four
// fold bits.  Host and device build the table from the same function.
__host__ __device__ __forceinline__ uint32_t vt_entry(uint32_t e, const uint16_t *code, uint32_t term_mask) {
  const uint32_t g = (e >> 3) & 3u, arm = e & 7u;
  const uint32_t c = code[arm], a = c & 7u, msg = c >> 9;
  uint32_t v = 0;
  if (g != 0 && a >= 1 && a <= 4) {
    const uint32_t fold = VT_HE | (((term_mask >> a) & 1u) ? VT_TERM : 0u) | (a != NXS_ACTION_RUNNING ? VT_FAIL : 0u);
    if (g == 1) v = a | msg << 9 | 1u << 12 | NXS_ERR_NO_CHECKPOINT << 15 | VT_ERR; // :265-273
    else if (g == 2) v = a | msg << 9 | 1u << 14 | fold;                             // :276-279
    else v = applied_verdict(a) | msg << 9 | fold;                                    // :283-370
  }
  return v;
}
// The per-record verdict as ONE byte: 4 * arm | 32 * gate | 128 * finished — the byte the record pass keeps per
// record anyway.  The host-pointer path downloads these bytes instead of the 32-bit verdicts when the link is
// the bottleneck and expands them with this function.
__host__ __device__ __forceinline__ uint32_t verdict_of_code(uint32_t byte, const uint16_t *code, uint32_t term_mask) {
  return vt_entry(byte >> 2, code, term_mask) & kVerdictMask;
}

struct ClassifyParams {
  const uint8_t *kind;
  const uint16_t *reason;
  const uint8_t *flags;
  const uint8_t *stage;
  const uint64_t *key;
  uint32_t *verdict;
  uint8_t *verdict_code; // instead of `verdict` (which is then null): one byte per record, see verdict_of_code()
  nxs_job_row *rows;
  uint64_t *n_jobs;
  TileDesc *desc;
  uint32_t *counters; // [0] next tile id, [1] CTAs done claiming, [2] watchdog flag
  uint32_t n;
  uint32_t n_tiles;
  uint32_t record_base; // added to winner_idx
  uint32_t finished_mask;
  uint32_t term_mask; // bit a: new_stage_of(a) is a finished stage
  uint32_t gen;
  uint8_t lut[128]; // (kind, reason) -> arm
  uint16_t code[8]; // arm -> code
#ifdef NXS_TRACE
  unsigned long long *trace; // [0] = number of events, then (time, cta<<40 | event<<32 | tile) pairs
  uint32_t trace_cap;
#endif
};

template <int TILE> constexpr size_t classify_smem_bytes() {
  // dynamic part: the ring of stages and the ring of tile descriptions (Ctl is a static __shared__ object)
  return sizeof(StageBuf<TILE>) * StagesFor<TILE>::value + sizeof(TileMeta<TILE>) * StagesFor<TILE>::metas;
}

#if defined(__CUDACC__)

#ifndef NXS_IDLE_NS
#define NXS_IDLE_NS 200
#endif
#ifndef NXS_CLASSIFY_UNROLL
#define NXS_CLASSIFY_UNROLL 4 /* of the 16 steps of a range; 2 measured 1 % slower, 1 another 1 % */
#endif
constexpr int kClassifyUnroll = NXS_CLASSIFY_UNROLL;

__device__ __forceinline__ uint64_t ld_desc_word(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_desc_word(uint64_t *p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

#ifdef NXS_TRACE
// Debug build only: CTAs 0, 97 and 211 log pipeline events with the global timer.
enum : uint32_t { EV_CLAIM = 1, EV_ISSUED, EV_COUNT_START, EV_COUNT_END, EV_LB_START, EV_LB_END, EV_P1_START,
                  EV_P1_END, EV_SUM_START, EV_SUM_END, EV_LB_POLL, EV_SLB_END, EV_STEP, EV_W_IDLE, EV_W_BUSY, EV_C_MARK, EV_C_LB, EV_C_IDLE };
__device__ __forceinline__ void trace_event(const ClassifyParams &p, uint32_t ev, uint32_t tile) {
  if (!p.trace || !(blockIdx.x == 0 || blockIdx.x == 97 || blockIdx.x == 211)) return;
  const unsigned long long i = atomicAdd(p.trace, 1ull);
  if (i >= p.trace_cap) return;
  p.trace[1 + 2 * i] = global_timer_ns();
  p.trace[2 + 2 * i] = (unsigned long long)blockIdx.x << 40 | (unsigned long long)ev << 32 | tile;
}
#ifdef NXS_TRACE_STATS // only the per-warp idle/busy totals (no per-tile events: they cost ~0.6 us each)
#define NXS_TRACE_EVENT(p, ev, tile) ((void)0)
#define NXS_TRACE_STAT(p, ev, v) trace_event(p, ev, v)
#else
#define NXS_TRACE_EVENT(p, ev, tile) trace_event(p, ev, tile)
#define NXS_TRACE_STAT(p, ev, v) ((void)0)
#endif
#else
#define NXS_TRACE_STAT(p, ev, v) ((void)0)
#define NXS_TRACE_EVENT(p, ev, tile) ((void)0)
#endif

// Naps of the service warps between two polls (0 = none).  The hardware scheduler prefers the
// warps with the higher ids — the service warps — so every poll they issue is taken from a worker.
#ifndef NXS_PRODUCER_NS
#define NXS_PRODUCER_NS 0
#endif
#ifndef NXS_COUNTER_NS
#define NXS_COUNTER_NS 0
#endif
#ifndef NXS_COUNT_LB_NS
#define NXS_COUNT_LB_NS 0
#endif
#ifndef NXS_SUMMARY_NS
#define NXS_SUMMARY_NS 0
#endif
#ifndef NXS_SUMM_LB_NS
#define NXS_SUMM_LB_NS 0
#endif
template <unsigned NS> __device__ __forceinline__ void aux_nap() {
  if (NS > 0) __nanosleep(NS);
}

struct Watchdog {
  uint64_t t0;
  uint32_t spins;
  __device__ __forceinline__ Watchdog() : t0(0), spins(0) {}
  __device__ __forceinline__ void tick(uint32_t *flag) {
    if ((++spins & 63u) == 0) {
      const uint64_t now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kWatchdogNs) {
        atomicExch(flag, 1u);
        __trap();
      }
    }
  }
  __device__ __forceinline__ void reset() { t0 = 0; spins = 0; }
};

__device__ __forceinline__ Summ shfl_summ(Summ s, int src) {
  return Summ{__shfl_sync(0xffffffffu, s.win_idx, src), __shfl_sync(0xffffffffu, s.f, src)};
}

// --------------------------------------------------------------------- producer
// Tile k of the CTA (k = 0, 1, ... in claim order) uses stage k % STAGES of the ring and slot
// k % (STAGES + 1) of the tile descriptions.  Its id is claimed one tile early: when tile k - 1 is
// sent on its way.  Every CTA does the same, so the ids are still claimed in the order in which
// the tiles will be loaded (which is what keeps the look-backs short), but the marker and look-back
// warps learn the id a whole tile time before the stage is free, and have the tile's job count and
// first row index ready when it lands.  A CTA never holds more than one id without a stage.
template <int TILE, int STAGES>
__device__ __forceinline__ void producer_loop(const ClassifyParams &p, StageBuf<TILE> *stages, TileMeta<TILE> *metas,
                                              Ctl<STAGES> &ctl) {
  constexpr uint32_t METAS = STAGES + 1;
  Watchdog wd;
  auto publish = [&](uint32_t k, uint32_t id) { // the id of tile k of the CTA: into its slot, and tell everybody
    metas[k % METAS].tile = id < p.n_tiles ? (int32_t)id : -1;
    mbar_arrive(&ctl.claimed[k % METAS]);
    // the tile's keys on their way into L2 at once: the marker warps read them next, the ring a tile time later
    if (id < p.n_tiles && p.n - id * TILE >= (uint32_t)TILE) {
      bulk_prefetch_l2(p.key + (size_t)id * TILE, 8u * TILE);
#ifndef NXS_NO_PREFETCH_COLS /* (the other columns too: their bulk copies then find them in L2; 0.4 % on the 1e8 launch) */
      bulk_prefetch_l2(p.reason + (size_t)id * TILE, 2u * TILE);
      bulk_prefetch_l2(p.kind + (size_t)id * TILE, TILE);
      bulk_prefetch_l2(p.flags + (size_t)id * TILE, TILE);
      bulk_prefetch_l2(p.stage + (size_t)id * TILE, TILE);
#endif
    }
    if (id < p.n_tiles) NXS_TRACE_EVENT(p, EV_CLAIM, id);
  };
  uint32_t id = atomicAdd(p.counters, 1u);
  publish(0, id);
  for (uint32_t k = 0; id < p.n_tiles; ++k) {
    const uint32_t s = k % STAGES, use = k / STAGES;
    if (use > 0) { // the workers and the summary warp are done with the stage's previous tile (tile k - STAGES)
      while (!mbar_try_wait(&ctl.p1done[s], (use - 1) & 1u, 20000u)) {
        aux_nap<NXS_PRODUCER_NS>();
        wd.tick(p.counters + 2);
      }
      while (!mbar_try_wait(&ctl.freeb[s], (use - 1) & 1u, 20000u)) wd.tick(p.counters + 2);
      wd.reset();
    }
    StageBuf<TILE> &sb = stages[s];
    const uint32_t start = id * TILE;
    if (p.n - start >= (uint32_t)TILE) {
      uint64_t *cbar = &ctl.cols[s];
      mbar_arrive_expect_tx(cbar, 5u * TILE);
      bulk_g2s(sb.reason, p.reason + start, 2u * TILE, cbar);
      bulk_g2s(sb.kind, p.kind + start, TILE, cbar);
      bulk_g2s(sb.flags, p.flags + start, TILE, cbar);
      bulk_g2s(sb.stage, p.stage + start, TILE, cbar);
      // (the keys are needed last: for the rows, in the job pass)
      uint64_t *bar = &ctl.full[s];
      mbar_arrive_expect_tx(bar, 8u * TILE);
      bulk_g2s(&sb.keyx[2], p.key + start, 8u * TILE, bar);
      NXS_TRACE_EVENT(p, EV_ISSUED, id);
    } else { // ragged last tile: the workers copy their ranges themselves
      mbar_arrive(&ctl.full[s]);
      mbar_arrive(&ctl.cols[s]);
    }
    // The next id.  Its slot was last used by tile k - STAGES, which the summary warp has just let go of.
    id = atomicAdd(p.counters, 1u);
    publish(k + 1, id);
  }
  // leave the counters zero for the next launch: the last CTA to stop claiming resets them
  __threadfence();
  if (atomicAdd(p.counters + 1, 1u) == gridDim.x - 1) {
    p.counters[0] = 0;
    p.counters[1] = 0;
  }
}

// ------------------------------------------------------- job boundaries (marker warps)
// As soon as a tile id is claimed the marker warps mark the tile, each its share of the worker ranges:
// per range (`range` = the worker warp that owns it) the "last record of its job" bit mask of every
// 32-record chunk, with the place in the job-end list where the chunk's first entry goes, and the
// number of jobs that end in the range; range 0 also notes whether the tile's first record begins a
// job.  The keys are read straight from global memory — the tile itself is loaded later — so nothing
// here waits for the ring: the tile's job count, which every later tile of the grid waits for, is
// known long before the tile's stage is free.  (The same bytes reach shared memory through L2 a few
// microseconds later; HBM delivers them once.)
template <int TILE>
__device__ __forceinline__ void mark_range(const ClassifyParams &p, TileMeta<TILE> &tm, uint32_t ends_a, uint32_t tile,
                                           uint32_t range, uint32_t lane) {
  constexpr uint32_t RANGE = TILE / kWorkerWarps, CPW = RANGE / 32;
  const uint32_t g0 = tile * TILE + range * RANGE; // first record of the range
  const uint64_t *const key = p.key + g0 + lane;
  uint32_t jobs = 0;
  if (g0 < p.n && p.n - g0 > RANGE) { // the range and the record behind it exist (no sum: g0 + RANGE may wrap at 2^32)
#pragma unroll
    for (uint32_t j = 0; j < CPW; ++j) {
      const bool tail = __ldg(key + j * 32) != __ldg(key + j * 32 + 1);
      const uint32_t T = __ballot_sync(0xffffffffu, tail);
      if (lane == 0) tm.tailinfo[range * CPW + j] = make_uint2(T, ends_a + 2u * jobs);
      jobs += __popc(T);
    }
  } else { // the end of the batch falls in (or right behind) this range
    for (uint32_t j = 0; j < CPW; ++j) {
      const uint32_t g = g0 + j * 32 + lane;
      bool tail = false;
      if (g < p.n) tail = g + 1 == p.n || p.key[g] != p.key[g + 1];
      const uint32_t T = __ballot_sync(0xffffffffu, tail);
      if (lane == 0) tm.tailinfo[range * CPW + j] = make_uint2(T, ends_a + 2u * jobs);
      jobs += __popc(T);
    }
  }
  if (lane == 0) {
    tm.range_cnt[range] = jobs;
    if (range == 0) tm.first_is_head = g0 == 0 ? 1u : (uint32_t)(p.key[g0] != p.key[g0 - 1]);
  }
}

template <int TILE, int STAGES>
__device__ __forceinline__ void marker_loop(const ClassifyParams &p, StageBuf<TILE> *stages, TileMeta<TILE> *metas,
                                            Ctl<STAGES> &ctl, uint32_t which, uint32_t lane) {
  static_assert(kWorkerWarps % kMarkerWarps == 0, "the worker ranges are split evenly among the marker warps");
  constexpr uint32_t METAS = STAGES + 1, RPP = kWorkerWarps / kMarkerWarps, RANGE = TILE / kWorkerWarps;
  Watchdog wd;
  for (uint32_t k = 0;; ++k) {
    const uint32_t m = k % METAS;
    while (!__all_sync(0xffffffffu, mbar_try_wait(&ctl.claimed[m], (k / METAS) & 1u, 2000u))) wd.tick(p.counters + 2);
    wd.reset();
    TileMeta<TILE> &tm = metas[m];
    const int32_t t = tm.tile;
    if (t >= 0) {
      if (which == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_COUNT_START, (uint32_t)t);
      StageBuf<TILE> &sb = stages[k % STAGES]; // (only its address is used: where the job-end lists will live)
#pragma unroll 1
      for (uint32_t r = which * RPP; r < (which + 1) * RPP; ++r)
        mark_range<TILE>(p, tm, smem_u32(sb.reason + r * RANGE), (uint32_t)t, r, lane);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&ctl.marked[m]);
    if (t < 0) break;
  }
}

// ----------------------------------------------------------------- look-back warps
struct CountLookback {
  int pos;       // nearest predecessor tile of the next window
  uint32_t excl; // job rows counted so far
};
#ifndef NXS_COUNT_WINDOWS
#define NXS_COUNT_WINDOWS 1 /* 8192-record tiles: 32 predecessors per poll reach far enough */
#endif
constexpr int kCountWindows = NXS_COUNT_WINDOWS;
// One non-blocking step of the count look-back: kCountWindows x 32 predecessor count
// words in one round trip (independent loads), consumed 32 at a time, nearest first.
// A window is consumed only when every word it needs has been published.
__device__ __forceinline__ bool count_lookback_poll(const ClassifyParams &p, CountLookback &st, uint32_t lane) {
  uint64_t c[kCountWindows];
#pragma unroll
  for (int k = 0; k < kCountWindows; ++k) {
    const int idx = st.pos - (int)lane - 32 * k;
    c[k] = idx >= 0 ? ld_desc_word(&p.desc[idx].c) : 0ull;
  }
#pragma unroll
  for (int k = 0; k < kCountWindows; ++k) {
    const bool valid = st.pos - (int)lane >= 0;
    const uint32_t cs = desc_c_status(c[k], p.gen);
    const uint32_t m_valid = __ballot_sync(0xffffffffu, valid);
    const uint32_t m_ok = __ballot_sync(0xffffffffu, valid && cs != DESC_INVALID);
    const uint32_t m_inc = __ballot_sync(0xffffffffu, valid && cs == DESC_INC);
    const uint32_t need = m_inc ? (2u << lo_bit(m_inc)) - 1u : m_valid;
    if (need & ~m_ok) return false;
    uint32_t contrib = ((need >> lane) & 1u) ? (uint32_t)c[k] : 0u;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
    st.excl += contrib;
    if (m_inc) return true;
    st.pos -= 32;
  }
  return false;
}

// The look-back warps of a CTA take its tiles in turn: warp `which` OWNS the tiles k = which, which +
// kCounterWarps, ... (k = order in which the CTA claimed them), so that several look-backs can be in
// flight per CTA.  When the marker warps are done with an own tile, the warp adds up the ranges' job
// counts, publishes the tile's job count — every later tile of every CTA waits for the counts — and
// then polls predecessor counts for the index of the tile's first job row, which the workers need for
// the job pass.  Each warp is an event loop: pass over the tiles of the CTA in order (publishing on
// the way), poll for the oldest own tile in between.  Every warp watches every use of every `marked`
// barrier in order; a slot cannot be marked twice more while a look-back for it is pending (its next
// use needs the slot's tile classified, which needs this look-back), so no parity wait can mistake a
// phase for the one two before.
template <int TILE, int STAGES>
__device__ __forceinline__ void lookback_loop(const ClassifyParams &p, TileMeta<TILE> *metas, Ctl<STAGES> &ctl,
                                              uint32_t which, uint32_t lane) {
  constexpr uint32_t METAS = STAGES + 1;
  uint32_t km = 0;              // next tile of the CTA whose marking to await (all of them, in order)
  uint32_t kl = which;          // next own tile to look back for
  uint32_t k_end = 0xFFFFFFFFu; // index of the end marker, once seen
  bool lb_active = false;
  CountLookback st{0, 0u};
  Watchdog wd;
  while (true) {
    const uint32_t mm = km % METAS;
    if (k_end == 0xFFFFFFFFu && __all_sync(0xffffffffu, mbar_test_wait(&ctl.marked[mm], (km / METAS) & 1u))) {
      TileMeta<TILE> &tm = metas[mm];
      const int32_t t = tm.tile;
      if (t < 0) {
        k_end = km;
      } else {
        if (km % kCounterWarps == which) { // own tile: job rows before each range, the tile's count
          const uint32_t tile = (uint32_t)t;
          const uint32_t c = lane < (uint32_t)kWorkerWarps ? tm.range_cnt[lane] : 0u;
          uint32_t incl = c;
#pragma unroll
          for (int d = 1; d < kWorkerWarps; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
            if ((int)lane >= d) incl += v;
          }
          const uint32_t total = __shfl_sync(0xffffffffu, incl, kWorkerWarps - 1);
          if (lane < (uint32_t)kWorkerWarps) tm.range_row[lane] = incl - c; // tile-local until the look-back is done
          if (lane == 0) {
            tm.total = total;
            st_desc_word(&p.desc[tile].c, desc_c(total, tile == 0 ? DESC_INC : DESC_AGG, p.gen));
            if (p.n_tiles == 1) *p.n_jobs = total;
            NXS_TRACE_EVENT(p, EV_COUNT_END, tile);
          }
          __syncwarp();
        }
        ++km;
      }
      wd.reset();
      continue;
    }
    if (kl < km) { // look up the first row index of the oldest own tile (one round trip per poll)
      const uint32_t ml = kl % METAS;
      TileMeta<TILE> &tm = metas[ml];
      const uint32_t tile = (uint32_t)tm.tile;
      if (!lb_active) {
        st = CountLookback{(int)tile - 1, 0u};
        lb_active = true;
        if (lane == 0) NXS_TRACE_EVENT(p, EV_LB_START, tile);
      }
      if (lane == 0) NXS_TRACE_EVENT(p, EV_LB_POLL, tile);
      if (tile == 0 || count_lookback_poll(p, st, lane)) {
        if (lane < (uint32_t)kWorkerWarps) tm.range_row[lane] += st.excl;
        if (lane == 0) {
          const uint32_t total = tm.total;
          if (tile != 0) st_desc_word(&p.desc[tile].c, desc_c(st.excl + total, DESC_INC, p.gen));
          if (tile == p.n_tiles - 1) *p.n_jobs = (uint64_t)st.excl + total;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&ctl.base_ready[ml]);
        if (lane == 0) NXS_TRACE_EVENT(p, EV_LB_END, tile);
        lb_active = false;
        kl += kCounterWarps;
        wd.reset();
        continue;
      }
      aux_nap<NXS_COUNT_LB_NS>();
      wd.tick(p.counters + 2);
      continue;
    }
    if (k_end != 0xFFFFFFFFu && kl >= k_end) break; // every own tile is done
    // nothing to do: wait for the next tile to be marked
    if (k_end == 0xFFFFFFFFu) {
      if (!__all_sync(0xffffffffu, mbar_try_wait(&ctl.marked[mm], (km / METAS) & 1u, 2000u))) aux_nap<NXS_COUNTER_NS>();
    }
    wd.tick(p.counters + 2);
  }
}

// ---------------------------------------------------------------------- workers
struct __align__(16) MaskRow { // ballots of one 32-record chunk (bit b = record b of the chunk)
  uint32_t he; // handled by a switch arm, guards passed, checkpoint exists
  uint32_t tm; // ... and the arm's new stage is a finished one (a stop)
  uint32_t fl; // ... and the arm is a failure action
  uint32_t er; // handled, guards passed, no checkpoint (:265-273)
};
#ifndef NXS_RECORD_GROUP
#define NXS_RECORD_GROUP 4 /* 32-record steps whose loads are issued together */
#endif
constexpr uint32_t kRecordGroup = NXS_RECORD_GROUP;

__device__ __forceinline__ uint32_t shr_clamp(uint32_t v, uint32_t n) { // n >= 32 gives 0
  uint32_t r;
  asm("shr.u32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));
  return r;
}

// A run of more than 32 records (a long job): word by word through the mask rows.  Rare, kept out of line.
template <int TILE, int STAGES>
__device__ __noinline__ Summ long_run(const StageBuf<TILE> &sb, const Ctl<STAGES> &ctl, uint32_t r0, uint32_t s,
                                      uint32_t e1, bool stop0, uint32_t gbase) {
  const uint8_t *const arms = sb.kind + r0;
  const MaskRow *const rows_m = reinterpret_cast<const MaskRow *>(sb.flags + r0);
  Summ S = summ_identity();
  for (uint32_t pos = s; pos < e1;) {
    const uint32_t ww = pos >> 5, ss = pos & 31u;
    const uint32_t take = min(32u - ss, e1 - pos);
    const uint32_t m = shr_clamp(0xFFFFFFFFu, 32u - take);
    const MaskRow r = rows_m[ww];
    uint32_t wm;
    bool liw;
    const uint32_t f = summ_run_flags((r.he >> ss) & m, (r.tm >> ss) & m, r.fl >> ss, (r.er >> ss) & m, pos == s && stop0, wm, liw);
    uint32_t codex = 0, wpos = pos;
    if (wm) {
      wpos = pos + hi_bit(wm);
      codex = ctl.codex[(arms[wpos] >> 2) & 7u];
    }
    S = summ_combine(S, Summ{gbase + wpos, summ_run_finish(f, codex, liw)});
    pos += take;
  }
  return S;
}

// One round of the job pass: lane `lane` takes run k = kb + lane of the range.  Run k = records
// [ends[k-1], ends[k]) (run 0 starts at the range's first record); the caller has put the end of the
// range behind the last job end, so the last run (k == njobs, `open_end`) is the job still open there.
// FIRST: this round holds run 0, the only one that can have begun before the range.
template <int TILE, bool FIRST, int STAGES>
__device__ __forceinline__ void job_round(const ClassifyParams &p, StageBuf<TILE> &sb, const Ctl<STAGES> &ctl,
                                          uint32_t warp, uint32_t r0, uint32_t k, uint32_t njobs, bool first_is_head,
                                          uint32_t gbase, nxs_job_row *prow) {
  const uint16_t *const ends = sb.reason + r0;
  const uint8_t *const arms = sb.kind + r0;
  const MaskRow *const rows_m = reinterpret_cast<const MaskRow *>(sb.flags + r0);
  uint32_t s = (ends - 1)[k];
  const uint32_t e1 = ends[k];
  if (FIRST && k == 0) s = 0;
  const uint32_t len = e1 - s;
  const bool anchored = !FIRST || k != 0 || first_is_head;
  // IsFinished() of the row the job's first record found (:276) — bit 7 of the byte the record pass
  // left for the record: a stop before the first record
  const bool stop0 = anchored && (arms[s] & 0x80u);
  Summ S;
  if (len <= 32u) {
    const uint32_t w = s >> 5, sh = s & 31u;
    const uint4 lo = *reinterpret_cast<const uint4 *>(rows_m + w), hi = *reinterpret_cast<const uint4 *>(rows_m + w + 1);
    const uint32_t m = shr_clamp(0xFFFFFFFFu, 32u - len);
    uint32_t wm;
    bool liw;
    const uint32_t f = summ_run_flags(__funnelshift_r(lo.x, hi.x, sh) & m, __funnelshift_r(lo.y, hi.y, sh) & m,
                                      __funnelshift_r(lo.z, hi.z, sh), __funnelshift_r(lo.w, hi.w, sh) & m, stop0, wm, liw);
    const uint32_t wpos = s + hi_bit(wm | 1u); // (the run's first record when there is no winner: dropped below)
    uint32_t codex = *reinterpret_cast<const uint32_t *>(reinterpret_cast<const uint8_t *>(ctl.codex) + (arms[wpos] & 0x1Cu));
    if (!wm) codex = 0;
    S = Summ{gbase + wpos, summ_run_finish(f, codex, liw)};
  } else {
    S = long_run(sb, ctl, r0, s, e1, stop0, gbase);
  }
  if (k == njobs) { // the job still open at the end of the range (this run exists only when there is one)
    if (anchored) S.f |= F_ANCHORED;
    sb.range[warp].suffix = summ_pack(S);
    sb.range[warp].flags = first_is_head ? RANGE_FIRST_IS_HEAD : 0u;
    if (njobs == 0 || first_is_head) sb.range[warp].has_unres = 0;
  } else {
    const uint64_t key = sb.keyx[2 + r0 + e1 - 1];
    if (anchored) {
      uint32_t verdict, winner;
      summ_resolve(S, verdict, winner);
      __stcs(reinterpret_cast<uint4 *>(prow), make_uint4((uint32_t)key, (uint32_t)(key >> 32), verdict, winner));
    } else { // the range's first job began before it: left to the summary warp
      sb.range[warp].unres = summ_pack(S);
      sb.range[warp].unres_key = key;
      sb.range[warp].has_unres = 1;
    }
  }
}

__device__ __forceinline__ void sts_u16(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((uint16_t)v) : "memory");
}

// ------------------------------------------------------------------ record pass
// onEvent: guards (:155,:161,:218) and the (kind, reason) switch arms, then superviseAction's order:
// no checkpoint (:265) -> finished (:276) -> switch (:283).  All of it is table lookups.  On the way
// the pass leaves what the job pass reads, in place of the columns it consumes (see StageBuf).
template <int TILE, bool FULL, int STAGES, bool CODES>
__device__ __forceinline__ void record_pass(const ClassifyParams &p, StageBuf<TILE> &sb, const TileMeta<TILE> &tm,
                                            const Ctl<STAGES> &ctl, uint32_t tile, uint32_t warp, uint32_t lane) {
  static_assert(TILE % (kWorkerWarps * 32 * kRecordGroup) == 0, "a range is a whole number of record groups");
  constexpr uint32_t RANGE = TILE / kWorkerWarps; // records per warp
  constexpr uint32_t CPW = RANGE / 32;            // 32-record chunks per warp
  const uint32_t start = tile * TILE;
  const uint32_t cnt = FULL ? (uint32_t)TILE : p.n - start;
  const uint32_t r0 = warp * RANGE;
  if (!FULL && r0 >= cnt) return; // nothing of the ragged tile falls in this range
  const uint32_t cnt_r = FULL ? RANGE : min(RANGE, cnt - r0); // records of the range that exist
  if (!FULL) { // a ragged last tile is not sent by the producer: every warp fetches its own range
    for (uint32_t i = lane; i < cnt_r; i += 32) {
      sb.keyx[2 + r0 + i] = p.key[start + r0 + i];
      sb.reason[r0 + i] = p.reason[start + r0 + i];
      sb.kind[r0 + i] = p.kind[start + r0 + i];
      sb.flags[r0 + i] = p.flags[start + r0 + i];
      sb.stage[r0 + i] = p.stage[start + r0 + i];
    }
    __syncwarp();
  }
  const uint32_t lt = (1u << lane) - 1u;
  uint8_t *const arms = sb.kind + r0;
  MaskRow *const rows_m = reinterpret_cast<MaskRow *>(sb.flags + r0);
  const bool write_verdict = p.verdict != nullptr; // (the per-record verdicts are optional: the job rows drive the I/O)
  uint32_t *pv = p.verdict + (size_t)start + r0 + lane;
  uint8_t *pc = p.verdict_code + (size_t)start + r0 + lane;
#pragma unroll 1
  for (uint32_t g = 0; g < CPW; g += kRecordGroup, pv += 32 * kRecordGroup, pc += 32 * kRecordGroup) {
    uint2 ti[kRecordGroup];
    uint32_t vt[kRecordGroup], off[kRecordGroup];
    // loads and lookups of the whole group first: the stores below land on bytes read here
#pragma unroll
    for (uint32_t u = 0; u < kRecordGroup; ++u) {
      const uint32_t i = (g + u) * 32 + lane;
      ti[u] = tm.tailinfo[r0 / 32 + g + u];
      vt[u] = 0;
      off[u] = 0;
      if (FULL || i < cnt_r) {
        const uint32_t kind = sb.kind[r0 + i], reason = sb.reason[r0 + i], fl = sb.flags[r0 + i], st = sb.stage[r0 + i];
        const uint32_t arm4 = ctl.lut[umin32(kind, 3u) * 33u + umin32(reason, 32u)];
        const uint32_t l2 = ctl.lut2[st * 8u + (fl & 7u)];
        off[u] = arm4 + l2; // 4 * arm | 32 * gate class | 128 * IsFinished(cur_stage): one byte
        vt[u] = *reinterpret_cast<const uint32_t *>(reinterpret_cast<const uint8_t *>(ctl.vt) + off[u]);
        if (CODES) pc[32 * u] = (uint8_t)off[u]; // (the host expands the byte: verdict_of_code)
        else if (write_verdict) __stcs(pv + 32 * u, vt[u] & kVerdictMask);
      }
    }
#pragma unroll
    for (uint32_t u = 0; u < kRecordGroup; ++u) {
      const uint32_t c = g + u, i = c * 32 + lane;
      const uint32_t HE = ballot_bit(vt[u], VT_HE), TM = ballot_bit(vt[u], VT_TERM);
      const uint32_t FL = ballot_bit(vt[u], VT_FAIL), ER = ballot_bit(vt[u], VT_ERR);
      if (FULL || i < cnt_r) arms[i] = (uint8_t)off[u];
      if (lane == 0) *reinterpret_cast<uint4 *>(rows_m + c) = make_uint4(HE, TM, FL, ER);
      // the list of the positions (+ 1) where jobs end, from the marker warps' bit masks
      const uint32_t T = ti[u].x;
      if ((T >> lane) & 1u) sts_u16(ti[u].y + 2u * __popc(T & lt), i + 1);
    }
  }
}

// --------------------------------------------------------------------- job pass (a lane per run)
template <int TILE, bool FULL, int STAGES>
__device__ __forceinline__ void job_pass(const ClassifyParams &p, StageBuf<TILE> &sb, const TileMeta<TILE> &tm,
                                         const Ctl<STAGES> &ctl, uint32_t tile, uint32_t warp, uint32_t lane) {
  constexpr uint32_t RANGE = TILE / kWorkerWarps;
  const uint32_t start = tile * TILE;
  const uint32_t cnt = FULL ? (uint32_t)TILE : p.n - start;
  const uint32_t r0 = warp * RANGE;
  if (!FULL && r0 >= cnt) { // nothing of the ragged tile falls in this range
    if (lane == 0) sb.range[warp] = RangeInfo{summ_pack(summ_identity()), 0ull, 0ull, 0u, 0u};
    return;
  }
  const uint32_t cnt_r = FULL ? RANGE : min(RANGE, cnt - r0);
  const uint32_t njobs = tm.range_cnt[warp];
  // Does the first record of the range begin a job, does its last record end one?  (warp-uniform)
  const bool first_is_head = (r0 == 0 ? tm.first_is_head : tm.tailinfo[r0 / 32 - 1].x >> 31) != 0;
  const bool closed = (tm.tailinfo[(r0 + cnt_r - 1) / 32].x >> ((cnt_r - 1) & 31u)) & 1u;
  // If the range does not end on a job boundary, the end of the range closes the job-end list: the run
  // behind the last job end is the job still open there.  (Then njobs < RANGE, so the entry exists.)
  if (lane == 0) {
    if (!closed) sb.reason[r0 + njobs] = (uint16_t)cnt_r;
    else { // nothing flows out of the range
      sb.range[warp].suffix = summ_pack(Summ{0u, F_ANCHORED});
      sb.range[warp].flags = RANGE_CLOSED | (first_is_head ? RANGE_FIRST_IS_HEAD : 0u);
      if (njobs == 0 || first_is_head) sb.range[warp].has_unres = 0;
    }
  }
  __syncwarp();
  const uint32_t nruns = njobs + (closed ? 0u : 1u);
  const uint32_t gbase = p.record_base + start + r0; // batch index of the range's first record
  nxs_job_row *prow = p.rows + ((size_t)tm.range_row[warp] + lane); // row of run `lane`
  if (lane < nruns) job_round<TILE, true, STAGES>(p, sb, ctl, warp, r0, lane, njobs, first_is_head, gbase, prow);
#pragma unroll 1
  for (uint32_t k = 32 + lane; k < nruns; k += 32) {
    prow += 32;
    job_round<TILE, false, STAGES>(p, sb, ctl, warp, r0, k, njobs, first_is_head, gbase, prow);
  }
  // the stage's columns were written here through the generic proxy; the next tile lands in them through the async proxy
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
}

template <int TILE, int STAGES>
__device__ __forceinline__ void worker_loop(const ClassifyParams &p, StageBuf<TILE> *stages, TileMeta<TILE> *metas,
                                            Ctl<STAGES> &ctl, uint32_t warp, uint32_t lane) {
  constexpr uint32_t METAS = STAGES + 1;
  uint32_t s = 0, par = 0, m = 0, mpar = 0; // next tile to classify: stage and slot of the description ring, with parities
  const uint32_t cols_addr = smem_u32_pinned(&ctl.cols[0]), full_addr = smem_u32_pinned(&ctl.full[0]);
  const uint32_t claimed_addr = smem_u32_pinned(&ctl.claimed[0]), marked_addr = smem_u32_pinned(&ctl.marked[0]);
  const uint32_t base_addr = smem_u32_pinned(&ctl.base_ready[0]), done_addr = smem_u32_pinned(&ctl.p1done[0]);
  Watchdog wd;
#ifdef NXS_TRACE_STATS
  uint64_t st_idle = 0, st_busy = 0, st_idle_base = 0;
#endif
  // An idle warp sleeps off the issue slots in a loop kept as short as possible (a barrier test, a
  // nap): every instruction an idle warp issues is taken from a busy one.
#define NXS_WORKER_WAIT(addr, parity)                                      \
  while (!__all_sync(0xffffffffu, mbar_test_wait_a((addr), (parity)))) {   \
    __nanosleep(NXS_IDLE_NS);                                              \
    wd.tick(p.counters + 2);                                               \
  }                                                                        \
  wd.reset()
  while (true) {
#ifdef NXS_TRACE_STATS
    const uint64_t t_a = global_timer_ns();
#endif
    NXS_WORKER_WAIT(claimed_addr + 8u * m, mpar);
    const TileMeta<TILE> &tm = metas[m];
    const int32_t t = tm.tile;
    if (t < 0) break; // (every role learns the end of the work from the description ring)
    StageBuf<TILE> &sb = stages[s];
    const bool full_tile = p.n - (uint32_t)t * TILE >= (uint32_t)TILE;
    // (1) the record pass: the tile's columns, and (long since there) the marker warps' bit masks
    NXS_WORKER_WAIT(cols_addr + 8u * s, par);
    NXS_WORKER_WAIT(marked_addr + 8u * m, mpar);
#ifdef NXS_TRACE_STATS
    uint64_t t_b = global_timer_ns();
    st_idle += t_b - t_a;
#endif
    if (warp == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_P1_START, (uint32_t)t);
    if (p.verdict_code == nullptr) {
      if (full_tile) record_pass<TILE, true, STAGES, false>(p, sb, tm, ctl, (uint32_t)t, warp, lane);
      else record_pass<TILE, false, STAGES, false>(p, sb, tm, ctl, (uint32_t)t, warp, lane);
    } else { // the host-pointer path, link-bound: one byte per record goes out instead of four
      if (full_tile) record_pass<TILE, true, STAGES, true>(p, sb, tm, ctl, (uint32_t)t, warp, lane);
      else record_pass<TILE, false, STAGES, true>(p, sb, tm, ctl, (uint32_t)t, warp, lane);
    }
    __syncwarp();
    // (2) the job pass: the index of the tile's first job row (known since before the tile landed) and its keys
#ifdef NXS_TRACE_STATS
    const uint64_t t_c = global_timer_ns();
#endif
    NXS_WORKER_WAIT(base_addr + 8u * m, mpar);
    NXS_WORKER_WAIT(full_addr + 8u * s, par);
#ifdef NXS_TRACE_STATS
    const uint64_t t_d = global_timer_ns();
    st_idle += t_d - t_c;
    st_idle_base += t_d - t_c;
    t_b += t_d - t_c;
#endif
    if (full_tile) job_pass<TILE, true, STAGES>(p, sb, tm, ctl, (uint32_t)t, warp, lane);
    else job_pass<TILE, false, STAGES>(p, sb, tm, ctl, (uint32_t)t, warp, lane);
    if (warp == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_P1_END, (uint32_t)t);
    __syncwarp();
#ifdef NXS_TRACE_STATS
    st_busy += global_timer_ns() - t_b;
#endif
    if (lane == 0) mbar_arrive_a(done_addr + 8u * s);
    if (++s == STAGES) s = 0, par ^= 1u;
    if (++m == METAS) m = 0, mpar ^= 1u;
  }
#undef NXS_WORKER_WAIT
#ifdef NXS_TRACE_STATS
  if (lane == 0) {
    NXS_TRACE_STAT(p, EV_W_IDLE, (uint32_t)(st_idle >> 4) | warp << 28);
    NXS_TRACE_STAT(p, EV_W_BUSY, (uint32_t)(st_busy >> 4) | warp << 28);
    NXS_TRACE_STAT(p, EV_C_IDLE, (uint32_t)(st_idle_base >> 4) | warp << 28);
  }
#endif
}

// ----------------------------------------------------------------- summary warp
struct SummLookback {
  int pos;
  Summ carry;
};
#ifndef NXS_SUMM_WINDOWS
#define NXS_SUMM_WINDOWS 2
#endif
constexpr int kSummWindows = NXS_SUMM_WINDOWS;
// One non-blocking step of the summary look-back: walks predecessor summary words,
// nearest first, until one is anchored (holds the first record of the job).
__device__ __forceinline__ bool summ_lookback_poll(const ClassifyParams &p, SummLookback &st, uint32_t lane) {
  uint64_t s[kSummWindows];
#pragma unroll
  for (int k = 0; k < kSummWindows; ++k) {
    const int idx = st.pos - (int)lane - 32 * k;
    s[k] = idx >= 0 ? ld_desc_word(&p.desc[idx].s) : 0ull;
  }
#pragma unroll
  for (int k = 0; k < kSummWindows; ++k) {
    const bool valid = st.pos - (int)lane >= 0;
    const uint32_t ss = desc_s_status(s[k], p.gen);
    const Summ mine = desc_s_summ(s[k]);
    const uint32_t m_valid = __ballot_sync(0xffffffffu, valid);
    const uint32_t m_ok = __ballot_sync(0xffffffffu, valid && ss != DESC_INVALID);
    const uint32_t m_anch = __ballot_sync(0xffffffffu, valid && ss != DESC_INVALID && (mine.f & F_ANCHORED));
    const uint32_t need = m_anch ? (2u << lo_bit(m_anch)) - 1u : m_valid;
    if (need & ~m_ok) return false;
    const uint32_t upto = hi_bit(need);
    for (uint32_t j = 0; j <= upto; ++j) st.carry = summ_combine(shfl_summ(mine, (int)j), st.carry);
    if (m_anch) return true;
    st.pos -= 32;
  }
  return false;
}

template <int TILE, int STAGES>
__device__ __forceinline__ void summary_loop(const ClassifyParams &p, StageBuf<TILE> *stages, TileMeta<TILE> *metas,
                                             Ctl<STAGES> &ctl, uint32_t lane) {
  constexpr uint32_t METAS = STAGES + 1;
  uint32_t s = 0, par = 0, m = 0, mpar = 0; // next tile: stage and slot of the description ring, with parities
  uint32_t q_head = 0, q_tail = 0; // pending tiles: ring [q_head, q_tail)
  bool end = false, lb_active = false;
  SummLookback st{0, summ_identity()};
  Watchdog wd;
  while (true) {
    // A finished tile comes first (frees its stage, publishes its summary) unless the pending ring is full.
    if (!end && __all_sync(0xffffffffu, mbar_test_wait(&ctl.claimed[m], mpar)) && metas[m].tile < 0) {
      end = true; // (the end of the work is learnt from the description ring)
      continue;
    }
    if (!end && q_tail - q_head < (uint32_t)kPendingSlots &&
        __all_sync(0xffffffffu, mbar_test_wait(&ctl.claimed[m], mpar)) &&
        __all_sync(0xffffffffu, mbar_test_wait(&ctl.p1done[s], par))) {
      StageBuf<TILE> &sb = stages[s];
      const TileMeta<TILE> &tm = metas[m];
      const uint32_t tile = (uint32_t)tm.tile;
      if (lane == 0) NXS_TRACE_EVENT(p, EV_SUM_START, tile);
      // lanes 0..7 take over what the workers left for their range, then the stage is free
      const bool mine = lane < (uint32_t)kWorkerWarps;
      const Summ suffix = mine ? summ_unpack(sb.range[lane].suffix) : summ_identity();
      const uint32_t flags = mine ? sb.range[lane].flags : 0u;
      const bool has_unres = mine && sb.range[lane].has_unres;
      const Summ unres = has_unres ? summ_unpack(sb.range[lane].unres) : summ_identity();
      const uint64_t ukey = has_unres ? sb.range[lane].unres_key : 0ull;
      const uint32_t rowpos = mine ? tm.range_row[lane] : 0u;
      const bool first_is_head = tm.first_is_head != 0;
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl.freeb[s]); // (stage and description slot)
      if (++s == STAGES) s = 0, par ^= 1u;
      if (++m == METAS) m = 0, mpar ^= 1u;

      // Which ranges hold the first record of the job they leave open (or end on a job boundary)?
      const uint32_t anch_mask = __ballot_sync(0xffffffffu, mine && (suffix.f & F_ANCHORED));
      const uint32_t closed_mask = __ballot_sync(0xffffffffu, mine && (flags & RANGE_CLOSED));
      // (1) the job open at the tile's end: the ranges from the last such range to the right
      //     (usually just the last range)
      Summ T = summ_identity();
      {
        const uint32_t stop = anch_mask | closed_mask;
        const int k0 = stop ? (int)hi_bit(stop) : 0;
        if ((closed_mask >> k0) & 1u) {
          T = Summ{0u, F_ANCHORED}; // nothing flows into the next tile (everything right of k0 is empty)
        } else {
          for (int w = k0; w < kWorkerWarps; ++w) T = summ_combine(T, shfl_summ(suffix, w));
        }
      }
      // (2) for every range whose first job began before it: what the earlier ranges of the tile add,
      //     nearest first, until one holds the job's first record (usually the range next to it)
      Summ acc = summ_identity();
      {
        bool need = has_unres && lane > 0;
        for (int d = 1; d < kWorkerWarps && __any_sync(0xffffffffu, need); ++d) {
          const Summ sw = Summ{__shfl_up_sync(0xffffffffu, suffix.win_idx, d), __shfl_up_sync(0xffffffffu, suffix.f, d)};
          if (need) {
            acc = summ_combine(sw, acc);
            if ((sw.f & F_ANCHORED) || (int)lane == d) need = false;
          }
        }
      }
      // (3) rows whose job began inside the tile can be closed now; the others wait for the carry
      bool waits = false;
      if (has_unres) {
        const Summ S = summ_combine(acc, unres);
        if (S.f & F_ANCHORED) {
          uint32_t verdict, winner;
          summ_resolve(S, verdict, winner);
          __stcs(reinterpret_cast<uint4 *>(p.rows + rowpos),
                 make_uint4((uint32_t)ukey, (uint32_t)(ukey >> 32), verdict, winner));
        } else {
          waits = true;
        }
      }
      const uint32_t wait_mask = __ballot_sync(0xffffffffu, waits);
      const bool t_anchored = (T.f & F_ANCHORED) != 0;
      if (tile == 0 || (first_is_head && wait_mask == 0)) {
        // nothing flows in from earlier tiles: the tile's summary is already inclusive
        if (lane == 0) st_desc_word(&p.desc[tile].s, desc_s(T, DESC_INC, p.gen));
      } else {
        // an anchored summary is already inclusive even while some rows of the tile still wait
        if (lane == 0) st_desc_word(&p.desc[tile].s, desc_s(T, t_anchored ? DESC_INC : DESC_AGG, p.gen));
        PendingTile &q = ctl.pending[q_tail % kPendingSlots];
        if (waits) {
          q.unres[lane] = summ_pack(summ_combine(acc, unres));
          q.key[lane] = ukey;
          q.rowpos[lane] = rowpos;
        }
        if (lane == 0) {
          q.mask = wait_mask;
          q.tile = tile;
          q.local_summ = summ_pack(T);
          q.need_inc = t_anchored ? 0u : 1u;
        }
        __syncwarp();
        ++q_tail;
      }
      if (lane == 0) NXS_TRACE_EVENT(p, EV_SUM_END, tile);
      wd.reset();
      continue;
    }
    if (q_head != q_tail) { // oldest pending tile: look up what earlier tiles carry into it
      PendingTile &q = ctl.pending[q_head % kPendingSlots];
      if (!lb_active) {
        st.pos = (int)q.tile - 1;
        st.carry = summ_identity();
        lb_active = true;
      }
      if (summ_lookback_poll(p, st, lane)) {
        const uint32_t tile = q.tile;
        if (lane == 0 && q.need_inc) // a tile with no job start of its own passes the carried summary on
          st_desc_word(&p.desc[tile].s, desc_s(summ_combine(st.carry, summ_unpack(q.local_summ)), DESC_INC, p.gen));
        if ((q.mask >> lane) & 1u) {
          uint32_t verdict, winner;
          summ_resolve(summ_combine(st.carry, summ_unpack(q.unres[lane])), verdict, winner);
          const uint64_t key = q.key[lane];
          __stcs(reinterpret_cast<uint4 *>(p.rows + q.rowpos[lane]),
                 make_uint4((uint32_t)key, (uint32_t)(key >> 32), verdict, winner));
        }
        __syncwarp();
        if (lane == 0) NXS_TRACE_EVENT(p, EV_SLB_END, tile);
        lb_active = false;
        ++q_head;
        wd.reset();
        continue;
      }
      aux_nap<NXS_SUMM_LB_NS>();
    } else if (end) {
      break;
    } else if (__all_sync(0xffffffffu, mbar_try_wait(&ctl.claimed[m], mpar, 2000u))) {
      if (metas[m].tile >= 0 && !__all_sync(0xffffffffu, mbar_try_wait(&ctl.p1done[s], par, 2000u))) aux_nap<NXS_SUMMARY_NS>();
    }
    wd.tick(p.counters + 2);
  }
}

#ifndef NXS_MAXNREG
#define NXS_MAXNREG 72 /* 832 threads x 72 registers fit the register file */
#endif
template <int TILE>
__global__ void __maxnreg__(NXS_MAXNREG) classify_kernel(const __grid_constant__ ClassifyParams p) {
  constexpr int STAGES = StagesFor<TILE>::value;
  // (no manual re-alignment: keeping the pointer derived from the array keeps it in the shared
  // address space, so accesses compile to LDS/STS rather than generic loads)
  extern __shared__ __align__(128) uint8_t smem_raw[];
  StageBuf<TILE> *stages = reinterpret_cast<StageBuf<TILE> *>(smem_raw);
  TileMeta<TILE> *metas = reinterpret_cast<TileMeta<TILE> *>(smem_raw + sizeof(StageBuf<TILE>) * STAGES);
  // static, so that the addresses of the barriers and tables are immediates in the hot loops
  __shared__ Ctl<STAGES> ctl;
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&ctl.full[s], 1);
      mbar_init(&ctl.cols[s], 1);
      mbar_init(&ctl.p1done[s], kWorkerWarps);
      mbar_init(&ctl.freeb[s], 1);
    }
    for (int m = 0; m < STAGES + 1; ++m) {
      mbar_init(&ctl.claimed[m], 1);
      mbar_init(&ctl.marked[m], kMarkerWarps);
      mbar_init(&ctl.base_ready[m], 1);
    }
    mbar_fence_init();
  }
  if (tid < 136) ctl.lut[tid] = (tid < 132 && tid % 33u < 32u) ? (uint8_t)(4u * p.lut[tid / 33u * 32u + tid % 33u]) : 0;
  for (uint32_t e = tid; e < 2048u; e += kThreads) {
    const uint32_t fl = e & 7u, st = (e >> 3) & 15u;
    const uint32_t fin = (p.finished_mask >> st) & 1u;
    ctl.lut2[e] = (uint8_t)(32u * ((fl & 3u) != 3u ? 0u : !(fl & NXS_FLAG_CHECKPOINT) ? 1u : fin ? 2u : 3u) + 128u * fin);
  }
  if (tid < 8) ctl.codex[tid] = p.code[tid] ? (p.code[tid] | F_HAS_WIN | F_HAS_APP) : 0u;
  if (tid >= 128 && tid < 192) {
    const uint32_t e = tid - 128;
    ctl.vt[e] = vt_entry(e, p.code, p.term_mask);
  }
  __syncthreads(); // the only CTA-wide barrier

  if (warp < (uint32_t)kWorkerWarps) worker_loop<TILE, STAGES>(p, stages, metas, ctl, warp, lane);
  else if (warp == (uint32_t)kProducerWarp) {
    if (lane == 0) producer_loop<TILE, STAGES>(p, stages, metas, ctl);
  } else if (warp == (uint32_t)kSummaryWarp) summary_loop<TILE, STAGES>(p, stages, metas, ctl, lane);
  else if (warp < (uint32_t)kMarkerWarp0) lookback_loop<TILE, STAGES>(p, metas, ctl, warp - kCounterWarp0, lane);
  else marker_loop<TILE, STAGES>(p, stages, metas, ctl, warp - kMarkerWarp0, lane);
}

#endif // __CUDACC__

} // namespace nxs
