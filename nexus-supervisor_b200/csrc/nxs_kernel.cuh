// nxs_kernel.cuh — the classification kernel (sm_100a).
//
// One persistent grid makes a single pass over HBM: it classifies every record
// (onEvent + superviseAction decision, services/supervisor.go:137-374) and folds
// the records of each job into one row.  Each CTA is a warp-specialised pipeline
// over a ring of shared-memory stages, driven by mbarriers only (no CTA-wide
// barrier):
//   producer warp  claims tile ids from a global counter and streams the five
//                  input columns of a tile into a stage with 1-D bulk async
//                  copies (TMA engine, SASS UBLKCP)
//   worker warps   phase 1 as soon as a tile has landed.  Each warp owns a
//                  contiguous range of the tile (TILE/8 records) and walks it 32
//                  records at a time, lane per record: table lookups -> packed
//                  verdict -> global; warp ballots -> bit masks kept in
//                  registers; the lane holding the LAST record of a job folds the
//                  job's masks into its row right away (codes of other lanes come
//                  by shuffle, the job left open by the previous 32 records is
//                  carried in registers) and parks key + packed row in shared
//                  memory, compacted per range.
//                  phase 2 once the tile's look-back is done: patch the one job
//                  per range that began before the range, then copy the parked
//                  rows to their final position in global memory
//   scan warp      event loop: (a) as soon as phase 1 of a tile is complete,
//                  publish its aggregate (job count + summary of the job left
//                  open at its end) — this never waits on another CTA;
//                  (b) poll predecessor descriptors (decoupled look-back) for
//                  the oldest pending tile; when complete, publish the inclusive
//                  descriptor and release phase 2
// Scratch: TileDesc[n_tiles] (16 B each) + 3 counters.
#pragma once
#include "nxs_device.cuh"

namespace nxs {

constexpr int kWorkerWarps = 8;
constexpr int kWorkers = kWorkerWarps * 32;
constexpr int kProducerWarp = kWorkerWarps;
constexpr int kScanWarp = kWorkerWarps + 1;
constexpr int kThreads = kWorkers + 64;
constexpr int kStages = 3;
constexpr uint64_t kWatchdogNs = 8000000000ull; // a stuck wait traps instead of hanging the GPU

// What a worker warp leaves behind for its range of the tile.
struct __align__(8) RangeInfo {
  uint64_t suffix; // summary of the job still open at the end of the range (restricted to the range)
  uint64_t unres;  // summary of the range's first job when it began before the range
  uint32_t count;  // jobs ending in the range
  uint32_t flags;
  uint32_t has_unres; // set by the lane that parked `unres`
  uint32_t pad;
};
enum : uint32_t { RANGE_CLOSED = 1u, RANGE_FIRST_IS_HEAD = 4u };

// Parked row (phase 1 -> phase 2): Summ.f[0,16) | job already finished at its first record[16] |
// winner index within the tile[17,..)
constexpr uint32_t kParkedFin0 = 1u << 16;
constexpr uint32_t kVerdictMask = 0x1FFFFu;

template <int TILE> struct __align__(128) StageBuf {
  // keyx[1] = key of the record before the tile, keyx[2..TILE+1] = the tile's
  // keys (16-byte aligned for the bulk copy), keyx[TILE+2] = key after the tile.
  // Phase 1 compacts the keys of each range's job-ending records in place.
  uint64_t keyx[TILE + 4];
  uint16_t reason[TILE];
  uint8_t kind[TILE];
  uint8_t flags[TILE];
  uint8_t stage[TILE];
  uint32_t packed[TILE]; // parked rows, compacted per range like the keys
  RangeInfo range[kWorkerWarps];
  uint32_t range_off[kWorkerWarps]; // jobs ending before each range, tile-local
  uint64_t carry;                   // summary carried into the tile's first job
  uint64_t local_summ;              // summary of the job left open at the tile's end
  uint32_t local_count;             // jobs ending in the tile
  uint32_t tile_base;               // job rows before the tile
  int32_t tile;                     // tile id, -1 = end of work
};

// Verdict table entry = packed verdict of the record plus the fold bits above it.
constexpr uint32_t VT_HE = 1u << 17, VT_TERM = 1u << 18, VT_FAIL = 1u << 19, VT_ERR = 1u << 20;

struct Ctl {
  uint64_t full[kStages];   // tile landed            (producer -> workers)
  uint64_t p1done[kStages]; // phase 1 complete       (workers -> scan warp)
  uint64_t lbdone[kStages]; // look-back complete     (scan warp -> workers)
  uint64_t freeb[kStages];  // phase 2 complete       (workers -> producer)
  uint32_t vt[1024];        // [gate class 0..3][code 0..255] -> verdict | fold bits
  uint8_t lut[128];         // (kind, reason) -> code (action | cause_tpl << 3 | msg_tpl << 5)
  uint16_t lut2[128];       // (flags & 7) | stage << 3 -> 256 * gate class (0 dropped, 1 no checkpoint, 2 finished, 3 apply)
};

// Tile descriptor: two self-validating 64-bit words, each published with one
// relaxed store, so no fence is needed on either side.
//   c: count[0,32) | status[32,34) | generation[34,41)
//   s: win_idx[0,32) | Summ.f[32,55) | status[55,57) | generation[57,64)
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

struct ClassifyParams {
  const uint8_t *kind;
  const uint16_t *reason;
  const uint8_t *flags;
  const uint8_t *stage;
  const uint64_t *key;
  uint32_t *verdict;
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
  uint8_t lut[128];
#ifdef NXS_TRACE
  unsigned long long *trace; // [0] = number of events, then (time, cta<<40 | event<<32 | tile) pairs
  uint32_t trace_cap;
#endif
};

template <int TILE> constexpr size_t classify_smem_bytes() {
  return sizeof(StageBuf<TILE>) * kStages + sizeof(Ctl);
}

#if defined(__CUDACC__)

__device__ __forceinline__ void ld_desc(const TileDesc *d, uint64_t &c, uint64_t &s) {
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(c), "=l"(s) : "l"(d) : "memory");
}
__device__ __forceinline__ void st_desc(TileDesc *d, uint64_t c, uint64_t s) {
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(d), "l"(c), "l"(s) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void named_barrier_workers() { asm volatile("bar.sync 1, %0;" ::"n"(kWorkers) : "memory"); }

#ifdef NXS_TRACE
// Debug build only: CTAs 0, 97 and 211 log pipeline events with the global timer.
enum : uint32_t { EV_CLAIM = 1, EV_ISSUED, EV_P1_START, EV_P1_END, EV_AGG_START, EV_AGG_END, EV_LB_START, EV_LB_END,
                  EV_P2_START, EV_P2_END, EV_LB_POLL };
__device__ __forceinline__ void trace_event(const ClassifyParams &p, uint32_t ev, uint32_t tile) {
  if (!p.trace || !(blockIdx.x == 0 || blockIdx.x == 97 || blockIdx.x == 211)) return;
  const unsigned long long i = atomicAdd(p.trace, 1ull);
  if (i >= p.trace_cap) return;
  p.trace[1 + 2 * i] = global_timer_ns();
  p.trace[2 + 2 * i] = (unsigned long long)blockIdx.x << 40 | (unsigned long long)ev << 32 | tile;
}
#define NXS_TRACE_EVENT(p, ev, tile) trace_event(p, ev, tile)
#else
#define NXS_TRACE_EVENT(p, ev, tile) ((void)0)
#endif

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

// --------------------------------------------------------------------- producer
template <int TILE>
__device__ __forceinline__ void producer_loop(const ClassifyParams &p, StageBuf<TILE> *stages, Ctl &ctl) {
  uint32_t next = atomicAdd(p.counters, 1u);
  Watchdog wd;
  for (uint32_t k = 0;; ++k) {
    const uint32_t s = k % kStages, use = k / kStages;
    if (use > 0) {
      while (!mbar_try_wait(&ctl.freeb[s], (use - 1) & 1u, 20000u)) wd.tick(p.counters + 2);
      wd.reset();
    }
    StageBuf<TILE> &sb = stages[s];
    uint64_t *bar = &ctl.full[s];
    const uint32_t id = next;
    if (id >= p.n_tiles) {
      sb.tile = -1;
      mbar_arrive_n(bar, 2);
      break;
    }
    NXS_TRACE_EVENT(p, EV_CLAIM, id);
    next = atomicAdd(p.counters, 1u); // consumed one iteration later: latency hidden
    sb.tile = (int32_t)id;
    const uint32_t start = id * TILE;
    if (p.n - start >= (uint32_t)TILE) {
      mbar_arrive_expect_tx(bar, 13u * TILE);
      bulk_g2s(&sb.keyx[2], p.key + start, 8u * TILE, bar);
      bulk_g2s(sb.reason, p.reason + start, 2u * TILE, bar);
      bulk_g2s(sb.kind, p.kind + start, TILE, bar);
      bulk_g2s(sb.flags, p.flags + start, TILE, bar);
      bulk_g2s(sb.stage, p.stage + start, TILE, bar);
      if (start > 0) cp_async_8(&sb.keyx[1], p.key + start - 1);
      if (start + TILE < p.n) cp_async_8(&sb.keyx[2 + TILE], p.key + start + TILE);
      mbar_arrive_after_cp_async(bar);
      NXS_TRACE_EVENT(p, EV_ISSUED, id);
    } else {
      mbar_arrive_n(bar, 2); // ragged last tile: the workers load it themselves
    }
  }
  // leave the counters zero for the next launch: the last CTA to stop claiming resets them
  __threadfence();
  if (atomicAdd(p.counters + 1, 1u) == gridDim.x - 1) {
    p.counters[0] = 0;
    p.counters[1] = 0;
  }
}

// ---------------------------------------------------------------------- workers
__device__ __forceinline__ Summ shfl_summ(Summ s, int src) {
  return Summ{__shfl_sync(0xffffffffu, s.win_idx, src), __shfl_sync(0xffffffffu, s.f, src)};
}
__device__ __forceinline__ Summ shfl_down_summ(Summ s, int d) {
  return Summ{__shfl_down_sync(0xffffffffu, s.win_idx, d), __shfl_down_sync(0xffffffffu, s.f, d)};
}

template <int TILE, bool FULL>
__device__ __forceinline__ void phase1(const ClassifyParams &p, StageBuf<TILE> &sb, const Ctl &ctl, uint32_t tile,
                                       uint32_t warp, uint32_t lane) {
  constexpr uint32_t RANGE = TILE / kWorkerWarps; // records per warp
  constexpr uint32_t CPW = RANGE / 32;            // 32-record chunks per warp
  const uint32_t start = tile * TILE;
  const uint32_t cnt = FULL ? (uint32_t)TILE : p.n - start;
  if (!FULL) { // ragged tile: guarded loads by the worker warps, then a worker-only barrier
    for (uint32_t i = warp * 32 + lane; i < cnt; i += kWorkers) {
      sb.keyx[2 + i] = p.key[start + i];
      sb.reason[i] = p.reason[start + i];
      sb.kind[i] = p.kind[start + i];
      sb.flags[i] = p.flags[start + i];
      sb.stage[i] = p.stage[start + i];
    }
    if (warp == 0 && lane == 0 && start > 0) sb.keyx[1] = p.key[start - 1];
    named_barrier_workers();
  }
  const uint32_t r0 = warp * RANGE;
  if (!FULL && r0 >= cnt) { // nothing of the ragged tile falls in this range
    if (lane == 0) sb.range[warp] = RangeInfo{summ_pack(summ_identity()), 0ull, 0u, 0u, 0u, 0u};
    return;
  }
  const bool last_tile = start + cnt == p.n;
  const uint32_t gbase = p.record_base + start;
  const uint32_t le = (2u << lane) - 1u, lt = le >> 1;
  // Does the first record of the range begin a job?  (warp-uniform)
  uint32_t prev_bit = (r0 == 0 && start == 0) ? 1u : (uint32_t)(sb.keyx[2 + r0] != sb.keyx[1 + r0]);
  const uint32_t first_is_head = prev_bit;
  Summ ctx = summ_identity(); // the job left open by the records of this range seen so far
  uint32_t wcount = 0, closed = 0;
  if (lane == 0) sb.range[warp].has_unres = 0;
  __syncwarp();

#pragma unroll
  for (uint32_t j = 0; j < CPW; ++j) {
    const uint32_t i = r0 + j * 32 + lane;
    const bool valid = FULL || i < cnt;
    uint32_t st = 0, vt = 0, code = 0;
    uint64_t key = 0;
    bool tail = false;
    if (valid) {
      const uint32_t kind = sb.kind[i], reason = sb.reason[i], fl = sb.flags[i];
      st = sb.stage[i] & 15u;
      key = sb.keyx[2 + i];
      const uint64_t nk = sb.keyx[3 + i];
      tail = key != nk;
      if (last_tile && i == cnt - 1) tail = true;
      // onEvent: guards (:155,:161,:218) and the (kind, reason) switch arms, then superviseAction's
      // order: no checkpoint (:265) -> finished (:276) -> switch (:283).  All three are table lookups.
      code = ctl.lut[min(kind, 3u) * 32u + (reason < 32u ? reason : 0u)];
      vt = ctl.vt[ctl.lut2[(fl & 7u) | st << 3] + code];
      __stcs(p.verdict + start + i, vt & kVerdictMask);
    }
    const uint32_t T = __ballot_sync(0xffffffffu, tail);
    const uint32_t H = T << 1 | prev_bit; // bit b: record b of the chunk begins a job
    prev_bit = T >> 31;
    const uint32_t HE = ballot_bit(vt, VT_HE), TM = ballot_bit(vt, VT_TERM);
    const uint32_t FL = ballot_bit(vt, VT_FAIL), ER = ballot_bit(vt, VT_ERR);

    // fold of the records from the job's first record in this chunk (or the chunk start) up to this lane
    const uint32_t h_in = H & le;
    const uint32_t hb = hi_bit(h_in | 1u); // h_in == 0 -> 0: the run starts at the chunk start
    const uint32_t seg = le & (~0u << hb);
    const uint32_t hes = HE & seg, tms = TM & seg;
    const uint32_t upto = tms ^ (tms - 1u); // through the first stop; all ones when there is none
    const uint32_t app = hes & upto;        // applied records (under the premise)
    const uint32_t wm = app & FL;           // ... with a failure action
    const uint32_t wp = hi_bit(wm | 1u);    // the winner: the last of them
    const uint32_t code_w = __shfl_sync(0xffffffffu, code, wp);
    const uint32_t st_head = __shfl_sync(0xffffffffu, st, hb);
    uint32_t f = 0;
    if (ER & seg) f |= F_ERR;
    if (tms) f |= F_STOPPED;
    if (hes & ~upto) f |= F_SKIP;
    if (hes) f |= F_HAS_APP | NXS_ACTION_RUNNING; // a non-failure handled record is a ToRunning
    if (wm) {
      f |= F_HAS_WIN | (code_w >> 3) << F_WIN_SHIFT;
      if ((app >> wp) == 1u) f = (f & ~F_ALAST_MASK) | (code_w & 7u); // nothing applied after the winner
    }
    Summ S{gbase + r0 + j * 32 + wp, f};
    if (h_in) S.f |= st_head << F_S0_SHIFT | F_ANCHORED; // the job begins in this chunk
    else S = summ_combine(ctx, S);                        // it continues the job left open before the chunk

    __syncwarp(); // every lane has read its keys: compaction may overwrite them now
    if (tail) {
      const uint32_t slot = r0 + wcount + __popc(T & lt);
      sb.keyx[2 + slot] = key;
      if (S.f & F_ANCHORED) {
        const uint32_t fin0 = (p.finished_mask >> ((S.f >> F_S0_SHIFT) & 15u)) & 1u;
        sb.packed[slot] = (S.f & 0xFFFFu) | fin0 << 16 | (S.win_idx - gbase) << 17;
      } else { // only the first job of a range can have begun before it
        sb.range[warp].unres = summ_pack(S);
        sb.range[warp].has_unres = 1;
      }
    }
    wcount += __popc(T);
    if (FULL || r0 + j * 32 + 31 < cnt) closed = T >> 31;
    else if (last_tile) closed = 1; // the ragged tile ends inside this chunk: its last record closes a job
    ctx = shfl_summ(S, 31);
  }
  __syncwarp();
  if (lane == 0) {
    sb.range[warp].suffix = summ_pack(ctx);
    sb.range[warp].count = wcount;
    sb.range[warp].flags = (closed ? RANGE_CLOSED : 0u) | (first_is_head ? RANGE_FIRST_IS_HEAD : 0u);
  }
}

template <int TILE>
__device__ __forceinline__ void phase2(const ClassifyParams &p, const StageBuf<TILE> &sb, uint32_t warp,
                                       uint32_t lane) {
  constexpr uint32_t RANGE = TILE / kWorkerWarps;
  const uint32_t count = sb.range[warp].count;
  if (count == 0) return;
  const uint32_t r0 = warp * RANGE;
  const uint32_t gbase = p.record_base + (uint32_t)sb.tile * TILE;
  nxs_job_row *rows = p.rows + sb.tile_base + sb.range_off[warp];
  uint32_t first = 0;
  if (sb.range[warp].has_unres) { // the range's first job began before the range: gather what came before
    if (lane == 0) {
      Summ acc = summ_identity();
      bool anchored = false;
      for (int w = (int)warp - 1; w >= 0 && !anchored; --w) {
        const Summ sfx = summ_unpack(sb.range[w].suffix);
        acc = summ_combine(sfx, acc);
        anchored = (sfx.f & F_ANCHORED) != 0;
      }
      if (!anchored) acc = summ_combine(summ_unpack(sb.carry), acc);
      uint32_t verdict, winner;
      summ_resolve(summ_combine(acc, summ_unpack(sb.range[warp].unres)), p.finished_mask, verdict, winner);
      const uint64_t key = sb.keyx[2 + r0];
      __stcs(reinterpret_cast<uint4 *>(rows), make_uint4((uint32_t)key, (uint32_t)(key >> 32), verdict, winner));
    }
    first = 1;
  }
  for (uint32_t j = first + lane; j < count; j += 32) {
    const uint64_t key = sb.keyx[2 + r0 + j];
    const uint32_t pk = sb.packed[r0 + j];
    // close the job (summ_resolve on the parked fields)
    const bool fin0 = (pk & kParkedFin0) != 0;
    uint32_t v = fin0 ? (((pk & F_HAS_APP) ? F_SKIP : 0u) | (pk & F_ERR))
                      : ((pk & F_VERDICT_MASK) | new_stage_of(pk & F_ALAST_MASK) << 3);
    v |= (pk & F_ERR) >> 3;
    const uint32_t winner = (!fin0 && (pk & F_HAS_WIN)) ? gbase + (pk >> 17) : NXS_NO_WINNER;
    __stcs(reinterpret_cast<uint4 *>(rows + j), make_uint4((uint32_t)key, (uint32_t)(key >> 32), v, winner));
  }
}

template <int TILE>
__device__ __forceinline__ void worker_loop(const ClassifyParams &p, StageBuf<TILE> *stages, Ctl &ctl,
                                            uint32_t warp, uint32_t lane) {
  uint32_t k1 = 0, s1 = 0, par1 = 0; // next tile for phase 1: ordinal, stage, parity
  uint32_t k2 = 0, s2 = 0, par2 = 0; // next tile for phase 2
  bool end = false;
  Watchdog wd;
  while (true) {
    // Phase 1 of a landed tile comes first (it feeds the aggregate every later tile waits for); a warp
    // with nothing to fold sleeps on the barrier it needs next instead of polling.
    bool landed = false;
    if (!end) {
      const bool p2_pending = k2 < k1;
      landed = __all_sync(0xffffffffu, p2_pending ? mbar_test_wait(&ctl.full[s1], par1)
                                                  : mbar_try_wait(&ctl.full[s1], par1, 2000u));
    }
    if (landed) {
      StageBuf<TILE> &sb = stages[s1];
      const int32_t t = sb.tile;
      if (t < 0) {
        end = true;
      } else {
        if (warp == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_P1_START, (uint32_t)t);
        if (p.n - (uint32_t)t * TILE >= (uint32_t)TILE) phase1<TILE, true>(p, sb, ctl, (uint32_t)t, warp, lane);
        else phase1<TILE, false>(p, sb, ctl, (uint32_t)t, warp, lane);
        if (warp == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_P1_END, (uint32_t)t);
        ++k1;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl.p1done[s1]);
      if (++s1 == kStages) s1 = 0, par1 ^= 1u;
      wd.reset();
      continue;
    }
    if (k2 < k1) {
      if (__all_sync(0xffffffffu, mbar_try_wait(&ctl.lbdone[s2], par2, end ? 2000u : 300u))) {
        if (warp == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_P2_START, (uint32_t)stages[s2].tile);
        phase2<TILE>(p, stages[s2], warp, lane);
        if (warp == 0 && lane == 0) NXS_TRACE_EVENT(p, EV_P2_END, (uint32_t)stages[s2].tile);
        __syncwarp();
        if (lane == 0) mbar_arrive(&ctl.freeb[s2]);
        ++k2;
        if (++s2 == kStages) s2 = 0, par2 ^= 1u;
        wd.reset();
        continue;
      }
    } else if (end) {
      break;
    }
    wd.tick(p.counters + 2);
  }
}

// -------------------------------------------------------------------- scan warp
// Tile-local job count, per-range row offsets and the summary of the job left
// open at the end of the tile; publishes the tile's aggregate descriptor.
template <int TILE>
__device__ __forceinline__ void aggregate_tile(const ClassifyParams &p, StageBuf<TILE> &sb, uint32_t tile,
                                               uint32_t lane) {
  const uint32_t mine = lane < (uint32_t)kWorkerWarps ? sb.range[lane].count : 0u;
  uint32_t incl = mine;
#pragma unroll
  for (int d = 1; d < kWorkerWarps; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if ((int)lane >= d) incl += o;
  }
  if (lane < (uint32_t)kWorkerWarps) sb.range_off[lane] = incl - mine;
  const uint32_t total = __shfl_sync(0xffffffffu, incl, kWorkerWarps - 1);
  if (lane == 0) {
    // the job open at the tile's end: walk the ranges from the right until one holds its first record
    Summ T = summ_identity();
    for (int w = kWorkerWarps - 1; w >= 0; --w) {
      if (sb.range[w].flags & RANGE_CLOSED) { // (everything right of a closed range is empty)
        T = Summ{0u, F_ANCHORED};             // nothing flows into the next tile
        break;
      }
      T = summ_combine(summ_unpack(sb.range[w].suffix), T);
      if (T.f & F_ANCHORED) break;
    }
    sb.local_count = total;
    sb.local_summ = summ_pack(T);
    if (tile == 0) {
      st_desc(p.desc, desc_c(total, DESC_INC, p.gen), desc_s(T, DESC_INC, p.gen));
      if (p.n_tiles == 1) *p.n_jobs = total;
    } else {
      st_desc(p.desc + tile, desc_c(total, DESC_AGG, p.gen), desc_s(T, DESC_AGG, p.gen));
    }
  }
  __syncwarp();
}

struct LookbackState {
  int pos;          // nearest predecessor tile of the next window
  uint32_t excl;    // job rows counted so far
  Summ carry;       // summary gathered so far (nearest tiles combined last)
  bool count_done;  // an inclusive count has been reached
  bool summ_done;   // an anchored summary has been reached (or none is needed)
};

// One non-blocking look-back step.  The warp reads kLookbackWindows x 32 predecessor descriptors
// in one round trip (the loads are independent, so they cost one latency) and consumes them 32 at
// a time, nearest first.  A window is consumed only when every descriptor it needs has been
// published; otherwise the walk stops there and the next poll resumes from it.  Returns true when
// both walks (job count, carried summary) are complete.
// Why several windows per round trip: tiles are published at up to ~100 per microsecond, and a walk
// that advanced only 32 tiles per L2 round trip could never catch up with the newest inclusive tile.
constexpr int kLookbackWindows = 8;
__device__ __forceinline__ bool lookback_poll(const ClassifyParams &p, LookbackState &st, uint32_t lane) {
  uint64_t c[kLookbackWindows], s[kLookbackWindows];
#pragma unroll
  for (int k = 0; k < kLookbackWindows; ++k) {
    const int idx = st.pos - (int)lane - 32 * k;
    c[k] = 0;
    s[k] = 0;
    if (idx >= 0) ld_desc(p.desc + idx, c[k], s[k]);
  }
#pragma unroll
  for (int k = 0; k < kLookbackWindows; ++k) {
    const bool valid = st.pos - (int)lane >= 0;
    const uint32_t cs = desc_c_status(c[k], p.gen), ss = desc_s_status(s[k], p.gen);
    const Summ mine = desc_s_summ(s[k]);
    const uint32_t m_valid = __ballot_sync(0xffffffffu, valid);
    const uint32_t m_cok = __ballot_sync(0xffffffffu, valid && cs != DESC_INVALID);
    const uint32_t m_cinc = __ballot_sync(0xffffffffu, valid && cs == DESC_INC);
    const uint32_t m_sok = __ballot_sync(0xffffffffu, valid && ss != DESC_INVALID);
    const uint32_t m_sanch = __ballot_sync(0xffffffffu, valid && ss != DESC_INVALID && (mine.f & F_ANCHORED));
    const uint32_t need_c = st.count_done ? 0u : (m_cinc ? (2u << lo_bit(m_cinc)) - 1u : m_valid);
    const uint32_t need_s = st.summ_done ? 0u : (m_sanch ? (2u << lo_bit(m_sanch)) - 1u : m_valid);
    if ((need_c & ~m_cok) | (need_s & ~m_sok)) return false;
    if (!st.count_done) {
      uint32_t contrib = ((need_c >> lane) & 1u) ? (uint32_t)c[k] : 0u;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
      st.excl += contrib;
      st.count_done = m_cinc != 0;
    }
    if (!st.summ_done) {
      const uint32_t upto = hi_bit(need_s);
      for (uint32_t j = 0; j <= upto; ++j) st.carry = summ_combine(shfl_summ(mine, (int)j), st.carry);
      st.summ_done = m_sanch != 0;
    }
    if (st.count_done && st.summ_done) return true;
    st.pos -= 32;
  }
  return false;
}

template <int TILE>
__device__ __forceinline__ void scan_loop(const ClassifyParams &p, StageBuf<TILE> *stages, Ctl &ctl, uint32_t lane) {
  uint32_t ka = 0, sa = 0, para = 0; // next tile to aggregate
  uint32_t kl = 0, sl = 0;           // next tile to look back for
  bool end = false, lb_active = false;
  LookbackState st{0, 0u, summ_identity(), false, false};
  Watchdog wd;
  while (true) {
    // Aggregates first: publishing one never waits on another CTA, and every later tile needs it.
    if (!end && __all_sync(0xffffffffu, mbar_test_wait(&ctl.p1done[sa], para))) {
      StageBuf<TILE> &sb = stages[sa];
      const int32_t t = sb.tile;
      if (t < 0) {
        end = true;
      } else {
        if (lane == 0) NXS_TRACE_EVENT(p, EV_AGG_START, (uint32_t)t);
        aggregate_tile<TILE>(p, sb, (uint32_t)t, lane);
        if (lane == 0) NXS_TRACE_EVENT(p, EV_AGG_END, (uint32_t)t);
        ++ka;
        if (++sa == kStages) sa = 0, para ^= 1u;
      }
      wd.reset();
      continue;
    }
    if (kl < ka) { // a tile waits for its look-back: poll its predecessors (each poll is one L2 round trip)
      StageBuf<TILE> &sb = stages[sl];
      const uint32_t tile = (uint32_t)sb.tile;
      if (!lb_active) {
        st.pos = (int)tile - 1;
        st.excl = 0;
        st.carry = summ_identity();
        st.count_done = tile == 0;
        st.summ_done = tile == 0 || (sb.range[0].flags & RANGE_FIRST_IS_HEAD); // nothing to carry
        lb_active = true;
        if (lane == 0) NXS_TRACE_EVENT(p, EV_LB_START, tile);
      }
      if (lane == 0) NXS_TRACE_EVENT(p, EV_LB_POLL, tile);
      if ((st.count_done && st.summ_done) || lookback_poll(p, st, lane)) {
        if (lane == 0) {
          const Summ T = summ_unpack(sb.local_summ);
          const uint32_t total = sb.local_count;
          if (tile != 0) {
            const Summ I = (T.f & F_ANCHORED) ? T : summ_combine(st.carry, T);
            st_desc(p.desc + tile, desc_c(st.excl + total, DESC_INC, p.gen), desc_s(I, DESC_INC, p.gen));
            if (tile == p.n_tiles - 1) *p.n_jobs = (uint64_t)st.excl + total;
          }
          sb.tile_base = st.excl;
          sb.carry = summ_pack(st.carry);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&ctl.lbdone[sl]);
        if (lane == 0) NXS_TRACE_EVENT(p, EV_LB_END, tile);
        lb_active = false;
        ++kl;
        if (++sl == kStages) sl = 0;
        wd.reset();
        continue;
      }
    } else if (end) {
      break;
    } else { // nothing pending: sleep until the workers finish phase 1 of the next tile
      (void)__all_sync(0xffffffffu, mbar_try_wait(&ctl.p1done[sa], para, 2000u));
    }
    wd.tick(p.counters + 2);
  }
}

template <int TILE>
__global__ void __launch_bounds__(kThreads, 2) classify_kernel(const __grid_constant__ ClassifyParams p) {
  // (no manual re-alignment: keeping the pointer derived from the array keeps it in the shared
  // address space, so accesses compile to LDS/STS rather than generic loads)
  extern __shared__ __align__(128) uint8_t smem_raw[];
  StageBuf<TILE> *stages = reinterpret_cast<StageBuf<TILE> *>(smem_raw);
  Ctl &ctl = *reinterpret_cast<Ctl *>(smem_raw + sizeof(StageBuf<TILE>) * kStages);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&ctl.full[s], 2);
      mbar_init(&ctl.p1done[s], kWorkerWarps);
      mbar_init(&ctl.lbdone[s], 1);
      mbar_init(&ctl.freeb[s], kWorkerWarps);
    }
    mbar_fence_init();
  }
  if (tid < 128) {
    ctl.lut[tid] = p.lut[tid];
    const uint32_t fl = tid & 7u, st = tid >> 3;
    ctl.lut2[tid] = (uint16_t)(256u * ((fl & 3u) != 3u ? 0u : !(fl & NXS_FLAG_CHECKPOINT) ? 1u : ((p.finished_mask >> st) & 1u) ? 2u : 3u));
  }
  for (uint32_t e = tid; e < 1024; e += kThreads) {
    const uint32_t g = e >> 8, code = e & 255u, a = code & 7u, msg = code >> 5;
    uint32_t v = 0;
    if (g != 0 && a >= 1 && a <= 4) {
      const uint32_t fold = VT_HE | (((p.term_mask >> a) & 1u) ? VT_TERM : 0u) | (a != NXS_ACTION_RUNNING ? VT_FAIL : 0u);
      if (g == 1) v = a | msg << 9 | 1u << 12 | NXS_ERR_NO_CHECKPOINT << 15 | VT_ERR; // :265-273
      else if (g == 2) v = a | msg << 9 | 1u << 14 | fold;                             // :276-279
      else v = applied_verdict(a) | msg << 9 | fold;                                    // :283-370
    }
    ctl.vt[e] = v;
  }
  __syncthreads(); // the only CTA-wide barrier

  if (warp < (uint32_t)kWorkerWarps) worker_loop<TILE>(p, stages, ctl, warp, lane);
  else if (warp == (uint32_t)kProducerWarp) {
    if (lane == 0) producer_loop<TILE>(p, stages, ctl);
  } else scan_loop<TILE>(p, stages, ctl, lane);
}

#endif // __CUDACC__

} // namespace nxs
