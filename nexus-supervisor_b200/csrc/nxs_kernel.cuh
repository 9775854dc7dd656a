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
//   worker warps   phase 1 as soon as a tile has landed: lane-per-record LUT
//                  lookup, guards, packed verdict -> global; warp ballots ->
//                  per-32-record bit masks in shared memory
//                  phase 2 once the tile's look-back is done: lanes holding the
//                  LAST record of a job resolve its masks into the 16-byte row
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

struct __align__(16) ChunkMasks {
  uint32_t head, tail, he, term; // bit b = record 32*chunk + b of the tile
  uint32_t fail, err, pad0, pad1;
};

template <int TILE> struct __align__(128) StageBuf {
  // keyx[1] = key of the record before the tile, keyx[2..TILE+1] = the tile's
  // keys (16-byte aligned for the bulk copy), keyx[TILE+2] = key after the tile
  uint64_t keyx[TILE + 4];
  uint16_t reason[TILE];
  uint8_t kind[TILE];
  uint8_t flags[TILE];
  uint8_t stage[TILE];
  uint8_t code[TILE]; // LUT code per record (phase 1 -> phase 2)
  ChunkMasks masks[TILE / 32];
  uint32_t chunk_off[TILE / 32]; // jobs ending before each chunk, tile-local
  uint64_t carry;                // summary carried into the tile's first job
  uint64_t local_summ;           // summary of the job left open at the tile's end
  uint32_t local_count;          // jobs ending in the tile
  uint32_t tile_base;            // job rows before the tile
  int32_t tile;                  // tile id, -1 = end of work
};

struct Ctl {
  uint64_t full[kStages];   // tile landed            (producer -> workers)
  uint64_t p1done[kStages]; // phase 1 complete       (workers -> scan warp)
  uint64_t lbdone[kStages]; // look-back complete     (scan warp -> workers)
  uint64_t freeb[kStages];  // phase 2 complete       (workers -> producer)
  uint32_t appl[8];         // applied_verdict(action)
  uint8_t lut[128];         // (kind, reason) -> code
};

// Tile descriptor: two self-validating 64-bit words, each published with one
// relaxed store, so no fence is needed on either side.
//   c: count[0,32) | status[32,34) | generation[34,43)
//   s: win_idx[0,32) | Summ.f[32,53) | status[53,55) | generation[55,64)
struct __align__(16) TileDesc {
  uint64_t c, s;
};
enum : uint32_t { DESC_INVALID = 0, DESC_AGG = 1, DESC_INC = 2 };
constexpr uint32_t kGenMax = 511;
__host__ __device__ __forceinline__ uint64_t desc_c(uint32_t count, uint32_t status, uint32_t gen) {
  return (uint64_t)count | (uint64_t)status << 32 | (uint64_t)gen << 34;
}
__host__ __device__ __forceinline__ uint64_t desc_s(Summ sm, uint32_t status, uint32_t gen) {
  return (uint64_t)sm.win_idx | (uint64_t)(sm.f & 0x1FFFFFu) << 32 | (uint64_t)status << 53 | (uint64_t)gen << 55;
}
__host__ __device__ __forceinline__ uint32_t desc_c_status(uint64_t c, uint32_t gen) {
  return ((uint32_t)(c >> 34) & 511u) == gen ? (uint32_t)(c >> 32) & 3u : DESC_INVALID;
}
__host__ __device__ __forceinline__ uint32_t desc_s_status(uint64_t s, uint32_t gen) {
  return (uint32_t)(s >> 55) == gen ? (uint32_t)(s >> 53) & 3u : DESC_INVALID;
}
__host__ __device__ __forceinline__ Summ desc_s_summ(uint64_t s) {
  return Summ{(uint32_t)s, (uint32_t)(s >> 32) & 0x1FFFFFu};
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
};

template <int TILE> constexpr size_t classify_smem_bytes() {
  return sizeof(StageBuf<TILE>) * kStages + sizeof(Ctl) + 128;
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

struct Watchdog {
  uint64_t t0;
  uint32_t spins;
  __device__ __forceinline__ Watchdog() : t0(0), spins(0) {}
  __device__ __forceinline__ void tick(uint32_t *flag) {
    if ((++spins & 4095u) == 0) {
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
      while (!mbar_try_wait(&ctl.freeb[s], (use - 1) & 1u)) wd.tick(p.counters + 2);
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
template <int TILE, bool FULL>
__device__ __forceinline__ void phase1(const ClassifyParams &p, StageBuf<TILE> &sb, const Ctl &ctl, uint32_t tile,
                                       uint32_t warp, uint32_t lane) {
  constexpr int CHUNKS = TILE / 32;
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
  const bool first_tile = start == 0;
  const bool last_tile = start + cnt == p.n;
  const uint32_t fin = p.finished_mask, tmask = p.term_mask;
#pragma unroll 2
  for (uint32_t c = warp; c < (uint32_t)CHUNKS; c += kWorkerWarps) {
    const uint32_t i = c * 32 + lane;
    bool head = false, tail = false, he = false, term = false, fail = false, err = false;
    if (FULL || i < cnt) {
      const uint32_t kind = sb.kind[i], reason = sb.reason[i], fl = sb.flags[i];
      const uint32_t st = sb.stage[i] & 15u;
      const uint64_t key = sb.keyx[2 + i], pk = sb.keyx[1 + i], nk = sb.keyx[3 + i];
      head = (key != pk) || (first_tile && i == 0);
      tail = (key != nk) || (last_tile && i == cnt - 1);
      // onEvent: guards (:155,:161,:218) then the (kind, reason) switch arms
      const uint32_t idx = min(kind, 3u) * 32u + (reason < 32u ? reason : 0u);
      const uint32_t code = ((fl & 3u) == 3u) ? ctl.lut[idx] : 0u;
      const uint32_t a = code & 7u;
      // superviseAction: no checkpoint (:265) -> finished (:276) -> switch (:283)
      const bool exists = (fl & NXS_FLAG_CHECKPOINT) != 0;
      const bool finished = (fin >> st) & 1u;
      const uint32_t base = exists ? (finished ? (a | 1u << 14) : ctl.appl[a])
                                   : (a | 1u << 12 | NXS_ERR_NO_CHECKPOINT << 15);
      const uint32_t v = a ? (base | (code >> 3) << 9) : 0u;
      __stcs(p.verdict + start + i, v);
      sb.code[i] = (uint8_t)code;
      he = a != 0 && exists;
      err = a != 0 && !exists;
      term = he && ((tmask >> a) & 1u);
      fail = he && a != NXS_ACTION_RUNNING;
    }
    ChunkMasks m;
    m.head = __ballot_sync(0xffffffffu, head);
    m.tail = __ballot_sync(0xffffffffu, tail);
    m.he = __ballot_sync(0xffffffffu, he);
    m.term = __ballot_sync(0xffffffffu, term);
    m.fail = __ballot_sync(0xffffffffu, fail);
    m.err = __ballot_sync(0xffffffffu, err);
    m.pad0 = m.pad1 = 0;
    if (lane == 0) sb.masks[c] = m;
  }
}

template <int TILE>
__device__ __forceinline__ void phase2(const ClassifyParams &p, const StageBuf<TILE> &sb, uint32_t warp,
                                       uint32_t lane) {
  constexpr int CHUNKS = TILE / 32;
  const uint32_t start = (uint32_t)sb.tile * TILE;
  const uint32_t tile_base = sb.tile_base;
  const Summ carry = summ_unpack(sb.carry);
  const uint32_t le = (2u << lane) - 1u;
  for (uint32_t c = warp; c < (uint32_t)CHUNKS; c += kWorkerWarps) {
    const ChunkMasks m = sb.masks[c];
    if (!((m.tail >> lane) & 1u)) continue;
    const uint32_t pos = tile_base + sb.chunk_off[c] + __popc(m.tail & (le >> 1));
    // first record of this job: nearest head at or before this lane
    uint32_t cc = c, hm = m.head & le;
    while (hm == 0 && cc > 0) {
      --cc;
      hm = sb.masks[cc].head;
    }
    uint32_t hbit = 0;
    Summ S = carry; // no head in the tile before this record: the job began earlier
    if (hm) {
      hbit = hi_bit(hm);
      S = summ_anchor(sb.stage[cc * 32 + hbit]);
    }
    for (uint32_t k = cc; k <= c; ++k) {
      uint32_t mk = ~0u;
      if (k == cc) mk <<= hbit;
      if (k == c) mk &= le;
      const ChunkMasks w = sb.masks[k];
      S = summ_combine(S, summ_piece(w.he & mk, w.term & mk, w.fail & mk, w.err & mk, sb.code, k * 32,
                                     p.record_base + start));
    }
    uint32_t verdict, winner;
    summ_resolve(S, p.finished_mask, verdict, winner);
    const uint64_t key = sb.keyx[2 + c * 32 + lane];
    __stcs(reinterpret_cast<uint4 *>(p.rows + pos),
           make_uint4((uint32_t)key, (uint32_t)(key >> 32), verdict, winner));
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
    bool progressed = false;
    if (!end && __all_sync(0xffffffffu, mbar_test_wait(&ctl.full[s1], par1))) {
      StageBuf<TILE> &sb = stages[s1];
      const int32_t t = sb.tile;
      if (t < 0) {
        end = true;
      } else {
        if (p.n - (uint32_t)t * TILE >= (uint32_t)TILE) phase1<TILE, true>(p, sb, ctl, (uint32_t)t, warp, lane);
        else phase1<TILE, false>(p, sb, ctl, (uint32_t)t, warp, lane);
        ++k1;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl.p1done[s1]);
      if (++s1 == kStages) s1 = 0, par1 ^= 1u;
      progressed = true;
    }
    if (!progressed && k2 < k1 && __all_sync(0xffffffffu, mbar_test_wait(&ctl.lbdone[s2], par2))) {
      phase2<TILE>(p, stages[s2], warp, lane);
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl.freeb[s2]);
      ++k2;
      if (++s2 == kStages) s2 = 0, par2 ^= 1u;
      progressed = true;
    }
    if (end && k2 == k1) break;
    if (progressed) wd.reset();
    else wd.tick(p.counters + 2);
  }
}

// -------------------------------------------------------------------- scan warp
__device__ __forceinline__ Summ shfl_summ(Summ s, int src) {
  return Summ{__shfl_sync(0xffffffffu, s.win_idx, src), __shfl_sync(0xffffffffu, s.f, src)};
}
__device__ __forceinline__ Summ shfl_down_summ(Summ s, int d) {
  return Summ{__shfl_down_sync(0xffffffffu, s.win_idx, d), __shfl_down_sync(0xffffffffu, s.f, d)};
}

// Tile-local job count, per-chunk row offsets and the summary of the job left
// open at the end of the tile; publishes the tile's aggregate descriptor.
template <int TILE>
__device__ __forceinline__ void aggregate_tile(const ClassifyParams &p, StageBuf<TILE> &sb, uint32_t tile,
                                               uint32_t lane) {
  constexpr int CHUNKS = TILE / 32;
  constexpr int CPL = CHUNKS / 32; // chunks per lane
  static_assert(CHUNKS % 32 == 0, "TILE must be a multiple of 1024");
  const uint32_t start = tile * TILE;
  const uint32_t cnt = min((uint32_t)TILE, p.n - start);

  uint32_t tails[CPL], running = 0;
  int last_head = -1;
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const uint32_t c = lane * CPL + j;
    const uint32_t t = sb.masks[c].tail, h = sb.masks[c].head;
    tails[j] = __popc(t);
    running += tails[j];
    if (h) last_head = (int)(c * 32 + hi_bit(h));
  }
  uint32_t incl = running;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if ((int)lane >= d) incl += o;
  }
  uint32_t off = incl - running;
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    sb.chunk_off[lane * CPL + j] = off;
    off += tails[j];
  }
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) last_head = max(last_head, __shfl_xor_sync(0xffffffffu, last_head, d));

  const uint32_t last = cnt - 1;
  const bool closed = (sb.masks[last >> 5].tail >> (last & 31)) & 1u;
  Summ T{0u, F_ANCHORED}; // closed: nothing flows into the next tile
  if (!closed) {          // warp-uniform
    const uint32_t from = last_head < 0 ? 0u : (uint32_t)last_head;
    const uint32_t c_from = from >> 5, c_last = last >> 5;
    Summ S = summ_identity();
    if (c_last - c_from < 2) { // usual case: the open job began in the last two chunks; every lane computes it
      for (uint32_t c = c_from; c <= c_last; ++c) {
        uint32_t mk = ~0u;
        if (c == c_from) mk <<= (from & 31);
        const ChunkMasks w = sb.masks[c];
        S = summ_combine(S, summ_piece(w.he & mk, w.term & mk, w.fail & mk, w.err & mk, sb.code, c * 32,
                                       p.record_base + start));
      }
    } else { // long open job: pieces per lane, then an ordered warp reduction
#pragma unroll
      for (int j = 0; j < CPL; ++j) {
        const uint32_t c = lane * CPL + j;
        if (c < c_from) continue;
        uint32_t mk = ~0u;
        if (c == c_from) mk <<= (from & 31);
        const ChunkMasks w = sb.masks[c];
        S = summ_combine(S, summ_piece(w.he & mk, w.term & mk, w.fail & mk, w.err & mk, sb.code, c * 32,
                                       p.record_base + start));
      }
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const Summ o = shfl_down_summ(S, d);
        if ((int)lane + d < 32) S = summ_combine(S, o);
      }
      S = shfl_summ(S, 0);
    }
    T = last_head < 0 ? S : summ_combine(summ_anchor(sb.stage[from]), S);
  }
  if (lane == 0) {
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

// One non-blocking look-back step over up to 32 predecessor descriptors.
// Returns true when both walks are complete.  A window is consumed only when
// every descriptor it needs has been published; otherwise nothing changes.
__device__ __forceinline__ bool lookback_poll(const ClassifyParams &p, LookbackState &st, uint32_t lane) {
  const int idx = st.pos - (int)lane;
  const bool valid = idx >= 0;
  uint64_t c = 0, s = 0;
  if (valid) ld_desc(p.desc + idx, c, s);
  const uint32_t cs = desc_c_status(c, p.gen), ss = desc_s_status(s, p.gen);
  const Summ mine = desc_s_summ(s);
  const uint32_t m_valid = __ballot_sync(0xffffffffu, valid);
  const uint32_t m_cok = __ballot_sync(0xffffffffu, valid && cs != DESC_INVALID);
  const uint32_t m_cinc = __ballot_sync(0xffffffffu, valid && cs == DESC_INC);
  const uint32_t m_sok = __ballot_sync(0xffffffffu, valid && ss != DESC_INVALID);
  const uint32_t m_sanch = __ballot_sync(0xffffffffu, valid && ss != DESC_INVALID && (mine.f & F_ANCHORED));
  const uint32_t need_c = st.count_done ? 0u : (m_cinc ? (2u << lo_bit(m_cinc)) - 1u : m_valid);
  const uint32_t need_s = st.summ_done ? 0u : (m_sanch ? (2u << lo_bit(m_sanch)) - 1u : m_valid);
  if ((need_c & ~m_cok) | (need_s & ~m_sok)) return false;
  if (!st.count_done) {
    uint32_t contrib = ((need_c >> lane) & 1u) ? (uint32_t)c : 0u;
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
    bool progressed = false;
    if (!end && __all_sync(0xffffffffu, mbar_test_wait(&ctl.p1done[sa], para))) {
      StageBuf<TILE> &sb = stages[sa];
      const int32_t t = sb.tile;
      if (t < 0) {
        end = true;
      } else {
        aggregate_tile<TILE>(p, sb, (uint32_t)t, lane);
        ++ka;
        if (++sa == kStages) sa = 0, para ^= 1u;
      }
      progressed = true;
    } else if (kl < ka) {
      StageBuf<TILE> &sb = stages[sl];
      const uint32_t tile = (uint32_t)sb.tile;
      if (!lb_active) {
        st.pos = (int)tile - 1;
        st.excl = 0;
        st.carry = summ_identity();
        st.count_done = tile == 0;
        st.summ_done = tile == 0 || (sb.masks[0].head & 1u); // first record starts a job: nothing to carry
        lb_active = true;
      }
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
        lb_active = false;
        ++kl;
        if (++sl == kStages) sl = 0;
        progressed = true;
      }
    }
    if (end && kl == ka) break;
    if (progressed) wd.reset();
    else wd.tick(p.counters + 2);
  }
}

template <int TILE>
__global__ void __launch_bounds__(kThreads) classify_kernel(const __grid_constant__ ClassifyParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *aligned = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  StageBuf<TILE> *stages = reinterpret_cast<StageBuf<TILE> *>(aligned);
  Ctl &ctl = *reinterpret_cast<Ctl *>(aligned + sizeof(StageBuf<TILE>) * kStages);
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
  if (tid < 128) ctl.lut[tid] = p.lut[tid];
  if (tid < 8) ctl.appl[tid] = (tid >= 1 && tid <= 4) ? applied_verdict(tid) : 0u;
  __syncthreads(); // the only CTA-wide barrier

  if (warp < (uint32_t)kWorkerWarps) worker_loop<TILE>(p, stages, ctl, warp, lane);
  else if (warp == (uint32_t)kProducerWarp) {
    if (lane == 0) producer_loop<TILE>(p, stages, ctl);
  } else scan_loop<TILE>(p, stages, ctl, lane);
}

#endif // __CUDACC__

} // namespace nxs
