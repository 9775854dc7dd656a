// nxs_kernel.cuh — the classification kernel (sm_100a).
//
// One persistent grid streams the five input columns through shared memory with
// 1-D bulk async copies (TMA engine) on an mbarrier ring, classifies every record
// (onEvent + superviseAction decision, services/supervisor.go:137-374) and folds
// the records of each job into one row, in a single pass over HBM:
//   phase 1  lane-per-record: LUT lookup, guards, packed verdict -> global;
//            warp ballots -> per-32-record bit masks in shared memory
//   aggregate/look-back (warp 0): job count + open-job summary of the tile are
//            published to a tile descriptor; predecessors are read back
//            (decoupled look-back) for the row offset and the carried-in summary
//   phase 2  lanes that hold the LAST record of a job resolve its masks and
//            store the 16-byte job row
// Scratch: TileDesc[n_tiles] + 2 counters (zero between launches).
#pragma once
#include "nxs_device.cuh"

namespace nxs {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxStages = 3;

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
};

struct Ctl {
  uint64_t full_bar[kMaxStages];
  uint64_t local_summ[kMaxStages]; // open-job summary of the tile in each stage
  uint64_t carry;                  // summary carried into the head tile
  int32_t tile[kMaxStages];        // tile id per stage, -1 = none
  uint32_t landed[kMaxStages];
  uint32_t local_count[kMaxStages];
  uint32_t tile_base; // job rows before the head tile
  uint32_t appl[8];   // applied_verdict(action)
  uint8_t lut[128];   // (kind, reason) -> code
};

struct __align__(32) TileDesc {
  uint64_t state; // count[0,32) | status[32,34) | generation[34,64)
  uint64_t agg;   // open-job summary of this tile alone
  uint64_t inc;   // ... including everything carried in
  uint64_t pad;
};
enum : uint32_t { DESC_INVALID = 0, DESC_AGG = 1, DESC_INC = 2 };
__host__ __device__ __forceinline__ uint64_t desc_state(uint32_t count, uint32_t status, uint32_t gen) {
  return (uint64_t)count | (uint64_t)status << 32 | (uint64_t)gen << 34;
}
constexpr uint32_t kGenMax = (1u << 30) - 1;

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
  uint32_t *counters; // [0] next tile id, [1] finished CTAs, [2] error flag
  uint32_t n;
  uint32_t n_tiles;
  uint32_t record_base; // added to winner_idx
  uint32_t finished_mask;
  uint32_t term_mask; // bit a: new_stage_of(a) is a finished stage
  uint32_t gen;
  uint32_t depth; // ring depth, 1..kMaxStages
  uint8_t lut[128];
};

template <int TILE> constexpr size_t classify_smem_bytes() {
  return sizeof(StageBuf<TILE>) * kMaxStages + sizeof(Ctl) + 128;
}

// --------------------------------------------------------------------------
template <int TILE>
__device__ __forceinline__ void issue_tile(const ClassifyParams &p, StageBuf<TILE> &sb, Ctl &ctl, int s,
                                           uint32_t id) {
  if (id >= p.n_tiles) {
    ctl.tile[s] = -1;
    return;
  }
  ctl.tile[s] = (int32_t)id;
  const uint32_t start = id * TILE;
  uint64_t *bar = &ctl.full_bar[s];
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
    mbar_arrive_n(bar, 2); // ragged last tile: loaded by all threads in phase 1
  }
}

template <int TILE>
__device__ __forceinline__ void phase1(const ClassifyParams &p, StageBuf<TILE> &sb, const Ctl &ctl,
                                       uint32_t tile) {
  constexpr int CHUNKS = TILE / 32;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t start = tile * TILE;
  const uint32_t cnt = min((uint32_t)TILE, p.n - start);
  if (cnt < (uint32_t)TILE) { // ragged tile (block-uniform)
    for (uint32_t i = threadIdx.x; i < cnt; i += kThreads) {
      sb.keyx[2 + i] = p.key[start + i];
      sb.reason[i] = p.reason[start + i];
      sb.kind[i] = p.kind[start + i];
      sb.flags[i] = p.flags[start + i];
      sb.stage[i] = p.stage[start + i];
    }
    if (threadIdx.x == 0 && start > 0) sb.keyx[1] = p.key[start - 1];
    __syncthreads();
  }
  const bool first_tile = start == 0;
  const bool last_tile = start + cnt == p.n;
  const uint32_t fin = p.finished_mask, tmask = p.term_mask;
#pragma unroll 2
  for (uint32_t c = warp; c < (uint32_t)CHUNKS; c += kWarps) {
    const uint32_t i = c * 32 + lane;
    bool head = false, tail = false, he = false, term = false, fail = false, err = false;
    if (i < cnt) {
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

__device__ __forceinline__ Summ shfl_summ(Summ s, int src) {
  return Summ{__shfl_sync(0xffffffffu, s.win_idx, src), __shfl_sync(0xffffffffu, s.f, src)};
}
__device__ __forceinline__ Summ shfl_down_summ(Summ s, int d) {
  return Summ{__shfl_down_sync(0xffffffffu, s.win_idx, d), __shfl_down_sync(0xffffffffu, s.f, d)};
}

// Warp 0: tile-local job count, per-chunk row offsets and the summary of the job
// left open at the end of the tile; publishes the tile's aggregate descriptor.
template <int TILE>
__device__ __forceinline__ void aggregate_tile(const ClassifyParams &p, StageBuf<TILE> &sb, Ctl &ctl, int s,
                                               uint32_t tile) {
  constexpr int CHUNKS = TILE / 32;
  constexpr int CPL = CHUNKS / 32; // chunks per lane
  static_assert(CHUNKS % 32 == 0, "TILE must be a multiple of 1024");
  const uint32_t lane = threadIdx.x & 31;
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
    Summ S = summ_identity();
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      const uint32_t c = lane * CPL + j;
      if (c * 32 + 31 < from) continue;
      uint32_t mk = ~0u;
      if (c == (from >> 5)) mk <<= (from & 31);
      const ChunkMasks w = sb.masks[c];
      S = summ_combine(S, summ_piece(w.he & mk, w.term & mk, w.fail & mk, w.err & mk, sb.code, c * 32,
                                     p.record_base + start));
    }
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { // ordered reduction, lane 0 ends with the whole run
      const Summ o = shfl_down_summ(S, d);
      if ((int)lane + d < 32) S = summ_combine(S, o);
    }
    T = last_head < 0 ? S : summ_combine(summ_anchor(sb.stage[from]), S);
  }
  if (lane == 0) {
    ctl.local_count[s] = total;
    ctl.local_summ[s] = summ_pack(T);
    TileDesc *d = p.desc + tile;
    if (tile == 0) {
      st_relaxed(&d->inc, summ_pack(T));
      st_release(&d->state, desc_state(total, DESC_INC, p.gen));
      if (p.n_tiles == 1) *p.n_jobs = total;
    } else {
      st_relaxed(&d->agg, summ_pack(T));
      st_release(&d->state, desc_state(total, DESC_AGG, p.gen));
    }
  }
}

// Warp 0: decoupled look-back over predecessor tiles.  Leaves the number of job
// rows before this tile and the summary carried into its first job in ctl, and
// publishes the tile's inclusive descriptor.
template <int TILE>
__device__ __forceinline__ void lookback_tile(const ClassifyParams &p, const StageBuf<TILE> &sb, Ctl &ctl,
                                              int s, uint32_t tile) {
  const uint32_t lane = threadIdx.x & 31;
  if (tile == 0) {
    if (lane == 0) {
      ctl.tile_base = 0;
      ctl.carry = summ_pack(summ_identity());
    }
    return;
  }
  const uint32_t total = ctl.local_count[s];
  const Summ T = summ_unpack(ctl.local_summ[s]);
  bool need = !(sb.masks[0].head & 1u); // first record continues a job of an earlier tile
  uint32_t excl = 0;
  Summ carry = summ_identity();
  int pos = (int)tile - 1;
  const uint64_t t0 = global_timer_ns();
  while (true) {
    const int idx = pos - (int)lane;
    const bool valid = idx >= 0;
    uint64_t st = 0;
    if (valid) {
      const uint64_t *sp = &p.desc[idx].state;
      uint32_t spins = 0;
      while (true) {
        st = ld_acquire(sp);
        if ((uint32_t)(st >> 34) == p.gen && ((st >> 32) & 3u) != DESC_INVALID) break;
        if ((++spins & 1023u) == 0 && global_timer_ns() - t0 > 4000000000ull) {
          atomicExch(p.counters + 2, 1u); // watchdog: never hang the GPU
          __trap();
        }
      }
    }
    const uint32_t status = valid ? (uint32_t)(st >> 32) & 3u : DESC_INVALID;
    const uint32_t inc_mask = __ballot_sync(0xffffffffu, status == DESC_INC);
    const uint32_t first_inc = inc_mask ? lo_bit(inc_mask) : 32u;
    uint32_t contrib = (valid && lane <= first_inc) ? (uint32_t)st : 0u;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
    excl += contrib;
    if (need) { // warp-uniform
      Summ mine = summ_identity();
      if (valid && lane <= first_inc)
        mine = summ_unpack(ld_relaxed(status == DESC_INC ? &p.desc[idx].inc : &p.desc[idx].agg));
      const uint32_t upto = min(first_inc, 31u);
      for (uint32_t j = 0; j <= upto && need; ++j) {
        const Summ sj = shfl_summ(mine, (int)j);
        carry = summ_combine(sj, carry);
        if (sj.f & F_ANCHORED) need = false;
      }
    }
    if (inc_mask) break;
    pos -= 32;
  }
  if (lane == 0) {
    const Summ I = (T.f & F_ANCHORED) ? T : summ_combine(carry, T);
    TileDesc *d = p.desc + tile;
    st_relaxed(&d->inc, summ_pack(I));
    st_release(&d->state, desc_state(excl + total, DESC_INC, p.gen));
    ctl.tile_base = excl;
    ctl.carry = summ_pack(carry);
    if (tile == p.n_tiles - 1) *p.n_jobs = (uint64_t)excl + total;
  }
}

template <int TILE>
__device__ __forceinline__ void phase2(const ClassifyParams &p, const StageBuf<TILE> &sb, const Ctl &ctl,
                                       uint32_t tile) {
  constexpr int CHUNKS = TILE / 32;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t start = tile * TILE;
  const uint32_t tile_base = ctl.tile_base;
  const Summ carry = summ_unpack(ctl.carry);
  const uint32_t le = (2u << lane) - 1u;
  for (uint32_t c = warp; c < (uint32_t)CHUNKS; c += kWarps) {
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
__global__ void __launch_bounds__(kThreads) classify_kernel(const __grid_constant__ ClassifyParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *aligned = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  StageBuf<TILE> *stages = reinterpret_cast<StageBuf<TILE> *>(aligned);
  Ctl &ctl = *reinterpret_cast<Ctl *>(aligned + sizeof(StageBuf<TILE>) * kMaxStages);
  const uint32_t tid = threadIdx.x, warp = tid >> 5;
  const int D = (int)p.depth;

  if (tid == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(&ctl.full_bar[s], 2);
      ctl.tile[s] = -1;
      ctl.landed[s] = 0;
    }
    mbar_fence_init();
  }
  if (tid < 128) ctl.lut[tid] = p.lut[tid];
  if (tid < 8) ctl.appl[tid] = (tid >= 1 && tid <= 4) ? applied_verdict(tid) : 0u;
  __syncthreads();

  uint32_t next_id = 0xFFFFFFFFu; // thread 0: tile id fetched one iteration ahead
  if (tid == 0) {
    const uint32_t first = atomicAdd(p.counters, (uint32_t)D);
    for (int s = 0; s < D; ++s) issue_tile<TILE>(p, stages[s], ctl, s, first + s);
    if (first + D < p.n_tiles) next_id = atomicAdd(p.counters, 1u);
  }

  uint32_t parity = 0, p1done = 0;
  int head = 0;
  while (true) {
    if (tid == 0) {
      for (int s = 0; s < D; ++s)
        if (s != head && ctl.tile[s] >= 0 && !((p1done >> s) & 1u))
          ctl.landed[s] = mbar_test_wait(&ctl.full_bar[s], (parity >> s) & 1u);
    }
    __syncthreads();
    const int t = ctl.tile[head];
    if (t < 0) break;
    uint32_t newly = 0;
    for (int k = 0; k < D; ++k) {
      int s = head + k;
      if (s >= D) s -= D;
      const int ts = ctl.tile[s];
      if (ts < 0 || ((p1done >> s) & 1u)) continue;
      if (k > 0 && !ctl.landed[s]) continue;
      mbar_wait(&ctl.full_bar[s], (parity >> s) & 1u);
      parity ^= 1u << s;
      phase1<TILE>(p, stages[s], ctl, (uint32_t)ts);
      p1done |= 1u << s;
      newly |= 1u << s;
    }
    __syncthreads();
    if (warp == 0) {
      for (int k = 0; k < D; ++k) {
        int s = head + k;
        if (s >= D) s -= D;
        if ((newly >> s) & 1u) aggregate_tile<TILE>(p, stages[s], ctl, s, (uint32_t)ctl.tile[s]);
      }
      __syncwarp();
      lookback_tile<TILE>(p, stages[head], ctl, head, (uint32_t)t);
    }
    __syncthreads();
    phase2<TILE>(p, stages[head], ctl, (uint32_t)t);
    p1done &= ~(1u << head);
    __syncthreads();
    if (tid == 0) {
      issue_tile<TILE>(p, stages[head], ctl, head, next_id);
      next_id = (next_id < p.n_tiles) ? atomicAdd(p.counters, 1u) : 0xFFFFFFFFu;
    }
    head = head + 1 == D ? 0 : head + 1;
  }

  // leave the counters zero for the next launch
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(p.counters + 1, 1u) == gridDim.x - 1) {
      p.counters[0] = 0;
      p.counters[1] = 0;
      __threadfence();
    }
  }
}

} // namespace nxs
