// nxs_device.cuh — device-side building blocks of the classification kernel:
// PTX wrappers (mbarrier, bulk async copy, acquire/release), the packed verdict
// arithmetic and the per-job fold algebra.  sm_100a only.
//
// The fold algebra restates, as an associative operator over runs of records of
// one job, what the reference does one event at a time in superviseAction
// (services/supervisor.go:261-374): read the row, return if IsFinished(), else
// apply the action's stage / cause / details.  See DESIGN.md "Per-job fold".
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "nxs_classify.h"

namespace nxs {

// ------------------------------------------------------------------ PTX glue
#if defined(__CUDACC__)
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// Shared-space address computed ONCE: the volatile asm keeps the compiler from re-deriving it (three
// instructions: CTA id in cluster, shift, add) at every use inside hot or idle loops.
__device__ __forceinline__ uint32_t smem_u32_pinned(const void *p) {
  uint32_t a;
  asm volatile("{\n\t.reg .u64 t;\n\tcvta.to.shared.u64 t, %1;\n\tcvt.u32.u64 %0, t;\n\t}" : "=r"(a) : "l"(p));
  return a;
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_n(uint64_t *bar, uint32_t n) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n) : "memory");
}
// Arrival that fires once every cp.async this thread issued so far has landed.
__device__ __forceinline__ void mbar_arrive_after_cp_async(uint64_t *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Suspends the thread until the phase completes or about `hint_ns` have passed.
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity, uint32_t hint_ns = 20000u) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(hint_ns)
      : "memory");
  return ok != 0;
}
// 32-bit unsigned min as one instruction (the compiler otherwise builds 16-bit pair tricks around byte loads).
__device__ __forceinline__ uint32_t umin32(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("min.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// Warp ballot of one bit of a word (one LOP3 + one VOTE).
__device__ __forceinline__ uint32_t ballot_bit(uint32_t word, uint32_t bit) {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
      "and.b32 t, %1, %2;\n\t"
      "setp.ne.u32 p, t, 0;\n\t"
      "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t}"
      : "=r"(r)
      : "r"(word), "r"(bit));
  return r;
}
__device__ __forceinline__ bool mbar_test_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Same on a precomputed shared-space address (saves the generic->shared conversion in hot loops).
__device__ __forceinline__ bool mbar_test_wait_a(uint32_t bar_addr, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar_addr), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_try_wait_a(uint32_t bar_addr, uint32_t parity, uint32_t hint_ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar_addr), "r"(parity), "r"(hint_ns)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// 1-D bulk async copy global -> shared (TMA engine; SASS UBLKCP).  dst, src and
// bytes are multiples of 16.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// Asks the TMA engine to bring [src, src + bytes) into L2 (multiples of 16); nothing waits for it.
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_8(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_relaxed(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(uint64_t *p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release(uint64_t *p, uint64_t v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Loads from tables in shared memory that do not change after start-up, by 32-bit shared address
// (the address arithmetic is then exactly what is written: one instruction per table index).
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

#endif // __CUDACC__

// ------------------------------------------------------- verdict arithmetic
// Code of a switch arm of onEvent = the low 12 bits of the verdict an applied record of that arm
// gets: action | new_stage << 3 | cause_tpl << 7 | msg_tpl << 9; 0 = `default`.
// applier switch (services/supervisor.go:283-370), indexed by action id
__host__ __device__ __forceinline__ uint32_t new_stage_of(uint32_t action) {
  // 1 -> SCHEDULING_FAILED(5) :297, 2 -> FAILED(4) :324, 3 -> DEADLINE_EXCEEDED(6) :349, 4 -> RUNNING(2) :362
  return (0x26450u >> (action * 4)) & 15u;
}
__host__ __device__ __forceinline__ uint32_t cause_tpl_of(uint32_t action) {
  // 1 -> T0 :298, 2 -> T1 :325, 3 -> message verbatim :350, 4 -> untouched :361-364
  return (0xE4u >> (action * 2)) & 3u;
}
__host__ __device__ __forceinline__ uint32_t make_code(uint32_t action, uint32_t msg_tpl) {
  return action | new_stage_of(action) << 3 | cause_tpl_of(action) << 7 | msg_tpl << 9;
}
// Verdict of an APPLIED record with this action (msg_tpl added by the caller).
__host__ __device__ __forceinline__ uint32_t applied_verdict(uint32_t action) {
  return NXS_V_PACK(action, new_stage_of(action), cause_tpl_of(action), 0, action != NXS_ACTION_RUNNING, 1,
                    0, 0);
}
constexpr uint32_t kCodeRunning = NXS_ACTION_RUNNING | NXS_STAGE_RUNNING << 3; // last applied record was a ToRunning

// --------------------------------------------------------------- fold algebra
// Summary of a run of consecutive records of ONE job, valid under the premise
// "the row was not finished when the run began".  (If it was, nothing applies
// and every handled record is a skip: the job's anchor then carries a stop, see summ_anchor.)
// The low 16 bits ARE the packed verdict fields, so closing a job is a mask.
struct Summ {
  uint32_t win_idx; // record index of the winner (if F_HAS_WIN)
  uint32_t f;
};
enum : uint32_t {
  F_LAST_MASK = 0x7Fu,    // bits 0-6: action and new stage of the last applied record (verdict.action/new_stage)
  F_WIN_MASK = 31u << 7,  // bits 7-8 cause_tpl, 9-11 msg_tpl of the winner (verdict.cause_tpl/msg_tpl)
  F_HAS_WIN = 1u << 12,   // some applied record had a failure action (verdict.delete_job)
  F_HAS_APP = 1u << 13,   // some handled record with a checkpoint: applied under the premise (verdict.upsert)
  F_SKIP = 1u << 14,      // some handled record came after the stop (verdict.skip_finished)
  F_ERR = 1u << 15,       // some record had no checkpoint (verdict.err = 1)
  F_STOPPED = 1u << 16,   // an applied record moved the row to a finished stage
  F_FIN0 = 1u << 17,      // the row was already finished at the job's first record (informational: it comes with F_STOPPED)
  F_ANCHORED = 1u << 18,  // the run starts at the job's first record (look-back stops here)
  F_BITS = 19
};

__host__ __device__ __forceinline__ Summ summ_identity() { return Summ{0u, 0u}; }
// A run that begins at the job's first record; fin0 = IsFinished() of the row it found (:276).
// A finished row is a stop before the job's first record: nothing after it applies, every handled
// record is a skip (:276-279) — which is exactly what summ_combine does to the right of a stop.
__host__ __device__ __forceinline__ Summ summ_anchor(bool fin0) {
  return Summ{0u, F_ANCHORED | (fin0 ? (F_FIN0 | F_STOPPED) : 0u)};
}
// anchor(fin0) combined with a piece computed under the premise, on the flags alone
__host__ __device__ __forceinline__ uint32_t summ_anchor_flags(uint32_t piece_f, uint32_t fin0_bit) {
  if (fin0_bit) piece_f = (piece_f & F_ERR) | F_STOPPED | ((piece_f & F_HAS_APP) << 1); // F_HAS_APP -> F_SKIP
  return piece_f | F_ANCHORED | fin0_bit;
}
__host__ __device__ __forceinline__ uint64_t summ_pack(Summ s) {
  return (uint64_t)s.f << 32 | s.win_idx;
}
__host__ __device__ __forceinline__ Summ summ_unpack(uint64_t v) {
  return Summ{(uint32_t)v, (uint32_t)(v >> 32)};
}

// L then R (L earlier in record order).  Associative.
__host__ __device__ __forceinline__ Summ summ_combine(Summ L, Summ R) {
  uint32_t f = (L.f | R.f) & F_ERR;
  f |= ((L.f & F_ANCHORED) ? L.f : R.f) & (F_FIN0 | F_ANCHORED);
  uint32_t win;
  if (L.f & F_STOPPED) {
    f |= L.f & (F_STOPPED | F_HAS_APP | F_LAST_MASK | F_HAS_WIN | F_WIN_MASK | F_SKIP);
    if (R.f & F_HAS_APP) f |= F_SKIP;
    win = L.win_idx;
  } else {
    f |= R.f & F_STOPPED;
    f |= (L.f | R.f) & (F_HAS_APP | F_SKIP);
    f |= ((R.f & F_HAS_APP) ? R.f : L.f) & F_LAST_MASK;
    const bool rw = (R.f & F_HAS_WIN) != 0;
    f |= (rw ? R.f : L.f) & (F_HAS_WIN | F_WIN_MASK);
    win = rw ? R.win_idx : L.win_idx;
  }
  return Summ{win, f};
}

__host__ __device__ __forceinline__ uint32_t hi_bit(uint32_t m) { // m != 0
#ifdef __CUDA_ARCH__
  uint32_t r;
  asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(m));
  return r;
#else
  return 31u - __builtin_clz(m);
#endif
}
__host__ __device__ __forceinline__ uint32_t lo_bit(uint32_t m) { // m != 0
#ifdef __CUDA_ARCH__
  return __ffs(m) - 1u;
#else
  return __builtin_ctz(m);
#endif
}

// Flag part of the summary of up to 32 consecutive records given as bit masks
// (bit b = record b of the chunk), already restricted to the run.  he: handled &
// checkpoint exists; term: he & the action's stage is finished; fail: he &
// failure action; err: handled & no checkpoint.  Returns the flags without the
// winner's code; win_pos = chunk position of the winner (meaningful iff F_HAS_WIN),
// last_is_win = the last applied record is the winner (else it is a ToRunning).
__host__ __device__ __forceinline__ uint32_t summ_piece_flags(uint32_t he, uint32_t term, uint32_t fail,
                                                              uint32_t err, uint32_t &win_pos, bool &last_is_win) {
  uint32_t f = err ? F_ERR : 0u;
  uint32_t app = he;
  if (term) {
    const uint32_t upto = (2u << lo_bit(term)) - 1u; // records up to and including the first stop
    app = he & upto;
    f |= F_STOPPED;
    if (he & ~upto) f |= F_SKIP;
  }
  if (he) f |= F_HAS_APP;
  const uint32_t wm = app & fail;
  win_pos = hi_bit(wm | 1u);
  last_is_win = wm != 0 && win_pos == hi_bit(app | 1u);
  if (wm) f |= F_HAS_WIN;
  return f;
}
// Completes the flags with the winner's code.
__host__ __device__ __forceinline__ uint32_t summ_piece_finish(uint32_t f, uint32_t code_w, bool last_is_win) {
  if (f & F_HAS_WIN) f |= code_w & F_WIN_MASK;
  if (f & F_HAS_APP) f |= last_is_win ? (code_w & F_LAST_MASK) : kCodeRunning;
  return f;
}
// Convenience form with the codes in memory (host walk-through).
__host__ __device__ __forceinline__ Summ summ_piece(uint32_t he, uint32_t term, uint32_t fail, uint32_t err,
                                                    const uint16_t *code, uint32_t base, uint32_t global_base) {
  uint32_t wp;
  bool liw;
  const uint32_t f = summ_piece_flags(he, term, fail, err, wp, liw);
  const uint32_t cw = (f & F_HAS_WIN) ? code[base + wp] : 0u;
  return Summ{(f & F_HAS_WIN) ? global_base + base + wp : 0u, summ_piece_finish(f, cw, liw)};
}

// ------------------------------------------------------------ one lane per job
// The job pass of the kernel gives every run of records (a job, or the part of a job that lies in a
// worker's range) to ONE lane.  The lane cuts the run's bits out of the per-chunk ballot words and
// summarises them here.  All masks are already restricted to the run (bit 0 = its first record).
//   hes: handled & checkpoint exists; tms: ... & the action's stage is finished; fl: failure action
//   (need not be restricted); er: handled & no checkpoint.
//   stop0: the run starts at the job's first record and IsFinished() held for the row it found
//   (:276) — the anchor's stop: nothing applies and every handled record is a skip.
// Returns the flags except the winner's code bits; wm = the applied failure records (winner = its
// highest bit), last_is_win = the last applied record is the winner.
__host__ __device__ __forceinline__ uint32_t summ_run_flags(uint32_t hes, uint32_t tms, uint32_t fl, uint32_t er,
                                                            bool stop0, uint32_t &wm, bool &last_is_win) {
  const uint32_t upto = tms ^ (tms - 1u); // through the first stop; all ones when there is none
  const uint32_t app = stop0 ? 0u : (hes & upto);
  wm = app & fl;
  last_is_win = wm > (app & ~fl); // the highest applied bit is a failure record
  uint32_t f = app ? (F_HAS_APP | kCodeRunning) : 0u;
  if (er) f |= F_ERR;
  if (tms) f |= F_STOPPED;
  if (stop0) f |= F_STOPPED | F_FIN0;
  if (hes != app) f |= F_SKIP; // handled records after the stop
  return f;
}
// codex = code of the winner's switch arm | F_HAS_WIN | F_HAS_APP, or 0 when there is no winner.
__host__ __device__ __forceinline__ uint32_t summ_run_finish(uint32_t f, uint32_t codex, bool last_is_win) {
  return last_is_win ? (codex | (f & ~(F_LAST_MASK | F_WIN_MASK))) : (f | (codex & ~F_LAST_MASK));
}

// Close a job (its summary must be anchored): emit (verdict, winner_idx).
__host__ __device__ __forceinline__ void summ_resolve(Summ s, uint32_t &verdict, uint32_t &winner) {
  verdict = (s.f & 0xFFFFu) | (s.f & F_ERR) >> 3; // a missing checkpoint also deletes the job (:268)
  winner = (s.f & F_HAS_WIN) ? s.win_idx : NXS_NO_WINNER;
}

} // namespace nxs
