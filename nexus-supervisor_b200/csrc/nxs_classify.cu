// nxs_classify.cu — context, launch logic and the C ABI declared in
// include/nxs_classify.h.  The GPU path is the only path: no CPU fallback.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "nxs_host.h"
#include "nxs_kernel.cuh"

using namespace nxs;

struct nxs_ctx {
  int device = 0;
  int num_sms = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  uint64_t max_records = 0;
  uint32_t finished_mask = 0;
  uint32_t term_mask = 0;
  int tile = 0;           // records per tile forced with NXS_TILE; 0 = chosen per launch from the batch size
  int ctas_per_sm = 1;    // resident CTAs (the same for the three tile sizes: the ring fills the shared memory)
  uint32_t max_tiles = 0;
  TileDesc *desc = nullptr;
  uint32_t *counters = nullptr;
  uint32_t gen = 0;
  uint64_t launches = 0;
  int last_tile = 0; // records per tile of the last classification launch
  uint8_t lut[128];
  uint16_t code[8];
  // staging for the host-pointer entry point (allocated on first use): two buffer sets, so the
  // upload of one slice overlaps the classification and the download of the previous one
  uint64_t staged_records = 0; // capacity of ONE set, in records
  uint8_t *d_in[2] = {nullptr, nullptr};
  uint32_t *d_verdict[2] = {nullptr, nullptr};
  nxs_job_row *d_rows[2] = {nullptr, nullptr};
  uint64_t wire_h2d = 0, wire_d2h = 0; // bytes the last host-pointer call moved over the link
  cudaEvent_t e_codes[2] = {nullptr, nullptr}; // a slice's verdict bytes have arrived in h_codes
  uint8_t *h_codes[2] = {nullptr, nullptr}; // page-locked: a slice's per-record verdicts as bytes, before their expansion
  uint32_t verdict_of[256];                 // verdict_of_code() of every byte
  uint16_t *h_attr[2] = {nullptr, nullptr}; // page-locked: the four small columns of a slice, packed for the upload
  uint64_t *d_njobs = nullptr; // 2 counters
  uint64_t *h_njobs = nullptr; // pinned, 2 counters
  // page-locked bounce buffers for callers whose own buffers are pageable (allocated on first such call)
  uint64_t bounce_records = 0; // capacity of ONE set, in records
  uint8_t *h_in[2] = {nullptr, nullptr};
  uint32_t *h_verdict[2] = {nullptr, nullptr};
  nxs_job_row *h_rows[2] = {nullptr, nullptr};
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t e_h2d[2] = {nullptr, nullptr}, e_kernel[2] = {nullptr, nullptr}, e_count[2] = {nullptr, nullptr},
              e_d2h[2] = {nullptr, nullptr};
  char err[256] = {0};
#ifdef NXS_TRACE
  unsigned long long *trace = nullptr;
#endif
};

namespace {

int fail_cuda(nxs_ctx *ctx, cudaError_t e, const char *what) {
  if (ctx) snprintf(ctx->err, sizeof ctx->err, "%s: %s", what, cudaGetErrorString(e));
  return NXS_ERR_CUDA;
}
#define NXS_CUDA(ctx, call)                                 \
  do {                                                      \
    cudaError_t e_ = (call);                                \
    if (e_ != cudaSuccess) return fail_cuda(ctx, e_, #call); \
  } while (0)

// Makes `device` current for the calling thread and puts the caller's device back on scope exit: a
// multi-GPU host (torch, a per-GPU worker pool) must not find its current device changed by a call.
struct DeviceGuard {
  int prev = -1;
  cudaError_t status;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    status = cudaSetDevice(device);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
  DeviceGuard(const DeviceGuard &) = delete;
  DeviceGuard &operator=(const DeviceGuard &) = delete;
};

int fail_arg(nxs_ctx *ctx, const char *msg) {
  if (ctx) snprintf(ctx->err, sizeof ctx->err, "%s", msg);
  return NXS_ERR_ARG;
}

template <int TILE> int configure_variant(nxs_ctx *ctx) {
  const size_t smem = classify_smem_bytes<TILE>();
  NXS_CUDA(ctx, cudaFuncSetAttribute(classify_kernel<TILE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  NXS_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, classify_kernel<TILE>, kThreads, smem));
  if (occ < 1) {
    snprintf(ctx->err, sizeof ctx->err, "classify kernel does not fit on an SM");
    return NXS_ERR_CUDA;
  }
  ctx->ctas_per_sm = occ;
  return NXS_OK;
}

template <int TILE> int launch_variant(nxs_ctx *ctx, const ClassifyParams &p, uint32_t grid) {
  classify_kernel<TILE><<<grid, kThreads, classify_smem_bytes<TILE>(), ctx->stream>>>(p);
  NXS_CUDA(ctx, cudaGetLastError());
  return NXS_OK;
}

// Tile size for a batch of n records when NXS_TILE does not force one.  Measured on a B200 with COLD
// inputs (scripts/small_batch.py rewrites 512 MB before every launch, so the batch comes from HBM, not
// from L2; us per launch, 2048 / 4096 / 8192-record tiles): 1e5 records 14.3 / 16.4 / 22.5, 3e5 16.4 /
// 18.4 / 29.7, 1e6 23.6 / 21.5 / 23.5, 4e6 44.0 / 41.0 / 41.0, 1e7 84.0 / 69.7 / 67.6, 4e7 279 / 220 /
// 197, 6.4e7 434 / 338 / 289 (profiles/r02_v11_small_batches.txt): small tiles spread a small batch
// over more SMs, large tiles mean fewer look-backs and hand-overs per record.
int tile_for_batch(uint64_t n) {
  if (n <= 500000u) return NXS_TILE_S;
  if (n <= 6000000u) return NXS_TILE_M;
  return NXS_TILE_L;
}

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// nxs_batch_packed on the device: attr (u16 per record) -> the four byte-wide columns the classification
// kernel reads.  `reason` may alias `attr` (same index, same width).  7 bytes of HBM traffic per record —
// next to nothing beside the PCIe transfer the packed layout exists to shorten.
__global__ void expand_attr_kernel(const uint16_t *attr, uint64_t n, uint16_t *reason, uint8_t *kind, uint8_t *flags,
                                   uint8_t *stage) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t a = attr[i];
    reason[i] = (uint16_t)NXS_ATTR_REASON(a);
    kind[i] = (uint8_t)NXS_ATTR_KIND(a);
    flags[i] = (uint8_t)NXS_ATTR_FLAGS(a);
    stage[i] = (uint8_t)NXS_ATTR_STAGE(a);
  }
}

} // namespace

extern "C" {

int nxs_abi_version(void) { return NXS_ABI_VERSION; }

const char *nxs_strerror(int status) {
  switch (status) {
  case NXS_OK: return "ok";
  case NXS_ERR_ARG: return "invalid argument";
  case NXS_ERR_CAPACITY: return "batch exceeds the context capacity";
  case NXS_ERR_CUDA: return "CUDA error";
  case NXS_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
  case NXS_ERR_NOMEM: return "out of memory";
  case NXS_ERR_UNKNOWN_ACTION: return "unknown analysis result action";
  case NXS_ERR_NOALLOC_LIMIT: return "batch exceeds NXS_NOALLOC_MAX_RECORDS (the static scratch of the no-alloc entry point)";
  default: return "unknown status";
  }
}

const char *nxs_last_error(const nxs_ctx *ctx) { return ctx ? ctx->err : ""; }

int nxs_create(int device, uint64_t max_records, uint16_t finished_mask, nxs_ctx **out) {
  if (!out) return NXS_ERR_ARG;
  *out = nullptr;
  if (max_records == 0 || max_records >= 0xFFFFFFFFull) return NXS_ERR_CAPACITY;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
    cudaGetLastError();
    return NXS_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= n_dev) return NXS_ERR_ARG;
  nxs_ctx *ctx = new (std::nothrow) nxs_ctx();
  if (!ctx) return NXS_ERR_NOMEM;
  ctx->device = device;
  ctx->max_records = max_records;
  ctx->finished_mask = finished_mask;
  for (uint32_t a = 1; a <= 4; ++a)
    if ((finished_mask >> new_stage_of(a)) & 1u) ctx->term_mask |= 1u << a;
  nxs_build_lut(ctx->lut, ctx->code);
  for (uint32_t b = 0; b < 256; ++b) ctx->verdict_of[b] = verdict_of_code(b, ctx->code, ctx->term_mask);
  int rc = NXS_OK;
  auto bail = [&](int code) {
    nxs_destroy(ctx);
    return code;
  };
  DeviceGuard guard(device);
  if (guard.status != cudaSuccess) return bail(NXS_ERR_CUDA);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail(NXS_ERR_CUDA);
  if (prop.major < 10) return bail(NXS_ERR_NO_DEVICE); // sm_100a code only
  ctx->num_sms = prop.multiProcessorCount;
  const char *tile_env = getenv("NXS_TILE");
  ctx->tile = tile_env ? atoi(tile_env) : 0;
  if (ctx->tile != 0 && ctx->tile != NXS_TILE_S && ctx->tile != NXS_TILE_M && ctx->tile != NXS_TILE_L) return bail(NXS_ERR_ARG);
  rc = configure_variant<NXS_TILE_S>(ctx);
  int occ_min = ctx->ctas_per_sm;
  if (rc == NXS_OK) rc = configure_variant<NXS_TILE_M>(ctx);
  occ_min = ctx->ctas_per_sm < occ_min ? ctx->ctas_per_sm : occ_min;
  if (rc == NXS_OK) rc = configure_variant<NXS_TILE_L>(ctx);
  ctx->ctas_per_sm = ctx->ctas_per_sm < occ_min ? ctx->ctas_per_sm : occ_min;
  if (rc != NXS_OK) return bail(rc);
  if (getenv("NXS_VERBOSE"))
    fprintf(stderr, "nxs: device %d, %d SMs, tile %d records (0 = by batch size), %d resident CTAs/SM x %d threads\n",
            device, ctx->num_sms, ctx->tile, ctx->ctas_per_sm, kThreads);
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(NXS_ERR_CUDA);
  ctx->own_stream = true;
  ctx->max_tiles = (uint32_t)((max_records + NXS_TILE_S - 1) / NXS_TILE_S); // descriptors for the smallest tile
  if (cudaMalloc(&ctx->desc, sizeof(TileDesc) * (size_t)ctx->max_tiles) != cudaSuccess) return bail(NXS_ERR_NOMEM);
  if (cudaMalloc(&ctx->counters, 64) != cudaSuccess) return bail(NXS_ERR_NOMEM);
  if (cudaMemset(ctx->desc, 0, sizeof(TileDesc) * (size_t)ctx->max_tiles) != cudaSuccess) return bail(NXS_ERR_CUDA);
  if (cudaMemset(ctx->counters, 0, 64) != cudaSuccess) return bail(NXS_ERR_CUDA);
  if (cudaMalloc(&ctx->d_njobs, 16) != cudaSuccess) return bail(NXS_ERR_NOMEM);
  if (cudaMallocHost(&ctx->h_njobs, 16) != cudaSuccess) return bail(NXS_ERR_NOMEM);
  *out = ctx;
  return NXS_OK;
}

void nxs_destroy(nxs_ctx *ctx) {
  if (!ctx) return;
  DeviceGuard guard(ctx->device);
  if (ctx->stream && ctx->own_stream) {
    cudaStreamSynchronize(ctx->stream);
    cudaStreamDestroy(ctx->stream);
  }
  cudaFree(ctx->desc);
  cudaFree(ctx->counters);
  for (int b = 0; b < 2; ++b) {
    cudaFree(ctx->d_in[b]);
    cudaFree(ctx->d_verdict[b]);
    cudaFree(ctx->d_rows[b]);
    if (ctx->h_attr[b]) cudaFreeHost(ctx->h_attr[b]);
    if (ctx->h_codes[b]) cudaFreeHost(ctx->h_codes[b]);
    if (ctx->e_h2d[b]) cudaEventDestroy(ctx->e_h2d[b]);
    if (ctx->e_codes[b]) cudaEventDestroy(ctx->e_codes[b]);
    if (ctx->e_kernel[b]) cudaEventDestroy(ctx->e_kernel[b]);
    if (ctx->e_count[b]) cudaEventDestroy(ctx->e_count[b]);
    if (ctx->e_d2h[b]) cudaEventDestroy(ctx->e_d2h[b]);
  }
  if (ctx->s_h2d) cudaStreamDestroy(ctx->s_h2d);
  if (ctx->s_d2h) cudaStreamDestroy(ctx->s_d2h);
  cudaFree(ctx->d_njobs);
  if (ctx->h_njobs) cudaFreeHost(ctx->h_njobs);
  for (int b = 0; b < 2; ++b) {
    if (ctx->h_in[b]) cudaFreeHost(ctx->h_in[b]);
    if (ctx->h_verdict[b]) cudaFreeHost(ctx->h_verdict[b]);
    if (ctx->h_rows[b]) cudaFreeHost(ctx->h_rows[b]);
  }
  cudaGetLastError();
  delete ctx;
}

int nxs_set_stream(nxs_ctx *ctx, void *cuda_stream) {
  if (!ctx) return NXS_ERR_ARG;
  DeviceGuard guard(ctx->device);
  // Launches of one ctx share its tile descriptors and counters: whatever the old stream still holds
  // must be done before work can be queued on another stream.
  if (ctx->launches > 0 || ctx->own_stream) NXS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  ctx->stream = static_cast<cudaStream_t>(cuda_stream);
  ctx->own_stream = false;
  return NXS_OK;
}

int nxs_sync(nxs_ctx *ctx) {
  if (!ctx) return NXS_ERR_ARG;
  DeviceGuard guard(ctx->device);
  NXS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return NXS_OK;
}

uint64_t nxs_launch_count(const nxs_ctx *ctx) { return ctx ? ctx->launches : 0; }
int nxs_last_tile(const nxs_ctx *ctx) { return ctx ? ctx->last_tile : 0; }

static int launch_classify(nxs_ctx *ctx, const nxs_batch *in, const nxs_out *out, uint32_t record_base,
                           uint8_t *verdict_code = nullptr);

int nxs_classify_device(nxs_ctx *ctx, const nxs_batch *in, const nxs_out *out) {
  return launch_classify(ctx, in, out, 0);
}

// Validates device pointers and enqueues one classification launch on the ctx stream.
// record_base is added to every winner_idx (used when a batch is processed in slices).
// verdict_code != nullptr (host-pointer path only): one byte per record instead of out->verdict, which is then null.
static int launch_classify(nxs_ctx *ctx, const nxs_batch *in, const nxs_out *out, uint32_t record_base,
                           uint8_t *verdict_code) {
  if (!ctx || !in || !out) return NXS_ERR_ARG;
  if (!out->n_jobs) return fail_arg(ctx, "out->n_jobs is null");
  if (in->n > ctx->max_records) {
    snprintf(ctx->err, sizeof ctx->err, "n=%llu exceeds max_records=%llu", (unsigned long long)in->n,
             (unsigned long long)ctx->max_records);
    return NXS_ERR_CAPACITY;
  }
  DeviceGuard guard(ctx->device);
  NXS_CUDA(ctx, guard.status);
  if (in->n == 0) {
    NXS_CUDA(ctx, cudaMemsetAsync(out->n_jobs, 0, sizeof(uint64_t), ctx->stream));
    return NXS_OK;
  }
  if (!in->kind || !in->reason_id || !in->flags || !in->cur_stage || !in->job_key || !out->job_rows)
    return fail_arg(ctx, "null column pointer");
  if (!aligned16(in->kind) || !aligned16(in->reason_id) || !aligned16(in->flags) || !aligned16(in->cur_stage) ||
      !aligned16(in->job_key) || !aligned16(out->job_rows) || (reinterpret_cast<uintptr_t>(out->verdict) & 3u))
    return fail_arg(ctx, "device columns must be 16-byte aligned");

  if (ctx->gen >= kGenMax) { // 7-bit descriptor generation wrapped: start over from zeroed scratch
    NXS_CUDA(ctx, cudaMemsetAsync(ctx->desc, 0, sizeof(TileDesc) * (size_t)ctx->max_tiles, ctx->stream));
    ctx->gen = 0;
  }
  ClassifyParams p;
  p.kind = in->kind;
  p.reason = in->reason_id;
  p.flags = in->flags;
  p.stage = in->cur_stage;
  p.key = in->job_key;
  p.verdict = out->verdict;
  p.verdict_code = verdict_code;
  p.rows = out->job_rows;
  p.n_jobs = out->n_jobs;
  p.desc = ctx->desc;
  p.counters = ctx->counters;
  p.n = (uint32_t)in->n;
  const int tile = ctx->tile ? ctx->tile : tile_for_batch(in->n);
  p.n_tiles = (uint32_t)((in->n + tile - 1) / tile);
  p.record_base = record_base;
  p.finished_mask = ctx->finished_mask;
  p.term_mask = ctx->term_mask;
  p.gen = ++ctx->gen;
  memcpy(p.lut, ctx->lut, sizeof p.lut);
  memcpy(p.code, ctx->code, sizeof p.code);
#ifdef NXS_TRACE
  if (!ctx->trace) cudaMalloc(&ctx->trace, 8 + 16 * (size_t)(1 << 20));
  cudaMemsetAsync(ctx->trace, 0, 8, ctx->stream);
  p.trace = ctx->trace;
  p.trace_cap = 1 << 20;
#endif
  const uint32_t max_grid = (uint32_t)(ctx->num_sms * ctx->ctas_per_sm);
  const uint32_t grid = p.n_tiles < max_grid ? p.n_tiles : max_grid;
  int rc;
  if (tile == NXS_TILE_S) rc = launch_variant<NXS_TILE_S>(ctx, p, grid);
  else if (tile == NXS_TILE_M) rc = launch_variant<NXS_TILE_M>(ctx, p, grid);
  else rc = launch_variant<NXS_TILE_L>(ctx, p, grid);
  if (rc == NXS_OK) {
    ctx->launches += 1;
    ctx->last_tile = tile;
  }
  return rc;
}

// Records per slice of the host-pointer path.  Large enough that a launch is far from
// launch-bound (thousands of tiles), small enough that upload, classification and download of
// neighbouring slices overlap well.
static const uint64_t kHostSliceRecords = 8u << 20;

static int ensure_staging(nxs_ctx *ctx, uint64_t slice) {
  if (!ctx->s_h2d) {
    if (cudaStreamCreateWithFlags(&ctx->s_h2d, cudaStreamNonBlocking) != cudaSuccess) return NXS_ERR_CUDA;
    if (cudaStreamCreateWithFlags(&ctx->s_d2h, cudaStreamNonBlocking) != cudaSuccess) return NXS_ERR_CUDA;
    for (int b = 0; b < 2; ++b) {
      if (cudaEventCreateWithFlags(&ctx->e_h2d[b], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&ctx->e_kernel[b], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&ctx->e_count[b], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&ctx->e_d2h[b], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&ctx->e_codes[b], cudaEventDisableTiming) != cudaSuccess)
        return NXS_ERR_CUDA;
    }
  }
  if (slice <= ctx->staged_records) return NXS_OK;
  const uint64_t cap = (slice + 1023) & ~1023ull; // keeps every column 16-byte aligned
  for (int b = 0; b < 2; ++b) {
    cudaFree(ctx->d_in[b]);
    cudaFree(ctx->d_verdict[b]);
    cudaFree(ctx->d_rows[b]);
    if (ctx->h_attr[b]) cudaFreeHost(ctx->h_attr[b]);
    if (ctx->h_codes[b]) cudaFreeHost(ctx->h_codes[b]);
    ctx->h_codes[b] = nullptr;
    ctx->d_in[b] = nullptr;
    ctx->d_verdict[b] = nullptr;
    ctx->d_rows[b] = nullptr;
    ctx->h_attr[b] = nullptr;
  }
  ctx->staged_records = 0;
  for (int b = 0; b < 2; ++b) {
    if (cudaMalloc(&ctx->d_in[b], cap * 13) != cudaSuccess) return NXS_ERR_NOMEM;
    if (cudaMalloc(&ctx->d_verdict[b], cap * 4) != cudaSuccess) return NXS_ERR_NOMEM;
    if (cudaMalloc(&ctx->d_rows[b], cap * sizeof(nxs_job_row)) != cudaSuccess) return NXS_ERR_NOMEM;
    if (cudaMallocHost(&ctx->h_attr[b], cap * sizeof(uint16_t)) != cudaSuccess) return NXS_ERR_NOMEM;
    if (cudaMallocHost(&ctx->h_codes[b], cap) != cudaSuccess) return NXS_ERR_NOMEM;
  }
  ctx->staged_records = cap;
  return NXS_OK;
}

// Host-pointer entry point.  The batch is cut into slices on job boundaries; slice k+1 is uploaded
// (copy stream 1) while slice k is classified (ctx stream) and slice k-1 is downloaded (copy stream
// 2), with two device buffer sets.  Job rows of consecutive slices are appended in the caller's
// buffer, so a slice's rows are downloaded once the job count of the previous one is known.
// Pageable caller memory.  cudaMemcpyAsync from or to ordinary (not page-locked) host memory is staged
// by the driver through a small buffer of its own and blocks the calling thread: ~10 GB/s, and upload,
// kernel and download stop overlapping (0.6e9 instead of 3.5e9 records/s on one B200).  For batches
// large enough to matter the library therefore bounces such buffers through page-locked ones of its own,
// copied by a few host threads: slice k+1 is copied in while slice k is on the GPU.
static const uint64_t kBounceMinRecords = 2u << 20;
static bool is_pageable(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}
static int ensure_bounce(nxs_ctx *ctx, uint64_t cap) {
  if (cap <= ctx->bounce_records) return NXS_OK;
  for (int b = 0; b < 2; ++b) {
    if (ctx->h_in[b]) cudaFreeHost(ctx->h_in[b]);
    if (ctx->h_verdict[b]) cudaFreeHost(ctx->h_verdict[b]);
    if (ctx->h_rows[b]) cudaFreeHost(ctx->h_rows[b]);
    ctx->h_in[b] = nullptr, ctx->h_verdict[b] = nullptr, ctx->h_rows[b] = nullptr;
  }
  ctx->bounce_records = 0;
  for (int b = 0; b < 2; ++b) {
    if (cudaMallocHost(&ctx->h_in[b], cap * 13) != cudaSuccess || cudaMallocHost(&ctx->h_verdict[b], cap * 4) != cudaSuccess ||
        cudaMallocHost(&ctx->h_rows[b], cap * sizeof(nxs_job_row)) != cudaSuccess) {
      cudaGetLastError();
      return NXS_ERR_NOMEM;
    }
  }
  ctx->bounce_records = cap;
  return NXS_OK;
}
struct CopySeg {
  void *dst;
  const void *src;
  size_t bytes;
};
#ifndef NXS_BOUNCE_THREADS
#define NXS_BOUNCE_THREADS 8
#endif
static const unsigned kBounceThreads = NXS_BOUNCE_THREADS;
// memcpy of a few segments, the bytes split evenly over up to kBounceThreads host threads.
static void par_copy(const CopySeg *seg, int n_seg) {
  size_t total = 0;
  for (int i = 0; i < n_seg; ++i) total += seg[i].bytes;
  const unsigned hw = std::thread::hardware_concurrency();
  const unsigned T = (unsigned)std::min<size_t>(std::min(kBounceThreads, hw ? hw : 4u), total / (4u << 20) + 1);
  auto work = [&](unsigned t) {
    size_t lo = total * t / T, hi = total * (t + 1) / T, off = 0;
    for (int i = 0; i < n_seg && off < hi; ++i) {
      const size_t a = std::max(lo, off), b = std::min(hi, off + seg[i].bytes);
      if (a < b) memcpy(static_cast<char *>(seg[i].dst) + (a - off), static_cast<const char *>(seg[i].src) + (a - off), b - a);
      off += seg[i].bytes;
    }
  };
  run_parallel(T, work);
}

// dst[i] = table[codes[i]]: the per-record verdicts of a slice from their one-byte form, over a few host threads.
static void expand_codes(const uint8_t *codes, uint64_t n, const uint32_t *table, uint32_t *dst) {
  const unsigned hw = std::thread::hardware_concurrency();
  const unsigned T = (unsigned)std::min<uint64_t>(std::min(kBounceThreads, hw ? hw : 4u), n / (1u << 20) + 1);
  auto work = [=](unsigned t) {
    for (uint64_t i = n * t / T, e = n * (t + 1) / T; i < e; ++i) dst[i] = table[codes[i]];
  };
  run_parallel(T, work);
}

// attr[i] = NXS_ATTR_PACK(...) for i in [0, n): a streaming loop (one thread moves ~1.4e9 records/s), split over a
// few host threads for large n.
static void pack_columns(const uint8_t *kind, const uint16_t *reason_id, const uint8_t *flags, const uint8_t *cur_stage,
                         uint64_t n, uint16_t *attr) {
  const unsigned hw = std::thread::hardware_concurrency();
  const unsigned T = (unsigned)std::min<uint64_t>(std::min(16u, hw ? hw : 4u), n / (1u << 20) + 1);
  auto work = [=](unsigned t) {
    for (uint64_t i = n * t / T, e = n * (t + 1) / T; i < e; ++i)
      attr[i] = (uint16_t)NXS_ATTR_PACK(kind[i], reason_id[i], flags[i], cur_stage[i]);
  };
  run_parallel(T, work);
}

// `attr` != nullptr: the batch is packed (job_key + attr); the four small columns of `in` are unused.
// Otherwise the library packs them itself, slice by slice, on their way to the GPU: the kernel reads only
// 6 + 2 + 3 + 4 bits of them (reasons above 32 and kinds above 2 are `default` to both switches), so 10
// instead of 13 bytes per record cross the link and nothing else changes (NXS_NO_AUTOPACK=1 turns it off).
static int classify_slices(nxs_ctx *ctx, const nxs_batch *in, const uint16_t *attr, const nxs_out *out,
                           uint64_t slice_cap) {
  const uint64_t n = in->n;
  const uint64_t cap = ctx->staged_records;
  cudaStream_t s_comp = ctx->stream;
  // bounce pageable caller buffers through page-locked ones (see above)
  const bool bounce_in = n >= kBounceMinRecords && is_pageable(in->job_key);
  const bool bounce_out = n >= kBounceMinRecords && is_pageable(out->job_rows);
  if (bounce_in || bounce_out) {
    const int brc = ensure_bounce(ctx, cap);
    if (brc != NXS_OK) {
      snprintf(ctx->err, sizeof ctx->err, "cannot allocate page-locked bounce buffers for slices of %llu records",
               (unsigned long long)cap);
      return brc;
    }
  }
  // multi-slice batches: the small columns travel packed (the packing of slice k+1 hides behind the upload of slice k)
  static const bool no_autopack = getenv("NXS_NO_AUTOPACK") != nullptr;
  const bool autopack = !attr && !no_autopack && n > slice_cap;
  const bool packed = attr != nullptr || autopack;
  // ... and the per-record verdicts come back as one byte each (4 * arm | 32 * gate | 128 * finished: what the
  // 32-bit verdict is a function of), expanded into the caller's array by the helper threads below
  // (NXS_NO_VERDICT_CODES=1 turns it off)
  static const bool no_codes = getenv("NXS_NO_VERDICT_CODES") != nullptr;
  const bool codes = out->verdict != nullptr && !no_codes && n > slice_cap;
  const bool helpers = bounce_out || codes; // results pass through host staging: a helper thread per slice finishes them
  ctx->wire_h2d = 0, ctx->wire_d2h = 0;
  const bool verbose = (bounce_in || helpers || autopack) && getenv("NXS_VERBOSE") != nullptr;
  double t_in = 0, t_wait_count = 0, t_wait_d2h = 0;
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  uint64_t lo = 0, rows_done = 0;
  uint64_t pend_lo[2] = {0, 0}, pend_n[2] = {0, 0}; // slices whose rows are still to be downloaded
  bool pending[2] = {false, false};
  // slices whose outputs sit in the bounce buffers and still have to be copied to the caller's
  bool out_wait[2] = {false, false};
  uint64_t out_lo[2] = {0, 0}, out_n[2] = {0, 0}, out_row0[2] = {0, 0}, out_jobs[2] = {0, 0};
  int rc = NXS_OK;
  auto finish_slice = [&](int b) -> int { // job count is (about to be) on the host: download the rows
    const double t0 = now();
    NXS_CUDA(ctx, cudaEventSynchronize(ctx->e_count[b]));
    t_wait_count += now() - t0;
    const uint64_t jobs = ctx->h_njobs[b];
    if (jobs > pend_n[b]) {
      snprintf(ctx->err, sizeof ctx->err, "internal error: %llu jobs in a slice of %llu records",
               (unsigned long long)jobs, (unsigned long long)pend_n[b]);
      return NXS_ERR_CUDA;
    }
    nxs_job_row *dst = bounce_out ? ctx->h_rows[b] : out->job_rows + rows_done;
    NXS_CUDA(ctx, cudaMemcpyAsync(dst, ctx->d_rows[b], jobs * sizeof(nxs_job_row), cudaMemcpyDeviceToHost, ctx->s_d2h));
    NXS_CUDA(ctx, cudaEventRecord(ctx->e_d2h[b], ctx->s_d2h));
    ctx->wire_d2h += jobs * sizeof(nxs_job_row) + 8;
    out_row0[b] = rows_done;
    out_jobs[b] = jobs;
    rows_done += jobs;
    pending[b] = false;
    return NXS_OK;
  };
  // Bounce buffers of set b -> the caller's buffers, once the set's downloads are done: a helper thread per
  // slice, started when the slice's row download has been queued, so that these copies run beside the
  // main thread's copies INTO the bounce buffers and its waits (joined before the set is used again).
  // The verdict bytes of a slice are expanded into the caller's array by a helper of their own, started the
  // moment their download is queued: it needs neither the job count nor the rows.
  std::thread out_thread[4]; // [0..1] rows (and 32-bit verdicts) out of the bounce buffers, [2..3] verdict bytes
  std::thread *const code_thread = out_thread + 2;
  int out_rc[2] = {NXS_OK, NXS_OK}, code_rc[2] = {NXS_OK, NXS_OK};
  double out_secs[2] = {0, 0}, code_secs[2] = {0, 0};
  struct JoinAll { // (no exit from this function may leave a helper thread running)
    std::thread *t;
    ~JoinAll() {
      for (int b = 0; b < 4; ++b)
        if (t[b].joinable()) t[b].join();
    }
  } join_all{out_thread};
  auto start_expand = [&](int b, uint64_t first, uint64_t count) {
    uint32_t *dst = out->verdict + first;
    auto body = [ctx, b, dst, count, &code_rc, &code_secs, now] {
      if (cudaSetDevice(ctx->device) != cudaSuccess || cudaEventSynchronize(ctx->e_codes[b]) != cudaSuccess) {
        code_rc[b] = NXS_ERR_CUDA;
        return;
      }
      const double t1 = now();
      expand_codes(ctx->h_codes[b], count, ctx->verdict_of, dst);
      code_secs[b] += now() - t1;
    };
    try {
      code_thread[b] = std::thread(body);
    } catch (...) { // no thread to be had: do it here (slower, not wrong)
      body();
    }
  };
  auto start_copy_out = [&](int b) {
    if (!out_wait[b] || !bounce_out) return;
    uint32_t *dst_v = out->verdict ? out->verdict + out_lo[b] : nullptr;
    nxs_job_row *dst_r = out->job_rows + out_row0[b];
    const uint64_t n_v = out_n[b], n_r = out_jobs[b];
    auto body = [ctx, b, dst_v, dst_r, n_v, n_r, &out_rc, &out_secs, now, codes, bounce_out] {
      if (cudaSetDevice(ctx->device) != cudaSuccess || cudaEventSynchronize(ctx->e_d2h[b]) != cudaSuccess) {
        out_rc[b] = NXS_ERR_CUDA;
        return;
      }
      const double t1 = now();
      CopySeg seg[2];
      int ns = 0;
      if (bounce_out && dst_v && !codes) seg[ns++] = CopySeg{dst_v, ctx->h_verdict[b], n_v * 4};
      if (bounce_out) seg[ns++] = CopySeg{dst_r, ctx->h_rows[b], n_r * sizeof(nxs_job_row)};
      if (ns) par_copy(seg, ns);
      out_secs[b] += now() - t1;
    };
    try {
      out_thread[b] = std::thread(body);
    } catch (...) {
      body();
    }
  };
  auto copy_out = [&](int b) -> int { // wait for the helpers of set b
    if (!out_wait[b]) return NXS_OK;
    const double t0 = now();
    if (out_thread[b].joinable()) out_thread[b].join();
    if (code_thread[b].joinable()) code_thread[b].join();
    t_wait_d2h += now() - t0;
    out_wait[b] = false;
    if (out_rc[b] != NXS_OK || code_rc[b] != NXS_OK) {
      snprintf(ctx->err, sizeof ctx->err, "copying a slice's results to the caller's buffers failed");
      const int r = out_rc[b] != NXS_OK ? out_rc[b] : code_rc[b];
      out_rc[b] = code_rc[b] = NXS_OK;
      return r;
    }
    return NXS_OK;
  };

  for (int k = 0; lo < n; ++k) {
    const int b = k & 1;
    uint64_t hi = n - lo <= slice_cap ? n : lo + kHostSliceRecords;
    while (hi < n && in->job_key[hi] == in->job_key[hi - 1] && hi - lo < cap) ++hi; // cut on a job boundary
    if (hi < n && in->job_key[hi] == in->job_key[hi - 1]) {
      snprintf(ctx->err, sizeof ctx->err, "a single job spans more than %llu records: pass the batch to "
               "nxs_classify_device instead", (unsigned long long)(cap - kHostSliceRecords));
      return NXS_ERR_CAPACITY;
    }
    const uint64_t m = hi - lo;
    if (pending[b]) { // buffer set b still holds the slice before last: get its rows out first
      rc = finish_slice(b);
      if (rc != NXS_OK) return rc;
      start_copy_out(b);
    }
    uint8_t *base = ctx->d_in[b];
    uint64_t *d_key = reinterpret_cast<uint64_t *>(base);
    uint16_t *d_reason = reinterpret_cast<uint16_t *>(base + cap * 8);
    uint8_t *d_kind = base + cap * 10, *d_flags = base + cap * 11, *d_stage = base + cap * 12;
    // where this slice's columns are read from: the caller's buffers, or page-locked copies of them
    const uint64_t *s_key = in->job_key + lo;
    const uint16_t *s_reason = attr ? attr + lo : in->reason_id + lo;
    const uint8_t *s_kind = packed ? nullptr : in->kind + lo, *s_flags = packed ? nullptr : in->flags + lo,
                  *s_stage = packed ? nullptr : in->cur_stage + lo;
    if (bounce_in || autopack)
      NXS_CUDA(ctx, cudaEventSynchronize(ctx->e_h2d[b])); // (the upload of the slice before last has long left the set's host staging)
    if (autopack) { // the four small columns of the slice -> one page-locked u16 column (while the previous slice is on the wire)
      const double t0 = now();
      pack_columns(in->kind + lo, in->reason_id + lo, in->flags + lo, in->cur_stage + lo, m, ctx->h_attr[b]);
      t_in += now() - t0;
      s_reason = ctx->h_attr[b];
    }
    if (bounce_in) {
      uint8_t *hb = ctx->h_in[b];
      CopySeg seg[5] = {{hb, s_key, m * 8}, {hb + cap * 8, s_reason, autopack ? 0 : m * 2}, {hb + cap * 10, s_kind, packed ? 0 : m},
                        {hb + cap * 11, s_flags, packed ? 0 : m}, {hb + cap * 12, s_stage, packed ? 0 : m}};
      const double t0 = now();
      par_copy(seg, 5);
      t_in += now() - t0;
      s_key = reinterpret_cast<const uint64_t *>(hb);
      if (!autopack) s_reason = reinterpret_cast<const uint16_t *>(hb + cap * 8);
      s_kind = hb + cap * 10, s_flags = hb + cap * 11, s_stage = hb + cap * 12;
    }
    // upload (after the previous user of this buffer set has been fully downloaded)
    NXS_CUDA(ctx, cudaStreamWaitEvent(ctx->s_h2d, ctx->e_d2h[b], 0));
    NXS_CUDA(ctx, cudaMemcpyAsync(d_key, s_key, m * 8, cudaMemcpyHostToDevice, ctx->s_h2d));
    NXS_CUDA(ctx, cudaMemcpyAsync(d_reason, s_reason, m * 2, cudaMemcpyHostToDevice, ctx->s_h2d)); // (attr when packed: 10 bytes per record on the wire instead of 13)
    if (!packed) {
      NXS_CUDA(ctx, cudaMemcpyAsync(d_kind, s_kind, m, cudaMemcpyHostToDevice, ctx->s_h2d));
      NXS_CUDA(ctx, cudaMemcpyAsync(d_flags, s_flags, m, cudaMemcpyHostToDevice, ctx->s_h2d));
      NXS_CUDA(ctx, cudaMemcpyAsync(d_stage, s_stage, m, cudaMemcpyHostToDevice, ctx->s_h2d));
    }
    NXS_CUDA(ctx, cudaEventRecord(ctx->e_h2d[b], ctx->s_h2d));
    ctx->wire_h2d += m * (packed ? 10u : 13u);
    // classify
    NXS_CUDA(ctx, cudaStreamWaitEvent(s_comp, ctx->e_h2d[b], 0));
    if (packed) {
      const unsigned blocks = (unsigned)std::min<uint64_t>((m + 255) / 256, (uint64_t)ctx->num_sms * 16);
      expand_attr_kernel<<<blocks, 256, 0, s_comp>>>(d_reason, m, d_reason, d_kind, d_flags, d_stage);
      NXS_CUDA(ctx, cudaGetLastError());
      ctx->launches += 1;
    }
    nxs_batch d_in{d_kind, d_reason, d_flags, d_stage, d_key, m};
    nxs_out d_out{out->verdict && !codes ? ctx->d_verdict[b] : nullptr, ctx->d_rows[b], ctx->d_njobs + b};
    rc = launch_classify(ctx, &d_in, &d_out, (uint32_t)lo, codes ? reinterpret_cast<uint8_t *>(ctx->d_verdict[b]) : nullptr);
    if (rc != NXS_OK) return rc;
    NXS_CUDA(ctx, cudaEventRecord(ctx->e_kernel[b], s_comp));
    // The previous slice's rows go out first (rows land in slice order, and the next upload into its
    // buffer set waits only for them), then this slice's job count and verdicts.
    if (pending[b ^ 1]) {
      rc = finish_slice(b ^ 1);
      if (rc != NXS_OK) return rc;
      start_copy_out(b ^ 1);
    }
    if (helpers) { // the set's previous results must have left its host staging before new ones are queued into it
      rc = copy_out(b);
      if (rc != NXS_OK) return rc;
    }
    NXS_CUDA(ctx, cudaStreamWaitEvent(ctx->s_d2h, ctx->e_kernel[b], 0));
    NXS_CUDA(ctx, cudaMemcpyAsync(ctx->h_njobs + b, ctx->d_njobs + b, 8, cudaMemcpyDeviceToHost, ctx->s_d2h));
    NXS_CUDA(ctx, cudaEventRecord(ctx->e_count[b], ctx->s_d2h));
    if (codes) {
      NXS_CUDA(ctx, cudaMemcpyAsync(ctx->h_codes[b], ctx->d_verdict[b], m, cudaMemcpyDeviceToHost, ctx->s_d2h));
      NXS_CUDA(ctx, cudaEventRecord(ctx->e_codes[b], ctx->s_d2h));
      start_expand(b, lo, m);
      ctx->wire_d2h += m;
    } else if (out->verdict) {
      NXS_CUDA(ctx, cudaMemcpyAsync(bounce_out ? ctx->h_verdict[b] : out->verdict + lo, ctx->d_verdict[b], m * 4,
                                    cudaMemcpyDeviceToHost, ctx->s_d2h));
      ctx->wire_d2h += m * 4;
    }
    pend_lo[b] = lo;
    pend_n[b] = m;
    pending[b] = true;
    if (helpers) out_wait[b] = true, out_lo[b] = lo, out_n[b] = m;
    lo = hi;
  }
  for (int b = 0; b < 2; ++b) {
    // (at most one slice is left, but keep the order general: the older one first)
    const int o = (pending[0] && pending[1] && pend_lo[1] < pend_lo[0]) ? (b ^ 1) : b;
    if (pending[o]) {
      rc = finish_slice(o);
      if (rc != NXS_OK) return rc;
      start_copy_out(o);
    }
  }
  for (int b = 0; b < 2; ++b) {
    rc = copy_out(b);
    if (rc != NXS_OK) return rc;
  }
  NXS_CUDA(ctx, cudaStreamSynchronize(ctx->s_d2h));
  NXS_CUDA(ctx, cudaStreamSynchronize(s_comp));
  if (verbose)
    fprintf(stderr, "nxs host path: pack / copy in %.1f ms, copy out (helper threads) %.1f ms, waiting for job counts %.1f ms, for the helpers %.1f ms\n",
            t_in * 1e3, (out_secs[0] + out_secs[1] + code_secs[0] + code_secs[1]) * 1e3, t_wait_count * 1e3, t_wait_d2h * 1e3);
  *out->n_jobs = rows_done;
  return NXS_OK;
}

static int classify_host(nxs_ctx *ctx, const nxs_batch *in, const uint16_t *attr, const nxs_out *out) {
  DeviceGuard guard(ctx->device);
  NXS_CUDA(ctx, guard.status);
  // A slice may grow past kHostSliceRecords to reach the end of its last job; leave room for that.
  const uint64_t slice_cap = in->n <= kHostSliceRecords ? in->n : kHostSliceRecords + (kHostSliceRecords >> 3);
  int rc = ensure_staging(ctx, slice_cap);
  if (rc != NXS_OK) {
    cudaGetLastError();
    snprintf(ctx->err, sizeof ctx->err, "cannot set up device staging for slices of %llu records",
             (unsigned long long)slice_cap);
    return rc;
  }
  // everything enqueued earlier on the ctx stream must be done before the copy streams touch the staging
  NXS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  rc = classify_slices(ctx, in, attr, out, slice_cap);
  if (rc != NXS_OK) {
    // The caller owns its buffers again the moment this returns: nothing queued for an earlier slice
    // (uploads reading the columns, downloads writing verdicts and rows) may outlive the call.
    cudaStreamSynchronize(ctx->s_h2d);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->s_d2h);
    cudaGetLastError();
  }
  return rc;
}

int nxs_classify(nxs_ctx *ctx, const nxs_batch *in, const nxs_out *out) {
  if (!ctx || !in || !out) return NXS_ERR_ARG;
  if (!out->n_jobs) return fail_arg(ctx, "out->n_jobs is null");
  if (in->n > ctx->max_records) return NXS_ERR_CAPACITY;
  if (in->n == 0) {
    *out->n_jobs = 0;
    return NXS_OK;
  }
  if (!in->kind || !in->reason_id || !in->flags || !in->cur_stage || !in->job_key || !out->job_rows)
    return fail_arg(ctx, "null column pointer");
  return classify_host(ctx, in, nullptr, out);
}

int nxs_classify_packed(nxs_ctx *ctx, const nxs_batch_packed *in, const nxs_out *out) {
  if (!ctx || !in || !out) return NXS_ERR_ARG;
  if (!out->n_jobs) return fail_arg(ctx, "out->n_jobs is null");
  if (in->n > ctx->max_records) return NXS_ERR_CAPACITY;
  if (in->n == 0) {
    *out->n_jobs = 0;
    return NXS_OK;
  }
  if (!in->job_key || !in->attr || !out->job_rows) return fail_arg(ctx, "null column pointer");
  const nxs_batch keys{nullptr, nullptr, nullptr, nullptr, in->job_key, in->n};
  return classify_host(ctx, &keys, in->attr, out);
}

void nxs_last_wire_bytes(const nxs_ctx *ctx, uint64_t *h2d, uint64_t *d2h) {
  if (h2d) *h2d = ctx ? ctx->wire_h2d : 0;
  if (d2h) *d2h = ctx ? ctx->wire_d2h : 0;
}

void nxs_pack_columns(const uint8_t *kind, const uint16_t *reason_id, const uint8_t *flags, const uint8_t *cur_stage,
                      uint64_t n, uint16_t *attr) {
  pack_columns(kind, reason_id, flags, cur_stage, n, attr);
}

int nxs_alloc_host(size_t bytes, void **out) {
  if (!out) return NXS_ERR_ARG;
  *out = nullptr;
  if (cudaMallocHost(out, bytes ? bytes : 1) != cudaSuccess) {
    cudaGetLastError();
    return NXS_ERR_NOMEM;
  }
  return NXS_OK;
}
void nxs_free_host(void *p) {
  if (p) cudaFreeHost(p);
}

} // extern "C"

// ---------------------------------------------------------------------------
// Allocation-free entry point.
//
// For callers that cannot let this library allocate device memory (their own
// pool owns the HBM budget, or a cudaMalloc/cudaFree stall is unacceptable on
// their hot loop).  The kernel's scratch — one 16-byte descriptor per tile plus
// the counters — lives in this file-scope __device__ array instead of a
// cudaMalloc'ed buffer: 8192-record tiles, so NXS_NOALLOC_MAX_RECORDS = 2^24
// records need 2048 descriptors (32 KiB) + 256 bytes for the counters.  Every
// GPU that loads the library carries the array until the process exits.
//
// `g_noalloc_scratch` as written in host code is only a host-side handle for the
// symbol, never a device address: the address each device gave the array is
// looked up once per device with cudaGetSymbolAddress and handed to the kernel.
constexpr int kNoallocTile = NXS_TILE_DEFAULT;
constexpr uint32_t kNoallocMaxTiles = (NXS_NOALLOC_MAX_RECORDS + kNoallocTile - 1) / kNoallocTile;
constexpr size_t kNoallocCounterBytes = 256;
constexpr size_t kNoallocScratchBytes = kNoallocCounterBytes + sizeof(TileDesc) * kNoallocMaxTiles;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocScratchBytes];

namespace {
constexpr int kNoallocMaxDevices = 64;
struct NoallocDevice {
  bool ready = false;
  unsigned char *scratch = nullptr; // device address of g_noalloc_scratch on this device
  uint32_t max_grid = 0;
};
NoallocDevice g_noalloc_dev[kNoallocMaxDevices];
std::mutex g_noalloc_mutex;

// Once per device, with that device current.
int noalloc_setup(int device, NoallocDevice **out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  NoallocDevice &d = g_noalloc_dev[device];
  if (!d.ready) {
    void *addr = nullptr;
    if (cudaGetSymbolAddress(&addr, g_noalloc_scratch) != cudaSuccess || addr == nullptr) {
      cudaGetLastError();
      return NXS_ERR_CUDA;
    }
    const size_t smem = classify_smem_bytes<kNoallocTile>();
    int sms = 0, occ = 0;
    if (cudaFuncSetAttribute(classify_kernel<kNoallocTile>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, classify_kernel<kNoallocTile>, kThreads, smem) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || occ < 1 || sms < 1) {
      cudaGetLastError();
      return NXS_ERR_CUDA;
    }
    d.scratch = static_cast<unsigned char *>(addr);
    d.max_grid = (uint32_t)(sms * occ);
    d.ready = true;
  }
  *out = &d;
  return NXS_OK;
}
} // namespace

extern "C" int nxs_classify_device_noalloc(int device, void *cuda_stream, uint16_t finished_mask,
                                           const nxs_batch *in, const nxs_out *out) {
  if (!in || !out || !out->n_jobs) return NXS_ERR_ARG;
  if (in->n > NXS_NOALLOC_MAX_RECORDS) return NXS_ERR_NOALLOC_LIMIT; // refused before anything is touched
  if (device < 0 || device >= kNoallocMaxDevices) return NXS_ERR_ARG;
  DeviceGuard guard(device); // (the caller's current device is put back on return)
  if (guard.status != cudaSuccess) {
    cudaGetLastError();
    return NXS_ERR_ARG;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  if (in->n == 0) return cudaMemsetAsync(out->n_jobs, 0, sizeof(uint64_t), stream) == cudaSuccess ? NXS_OK : NXS_ERR_CUDA;
  if (!in->kind || !in->reason_id || !in->flags || !in->cur_stage || !in->job_key || !out->job_rows)
    return NXS_ERR_ARG;
  if (!aligned16(in->kind) || !aligned16(in->reason_id) || !aligned16(in->flags) || !aligned16(in->cur_stage) ||
      !aligned16(in->job_key) || !aligned16(out->job_rows) || (reinterpret_cast<uintptr_t>(out->verdict) & 3u))
    return NXS_ERR_ARG;
  NoallocDevice *dev = nullptr;
  int rc = noalloc_setup(device, &dev);
  if (rc != NXS_OK) return rc;

  ClassifyParams p;
  p.kind = in->kind;
  p.reason = in->reason_id;
  p.flags = in->flags;
  p.stage = in->cur_stage;
  p.key = in->job_key;
  p.verdict = out->verdict;
  p.verdict_code = nullptr;
  p.rows = out->job_rows;
  p.n_jobs = out->n_jobs;
  p.counters = reinterpret_cast<uint32_t *>(dev->scratch);
  p.desc = reinterpret_cast<TileDesc *>(dev->scratch + kNoallocCounterBytes);
  p.n = (uint32_t)in->n;
  p.n_tiles = (uint32_t)((in->n + kNoallocTile - 1) / kNoallocTile);
  p.record_base = 0;
  p.finished_mask = finished_mask;
  p.term_mask = 0;
  for (uint32_t a = 1; a <= 4; ++a)
    if ((finished_mask >> new_stage_of(a)) & 1u) p.term_mask |= 1u << a;
  p.gen = 1;
#ifdef NXS_TRACE
  p.trace = nullptr;
  p.trace_cap = 0;
#endif
  nxs_build_lut(p.lut, p.code);
  // The array still holds the previous call's descriptors and counters: the counters and every
  // descriptor this launch will read must start at zero.
  if (cudaMemsetAsync(dev->scratch, 0, kNoallocCounterBytes + sizeof(TileDesc) * (size_t)p.n_tiles, stream) != cudaSuccess) {
    cudaGetLastError();
    return NXS_ERR_CUDA;
  }
  const uint32_t grid = p.n_tiles < dev->max_grid ? p.n_tiles : dev->max_grid;
  classify_kernel<kNoallocTile><<<grid, kThreads, classify_smem_bytes<kNoallocTile>(), stream>>>(p);
  if (cudaGetLastError() != cudaSuccess) return NXS_ERR_CUDA;
  return NXS_OK;
}

#ifdef NXS_TRACE
// Debug build only: copies the event log of the last launch to the host.  Returns the event count.
extern "C" long long nxs_debug_dump_trace(nxs_ctx *ctx, unsigned long long *host, unsigned long long max_events) {
  if (!ctx || !ctx->trace) return -1;
  cudaStreamSynchronize(ctx->stream);
  unsigned long long n = 0;
  cudaMemcpy(&n, ctx->trace, 8, cudaMemcpyDeviceToHost);
  if (n > (1ull << 20)) n = 1ull << 20;
  if (n > max_events) n = max_events;
  cudaMemcpy(host, ctx->trace + 1, n * 16, cudaMemcpyDeviceToHost);
  return (long long)n;
}
#endif

// Internal hooks for the other translation units of the library (not declared
// in the public header).
extern "C" int nxs_internal_ctx_info(nxs_ctx *ctx, int *device, void **stream, char **err, size_t *err_cap) {
  if (!ctx) return NXS_ERR_ARG;
  *device = ctx->device;
  *stream = ctx->stream;
  *err = ctx->err;
  *err_cap = sizeof ctx->err;
  return NXS_OK;
}
extern "C" void nxs_internal_count_launches(nxs_ctx *ctx, uint64_t n) {
  if (ctx) ctx->launches += n;
}
