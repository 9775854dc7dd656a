// nxs_generate.cu — device-side synthetic event-record streams (bench and test
// input, never on the classification path).  Counter-based: record i of a stream
// is a pure function of (workload, seed, i), so a GPU can materialise any range
// or any job-hash shard of it without touching PCIe, and the CPU oracle can
// regenerate the identical records.  The stream definitions are in DESIGN.md
// "Synthetic workloads" (they follow SURVEY.md §8(d)).
#include <cstdio>
#include <vector>

#include "nxs_host.h"

namespace {

struct Rec {
  uint64_t key;
  uint16_t reason;
  uint8_t kind, flags, stage;
};

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// NXS_WORKLOAD_UNIFORM: the 7 handled (kind, reason) pairs + Pod/Scheduled, 1/8
// each; every guard passes; stage BUFFERED or RUNNING; one record per job.
__device__ Rec gen_uniform(uint64_t seed, uint64_t i) {
  const uint64_t r = mix64(seed ^ i);
  const uint32_t pair = (uint32_t)r & 7u;
  Rec o;
  o.kind = pair < 4 ? NXS_KIND_JOB : NXS_KIND_POD;
  o.reason = (uint16_t)(pair + 1);
  o.flags = NXS_FLAG_SUPERVISED | NXS_FLAG_FOUND | NXS_FLAG_CHECKPOINT;
  o.stage = ((r >> 8) & 1) ? NXS_STAGE_RUNNING : NXS_STAGE_BUFFERED;
  o.key = mix64(r);
  return o;
}

// NXS_WORKLOAD_SKEWED: jobs of 1 + Geometric(0.4) records capped at 8 (a new job
// starts at a record with probability 0.4, and at every multiple of 61 so the
// search for a job's first record is bounded); 90 % healthy records, 10 %
// failures with a Zipf(1.2) tail; 2 % of jobs already finished; 0.1 % of jobs
// without checkpoint; 0.1 % stale events.
__device__ __forceinline__ bool starts_job(uint64_t seed, uint64_t i) {
  return (i % 61u) == 0 || (uint32_t)(mix64(seed ^ i) & 0xFFFFu) < 26214u;
}
__device__ Rec gen_skewed(uint64_t seed, uint64_t i) {
  const uint64_t c = mix64(seed ^ i);
  uint64_t j = i;
  while (!starts_job(seed, j)) --j;
  const uint64_t first = j + ((i - j) & ~7ull);
  const uint64_t key = mix64(mix64(seed ^ first) ^ 0xA5A5A5A5A5A5A5A5ull);
  Rec o;
  o.key = key;
  const uint32_t u = (uint32_t)(c >> 16) & 0xFFFFu, u2 = (uint32_t)(c >> 32) & 0xFFFFu, u3 = (uint32_t)(c >> 48);
  if (u < 58982u) {
    if (u2 < 21845u) {
      o.kind = NXS_KIND_POD;
      o.reason = NXS_REASON_STARTED;
    } else {
      o.reason = (uint16_t)(NXS_REASON_SCHEDULED + u2 % 7u);
      o.kind = o.reason >= NXS_REASON_SUCCESSFUL_CREATE ? NXS_KIND_JOB : NXS_KIND_POD;
    }
  } else if (u3 < 30427u) {
    o.kind = NXS_KIND_POD, o.reason = NXS_REASON_BACKOFF;
  } else if (u3 < 43672u) {
    o.kind = NXS_KIND_POD, o.reason = NXS_REASON_FAILED;
  } else if (u3 < 51813u) {
    o.kind = NXS_KIND_JOB, o.reason = NXS_REASON_POD_FAILURE_POLICY;
  } else if (u3 < 57580u) {
    o.kind = NXS_KIND_JOB, o.reason = NXS_REASON_BACKOFF_LIMIT_EXCEEDED;
  } else if (u3 < 61991u) {
    o.kind = NXS_KIND_JOB, o.reason = NXS_REASON_DEADLINE_EXCEEDED;
  } else {
    o.kind = NXS_KIND_JOB, o.reason = NXS_REASON_FAILED_CREATE;
  }
  if (((uint32_t)(key >> 8) & 0xFFFFu) < 1311u) {
    const uint32_t pick = (uint32_t)((key >> 24) % 5u);
    o.stage = pick < 4 ? (uint8_t)(NXS_STAGE_COMPLETED + pick) : (uint8_t)NXS_STAGE_CANCELLED;
  } else {
    o.stage = ((key >> 32) & 1) ? NXS_STAGE_RUNNING : NXS_STAGE_BUFFERED;
  }
  uint8_t fl = NXS_FLAG_SUPERVISED | NXS_FLAG_FOUND | NXS_FLAG_CHECKPOINT;
  if (((key >> 40) & 0x3FFu) == 0) fl &= (uint8_t)~NXS_FLAG_CHECKPOINT;
  if (((c >> 40) & 0x3FFu) == 0) fl &= (uint8_t)~NXS_FLAG_FOUND;
  o.flags = fl;
  return o;
}

__device__ __forceinline__ Rec gen(int workload, uint64_t seed, uint64_t i) {
  return workload == NXS_WORKLOAD_UNIFORM ? gen_uniform(seed, i) : gen_skewed(seed, i);
}
__device__ __forceinline__ void put(const nxs_columns &c, uint64_t at, const Rec &r) {
  c.kind[at] = r.kind;
  c.reason_id[at] = r.reason;
  c.flags[at] = r.flags;
  c.cur_stage[at] = r.stage;
  c.job_key[at] = r.key;
}

__global__ void gen_all_kernel(int workload, uint64_t seed, uint64_t first, uint64_t n, nxs_columns cols) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    put(cols, i, gen(workload, seed, first + i));
}

constexpr int kGenThreads = 256;
constexpr int kGenItems = 16; // records per thread, so a block owns 4096 consecutive records

// Sharded generation, pass 1: how many records of each block's range fall in `shard`.
__global__ void gen_count_kernel(int workload, uint64_t seed, uint64_t first, uint64_t n, uint32_t shard,
                                 uint32_t n_shards, uint32_t *block_counts) {
  __shared__ uint32_t total;
  if (threadIdx.x == 0) total = 0;
  __syncthreads();
  const uint64_t base = (uint64_t)blockIdx.x * (kGenThreads * kGenItems);
  uint32_t mine = 0;
  for (int it = 0; it < kGenItems; ++it) {
    const uint64_t i = base + (uint64_t)it * kGenThreads + threadIdx.x;
    if (i < n) mine += (mix64(gen(workload, seed, first + i).key) % n_shards) == shard;
  }
  atomicAdd(&total, mine);
  __syncthreads();
  if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

// pass 2: regenerate and write the kept records at block_offsets[block] + rank
// in stream order.
__global__ void gen_shard_kernel(int workload, uint64_t seed, uint64_t first, uint64_t n, uint32_t shard,
                                 uint32_t n_shards, const uint64_t *block_offsets, nxs_columns cols) {
  __shared__ uint32_t warp_counts[kGenThreads / 32];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t base = (uint64_t)blockIdx.x * (kGenThreads * kGenItems);
  uint64_t at = block_offsets[blockIdx.x];
  for (int it = 0; it < kGenItems; ++it) {
    const uint64_t i = base + (uint64_t)it * kGenThreads + threadIdx.x;
    Rec r;
    bool keep = false;
    if (i < n) {
      r = gen(workload, seed, first + i);
      keep = (mix64(r.key) % n_shards) == shard;
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_counts[warp] = __popc(ballot);
    __syncthreads();
    uint32_t before = 0, all = 0;
    for (uint32_t w = 0; w < kGenThreads / 32; ++w) {
      const uint32_t cnt = warp_counts[w];
      before += w < warp ? cnt : 0;
      all += cnt;
    }
    if (keep && at + before + __popc(ballot & ((1u << lane) - 1u)) < cols.capacity)
      put(cols, at + before + __popc(ballot & ((1u << lane) - 1u)), r);
    at += all;
    __syncthreads();
  }
}

} // namespace

// Defined in nxs_classify.cu
extern "C" int nxs_internal_ctx_info(nxs_ctx *ctx, int *device, void **stream, char **err, size_t *err_cap);
extern "C" void nxs_internal_count_launches(nxs_ctx *ctx, uint64_t n);

extern "C" int nxs_generate_device(nxs_ctx *ctx, int workload, uint64_t seed, uint64_t first, uint64_t n,
                                   uint32_t shard, uint32_t n_shards, const nxs_columns *d_cols,
                                   uint64_t *n_kept) {
  int device;
  void *stream_v;
  char *err;
  size_t err_cap;
  if (!ctx || !d_cols || nxs_internal_ctx_info(ctx, &device, &stream_v, &err, &err_cap) != NXS_OK) return NXS_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  auto fail = [&](cudaError_t e, const char *what) {
    snprintf(err, err_cap, "%s: %s", what, cudaGetErrorString(e));
    return NXS_ERR_CUDA;
  };
  if (workload != NXS_WORKLOAD_UNIFORM && workload != NXS_WORKLOAD_SKEWED) return NXS_ERR_ARG;
  if (n_shards == 0) n_shards = 1;
  if (shard >= n_shards) return NXS_ERR_ARG;
  struct Restore { // the caller's current device is put back on return
    int prev = -1;
    Restore() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
    ~Restore() { if (prev >= 0) cudaSetDevice(prev); }
  } restore;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(e, "cudaSetDevice");
  if (n == 0) {
    if (n_kept) *n_kept = 0;
    return NXS_OK;
  }
  if (n_shards == 1) {
    if (n > d_cols->capacity) return NXS_ERR_CAPACITY;
    const uint64_t want = (n + kGenThreads - 1) / kGenThreads;
    const unsigned blocks = (unsigned)(want < 148ull * 64 ? want : 148ull * 64);
    gen_all_kernel<<<blocks, kGenThreads, 0, stream>>>(workload, seed, first, n, *d_cols);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "gen_all_kernel");
    nxs_internal_count_launches(ctx, 1);
    if (n_kept) *n_kept = n;
    return NXS_OK;
  }
  const uint64_t per_block = (uint64_t)kGenThreads * kGenItems;
  const uint64_t blocks64 = (n + per_block - 1) / per_block;
  if (blocks64 > 0x7FFFFFFFull) return NXS_ERR_CAPACITY;
  const unsigned blocks = (unsigned)blocks64;
  uint32_t *d_counts = nullptr;
  uint64_t *d_offsets = nullptr;
  if ((e = cudaMalloc(&d_counts, blocks * sizeof(uint32_t))) != cudaSuccess) return fail(e, "cudaMalloc");
  if ((e = cudaMalloc(&d_offsets, blocks * sizeof(uint64_t))) != cudaSuccess) {
    cudaFree(d_counts);
    return fail(e, "cudaMalloc");
  }
  int rc = NXS_OK;
  std::vector<uint32_t> counts(blocks);
  std::vector<uint64_t> offsets(blocks);
  do {
    gen_count_kernel<<<blocks, kGenThreads, 0, stream>>>(workload, seed, first, n, shard, n_shards, d_counts);
    if ((e = cudaGetLastError()) != cudaSuccess) { rc = fail(e, "gen_count_kernel"); break; }
    if ((e = cudaMemcpyAsync(counts.data(), d_counts, blocks * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream)) != cudaSuccess) { rc = fail(e, "copy counts"); break; }
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) { rc = fail(e, "sync"); break; }
    uint64_t total = 0;
    for (unsigned b = 0; b < blocks; ++b) {
      offsets[b] = total;
      total += counts[b];
    }
    if (n_kept) *n_kept = total;
    if (total > d_cols->capacity) { rc = NXS_ERR_CAPACITY; break; }
    if ((e = cudaMemcpyAsync(d_offsets, offsets.data(), blocks * sizeof(uint64_t), cudaMemcpyHostToDevice, stream)) != cudaSuccess) { rc = fail(e, "copy offsets"); break; }
    gen_shard_kernel<<<blocks, kGenThreads, 0, stream>>>(workload, seed, first, n, shard, n_shards, d_offsets, *d_cols);
    if ((e = cudaGetLastError()) != cudaSuccess) { rc = fail(e, "gen_shard_kernel"); break; }
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) { rc = fail(e, "sync"); break; }
    nxs_internal_count_launches(ctx, 2);
  } while (false);
  cudaFree(d_counts);
  cudaFree(d_offsets);
  return rc;
}
