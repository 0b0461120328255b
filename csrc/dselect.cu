// Distributed exact top-k: every rank holds one slice of a vector (its shard of the true_topk error vector,
// or the median estimates of its coordinate range of the sketch); the k largest-magnitude entries of the WHOLE
// vector are found without ever gathering it.
//
//   per radix pass (12 + 11 + 8 bits of |v|): each rank histograms its slice; the last block of that kernel
//   publishes the histogram in this rank's symmetric buffer, crosses a flag barrier in peer memory, sums
//   the W histograms (peer loads, W x 16 KB) and runs the suffix scan -- every rank arrives at the same
//   (bin, remainder) without a host round trip or a collective library call;
//   then each rank counts its entries above / equal to the threshold, the counts are exchanged the same way
//   (rank r's entries are placed after those of ranks < r: the selection is a function of the data alone),
//   and every rank writes its winners straight into EVERY rank's (idx, val) arrays with peer stores.
//
// The result is the same list, bit for bit, on all ranks (each element is written by exactly one owner).
// Reference rule: the server-side top-k of /root/reference/CommEfficient/fed_aggregator.py:511-542 (true_topk)
// and :584-611 (sketch), computed there on one GPU over the full vector.
#include "peer.cuh"
#include "select.cuh"

constexpr int kDselPhases = 5;      // barrier slots: three histogram passes, the counts, the final "lists complete"

struct CebDselExch {                // one per rank, in symmetric memory; written by its owner, read by peers
  uint32_t hist[3][4096];
  uint32_t gt, eq;
  uint32_t pad[62];
};

struct DselCtx {
  CebDselExch* exch[kMaxPeers];
  uint32_t* flags[kMaxPeers];       // [kDselPhases][kFlagStride]
  int32_t* out_idx[kMaxPeers];
  float* out_val[kMaxPeers];
  uint32_t* err;                    // local error word (see peer.cuh)
  int rank, world;
  // The barrier epoch lives in DEVICE memory (bumped by the first kernel of every call) so that a call
  // captured in a CUDA graph crosses fresh barriers on every replay; all ranks make the same calls in the
  // same order, so their counters agree.
  const uint32_t* epoch_dev;
};

__global__ void dsel_bump_kernel(uint32_t* epoch_dev) { *epoch_dev += 1u; }

__device__ __forceinline__ void dsel_barrier(const DselCtx& c, int phase) {
  // executed by one whole block
  __threadfence_system();
  __syncthreads();
  const uint32_t epoch = *c.epoch_dev;
  if ((int)threadIdx.x < c.world) st_release_sys(c.flags[threadIdx.x] + phase * kFlagStride + c.rank, epoch);
  if ((int)threadIdx.x < c.world) {
    const uint32_t* slot = c.flags[c.rank] + phase * kFlagStride + threadIdx.x;
    unsigned long long polls = 0;
    const unsigned long long t0 = peer_now_ns();
    while (!epoch_reached(ld_acquire_sys(slot), epoch)) {
      __nanosleep(20);
      if (peer_spin_expired(polls, t0)) {
        atomicExch(c.err, 0x40000000u | ((uint32_t)phase << 8) | (uint32_t)threadIdx.x);
        break;
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ uint32_t ld_peer_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Executed by one whole block once this rank's histogram of `PASS` is complete in s->hist*: publish it, cross
// the barrier, replace it by the sum over ranks, scan.
template <int PASS>
__device__ void dsel_merge_block(CebSelectScratch* __restrict__ s, const DselCtx& c, uint32_t k) {
  constexpr int NB = PASS == 1 ? 4096 : (PASS == 2 ? 2048 : 256);
  uint32_t* gh = PASS == 1 ? s->hist1 : (PASS == 2 ? s->hist2 : s->hist3);
  uint32_t* mine = c.exch[c.rank]->hist[PASS - 1];
  for (int j = threadIdx.x; j < NB; j += blockDim.x) mine[j] = __ldcg(gh + j);
  dsel_barrier(c, PASS - 1);
  for (int j = threadIdx.x; j < NB; j += blockDim.x) {
    uint32_t sum = 0;
    for (int q = 0; q < c.world; ++q) sum += ld_peer_u32(c.exch[q]->hist[PASS - 1] + j);
    gh[j] = sum;
  }
  __threadfence();
  __syncthreads();
  if (PASS == 1) sel_scan_block<4096>(s->hist1, k, &s->b1, &s->k1);
  if (PASS == 2) sel_scan_block<2048>(s->hist2, s->k1, &s->b2, &s->k2);
  if (PASS == 3) sel_scan_block<256>(s->hist3, s->k2, &s->b3, &s->k3);
}

// Pass-1 merge when a fused producer (the range unsketch) already accumulated this rank's digit-1 histogram.
__global__ void __launch_bounds__(kSelThreads)
dsel_merge1_kernel(CebSelectScratch* __restrict__ s, DselCtx c, uint32_t k) { dsel_merge_block<1>(s, c, k); }

// Histogram pass over this rank's slice; the last block to finish merges across ranks.
template <int PASS>
__global__ void __launch_bounds__(kSelThreads)
dsel_hist_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s, DselCtx c, uint32_t k) {
  constexpr int NB = PASS == 1 ? 4096 : (PASS == 2 ? 2048 : 256);
  __shared__ uint32_t sh[NB];
  __shared__ bool is_last;
  for (int j = threadIdx.x; j < NB; j += blockDim.x) sh[j] = 0;
  uint32_t b1 = 0, b2 = 0;
  if (PASS >= 2) b1 = s->b1;
  if (PASS >= 3) b2 = s->b2;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t key = abs_key(__ldg(v + i));
    if (PASS == 1) atomicAdd(&sh[sel_d1(key)], 1u);
    else if (PASS == 2) { if (sel_d1(key) == b1) atomicAdd(&sh[sel_d2(key)], 1u); }
    else { if (sel_d1(key) == b1 && sel_d2(key) == b2) atomicAdd(&sh[sel_d3(key)], 1u); }
  }
  __syncthreads();
  uint32_t* gh = PASS == 1 ? s->hist1 : (PASS == 2 ? s->hist2 : s->hist3);
  for (int j = threadIdx.x; j < NB; j += blockDim.x) {
    const uint32_t cnt = sh[j];
    if (cnt) atomicAdd(gh + j, cnt);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&s->done[PASS], 1u) == gridDim.x - 1;
  __syncthreads();
  if (is_last) {
    __threadfence();
    dsel_merge_block<PASS>(s, c, k);
  }
}

// Entries above / equal to the threshold per block of this rank's slice; the last block exchanges the rank
// totals and leaves this rank's global bases in s->out_count (above) / s->tie_taken (ties).
__global__ void __launch_bounds__(kSelThreads)
dsel_count_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s, DselCtx c) {
  __shared__ uint32_t sh_gt, sh_eq;
  __shared__ bool is_last;
  const uint32_t T = (s->b1 << 19) | (s->b2 << 8) | s->b3;
  const SelSlice sl = sel_slice(n);
  const bool vec = (((uintptr_t)v) & 15) == 0;
  if (threadIdx.x == 0) { sh_gt = 0; sh_eq = 0; }
  __syncthreads();
  uint32_t gt = 0, eq = 0;
  for (int64_t i = sl.begin + (int64_t)threadIdx.x * 4; i < sl.end; i += (int64_t)blockDim.x * 4) {
    float x[4];
    sel_load4(v, i, sl.end, vec, x);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (i + j < sl.end) {
        const uint32_t key = abs_key(x[j]);
        gt += key > T;
        eq += key == T;
      }
    }
  }
  gt = __reduce_add_sync(0xffffffffu, gt);
  eq = __reduce_add_sync(0xffffffffu, eq);
  if ((threadIdx.x & 31) == 0) {
    if (gt) atomicAdd(&sh_gt, gt);
    if (eq) atomicAdd(&sh_eq, eq);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    s->blk_gt[blockIdx.x] = sh_gt;
    s->blk_eq[blockIdx.x] = sh_eq;
    __threadfence();
    is_last = atomicAdd(&s->done[0], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (threadIdx.x < 32) {
    uint32_t g = 0, e = 0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += 32) { g += __ldcg(s->blk_gt + b); e += __ldcg(s->blk_eq + b); }
    g = __reduce_add_sync(0xffffffffu, g);
    e = __reduce_add_sync(0xffffffffu, e);
    if (threadIdx.x == 0) { c.exch[c.rank]->gt = g; c.exch[c.rank]->eq = e; }
  }
  dsel_barrier(c, 3);
  if (threadIdx.x == 0) {
    uint32_t bg = 0, be = 0;
    for (int q = 0; q < c.rank; ++q) { bg += ld_peer_u32(&c.exch[q]->gt); be += ld_peer_u32(&c.exch[q]->eq); }
    s->out_count = bg;
    s->tie_taken = be;
  }
}

__device__ __forceinline__ void st_peer_u32(void* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Ordered emit into every rank's lists: entries above the threshold at [0, k - k3) in (rank, index) order,
// then the k3 lowest-(rank, index) entries equal to it.
__global__ void __launch_bounds__(kSelThreads)
dsel_emit_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s, DselCtx c, uint32_t k,
                 int64_t idx_offset) {
  __shared__ uint32_t base_sh[2];
  __shared__ uint32_t wtot[2][kSelThreads / 32];
  __shared__ bool is_last;
  const uint32_t T = (s->b1 << 19) | (s->b2 << 8) | s->b3;
  const uint32_t need = s->k3;
  const uint32_t first_tie_pos = k - need;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const SelSlice sl = sel_slice(n);
  const bool vec = (((uintptr_t)v) & 15) == 0;
  if (warp == 0) {
    uint32_t g = 0, e = 0;
    for (int b = lane; b < (int)blockIdx.x; b += 32) { g += s->blk_gt[b]; e += s->blk_eq[b]; }
    g = __reduce_add_sync(0xffffffffu, g);
    e = __reduce_add_sync(0xffffffffu, e);
    if (lane == 0) { base_sh[0] = g + s->out_count; base_sh[1] = e + s->tie_taken; }
  }
  __syncthreads();
  uint32_t run_gt = base_sh[0], run_eq = base_sh[1];
  int buf = 0;
  for (int64_t i0 = sl.begin; i0 < sl.end; i0 += (int64_t)blockDim.x * 4, buf ^= 1) {
    const int64_t i = i0 + (int64_t)threadIdx.x * 4;
    float x[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t fg = 0, fe = 0;
    if (i < sl.end) {
      sel_load4(v, i, sl.end, vec, x);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (i + j < sl.end) {
          const uint32_t key = abs_key(x[j]);
          fg |= (uint32_t)(key > T) << j;
          fe |= (uint32_t)(key == T) << j;
        }
      }
    }
    const uint32_t mine = (uint32_t)__popc(fg) | ((uint32_t)__popc(fe) << 16);
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    if (lane == 31) wtot[buf][warp] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w2 = 0; w2 < kSelThreads / 32; ++w2) {
      const uint32_t t = wtot[buf][w2];
      total += t;
      if (w2 < warp) before += t;
    }
    const uint32_t excl = before + incl - mine;
    uint32_t pg = run_gt + (excl & 0xffffu), pe = run_eq + (excl >> 16);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t pos = 0xffffffffu;
      float val = x[j];
      if (fg & (1u << j)) {
        pos = pg++;
      } else if (fe & (1u << j)) {
        if (pe < need) { pos = first_tie_pos + pe; if (T == 0u) val = 0.f; }
        ++pe;
      }
      if (pos < k) {
        const uint32_t gi = (uint32_t)(i + j + idx_offset);
        for (int q = 0; q < c.world; ++q) {
          st_peer_u32(c.out_idx[q] + pos, gi);
          st_peer_u32(c.out_val[q] + pos, __float_as_uint(val));
        }
      }
    }
    run_gt += total & 0xffffu;
    run_eq += total >> 16;
  }
  // the lists are complete on every rank once every rank's stores have landed
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&s->pad[0], 1u) == gridDim.x - 1;     // (done[1..3]: the histogram passes)
  __syncthreads();
  if (is_last) dsel_barrier(c, 4);
}

CEB_API int64_t ceb_dsel_exch_bytes() { return (int64_t)sizeof(CebDselExch); }
CEB_API int ceb_dsel_flag_words() { return kDselPhases * kFlagStride; }

// v: this rank's slice (n_local >= 0 values; the global index of v[0] is idx_offset); k: how many entries of
// the whole vector to select (<= total length, which the caller guarantees).  exch / flags / out_idx / out_val:
// host arrays of `world` peer-mapped device pointers (symmetric allocations: ceb_dsel_exch_bytes() bytes,
// ceb_dsel_flag_words() uint32, k int32, k floats).  scratch: local, ceb_select_scratch_bytes().  err: local
// uint32 error word.  pass1_done != 0: scratch->hist1 already holds this rank's digit-1 histogram (and the rest
// of the scratch was zeroed before it was built).  epoch_dev: local device uint32, zero at start, owned by this
// entry point (it adds one per call; every rank must make the same sequence of calls).
CEB_API int ceb_dsel_run(const float* v, int64_t n_local, int64_t idx_offset, int64_t k, void* scratch,
                         int64_t scratch_bytes, const void* const* exch, const void* const* flags,
                         const void* const* out_idx, const void* const* out_val, void* err, int rank, int world,
                         void* epoch_dev, int pass1_done, void* stream) {
  CEB_REQUIRE(scratch && exch && flags && out_idx && out_val && err && epoch_dev, "ceb_dsel_run: null pointer");
  CEB_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "ceb_dsel_run: rank %d / world %d", rank, world);
  CEB_REQUIRE(n_local >= 0 && n_local < ((int64_t)1 << 31) && (n_local == 0 || v) && k > 0 && k < ((int64_t)1 << 31),
              "ceb_dsel_run: bad sizes (n_local=%lld k=%lld)", (long long)n_local, (long long)k);
  CEB_REQUIRE(scratch_bytes >= (int64_t)sizeof(CebSelectScratch), "ceb_dsel_run: scratch too small");
  DselCtx c;
  for (int q = 0; q < kMaxPeers; ++q) {
    c.exch[q] = q < world ? (CebDselExch*)exch[q] : nullptr;
    c.flags[q] = q < world ? (uint32_t*)flags[q] : nullptr;
    c.out_idx[q] = q < world ? (int32_t*)out_idx[q] : nullptr;
    c.out_val[q] = q < world ? (float*)out_val[q] : nullptr;
  }
  c.err = (uint32_t*)err; c.rank = rank; c.world = world; c.epoch_dev = (const uint32_t*)epoch_dev;
  cudaStream_t st = (cudaStream_t)stream;
  dsel_bump_kernel<<<1, 1, 0, st>>>((uint32_t*)epoch_dev);
  CebSelectScratch* s = (CebSelectScratch*)scratch;
  const int64_t n = n_local > 0 ? n_local : 1;        // (an empty slice still takes part in every barrier)
  int grid = ceb_grid_for(n, kSelThreads, 4);
  if (grid > kSelMaxBlocks) grid = kSelMaxBlocks;
  if (!pass1_done) {
    CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), st));
    dsel_hist_kernel<1><<<grid, kSelThreads, 0, st>>>(v, n_local, s, c, (uint32_t)k);
  } else {
    dsel_merge1_kernel<<<1, kSelThreads, 0, st>>>(s, c, (uint32_t)k);
  }
  dsel_hist_kernel<2><<<grid, kSelThreads, 0, st>>>(v, n_local, s, c, (uint32_t)k);
  dsel_hist_kernel<3><<<grid, kSelThreads, 0, st>>>(v, n_local, s, c, (uint32_t)k);
  int egrid = (int)ceb_cdiv(n, kSelThreads * 4);
  if (egrid > grid) egrid = grid;
  dsel_count_kernel<<<egrid, kSelThreads, 0, st>>>(v, n_local, s, c);
  dsel_emit_kernel<<<egrid, kSelThreads, 0, st>>>(v, n_local, s, c, (uint32_t)k, idx_offset);
  CEB_COUNT_LAUNCH(6);
  CEB_CHECK_LAUNCH();
  return 0;
}
