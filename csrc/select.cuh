// Exact top-k by magnitude: three-digit (12+11+8 bit) MSD radix select over
// key = bits(|v|), entirely on the device (no host round trip, fixed launch
// count).  Replaces the reference's torch.topk(vec**2) + zero-init + two index
// ops (/root/reference/CommEfficient/utils.py:232-252).
//
// Scratch (global memory, caller- or library-provided, must be zeroed at the
// start of each select -- ceb does it with one cudaMemsetAsync):
#pragma once
#include "common.cuh"

constexpr int kSelMaxBlocks = 1024;

struct CebSelectScratch {
  uint32_t hist1[4096];   // digit 1: key >> 19
  uint32_t hist2[2048];   // digit 2: (key >> 8) & 0x7ff, only keys whose digit 1 == b1
  uint32_t hist3[256];    // digit 3: key & 0xff, only keys whose digits 1,2 == b1,b2
  uint32_t done[4];       // finished-block tickets per pass
  uint32_t b1, k1;        // chosen bin of pass 1, how many still to find inside it
  uint32_t b2, k2;
  uint32_t b3, k3;        // k3 = number of keys equal to the threshold to take
  uint32_t out_count;     // (unused by the ordered emit; kept for layout stability)
  uint32_t tie_taken;
  uint32_t pad[4];
  // ordered emit: entries above / equal to the threshold in every block's slice of the vector
  uint32_t blk_gt[kSelMaxBlocks];
  uint32_t blk_eq[kSelMaxBlocks];
};

static_assert(sizeof(CebSelectScratch) % 16 == 0, "scratch must stay 16B aligned");

constexpr int kSelThreads = 512;

__device__ __forceinline__ uint32_t sel_d1(uint32_t key) { return key >> 19; }
__device__ __forceinline__ uint32_t sel_d2(uint32_t key) { return (key >> 8) & 0x7ffu; }
__device__ __forceinline__ uint32_t sel_d3(uint32_t key) { return key & 0xffu; }

// Executed by ONE whole block (the last one to finish a histogram pass): find the
// bin where the suffix count crosses `want`, write it and the remainder.
template <int NBINS>
__device__ void sel_scan_block(const uint32_t* __restrict__ hist, uint32_t want,
                               uint32_t* out_bin, uint32_t* out_rem) {
  __shared__ uint32_t part[kSelThreads];
  constexpr int PER = (NBINS + kSelThreads - 1) / kSelThreads;
  const int tid = threadIdx.x;
  uint32_t local[PER];
  uint32_t sum = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    int bin = tid * PER + j;
    local[j] = bin < NBINS ? __ldcg(hist + bin) : 0u;
    sum += local[j];
  }
  part[tid] = sum;
  __syncthreads();
  // suffix sums over threads (exclusive: counts in strictly higher thread ranges)
  if (tid < 32) {
    // 512 partials -> each of 32 lanes owns 16 consecutive partials
    constexpr int CH = kSelThreads / 32;
    uint32_t s = 0;
    for (int j = 0; j < CH; ++j) s += part[tid * CH + j];
    // exclusive suffix scan across lanes
    uint32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t up = __shfl_down_sync(0xffffffffu, incl, o);
      if (tid + o < 32) incl += up;
    }
    uint32_t above = incl - s;  // sum over higher lanes
    for (int j = CH - 1; j >= 0; --j) {
      uint32_t mine = part[tid * CH + j];
      part[tid * CH + j] = above;  // exclusive suffix for that thread
      above += mine;
    }
  }
  __syncthreads();
  uint32_t above = part[tid];
  if (above < want && above + sum >= want) {
    // crossing bin is inside this thread's range; walk from the top
    uint32_t acc = above;
#pragma unroll
    for (int j = PER - 1; j >= 0; --j) {
      if (acc < want && acc + local[j] >= want) {
        *out_bin = (uint32_t)(tid * PER + j);
        *out_rem = want - acc;
      }
      acc += local[j];
    }
  }
}

// Merge a block-private histogram into the global one; the last block to arrive
// (ticket counter) runs the suffix scan for this pass.  Must be called by every
// thread of every block, after a __syncthreads() that completes `sh`.
template <int PASS>
__device__ __forceinline__ void sel_merge_and_scan(const uint32_t* sh, CebSelectScratch* __restrict__ s,
                                                   uint32_t k) {
  constexpr int NB = PASS == 1 ? 4096 : (PASS == 2 ? 2048 : 256);
  __shared__ bool is_last;
  uint32_t* gh = PASS == 1 ? s->hist1 : (PASS == 2 ? s->hist2 : s->hist3);
  for (int j = threadIdx.x; j < NB; j += blockDim.x) {
    uint32_t c = sh[j];
    if (c) atomicAdd(gh + j, c);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t nblocks = gridDim.x * gridDim.y * gridDim.z;
    uint32_t ticket = atomicAdd(&s->done[PASS], 1u);
    is_last = (ticket == nblocks - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    if (PASS == 1) sel_scan_block<4096>(s->hist1, k, &s->b1, &s->k1);
    if (PASS == 2) sel_scan_block<2048>(s->hist2, s->k1, &s->b2, &s->k2);
    if (PASS == 3) sel_scan_block<256>(s->hist3, s->k2, &s->b3, &s->k3);
  }
}

// Histogram pass `PASS` (1, 2 or 3) over a dense fp32 vector.
template <int PASS>
__global__ void __launch_bounds__(kSelThreads)
sel_hist_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s, uint32_t k) {
  constexpr int NB = PASS == 1 ? 4096 : (PASS == 2 ? 2048 : 256);
  __shared__ uint32_t sh[NB];
  for (int j = threadIdx.x; j < NB; j += blockDim.x) sh[j] = 0;
  uint32_t b1 = 0, b2 = 0;
  if (PASS >= 2) b1 = s->b1;
  if (PASS >= 3) b2 = s->b2;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t key = abs_key(__ldg(v + i));
    if (PASS == 1) {
      atomicAdd(&sh[sel_d1(key)], 1u);
    } else if (PASS == 2) {
      if (sel_d1(key) == b1) atomicAdd(&sh[sel_d2(key)], 1u);
    } else {
      if (sel_d1(key) == b1 && sel_d2(key) == b2) atomicAdd(&sh[sel_d3(key)], 1u);
    }
  }
  __syncthreads();
  sel_merge_and_scan<PASS>(sh, s, k);
}

// Contiguous per-block slices of the vector (the ordered emit and its counting pass must agree on them).
struct SelSlice { int64_t begin, end; };
__device__ __forceinline__ SelSlice sel_slice(int64_t n) {
  const int64_t per_block = ((n + gridDim.x - 1) / gridDim.x + 3) & ~(int64_t)3;
  SelSlice s;
  s.begin = (int64_t)blockIdx.x * per_block;
  s.end = s.begin + per_block < n ? s.begin + per_block : n;
  if (s.begin > n) s.begin = n;
  return s;
}
__device__ __forceinline__ void sel_load4(const float* __restrict__ v, int64_t i, int64_t end, bool vec, float (&x)[4]) {
  if (vec && i + 3 < end) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(v + i));
    x[0] = t.x; x[1] = t.y; x[2] = t.z; x[3] = t.w;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = (i + j < end) ? __ldg(v + i + j) : 0.f;
  }
}

// Host-side driver shared by ceb_topk_select and the unsketch path.  When
// `first_pass_done` the caller already accumulated hist1 (fused producer) and ran
// the pass-1 scan itself; otherwise all passes run here.
int ceb_select_run(const float* v, int64_t n, int64_t k, CebSelectScratch* s,
                   int32_t* out_idx, float* out_val, bool skip_memset_and_pass1,
                   cudaStream_t stream);
