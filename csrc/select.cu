// Host drivers for the radix select (see select.cuh) and the error plumbing.
#include "select.cuh"
#include <cstdarg>
#include <string>

unsigned long long g_ceb_launches = 0;
static thread_local std::string g_ceb_error;

void ceb_set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_ceb_error = buf;
}

CEB_API const char* ceb_last_error() { return g_ceb_error.c_str(); }
CEB_API unsigned long long ceb_launch_count() { return g_ceb_launches; }
CEB_API void ceb_reset_launch_count() { g_ceb_launches = 0; }
// Kernels replayed from a captured CUDA graph are not launched through this library again;
// the caller that replays the graph credits them here (n = launches recorded at capture).
CEB_API void ceb_add_launch_count(unsigned long long n) { g_ceb_launches += n; }
CEB_API int ceb_version() { return 1; }
CEB_API int64_t ceb_select_scratch_bytes() { return (int64_t)sizeof(CebSelectScratch); }

// Final pass, deterministic: the output is the k selected entries in ASCENDING INDEX ORDER -- first
// every entry above the threshold (k - k3 of them), then the k3 LOWEST-INDEX entries equal to it.
// Which of several equal-magnitude entries at the threshold win is therefore a function of the data
// alone, never of thread timing: replicas that run the same select on the same vector pick the same
// set (the multi-GPU design relies on that; torch.topk leaves it unspecified).
// Two kernels over identical contiguous per-block slices: (1) count entries above / equal to the
// threshold per block, (2) every block sums the counts of the blocks before it and writes its entries
// at their exact ranks.  Loads are 16-byte vectors; in-block ranks come from a warp scan + a 16-entry
// warp-total table per iteration.
__global__ void __launch_bounds__(kSelThreads)
sel_count_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s) {
  __shared__ uint32_t sh_gt, sh_eq;
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
  if (threadIdx.x == 0) { s->blk_gt[blockIdx.x] = sh_gt; s->blk_eq[blockIdx.x] = sh_eq; }
}

__global__ void __launch_bounds__(kSelThreads)
sel_emit_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s,
                int32_t* __restrict__ out_idx, float* __restrict__ out_val, uint32_t k, int64_t idx_offset) {
  __shared__ uint32_t base_sh[2];
  __shared__ uint32_t wtot[2][kSelThreads / 32];      // per-warp (gt | eq << 16) totals, double buffered
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
    if (lane == 0) { base_sh[0] = g; base_sh[1] = e; }
  }
  __syncthreads();
  uint32_t run_gt = base_sh[0], run_eq = base_sh[1];
  int buf = 0;
  // uniform trip count: every thread of the block takes part in every iteration's barrier
  for (int64_t i0 = sl.begin; i0 < sl.end; i0 += (int64_t)blockDim.x * 4, buf ^= 1) {
    const int64_t i = i0 + (int64_t)threadIdx.x * 4;
    float x[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t fg = 0, fe = 0;                           // 4-bit flags
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
    uint32_t excl = before + incl - mine;
    uint32_t pg = run_gt + (excl & 0xffffu), pe = run_eq + (excl >> 16);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (fg & (1u << j)) {
        if (pg < k) { out_idx[pg] = (int32_t)(i + j + idx_offset); out_val[pg] = x[j]; }
        ++pg;
      } else if (fe & (1u << j)) {
        if (pe < need) {
          const uint32_t pos = first_tie_pos + pe;
          if (pos < k) { out_idx[pos] = (int32_t)(i + j + idx_offset); out_val[pos] = (T == 0u) ? 0.f : x[j]; }
        }
        ++pe;
      }
    }
    run_gt += total & 0xffffu;
    run_eq += total >> 16;
  }
}

int ceb_select_run(const float* v, int64_t n, int64_t k, CebSelectScratch* s, int32_t* out_idx,
                   float* out_val, bool pass1_done, cudaStream_t stream) {
  int grid = ceb_grid_for(n, kSelThreads, 4);
  if (grid > kSelMaxBlocks) grid = kSelMaxBlocks;
  if (!pass1_done) {
    CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), stream));
    sel_hist_kernel<1><<<grid, kSelThreads, 0, stream>>>(v, n, s, (uint32_t)k);
    CEB_COUNT_LAUNCH(1);
  }
  sel_hist_kernel<2><<<grid, kSelThreads, 0, stream>>>(v, n, s, (uint32_t)k);
  sel_hist_kernel<3><<<grid, kSelThreads, 0, stream>>>(v, n, s, (uint32_t)k);
  const int egrid = (int)(ceb_cdiv(n, kSelThreads * 4) < grid ? ceb_cdiv(n, kSelThreads * 4) : grid);
  sel_count_kernel<<<egrid, kSelThreads, 0, stream>>>(v, n, s);
  sel_emit_kernel<<<egrid, kSelThreads, 0, stream>>>(v, n, s, out_idx, out_val, (uint32_t)k, 0);
  CEB_COUNT_LAUNCH(4);
  CEB_CHECK_LAUNCH();
  return 0;
}

// out_idx/out_val receive exactly k entries (k <= n), unordered.
CEB_API int ceb_topk_select(const float* vec, int64_t n, int64_t k, void* scratch,
                            int64_t scratch_bytes, int32_t* out_idx, float* out_val, void* stream) {
  CEB_REQUIRE(vec && out_idx && out_val && scratch, "ceb_topk_select: null pointer");
  CEB_REQUIRE(n > 0 && n < (int64_t)1 << 31, "ceb_topk_select: n=%lld out of range", (long long)n);
  CEB_REQUIRE(k > 0 && k <= n, "ceb_topk_select: need 0 < k <= n (k=%lld n=%lld)", (long long)k, (long long)n);
  CEB_REQUIRE(scratch_bytes >= (int64_t)sizeof(CebSelectScratch),
              "ceb_topk_select: scratch too small (%lld < %lld bytes)", (long long)scratch_bytes,
              (long long)sizeof(CebSelectScratch));
  return ceb_select_run(vec, n, k, (CebSelectScratch*)scratch, out_idx, out_val, false,
                        (cudaStream_t)stream);
}
