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

// Final pass: write every entry above the threshold, plus exactly k3 entries equal to it (first
// come first served -- which ties win is unspecified, as with torch.topk).  Two sweeps over the
// block's slice: count, reserve a contiguous output range with ONE global atomic per block, then
// write (a single output counter hammered by every warp serialises at the L2).
__global__ void __launch_bounds__(kSelThreads)
sel_emit_kernel(const float* __restrict__ v, int64_t n, CebSelectScratch* __restrict__ s,
                int32_t* __restrict__ out_idx, float* __restrict__ out_val, uint32_t k) {
  __shared__ uint32_t blk_count, blk_base, blk_fill;
  const uint32_t T = (s->b1 << 19) | (s->b2 << 8) | s->b3;
  const uint32_t need = s->k3;
  const int lane = threadIdx.x & 31;
  // contiguous slice per block so both sweeps hit the same (L2-resident) data
  const int64_t per_block = ((n + gridDim.x - 1) / gridDim.x + 31) & ~(int64_t)31;
  const int64_t begin = (int64_t)blockIdx.x * per_block;
  const int64_t end = begin + per_block < n ? begin + per_block : n;
  if (threadIdx.x == 0) { blk_count = 0; blk_fill = 0; }
  __syncthreads();
  uint32_t mine = 0;
  for (int64_t i = begin + threadIdx.x; i < end; i += blockDim.x) mine += abs_key(__ldg(v + i)) > T;
  mine = __reduce_add_sync(0xffffffffu, mine);
  if (lane == 0 && mine) atomicAdd(&blk_count, mine);
  __syncthreads();
  if (threadIdx.x == 0) blk_base = blk_count ? atomicAdd(&s->out_count, blk_count) : 0u;
  __syncthreads();
  const uint32_t base = blk_base;
  for (int64_t i0 = begin + (threadIdx.x - lane); i0 < end; i0 += blockDim.x) {
    const int64_t i = i0 + lane;
    float val = 0.f;
    uint32_t key = 0;
    const bool valid = i < end;
    if (valid) {
      val = __ldg(v + i);
      key = abs_key(val);
    }
    const bool gt = valid && key > T;
    const unsigned m = __ballot_sync(0xffffffffu, gt);
    if (m) {
      uint32_t wbase = 0;
      const int leader = __ffs(m) - 1;
      if (lane == leader) wbase = atomicAdd(&blk_fill, (uint32_t)__popc(m));
      wbase = __shfl_sync(0xffffffffu, wbase, leader);
      if (gt) {
        const uint32_t pos = base + wbase + __popc(m & ((1u << lane) - 1u));
        if (pos < k) { out_idx[pos] = (int32_t)i; out_val[pos] = val; }
      }
    }
    // ties at the threshold: few (or all zeros); each takes a ticket until the quota is gone
    if (valid && key == T && *((volatile uint32_t*)&s->tie_taken) < need) {
      const uint32_t ticket = atomicAdd(&s->tie_taken, 1u);
      if (ticket < need) {
        const uint32_t pos = atomicAdd(&s->out_count, 1u);
        if (pos < k) { out_idx[pos] = (int32_t)i; out_val[pos] = (key == 0u) ? 0.f : val; }
      }
    }
  }
}

int ceb_select_run(const float* v, int64_t n, int64_t k, CebSelectScratch* s, int32_t* out_idx,
                   float* out_val, bool pass1_done, cudaStream_t stream) {
  const int grid = ceb_grid_for(n, kSelThreads, 4);
  if (!pass1_done) {
    CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), stream));
    sel_hist_kernel<1><<<grid, kSelThreads, 0, stream>>>(v, n, s, (uint32_t)k);
    CEB_COUNT_LAUNCH(1);
  }
  sel_hist_kernel<2><<<grid, kSelThreads, 0, stream>>>(v, n, s, (uint32_t)k);
  sel_hist_kernel<3><<<grid, kSelThreads, 0, stream>>>(v, n, s, (uint32_t)k);
  sel_emit_kernel<<<grid, kSelThreads, 0, stream>>>(v, n, s, out_idx, out_val, (uint32_t)k);
  CEB_COUNT_LAUNCH(3);
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
