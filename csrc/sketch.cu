// Count-sketch kernels (sm_100a).  Replace the external `csvec.CSVec` the reference
// calls at /root/reference/CommEfficient/fed_worker.py:312-320 (accumulateVec,
// sketch-space clip) and fed_aggregator.py:568-613 (momentum/error in sketch space,
// unSketch top-k, bucket zeroing).  Hash layout: see commefficient_b200/ops/hashing.py.
//
//   packed[row][t] : int32, low 31 bits = bucket_r(t), bit 31 = 1 when sign_r(t) = -1
//   blk[b]         : int32, low 31 bits = column offset of hash block b, bit 31 = block sign
//   coordinate i   : b = i / n, t = i - b*n
//
// One thread owns one token t and walks the hash blocks, so the r packed words
// are loaded once per token and reused for every block (20x fewer table reads at
// the default num_blocks=20); gradient reads stay coalesced because consecutive
// threads own consecutive tokens.
#include "select.cuh"

constexpr int kMaxRows = 16;
constexpr int kSketchThreads = 256;

// ------------------------------------------------------------------ accumulate
template <int R>
__global__ void __launch_bounds__(kSketchThreads)
sketch_accum_kernel(const float* __restrict__ g, const float* __restrict__ w, float wd, float scale,
                    int64_t d, const int32_t* __restrict__ packed, int64_t n,
                    const int32_t* __restrict__ blk, int num_blocks, float* __restrict__ table,
                    uint32_t c) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += stride) {
    uint32_t hp[R];
#pragma unroll
    for (int r = 0; r < R; ++r) hp[r] = (uint32_t)__ldg(packed + (int64_t)r * n + t);
    for (int b = blockIdx.y; b < num_blocks; b += gridDim.y) {
      const int64_t i = (int64_t)b * n + t;
      if (i >= d) break;
      float v = __ldg(g + i);
      if (w != nullptr) v = fmaf(wd, __ldg(w + i), v);
      v *= scale;
      if (v == 0.f) continue;
      const uint32_t bp = (uint32_t)__ldg(blk + b);
      const uint32_t off = bp & 0x7fffffffu;
      const uint32_t vb = __float_as_uint(v) ^ (bp & 0x80000000u);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        uint32_t col = (hp[r] & 0x7fffffffu) + off;
        if (col >= c) col -= c;
        red_add_f32(table + (int64_t)r * c + col, __uint_as_float(vb ^ (hp[r] & 0x80000000u)));
      }
    }
  }
}

static inline dim3 token_grid(int64_t n, int num_blocks, int threads) {
  int64_t gx = ceb_cdiv(n, threads);
  int64_t cap = (int64_t)ceb_num_sms() * 16;
  int gy = 1;
  if (gx > cap) gx = cap;
  // small token counts: spread hash blocks over grid.y to keep the SMs busy
  while (gx * gy < (int64_t)ceb_num_sms() * 4 && gy < num_blocks) gy *= 2;
  if (gy > num_blocks) gy = num_blocks;
  return dim3((unsigned)gx, (unsigned)gy, 1);
}

#define CEB_DISPATCH_ROWS(R_, ...)                     \
  switch (R_) {                                        \
    case 1: { constexpr int R = 1; __VA_ARGS__; } break;   \
    case 2: { constexpr int R = 2; __VA_ARGS__; } break;   \
    case 3: { constexpr int R = 3; __VA_ARGS__; } break;   \
    case 4: { constexpr int R = 4; __VA_ARGS__; } break;   \
    case 5: { constexpr int R = 5; __VA_ARGS__; } break;   \
    case 6: { constexpr int R = 6; __VA_ARGS__; } break;   \
    case 7: { constexpr int R = 7; __VA_ARGS__; } break;   \
    case 8: { constexpr int R = 8; __VA_ARGS__; } break;   \
    case 9: { constexpr int R = 9; __VA_ARGS__; } break;   \
    case 10: { constexpr int R = 10; __VA_ARGS__; } break; \
    case 11: { constexpr int R = 11; __VA_ARGS__; } break; \
    case 12: { constexpr int R = 12; __VA_ARGS__; } break; \
    case 13: { constexpr int R = 13; __VA_ARGS__; } break; \
    case 14: { constexpr int R = 14; __VA_ARGS__; } break; \
    case 15: { constexpr int R = 15; __VA_ARGS__; } break; \
    case 16: { constexpr int R = 16; __VA_ARGS__; } break; \
    default: break;                                    \
  }

static int check_sketch_dims(const char* who, int r, int64_t c, int64_t d, int64_t n, int num_blocks) {
  CEB_REQUIRE(r >= 1 && r <= kMaxRows, "%s: num_rows=%d outside [1,%d]", who, r, kMaxRows);
  CEB_REQUIRE(c >= 1 && c < ((int64_t)1 << 30), "%s: num_cols=%lld outside [1,2^30)", who, (long long)c);
  CEB_REQUIRE(d >= 1 && d < ((int64_t)1 << 31), "%s: d=%lld outside [1,2^31)", who, (long long)d);
  CEB_REQUIRE(num_blocks >= 1 && n >= 1 && n * num_blocks >= d,
              "%s: n_tokens*num_blocks (%lld*%d) does not cover d=%lld", who, (long long)n, num_blocks, (long long)d);
  return 0;
}

// table[r][bucket(i)] += sign(i) * scale * (g[i] + wd * w[i])      (w may be null)
CEB_API int ceb_sketch_accumulate(const float* g, const float* w, float wd, float scale, int64_t d,
                                  const int32_t* packed, int64_t n_tokens, const int32_t* blk,
                                  int num_blocks, float* table, int r, int64_t c, void* stream) {
  CEB_REQUIRE(g && packed && blk && table, "ceb_sketch_accumulate: null pointer");
  if (int e = check_sketch_dims("ceb_sketch_accumulate", r, c, d, n_tokens, num_blocks)) return e;
  dim3 grid = token_grid(n_tokens, num_blocks, kSketchThreads);
  CEB_DISPATCH_ROWS(r, (sketch_accum_kernel<R><<<grid, kSketchThreads, 0, (cudaStream_t)stream>>>(
                           g, w, wd, scale, d, packed, n_tokens, blk, num_blocks, table, (uint32_t)c)));
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Coordinates [lo, hi) only, one thread per coordinate (the hash words of its token are loaded per
// coordinate instead of once per 20 hash blocks): lets a caller sketch the part of the gradient that is
// already final on a second stream while the rest of the backward pass is still running -- the
// accumulate is bound by L2 atomics, the backward by the tensor pipe, so they overlap well.
template <int R>
__global__ void __launch_bounds__(kSketchThreads)
sketch_accum_range_kernel(const float* __restrict__ g, const float* __restrict__ w, float wd, float scale,
                          int64_t lo, int64_t hi, const int32_t* __restrict__ packed, int64_t n,
                          const int32_t* __restrict__ blk, float* __restrict__ table, uint32_t c) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
    float v = __ldg(g + i);
    if (w != nullptr) v = fmaf(wd, __ldg(w + i), v);
    v *= scale;
    if (v == 0.f) continue;
    const int b = (int)(i / n);
    const int64_t t = i - (int64_t)b * n;
    const uint32_t bp = (uint32_t)__ldg(blk + b);
    const uint32_t off = bp & 0x7fffffffu;
    const uint32_t vb = __float_as_uint(v) ^ (bp & 0x80000000u);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t hp = (uint32_t)__ldg(packed + (int64_t)r * n + t);
      uint32_t col = (hp & 0x7fffffffu) + off;
      if (col >= c) col -= c;
      red_add_f32(table + (int64_t)r * c + col, __uint_as_float(vb ^ (hp & 0x80000000u)));
    }
  }
}

// table += S(scale * (g + wd * w)) restricted to coordinates lo <= i < hi (g, w indexed by the GLOBAL i).
CEB_API int ceb_sketch_accumulate_range(const float* g, const float* w, float wd, float scale, int64_t d, int64_t lo,
                                        int64_t hi, const int32_t* packed, int64_t n_tokens, const int32_t* blk,
                                        int num_blocks, float* table, int r, int64_t c, void* stream) {
  CEB_REQUIRE(g && packed && blk && table, "ceb_sketch_accumulate_range: null pointer");
  CEB_REQUIRE(0 <= lo && lo <= hi && hi <= d, "ceb_sketch_accumulate_range: bad range [%lld, %lld) for d=%lld", (long long)lo,
              (long long)hi, (long long)d);
  if (int e = check_sketch_dims("ceb_sketch_accumulate_range", r, c, d, n_tokens, num_blocks)) return e;
  if (hi == lo) return 0;
  int64_t gx = ceb_cdiv(hi - lo, kSketchThreads);
  const int64_t cap = (int64_t)ceb_num_sms() * 16;
  if (gx > cap) gx = cap;
  CEB_DISPATCH_ROWS(r, (sketch_accum_range_kernel<R><<<(unsigned)gx, kSketchThreads, 0, (cudaStream_t)stream>>>(
                           g, w, wd, scale, lo, hi, packed, n_tokens, blk, table, (uint32_t)c)));
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ l2 estimate / clip
__global__ void __launch_bounds__(256)
rowsumsq_kernel(const float* __restrict__ table, int64_t c, float* __restrict__ out) {
  const int row = blockIdx.y;
  const float* p = table + (int64_t)row * c;
  float acc = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < c; j += stride) {
    float v = __ldg(p + j);
    acc = fmaf(v, v, acc);
  }
  __shared__ float ws[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? ws[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(out + row, v);
  }
}

// out[0] = min(1, clip / l2) (1 when clip <= 0), out[1] = l2 = sqrt(median_r rowsq)
__global__ void l2_scale_kernel(const float* __restrict__ rowsq, int r, float clip, float* __restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float v[kMaxRows];
  for (int i = 0; i < r; ++i) v[i] = rowsq[i];
  for (int i = 1; i < r; ++i) {
    float x = v[i];
    int j = i - 1;
    while (j >= 0 && v[j] > x) { v[j + 1] = v[j]; --j; }
    v[j + 1] = x;
  }
  float l2 = sqrtf(v[(r - 1) / 2]);
  out[1] = l2;
  out[0] = (clip > 0.f && l2 >= clip) ? clip / l2 : 1.f;
}

// rowsq_tmp: r floats of scratch; out2: two floats {scale, l2}.
CEB_API int ceb_sketch_l2_scale(const float* table, int r, int64_t c, float clip, float* rowsq_tmp,
                                float* out2, void* stream) {
  CEB_REQUIRE(table && rowsq_tmp && out2, "ceb_sketch_l2_scale: null pointer");
  CEB_REQUIRE(r >= 1 && r <= kMaxRows, "ceb_sketch_l2_scale: num_rows=%d outside [1,%d]", r, kMaxRows);
  cudaStream_t st = (cudaStream_t)stream;
  CEB_CHECK_CUDA(cudaMemsetAsync(rowsq_tmp, 0, sizeof(float) * r, st));
  int gx = (int)ceb_cdiv(c, 256 * 8);
  if (gx < 1) gx = 1;
  if (gx > 64) gx = 64;
  rowsumsq_kernel<<<dim3(gx, r, 1), 256, 0, st>>>(table, c, rowsq_tmp);
  l2_scale_kernel<<<1, 32, 0, st>>>(rowsq_tmp, r, clip, out2);
  CEB_COUNT_LAUNCH(2);
  CEB_CHECK_LAUNCH();
  return 0;
}

// y += alpha * (dev_scale ? *dev_scale : 1) * x ; optionally zero x afterwards
__global__ void __launch_bounds__(256)
axpy_dev_kernel(float* __restrict__ y, float* __restrict__ x, int64_t n, float alpha,
                const float* __restrict__ dev_scale, int zero_x) {
  const float a = alpha * (dev_scale ? __ldg(dev_scale) : 1.f);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    y[i] = fmaf(a, x[i], y[i]);
    if (zero_x) x[i] = 0.f;
  }
}

CEB_API int ceb_axpy_dev(float* y, float* x, int64_t n, float alpha, const float* dev_scale,
                         int zero_x, void* stream) {
  CEB_REQUIRE(y && x && n >= 0, "ceb_axpy_dev: bad arguments");
  if (n == 0) return 0;
  axpy_dev_kernel<<<ceb_grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(y, x, n, alpha, dev_scale, zero_x);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ server momentum / error
// Vv = rho*Vv + g*inv_n ; error_mode 2 (virtual): Ve += Vv ; 1 (local): Ve = Vv ; 0: untouched
__global__ void __launch_bounds__(256)
sketch_server_update_kernel(float* __restrict__ Vv, float* __restrict__ Ve, const float* __restrict__ g,
                            float inv_n, float rho, int error_mode, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v = fmaf(rho, Vv[i], __ldg(g + i) * inv_n);
    Vv[i] = v;
    if (error_mode == 2) Ve[i] += v;
    else if (error_mode == 1) Ve[i] = v;
  }
}

CEB_API int ceb_sketch_server_update(float* Vv, float* Ve, const float* g, float inv_n, float rho,
                                     int error_mode, int64_t n, void* stream) {
  CEB_REQUIRE(Vv && g && (error_mode == 0 || Ve), "ceb_sketch_server_update: null pointer");
  CEB_REQUIRE(error_mode >= 0 && error_mode <= 2, "ceb_sketch_server_update: error_mode=%d", error_mode);
  sketch_server_update_kernel<<<ceb_grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(
      Vv, Ve, g, inv_n, rho, error_mode, n);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ unsketch: median estimate + histogram
template <int R>
__device__ __forceinline__ float lower_median(float (&v)[R]) {
#pragma unroll
  for (int i = 1; i < R; ++i) {
#pragma unroll
    for (int j = i; j > 0; --j) {
      float a = v[j - 1], b = v[j];
      v[j - 1] = fminf(a, b);
      v[j] = fmaxf(a, b);
    }
  }
  return v[(R - 1) / 2];
}

template <int R>
__global__ void __launch_bounds__(kSelThreads)
unsketch_est_kernel(const float* __restrict__ table, uint32_t c, const int32_t* __restrict__ packed,
                    int64_t n, const int32_t* __restrict__ blk, int num_blocks, int64_t d,
                    float* __restrict__ est, CebSelectScratch* __restrict__ s, uint32_t k) {
  __shared__ uint32_t sh[4096];
  for (int j = threadIdx.x; j < 4096; j += blockDim.x) sh[j] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += stride) {
    uint32_t hp[R];
#pragma unroll
    for (int r = 0; r < R; ++r) hp[r] = (uint32_t)__ldg(packed + (int64_t)r * n + t);
    for (int b = blockIdx.y; b < num_blocks; b += gridDim.y) {
      const int64_t i = (int64_t)b * n + t;
      if (i >= d) break;
      const uint32_t bp = (uint32_t)__ldg(blk + b);
      const uint32_t off = bp & 0x7fffffffu;
      const uint32_t bs = bp & 0x80000000u;
      float vals[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        uint32_t col = (hp[r] & 0x7fffffffu) + off;
        if (col >= c) col -= c;
        float x = __ldg(table + (int64_t)r * c + col);
        vals[r] = __uint_as_float(__float_as_uint(x) ^ ((hp[r] & 0x80000000u) ^ bs));
      }
      float m = lower_median<R>(vals);
      est[i] = m;
      atomicAdd(&sh[sel_d1(abs_key(m))], 1u);
    }
  }
  __syncthreads();
  sel_merge_and_scan<1>(sh, s, k);
}

// est (d floats) receives every coordinate's median estimate; out_idx/out_val the k
// largest-magnitude ones (unordered).
CEB_API int ceb_sketch_unsketch_topk(const float* table, const int32_t* packed, int64_t n_tokens,
                                     const int32_t* blk, int num_blocks, int r, int64_t c, int64_t d,
                                     int64_t k, float* est, void* scratch, int64_t scratch_bytes,
                                     int32_t* out_idx, float* out_val, void* stream) {
  CEB_REQUIRE(table && packed && blk && est && scratch && out_idx && out_val,
              "ceb_sketch_unsketch_topk: null pointer");
  if (int e = check_sketch_dims("ceb_sketch_unsketch_topk", r, c, d, n_tokens, num_blocks)) return e;
  CEB_REQUIRE(k > 0 && k <= d, "ceb_sketch_unsketch_topk: need 0 < k <= d (k=%lld d=%lld)", (long long)k, (long long)d);
  CEB_REQUIRE(scratch_bytes >= (int64_t)sizeof(CebSelectScratch), "ceb_sketch_unsketch_topk: scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  CebSelectScratch* s = (CebSelectScratch*)scratch;
  CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), st));
  dim3 grid = token_grid(n_tokens, num_blocks, kSelThreads);
  CEB_DISPATCH_ROWS(r, (unsketch_est_kernel<R><<<grid, kSelThreads, 0, st>>>(
                           table, (uint32_t)c, packed, n_tokens, blk, num_blocks, d, est, s, (uint32_t)k)));
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return ceb_select_run(est, d, k, s, out_idx, out_val, true, st);
}

// Range variant for the sharded server tail: one thread per coordinate i in [lo, hi) (consecutive threads read
// consecutive packed hash words), estimates written compactly to est[i - lo]; digit-1 histogram of this
// range accumulated into the scratch (the cross-rank merge is ceb_dsel_run's first step).
template <int R>
__global__ void __launch_bounds__(kSelThreads)
unsketch_est_range_kernel(const float* __restrict__ table, uint32_t c, const int32_t* __restrict__ packed,
                          int64_t n, const int32_t* __restrict__ blk, int64_t lo, int64_t hi,
                          float* __restrict__ est, CebSelectScratch* __restrict__ s, uint32_t k) {
  __shared__ uint32_t sh[4096];
  for (int j = threadIdx.x; j < 4096; j += blockDim.x) sh[j] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
    const int64_t b = i / n, t = i - b * n;
    const uint32_t bp = (uint32_t)__ldg(blk + b);
    const uint32_t off = bp & 0x7fffffffu;
    const uint32_t bs = bp & 0x80000000u;
    float vals[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t hp = (uint32_t)__ldg(packed + (int64_t)r * n + t);
      uint32_t col = (hp & 0x7fffffffu) + off;
      if (col >= c) col -= c;
      const float x = __ldg(table + (int64_t)r * c + col);
      vals[r] = __uint_as_float(__float_as_uint(x) ^ ((hp & 0x80000000u) ^ bs));
    }
    const float m = lower_median<R>(vals);
    est[i - lo] = m;
    atomicAdd(&sh[sel_d1(abs_key(m))], 1u);
  }
  __syncthreads();
  sel_merge_and_scan<1>(sh, s, k);
}

// est (hi - lo floats) receives the median estimates of coordinates [lo, hi); the scratch is zeroed and left
// with this range's digit-1 histogram (pass1_done = 1 for ceb_dsel_run).
CEB_API int ceb_sketch_unsketch_est_range(const float* table, const int32_t* packed, int64_t n_tokens,
                                          const int32_t* blk, int num_blocks, int r, int64_t c, int64_t d,
                                          int64_t lo, int64_t hi, float* est, void* scratch, int64_t scratch_bytes,
                                          void* stream) {
  CEB_REQUIRE(table && packed && blk && scratch, "ceb_sketch_unsketch_est_range: null pointer");
  if (int e = check_sketch_dims("ceb_sketch_unsketch_est_range", r, c, d, n_tokens, num_blocks)) return e;
  CEB_REQUIRE(lo >= 0 && lo <= hi && hi <= d && (hi == lo || est), "ceb_sketch_unsketch_est_range: bad range [%lld, %lld) of %lld",
              (long long)lo, (long long)hi, (long long)d);
  CEB_REQUIRE(scratch_bytes >= (int64_t)sizeof(CebSelectScratch), "ceb_sketch_unsketch_est_range: scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  CebSelectScratch* s = (CebSelectScratch*)scratch;
  CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), st));
  const int grid = ceb_grid_for(hi - lo > 0 ? hi - lo : 1, kSelThreads, 4);
  CEB_DISPATCH_ROWS(r, (unsketch_est_range_kernel<R><<<grid, kSelThreads, 0, st>>>(
                           table, (uint32_t)c, packed, n_tokens, blk, lo, hi, est, s, 1u)));
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Estimates only (CSVec.unSketch over all coordinates / tests).
template <int R>
__global__ void __launch_bounds__(256)
sketch_query_kernel(const float* __restrict__ table, uint32_t c, const int32_t* __restrict__ packed,
                    int64_t n, const int32_t* __restrict__ blk, const int32_t* __restrict__ idx,
                    int64_t count, float* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += stride) {
    const int64_t i = idx ? (int64_t)__ldg(idx + j) : j;
    const int64_t b = i / n, t = i - b * n;
    const uint32_t bp = (uint32_t)__ldg(blk + b);
    float vals[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      uint32_t h = (uint32_t)__ldg(packed + (int64_t)r * n + t);
      uint32_t col = (h & 0x7fffffffu) + (bp & 0x7fffffffu);
      if (col >= c) col -= c;
      float x = __ldg(table + (int64_t)r * c + col);
      vals[r] = __uint_as_float(__float_as_uint(x) ^ ((h ^ bp) & 0x80000000u));
    }
    out[j] = lower_median<R>(vals);
  }
}

// out[j] = estimate of coordinate idx[j] (idx null: coordinates 0..count-1)
CEB_API int ceb_sketch_query(const float* table, const int32_t* packed, int64_t n_tokens, const int32_t* blk,
                             int num_blocks, int r, int64_t c, int64_t d, const int32_t* idx, int64_t count,
                             float* out, void* stream) {
  CEB_REQUIRE(table && packed && blk && out, "ceb_sketch_query: null pointer");
  if (int e = check_sketch_dims("ceb_sketch_query", r, c, d, n_tokens, num_blocks)) return e;
  if (count <= 0) return 0;
  CEB_DISPATCH_ROWS(r, (sketch_query_kernel<R><<<ceb_grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(
                           table, (uint32_t)c, packed, n_tokens, blk, idx, count, out)));
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ post-select: zero buckets + apply
// For every selected (idx,val) with val != 0: zero its r buckets in Vv (and Ve when
// given), and apply w[idx] -= lr * val (lr_vec[idx] when given).  `changed` (bytes,
// nullable) is set to 1 at idx for download accounting.
__global__ void __launch_bounds__(256)
sketch_apply_zero_kernel(const int32_t* __restrict__ idx, const float* __restrict__ val, int64_t k,
                         const int32_t* __restrict__ packed, int64_t n, const int32_t* __restrict__ blk,
                         int r, uint32_t c, float* __restrict__ Vv, float* __restrict__ Ve,
                         float* __restrict__ w, float lr, const float* __restrict__ lr_vec,
                         uint8_t* __restrict__ changed, const float* __restrict__ lr_dev) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  if (lr_dev) lr = __ldg(lr_dev);          // learning rate read from device memory (CUDA-graph replays)
  const float v = __ldg(val + j);
  if (v == 0.f) return;
  const int64_t i = __ldg(idx + j);
  const int64_t b = i / n, t = i - b * n;
  const uint32_t off = (uint32_t)__ldg(blk + b) & 0x7fffffffu;
  for (int row = 0; row < r; ++row) {
    uint32_t col = ((uint32_t)__ldg(packed + (int64_t)row * n + t) & 0x7fffffffu) + off;
    if (col >= c) col -= c;
    if (Vv) Vv[(int64_t)row * c + col] = 0.f;
    if (Ve) Ve[(int64_t)row * c + col] = 0.f;
  }
  if (w) {
    const float step = (lr_vec ? __ldg(lr_vec + i) : lr) * v;
    w[i] -= step;
    if (changed && step != 0.f) changed[i] = 1;
  }
}

static int apply_zero_impl(const int32_t* idx, const float* val, int64_t k, const int32_t* packed, int64_t n_tokens,
                           const int32_t* blk, int r, int64_t c, float* Vv, float* Ve, float* w, float lr,
                           const float* lr_vec, uint8_t* changed, const float* lr_dev, void* stream) {
  CEB_REQUIRE(idx && val && packed && blk, "ceb_sketch_apply_zero: null pointer");
  CEB_REQUIRE(r >= 1 && r <= kMaxRows, "ceb_sketch_apply_zero: num_rows=%d", r);
  if (k <= 0) return 0;
  sketch_apply_zero_kernel<<<(unsigned)ceb_cdiv(k, 256), 256, 0, (cudaStream_t)stream>>>(
      idx, val, k, packed, n_tokens, blk, r, (uint32_t)c, Vv, Ve, w, lr, lr_vec, changed, lr_dev);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_sketch_apply_zero(const int32_t* idx, const float* val, int64_t k, const int32_t* packed,
                                  int64_t n_tokens, const int32_t* blk, int r, int64_t c, float* Vv,
                                  float* Ve, float* w, float lr, const float* lr_vec, uint8_t* changed,
                                  void* stream) {
  return apply_zero_impl(idx, val, k, packed, n_tokens, blk, r, c, Vv, Ve, w, lr, lr_vec, changed, nullptr, stream);
}

// Same with the scalar learning rate read from device memory at run time, so the launch can be
// replayed from a CUDA graph while the schedule changes the rate every round.
CEB_API int ceb_sketch_apply_zero_dev(const int32_t* idx, const float* val, int64_t k, const int32_t* packed,
                                      int64_t n_tokens, const int32_t* blk, int r, int64_t c, float* Vv, float* Ve,
                                      float* w, const float* lr_dev, uint8_t* changed, void* stream) {
  CEB_REQUIRE(lr_dev != nullptr, "ceb_sketch_apply_zero_dev: lr_dev is NULL");
  return apply_zero_impl(idx, val, k, packed, n_tokens, blk, r, c, Vv, Ve, w, 0.f, nullptr, changed, lr_dev, stream);
}

// ------------------------------------------------------------------ heavy-hitter recovery without any allocation
// ceb_sketch_unsketch_topk needs two pieces of scratch from its caller: the d-float
// estimate vector and the radix-select block.  Some hosts of this library cannot supply
// them: an aggregation service whose device memory was carved up once at start-up and
// handed to other owners, a plug-in living inside another framework's memory pool with
// no right to ask it for more, or a latency-critical loop that must not risk the
// device-wide synchronisation a cudaMalloc can cause.  For them the variant below keeps
// its scratch in one file-scope device array that is part of the loaded module.
//
// Cost and limits, stated plainly:
//   * the array is CEB_STATIC_UNSKETCH_MAX_D floats (8 Mi coordinates = 32 MiB) plus the
//     ~34 KB select block; every GPU on which this library is loaded holds it until the
//     process exits, whether or not this entry point is ever called;
//   * d above that maximum is refused with an error (use ceb_sketch_unsketch_topk);
//   * all calls on a device share the one array: it still holds the previous call's
//     contents, so the select block is re-zeroed here on every call, and two calls that
//     overlap on the GPU (different streams, no ordering between them) race on it.
#include <mutex>

constexpr int64_t kStaticMaxCoords = (int64_t)CEB_STATIC_UNSKETCH_MAX_D;
constexpr size_t kStaticSelectBytes = (sizeof(CebSelectScratch) + 255) / 256 * 256;
static __device__ __align__(256) unsigned char
    g_static_unsketch_scratch[kStaticSelectBytes + (size_t)kStaticMaxCoords * sizeof(float)];

// In host code the array's name is only a handle for the runtime, not an address a kernel
// can use; the real device address is looked up once per device (that device current).
static unsigned char* g_static_scratch_addr[64] = {nullptr};
static std::mutex g_static_scratch_mutex;

static int static_scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  CEB_CHECK_CUDA(cudaGetDevice(&dev));
  CEB_REQUIRE(dev >= 0 && dev < 64, "static scratch: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_addr[dev] == nullptr) {
    unsigned char* p = nullptr;
    CEB_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_static_unsketch_scratch));
    CEB_REQUIRE(p != nullptr, "static scratch: cudaGetSymbolAddress returned a null address");
    g_static_scratch_addr[dev] = p;
  }
  *out = g_static_scratch_addr[dev];
  return 0;
}

// Same result as ceb_sketch_unsketch_topk (out_idx/out_val: the k largest-magnitude median
// estimates, unordered) with no scratch arguments and no allocation of any kind.  The first
// call on each device resolves the scratch address (cudaGetSymbolAddress); make that first
// call outside any stream capture.
CEB_API int ceb_sketch_unsketch_topk_static(const float* table, const int32_t* packed, int64_t n_tokens,
                                            const int32_t* blk, int num_blocks, int r, int64_t c,
                                            int64_t d, int64_t k, int32_t* out_idx, float* out_val,
                                            void* stream) {
  CEB_REQUIRE(d <= kStaticMaxCoords,
              "ceb_sketch_unsketch_topk_static: d=%lld exceeds the built-in scratch (%lld coordinates); "
              "use ceb_sketch_unsketch_topk with caller-provided scratch",
              (long long)d, (long long)kStaticMaxCoords);
  CEB_REQUIRE(table && packed && blk && out_idx && out_val, "ceb_sketch_unsketch_topk_static: null pointer");
  if (int e = check_sketch_dims("ceb_sketch_unsketch_topk_static", r, c, d, n_tokens, num_blocks)) return e;
  CEB_REQUIRE(k > 0 && k <= d, "ceb_sketch_unsketch_topk_static: need 0 < k <= d (k=%lld d=%lld)",
              (long long)k, (long long)d);
  unsigned char* base = nullptr;
  if (int e = static_scratch_for_current_device(&base)) return e;
  CebSelectScratch* s = (CebSelectScratch*)base;
  float* est = (float*)(base + kStaticSelectBytes);
  cudaStream_t st = (cudaStream_t)stream;
  // histograms, tickets and counters must start at zero; est is fully overwritten for i < d
  CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), st));
  dim3 grid = token_grid(n_tokens, num_blocks, kSelThreads);
  CEB_DISPATCH_ROWS(r, (unsketch_est_kernel<R><<<grid, kSelThreads, 0, st>>>(
                           table, (uint32_t)c, packed, n_tokens, blk, num_blocks, d, est, s, (uint32_t)k)));
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return ceb_select_run(est, d, k, s, out_idx, out_val, true, st);
}
