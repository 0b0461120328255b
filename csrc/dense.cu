// Dense-vector kernels for the non-sketch modes and the per-client state updates
// (sm_100a).  Reference call sites: server rules
// /root/reference/CommEfficient/fed_aggregator.py:483-566, local momentum / error /
// masking fed_worker.py:190-216, fedavg local SGD fed_worker.py:82-108, DP clip+noise
// fed_worker.py:304-309, download accounting fed_aggregator.py:244-250.
#include "select.cuh"
#include <curand_kernel.h>


// ------------------------------------------------------------------ server: momentum (+error) (+apply)
// Vv = rho*Vv + g*inv_n (+ N(0, noise_std^2) when noise_std > 0)
//   apply != 0 : w -= lr * Vv                       (uncompressed / local_topk / fedavg)
//   Ve != null : Ve += Vv and histogram digit 1 of |Ve| into the select scratch (true_topk)
__global__ void __launch_bounds__(kSelThreads)
dense_server_kernel(float* __restrict__ Vv, float* __restrict__ Ve, const float* __restrict__ g,
                    float inv_n, float rho, int64_t n, float* __restrict__ w, float lr,
                    const float* __restrict__ lr_vec, int apply, float noise_std,
                    unsigned long long seed, unsigned long long noise_offset,
                    uint8_t* __restrict__ changed, CebSelectScratch* __restrict__ s, uint32_t k) {
  __shared__ uint32_t sh[4096];
  const bool do_hist = (Ve != nullptr) && (s != nullptr);
  if (do_hist) {
    for (int j = threadIdx.x; j < 4096; j += blockDim.x) sh[j] = 0;
    __syncthreads();
  }
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  curandStatePhilox4_32_10_t rng;
  if (noise_std > 0.f) curand_init(seed, (unsigned long long)tid, noise_offset, &rng);
  for (int64_t i = tid; i < n; i += stride) {
    float v = fmaf(rho, Vv[i], __ldg(g + i) * inv_n);
    Vv[i] = v;
    float u = v;
    if (noise_std > 0.f) u = fmaf(noise_std, curand_normal(&rng), v);
    if (apply) {
      const float step = (lr_vec ? __ldg(lr_vec + i) : lr) * u;
      w[i] -= step;
      if (changed && step != 0.f) changed[i] = 1;
    }
    if (Ve) {
      float e = Ve[i] + v;
      Ve[i] = e;
      if (do_hist) atomicAdd(&sh[sel_d1(abs_key(e))], 1u);
    }
  }
  if (do_hist) {
    __syncthreads();
    sel_merge_and_scan<1>(sh, s, k);
  }
}

CEB_API int ceb_dense_server_update(float* Vv, const float* g, float inv_n, float rho, int64_t n, float* w,
                                    float lr, const float* lr_vec, float noise_std,
                                    unsigned long long seed, unsigned long long noise_offset,
                                    uint8_t* changed, void* stream) {
  CEB_REQUIRE(Vv && g && w && n > 0, "ceb_dense_server_update: bad arguments");
  dense_server_kernel<<<ceb_grid_for(n, kSelThreads, 4), kSelThreads, 0, (cudaStream_t)stream>>>(
      Vv, nullptr, g, inv_n, rho, n, w, lr, lr_vec, 1, noise_std, seed, noise_offset, changed, nullptr, 0);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Post-select for true_topk: Ve[idx] = 0, Vv[idx] = 0, w[idx] -= lr*val; optionally also
// zero those coordinates in `n_vel_rows` per-client velocity rows (momentum factor
// masking of the participating clients, fed_aggregator.py:526-533).
__global__ void __launch_bounds__(256)
dense_apply_zero_kernel(const int32_t* __restrict__ idx, const float* __restrict__ val, int64_t k,
                        float* __restrict__ Vv, float* __restrict__ Ve, float* __restrict__ w, float lr,
                        const float* __restrict__ lr_vec, uint8_t* __restrict__ changed,
                        float* __restrict__ client_vel, const int64_t* __restrict__ vel_rows,
                        int n_vel_rows, int64_t d) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  const float v = __ldg(val + j);
  if (v == 0.f) return;
  const int64_t i = __ldg(idx + j);
  if (Ve) Ve[i] = 0.f;
  if (Vv) Vv[i] = 0.f;
  if (w) {
    const float step = (lr_vec ? __ldg(lr_vec + i) : lr) * v;
    w[i] -= step;
    if (changed && step != 0.f) changed[i] = 1;
  }
  for (int r = 0; r < n_vel_rows; ++r) client_vel[__ldg(vel_rows + r) * d + i] = 0.f;
}

CEB_API int ceb_dense_apply_zero(const int32_t* idx, const float* val, int64_t k, float* Vv, float* Ve,
                                 float* w, float lr, const float* lr_vec, uint8_t* changed,
                                 float* client_vel, const int64_t* vel_rows, int n_vel_rows, int64_t d,
                                 void* stream) {
  CEB_REQUIRE(idx && val, "ceb_dense_apply_zero: null pointer");
  if (k <= 0) return 0;
  dense_apply_zero_kernel<<<(unsigned)ceb_cdiv(k, 256), 256, 0, (cudaStream_t)stream>>>(
      idx, val, k, Vv, Ve, w, lr, lr_vec, changed, client_vel, vel_rows, n_vel_rows, d);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Sharded form (multi-GPU true_topk): Vv / Ve are this rank's coordinate shard [lo, hi) (indexed from lo); every
// rank applies the whole update to its weight replica, only the owner of a coordinate zeroes its state.
__global__ void __launch_bounds__(256)
dense_apply_zero_shard_kernel(const int32_t* __restrict__ idx, const float* __restrict__ val, int64_t k,
                              float* __restrict__ Vv, float* __restrict__ Ve, int64_t lo, int64_t hi,
                              float* __restrict__ w, float lr, const float* __restrict__ lr_vec) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  const float v = __ldg(val + j);
  if (v == 0.f) return;
  const int64_t i = __ldg(idx + j);
  if (i >= lo && i < hi) { Ve[i - lo] = 0.f; Vv[i - lo] = 0.f; }
  w[i] -= (lr_vec ? __ldg(lr_vec + i) : lr) * v;
}

CEB_API int ceb_dense_apply_zero_shard(const int32_t* idx, const float* val, int64_t k, float* Vv_shard, float* Ve_shard,
                                       int64_t lo, int64_t hi, float* w, float lr, const float* lr_vec, void* stream) {
  CEB_REQUIRE(idx && val && Vv_shard && Ve_shard && w && lo <= hi, "ceb_dense_apply_zero_shard: bad arguments");
  if (k <= 0) return 0;
  dense_apply_zero_shard_kernel<<<(unsigned)ceb_cdiv(k, 256), 256, 0, (cudaStream_t)stream>>>(
      idx, val, k, Vv_shard, Ve_shard, lo, hi, w, lr, lr_vec);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// true_topk in one call: momentum + error + exact top-k of the error vector + zeroing + apply.
CEB_API int ceb_true_topk_step(float* Vv, float* Ve, const float* g, float inv_n, float rho, int64_t n,
                               int64_t k, float* w, float lr, const float* lr_vec, uint8_t* changed,
                               void* scratch, int64_t scratch_bytes, int32_t* out_idx, float* out_val,
                               float* client_vel, const int64_t* vel_rows, int n_vel_rows, void* stream) {
  CEB_REQUIRE(Vv && Ve && g && w && scratch && out_idx && out_val, "ceb_true_topk_step: null pointer");
  CEB_REQUIRE(n > 0 && n < ((int64_t)1 << 31) && k > 0 && k <= n, "ceb_true_topk_step: need 0 < k <= n < 2^31");
  CEB_REQUIRE(scratch_bytes >= (int64_t)sizeof(CebSelectScratch), "ceb_true_topk_step: scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  CebSelectScratch* s = (CebSelectScratch*)scratch;
  CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), st));
  dense_server_kernel<<<ceb_grid_for(n, kSelThreads, 4), kSelThreads, 0, st>>>(
      Vv, Ve, g, inv_n, rho, n, nullptr, 0.f, nullptr, 0, 0.f, 0ull, 0ull, nullptr, s, (uint32_t)k);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  if (int e = ceb_select_run(Ve, n, k, s, out_idx, out_val, true, st)) return e;
  return ceb_dense_apply_zero(out_idx, out_val, k, Vv, Ve, w, lr, lr_vec, changed, client_vel, vel_rows,
                              n_vel_rows, n, stream);
}

// ------------------------------------------------------------------ client state (local momentum / error)
// src = scale*g ; vel != null: vel = rho*vel + src, src = vel ; err != null: err += src, src = err.
// The transmit vector ends up in err, else vel, else g (scaled in place).  With a scratch the
// digit-1 histogram of |transmit| is accumulated for a following top-k (local_topk).
__global__ void __launch_bounds__(kSelThreads)
client_state_kernel(float* __restrict__ g, float scale, float* __restrict__ vel, float rho,
                    float* __restrict__ err, int64_t n, CebSelectScratch* __restrict__ s, uint32_t k) {
  __shared__ uint32_t sh[4096];
  if (s) {
    for (int j = threadIdx.x; j < 4096; j += blockDim.x) sh[j] = 0;
    __syncthreads();
  }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float x = g[i] * scale;
    if (vel) { x = fmaf(rho, vel[i], x); vel[i] = x; }
    if (err) { x += err[i]; err[i] = x; }
    if (!vel && !err) g[i] = x;
    if (s) atomicAdd(&sh[sel_d1(abs_key(x))], 1u);
  }
  if (s) {
    __syncthreads();
    sel_merge_and_scan<1>(sh, s, k);
  }
}

// acc[idx] += val ; err[idx] = 0 ; vel[idx] = 0   (indices unique within one call)
__global__ void __launch_bounds__(256)
scatter_add_zero_kernel(const int32_t* __restrict__ idx, const float* __restrict__ val, int64_t k,
                        float* __restrict__ acc, float* __restrict__ err, float* __restrict__ vel) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  const float v = __ldg(val + j);
  if (v == 0.f) return;
  const int64_t i = __ldg(idx + j);
  if (acc) acc[i] += v;
  if (err) err[i] = 0.f;
  if (vel) vel[i] = 0.f;
}

CEB_API int ceb_scatter_add_zero(const int32_t* idx, const float* val, int64_t k, float* acc, float* err,
                                 float* vel, void* stream) {
  CEB_REQUIRE(idx && val, "ceb_scatter_add_zero: null pointer");
  if (k <= 0) return 0;
  scatter_add_zero_kernel<<<(unsigned)ceb_cdiv(k, 256), 256, 0, (cudaStream_t)stream>>>(idx, val, k, acc, err, vel);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// One client's local step after backward.  k == 0: dense transmit, acc += transmit.
// k > 0 (local_topk): acc[idx] += topk(transmit), then error feedback / momentum masking.
CEB_API int ceb_client_step(float* g, float scale, float* vel, float rho, float* err, int64_t n, int64_t k,
                            float* acc, void* scratch, int64_t scratch_bytes, int32_t* out_idx,
                            float* out_val, void* stream) {
  CEB_REQUIRE(g && acc && n > 0, "ceb_client_step: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  float* transmit = err ? err : (vel ? vel : g);
  if (k <= 0) {
    client_state_kernel<<<ceb_grid_for(n, kSelThreads, 4), kSelThreads, 0, st>>>(g, scale, vel, rho, err, n, nullptr, 0);
    CEB_COUNT_LAUNCH(1);
    CEB_CHECK_LAUNCH();
    return ceb_axpy_dev(acc, transmit, n, 1.f, nullptr, 0, stream);
  }
  CEB_REQUIRE(k <= n && n < ((int64_t)1 << 31), "ceb_client_step: need k <= n < 2^31");
  CEB_REQUIRE(scratch && out_idx && out_val && scratch_bytes >= (int64_t)sizeof(CebSelectScratch),
              "ceb_client_step: top-k needs scratch and output buffers");
  CebSelectScratch* s = (CebSelectScratch*)scratch;
  CEB_CHECK_CUDA(cudaMemsetAsync(s, 0, sizeof(CebSelectScratch), st));
  client_state_kernel<<<ceb_grid_for(n, kSelThreads, 4), kSelThreads, 0, st>>>(g, scale, vel, rho, err, n, s, (uint32_t)k);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  if (int e = ceb_select_run(transmit, n, k, s, out_idx, out_val, true, st)) return e;
  return ceb_scatter_add_zero(out_idx, out_val, k, acc, err, vel, stream);
}

// ------------------------------------------------------------------ norms, clipping, noise
__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ x, int64_t n, float* __restrict__ out) {
  float acc = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v = __ldg(x + i);
    acc = fmaf(v, v, acc);
  }
  __shared__ float ws[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? ws[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(out, v);
  }
}

// x *= coef, coef = min(1, clip / (sqrt(*sumsq) + eps))  [eps > 0: torch clip_grad_norm_ rule]
//             or    (sqrt(*sumsq) >= clip ? clip / sqrt(*sumsq) : 1)            [eps == 0]
__global__ void __launch_bounds__(256)
clip_scale_kernel(float* __restrict__ x, int64_t n, const float* __restrict__ sumsq, float clip, float eps) {
  const float norm = sqrtf(__ldg(sumsq));
  float coef;
  if (eps > 0.f) coef = fminf(1.f, clip / (norm + eps));
  else coef = norm >= clip ? clip / norm : 1.f;
  if (coef == 1.f) return;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] *= coef;
}

// sumsq_out: one float of scratch (receives ||x||^2)
CEB_API int ceb_clip_by_norm(float* x, int64_t n, float clip, float eps, float* sumsq_out, void* stream) {
  CEB_REQUIRE(x && sumsq_out && n > 0, "ceb_clip_by_norm: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  CEB_CHECK_CUDA(cudaMemsetAsync(sumsq_out, 0, sizeof(float), st));
  sumsq_kernel<<<ceb_grid_for(n, 256, 4), 256, 0, st>>>(x, n, sumsq_out);
  clip_scale_kernel<<<ceb_grid_for(n, 256), 256, 0, st>>>(x, n, sumsq_out, clip, eps);
  CEB_COUNT_LAUNCH(2);
  CEB_CHECK_LAUNCH();
  return 0;
}

__global__ void __launch_bounds__(256)
gaussian_noise_kernel(float* __restrict__ x, int64_t n, float std, unsigned long long seed,
                      unsigned long long offset) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  curandStatePhilox4_32_10_t rng;
  curand_init(seed, (unsigned long long)tid, offset, &rng);
  // each thread owns groups of 4 consecutive elements
  for (int64_t base = tid * 4; base < n; base += nthreads * 4) {
    float4 z = curand_normal4(&rng);
    float zs[4] = {z.x, z.y, z.z, z.w};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (base + j < n) x[base + j] = fmaf(std, zs[j], x[base + j]);
  }
}

// x += N(0, std^2) (Philox4x32-10 counter RNG, stream position = offset)
CEB_API int ceb_add_gaussian_noise(float* x, int64_t n, float std, unsigned long long seed,
                                   unsigned long long offset, void* stream) {
  CEB_REQUIRE(x && n > 0, "ceb_add_gaussian_noise: bad arguments");
  if (std == 0.f) return 0;
  gaussian_noise_kernel<<<ceb_grid_for(ceb_cdiv(n, 4), 256, 4), 256, 0, (cudaStream_t)stream>>>(x, n, std, seed, offset);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ fedavg helpers
// acc += weight * (w0 - wloc)
__global__ void __launch_bounds__(256)
fedavg_delta_kernel(float* __restrict__ acc, const float* __restrict__ w0, const float* __restrict__ wl,
                    float weight, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    acc[i] = fmaf(weight, __ldg(w0 + i) - __ldg(wl + i), acc[i]);
}

CEB_API int ceb_fedavg_delta(float* acc, const float* w0, const float* wloc, float weight, int64_t n, void* stream) {
  CEB_REQUIRE(acc && w0 && wloc && n > 0, "ceb_fedavg_delta: bad arguments");
  fedavg_delta_kernel<<<ceb_grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(acc, w0, wloc, weight, n);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ download accounting
// mask[i] |= (w[i] != prev[i]) ; prev[i] = w[i] ; *count += popcount(mask)
__global__ void __launch_bounds__(256)
count_changed_kernel(const float* __restrict__ w, float* __restrict__ prev, uint8_t* __restrict__ mask,
                     int64_t n, unsigned long long* __restrict__ count) {
  unsigned int local = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float a = __ldg(w + i);
    uint8_t m = mask ? mask[i] : 0;
    if (prev) {
      if (a != prev[i]) m = 1;
      prev[i] = a;
    }
    if (mask) mask[i] = m;
    local += m;
  }
  local = __reduce_add_sync(0xffffffffu, local);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, (unsigned long long)local);
}

CEB_API int ceb_count_changed(const float* w, float* prev, uint8_t* mask, int64_t n,
                              unsigned long long* count, void* stream) {
  CEB_REQUIRE(w && count && n > 0, "ceb_count_changed: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  CEB_CHECK_CUDA(cudaMemsetAsync(count, 0, sizeof(unsigned long long), st));
  count_changed_kernel<<<ceb_grid_for(n, 256, 4), 256, 0, st>>>(w, prev, mask, n, count);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}


// ------------------------------------------------------------------ download-byte tracker
// State (commefficient_b200/fed_aggregator.py:_DownloadTracker): last[d] = server round of every
// coordinate's last change, hist[r] = number of coordinates whose last change was round r.
// record: the k coordinates in idx move from their old bucket to bucket `round`.  Most of them share a
// handful of old buckets, so decrements are aggregated per warp (match.any) before the atomic.
__global__ void __launch_bounds__(256)
tracker_record_kernel(const int* __restrict__ idx, int64_t k, long long* __restrict__ last,
                      unsigned long long* __restrict__ hist, long long round,
                      const long long* __restrict__ round_dev) {
  if (round_dev) round = *round_dev;       // round number kept on the device (CUDA-graph replays)
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = i < k;
  long long old = -1;
  if (active) {
    const int c = idx[i];
    old = last[c];
    last[c] = round;
  }
  const unsigned peers = __match_any_sync(0xffffffffu, old);
  const int leader = __ffs(peers) - 1;
  if (active && (int)(threadIdx.x & 31) == leader)
    atomicAdd(hist + old, (unsigned long long)(-(long long)__popc(peers)));
  if (i == 0) atomicAdd(hist + round, (unsigned long long)k);
}

CEB_API int ceb_tracker_record(const int* idx, int64_t k, int64_t* last, int64_t* hist, int64_t round, void* stream) {
  CEB_REQUIRE(idx && last && hist && k > 0 && round > 0, "ceb_tracker_record: bad arguments");
  tracker_record_kernel<<<(unsigned)ceb_cdiv(k, 256), 256, 0, (cudaStream_t)stream>>>(
      idx, k, (long long*)last, (unsigned long long*)hist, (long long)round, nullptr);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Same with the round number read from device memory (*round_dev > 0 and < the length of hist).
CEB_API int ceb_tracker_record_dev(const int* idx, int64_t k, int64_t* last, int64_t* hist, const int64_t* round_dev,
                                   void* stream) {
  CEB_REQUIRE(idx && last && hist && round_dev && k > 0, "ceb_tracker_record_dev: bad arguments");
  tracker_record_kernel<<<(unsigned)ceb_cdiv(k, 256), 256, 0, (cudaStream_t)stream>>>(
      idx, k, (long long*)last, (unsigned long long*)hist, 0, (const long long*)round_dev);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// out[j] = sum_{r >= start[j]} hist[r]  (one warp per query; n is the number of rounds so far, small)
__global__ void __launch_bounds__(256)
tracker_counts_kernel(const long long* __restrict__ hist, int n, const long long* __restrict__ start, int m,
                      float* __restrict__ out) {
  const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (j >= m) return;
  const int lane = threadIdx.x & 31;
  long long s = 0;
  for (long long r = start[j] + lane; r < n; r += 32) s += hist[r];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[j] = (float)s;
}

CEB_API int ceb_tracker_counts(const int64_t* hist, int n, const int64_t* start, int m, float* out, void* stream) {
  CEB_REQUIRE(hist && start && out && n > 0 && m > 0, "ceb_tracker_counts: bad arguments");
  tracker_counts_kernel<<<(unsigned)ceb_cdiv(m, 8), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)hist, n, (const long long*)start, m, out);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}
