// Transformer-side kernels (sm_100a): fused softmax cross-entropy over a large vocabulary,
// LayerNorm forward/backward, tanh-GELU forward/backward.  They replace the ATen op chains the
// reference runs inside pytorch_transformers' GPT-2 (/root/reference/CommEfficient/gpt2_train.py:76-99:
// shifted LM cross-entropy with ignore_index -1 over 50 262 classes; LayerNorm / GELU in every block).
#include <cuda_bf16.h>
#include "common.cuh"

CEB_API int ceb_add_layernorm_drop_fwd(const float* x, const void* branch_bf16, const float* gamma, const float* beta,
                                       float* resid_out, void* y_bf16, float* mean, float* rstd, int64_t rows, int E,
                                       float eps, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream);
CEB_API int ceb_add_layernorm_drop_bwd(const void* dy_bf16, const float* resid, const float* gamma, const float* mean,
                                       const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                                       float* dbeta, int64_t rows, int E, float p_drop, const uint64_t* seed_dev,
                                       uint32_t salt, void* stream);

// ------------------------------------------------------------------ softmax cross-entropy
// One block per row.  Forward: loss[row] = logsumexp(row) - row[label] (0 when label == ignore),
// lse[row] saved.  Backward (in place allowed): d[row][j] = g[row] * (exp(l[j] - lse) - [j == label]).
// Rows are read as bf16 pairs: a 50 262-wide bf16 row is only 4-byte aligned.
constexpr int CE_THREADS = 512;

__device__ __forceinline__ float block_reduce_max(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : -INFINITY;
  if (threadIdx.x < 32) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (threadIdx.x == 0) sh[0] = v;
  }
  __syncthreads();
  v = sh[0];
  __syncthreads();
  return v;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* sh) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
  if (threadIdx.x < 32) {
    v = warp_sum(v);
    if (threadIdx.x == 0) sh[0] = v;
  }
  __syncthreads();
  v = sh[0];
  __syncthreads();
  return v;
}

__global__ void __launch_bounds__(CE_THREADS)
ce_fwd_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ld, int V, const int64_t* __restrict__ labels,
              int64_t ignore_index, float* __restrict__ loss, float* __restrict__ lse_out) {
  __shared__ float sh[32];
  const int64_t row = blockIdx.x;
  const __nv_bfloat16* p = logits + row * ld;
  const int64_t label = labels[row];
  if (label == ignore_index) {          // uniform per block
    if (threadIdx.x == 0) { loss[row] = 0.f; lse_out[row] = 0.f; }
    return;
  }
  const __nv_bfloat162* p2 = reinterpret_cast<const __nv_bfloat162*>(p);
  const int pairs = V >> 1;
  float m = -INFINITY, s = 0.f;          // online softmax: running max and sum of exp(x - m)
  // 16-byte loads when every row starts 16-byte aligned (row-padded logits buffers), bf16 pairs otherwise
  const bool vec = (ld % 8 == 0) && ((((uintptr_t)logits) & 15) == 0);
  const int nvec = vec ? (V >> 3) : 0;
  const uint4* p4 = reinterpret_cast<const uint4*>(p);
  for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
    const uint4 u = p4[j];
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __bfloat1622float2(*(const __nv_bfloat162*)&w[i]);
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
    float mx = f[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, f[i]);
    if (mx > m) { s *= __expf(m - mx); m = mx; }
#pragma unroll
    for (int i = 0; i < 8; ++i) s += __expf(f[i] - m);
  }
  for (int j = nvec * 4 + threadIdx.x; j < pairs; j += blockDim.x) {
    const float2 v = __bfloat1622float2(p2[j]);
    const float mx = fmaxf(v.x, v.y);
    if (mx > m) { s *= __expf(m - mx); m = mx; }
    s += __expf(v.x - m) + __expf(v.y - m);
  }
  if ((V & 1) && threadIdx.x == 0) {
    const float x = __bfloat162float(p[V - 1]);
    if (x > m) { s *= __expf(m - x); m = x; }
    s += __expf(x - m);
  }
  const float gm = block_reduce_max(m, sh);
  s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
  const float gs = block_reduce_sum(s, sh);
  if (threadIdx.x == 0) {
    const float lse = gm + __logf(gs);
    lse_out[row] = lse;
    loss[row] = lse - __bfloat162float(p[label]);
  }
}

__global__ void __launch_bounds__(CE_THREADS)
ce_bwd_kernel(const __nv_bfloat16* __restrict__ logits, __nv_bfloat16* __restrict__ dlogits, int64_t ld, int V,
              const int64_t* __restrict__ labels, int64_t ignore_index, const float* __restrict__ lse,
              const float* __restrict__ grow, const int* __restrict__ row_limit) {
  const int64_t row = blockIdx.x;
  // compacted rows: nothing past the 128-row tile that holds the last valid row is ever read
  if (row_limit && row >= (((int64_t)__ldg(row_limit) + 127) & ~(int64_t)127)) return;
  const int64_t label = labels[row];
  const bool ignored = label == ignore_index;
  const float g = ignored ? 0.f : grow[row];
  const float l = lse[row];
  const __nv_bfloat162* p2 = reinterpret_cast<const __nv_bfloat162*>(logits + row * ld);
  __nv_bfloat162* d2 = reinterpret_cast<__nv_bfloat162*>(dlogits + row * ld);
  const int pairs = V >> 1;
  const bool vec = (ld % 8 == 0) && (((((uintptr_t)logits) | ((uintptr_t)dlogits)) & 15) == 0);
  const int nvec = vec ? (V >> 3) : 0;
  const uint4* p4 = reinterpret_cast<const uint4*>(p2);
  uint4* d4 = reinterpret_cast<uint4*>(d2);
  for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (!ignored) {
      const uint4 u = p4[j];
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
      uint32_t r[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(*(const __nv_bfloat162*)&w[i]);
        const int c = 8 * j + 2 * i;
        t.x = g * (__expf(t.x - l) - (c == label ? 1.f : 0.f));
        t.y = g * (__expf(t.y - l) - ((c + 1) == label ? 1.f : 0.f));
        const __nv_bfloat162 h = __floats2bfloat162_rn(t.x, t.y);
        r[i] = *(const uint32_t*)&h;
      }
      o = make_uint4(r[0], r[1], r[2], r[3]);
    }
    d4[j] = o;
  }
  for (int j = nvec * 4 + threadIdx.x; j < pairs; j += blockDim.x) {
    float2 v = __bfloat1622float2(p2[j]);
    v.x = ignored ? 0.f : g * (__expf(v.x - l) - ((2 * j) == label ? 1.f : 0.f));
    v.y = ignored ? 0.f : g * (__expf(v.y - l) - ((2 * j + 1) == label ? 1.f : 0.f));
    d2[j] = __floats2bfloat162_rn(v.x, v.y);
  }
  if ((V & 1) && threadIdx.x == 0) {
    const float x = __bfloat162float(logits[row * ld + V - 1]);
    dlogits[row * ld + V - 1] = __float2bfloat16(ignored ? 0.f : g * (__expf(x - l) - ((V - 1) == label ? 1.f : 0.f)));
  }
}

CEB_API int ceb_softmax_ce_fwd(const void* logits_bf16, int64_t ld, int64_t rows, int V, const int64_t* labels,
                               int64_t ignore_index, float* loss, float* lse, void* stream) {
  CEB_REQUIRE(logits_bf16 && labels && loss && lse && rows > 0 && V > 1 && ld % 2 == 0, "ceb_softmax_ce_fwd: bad arguments");
  ce_fwd_kernel<<<(unsigned)rows, CE_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits_bf16, ld, V, labels,
                                                                      ignore_index, loss, lse);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

static int ce_bwd_impl(const void* logits_bf16, void* dlogits_bf16, int64_t ld, int64_t rows, int V,
                       const int64_t* labels, int64_t ignore_index, const float* lse, const float* grad_rows,
                       const int* row_limit, void* stream) {
  CEB_REQUIRE(logits_bf16 && dlogits_bf16 && labels && lse && grad_rows && rows > 0 && ld % 2 == 0, "ceb_softmax_ce_bwd: bad arguments");
  ce_bwd_kernel<<<(unsigned)rows, CE_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits_bf16,
      (__nv_bfloat16*)dlogits_bf16, ld, V, labels, ignore_index, lse, grad_rows, row_limit);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_softmax_ce_bwd(const void* logits_bf16, void* dlogits_bf16, int64_t ld, int64_t rows, int V,
                               const int64_t* labels, int64_t ignore_index, const float* lse, const float* grad_rows,
                               void* stream) {
  return ce_bwd_impl(logits_bf16, dlogits_bf16, ld, rows, V, labels, ignore_index, lse, grad_rows, nullptr, stream);
}

// Rows are compacted (valid labels first): rows past the 128-row tile containing row *row_limit_dev - 1
// are skipped entirely (neither read nor written).  dlogits may alias logits (in place).
CEB_API int ceb_softmax_ce_bwd_rows(const void* logits_bf16, void* dlogits_bf16, int64_t ld, int64_t rows, int V,
                                    const int64_t* labels, int64_t ignore_index, const float* lse,
                                    const float* grad_rows, const int* row_limit_dev, void* stream) {
  CEB_REQUIRE(row_limit_dev != nullptr, "ceb_softmax_ce_bwd_rows: row_limit_dev is NULL");
  return ce_bwd_impl(logits_bf16, dlogits_bf16, ld, rows, V, labels, ignore_index, lse, grad_rows, row_limit_dev, stream);
}

// ------------------------------------------------------------------ LayerNorm (one warp per row)
// y (bf16) = (x - mean) * rstd * gamma + beta ; x fp32 [rows][E]; mean / rstd saved for backward.
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              __nv_bfloat16* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, int64_t rows,
              int E, float eps, const __nv_bfloat16* __restrict__ branch, float* __restrict__ resid_out,
              uint32_t drop_thresh = 0, float inv_keep = 1.f, const unsigned long long* __restrict__ seed_dev = nullptr,
              uint32_t salt = 0) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + row * E;
  if (branch) {
    // fused residual add: resid_out = x + dropout(branch) (the new fp32 residual stream), normalised below
    const __nv_bfloat16* br = branch + row * E;
    float* ro = resid_out + row * E;
    if (drop_thresh) {
      const uint32_t rowkey = ceb_drop_rowkey(seed_dev, salt, (uint32_t)row);
      for (int j = lane; j < E; j += 32) {
        const float keep = ceb_hash32(rowkey + (uint32_t)j) >= drop_thresh ? inv_keep : 0.f;
        ro[j] = fmaf(__bfloat162float(br[j]), keep, xr[j]);
      }
    } else {
      for (int j = lane; j < E; j += 32) ro[j] = xr[j] + __bfloat162float(br[j]);
    }
    __syncwarp();
    xr = ro;
  }
  float s = 0.f, ss = 0.f;
  for (int j = lane; j < E; j += 32) { const float v = xr[j]; s += v; ss = fmaf(v, v, ss); }
  s = warp_sum(s); ss = warp_sum(ss);
  const float mu = s / E;
  const float rs = rsqrtf(fmaxf(ss / E - mu * mu, 0.f) + eps);
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
  for (int j = lane; j < E; j += 32)
    y[row * E + j] = __float2bfloat16((xr[j] - mu) * rs * __ldg(gamma + j) + __ldg(beta + j));
}

// Vector variant for E = 128 * G and 16-byte aligned rows: a lane owns G float4 column groups, the row stays in
// registers between the statistics and the normalisation (one global read of x / branch, 16-byte accesses).
template <int G>
__global__ void __launch_bounds__(256)
ln_fwd_vec_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                  __nv_bfloat16* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, int64_t rows,
                  float eps, const __nv_bfloat16* __restrict__ branch, float* __restrict__ resid_out,
                  uint32_t drop_thresh, float inv_keep, const unsigned long long* __restrict__ seed_dev, uint32_t salt) {
  constexpr int E = 128 * G;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float4 v[G];
#pragma unroll
  for (int g = 0; g < G; ++g) v[g] = *(const float4*)(x + row * E + g * 128 + lane * 4);
  if (branch) {
    const uint32_t rowkey = drop_thresh ? ceb_drop_rowkey(seed_dev, salt, (uint32_t)row) : 0u;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const uint2 raw = *(const uint2*)(branch + row * E + g * 128 + lane * 4);
      const float2 b01 = __bfloat1622float2(*(const __nv_bfloat162*)&raw.x);
      const float2 b23 = __bfloat1622float2(*(const __nv_bfloat162*)&raw.y);
      float k0 = 1.f, k1 = 1.f, k2 = 1.f, k3 = 1.f;
      if (drop_thresh) {
        const uint32_t key = rowkey + (uint32_t)(g * 128 + lane * 4);
        k0 = ceb_hash32(key) >= drop_thresh ? inv_keep : 0.f;
        k1 = ceb_hash32(key + 1) >= drop_thresh ? inv_keep : 0.f;
        k2 = ceb_hash32(key + 2) >= drop_thresh ? inv_keep : 0.f;
        k3 = ceb_hash32(key + 3) >= drop_thresh ? inv_keep : 0.f;
      }
      v[g].x = fmaf(b01.x, k0, v[g].x); v[g].y = fmaf(b01.y, k1, v[g].y);
      v[g].z = fmaf(b23.x, k2, v[g].z); v[g].w = fmaf(b23.y, k3, v[g].w);
      *(float4*)(resid_out + row * E + g * 128 + lane * 4) = v[g];
    }
  }
  float s = 0.f, ss = 0.f;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    s += (v[g].x + v[g].y) + (v[g].z + v[g].w);
    ss = fmaf(v[g].x, v[g].x, fmaf(v[g].y, v[g].y, fmaf(v[g].z, v[g].z, fmaf(v[g].w, v[g].w, ss))));
  }
  s = warp_sum(s); ss = warp_sum(ss);
  const float mu = s * (1.f / E);
  const float rs = rsqrtf(fmaxf(ss * (1.f / E) - mu * mu, 0.f) + eps);
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const float4 gm = __ldg((const float4*)(gamma + g * 128 + lane * 4));
    const float4 bt = __ldg((const float4*)(beta + g * 128 + lane * 4));
    const __nv_bfloat162 lo = __floats2bfloat162_rn(fmaf((v[g].x - mu) * rs, gm.x, bt.x), fmaf((v[g].y - mu) * rs, gm.y, bt.y));
    const __nv_bfloat162 hi = __floats2bfloat162_rn(fmaf((v[g].z - mu) * rs, gm.z, bt.z), fmaf((v[g].w - mu) * rs, gm.w, bt.w));
    uint2 pk; pk.x = *(const uint32_t*)&lo; pk.y = *(const uint32_t*)&hi;
    *(uint2*)(y + row * E + g * 128 + lane * 4) = pk;
  }
}

static bool ln_vec_width(int E) { return E == 128 || E == 256 || E == 512 || E == 768 || E == 1024 || E == 1280; }

// launches the vector kernel when the width and the alignment allow it, else the scalar one
static void launch_ln_fwd(const float* x, const float* gamma, const float* beta, __nv_bfloat16* y, float* mean, float* rstd,
                          int64_t rows, int E, float eps, const __nv_bfloat16* branch, float* resid_out, uint32_t th,
                          float inv_keep, const unsigned long long* seed_dev, uint32_t salt, cudaStream_t st) {
  const unsigned grid = (unsigned)ceb_cdiv(rows, 8);
  const bool aligned = ((((uintptr_t)x | (uintptr_t)gamma | (uintptr_t)beta | (uintptr_t)y | (uintptr_t)branch |
                          (uintptr_t)resid_out) & 15) == 0);
#define CEB_LN_FWD_CASE(G_)                                                                                          \
  case G_: ln_fwd_vec_kernel<G_><<<grid, 256, 0, st>>>(x, gamma, beta, y, mean, rstd, rows, eps, branch, resid_out, \
                                                       th, inv_keep, seed_dev, salt); return;
  if (aligned && ln_vec_width(E)) {
    switch (E / 128) {
      CEB_LN_FWD_CASE(1) CEB_LN_FWD_CASE(2) CEB_LN_FWD_CASE(4) CEB_LN_FWD_CASE(6) CEB_LN_FWD_CASE(8) CEB_LN_FWD_CASE(10)
    }
  }
#undef CEB_LN_FWD_CASE
  ln_fwd_kernel<<<grid, 256, 0, st>>>(x, gamma, beta, y, mean, rstd, rows, E, eps, branch, resid_out, th, inv_keep, seed_dev,
                                      salt);
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)), g = dy * gamma ; dgamma += dy * xhat ; dbeta += dy
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
              const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dx,
              float* __restrict__ dgamma, float* __restrict__ dbeta, int64_t rows, int E) {
  extern __shared__ float sh[];                 // [2][E] block partials of dgamma / dbeta
  float* sg = sh;
  float* sb = sh + E;
  for (int j = threadIdx.x; j < 2 * E; j += blockDim.x) sh[j] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  for (int64_t row = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += (int64_t)gridDim.x * wpb) {
    const float mu = mean[row], rs = rstd[row];
    const float* xr = x + row * E;
    const __nv_bfloat16* gr = dy + row * E;
    float a = 0.f, b = 0.f;
    for (int j = lane; j < E; j += 32) {
      const float g = __bfloat162float(gr[j]) * __ldg(gamma + j);
      const float xh = (xr[j] - mu) * rs;
      a += g; b = fmaf(g, xh, b);
    }
    a = warp_sum(a) / E; b = warp_sum(b) / E;
    for (int j = lane; j < E; j += 32) {
      const float d = __bfloat162float(gr[j]);
      const float xh = (xr[j] - mu) * rs;
      dx[row * E + j] = rs * (d * __ldg(gamma + j) - a - xh * b);
      atomicAdd(sg + j, d * xh);
      atomicAdd(sb + j, d);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < E; j += blockDim.x) {
    if (sg[j] != 0.f) atomicAdd(dgamma + j, sg[j]);
    if (sb[j] != 0.f) atomicAdd(dbeta + j, sb[j]);
  }
}

CEB_API int ceb_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y_bf16, float* mean,
                              float* rstd, int64_t rows, int E, float eps, void* stream) {
  CEB_REQUIRE(x && gamma && beta && y_bf16 && mean && rstd && rows > 0 && E > 0, "ceb_layernorm_fwd: bad arguments");
  launch_ln_fwd(x, gamma, beta, (__nv_bfloat16*)y_bf16, mean, rstd, rows, E, eps, nullptr, nullptr, 0, 1.f, nullptr, 0,
                (cudaStream_t)stream);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Residual add fused into the LayerNorm that follows it: resid_out (fp32) = x (fp32) + branch (bf16),
// y (bf16) = LayerNorm(resid_out).  One pass instead of an add kernel plus a LayerNorm kernel.
CEB_API int ceb_add_layernorm_fwd(const float* x, const void* branch_bf16, const float* gamma, const float* beta,
                                  float* resid_out, void* y_bf16, float* mean, float* rstd, int64_t rows, int E,
                                  float eps, void* stream) {
  return ceb_add_layernorm_drop_fwd(x, branch_bf16, gamma, beta, resid_out, y_bf16, mean, rstd, rows, E, eps, 0.f,
                                    nullptr, 0, stream);
}

// Same with dropout on the branch before the add: resid_out = x + dropout_p(branch).  The mask is a pure
// function of (*seed_dev, salt, row, column) -- see common.cuh -- so nothing is stored for the backward.
CEB_API int ceb_add_layernorm_drop_fwd(const float* x, const void* branch_bf16, const float* gamma, const float* beta,
                                       float* resid_out, void* y_bf16, float* mean, float* rstd, int64_t rows, int E,
                                       float eps, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream) {
  CEB_REQUIRE(x && branch_bf16 && gamma && beta && resid_out && y_bf16 && mean && rstd && rows > 0 && E > 0,
              "ceb_add_layernorm_fwd: bad arguments");
  CEB_REQUIRE(p_drop >= 0.f && p_drop < 1.f && rows < (1ll << 32), "ceb_add_layernorm_fwd: p_drop=%f rows=%lld", p_drop, (long long)rows);
  const uint32_t th = ceb_drop_thresh(p_drop);
  launch_ln_fwd(x, gamma, beta, (__nv_bfloat16*)y_bf16, mean, rstd, rows, E, eps, (const __nv_bfloat16*)branch_bf16,
                resid_out, th, th ? 1.f / (1.f - p_drop) : 1.f, (const unsigned long long*)seed_dev, salt,
                (cudaStream_t)stream);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Register-accumulating variant for E = 128 * G (GPT-2: 768, 1024): one warp per row, every lane owns
// G float4 column groups, keeps x-hat / dy in registers between the two passes and accumulates its
// dgamma / dbeta partials in registers over all rows the warp walks; the block combines its 8 warps
// in shared memory once and issues one global atomic per column.
template <int G>
__global__ void __launch_bounds__(256, 2)
ln_bwd_vec_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
                  const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ dx,
                  float* __restrict__ dgamma, float* __restrict__ dbeta, int64_t rows,
                  const float* __restrict__ dres, __nv_bfloat16* __restrict__ dx_bf16, uint32_t drop_thresh = 0,
                  float inv_keep = 1.f, const unsigned long long* __restrict__ seed_dev = nullptr, uint32_t salt = 0) {
  // Two passes over the row: the first only accumulates the two row sums, the second RELOADS x / dy (they are
  // still in L1: a row is 4.5 KB) instead of keeping them in registers -- 212 registers and one block per SM
  // (12 % of the warp slots, latency-bound at a quarter of the DRAM bandwidth) became ~100 and two blocks.
  constexpr int E = 128 * G;
  __shared__ float sh[2 * E];
  for (int j = threadIdx.x; j < 2 * E; j += blockDim.x) sh[j] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  float4 ag[G], ab[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    ag[g] = make_float4(0.f, 0.f, 0.f, 0.f);
    ab[g] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  auto load_d = [&](const __nv_bfloat16* gr, int g) {
    const uint2 raw = *(const uint2*)(gr + g * 128 + lane * 4);
    const float2 d01 = __bfloat1622float2(*(const __nv_bfloat162*)&raw.x);
    const float2 d23 = __bfloat1622float2(*(const __nv_bfloat162*)&raw.y);
    return make_float4(d01.x, d01.y, d23.x, d23.y);
  };
  for (int64_t row = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += (int64_t)gridDim.x * wpb) {
    const float mu = mean[row], rs = rstd[row];
    const float* xr = x + row * E;
    const __nv_bfloat16* gr = dy + row * E;
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float4 xv = *(const float4*)(xr + g * 128 + lane * 4);
      const float4 d = load_d(gr, g);
      const float4 gm = __ldg((const float4*)(gamma + g * 128 + lane * 4));
      const float g0 = d.x * gm.x, g1 = d.y * gm.y, g2 = d.z * gm.z, g3 = d.w * gm.w;
      a += g0 + g1 + g2 + g3;
      b = fmaf(g0, (xv.x - mu) * rs, fmaf(g1, (xv.y - mu) * rs, fmaf(g2, (xv.z - mu) * rs, fmaf(g3, (xv.w - mu) * rs, b))));
    }
    a = warp_sum(a) * (1.f / E); b = warp_sum(b) * (1.f / E);
    const uint32_t rowkey = drop_thresh ? ceb_drop_rowkey(seed_dev, salt, (uint32_t)row) : 0u;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float4 xv = *(const float4*)(xr + g * 128 + lane * 4);
      const float4 d = load_d(gr, g);
      const float4 gm = __ldg((const float4*)(gamma + g * 128 + lane * 4));
      const float4 xh = make_float4((xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs);
      float4 o;
      o.x = rs * (d.x * gm.x - a - xh.x * b);
      o.y = rs * (d.y * gm.y - a - xh.y * b);
      o.z = rs * (d.z * gm.z - a - xh.z * b);
      o.w = rs * (d.w * gm.w - a - xh.w * b);
      if (dres) {      // gradient arriving through the residual stream joins here (fused add)
        const float4 r4 = *(const float4*)(dres + row * E + g * 128 + lane * 4);
        o.x += r4.x; o.y += r4.y; o.z += r4.z; o.w += r4.w;
      }
      *(float4*)(dx + row * E + g * 128 + lane * 4) = o;
      if (dx_bf16) {   // the same gradient for the bf16 branch that was added to the stream (through its dropout mask)
        if (drop_thresh) {
          const uint32_t key = rowkey + (uint32_t)(g * 128 + lane * 4);
          o.x *= ceb_hash32(key) >= drop_thresh ? inv_keep : 0.f;
          o.y *= ceb_hash32(key + 1) >= drop_thresh ? inv_keep : 0.f;
          o.z *= ceb_hash32(key + 2) >= drop_thresh ? inv_keep : 0.f;
          o.w *= ceb_hash32(key + 3) >= drop_thresh ? inv_keep : 0.f;
        }
        const __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
        uint2 pk; pk.x = *(const uint32_t*)&lo; pk.y = *(const uint32_t*)&hi;
        *(uint2*)(dx_bf16 + row * E + g * 128 + lane * 4) = pk;
      }
      ag[g].x = fmaf(d.x, xh.x, ag[g].x); ag[g].y = fmaf(d.y, xh.y, ag[g].y);
      ag[g].z = fmaf(d.z, xh.z, ag[g].z); ag[g].w = fmaf(d.w, xh.w, ag[g].w);
      ab[g].x += d.x; ab[g].y += d.y; ab[g].z += d.z; ab[g].w += d.w;
    }
  }
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int j = g * 128 + lane * 4;
    atomicAdd(sh + j, ag[g].x); atomicAdd(sh + j + 1, ag[g].y); atomicAdd(sh + j + 2, ag[g].z); atomicAdd(sh + j + 3, ag[g].w);
    atomicAdd(sh + E + j, ab[g].x); atomicAdd(sh + E + j + 1, ab[g].y);
    atomicAdd(sh + E + j + 2, ab[g].z); atomicAdd(sh + E + j + 3, ab[g].w);
  }
  __syncthreads();
  for (int j = threadIdx.x; j < E; j += blockDim.x) {
    atomicAdd(dgamma + j, sh[j]);
    atomicAdd(dbeta + j, sh[E + j]);
  }
}

template <int G>
static void launch_ln_bwd_vec(const void* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                              float* dx, float* dgamma, float* dbeta, int64_t rows, cudaStream_t st,
                              const float* dres = nullptr, void* dx_bf16 = nullptr, uint32_t drop_thresh = 0,
                              float inv_keep = 1.f, const unsigned long long* seed_dev = nullptr, uint32_t salt = 0) {
  int blocks = ceb_num_sms() * 4;
  if (blocks > ceb_cdiv(rows, 8)) blocks = (int)ceb_cdiv(rows, 8);
  ln_bwd_vec_kernel<G><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows,
                                              dres, (__nv_bfloat16*)dx_bf16, drop_thresh, inv_keep, seed_dev, salt);
}


template <typename... A>
static void dispatch_ln_bwd_vec(int E, A... args) {
  switch (E / 128) {
    case 1: launch_ln_bwd_vec<1>(args...); break;
    case 2: launch_ln_bwd_vec<2>(args...); break;
    case 4: launch_ln_bwd_vec<4>(args...); break;
    case 6: launch_ln_bwd_vec<6>(args...); break;
    case 8: launch_ln_bwd_vec<8>(args...); break;
    default: launch_ln_bwd_vec<10>(args...); break;
  }
}

// Backward of ceb_add_layernorm_fwd: dx (fp32) = LayerNorm-backward(dy) + dres (the gradient that reaches the
// residual stream from later layers; may be NULL), dx_bf16 = the same values in bf16 (gradient of the added
// branch); dgamma / dbeta accumulated.  E must be one of 128, 256, 512, 768, 1024, 1280.
CEB_API int ceb_add_layernorm_bwd(const void* dy_bf16, const float* resid, const float* gamma, const float* mean,
                                  const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                                  float* dbeta, int64_t rows, int E, void* stream) {
  return ceb_add_layernorm_drop_bwd(dy_bf16, resid, gamma, mean, rstd, dres, dx, dx_bf16, dgamma, dbeta, rows, E, 0.f,
                                    nullptr, 0, stream);
}

// Backward of ceb_add_layernorm_drop_fwd: dx_bf16 (the branch's gradient) additionally passes through the
// regenerated dropout mask; dx (the stream's gradient) does not.
CEB_API int ceb_add_layernorm_drop_bwd(const void* dy_bf16, const float* resid, const float* gamma, const float* mean,
                                       const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                                       float* dbeta, int64_t rows, int E, float p_drop, const uint64_t* seed_dev,
                                       uint32_t salt, void* stream) {
  CEB_REQUIRE(dy_bf16 && resid && gamma && mean && rstd && dx && dgamma && dbeta && rows > 0, "ceb_add_layernorm_bwd: bad arguments");
  CEB_REQUIRE(ln_vec_width(E), "ceb_add_layernorm_bwd: unsupported width E=%d", E);
  CEB_REQUIRE(((((uintptr_t)dy_bf16) | ((uintptr_t)resid) | ((uintptr_t)dx) | ((uintptr_t)gamma) | ((uintptr_t)dres) |
                ((uintptr_t)dx_bf16)) & 15) == 0, "ceb_add_layernorm_bwd: tensors must be 16-byte aligned");
  CEB_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "ceb_add_layernorm_bwd: p_drop=%f", p_drop);
  const uint32_t th = ceb_drop_thresh(p_drop);
  dispatch_ln_bwd_vec(E, dy_bf16, resid, gamma, mean, rstd, dx, dgamma, dbeta, rows, (cudaStream_t)stream, dres, dx_bf16,
                      th, th ? 1.f / (1.f - p_drop) : 1.f, (const unsigned long long*)seed_dev, salt);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dgamma / dbeta are accumulated (+=)
CEB_API int ceb_layernorm_bwd(const void* dy_bf16, const float* x, const float* gamma, const float* mean,
                              const float* rstd, float* dx, float* dgamma, float* dbeta, int64_t rows, int E,
                              void* stream) {
  CEB_REQUIRE(dy_bf16 && x && gamma && mean && rstd && dx && dgamma && dbeta && rows > 0, "ceb_layernorm_bwd: bad arguments");
  CEB_REQUIRE(E <= 4096, "ceb_layernorm_bwd: E=%d too wide for the shared-memory partials", E);
  const bool aligned = ((((uintptr_t)dy_bf16 | (uintptr_t)x | (uintptr_t)dx | (uintptr_t)gamma) & 15) == 0);
  if (aligned && (E == 768 || E == 1024 || E == 1280 || E == 512 || E == 256 || E == 128)) {
    cudaStream_t st = (cudaStream_t)stream;
    switch (E / 128) {
      case 1: launch_ln_bwd_vec<1>(dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, st); break;
      case 2: launch_ln_bwd_vec<2>(dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, st); break;
      case 4: launch_ln_bwd_vec<4>(dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, st); break;
      case 6: launch_ln_bwd_vec<6>(dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, st); break;
      case 8: launch_ln_bwd_vec<8>(dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, st); break;
      default: launch_ln_bwd_vec<10>(dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, st); break;
    }
    CEB_COUNT_LAUNCH(1);
    CEB_CHECK_LAUNCH();
    return 0;
  }
  int blocks = ceb_num_sms() * 2;
  if (blocks > ceb_cdiv(rows, 8)) blocks = (int)ceb_cdiv(rows, 8);
  ln_bwd_kernel<<<blocks, 256, 2 * E * sizeof(float), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, E);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ tanh-GELU (bf16)
// tanh.approx.f32 (one MUFU op, relative error ~2^-11) is below bf16's 2^-8 output precision.
// (ceb_gelu_tanh / ceb_gelu_tanh_grad live in common.cuh: the GEMM epilogues apply them too)
__device__ __forceinline__ float gelu_tanh(float x) { return ceb_gelu_tanh(x); }
__device__ __forceinline__ float gelu_tanh_grad(float x) { return ceb_gelu_tanh_grad(x); }
__device__ __forceinline__ uint32_t gelu_pair(uint32_t xr, uint32_t gr, bool bwd) {
  const float2 v = __bfloat1622float2(*(const __nv_bfloat162*)&xr);
  float2 r;
  if (bwd) {
    const float2 g = __bfloat1622float2(*(const __nv_bfloat162*)&gr);
    r.x = g.x * gelu_tanh_grad(v.x); r.y = g.y * gelu_tanh_grad(v.y);
  } else {
    r.x = gelu_tanh(v.x); r.y = gelu_tanh(v.y);
  }
  const __nv_bfloat162 h = __floats2bfloat162_rn(r.x, r.y);
  return *(const uint32_t*)&h;
}
// VEC = 4: 16-byte accesses (8 bf16) when all pointers are 16-byte aligned; VEC = 1: bf16 pairs.
template <int VEC>
__global__ void __launch_bounds__(256)
gelu_kernel(const uint32_t* __restrict__ x, const uint32_t* __restrict__ dy, uint32_t* __restrict__ out, int64_t n2) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * VEC;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC; i < n2; i += stride) {
    if (VEC == 4 && i + 4 <= n2) {
      const uint4 xv = *(const uint4*)(x + i);
      uint4 gv = make_uint4(0, 0, 0, 0);
      if (dy) gv = *(const uint4*)(dy + i);
      uint4 o;
      o.x = gelu_pair(xv.x, gv.x, dy != nullptr); o.y = gelu_pair(xv.y, gv.y, dy != nullptr);
      o.z = gelu_pair(xv.z, gv.z, dy != nullptr); o.w = gelu_pair(xv.w, gv.w, dy != nullptr);
      *(uint4*)(out + i) = o;
    } else {
      for (int64_t j = i; j < n2 && j < i + VEC; ++j) out[j] = gelu_pair(x[j], dy ? dy[j] : 0u, dy != nullptr);
    }
  }
}

// out = gelu(x) when dy is NULL, else out = dy * gelu'(x); bf16, n even
CEB_API int ceb_gelu_bf16(const void* x, const void* dy, void* out, int64_t n, void* stream) {
  CEB_REQUIRE(x && out && n > 0 && n % 2 == 0, "ceb_gelu_bf16: need n even");
  const bool aligned = ((((uintptr_t)x | (uintptr_t)dy | (uintptr_t)out) & 15) == 0);
  if (aligned)
    gelu_kernel<4><<<ceb_grid_for(ceb_cdiv(n / 2, 4), 256), 256, 0, (cudaStream_t)stream>>>(
        (const uint32_t*)x, (const uint32_t*)dy, (uint32_t*)out, n / 2);
  else
    gelu_kernel<1><<<ceb_grid_for(n / 2, 256), 256, 0, (cudaStream_t)stream>>>(
        (const uint32_t*)x, (const uint32_t*)dy, (uint32_t*)out, n / 2);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ column sums (bias gradients)
// out[c] += sum_r x[r][c] for bf16 x [rows][ld]; cols % 8 == 0, 16-byte aligned rows.  Replaces the
// `.float().sum(0)` pair (an fp32 copy of the whole gradient plus a reduction) behind every
// bias gradient.  Block = 32 column groups (8 columns = one 16-byte load each) x 8 row lanes.
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int64_t rows, int cols, float* __restrict__ out) {
  __shared__ float sh[8][256 + 8];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + tx) * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c0 < cols) {
    for (int64_t r = (int64_t)blockIdx.y * 8 + ty; r < rows; r += (int64_t)gridDim.y * 8) {
      const uint4 v = *(const uint4*)(x + r * ld + c0);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(*(const __nv_bfloat162*)&w[j]);
        acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) sh[ty][tx * 8 + j] = acc[j];
  __syncthreads();
  const int c = threadIdx.x;                 // 256 columns per block
  float s = 0.f;
#pragma unroll
  for (int y = 0; y < 8; ++y) s += sh[y][c];
  const int col = blockIdx.x * 256 + c;
  if (col < cols) atomicAdd(out + col, s);
}

CEB_API int ceb_colsum_bf16(const void* x_bf16, int64_t ld, int64_t rows, int cols, float* out, void* stream) {
  CEB_REQUIRE(x_bf16 && out && rows > 0 && cols > 0, "ceb_colsum_bf16: bad arguments");
  CEB_REQUIRE(cols % 8 == 0 && ld % 8 == 0 && (((uintptr_t)x_bf16) & 15) == 0,
              "ceb_colsum_bf16: need cols %% 8 == 0, ld %% 8 == 0 and a 16-byte aligned base (cols=%d ld=%lld)", cols,
              (long long)ld);
  const int gx = (int)ceb_cdiv(cols, 256);
  int gy = (int)ceb_cdiv(4 * ceb_num_sms(), gx);
  if (gy > ceb_cdiv(rows, 8)) gy = (int)ceb_cdiv(rows, 8);
  if (gy < 1) gy = 1;
  colsum_bf16_kernel<<<dim3(gx, gy), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x_bf16, ld, rows, cols, out);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}


// ------------------------------------------------------------------ classifier head + cross-entropy
// logits = scale * feat @ W^T  (feat bf16 [N][F], W fp32 [C][F], C <= 32, F % 64 == 0, F <= 2048),
// per-sample cross-entropy and top-1 correctness; with g != NULL the backward of sum_s g[s] * loss[s]:
// dfeat (bf16) and dW (+=, fp32).  One warp per sample, lane c owns logit c, W (and the dW partials)
// live in shared memory.  Replaces ~25 library launches of 2-7 us each (bf16 GEMM, scale, softmax,
// nll_loss, argmax, compare, their backward, two more GEMMs) behind the reference's
// `loss = criterion(model(x), y); acc = accuracy(...)` (cv_train.py:67-72) on a [N, 10] problem.
__device__ __forceinline__ float2 head_ld2(const __nv_bfloat16* p) { return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p)); }
__device__ __forceinline__ float2 head_ld2(const float* p) { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ void head_st2(__nv_bfloat16* p, float a, float b) { *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b); }
__device__ __forceinline__ void head_st2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }

// FT = feature / feature-gradient element type: bf16 (fast path) or float (TF32 / matched-precision path; the
// head itself is fp32 FMA either way, like the reference's cuBLAS SGEMM for this 512 -> C layer).
// Classes are distributed over lanes: lane l owns classes l, l + 32, ... (CL = ceil(C / 32) <= 4 per lane).
// dW_in_smem: the block accumulates its weight gradient in shared memory and adds it to the global gradient
// once (small heads); otherwise (the two [C][F] fp32 tables exceed shared memory) every lane reduces straight
// into global memory.
template <typename FT, int CL>
__global__ void __launch_bounds__(256)
linear_ce_head_kernel(const FT* __restrict__ feat, const float* __restrict__ W, float scale,
                      const int64_t* __restrict__ labels, int N, int F, int C, const float* __restrict__ g,
                      float* __restrict__ loss, float* __restrict__ correct, FT* __restrict__ dfeat,
                      float* __restrict__ dW, int dW_in_smem) {
  extern __shared__ float hs[];
  float* Ws = hs;                       // [C][F]
  float* dWs = hs + (size_t)C * F;      // [C][F], backward with dW_in_smem only
  const bool bwd = g != nullptr;
  for (int i = threadIdx.x; i < C * F; i += blockDim.x) {
    Ws[i] = W[i];
    if (bwd && dW_in_smem) dWs[i] = 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int J = F >> 6;                 // element pairs per lane
  for (int s = blockIdx.x * wpb + warp; s < N; s += gridDim.x * wpb) {
    float x[64];                        // F <= 2048 -> up to 32 pairs
    const FT* fr = feat + (size_t)s * F;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (j < J) {
        const float2 v = head_ld2(fr + (j * 32 + lane) * 2);
        x[2 * j] = v.x; x[2 * j + 1] = v.y;
      }
    }
    float mylogit[CL];
#pragma unroll
    for (int q = 0; q < CL; ++q) mylogit[q] = -INFINITY;
    for (int c = 0; c < C; ++c) {
      const float* wr = Ws + (size_t)c * F;
      float a = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < J) a = fmaf(x[2 * j], wr[j * 64 + 2 * lane], fmaf(x[2 * j + 1], wr[j * 64 + 2 * lane + 1], a));
      a = warp_sum(a) * scale;
#pragma unroll
      for (int q = 0; q < CL; ++q)
        if (c == q * 32 + lane) mylogit[q] = a;
    }
    // softmax statistics and argmax (first maximal index, like torch.argmax)
    float mx = -INFINITY;
    int am = 0x7fffffff;
#pragma unroll
    for (int q = 0; q < CL; ++q) {
      const int c = q * 32 + lane;
      if (c < C && (mylogit[q] > mx || (mylogit[q] == mx && c < am))) { mx = mylogit[q]; am = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oi = __shfl_xor_sync(0xffffffffu, am, o);
      if (om > mx || (om == mx && oi < am)) { mx = om; am = oi; }
    }
    float e = 0.f;
#pragma unroll
    for (int q = 0; q < CL; ++q)
      if (q * 32 + lane < C) e += __expf(mylogit[q] - mx);
    const float lse = mx + __logf(warp_sum(e));
    const int label = (int)labels[s];
    float at_label = 0.f;
#pragma unroll
    for (int q = 0; q < CL; ++q) {
      const float v = __shfl_sync(0xffffffffu, mylogit[q], label & 31);
      if ((label >> 5) == q) at_label = v;
    }
    if (!bwd) {
      if (lane == 0) {
        loss[s] = lse - at_label;
        correct[s] = (am == label) ? 1.f : 0.f;
      }
      continue;
    }
    // d raw logit (before the scale) of this lane's classes
    float dl[CL];
#pragma unroll
    for (int q = 0; q < CL; ++q) {
      const int c = q * 32 + lane;
      dl[q] = c < C ? g[s] * (__expf(mylogit[q] - lse) - (c == label ? 1.f : 0.f)) * scale : 0.f;
    }
    float dx[64];
#pragma unroll
    for (int j = 0; j < 64; ++j) dx[j] = 0.f;
    for (int c = 0; c < C; ++c) {
      float d = 0.f;
#pragma unroll
      for (int q = 0; q < CL; ++q) {
        const float v = __shfl_sync(0xffffffffu, dl[q], c & 31);
        if ((c >> 5) == q) d = v;
      }
      const float* wr = Ws + (size_t)c * F;
      float* dwr = (dW_in_smem ? dWs : dW) + (size_t)c * F;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if (j < J) {
          const int f = j * 64 + 2 * lane;
          dx[2 * j] = fmaf(d, wr[f], dx[2 * j]);
          dx[2 * j + 1] = fmaf(d, wr[f + 1], dx[2 * j + 1]);
          atomicAdd(dwr + f, d * x[2 * j]);
          atomicAdd(dwr + f + 1, d * x[2 * j + 1]);
        }
      }
    }
    FT* dr = dfeat + (size_t)s * F;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < J) head_st2(dr + (j * 32 + lane) * 2, dx[2 * j], dx[2 * j + 1]);
  }
  if (bwd && dW_in_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < C * F; i += blockDim.x) {
      const float v = dWs[i];
      if (v != 0.f) atomicAdd(dW + i, v);
    }
  }
}

template <typename FT, int CL>
static int linear_ce_head_launch(const FT* feat, const float* W, float scale, const int64_t* labels, int N, int F,
                                 int C, const float* g, float* loss, float* correct, FT* dfeat, float* dW, void* stream) {
  const size_t one = (size_t)C * F * sizeof(float);
  const int dW_in_smem = (g && 2 * one <= 200 * 1024) ? 1 : 0;
  const size_t smem = one * ((g && dW_in_smem) ? 2 : 1);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    CEB_CHECK_CUDA(cudaFuncSetAttribute(linear_ce_head_kernel<FT, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  int blocks = (int)ceb_cdiv(N, 8);
  if (blocks > ceb_num_sms()) blocks = ceb_num_sms();
  linear_ce_head_kernel<FT, CL><<<blocks, 256, smem, (cudaStream_t)stream>>>(feat, W, scale, labels, N, F, C, g, loss, correct,
                                                                          dfeat, dW, dW_in_smem);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

template <typename FT>
static int linear_ce_head_impl(const FT* feat, const float* W, float scale, const int64_t* labels, int N, int F,
                               int C, const float* g, float* loss, float* correct, FT* dfeat, float* dW, void* stream) {
  CEB_REQUIRE(feat && W && labels && N > 0, "ceb_linear_ce_head: bad arguments");
  CEB_REQUIRE(C >= 1 && C <= 128 && F % 64 == 0 && F >= 64 && F <= 2048 && (size_t)C * F * 4 <= 224 * 1024,
              "ceb_linear_ce_head: need C <= 128, F %% 64 == 0, F <= 2048, C * F * 4 <= 224 KB (C=%d F=%d)", C, F);
  CEB_REQUIRE(g ? (dfeat && dW) : (loss && correct), "ceb_linear_ce_head: missing outputs");
  switch ((C + 31) / 32) {
    case 1: return linear_ce_head_launch<FT, 1>(feat, W, scale, labels, N, F, C, g, loss, correct, dfeat, dW, stream);
    case 2: return linear_ce_head_launch<FT, 2>(feat, W, scale, labels, N, F, C, g, loss, correct, dfeat, dW, stream);
    case 3: return linear_ce_head_launch<FT, 3>(feat, W, scale, labels, N, F, C, g, loss, correct, dfeat, dW, stream);
    default: return linear_ce_head_launch<FT, 4>(feat, W, scale, labels, N, F, C, g, loss, correct, dfeat, dW, stream);
  }
}

CEB_API int ceb_linear_ce_head(const void* feat_bf16, const float* W, float scale, const int64_t* labels, int N, int F,
                               int C, const float* g, float* loss, float* correct, void* dfeat_bf16, float* dW,
                               void* stream) {
  return linear_ce_head_impl<__nv_bfloat16>((const __nv_bfloat16*)feat_bf16, W, scale, labels, N, F, C, g, loss, correct,
                                            (__nv_bfloat16*)dfeat_bf16, dW, stream);
}

CEB_API int ceb_linear_ce_head_f32(const float* feat, const float* W, float scale, const int64_t* labels, int N, int F,
                                   int C, const float* g, float* loss, float* correct, float* dfeat, float* dW,
                                   void* stream) {
  return linear_ce_head_impl<float>(feat, W, scale, labels, N, F, C, g, loss, correct, dfeat, dW, stream);
}

// ------------------------------------------------------------------ GPT-2 input embedding (gather + add + dropout)
// x[row][E] (fp32) = dropout_p( wte[ids[row]] + wpe[row % T] + (tt ? wte[tt[row]] : 0) ), one warp per row.
// Replaces nn.Embedding x3, two adds and nn.Dropout of the reference model's GPT2Model.forward
// (pytorch_transformers, driven from /root/reference/CommEfficient/gpt2_train.py:88-99).
__global__ void __launch_bounds__(256)
embed_fwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ tt, const float* __restrict__ wte,
                 const float* __restrict__ wpe, float* __restrict__ x, int64_t rows, int T, int E, uint32_t drop_thresh,
                 float inv_keep, const unsigned long long* __restrict__ seed_dev, uint32_t salt) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float4* a = reinterpret_cast<const float4*>(wte + ids[row] * E);
  const float4* b = reinterpret_cast<const float4*>(wpe + (row % T) * E);
  const float4* c = tt ? reinterpret_cast<const float4*>(wte + tt[row] * E) : nullptr;
  float4* o = reinterpret_cast<float4*>(x + row * E);
  const uint32_t rowkey = drop_thresh ? ceb_drop_rowkey(seed_dev, salt, (uint32_t)row) : 0u;
  for (int j = lane; j < E / 4; j += 32) {
    float4 v = __ldg(a + j);
    const float4 pb = __ldg(b + j);
    v.x += pb.x; v.y += pb.y; v.z += pb.z; v.w += pb.w;
    if (c) { const float4 pc = __ldg(c + j); v.x += pc.x; v.y += pc.y; v.z += pc.z; v.w += pc.w; }
    if (drop_thresh) {
      const uint32_t key = rowkey + (uint32_t)(j * 4);
      v.x *= ceb_hash32(key) >= drop_thresh ? inv_keep : 0.f;
      v.y *= ceb_hash32(key + 1) >= drop_thresh ? inv_keep : 0.f;
      v.z *= ceb_hash32(key + 2) >= drop_thresh ? inv_keep : 0.f;
      v.w *= ceb_hash32(key + 3) >= drop_thresh ? inv_keep : 0.f;
    }
    o[j] = v;
  }
}

// dwte[ids[row]] += g, dwpe[row % T] += g, dwte[tt[row]] += g with g = dx[row] through the regenerated mask:
// 16-byte vector reductions straight into the flat fp32 gradient (no sort, no dense [rows][E] temporaries).
__global__ void __launch_bounds__(256)
embed_bwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ tt, const float* __restrict__ dx,
                 float* __restrict__ dwte, float* __restrict__ dwpe, int64_t rows, int T, int E, uint32_t drop_thresh,
                 float inv_keep, const unsigned long long* __restrict__ seed_dev, uint32_t salt) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float* a = dwte + ids[row] * E;
  float* b = dwpe + (row % T) * E;
  float* c = tt ? dwte + tt[row] * E : nullptr;
  const float4* g4 = reinterpret_cast<const float4*>(dx + row * E);
  const uint32_t rowkey = drop_thresh ? ceb_drop_rowkey(seed_dev, salt, (uint32_t)row) : 0u;
  for (int j = lane; j < E / 4; j += 32) {
    float4 v = __ldg(g4 + j);
    if (drop_thresh) {
      const uint32_t key = rowkey + (uint32_t)(j * 4);
      v.x *= ceb_hash32(key) >= drop_thresh ? inv_keep : 0.f;
      v.y *= ceb_hash32(key + 1) >= drop_thresh ? inv_keep : 0.f;
      v.z *= ceb_hash32(key + 2) >= drop_thresh ? inv_keep : 0.f;
      v.w *= ceb_hash32(key + 3) >= drop_thresh ? inv_keep : 0.f;
    }
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a + j * 4), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(b + j * 4), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    if (c) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(c + j * 4), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  }
}

static int embed_check(const char* who, const void* ids, const void* w1, const void* w2, const void* x, int64_t rows, int T, int E) {
  CEB_REQUIRE(ids && w1 && w2 && x && rows > 0 && T > 0 && rows < (1ll << 32), "%s: bad arguments", who);
  CEB_REQUIRE(E % 4 == 0 && ((((uintptr_t)w1) | ((uintptr_t)w2) | ((uintptr_t)x)) & 15) == 0,
              "%s: E %% 4 == 0 and 16-byte aligned tables needed (E=%d)", who, E);
  return 0;
}

CEB_API int ceb_embed_fwd(const int64_t* ids, const int64_t* token_types, const float* wte, const float* wpe, float* x,
                          int64_t rows, int T, int E, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream) {
  if (int e = embed_check("ceb_embed_fwd", ids, wte, wpe, x, rows, T, E)) return e;
  const uint32_t th = ceb_drop_thresh(p_drop);
  embed_fwd_kernel<<<(unsigned)ceb_cdiv(rows, 8), 256, 0, (cudaStream_t)stream>>>(
      ids, token_types, wte, wpe, x, rows, T, E, th, th ? 1.f / (1.f - p_drop) : 1.f, (const unsigned long long*)seed_dev, salt);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dwte / dwpe are accumulated (+=): pass the flat gradient views.
CEB_API int ceb_embed_bwd(const int64_t* ids, const int64_t* token_types, const float* dx, float* dwte, float* dwpe,
                          int64_t rows, int T, int E, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream) {
  if (int e = embed_check("ceb_embed_bwd", ids, dwte, dwpe, dx, rows, T, E)) return e;
  const uint32_t th = ceb_drop_thresh(p_drop);
  embed_bwd_kernel<<<(unsigned)ceb_cdiv(rows, 8), 256, 0, (cudaStream_t)stream>>>(
      ids, token_types, dx, dwte, dwpe, rows, T, E, th, th ? 1.f / (1.f - p_drop) : 1.f, (const unsigned long long*)seed_dev, salt);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ segment-wise BatchNorm (NHWC bf16)
// BatchNorm whose statistics are per (client segment, channel): sample n belongs to segment seg[n], and
// every segment is normalised with the mean / variance of its own samples only -- the semantics of the
// reference's one-forward-per-client loop (models/resnet9.py:17-29, fixup_resnet18.py:143-146) inside ONE
// batched forward.  x, y, dy, dx: [N][L][C] bf16 (channels_last), C % 8 == 0; statistics fp32 [S][C].
// Thread = 8 channels of one pixel (16-byte accesses); block = 32 pixel lanes x 8 channel groups
// (64 channels); a block stays inside one sample, so its partial sums go to one segment.
constexpr int SEGBN_PIX = 256;          // pixels of one sample handled by a block

__device__ __forceinline__ void segbn_unpack(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ uint4 segbn_pack(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}

// 8 consecutive channels of one pixel: one 16-byte access for bf16, two for fp32 (the tf32 / fp32 paths keep
// fp32 activations)
__device__ __forceinline__ void segbn_load8(const __nv_bfloat16* p, float (&f)[8]) {
  segbn_unpack(*reinterpret_cast<const uint4*>(p), f);
}
__device__ __forceinline__ void segbn_load8(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void segbn_store8(__nv_bfloat16* p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = segbn_pack(f);
}
__device__ __forceinline__ void segbn_store8(float* p, const float (&f)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
}

// MODE 0: a[s][c] += sum x, b[s][c] += sum x^2.   MODE 1 (backward): a += sum dy, b += sum dy * xhat.
template <int MODE, typename T>
__global__ void __launch_bounds__(256)
segbn_sums_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                  const int64_t* __restrict__ seg, int L, int C, int chunks, const float* __restrict__ mean,
                  const float* __restrict__ rstd, float* __restrict__ a, float* __restrict__ b) {
  __shared__ float sa[32][65], sb[32][65];
  const int n = blockIdx.x / chunks, chunk = blockIdx.x - n * chunks;
  const int cg = threadIdx.x & 7, pl = threadIdx.x >> 3;
  const int c0 = blockIdx.y * 64 + cg * 8;
  const int s = (int)seg[n];
  float acc_a[8], acc_b[8], mu[8], rs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc_a[j] = 0.f; acc_b[j] = 0.f; mu[j] = 0.f; rs[j] = 1.f; }
  if (c0 < C) {
    if (MODE == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { mu[j] = mean[(size_t)s * C + c0 + j]; rs[j] = rstd[(size_t)s * C + c0 + j]; }
    }
    const int p_end = min(L, (chunk + 1) * SEGBN_PIX);
    for (int p = chunk * SEGBN_PIX + pl; p < p_end; p += 32) {
      const size_t off = ((size_t)n * L + p) * C + c0;
      float xv[8];
      segbn_load8(x + off, xv);
      if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc_a[j] += xv[j]; acc_b[j] = fmaf(xv[j], xv[j], acc_b[j]); }
      } else {
        float dv[8];
        segbn_load8(dy + off, dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc_a[j] += dv[j]; acc_b[j] = fmaf(dv[j], (xv[j] - mu[j]) * rs[j], acc_b[j]); }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { sa[pl][cg * 8 + j] = acc_a[j]; sb[pl][cg * 8 + j] = acc_b[j]; }
  __syncthreads();
  if (threadIdx.x < 128) {
    const int c = threadIdx.x & 63;
    const bool second = threadIdx.x >= 64;
    float t = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) t += second ? sb[r][c] : sa[r][c];
    const int cc = blockIdx.y * 64 + c;
    if (cc < C) atomicAdd((second ? b : a) + (size_t)s * C + cc, t);
  }
}

// mean = a / cnt, rstd = 1 / sqrt(b / cnt - mean^2 + eps); var (biased) also returned; cnt = samples[s] * L
__global__ void segbn_finalize_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                      const float* __restrict__ samples, int L, int S, int C, float eps,
                                      float* __restrict__ mean, float* __restrict__ var, float* __restrict__ rstd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= S * C) return;
  const float cnt = fmaxf(samples[i / C] * (float)L, 1.f);
  const float m = a[i] / cnt;
  const float v = fmaxf(b[i] / cnt - m * m, 0.f);
  mean[i] = m; var[i] = v; rstd[i] = rsqrtf(v + eps);
}

// MODE 0: y = (x - mean) * rstd * gamma + beta.
// MODE 1: dx = gamma * rstd * (dy - A/cnt - xhat * B/cnt)   (A, B from segbn_sums_kernel<1>)
template <int MODE, typename T>
__global__ void __launch_bounds__(256)
segbn_apply_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                   const int64_t* __restrict__ seg, int L, int C, int64_t total8, const float* __restrict__ mean,
                   const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                   const float* __restrict__ A, const float* __restrict__ B, const float* __restrict__ samples,
                   T* __restrict__ out) {
  const int c8 = C >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total8; i += stride) {
    const int c0 = (int)(i % c8) * 8;
    const int64_t pix = i / c8;
    const int n = (int)(pix / L);
    const int s = (int)seg[n];
    const size_t off = (size_t)pix * C + c0, so = (size_t)s * C + c0;
    float xv[8], o[8];
    segbn_load8(x + off, xv);
    if (MODE == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        o[j] = fmaf((xv[j] - mean[so + j]) * rstd[so + j], __ldg(gamma + c0 + j), __ldg(beta + c0 + j));
    } else {
      float dv[8];
      segbn_load8(dy + off, dv);
      const float inv = 1.f / fmaxf(samples[s] * (float)L, 1.f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float rs = rstd[so + j];
        const float xh = (xv[j] - mean[so + j]) * rs;
        o[j] = __ldg(gamma + c0 + j) * rs * (dv[j] - A[so + j] * inv - xh * B[so + j] * inv);
      }
    }
    segbn_store8(out + off, o);
  }
}

static int segbn_check(const char* who, const void* x, int64_t N, int L, int C, int S) {
  CEB_REQUIRE(x && N > 0 && L > 0 && S > 0, "%s: bad arguments", who);
  CEB_REQUIRE(C % 8 == 0 && (((uintptr_t)x) & 15) == 0, "%s: need C %% 8 == 0 and 16-byte aligned tensors (C=%d)", who, C);
  return 0;
}

// Forward.  sums: [2][S][C] fp32 scratch (zeroed here); mean / var / rstd: [S][C] outputs; samples: [S]
// fp32 samples per segment; seg: [N] int64 segment of every sample.
template <typename T>
static int segbn_fwd_impl(const char* who, const T* x, const int64_t* seg, const float* samples, int64_t N, int L, int C,
                          int S, const float* gamma, const float* beta, float eps, float* sums, float* mean, float* var,
                          float* rstd, T* y, void* stream) {
  if (int e = segbn_check(who, x, N, L, C, S)) return e;
  CEB_REQUIRE(seg && samples && gamma && beta && sums && mean && var && rstd && y, "%s: null pointer", who);
  cudaStream_t st = (cudaStream_t)stream;
  const int chunks = (int)ceb_cdiv(L, SEGBN_PIX);
  CEB_CHECK_CUDA(cudaMemsetAsync(sums, 0, (size_t)2 * S * C * sizeof(float), st));
  dim3 grid((unsigned)(N * chunks), (unsigned)ceb_cdiv(C, 64));
  segbn_sums_kernel<0, T><<<grid, 256, 0, st>>>(x, nullptr, seg, L, C, chunks, nullptr, nullptr, sums, sums + (size_t)S * C);
  segbn_finalize_kernel<<<(unsigned)ceb_cdiv((int64_t)S * C, 256), 256, 0, st>>>(sums, sums + (size_t)S * C, samples, L, S,
                                                                               C, eps, mean, var, rstd);
  const int64_t total8 = N * L * (C / 8);
  segbn_apply_kernel<0, T><<<ceb_grid_for(total8, 256), 256, 0, st>>>(x, nullptr, seg, L, C, total8, mean, rstd, gamma, beta,
                                                                     nullptr, nullptr, samples, y);
  CEB_COUNT_LAUNCH(3);
  CEB_CHECK_LAUNCH();
  return 0;
}

// Backward.  sums: [2][S][C] scratch; on return sums[0] = per-segment sum of dy (dbeta = its sum over
// segments), sums[1] = per-segment sum of dy * xhat (dgamma likewise).
template <typename T>
static int segbn_bwd_impl(const char* who, const T* dy, const T* x, const int64_t* seg, const float* samples, int64_t N,
                          int L, int C, int S, const float* gamma, const float* mean, const float* rstd, float* sums, T* dx,
                          void* stream) {
  if (int e = segbn_check(who, x, N, L, C, S)) return e;
  CEB_REQUIRE(dy && seg && samples && gamma && mean && rstd && sums && dx && (((uintptr_t)dy) & 15) == 0,
              "%s: null or misaligned pointer", who);
  cudaStream_t st = (cudaStream_t)stream;
  const int chunks = (int)ceb_cdiv(L, SEGBN_PIX);
  CEB_CHECK_CUDA(cudaMemsetAsync(sums, 0, (size_t)2 * S * C * sizeof(float), st));
  dim3 grid((unsigned)(N * chunks), (unsigned)ceb_cdiv(C, 64));
  segbn_sums_kernel<1, T><<<grid, 256, 0, st>>>(x, dy, seg, L, C, chunks, mean, rstd, sums, sums + (size_t)S * C);
  const int64_t total8 = N * L * (C / 8);
  segbn_apply_kernel<1, T><<<ceb_grid_for(total8, 256), 256, 0, st>>>(x, dy, seg, L, C, total8, mean, rstd, gamma, nullptr,
                                                                     sums, sums + (size_t)S * C, samples, dx);
  CEB_COUNT_LAUNCH(2);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_segment_bn_fwd(const void* x_bf16, const int64_t* seg, const float* samples, int64_t N, int L, int C,
                               int S, const float* gamma, const float* beta, float eps, float* sums, float* mean,
                               float* var, float* rstd, void* y_bf16, void* stream) {
  return segbn_fwd_impl("ceb_segment_bn_fwd", (const __nv_bfloat16*)x_bf16, seg, samples, N, L, C, S, gamma, beta, eps,
                        sums, mean, var, rstd, (__nv_bfloat16*)y_bf16, stream);
}
CEB_API int ceb_segment_bn_bwd(const void* dy_bf16, const void* x_bf16, const int64_t* seg, const float* samples,
                               int64_t N, int L, int C, int S, const float* gamma, const float* mean, const float* rstd,
                               float* sums, void* dx_bf16, void* stream) {
  return segbn_bwd_impl("ceb_segment_bn_bwd", (const __nv_bfloat16*)dy_bf16, (const __nv_bfloat16*)x_bf16, seg, samples, N,
                        L, C, S, gamma, mean, rstd, sums, (__nv_bfloat16*)dx_bf16, stream);
}
// fp32 activations (the tf32 / fp32 precision modes): same kernels, two 16-byte accesses per thread
CEB_API int ceb_segment_bn_f32_fwd(const float* x, const int64_t* seg, const float* samples, int64_t N, int L, int C,
                                   int S, const float* gamma, const float* beta, float eps, float* sums, float* mean,
                                   float* var, float* rstd, float* y, void* stream) {
  return segbn_fwd_impl("ceb_segment_bn_f32_fwd", x, seg, samples, N, L, C, S, gamma, beta, eps, sums, mean, var, rstd, y,
                        stream);
}
CEB_API int ceb_segment_bn_f32_bwd(const float* dy, const float* x, const int64_t* seg, const float* samples, int64_t N,
                                   int L, int C, int S, const float* gamma, const float* mean, const float* rstd,
                                   float* sums, float* dx, void* stream) {
  return segbn_bwd_impl("ceb_segment_bn_f32_bwd", dy, x, seg, samples, N, L, C, S, gamma, mean, rstd, sums, dx, stream);
}
