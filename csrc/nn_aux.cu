// Transformer-side kernels (sm_100a): fused softmax cross-entropy over a large vocabulary,
// LayerNorm forward/backward, tanh-GELU forward/backward.  They replace the ATen op chains the
// reference runs inside pytorch_transformers' GPT-2 (/root/reference/CommEfficient/gpt2_train.py:76-99:
// shifted LM cross-entropy with ignore_index -1 over 50 262 classes; LayerNorm / GELU in every block).
#include <cuda_bf16.h>
#include "common.cuh"

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
  for (int j = threadIdx.x; j < pairs; j += blockDim.x) {
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
              const float* __restrict__ grow) {
  const int64_t row = blockIdx.x;
  const int64_t label = labels[row];
  const bool ignored = label == ignore_index;
  const float g = ignored ? 0.f : grow[row];
  const float l = lse[row];
  const __nv_bfloat162* p2 = reinterpret_cast<const __nv_bfloat162*>(logits + row * ld);
  __nv_bfloat162* d2 = reinterpret_cast<__nv_bfloat162*>(dlogits + row * ld);
  const int pairs = V >> 1;
  for (int j = threadIdx.x; j < pairs; j += blockDim.x) {
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

CEB_API int ceb_softmax_ce_bwd(const void* logits_bf16, void* dlogits_bf16, int64_t ld, int64_t rows, int V,
                               const int64_t* labels, int64_t ignore_index, const float* lse, const float* grad_rows,
                               void* stream) {
  CEB_REQUIRE(logits_bf16 && dlogits_bf16 && labels && lse && grad_rows && rows > 0 && ld % 2 == 0, "ceb_softmax_ce_bwd: bad arguments");
  ce_bwd_kernel<<<(unsigned)rows, CE_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits_bf16,
      (__nv_bfloat16*)dlogits_bf16, ld, V, labels, ignore_index, lse, grad_rows);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ LayerNorm (one warp per row)
// y (bf16) = (x - mean) * rstd * gamma + beta ; x fp32 [rows][E]; mean / rstd saved for backward.
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              __nv_bfloat16* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, int64_t rows,
              int E, float eps) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + row * E;
  float s = 0.f, ss = 0.f;
  for (int j = lane; j < E; j += 32) { const float v = xr[j]; s += v; ss = fmaf(v, v, ss); }
  s = warp_sum(s); ss = warp_sum(ss);
  const float mu = s / E;
  const float rs = rsqrtf(fmaxf(ss / E - mu * mu, 0.f) + eps);
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
  for (int j = lane; j < E; j += 32)
    y[row * E + j] = __float2bfloat16((xr[j] - mu) * rs * __ldg(gamma + j) + __ldg(beta + j));
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
  ln_fwd_kernel<<<(unsigned)ceb_cdiv(rows, 8), 256, 0, (cudaStream_t)stream>>>(x, gamma, beta, (__nv_bfloat16*)y_bf16,
                                                                                mean, rstd, rows, E, eps);
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
  int blocks = ceb_num_sms() * 2;
  if (blocks > ceb_cdiv(rows, 8)) blocks = (int)ceb_cdiv(rows, 8);
  ln_bwd_kernel<<<blocks, 256, 2 * E * sizeof(float), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy_bf16, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, E);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ tanh-GELU (bf16)
__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
__global__ void __launch_bounds__(256)
gelu_kernel(const __nv_bfloat162* __restrict__ x, const __nv_bfloat162* __restrict__ dy, __nv_bfloat162* __restrict__ out,
            int64_t n2) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    const float2 v = __bfloat1622float2(x[i]);
    float2 r;
    if (dy) {
      const float2 g = __bfloat1622float2(dy[i]);
      r.x = g.x * gelu_tanh_grad(v.x); r.y = g.y * gelu_tanh_grad(v.y);
    } else {
      r.x = gelu_tanh(v.x); r.y = gelu_tanh(v.y);
    }
    out[i] = __floats2bfloat162_rn(r.x, r.y);
  }
}

// out = gelu(x) when dy is NULL, else out = dy * gelu'(x); bf16, n even
CEB_API int ceb_gelu_bf16(const void* x, const void* dy, void* out, int64_t n, void* stream) {
  CEB_REQUIRE(x && out && n > 0 && n % 2 == 0, "ceb_gelu_bf16: need n even");
  gelu_kernel<<<ceb_grid_for(n / 2, 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat162*)x, (const __nv_bfloat162*)dy,
                                                                         (__nv_bfloat162*)out, n / 2);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}
