// Small bf16 NHWC helpers around the tensor-core layers (sm_100a): weight casts / re-layouts,
// fused ReLU + 2x2 / 4x4 max-pool forward and its backward (which also applies the ReLU mask).
// Reference ops: nn.MaxPool2d + ReLU in /root/reference/CommEfficient/models/resnet9.py:48-56,95.
#include <cuda_bf16.h>
#include "common.cuh"

#define CEB_WPREP_MAX 16

// ------------------------------------------------------------------ casts / weight layouts
__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
  // parameters sit at arbitrary offsets of the flat buffer: vector path only when both sides allow it
  const bool aligned = (((uintptr_t)src & 15) == 0) && (((uintptr_t)dst & 7) == 0);
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
    if (aligned && i + 3 < n) {
      float4 v = *reinterpret_cast<const float4*>(src + i);
      __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
      uint2 pk; pk.x = *(uint32_t*)&a; pk.y = *(uint32_t*)&b;
      *reinterpret_cast<uint2*>(dst + i) = pk;
    } else {
      for (int64_t j = i; j < n && j < i + 4; ++j) dst[j] = __float2bfloat16(src[j]);
    }
  }
}

// dst (bf16) = src (fp32); n elements; src 16-byte aligned.
CEB_API int ceb_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
  CEB_REQUIRE(src && dst && n > 0, "ceb_cast_f32_bf16: bad arguments");
  cast_f32_bf16_kernel<<<ceb_grid_for(ceb_cdiv(n, 4), 256), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dst[c][r] (bf16, row stride ldo >= R) = src[r][c] (fp32 [R][C] row-major): the K-major copy of a
// weight that is stored MN-major for the GEMM at hand (GPT-2 Conv1D [in][out] in the forward pass,
// the tied LM-head / embedding matrix [V][E] in its data-gradient GEMM).  64 x 64 tiles through
// shared memory; loads are 256-byte row pieces, stores 128-byte row pieces of bf16 pairs.
__global__ void __launch_bounds__(256)
cast_transpose_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int R, int C, int64_t ldo) {
  __shared__ float tile[64][65];
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = (t >> 4) + 16 * i, c = (t & 15) * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      tile[r][c + j] = (r0 + r < R && c0 + c + j < C) ? src[(int64_t)(r0 + r) * C + c0 + c + j] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = (t >> 5) + 8 * i, r = (t & 31) * 2;
    if (c0 + c >= C) continue;
    __nv_bfloat16* o = dst + (int64_t)(c0 + c) * ldo + r0 + r;
    if (r0 + r + 1 < R) {
      *(__nv_bfloat162*)o = __floats2bfloat162_rn(tile[r][c], tile[r + 1][c]);
    } else if (r0 + r < R) {
      *o = __float2bfloat16(tile[r][c]);
    }
  }
}

CEB_API int ceb_cast_transpose_f32_bf16(const float* src, void* dst_bf16, int R, int C, int64_t ldo, void* stream) {
  CEB_REQUIRE(src && dst_bf16 && R > 0 && C > 0, "ceb_cast_transpose_f32_bf16: bad arguments");
  CEB_REQUIRE(ldo >= R && ldo % 2 == 0 && (((uintptr_t)dst_bf16) & 3) == 0,
              "ceb_cast_transpose_f32_bf16: need an even ldo >= R (ldo=%lld R=%d)", (long long)ldo, R);
  dim3 grid((unsigned)ceb_cdiv(C, 64), (unsigned)ceb_cdiv(R, 64));
  cast_transpose_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst_bf16, R, C, ldo);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dgrad weights: dst[ci][2-r][2-s][co] = src[co][r][s][ci]  (fp32 or bf16 source -> bf16)
template <typename T>
__global__ void __launch_bounds__(256)
w_rot180_t_kernel(const T* __restrict__ src, __nv_bfloat16* __restrict__ dst, int Cout, int Cin) {
  const int64_t total = (int64_t)Cout * 9 * Cin;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
    // o indexes dst: ((ci*3 + r2)*3 + s2)*Cout + co
    const int co = (int)(o % Cout);
    int64_t t = o / Cout;
    const int s2 = (int)(t % 3); t /= 3;
    const int r2 = (int)(t % 3);
    const int ci = (int)(t / 3);
    const int64_t si = (((int64_t)co * 3 + (2 - r2)) * 3 + (2 - s2)) * Cin + ci;
    dst[o] = __float2bfloat16((float)src[si]);
  }
}

CEB_API int ceb_conv_w_dgrad_layout(const float* w_krsc, void* dst_bf16, int Cout, int Cin, void* stream) {
  CEB_REQUIRE(w_krsc && dst_bf16 && Cout > 0 && Cin > 0, "ceb_conv_w_dgrad_layout: bad arguments");
  const int64_t total = (int64_t)Cout * 9 * Cin;
  w_rot180_t_kernel<float><<<ceb_grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w_krsc, (__nv_bfloat16*)dst_bf16, Cout, Cin);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ one-launch weight preparation
// For up to CEB_WPREP_MAX 3x3 conv layers at once: the bf16 K-major copy used by the forward conv
// and (optionally) the bf16 rotated / transposed copy used by the data-gradient conv -- what
// ceb_cast_f32_bf16 + ceb_conv_w_dgrad_layout do per layer, in ONE launch instead of two per layer
// (a 100-client ResNet-9 round is launch-latency bound on exactly such 3-5 us kernels).
__device__ __forceinline__ void store_cvt(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }
__device__ __forceinline__ void store_cvt(float* p, float v) { *p = v; }

struct WPrepTable {
  const float* src[CEB_WPREP_MAX];
  void* cast[CEB_WPREP_MAX];
  void* rot[CEB_WPREP_MAX];
  int cout[CEB_WPREP_MAX];
  int cin[CEB_WPREP_MAX];
  int blk_start[CEB_WPREP_MAX + 1];     // first block of every layer (prefix sum of its 32x32 tiles x 9 taps)
  int n;
};

// One block = one 32 (co) x 32 (ci) tile of one filter tap: read coalesced along ci, write the cast
// copy in place and the transposed (rotated) copy coalesced along co through shared memory.
template <typename OutT>
__global__ void __launch_bounds__(256)
wprep_kernel(WPrepTable t) {
  __shared__ float tile[32][33];
  int layer = 0;
  while (layer + 1 < t.n && (int)blockIdx.x >= t.blk_start[layer + 1]) ++layer;
  const float* __restrict__ src = t.src[layer];
  OutT* __restrict__ cast = (OutT*)t.cast[layer];
  OutT* __restrict__ rot = (OutT*)t.rot[layer];
  const int Cout = t.cout[layer], Cin = t.cin[layer];
  const int ci_tiles = (Cin + 31) / 32, co_tiles = (Cout + 31) / 32;
  int b = blockIdx.x - t.blk_start[layer];
  const int ci_t = b % ci_tiles; b /= ci_tiles;
  const int co_t = b % co_tiles;
  const int tap = b / co_tiles;                       // r*3 + s
  const int x = threadIdx.x & 31, y0 = threadIdx.x >> 5;
#pragma unroll
  for (int yy = 0; yy < 4; ++yy) {
    const int y = y0 + 8 * yy;
    const int co = co_t * 32 + y, ci = ci_t * 32 + x;
    float v = 0.f;
    if (co < Cout && ci < Cin) {
      const int64_t i = ((int64_t)co * 9 + tap) * Cin + ci;
      v = src[i];
      if (cast) store_cvt(cast + i, v);
    }
    tile[y][x] = v;
  }
  if (!rot) return;
  __syncthreads();
  const int tap_rot = 8 - tap;                        // (2-r)*3 + (2-s)
#pragma unroll
  for (int yy = 0; yy < 4; ++yy) {
    const int y = y0 + 8 * yy;
    const int ci = ci_t * 32 + y, co = co_t * 32 + x;
    if (co < Cout && ci < Cin) store_cvt(rot + ((int64_t)ci * 9 + tap_rot) * Cout + co, tile[x][y]);
  }
}

static int wprep_impl(int n_layers, const float* const* src, void* const* cast_bf16, void* const* rot_bf16,
                      const int* cout, const int* cin, void* stream, int f32) {
  CEB_REQUIRE(n_layers > 0 && n_layers <= CEB_WPREP_MAX && src && cout && cin, "ceb_conv_weights_prepare: 1..%d layers", CEB_WPREP_MAX);
  WPrepTable t{};
  t.n = n_layers;
  int blocks = 0;
  for (int i = 0; i < n_layers; ++i) {
    CEB_REQUIRE(src[i] && cout[i] > 0 && cin[i] > 0, "ceb_conv_weights_prepare: bad layer %d", i);
    t.src[i] = src[i];
    t.cast[i] = cast_bf16 ? cast_bf16[i] : nullptr;
    t.rot[i] = rot_bf16 ? rot_bf16[i] : nullptr;
    t.cout[i] = cout[i];
    t.cin[i] = cin[i];
    t.blk_start[i] = blocks;
    blocks += 9 * (int)ceb_cdiv(cout[i], 32) * (int)ceb_cdiv(cin[i], 32);
  }
  t.blk_start[n_layers] = blocks;
  if (f32) wprep_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(t);
  else wprep_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>(t);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_conv_weights_prepare(int n_layers, const float* const* src, void* const* cast_bf16, void* const* rot_bf16,
                                     const int* cout, const int* cin, void* stream) {
  return wprep_impl(n_layers, src, cast_bf16, rot_bf16, cout, cin, stream, 0);
}

// fp32 outputs (TF32 path): `copy` (nullable per layer) = plain copies, `rot` = data-gradient layouts.
CEB_API int ceb_conv_weights_prepare_f32(int n_layers, const float* const* src, void* const* copy_f32, void* const* rot_f32,
                                         const int* cout, const int* cin, void* stream) {
  return wprep_impl(n_layers, src, copy_f32, rot_f32, cout, cin, stream, 1);
}

// ------------------------------------------------------------------ max pooling (k = 2 or 4), 8 channels per thread
struct bf16x8 { uint4 v; };
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162 h0 = __floats2bfloat162_rn(f[0], f[1]), h1 = __floats2bfloat162_rn(f[2], f[3]);
  __nv_bfloat162 h2 = __floats2bfloat162_rn(f[4], f[5]), h3 = __floats2bfloat162_rn(f[6], f[7]);
  u.x = *(uint32_t*)&h0; u.y = *(uint32_t*)&h1; u.z = *(uint32_t*)&h2; u.w = *(uint32_t*)&h3;
  return u;
}

// Vector traits: one 16-byte access = 8 bf16 channels or 4 fp32 channels.
struct VecBF {
  static constexpr int kN = 8;
  __device__ static void load(const uint4* p, float (&f)[8]) { unpack8(__ldg(p), f); }
  __device__ static void store(uint4* p, const float (&f)[8]) { *p = pack8(f); }
};
struct VecF {
  static constexpr int kN = 4;
  __device__ static void load(const uint4* p, float (&f)[4]) {
    const uint4 u = __ldg(p);
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
  }
  __device__ static void store(uint4* p, const float (&f)[4]) {
    uint4 u; u.x = __float_as_uint(f[0]); u.y = __float_as_uint(f[1]); u.z = __float_as_uint(f[2]); u.w = __float_as_uint(f[3]);
    *p = u;
  }
};

template <int KP, typename V = VecBF>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const uint4* __restrict__ y, uint4* __restrict__ out, int N, int H, int W, int C8) {
  constexpr int VN = V::kN;
  const int Ho = H / KP, Wo = W / KP;
  const int64_t total = (int64_t)N * Ho * Wo * C8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
    const int c = (int)(o % C8);
    int64_t t = o / C8;
    const int wo = (int)(t % Wo); t /= Wo;
    const int ho = (int)(t % Ho);
    const int n = (int)(t / Ho);
    float best[VN];
#pragma unroll
    for (int i = 0; i < VN; ++i) best[i] = -INFINITY;
#pragma unroll
    for (int dy = 0; dy < KP; ++dy)
#pragma unroll
      for (int dx = 0; dx < KP; ++dx) {
        float f[VN];
        V::load(y + (((int64_t)n * H + ho * KP + dy) * W + wo * KP + dx) * C8 + c, f);
#pragma unroll
        for (int i = 0; i < VN; ++i) best[i] = fmaxf(best[i], f[i]);
      }
    V::store(out + o, best);
  }
}

// dz = route dp to the first arg-max of each KPxKP window of y; with relu_mask also zero where y <= 0
template <int KP, typename V = VecBF>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const uint4* __restrict__ y, const uint4* __restrict__ dp, uint4* __restrict__ dz,
                   int N, int H, int W, int C8, int relu_mask) {
  constexpr int VN = V::kN;
  const int Ho = H / KP, Wo = W / KP;
  const int64_t total = (int64_t)N * Ho * Wo * C8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
    const int c = (int)(o % C8);
    int64_t t = o / C8;
    const int wo = (int)(t % Wo); t /= Wo;
    const int ho = (int)(t % Ho);
    const int n = (int)(t / Ho);
    float best[VN];
    int arg[VN];
#pragma unroll
    for (int i = 0; i < VN; ++i) { best[i] = -INFINITY; arg[i] = 0; }
#pragma unroll
    for (int dy = 0; dy < KP; ++dy)
#pragma unroll
      for (int dx = 0; dx < KP; ++dx) {
        float f[VN];
        V::load(y + (((int64_t)n * H + ho * KP + dy) * W + wo * KP + dx) * C8 + c, f);
#pragma unroll
        for (int i = 0; i < VN; ++i)
          if (f[i] > best[i]) { best[i] = f[i]; arg[i] = dy * KP + dx; }
      }
    float g[VN];
    V::load(dp + o, g);
#pragma unroll
    for (int dy = 0; dy < KP; ++dy)
#pragma unroll
      for (int dx = 0; dx < KP; ++dx) {
        float outv[VN];
#pragma unroll
        for (int i = 0; i < VN; ++i)
          outv[i] = (arg[i] == dy * KP + dx && !(relu_mask && best[i] <= 0.f)) ? g[i] : 0.f;
        V::store(dz + (((int64_t)n * H + ho * KP + dy) * W + wo * KP + dx) * C8 + c, outv);
      }
  }
}

// Windows tile the map from the top-left corner; rows / columns beyond the last whole window are ignored
// (torch.nn.MaxPool2d's floor mode): Ho = H / k, Wo = W / k.
static int check_pool(const char* who, int N, int H, int W, int C, int k) {
  CEB_REQUIRE(N > 0 && C % 8 == 0 && (k == 2 || k == 3 || k == 4) && H >= k && W >= k,
              "%s: need C %% 8 == 0, k in {2,3,4}, H, W >= k (N=%d H=%d W=%d C=%d k=%d)", who, N, H, W, C, k);
  return 0;
}

template <typename V>
static void launch_pool_fwd(const void* y, void* out, int N, int H, int W, int Cv, int k, cudaStream_t st) {
  const int64_t total = (int64_t)N * (H / k) * (W / k) * Cv;
  const int grid = ceb_grid_for(total, 256);
  if (k == 2) maxpool_fwd_kernel<2, V><<<grid, 256, 0, st>>>((const uint4*)y, (uint4*)out, N, H, W, Cv);
  else if (k == 3) maxpool_fwd_kernel<3, V><<<grid, 256, 0, st>>>((const uint4*)y, (uint4*)out, N, H, W, Cv);
  else maxpool_fwd_kernel<4, V><<<grid, 256, 0, st>>>((const uint4*)y, (uint4*)out, N, H, W, Cv);
}
template <typename V>
static int launch_pool_bwd(const void* y, const void* dp, void* dz, int N, int H, int W, int Cv, int k, int relu_mask,
                           size_t elem_bytes_per_vec, cudaStream_t st) {
  if (H % k != 0 || W % k != 0)      // pixels outside every window get no gradient
    CEB_CHECK_CUDA(cudaMemsetAsync(dz, 0, (size_t)N * H * W * Cv * elem_bytes_per_vec, st));
  const int64_t total = (int64_t)N * (H / k) * (W / k) * Cv;
  const int grid = ceb_grid_for(total, 256);
  if (k == 2) maxpool_bwd_kernel<2, V><<<grid, 256, 0, st>>>((const uint4*)y, (const uint4*)dp, (uint4*)dz, N, H, W, Cv, relu_mask);
  else if (k == 3) maxpool_bwd_kernel<3, V><<<grid, 256, 0, st>>>((const uint4*)y, (const uint4*)dp, (uint4*)dz, N, H, W, Cv, relu_mask);
  else maxpool_bwd_kernel<4, V><<<grid, 256, 0, st>>>((const uint4*)y, (const uint4*)dp, (uint4*)dz, N, H, W, Cv, relu_mask);
  return 0;
}

// out[N,H/k,W/k,C] = maxpool_k(y[N,H,W,C]), bf16 NHWC
CEB_API int ceb_maxpool_nhwc_bf16(const void* y, void* out, int N, int H, int W, int C, int k, void* stream) {
  CEB_REQUIRE(y && out, "ceb_maxpool_nhwc_bf16: null pointer");
  if (int e = check_pool("ceb_maxpool_nhwc_bf16", N, H, W, C, k)) return e;
  launch_pool_fwd<VecBF>(y, out, N, H, W, C / 8, k, (cudaStream_t)stream);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dz[N,H,W,C] = maxpool_k backward of dp[N,H/k,W/k,C] through y (and through ReLU when relu_mask)
CEB_API int ceb_maxpool_bwd_nhwc_bf16(const void* y, const void* dp, void* dz, int N, int H, int W, int C, int k,
                                      int relu_mask, void* stream) {
  CEB_REQUIRE(y && dp && dz, "ceb_maxpool_bwd_nhwc_bf16: null pointer");
  if (int e = check_pool("ceb_maxpool_bwd_nhwc_bf16", N, H, W, C, k)) return e;
  if (int e = launch_pool_bwd<VecBF>(y, dp, dz, N, H, W, C / 8, k, relu_mask, 16, (cudaStream_t)stream)) return e;
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// fp32 NHWC variants (TF32 path): C % 4 == 0
CEB_API int ceb_maxpool_nhwc_f32(const float* y, float* out, int N, int H, int W, int C, int k, void* stream) {
  CEB_REQUIRE(y && out, "ceb_maxpool_nhwc_f32: null pointer");
  CEB_REQUIRE(N > 0 && C % 4 == 0 && (k == 2 || k == 3 || k == 4) && H >= k && W >= k, "ceb_maxpool_nhwc_f32: bad shape");
  launch_pool_fwd<VecF>(y, out, N, H, W, C / 4, k, (cudaStream_t)stream);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_maxpool_bwd_nhwc_f32(const float* y, const float* dp, float* dz, int N, int H, int W, int C, int k,
                                     int relu_mask, void* stream) {
  CEB_REQUIRE(y && dp && dz, "ceb_maxpool_bwd_nhwc_f32: null pointer");
  CEB_REQUIRE(N > 0 && C % 4 == 0 && (k == 2 || k == 3 || k == 4) && H >= k && W >= k, "ceb_maxpool_bwd_nhwc_f32: bad shape");
  if (int e = launch_pool_bwd<VecF>(y, dp, dz, N, H, W, C / 4, k, relu_mask, 16, (cudaStream_t)stream)) return e;
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dz = dy where y > 0 else 0 (ReLU backward), bf16, n % 8 == 0
template <typename V = VecBF>
__global__ void __launch_bounds__(256)
relu_bwd_kernel(const uint4* __restrict__ y, const uint4* __restrict__ dy, uint4* __restrict__ dz, int64_t n8) {
  constexpr int VN = V::kN;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float a[VN], g[VN];
    V::load(y + i, a);
    V::load(dy + i, g);
#pragma unroll
    for (int j = 0; j < VN; ++j) g[j] = a[j] > 0.f ? g[j] : 0.f;
    V::store(dz + i, g);
  }
}

CEB_API int ceb_relu_bwd_f32(const float* y, const float* dy, float* dz, int64_t n, void* stream) {
  CEB_REQUIRE(y && dy && dz && n > 0 && n % 4 == 0, "ceb_relu_bwd_f32: need n %% 4 == 0");
  relu_bwd_kernel<VecF><<<ceb_grid_for(n / 4, 256), 256, 0, (cudaStream_t)stream>>>((const uint4*)y, (const uint4*)dy, (uint4*)dz, n / 4);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_relu_bwd_bf16(const void* y, const void* dy, void* dz, int64_t n, void* stream) {
  CEB_REQUIRE(y && dy && dz && n > 0 && n % 8 == 0, "ceb_relu_bwd_bf16: need n %% 8 == 0");
  relu_bwd_kernel<VecBF><<<ceb_grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>((const uint4*)y, (const uint4*)dy, (uint4*)dz, n / 8);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ 3-channel stem (first layer)
// K = 27 is too thin for an implicit-GEMM tile, so the stem is lowered explicitly: one pass writes
// the im2col matrix A[pixel][64] (27 taps*channels, zero padded to one 128-byte swizzle row, fused
// with the fp32 -> bf16 conversion of the input) and the stem then IS a tensor-core GEMM with
// K = 64: forward  y = relu(A * W0pad^T)  (ceb_gemm_bf16_tn), weight gradient  dW0pad += dZ^T * A
// (ceb_gemm_bf16_mn_acc).  8 threads per pixel, 8 columns each: a warp writes four full rows.
//   x: [N,H,W,3] fp32 (channels_last view of the NCHW batch); column t = (r*3+s)*3 + ci.
// Block = 32 consecutive pixels of one image row x 8 column groups.  The three input rows the pixels'
// windows touch (with a one-pixel halo, zero outside the image) are staged in shared memory with
// coalesced loads; every thread then assembles its 8 output columns from there.
template <int CIN, typename OutT>
__global__ void __launch_bounds__(256)
im2col3x3_kernel(const float* __restrict__ x, OutT* __restrict__ A, int N, int H, int W) {
  constexpr int KPAD = sizeof(OutT) == 2 ? 64 : 32;        // one 128-byte swizzle row
  constexpr int PER = KPAD / 8;                            // columns per thread (16 bytes)
  __shared__ float rows[3][(32 + 2) * CIN];
  const int segs = (W + 31) / 32;                          // 32-pixel segments per image row
  const int64_t blk = blockIdx.x;
  const int seg = (int)(blk % segs);
  const int hq = (int)((blk / segs) % H);
  const int64_t nb = blk / ((int64_t)segs * H);
  const int w0 = seg * 32;
  for (int i = threadIdx.x; i < 3 * 34 * CIN; i += blockDim.x) {
    const int r = i / (34 * CIN), rem = i - r * (34 * CIN);
    const int px = rem / CIN, ci = rem - px * CIN;
    const int hh = hq + r - 1, wc = w0 + px - 1;
    float v = 0.f;
    if (hh >= 0 && hh < H && wc >= 0 && wc < W) v = __ldg(x + ((nb * H + hh) * W + wc) * CIN + ci);
    rows[r][rem] = v;
  }
  __syncthreads();
  const int g = threadIdx.x & 7, pl = threadIdx.x >> 3;
  const int wq = w0 + pl;
  if (wq >= W) return;
  float v[PER];
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int t = g * PER + j;                             // column (r*3 + s)*CIN + ci
    float val = 0.f;
    if (t < 9 * CIN) {
      const int r = t / (3 * CIN), rem = t - r * (3 * CIN);   // rem = s*CIN + ci: consecutive floats of the row
      val = rows[r][pl * CIN + rem];
    }
    v[j] = val;
  }
  const int64_t p = (nb * H + hq) * W + wq;
  if constexpr (sizeof(OutT) == 2) {
    *reinterpret_cast<uint4*>(A + p * KPAD + g * PER) = pack8(v);
  } else {
    *reinterpret_cast<float4*>(A + p * KPAD + g * PER) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// A[N*H*W][64] (bf16) = im2col of the 3x3 / pad 1 neighbourhoods of x[N,H,W,3] (fp32)
CEB_API int ceb_im2col3x3_c3(const float* x_nhwc, void* A_bf16, int N, int H, int W, void* stream) {
  CEB_REQUIRE(x_nhwc && A_bf16 && N > 0 && H > 0 && W > 0, "ceb_im2col3x3_c3: bad arguments");
  const int64_t blocks = (int64_t)N * H * ((W + 31) / 32);
  CEB_REQUIRE(blocks < (1ll << 31), "ceb_im2col3x3_c3: batch too large");
  im2col3x3_kernel<3, __nv_bfloat16><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x_nhwc, (__nv_bfloat16*)A_bf16, N, H, W);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// General stem lowering: x[N,H,W,Cin] fp32 with Cin in {1, 3}; out_f32 = 0: A[N*H*W][64] bf16, 1: A[N*H*W][32] fp32
// (9*Cin valid columns, zero padded to one 128-byte row either way).
CEB_API int ceb_im2col3x3(const float* x_nhwc, void* A, int N, int H, int W, int Cin, int out_f32, void* stream) {
  CEB_REQUIRE(x_nhwc && A && N > 0 && H > 0 && W > 0, "ceb_im2col3x3: bad arguments");
  CEB_REQUIRE(Cin == 1 || Cin == 3, "ceb_im2col3x3: Cin must be 1 or 3 (got %d)", Cin);
  const int64_t blocks = (int64_t)N * H * ((W + 31) / 32);
  CEB_REQUIRE(blocks < (1ll << 31), "ceb_im2col3x3: batch too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (Cin == 3 && out_f32) im2col3x3_kernel<3, float><<<(unsigned)blocks, 256, 0, st>>>(x_nhwc, (float*)A, N, H, W);
  else if (Cin == 3) im2col3x3_kernel<3, __nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(x_nhwc, (__nv_bfloat16*)A, N, H, W);
  else if (out_f32) im2col3x3_kernel<1, float><<<(unsigned)blocks, 256, 0, st>>>(x_nhwc, (float*)A, N, H, W);
  else im2col3x3_kernel<1, __nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(x_nhwc, (__nv_bfloat16*)A, N, H, W);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}


// ------------------------------------------------------------------ stride-2 helpers (any 16-byte-vector dtype)
// sub[n][ho][wo][:] = x[n][2 ho][2 wo][:]  (input of a 1x1 / stride-2 shortcut as a dense matrix)
__global__ void __launch_bounds__(256)
subsample2_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int N, int H, int W, int Ho, int Wo, int Cv) {
  const int64_t total = (int64_t)N * Ho * Wo * Cv;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
    const int c = (int)(o % Cv);
    int64_t t = o / Cv;
    const int wo = (int)(t % Wo); t /= Wo;
    const int ho = (int)(t % Ho);
    const int64_t n = t / Ho;
    out[o] = __ldg(x + ((n * H + 2 * ho) * W + 2 * wo) * Cv + c);
  }
}
// up[n][h][w][:] = g[n][h/2][w/2][:] where h, w are both even (and inside g), 0 elsewhere: the gradient of a
// stride-2 layer spread back onto the input grid, after which its data gradient is a stride-1 convolution
__global__ void __launch_bounds__(256)
upsample2_zero_kernel(const uint4* __restrict__ g, uint4* __restrict__ up, int N, int H, int W, int Ho, int Wo, int Cv) {
  const int64_t total = (int64_t)N * H * W * Cv;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
    const int c = (int)(o % Cv);
    int64_t t = o / Cv;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const int64_t n = t / H;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (!(h & 1) && !(w & 1) && (h >> 1) < Ho && (w >> 1) < Wo) v = __ldg(g + ((n * Ho + (h >> 1)) * Wo + (w >> 1)) * Cv + c);
    up[o] = v;
  }
}

// row_bytes = C * sizeof(element), a multiple of 16
CEB_API int ceb_subsample2_nhwc(const void* x, void* out, int N, int H, int W, int row_bytes, void* stream) {
  CEB_REQUIRE(x && out && N > 0 && H > 0 && W > 0 && row_bytes % 16 == 0, "ceb_subsample2_nhwc: bad arguments");
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2, Cv = row_bytes / 16;
  subsample2_kernel<<<ceb_grid_for((int64_t)N * Ho * Wo * Cv, 256), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)x, (uint4*)out, N, H, W, Ho, Wo, Cv);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

CEB_API int ceb_upsample2_zero_nhwc(const void* g, void* up, int N, int H, int W, int Ho, int Wo, int row_bytes, void* stream) {
  CEB_REQUIRE(g && up && N > 0 && H > 0 && W > 0 && row_bytes % 16 == 0, "ceb_upsample2_zero_nhwc: bad arguments");
  CEB_REQUIRE(2 * (Ho - 1) < H && 2 * (Wo - 1) < W, "ceb_upsample2_zero_nhwc: %dx%d does not fit %dx%d", Ho, Wo, H, W);
  const int Cv = row_bytes / 16;
  upsample2_zero_kernel<<<ceb_grid_for((int64_t)N * H * W * Cv, 256), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)g, (uint4*)up, N, H, W, Ho, Wo, Cv);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}
