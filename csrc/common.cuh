// Shared helpers for the commefficient_b200 native library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include "../include/commefficient_b200.h"

#define CEB_API extern "C" __attribute__((visibility("default")))

// Error plumbing: every entry point returns 0 on success, non-zero otherwise and
// leaves a human-readable message retrievable with ceb_last_error().
void ceb_set_error(const char* fmt, ...);

#define CEB_CHECK_CUDA(expr)                                                        \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      ceb_set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,            \
                    cudaGetErrorString(_e));                                        \
      return 1;                                                                     \
    }                                                                               \
  } while (0)

#define CEB_REQUIRE(cond, ...)                                                      \
  do {                                                                              \
    if (!(cond)) {                                                                  \
      ceb_set_error(__VA_ARGS__);                                                   \
      return 2;                                                                     \
    }                                                                               \
  } while (0)

#define CEB_CHECK_LAUNCH() CEB_CHECK_CUDA(cudaGetLastError())

// Launch counter: every kernel launch made by this library bumps it, so callers
// (bench.py "gpu_launches") can report how many of OUR kernels ran.
extern unsigned long long g_ceb_launches;
#define CEB_COUNT_LAUNCH(n) (g_ceb_launches += (n))

static inline int ceb_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

static inline int64_t ceb_cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Grid for a grid-stride loop: enough CTAs to fill the machine a few times over
// (148 SMs x resident CTAs), never more than the work needs.
static inline int ceb_grid_for(int64_t work_items, int threads, int ctas_per_sm = 8) {
  int64_t need = ceb_cdiv(work_items, threads);
  int64_t cap = (int64_t)ceb_num_sms() * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void red_add_f32(float* addr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}

// ---- counter-hash dropout shared by the attention, embedding and add+LayerNorm kernels ----------------
// An element keeps its value iff ceb_hash32(rowkey + column) >= thresh, thresh = p * 2^32; rowkey mixes the
// device-resident seed, the call site's salt and the row.  Pure function of its inputs: the backward pass
// regenerates the mask instead of storing it.
__device__ __forceinline__ uint32_t ceb_hash32(uint32_t x) {        // "lowbias32" integer finaliser
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t ceb_drop_rowkey(const unsigned long long* seed_dev, uint32_t salt, uint32_t row) {
  const unsigned long long seed = seed_dev ? __ldg(seed_dev) : 0ull;
  return ceb_hash32((uint32_t)seed ^ ceb_hash32((uint32_t)(seed >> 32) ^ salt ^ ceb_hash32(row)));
}
static inline uint32_t ceb_drop_thresh(float p_drop) {
  if (!(p_drop > 0.f)) return 0u;
  const double th = (double)p_drop * 4294967296.0;
  const uint32_t t = th >= 4294967295.0 ? 0xffffffffu : (uint32_t)th;
  return t == 0 ? 1u : t;
}

// tanh-GELU (GPT-2's activation).  tanh.approx.f32 is one MUFU op with relative error ~2^-11, below bf16's
// 2^-8 output precision.
__device__ __forceinline__ float ceb_tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ceb_gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + ceb_tanh_fast(u));
}
__device__ __forceinline__ float ceb_gelu_tanh_grad(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = ceb_tanh_fast(u);
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}

__device__ __forceinline__ uint32_t abs_key(float v) {
  // Monotone key of |v|; NaN maps to 0 so it is never preferred by top-k.
  uint32_t b = __float_as_uint(v) & 0x7fffffffu;
  return b > 0x7f800000u ? 0u : b;
}
