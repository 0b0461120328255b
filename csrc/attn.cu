// Causal multi-head self-attention (head dim 64) on tcgen05 tensor cores, forward and backward, with
// dropout on the attention probabilities regenerated from a counter hash (no mask tensor).
//
//   forward :  O = dropout(softmax(Q K^T / 8 + causal mask)) V          per (batch, head)
//   backward:  dQ, dK, dV from dO (flash-attention recomputation: only O and the row log-sum-exp are kept)
//
// Inputs are read straight out of the fused projection output qkv[B*T][3E] (row = token, columns
// [q | k | v], head h at columns h*64 .. h*64+63 of each third) through ONE 2-D TMA map; the output is
// written as out[B*T][E] -- the layout the following projection GEMM reads -- so no head transpose is ever
// materialised.  Per 128 x 128 (query x key) tile:
//   S  = Q K^T              tcgen05.mma, both operands K-major from 128-byte-swizzled shared memory -> TMEM
//   P  = softmax / dropout  one thread per query row reads its TMEM lane (tcgen05.ld), online softmax in
//                           registers (exp2, running max / sum), bf16 P written back to swizzled shared memory
//   O += P V                tcgen05.mma with V as an MN-major operand (no transposed copy of V)
// The backward adds dP = dO V^T, dV += P^T dO, dK += dS^T Q, dQ += dS K with P^T / dS^T read as MN-major
// operands from the same shared-memory tiles that serve as K-major operands for dQ.
//
// Replaces F.scaled_dot_product_attention / the unfused matmul-softmax-dropout-matmul inside
// pytorch_transformers' GPT-2 block that /root/reference/CommEfficient/gpt2_train.py:88-99,262-273 trains.
#include "tc_common.cuh"

using namespace tc;

constexpr int AT_TILE = 128;
constexpr int AT_D = 64;
constexpr int AT_THREADS = 160;            // warps 0-3: one thread per tile row; warp 4: TMA + MMA issue
constexpr int AT_TILE_BYTES = AT_TILE * AT_D * 2;        // [128][64] bf16 = 16 KB (128-byte rows)
constexpr int AT_P_BYTES = AT_TILE * AT_TILE * 2;        // [128][128] bf16 = 2 x 16 KB sub-tiles of 64 columns

struct AttnParams {
  __nv_bfloat16* out;        // fwd: O [B*T][E];   bwd: dqkv [B*T][3E]
  const __nv_bfloat16* o;    // bwd: O   [B*T][E]
  const __nv_bfloat16* dout; // bwd: dO  [B*T][E]
  float* lse;                // [B*H][T], log2 domain (fwd writes, bwd reads)
  float* dq_acc;             // bwd, T > 128 only: fp32 [B*T][E] scratch owned tile-wise by one CTA
  int B, T, H, E;
  float scale;               // 1 / sqrt(64)
  float scale_log2;          // scale * log2(e)
  uint32_t drop_thresh;      // keep an attention probability iff hash >= drop_thresh (0: no dropout)
  float inv_keep;            // 1 / (1 - p)
  const uint64_t* seed_dev;  // device-resident seed (nullable -> 0): graph replays see the current value
  uint32_t salt;             // call-site id (layer), so layers draw different masks
};

__device__ __forceinline__ uint32_t at_hash(uint32_t x) { return ceb_hash32(x); }
// per-row key; element (row, kv) keeps its probability iff at_hash(rowkey + kv) >= thresh
__device__ __forceinline__ uint32_t at_rowkey(const AttnParams& p, int bh, int q) {
  return ceb_drop_rowkey((const unsigned long long*)p.seed_dev, p.salt, (uint32_t)(bh * p.T + q));
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// byte offset of the 16-byte chunk holding columns [col8*8, col8*8+8) of row r in a [128][128] bf16 tile made
// of two 128-byte-swizzled [128][64] sub-tiles
__device__ __forceinline__ uint32_t p_chunk_off(int r, int col8) {
  const int sub = col8 >> 3, ch = col8 & 7;
  return (uint32_t)(sub * AT_TILE_BYTES + r * 128 + ((ch ^ (r & 7)) << 4));
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *(uint32_t*)&h;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// K-major operand: 128 rows x 64 columns per sub-tile, k-th 16-column slice
__device__ __forceinline__ uint64_t desc_k(uint32_t base, int k) { return smem_desc_sw128(base + k * 32, 16, 1024); }
// MN-major operand whose reduction index is the ROW of a [rows][64] tile: k-th 16-row slice; `lbo` = bytes
// between 64-column chunks along M / N (only used when the operand is wider than 64)
__device__ __forceinline__ uint64_t desc_mn(uint32_t base, int k, uint32_t lbo) {
  return smem_desc_sw128(base + k * 16 * 128, lbo, 1024);
}

// ------------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(AT_THREADS, 3)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  // [Q 16K][K 16K][pad 16K][V 16K]; the bf16 probability tile P (32 KB) overwrites K + pad once S = Q K^T has
  // retired (K is dead by then), which keeps a CTA at 64 KB so three fit on an SM
  uint8_t* sQ = smem;
  uint8_t* sK = smem + AT_TILE_BYTES;
  uint8_t* sP = sK;
  uint8_t* sV = smem + 3 * AT_TILE_BYTES;
  uint64_t* bar_q = (uint64_t*)(smem + 4 * AT_TILE_BYTES);
  uint64_t* bar_kv = bar_q + 1;
  uint64_t* bar_s = bar_q + 2;
  uint64_t* bar_p = bar_q + 3;
  uint64_t* bar_o = bar_q + 4;
  uint32_t* tmem_slot = (uint32_t*)(bar_q + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.x;
  const int qb = (int)gridDim.y - 1 - (int)blockIdx.y;         // long rows (many key blocks) first
  const int b = bh / p.H, h = bh - b * p.H;
  const int E = p.E;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(bar_q, 1); mbar_init(bar_kv, 1); mbar_init(bar_s, 1); mbar_init(bar_p, AT_TILE); mbar_init(bar_o, 1);
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // the P V product lands on the first 64 columns of S: every thread has consumed its S row before P V is issued
  const uint32_t tmem_S = tmem_base, tmem_O = tmem_base;

  if (warp == 4) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = idesc_bf16(AT_TILE, AT_TILE, 0, 0);
      constexpr uint32_t idesc_pv = idesc_bf16(AT_TILE, AT_D, 0, 1);
      const int row0 = b * p.T;
      mbar_expect_tx(bar_q, AT_TILE_BYTES);
      tma_load_2d(sQ, &tmQKV, bar_q, h * AT_D, row0 + qb * AT_TILE);
      const uint32_t aQ = smem_u32(sQ), aK = smem_u32(sK), aV = smem_u32(sV), aP = smem_u32(sP);
      for (int j = 0; j <= qb; ++j) {
        if (j > 0) mbar_wait(bar_o, (j - 1) & 1);              // the previous P V has finished reading sV / sP
        mbar_expect_tx(bar_kv, 2 * AT_TILE_BYTES);
        tma_load_2d(sK, &tmQKV, bar_kv, E + h * AT_D, row0 + j * AT_TILE);
        tma_load_2d(sV, &tmQKV, bar_kv, 2 * E + h * AT_D, row0 + j * AT_TILE);
        if (j == 0) mbar_wait(bar_q, 0);
        mbar_wait(bar_kv, j & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tmem_S, desc_k(aQ, k), desc_k(aK, k), idesc_s, k != 0);
        umma_commit(bar_s);
        mbar_wait(bar_p, j & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16(tmem_O, desc_k(aP + (k >> 2) * AT_TILE_BYTES, k & 3), desc_mn(aV, k, 16), idesc_pv, k != 0);
        umma_commit(bar_o);
      }
    }
    __syncwarp();
  } else {
    const int r = threadIdx.x;                                // tile row == TMEM lane
    const int q_abs = qb * AT_TILE + r;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    const uint32_t rowkey = p.drop_thresh ? at_rowkey(p, bh, q_abs) : 0u;
    const uint32_t aP = smem_u32(sP);
    float m = -INFINITY, l = 0.f;
    float o[AT_D];
#pragma unroll
    for (int d = 0; d < AT_D; ++d) o[d] = 0.f;
    for (int j = 0; j <= qb; ++j) {
      mbar_wait(bar_s, j & 1);
      tc_fence_after();
      const int kv0 = j * AT_TILE;
      const bool diag = j == qb;
      // pass 1: running maximum (scores scaled into the log2 domain)
      float mx = m;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld_32x32(tmem_S + lane_base + c * 32, v);
#pragma unroll
        for (int t = 0; t < 32; ++t) {
          const bool valid = !diag || (c * 32 + t <= r);
          if (valid) mx = fmaxf(mx, v[t] * p.scale_log2);
        }
      }
      const float alpha = exp2f(m - mx);                      // m = -inf on the first block -> 0
      l *= alpha;
      // pass 2: probabilities, row sum, dropout, bf16 P into swizzled shared memory
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld_32x32(tmem_S + lane_base + c * 32, v);
#pragma unroll
        for (int t = 0; t < 32; ++t) {
          const bool valid = !diag || (c * 32 + t <= r);
          float pr = valid ? exp2f(v[t] * p.scale_log2 - mx) : 0.f;
          l += pr;
          if (p.drop_thresh) pr = (at_hash(rowkey + (uint32_t)(kv0 + c * 32 + t)) >= p.drop_thresh) ? pr * p.inv_keep : 0.f;
          v[t] = pr;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(aP + p_chunk_off(r, c * 4 + g), pack_bf16(v[g * 8], v[g * 8 + 1]), pack_bf16(v[g * 8 + 2], v[g * 8 + 3]),
                       pack_bf16(v[g * 8 + 4], v[g * 8 + 5]), pack_bf16(v[g * 8 + 6], v[g * 8 + 7]));
      }
      m = mx;
      fence_async_smem();                                     // generic-proxy writes -> visible to the tensor core
      tc_fence_before();
      mbar_arrive(bar_p);
      mbar_wait(bar_o, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tmem_ld_32x32(tmem_O + lane_base + c * 32, v);
#pragma unroll
        for (int t = 0; t < 32; ++t) o[c * 32 + t] = fmaf(o[c * 32 + t], alpha, v[t]);
      }
      tc_fence_before();
    }
    if (q_abs < p.T) {
      const float inv_l = 1.f / l;
      __nv_bfloat16* dst = p.out + (size_t)(b * p.T + q_abs) * E + h * AT_D;
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        uint4 pk;
        pk.x = pack_bf16(o[g * 8] * inv_l, o[g * 8 + 1] * inv_l);
        pk.y = pack_bf16(o[g * 8 + 2] * inv_l, o[g * 8 + 3] * inv_l);
        pk.z = pack_bf16(o[g * 8 + 4] * inv_l, o[g * 8 + 5] * inv_l);
        pk.w = pack_bf16(o[g * 8 + 6] * inv_l, o[g * 8 + 7] * inv_l);
        *reinterpret_cast<uint4*>(dst + g * 8) = pk;
      }
      p.lse[(size_t)bh * p.T + q_abs] = m + log2f(l);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<128>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ backward
// One CTA per (batch, head).  Outer loop: key block j; inner loop: query blocks i >= j.
//   MULTI (T > 128): TMEM columns S [0,128) dP [128,256) dV [256,320) dK [320,384) dQ tile [384,448); dV / dK
//     accumulate over the inner loop; 128 KB of shared memory, one CTA per SM.
//   single tile (T <= 128): nothing accumulates across steps, so dV / dK / dQ reuse the S / dP columns (256
//     TMEM columns) and dS overwrites V (+16 KB) once dP = dO V^T has retired: 112 KB, two CTAs per SM -- the
//     kernel is a chain of short dependent phases (TMA, MMA, softmax math, MMA, stores), which only a second
//     resident CTA can overlap.
template <bool MULTI>
__global__ void __launch_bounds__(AT_THREADS, MULTI ? 1 : 2)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO, AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  if (!MULTI && smem != smem_raw) __trap();             // the single-tile variant is sized without alignment slack
  uint8_t* sQ = smem;
  uint8_t* sDO = smem + AT_TILE_BYTES;
  uint8_t* sK = smem + 2 * AT_TILE_BYTES;
  uint8_t* sV = smem + 3 * AT_TILE_BYTES;
  uint8_t* sDS = MULTI ? smem + 4 * AT_TILE_BYTES + AT_P_BYTES : sV;                 // dS * scale   [q][kv]
  uint8_t* sAd = MULTI ? smem + 4 * AT_TILE_BYTES : smem + 5 * AT_TILE_BYTES;        // dropped probabilities [q][kv]
  uint64_t* bar_kv = (uint64_t*)(smem + (MULTI ? 4 * AT_TILE_BYTES + 2 * AT_P_BYTES : 5 * AT_TILE_BYTES + AT_P_BYTES));
  constexpr int kTmemCols = MULTI ? 512 : 256;
  uint64_t* bar_qd = bar_kv + 1;
  uint64_t* bar_s = bar_kv + 2;
  uint64_t* bar_p = bar_kv + 3;
  uint64_t* bar_o = bar_kv + 4;
  uint32_t* tmem_slot = (uint32_t*)(bar_kv + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.x;
  const int b = bh / p.H, h = bh - b * p.H;
  const int E = p.E, T = p.T;
  const int nblk = (T + AT_TILE - 1) / AT_TILE;
  const int row0 = b * T;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    mbar_init(bar_kv, 1); mbar_init(bar_qd, 1); mbar_init(bar_s, 1); mbar_init(bar_p, AT_TILE); mbar_init(bar_o, 1);
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_dP = tmem_base + 128, tmem_dV = tmem_base + (MULTI ? 256 : 0),
                 tmem_dK = tmem_base + (MULTI ? 320 : 64), tmem_dQ = tmem_base + (MULTI ? 384 : 128);

  if (warp == 4) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = idesc_bf16(AT_TILE, AT_TILE, 0, 0);       // S, dP: both operands K-major
      constexpr uint32_t idesc_kv = idesc_bf16(AT_TILE, AT_D, 1, 1);         // dV, dK: A = tile^T (MN-major), B MN-major
      constexpr uint32_t idesc_dq = idesc_bf16(AT_TILE, AT_D, 0, 1);         // dQ: A K-major, B MN-major
      const uint32_t aQ = smem_u32(sQ), aDO = smem_u32(sDO), aK = smem_u32(sK), aV = smem_u32(sV), aAd = smem_u32(sAd),
                     aDS = smem_u32(sDS);
      uint32_t step = 0;
      for (int j = 0; j < nblk; ++j) {
        // (the last MMAs that read sK / sV were committed to bar_o of the previous step and waited for below)
        mbar_expect_tx(bar_kv, 2 * AT_TILE_BYTES);
        tma_load_2d(sK, &tmQKV, bar_kv, E + h * AT_D, row0 + j * AT_TILE);
        tma_load_2d(sV, &tmQKV, bar_kv, 2 * E + h * AT_D, row0 + j * AT_TILE);
        for (int i = j; i < nblk; ++i, ++step) {
          mbar_expect_tx(bar_qd, 2 * AT_TILE_BYTES);
          tma_load_2d(sQ, &tmQKV, bar_qd, h * AT_D, row0 + i * AT_TILE);
          tma_load_2d(sDO, &tmDO, bar_qd, h * AT_D, row0 + i * AT_TILE);
          if (i == j) mbar_wait(bar_kv, j & 1);
          mbar_wait(bar_qd, step & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_S, desc_k(aQ, k), desc_k(aK, k), idesc_s, k != 0);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_dP, desc_k(aDO, k), desc_k(aV, k), idesc_s, k != 0);
          umma_commit(bar_s);
          mbar_wait(bar_p, step & 1);
          tc_fence_after();
          const uint32_t acc = i != j;
#pragma unroll
          for (int k = 0; k < 8; ++k) {                       // reduction over the 128 query rows, 16 per instruction
            umma_bf16(tmem_dV, desc_mn(aAd, k, AT_TILE_BYTES), desc_mn(aDO, k, 16), idesc_kv, acc | (k != 0));
            umma_bf16(tmem_dK, desc_mn(aDS, k, AT_TILE_BYTES), desc_mn(aQ, k, 16), idesc_kv, acc | (k != 0));
          }
#pragma unroll
          for (int k = 0; k < 8; ++k)                         // reduction over the 128 key rows
            umma_bf16(tmem_dQ, desc_k(aDS + (k >> 2) * AT_TILE_BYTES, k & 3), desc_mn(aK, k, 16), idesc_dq, k != 0);
          umma_commit(bar_o);
          mbar_wait(bar_o, step & 1);                         // sQ / sDO / sK / sV are free again
        }
      }
    }
    __syncwarp();
  } else {
    const int r = threadIdx.x;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    const uint32_t aAd = smem_u32(sAd), aDS = smem_u32(sDS);
    uint32_t step = 0;
    for (int j = 0; j < nblk; ++j) {
      const int kv0 = j * AT_TILE;
      for (int i = j; i < nblk; ++i, ++step) {
        const int q_abs = i * AT_TILE + r;
        const bool q_ok = q_abs < T;
        const float L = q_ok ? __ldg(p.lse + (size_t)bh * T + q_abs) : 0.f;
        // delta = sum_d dO[q][d] * O[q][d]  (= row sum of dP * P after dropout); overlaps the TMA / MMA latency
        float delta = 0.f;
        if (q_ok) {
          const uint4* po = reinterpret_cast<const uint4*>(p.o + (size_t)(row0 + q_abs) * E + h * AT_D);
          const uint4* pd = reinterpret_cast<const uint4*>(p.dout + (size_t)(row0 + q_abs) * E + h * AT_D);
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const uint4 a = __ldg(po + g), d = __ldg(pd + g);
            const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
            const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&d);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 fa = __bfloat1622float2(ha[t]), fd = __bfloat1622float2(hd[t]);
              delta = fmaf(fa.x, fd.x, fmaf(fa.y, fd.y, delta));
            }
          }
        }
        const uint32_t rowkey = p.drop_thresh ? at_rowkey(p, bh, q_abs) : 0u;
        mbar_wait(bar_s, step & 1);
        tc_fence_after();
        const bool diag = i == j;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          float s[32], dp[32];
          tmem_ld_32x32(tmem_S + lane_base + c * 32, s);
          tmem_ld_32x32(tmem_dP + lane_base + c * 32, dp);
#pragma unroll
          for (int t = 0; t < 32; ++t) {
            const bool valid = q_ok && (!diag || (c * 32 + t <= r));
            const float pr = valid ? exp2f(s[t] * p.scale_log2 - L) : 0.f;
            float keep = 1.f;
            if (p.drop_thresh) keep = (at_hash(rowkey + (uint32_t)(kv0 + c * 32 + t)) >= p.drop_thresh) ? p.inv_keep : 0.f;
            s[t] = pr * keep;                                          // dropped probability (operand of dV)
            dp[t] = pr * (dp[t] * keep - delta) * p.scale;             // dS (operand of dK, dQ), scale folded in
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t off = p_chunk_off(r, c * 4 + g);
            st_shared_v4(aAd + off, pack_bf16(s[g * 8], s[g * 8 + 1]), pack_bf16(s[g * 8 + 2], s[g * 8 + 3]),
                         pack_bf16(s[g * 8 + 4], s[g * 8 + 5]), pack_bf16(s[g * 8 + 6], s[g * 8 + 7]));
            st_shared_v4(aDS + off, pack_bf16(dp[g * 8], dp[g * 8 + 1]), pack_bf16(dp[g * 8 + 2], dp[g * 8 + 3]),
                         pack_bf16(dp[g * 8 + 4], dp[g * 8 + 5]), pack_bf16(dp[g * 8 + 6], dp[g * 8 + 7]));
          }
        }
        fence_async_smem();
        tc_fence_before();
        mbar_arrive(bar_p);
        mbar_wait(bar_o, step & 1);
        tc_fence_after();
        // dQ tile of query block i: the last key block that touches it is the diagonal one (j == i)
        {
          float* acc_row = (p.dq_acc && q_ok) ? p.dq_acc + (size_t)(row0 + q_abs) * E + h * AT_D : nullptr;
          __nv_bfloat16* dst = p.out + (size_t)(row0 + q_abs) * 3 * E + h * AT_D;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float v[32];
            tmem_ld_32x32(tmem_dQ + lane_base + c * 32, v);
            if (!q_ok) continue;
            if (j > 0) {
#pragma unroll
              for (int t = 0; t < 32; t += 4) {
                const float4 a = *reinterpret_cast<const float4*>(acc_row + c * 32 + t);
                v[t] += a.x; v[t + 1] += a.y; v[t + 2] += a.z; v[t + 3] += a.w;
              }
            }
            if (diag) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                uint4 pk;
                pk.x = pack_bf16(v[g * 8], v[g * 8 + 1]); pk.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
                pk.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]); pk.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
                *reinterpret_cast<uint4*>(dst + c * 32 + g * 8) = pk;
              }
            } else {
#pragma unroll
              for (int t = 0; t < 32; t += 4)
                *reinterpret_cast<float4*>(acc_row + c * 32 + t) = make_float4(v[t], v[t + 1], v[t + 2], v[t + 3]);
            }
          }
        }
        tc_fence_before();
      }
      // dK, dV of key block j are complete: lane == key row
      {
        const int kv_abs = kv0 + r;
        __nv_bfloat16* dk = p.out + (size_t)(row0 + kv_abs) * 3 * E + E + h * AT_D;
        __nv_bfloat16* dv = dk + E;
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float vk[32], vv[32];
          tmem_ld_32x32(tmem_dK + lane_base + c * 32, vk);
          tmem_ld_32x32(tmem_dV + lane_base + c * 32, vv);
          if (kv_abs < T) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 pk;
              pk.x = pack_bf16(vk[g * 8], vk[g * 8 + 1]); pk.y = pack_bf16(vk[g * 8 + 2], vk[g * 8 + 3]);
              pk.z = pack_bf16(vk[g * 8 + 4], vk[g * 8 + 5]); pk.w = pack_bf16(vk[g * 8 + 6], vk[g * 8 + 7]);
              *reinterpret_cast<uint4*>(dk + c * 32 + g * 8) = pk;
              pk.x = pack_bf16(vv[g * 8], vv[g * 8 + 1]); pk.y = pack_bf16(vv[g * 8 + 2], vv[g * 8 + 3]);
              pk.z = pack_bf16(vv[g * 8 + 4], vv[g * 8 + 5]); pk.w = pack_bf16(vv[g * 8 + 6], vv[g * 8 + 7]);
              *reinterpret_cast<uint4*>(dv + c * 32 + g * 8) = pk;
            }
          }
        }
        tc_fence_before();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ host
constexpr int kFwdSmem = 4 * AT_TILE_BYTES + 1024 + 128;
constexpr int kBwdSmemMulti = 4 * AT_TILE_BYTES + 2 * AT_P_BYTES + 1024 + 128;
// no alignment slack: two CTAs (2 x (112 KB + 128 B) + 2 x 1 KB reserved) must fit the SM's 228 KB; the dynamic
// window of a kernel without static shared memory starts 1 KB-aligned (checked in the kernel)
constexpr int kBwdSmemSingle = 5 * AT_TILE_BYTES + AT_P_BYTES + 128;

static int attn_check(const char* who, int B, int T, int H, int E) {
  CEB_REQUIRE(B > 0 && T > 0 && H > 0 && E == H * AT_D, "%s: need E == H * 64 (B=%d T=%d H=%d E=%d)", who, B, T, H, E);
  CEB_REQUIRE(T <= 1024, "%s: sequences up to 1024 tokens (T=%d)", who, T);
  CEB_REQUIRE((int64_t)B * H < (1ll << 31) && (int64_t)B * H * T < (1ll << 31), "%s: batch too large", who);
  return 0;
}

static void attn_fill(AttnParams& p, int B, int T, int H, int E, float p_drop, const uint64_t* seed_dev, uint32_t salt) {
  p.B = B; p.T = T; p.H = H; p.E = E;
  p.scale = 0.125f;
  p.scale_log2 = 0.125f * 1.4426950408889634f;
  p.drop_thresh = ceb_drop_thresh(p_drop);
  p.inv_keep = p.drop_thresh ? 1.f / (1.f - p_drop) : 1.f;
  p.seed_dev = seed_dev; p.salt = salt;
}

// out[B*T][E] (bf16) = causal attention over qkv[B*T][3E] (bf16); lse[B*H][T] (fp32) is kept for the backward.
// p_drop: dropout probability on the attention weights (0 in eval); the mask is a pure function of
// (*seed_dev, salt, batch, head, query, key), so the backward regenerates it.
CEB_API int ceb_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int E, float p_drop,
                              const uint64_t* seed_dev, uint32_t salt, void* stream) {
  CEB_REQUIRE(qkv && out && lse, "ceb_attention_fwd: null pointer");
  if (int e = attn_check("ceb_attention_fwd", B, T, H, E)) return e;
  CEB_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "ceb_attention_fwd: p_drop=%f", p_drop);
  CEB_REQUIRE((((uintptr_t)qkv | (uintptr_t)out) & 15) == 0, "ceb_attention_fwd: tensors must be 16-byte aligned");
  CUtensorMap tm;
  uint64_t dims[2] = {(uint64_t)3 * E, (uint64_t)B * T}, str[1] = {(uint64_t)3 * E * 2};
  uint32_t box[2] = {AT_D, AT_TILE};
  if (int e = ceb_make_tmap(&tm, 2, qkv, 2, dims, str, box, nullptr)) return e;
  AttnParams p{};
  attn_fill(p, B, T, H, E, p_drop, seed_dev, salt);
  p.out = (__nv_bfloat16*)out; p.lse = lse;
  static bool configured = false;
  if (!configured) {
    CEB_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
    configured = true;
  }
  dim3 grid((unsigned)(B * H), (unsigned)ceb_cdiv(T, AT_TILE));
  attn_fwd_kernel<<<grid, AT_THREADS, kFwdSmem, (cudaStream_t)stream>>>(tm, p);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// dqkv[B*T][3E] (bf16) = gradient of the attention above w.r.t. qkv, given dout[B*T][E], the forward output
// out[B*T][E] and lse.  dq_scratch: fp32 [B*T][E], required only when T > 128 (may be NULL otherwise).
CEB_API int ceb_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                              float* dq_scratch, int B, int T, int H, int E, float p_drop, const uint64_t* seed_dev,
                              uint32_t salt, void* stream) {
  CEB_REQUIRE(qkv && out && dout && lse && dqkv, "ceb_attention_bwd: null pointer");
  if (int e = attn_check("ceb_attention_bwd", B, T, H, E)) return e;
  CEB_REQUIRE(T <= AT_TILE || dq_scratch != nullptr, "ceb_attention_bwd: T=%d > 128 needs dq_scratch", T);
  CEB_REQUIRE((((uintptr_t)qkv | (uintptr_t)out | (uintptr_t)dout | (uintptr_t)dqkv | (uintptr_t)dq_scratch) & 15) == 0,
              "ceb_attention_bwd: tensors must be 16-byte aligned");
  CUtensorMap tm, tmdo;
  uint64_t dims[2] = {(uint64_t)3 * E, (uint64_t)B * T}, str[1] = {(uint64_t)3 * E * 2};
  uint32_t box[2] = {AT_D, AT_TILE};
  if (int e = ceb_make_tmap(&tm, 2, qkv, 2, dims, str, box, nullptr)) return e;
  uint64_t dims2[2] = {(uint64_t)E, (uint64_t)B * T}, str2[1] = {(uint64_t)E * 2};
  if (int e = ceb_make_tmap(&tmdo, 2, dout, 2, dims2, str2, box, nullptr)) return e;
  AttnParams p{};
  attn_fill(p, B, T, H, E, p_drop, seed_dev, salt);
  p.out = (__nv_bfloat16*)dqkv; p.o = (const __nv_bfloat16*)out; p.dout = (const __nv_bfloat16*)dout;
  p.lse = (float*)lse; p.dq_acc = T > AT_TILE ? dq_scratch : nullptr;
  static bool configured = false;
  if (!configured) {
    CEB_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmemMulti));
    CEB_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmemSingle));
    configured = true;
  }
  if (T > AT_TILE) attn_bwd_kernel<true><<<(unsigned)(B * H), AT_THREADS, kBwdSmemMulti, (cudaStream_t)stream>>>(tm, tmdo, p);
  else attn_bwd_kernel<false><<<(unsigned)(B * H), AT_THREADS, kBwdSmemSingle, (cudaStream_t)stream>>>(tm, tmdo, p);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}
