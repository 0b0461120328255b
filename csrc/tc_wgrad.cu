// Weight gradient of a 3x3 / stride 1 / pad 1 convolution on tcgen05 tensor cores (sm_100a).
//
//   dW[co][r][s][ci] += sum over pixels p of dZ[p][co] * X[p + (r-1, s-1)][ci]      (NHWC, KRSC)
//
// This is a GEMM whose reduction dimension is the pixel index, i.e. BOTH operands are
// "MN-major" (channels contiguous, pixels strided): A = dZ^T (M = co), B = shifted X (N = ci).
// tcgen05.mma reads MN-major operands directly from 128-byte-swizzled shared memory (instruction
// descriptor bits 15/16), so no transposed copy of the activations is ever made.  One CTA owns a
// (co tile, kernel row r, ci tile) triple and a slice of the pixels (split-K): per 64-pixel step it
// loads ONE dZ tile and the THREE horizontally shifted X tiles (s = 0,1,2; TMA zero fill = padding)
// and accumulates three 128 x BN fp32 tiles in TMEM.  The epilogue adds them into the fp32 flat
// gradient buffer (KRSC, exactly the layout of utils/flat.py "nhwc") with 16-byte vector atomics.
// Replaces cuDNN's wgrad behind autograd of nn.Conv2d
// (/root/reference/CommEfficient/models/resnet9.py:36-37, fed_worker.py:286 loss.backward()).
#include "tc_common.cuh"
#include <cstdlib>

using namespace tc;

constexpr int WG_M = 128;          // output channels per tile
constexpr int WG_THREADS = 192;
// TF = 0: bf16 operands, 64 pixels per k-step, 64-channel (128-byte) MN chunks, 16 pixels per MMA.
// TF = 1: fp32 operands read as TF32 (kind::tf32): 32-channel chunks and 8 pixels per MMA; the k-step
//         is halved to 32 pixels so a stage keeps its size in bytes.
template <int TF> struct WgGeom {
  static constexpr int kPix = TF ? 32 : 64;          // pixels per k-step
  static constexpr int kChE = TF ? 32 : 64;          // channels per 128-byte chunk row
  static constexpr int kUk = TF ? 8 : 16;            // pixels per MMA instruction
  static constexpr int kEsz = TF ? 4 : 2;
};

struct WgParams {
  float* dW;             // [Cout][9*Cin] fp32, accumulated
  int Cout, Cin;
  int H, W;
  int k_steps_total;     // ceil(Npix / 64)
  int k_per_cta;         // k-steps handled by one CTA
  int ci_tiles;
  int tile_h, tile_n;    // box geometry of a 64-pixel step (tile_w == W)
  int ldw;               // row stride of dW (9*Cin for a convolution, ldc for a plain GEMM)
  int vec_ok;            // dW is 16-byte aligned (it is a slice of the flat gradient buffer): vector atomics allowed
  const int* k_limit;    // optional device-side count of reduction rows (see ceb_gemm_bf16_mn_acc_rows)
  // general convolution geometry: H, W above are the OUTPUT map; a k-step covers kpix (<= the template's k-step)
  // output pixels -- whole rows of the output map -- and transfers stage_bytes; the rows of a stage beyond kpix
  // are never written by TMA and are zeroed once at kernel start (zero pixels add nothing to the reduction)
  int stride, pad, kpix, stage_bytes;
};

// Shared-memory ring.  TAPS weight-tap tiles per stage (3 for a convolution row, 1 for a plain GEMM);
// a CTA of a pair (PAIR = 2) stages only its half of every tap's BN input channels.
template <int BN, int TAPS, int PAIR, int TF = 0>
struct WgSmem {
  using G = WgGeom<TF>;
  static constexpr int kChunk = G::kPix * 128;                    // [pixels][one 128-byte channel row]
  static constexpr int kAChunks = WG_M / G::kChE;
  static constexpr int kABytes = kAChunks * kChunk;               // this CTA's 128 output channels
  static constexpr int kCh = BN / G::kChE / PAIR;                 // channel chunks per tap staged by this CTA
  static constexpr int kBChunks = TAPS * kCh;
  static constexpr int kBBytes = kBChunks * kChunk;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesFit = (200 * 1024) / kStageBytes;
  static constexpr int kStages = kStagesFit > 8 ? 8 : kStagesFit;
  static constexpr int kTotal = kStages * kStageBytes + 1024 + 256;
  // One MMA instruction covers up to 256 accumulator columns: the first kC1 chunks of each CTA go
  // into instruction 1 (N1 columns over the pair), the rest into instruction 2.
  static constexpr int kC1 = kBChunks < 256 / (G::kChE * PAIR) ? kBChunks : 256 / (G::kChE * PAIR);
  static constexpr int kC2 = kBChunks - kC1;
  static constexpr int kN1 = kC1 * G::kChE * PAIR;
  static constexpr int kN2 = kC2 * G::kChE * PAIR;
  static constexpr int kNTot = kN1 + kN2;                         // accumulator columns (= TAPS * BN)
  static constexpr int kTmemCols = kNTot <= 128 ? 128 : (kNTot <= 256 ? 256 : 512);
  static_assert(kC2 * G::kChE * PAIR <= 256 && kNTot <= 512, "tile does not fit two MMA instructions / TMEM");
  static_assert(BN % (G::kChE * PAIR) == 0, "every CTA needs whole channel chunks");
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// CONV = false turns the kernel into the plain "both operands MN-major" GEMM
//   C[M][N] (fp32) += A[K][M]^T * B[K][N]        (A, B bf16 row-major, K = reduction)
// used for linear-layer weight gradients (dW = dY^T X) and the im2col'd stem: one tap, 2-D B.
// PAIR = 2: two CTAs of a 2-CTA cluster share every MMA (tcgen05.mma.cta_group::2, M = 256 output
// channels): each stages its own 128 channels of dZ and half of the BN input channels of every tap,
// so the shared-memory traffic per FLOP -- the bound of this kernel -- drops by ~40 %.
template <int BN, bool CONV, int PAIR, int TF = 0>
__global__ void __launch_bounds__(WG_THREADS, 1)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap tmapDZ, const __grid_constant__ CUtensorMap tmapX, WgParams p) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int TAPS = CONV ? 3 : 1;
  using L = WgSmem<BN, TAPS, PAIR, TF>;
  using G = WgGeom<TF>;
  constexpr int WG_KPIX = G::kPix;
  constexpr int CHE = G::kChE;
  constexpr int kStages = L::kStages;
  constexpr int CH = L::kCh;
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + kStages * L::kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = PAIR == 2 ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
  // unit = blockIdx.x / PAIR = ((co_unit * 3 + r) * ci_tiles + ci_tile), blockIdx.y = split
  const int unit = blockIdx.x / PAIR;
  const int ci_tile = unit % p.ci_tiles;
  const int r = CONV ? (unit / p.ci_tiles) % 3 : 0;
  const int co_tile = (CONV ? unit / (p.ci_tiles * 3) : unit / p.ci_tiles) * PAIR + (int)cta_rank;
  const int ks_begin = blockIdx.y * p.k_per_cta;
  int ks_total = p.k_steps_total;
  if (p.k_limit) ks_total = min(ks_total, (__ldg(p.k_limit) + p.kpix - 1) / p.kpix);
  const int ks_end = min(ks_begin + p.k_per_cta, ks_total);
  const int n_steps = ks_end - ks_begin;
  constexpr int TMEM_COLS = L::kTmemCols;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmapDZ);
    tma_prefetch_desc(&tmapX);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    mbar_init(tmem_full_bar, 1);
    mbar_fence_init();
  }
  if (p.kpix < WG_KPIX) {
    // partial k-steps: rows [kpix, k-step) of every chunk stay untouched by TMA -- make them zero
    uint4* z = reinterpret_cast<uint4*>(smem);
    for (int i = threadIdx.x; i < kStages * L::kStageBytes / 16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    if (PAIR == 2) tmem_alloc_pair<TMEM_COLS>(tmem_slot);
    else tmem_alloc<TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  if (PAIR == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (n_steps > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const int hw = p.H * p.W;
        const int ci_base = ci_tile * BN + (int)cta_rank * (BN / PAIR);   // this CTA's input channels of the tile
        for (int i = 0; i < n_steps; ++i) {
          const int s = i % kStages;
          const uint32_t phase = (i / kStages) & 1;
          mbar_wait(empty_bar + s, phase ^ 1);
          uint8_t* a_dst = smem + s * L::kStageBytes;
          uint8_t* b_dst = a_dst + L::kABytes;
          const int pix0 = (ks_begin + i) * p.kpix;
          if (PAIR == 2) {
            // both CTAs' bytes are counted on the leader's barrier, which only the leader arms
            const uint32_t fbar = mapa_u32(smem_u32(full_bar + s), 0);
            if (leader) mbar_expect_tx(full_bar + s, 2u * (uint32_t)p.stage_bytes);
#pragma unroll
            for (int ac = 0; ac < L::kAChunks; ++ac)
              tma_load_2d_pair(a_dst + ac * L::kChunk, &tmapDZ, fbar, co_tile * WG_M + ac * CHE, pix0);
            if (CONV) {
              const int n0 = pix0 / hw;
              const int h0 = (pix0 - n0 * hw) / p.W;
#pragma unroll
              for (int sft = 0; sft < 3; ++sft)
#pragma unroll
                for (int ch = 0; ch < CH; ++ch)
                  tma_load_4d_pair(b_dst + (sft * CH + ch) * L::kChunk, &tmapX, fbar, ci_base + ch * CHE, sft - p.pad,
                                   h0 * p.stride + r - p.pad, n0);
            } else {
#pragma unroll
              for (int ch = 0; ch < CH; ++ch)
                tma_load_2d_pair(b_dst + ch * L::kChunk, &tmapX, fbar, ci_base + ch * CHE, pix0);
            }
          } else {
            mbar_expect_tx(full_bar + s, (uint32_t)p.stage_bytes);
            // dZ: 2-D [pixels][Cout]: two 64-channel chunks
#pragma unroll
            for (int ac = 0; ac < L::kAChunks; ++ac)
              tma_load_2d(a_dst + ac * L::kChunk, &tmapDZ, full_bar + s, co_tile * WG_M + ac * CHE, pix0);
            // X: 4-D [N][H][W][Cin], shifted by (r-1, s-1)
            if (CONV) {
              const int n0 = pix0 / hw;
              const int h0 = (pix0 - n0 * hw) / p.W;
#pragma unroll
              for (int sft = 0; sft < 3; ++sft)
#pragma unroll
                for (int ch = 0; ch < CH; ++ch)
                  tma_load_4d(b_dst + (sft * CH + ch) * L::kChunk, &tmapX, full_bar + s, ci_base + ch * CHE, sft - p.pad,
                              h0 * p.stride + r - p.pad, n0);
            } else {
#pragma unroll
              for (int ch = 0; ch < CH; ++ch)
                tma_load_2d(b_dst + ch * L::kChunk, &tmapX, full_bar + s, ci_base + ch * CHE, pix0);
            }
          }
        }
      }
    } else if (warp == 1) {
      // The tap tiles of a stage are consecutive 64-channel MN-major chunks and their accumulators
      // consecutive TMEM columns, so they are ONE wide-N operand: issue N up to 256 per instruction
      // instead of one N = BN MMA per tap -- the dZ tile is then read from shared memory once (or
      // twice) per k-slice instead of three times, and shared-memory bandwidth bounds this kernel.
      constexpr uint32_t idesc1 = idesc_any<TF>(WG_M * PAIR, L::kN1, 1, 1);     // both operands MN-major
      constexpr uint32_t idesc2 = idesc_any<TF>(WG_M * PAIR, L::kN2 > 0 ? L::kN2 : 16, 1, 1);
      for (int i = 0; leader && i < n_steps; ++i) {
        const int s = i % kStages;
        const uint32_t phase = (i / kStages) & 1;
        mbar_wait(full_bar + s, phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + s * L::kStageBytes);
          const uint32_t b_addr = a_addr + L::kABytes;
#pragma unroll
          for (int k = 0; k < WG_KPIX / G::kUk; ++k) {
            // one instruction = kUk pixels = kUk rows of 128 B; MN chunks (one 128-byte channel row wide) are
            // L::kChunk apart
            const uint64_t da = smem_desc_mn<TF>(a_addr + k * G::kUk * 128, L::kChunk);
            const uint64_t db = smem_desc_mn<TF>(b_addr + k * G::kUk * 128, L::kChunk);
            umma_any<TF, PAIR>(tmem_base, da, db, idesc1, (i | k) != 0);
            if (L::kN2 > 0) {
              const uint64_t db2 = smem_desc_mn<TF>(b_addr + L::kC1 * L::kChunk + k * G::kUk * 128, L::kChunk);
              umma_any<TF, PAIR>(tmem_base + L::kN1, da, db2, idesc2, (i | k) != 0);
            }
          }
          if (PAIR == 2) {
            umma_commit_pair(empty_bar + s);
            if (i == n_steps - 1) umma_commit_pair(tmem_full_bar);
          } else {
            umma_commit(empty_bar + s);
            if (i == n_steps - 1) umma_commit(tmem_full_bar);
          }
        }
        __syncwarp();
      }
    } else {
      const int q = warp & 3;
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
      const int co = co_tile * WG_M + q * 32 + lane;
      const int ldw = p.ldw;
#pragma unroll 1
      for (int cc = 0; cc < L::kNTot / 32; ++cc) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + cc * 32, v);
        // accumulator column -> (source CTA of the B chunk, chunk within that CTA) -> (tap, channel)
        const int col = cc * 32;
        const bool second = col >= L::kN1;
        const int colm = second ? col - L::kN1 : col;
        const int per_cta = (second ? L::kN2 : L::kN1) / PAIR;
        const int src = colm / per_cta, local = colm - src * per_cta;
        const int g = (second ? L::kC1 : 0) + local / CHE;                 // chunk index inside the source CTA's stage
        const int sft = g / CH, ch = g - sft * CH;
        const int ci0 = ci_tile * BN + src * (BN / PAIR) + ch * CHE + (local % CHE);
        if (co < p.Cout && ci0 < p.Cin) {
          float* dst = p.dW + (int64_t)co * ldw + (CONV ? (r * 3 + sft) * p.Cin : 0) + ci0;
          if (p.vec_ok) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) red_add_v4(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) red_add_f32(dst + j, v[j]);
          }
        }
      }
      tc_fence_before();
    }
  }
  if (PAIR == 2) { tc_fence_before(); cluster_sync_all(); } else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR == 2) tmem_dealloc_pair<TMEM_COLS>(tmem_base);
    else tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// Split-K factor: minimise (number of CTA waves) x (k-steps per CTA + fixed per-CTA cost).  A grid
// of 150 CTAs on 148 SMs takes two waves; one k-step less per CTA and 147 CTAs take one.
static int pick_split(int base_ctas, int k_steps_total) {
  const int sms = ceb_num_sms();
  const int kOverheadSteps = 10;         // pipeline fill + TMEM alloc + atomic epilogue, in k-step units
  int best = 1;
  int64_t best_cost = -1;
  for (int sp = 1; sp <= 128 && sp <= k_steps_total; ++sp) {
    const int per = (int)ceb_cdiv(k_steps_total, sp);
    const int eff = (int)ceb_cdiv(k_steps_total, per);          // CTAs along K that actually get work
    const int64_t waves = ceb_cdiv((int64_t)base_ctas * eff, sms);
    const int64_t cost = waves * (per + kOverheadSteps);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = eff; }
  }
  return best;
}

template <int BN, bool CONV, int PAIR, int TF = 0>
static int launch_wgrad(const CUtensorMap& tdz, const CUtensorMap& tx, WgParams p, int co_units, int split, cudaStream_t st) {
  using L = WgSmem<BN, CONV ? 3 : 1, PAIR, TF>;
  static_assert(L::kTotal <= 232448, "wgrad stage ring exceeds the 227 KB shared-memory limit");
  static bool configured = false;
  if (!configured) {
    CEB_CHECK_CUDA(cudaFuncSetAttribute(tc_wgrad_kernel<BN, CONV, PAIR, TF>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
    configured = true;
  }
  if (p.kpix == 0) { p.kpix = WgGeom<TF>::kPix; p.stride = 1; p.pad = 1; }
  if (p.stage_bytes == 0) p.stage_bytes = L::kStageBytes;
  dim3 grid((unsigned)(PAIR * co_units * (CONV ? 3 : 1) * p.ci_tiles), (unsigned)split, 1);
  if (PAIR == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(WG_THREADS, 1, 1);
    cfg.dynamicSmemBytes = L::kTotal;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, tc_wgrad_kernel<BN, CONV, PAIR, TF>, tdz, tx, p));
    CEB_COUNT_LAUNCH(1);
    return 0;
  }
  tc_wgrad_kernel<BN, CONV, PAIR, TF><<<grid, WG_THREADS, L::kTotal, st>>>(tdz, tx, p);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// CEB_WG_PAIR=0 disables / =1 forces the CTA-pair weight-gradient kernels where their shape rules allow.
static int wg_pair_env() {
  static const int v = [] { const char* e = getenv("CEB_WG_PAIR"); return e ? atoi(e) : -1; }();
  return v;
}

// dW[Cout][3][3][Cin] (fp32) += wgrad(dz[N,H,W,Cout], x[N,H,W,Cin]); operands bf16 (tf = 0) or fp32 read as
// TF32 (tf = 1).  Cin, Cout multiples of 64, W in {4,8,16,32(,64 bf16)}, H*W a multiple of the k-step or a
// divisor of it.
static int conv_wgrad_impl(const void* dz, const void* x, float* dW, int Nimg, int H, int W, int Cin, int Cout,
                           void* stream, int tf, int stride = 1) {
  const int esz = tf ? 4 : 2, che = 128 / esz, kmax = tf ? 32 : 64;
  CEB_REQUIRE(dz && x && dW && Nimg > 0, "ceb_conv3x3_wgrad_nhwc: bad arguments");
  CEB_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0, "ceb_conv3x3_wgrad_nhwc: Cin, Cout must be multiples of 64 (Cin=%d Cout=%d)", Cin, Cout);
  CEB_REQUIRE(stride == 1 || stride == 2, "ceb_conv3x3_wgrad_nhwc: stride %d", stride);
  // H, W: INPUT map; the reduction runs over the pixels of the OUTPUT map (3x3, pad 1)
  const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
  CEB_REQUIRE(Wo >= 1 && Wo <= kmax, "ceb_conv3x3_wgrad_nhwc: output width %d exceeds the %d-pixel k-step", Wo, kmax);
  // a k-step = whole rows of the output map (or whole images)
  int tile_h = 1, tile_n = 1;
  for (int th = 1; th <= Ho; ++th)
    if (Ho % th == 0 && th * Wo <= kmax) tile_h = th;
  if (tile_h == Ho) tile_n = kmax / (Ho * Wo);
  if (tile_n < 1) tile_n = 1;
  if (tile_n > Nimg) tile_n = Nimg;
  const int kpix = tile_h * Wo * tile_n;
  const int64_t npix = (int64_t)Nimg * Ho * Wo;
  CUtensorMap tdz, tx;
  uint64_t dd[2] = {(uint64_t)Cout, (uint64_t)npix}, sd[1] = {(uint64_t)Cout * esz};
  uint32_t bd[2] = {(uint32_t)che, (uint32_t)kpix};
  if (int e = ceb_make_tmap(&tdz, esz, dz, 2, dd, sd, bd, nullptr, tf != 0)) return e;
  uint64_t dx[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)Nimg};
  uint64_t sx[3] = {(uint64_t)Cin * esz, (uint64_t)W * Cin * esz, (uint64_t)H * W * Cin * esz};
  uint32_t bx[4] = {(uint32_t)che, (uint32_t)(Wo * stride), (uint32_t)(tile_h * stride), (uint32_t)tile_n};
  uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
  if (int e = ceb_make_tmap(&tx, esz, x, 4, dx, sx, bx, es, tf != 0)) return e;

  const int bn = (Cin % 128 == 0) ? 128 : 64;
  // CTA pairs: 256 output channels per pair, each CTA stages 64 of the tile's 128 input channels
  const bool pair = bn == 128 && Cout % 256 == 0 && (wg_pair_env() >= 0 ? wg_pair_env() != 0 : true);
  WgParams p{};
  p.dW = dW; p.Cout = Cout; p.Cin = Cin; p.H = Ho; p.W = Wo;
  p.k_steps_total = (int)ceb_cdiv(npix, kpix);
  p.ci_tiles = Cin / bn;
  p.tile_h = tile_h; p.tile_n = tile_n; p.ldw = 9 * Cin;
  p.vec_ok = (((uintptr_t)dW & 15) == 0) ? 1 : 0;
  p.stride = stride; p.pad = 1; p.kpix = kpix;
  {
    const int a_chunks = WG_M / che, b_chunks = 3 * (bn / che) / (pair ? 2 : 1);
    p.stage_bytes = (a_chunks + b_chunks) * kpix * 128;
  }
  const int co_units = (int)ceb_cdiv(Cout, WG_M * (pair ? 2 : 1));
  const int base_ctas = (pair ? 2 : 1) * co_units * 3 * p.ci_tiles;
  int split = pick_split(base_ctas, p.k_steps_total);
  p.k_per_cta = (int)ceb_cdiv(p.k_steps_total, split);
  split = (int)ceb_cdiv(p.k_steps_total, p.k_per_cta);
  cudaStream_t st = (cudaStream_t)stream;
  if (tf) {
    if (pair) return launch_wgrad<128, true, 2, 1>(tdz, tx, p, co_units, split, st);
    if (bn == 128) return launch_wgrad<128, true, 1, 1>(tdz, tx, p, co_units, split, st);
    return launch_wgrad<64, true, 1, 1>(tdz, tx, p, co_units, split, st);
  }
  if (pair) return launch_wgrad<128, true, 2>(tdz, tx, p, co_units, split, st);
  if (bn == 128) return launch_wgrad<128, true, 1>(tdz, tx, p, co_units, split, st);
  return launch_wgrad<64, true, 1>(tdz, tx, p, co_units, split, st);
}

// General form: 3x3 / pad 1 convolution with stride 1 or 2 over an INPUT map H x W (any size whose output
// width fits a k-step: <= 64 pixels bf16, <= 32 TF32); dz is [N, Ho, Wo, Cout].  f32 = 1: fp32 operands (TF32).
CEB_API int ceb_conv3x3_wgrad_nhwc(const void* dz, const void* x, float* dW, int Nimg, int H, int W, int Cin, int Cout,
                                   int stride, int f32, void* stream) {
  return conv_wgrad_impl(dz, x, dW, Nimg, H, W, Cin, Cout, stream, f32, stride);
}

CEB_API int ceb_conv3x3_wgrad_bf16_nhwc(const void* dz, const void* x, float* dW, int Nimg, int H, int W, int Cin,
                                        int Cout, void* stream) {
  return conv_wgrad_impl(dz, x, dW, Nimg, H, W, Cin, Cout, stream, 0);
}

CEB_API int ceb_conv3x3_wgrad_tf32_nhwc(const float* dz, const float* x, float* dW, int Nimg, int H, int W, int Cin,
                                        int Cout, void* stream) {
  return conv_wgrad_impl(dz, x, dW, Nimg, H, W, Cin, Cout, stream, 1);
}

// C[M][N] (fp32, row stride ldc) += A[K][M]^T * B[K][N]; A (row stride lda), B (row stride ldb) bf16 (or fp32
// read as TF32) with the reduction running over ROWS.  N % 64 == 0; M is arbitrary (TMA zero-fills the tail
// of the last 128-row tile and the epilogue bounds its rows); lda, ldb multiples of 16 bytes.
static int mn_acc_impl(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int M, int N, int K, int ldc,
                       const int* k_limit, void* stream, int tf = 0);

CEB_API int ceb_gemm_bf16_mn_acc_ld(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int M, int N,
                                    int K, int ldc, void* stream) {
  return mn_acc_impl(A, lda, B, ldb, C, M, N, K, ldc, nullptr, stream);
}

CEB_API int ceb_gemm_tf32_mn_acc_ld(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int M, int N,
                                    int K, int ldc, void* stream) {
  return mn_acc_impl(A, lda, B, ldb, C, M, N, K, ldc, nullptr, stream, 1);
}

// Same, reducing over rows [0, *k_limit_dev) only (rounded up to whole 64-row steps, never past K:
// the rows between the limit and the end of its step must hold zeros or finite values the caller
// wants included).  k_limit_dev is a device int read by the kernel (CUDA-graph friendly).
CEB_API int ceb_gemm_bf16_mn_acc_rows(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int M, int N,
                                      int K, int ldc, const int* k_limit_dev, void* stream) {
  CEB_REQUIRE(k_limit_dev != nullptr, "ceb_gemm_bf16_mn_acc_rows: k_limit_dev is NULL");
  return mn_acc_impl(A, lda, B, ldb, C, M, N, K, ldc, k_limit_dev, stream);
}

static int mn_acc_impl(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int M, int N, int K, int ldc,
                       const int* k_limit, void* stream, int tf) {
  const int esz = tf ? 4 : 2, che = 128 / esz, kpix = tf ? 32 : 64, al = 16 / esz;
  CEB_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "ceb_gemm_mn_acc_ld: bad arguments");
  CEB_REQUIRE(N % che == 0, "ceb_gemm_mn_acc_ld: N must be a multiple of %d (N=%d)", che, N);
  CEB_REQUIRE(lda % al == 0 && ldb % al == 0 && lda >= M && ldb >= N && ldc >= N,
              "ceb_gemm_mn_acc_ld: need lda, ldb multiples of %d with lda >= M, ldb >= N, ldc >= N "
              "(lda=%lld ldb=%lld ldc=%d)", al, (long long)lda, (long long)ldb, ldc);
  CEB_REQUIRE((((uintptr_t)A | (uintptr_t)B) & 15) == 0, "ceb_gemm_mn_acc_ld: operands must be 16-byte aligned");
  CUtensorMap ta, tb;
  uint64_t da[2] = {(uint64_t)M, (uint64_t)K}, sa[1] = {(uint64_t)lda * esz};
  uint32_t ba[2] = {(uint32_t)che, (uint32_t)kpix};
  if (int e = ceb_make_tmap(&ta, esz, A, 2, da, sa, ba, nullptr, tf != 0)) return e;
  uint64_t db[2] = {(uint64_t)N, (uint64_t)K}, sb[1] = {(uint64_t)ldb * esz};
  if (int e = ceb_make_tmap(&tb, esz, B, 2, db, sb, ba, nullptr, tf != 0)) return e;
  // CTA pairs: 256 x 256 output tile per pair (each CTA: 128 rows of A, 128 columns of B)
  const bool pair = N % 256 == 0 && M >= 256 && (wg_pair_env() >= 0 ? wg_pair_env() != 0 : true);
  const int bn = pair ? 256 : ((N % 128 == 0) ? 128 : 64);
  WgParams p{};
  p.dW = C; p.Cout = M; p.Cin = N; p.H = 1; p.W = 1; p.ldw = ldc; p.k_limit = k_limit;
  p.vec_ok = (((uintptr_t)C & 15) == 0 && ldc % 4 == 0) ? 1 : 0;
  p.k_steps_total = (int)ceb_cdiv(K, kpix);
  p.ci_tiles = (int)ceb_cdiv(N, bn);                  // (a last partial tile is zero-filled by TMA and bounded by the epilogue)
  const int co_units = (int)ceb_cdiv(M, WG_M * (pair ? 2 : 1));
  const int base_ctas = (pair ? 2 : 1) * co_units * p.ci_tiles;
  int split = pick_split(base_ctas, p.k_steps_total);
  p.k_per_cta = (int)ceb_cdiv(p.k_steps_total, split);
  split = (int)ceb_cdiv(p.k_steps_total, p.k_per_cta);
  cudaStream_t st = (cudaStream_t)stream;
  if (tf) {
    if (pair) return launch_wgrad<256, false, 2, 1>(ta, tb, p, co_units, split, st);
    if (bn == 128) return launch_wgrad<128, false, 1, 1>(ta, tb, p, co_units, split, st);
    return launch_wgrad<64, false, 1, 1>(ta, tb, p, co_units, split, st);
  }
  if (pair) return launch_wgrad<256, false, 2>(ta, tb, p, co_units, split, st);
  if (bn == 128) return launch_wgrad<128, false, 1>(ta, tb, p, co_units, split, st);
  return launch_wgrad<64, false, 1>(ta, tb, p, co_units, split, st);
}

// Dense form (lda = M, ldb = N): M % 64 == N % 64 == 0.
CEB_API int ceb_gemm_bf16_mn_acc(const void* A, const void* B, float* C, int M, int N, int K, int ldc, void* stream) {
  CEB_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "ceb_gemm_bf16_mn_acc: bad arguments");
  CEB_REQUIRE(M % 64 == 0 && N % 64 == 0, "ceb_gemm_bf16_mn_acc: M, N must be multiples of 64 (M=%d N=%d)", M, N);
  return ceb_gemm_bf16_mn_acc_ld(A, M, B, N, C, M, N, K, ldc, stream);
}
