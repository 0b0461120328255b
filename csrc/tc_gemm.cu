// tcgen05 / TMEM / TMA tile kernel (sm_100a): one 128 x BN output tile per CTA,
//   D[128 x BN] (fp32, TMEM) = sum over k-steps of A_tile[128 x 64] * B_tile[BN x 64]^T   (bf16)
// A and B tiles are staged by TMA (128-byte swizzle) through a 4-deep mbarrier ring; one elected
// thread issues tcgen05.mma; four epilogue warps read the accumulator with tcgen05.ld and fuse
// bias / ReLU / dtype conversion into the store.  The same kernel is
//   * a plain "TN" GEMM  C[M,N] = A[M,K] * B[N,K]^T      (linear layers; A tile = 2-D TMA box), and
//   * a 3x3 / stride 1 / pad 1 implicit-GEMM convolution over NHWC activations: the A tile of
//     k-step (r, s, c0) is a 4-D TMA box (64 channels x tile_w x tile_h x tile_n) fetched at the
//     shifted coordinate (c0, s-1, h0+r-1, n0); TMA zero-fills out-of-bounds elements, which IS
//     the padding.  Weights are K-major "KRSC" ([Cout][3][3][Cin]) -- exactly how the flat
//     parameter buffer stores them (utils/flat.py, layout="nhwc").
// Replaces the cuDNN / cuBLAS calls behind the reference's nn.Conv2d / nn.Linear layers
// (/root/reference/CommEfficient/models/resnet9.py:36-37, fixup_resnet18.py:144-154).
#include "tc_common.cuh"
#include <cstdlib>

using namespace tc;

constexpr int BM = 128;
constexpr int BK = 64;          // bf16 elements per k-step = one 128-byte swizzle row
// 3 stages of (16 KB + BN*128 B): two CTAs fit in one SM's 227 KB for BN <= 128, so one CTA's
// epilogue overlaps the other's main loop without a persistent scheduler.
// Operand ring depth: as many 16 KB + BN*128 B stages as fit beside the epilogue staging area
// (~200 KB per SM, one persistent CTA per SM).
template <int BN> struct StagesFor { static constexpr int value = BN <= 64 ? 8 : (BN <= 128 ? 6 : 4); };
constexpr int kEpiWarps = 8;     // two per TMEM lane quarter, each draining half of the tile's columns
constexpr int kTcThreads = 64 + 32 * kEpiWarps; // warp 0: TMA, warp 1: MMA + TMEM alloc, warps 2..: epilogue
constexpr int kEpiStageBytes = kEpiWarps * 32 * 17 * 4;   // per epilogue warp: [32 rows][16 words + 1 pad]

struct TcParams {
  void* C;               // output [M][ldc]
  const float* bias;     // per output column (nullable)
  const void* mask;      // bf16 [M][ldc] (nullable): output is zeroed where mask <= 0 (ReLU backward)
  int M, N, ldc;
  int k_steps;
  int relu;              // 0: none, 1: ReLU, 2: tanh-GELU with both outputs (C = pre-activation, C2 = GELU of it),
                         // 3: multiply by GELU'(mask) (mask = the saved pre-activation); 2 and 3: bf16 direct path only
  int out_fp32;          // 0: bf16 store, 1: fp32 store, 2: fp32 atomic add
  int m_tiles, n_tiles;
  // convolution geometry
  int conv;
  int H, W, cin_steps;   // OUTPUT map height / width; cin_steps = Cin / (elements per 128-byte row)
  int tile_h;            // rows of the output map covered by one M tile (tile_w == W)
  int halo_bytes;        // MODE 1: bytes of one activation halo box
  // general convolution geometry (MODE 0): filter size 1 or 3, stride, padding; an M tile covers tile_pix
  // (<= 128) output pixels -- whole rows of the output map, so maps whose width does not divide 128 (28, 14,
  // 7, 3) leave the last rows of the 128-row MMA tile unused -- and its activation box holds a_bytes
  int ks, stride, pad, tile_pix, a_bytes;
  // split-K (atomic-accumulate outputs only): the k-steps are divided into k_split ranges of k_steps each
  // (k_total in all); a work item is (split, n tile, m tile).  Fills the machine when M x N alone gives fewer
  // tiles than SMs and K is long (the LM-head data gradient: 96 tiles, 785 k-steps).
  int k_split, k_total;
  int auto_split;        // the row count is only known on the device (m_limit): choose the split there
  // optional device-side row count: only the M tiles covering rows [0, *m_limit) are computed
  // (rows are compacted by the caller; lets a CUDA-graph-captured launch follow a data-dependent
  // row count without a host round trip)
  const int* m_limit;
  // optional residual (bf16 [M][ldc], direct-store path only): C2 == NULL: C = out + resid;
  // C2 != NULL: C = out (kept for the ReLU mask of the backward pass) and C2 = out + resid
  const void* resid;
  void* C2;
};

// MODE 0: generic tile (GEMM, or convolution with one TMA box per filter tap).
// MODE 1: "halo" convolution for wide feature maps (W = 16 or 32, Cout tile <= 128): the CTA owns
//   256 output pixels (two 128-row accumulators) and, per (horizontal tap s, 64-channel chunk),
//   loads ONE activation box that is two image rows taller than the tile.  The three vertical taps
//   r = 0,1,2 then read the same shared-memory box at row offsets r*W*128 B -- a multiple of the
//   1024-byte swizzle atom, so the UMMA descriptor stays canonical -- and both accumulators share
//   the three weight tiles.  Shared-memory fill per output pixel drops ~2.2x against MODE 0
//   (which is bound by L2 -> SM operand traffic on these layers, see profiles/).
// PAIR = 2: the CTA is half of a cta_group::2 pair and stages only BN/2 rows of every B tile.
template <int BN, int MODE, int PAIR = 1>
struct SmemLayout {
  static constexpr int kStages = PAIR == 2 ? (MODE == 1 ? 3 : (BN <= 128 ? 8 : 6))
                                           : (MODE == 1 ? (BN <= 64 ? 3 : 2) : StagesFor<BN>::value);
  static constexpr int kABytes = MODE == 1 ? 40960 : BM * BK * 2;     // halo box: (256/W + 2) rows x W x 128 B <= 40 KB
  static constexpr int kBTileBytes = (BN / PAIR) * BK * 2;            // one weight tile as staged by this CTA
  static constexpr int kBBytes = (MODE == 1 ? 3 : 1) * kBTileBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTotal = kStages * kStageBytes + kEpiStageBytes + 1024 /* align */ + 256 /* barriers */;
  static constexpr int kSubTiles = MODE == 1 ? 2 : 1;
  static constexpr int kTmemCols = 2 * kSubTiles * BN;
};

// Persistent: grid = min(#tiles, #SMs).  Tiles are walked M-fastest so CTAs running at the same
// time share the B (weight) tile in L2.  The accumulator is double buffered in TMEM (2 x BN
// columns): the epilogue warps drain tile i while the MMA thread already accumulates tile i+1,
// and the TMA producer runs ahead across tile boundaries, so operand latency is hidden even
// for short K loops (3x3 conv with 64 input channels has only 9 k-steps).
// BMAJ = 1: the B operand is "MN-major" in memory (B[K][N] row-major, e.g. GPT-2's Conv1D weights
// [in][out] or W itself in a data-gradient GEMM dX = dY * W): it is staged as 64-column chunks
// [64 k][64 n] and read by tcgen05.mma through an MN-major descriptor -- no transposed copy.
// PAIR = 2 (launched as 2-CTA clusters): two CTAs on the SMs of one TPC share every MMA --
// tcgen05.mma.cta_group::2 with M = 256, issued by the leader (cluster rank 0) only.  Each CTA stages
// its own 128 (or, MODE 1, 256) rows of A and HALF of each weight tile, so the shared-memory fill per
// FLOP drops by a third to a half against two independent CTAs, which is what bounds the short-K
// layers.  TMA completion bytes of both CTAs are counted on the leader's "full" barrier; the MMA
// commits are multicast to the "empty" / "accumulator full" barriers of both; the peer's epilogue
// warps release the accumulator with a remote arrive on the leader's barrier.
// TF = 1: fp32 operands read by the tensor cores as TF32 (tcgen05.mma.kind::tf32), fp32 output / mask /
// residual -- the matched-precision path (the reference's fp32 convolutions run TF32 inside cuDNN).  In
// BYTES nothing changes (a k-step is still one 128-byte swizzle row per operand row, an instruction still
// consumes 32 bytes of K), so the ring, the descriptors' advance and the TMEM layout are shared; only the
// element counts (32 per row, 8 per instruction), the instruction kind and the epilogue's element size differ.
template <int BN, int MODE, int BMAJ = 0, int PAIR = 1, int TF = 0>
__global__ void __launch_bounds__(kTcThreads, 1)
tc_tile_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB, TcParams p) {
  static_assert(PAIR == 1 || BMAJ == 0, "the CTA-pair variant stages K-major B only");
  constexpr int BKE = TF ? 32 : 64;                    // elements per 128-byte operand row
  constexpr int UKE = TF ? 8 : 16;                     // K elements per MMA instruction (32 bytes)
  extern __shared__ uint8_t smem_raw[];
  using L = SmemLayout<BN, MODE, PAIR>;
  constexpr int kStages = L::kStages;
  constexpr int SUBS = L::kSubTiles;
  constexpr int BMT = BM * SUBS;                       // output rows per CTA tile
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);   // SWIZZLE_128B needs 1024 B
  uint8_t* epi_smem = smem + kStages * L::kStageBytes;
  uint64_t* full_bar = (uint64_t*)(epi_smem + kEpiStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;      // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;       // [2]
  uint32_t* tmem_slot = (uint32_t*)(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = PAIR == 2 ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
  int m_tiles_rt = p.m_tiles;
  if (p.m_limit) {
    const int lim = (__ldg(p.m_limit) + p.tile_pix - 1) / p.tile_pix;
    m_tiles_rt = lim < m_tiles_rt ? lim : m_tiles_rt;
  }
  // work is handed out in units of PAIR vertically adjacent CTA tiles; a CTA of a pair owns
  // CTA tile (unit * PAIR + rank), which may lie wholly past M (TMA zero-fills, stores are bounded)
  const int m_tiles = (m_tiles_rt + PAIR - 1) / PAIR;
  const int mn_tiles = m_tiles * p.n_tiles;
  if (p.auto_split && p.m_limit) {
    // same cost model as the host's (waves x (k-steps per item + fixed cost)), evaluated with the real tile count
    const int slots = (int)gridDim.x / PAIR;
    const int kt = p.k_total;
    int best_steps = kt;
    long long best_cost = (long long)((mn_tiles + slots - 1) / slots) * (kt + 8);
    for (int sp = 2; sp <= 16 && sp * 16 <= kt; ++sp) {
      const int per = (kt + sp - 1) / sp;
      const int eff = (kt + per - 1) / per;
      const long long cost = (long long)((mn_tiles * eff + slots - 1) / slots) * (per + 8);
      if (cost < best_cost) { best_cost = cost; best_steps = per; }
    }
    p.k_steps = best_steps;
    p.k_split = (kt + best_steps - 1) / best_steps;
  }
  const int num_tiles = mn_tiles * p.k_split;
  const int first_tile = blockIdx.x / PAIR, tile_step = gridDim.x / PAIR;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmapA);
    tma_prefetch_desc(&tmapB);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tmem_full_bar + a, 1);
      mbar_init(tmem_empty_bar + a, PAIR * kEpiWarps);   // one arrive per epilogue warp (of both CTAs)
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    if (PAIR == 2) tmem_alloc_pair<L::kTmemCols>(tmem_slot);
    else tmem_alloc<L::kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  if (PAIR == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;                                   // global k-step counter (ring position)
      for (int t = first_tile; t < num_tiles; t += tile_step) {
        const int split = t / mn_tiles, tt = t - split * mn_tiles;
        const int m_tile = (tt % m_tiles) * PAIR + (int)cta_rank, n_tile = tt / m_tiles;
        const int ks0 = split * p.k_steps, ks1 = min(ks0 + p.k_steps, p.k_total);
        const int b_row0 = n_tile * BN + (int)cta_rank * (BN / PAIR);   // this CTA's share of the weight tile
        int n0 = 0, h0 = 0;
        if (p.conv) {
          const int pix0 = m_tile * p.tile_pix;
          const int hw = p.H * p.W;
          n0 = pix0 / hw;
          h0 = (pix0 - n0 * hw) / p.W;
        }
        for (int ks = ks0; ks < ks1; ++ks, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          mbar_wait(empty_bar + s, phase ^ 1);
          uint8_t* a_dst = smem + s * L::kStageBytes;
          uint8_t* b_dst = a_dst + L::kABytes;
          if (PAIR == 2) {
            // both CTAs' bytes land on the LEADER's full barrier; only the leader arms it
            const uint32_t fbar = mapa_u32(smem_u32(full_bar + s), 0);
            if (MODE == 1) {
              const int sft = ks / p.cin_steps;
              const int c0 = (ks - sft * p.cin_steps) * BKE;
              if (leader) mbar_expect_tx(full_bar + s, 2u * ((uint32_t)p.halo_bytes + L::kBBytes));
              tma_load_4d_pair(a_dst, &tmapA, fbar, c0, sft - 1, h0 - 1, n0);
#pragma unroll
              for (int r = 0; r < 3; ++r)
                tma_load_2d_pair(b_dst + r * L::kBTileBytes, &tmapB, fbar, ((r * 3 + sft) * p.cin_steps) * BKE + c0, b_row0);
            } else {
              if (leader) mbar_expect_tx(full_bar + s, 2u * ((uint32_t)p.a_bytes + L::kBBytes));
              if (p.conv) {
                const int tap = ks / p.cin_steps;
                const int c0 = (ks - tap * p.cin_steps) * BKE;
                const int r = tap / p.ks, sft = tap - r * p.ks;
                tma_load_4d_pair(a_dst, &tmapA, fbar, c0, sft - p.pad, h0 * p.stride + r - p.pad, n0);
              } else {
                tma_load_2d_pair(a_dst, &tmapA, fbar, ks * BKE, m_tile * BM);
              }
              tma_load_2d_pair(b_dst, &tmapB, fbar, ks * BKE, b_row0);
            }
          } else if (MODE == 1) {
            // k-step = (horizontal tap sft, channel chunk c0): one halo box + the 3 weight tiles r = 0..2
            const int sft = ks / p.cin_steps;
            const int c0 = (ks - sft * p.cin_steps) * BKE;
            mbar_expect_tx(full_bar + s, (uint32_t)p.halo_bytes + L::kBBytes);
            tma_load_4d(a_dst, &tmapA, full_bar + s, c0, sft - 1, h0 - 1, n0);
#pragma unroll
            for (int r = 0; r < 3; ++r)
              tma_load_2d(b_dst + r * (BN * BK * 2), &tmapB, full_bar + s,
                          ((r * 3 + sft) * p.cin_steps) * BKE + c0, n_tile * BN);
          } else {
            mbar_expect_tx(full_bar + s, (uint32_t)p.a_bytes + L::kBBytes);
            if (p.conv) {
              const int tap = ks / p.cin_steps;
              const int c0 = (ks - tap * p.cin_steps) * BKE;
              const int r = tap / p.ks, sft = tap - r * p.ks;
              tma_load_4d(a_dst, &tmapA, full_bar + s, c0, sft - p.pad, h0 * p.stride + r - p.pad, n0);
            } else {
              tma_load_2d(a_dst, &tmapA, full_bar + s, ks * BKE, m_tile * BM);
            }
            if (BMAJ == 1) {
#pragma unroll
              for (int ch = 0; ch < BN / BKE; ++ch)
                tma_load_2d(b_dst + ch * (BKE * 128), &tmapB, full_bar + s, n_tile * BN + ch * BKE, ks * BKE);
            } else {
              tma_load_2d(b_dst, &tmapB, full_bar + s, ks * BKE, n_tile * BN);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one elected thread; in a pair only the leader CTA issues) =====
    constexpr uint32_t idesc = idesc_any<TF>(BM * PAIR, BN, 0, BMAJ);
    uint32_t it = 0, local = 0;
    for (int t = first_tile; leader && t < num_tiles; t += tile_step, ++local) {
      const int acc = local & 1;
      mbar_wait(tmem_empty_bar + acc, ((local >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * (SUBS * BN);
      const int split = t / mn_tiles;
      const int ks0 = split * p.k_steps, ks1 = min(ks0 + p.k_steps, p.k_total);
      for (int ks = ks0; ks < ks1; ++ks, ++it) {
        const int s = it % kStages;
        const uint32_t phase = (it / kStages) & 1;
        mbar_wait(full_bar + s, phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + s * L::kStageBytes);
          const uint32_t b_addr = a_addr + L::kABytes;
          if (MODE == 1) {
            const uint32_t row_bytes = (uint32_t)p.W * 128u;          // one image row inside the halo box
            const uint32_t sub_rows = (uint32_t)(BM / p.W);           // image rows per 128-pixel accumulator
#pragma unroll
            for (int sub = 0; sub < 2; ++sub)
#pragma unroll
              for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const uint64_t da = smem_desc_sw128(a_addr + (sub * sub_rows + r) * row_bytes + k * 32, 16, 1024);
                  const uint64_t db = smem_desc_sw128(b_addr + r * L::kBTileBytes + k * 32, 16, 1024);
                  umma_any<TF, PAIR>(tmem_d + sub * BN, da, db, idesc, ((ks - ks0) | r | k) != 0);
                }
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t da = smem_desc_sw128(a_addr + k * 32, 16, 1024);
              const uint64_t db = BMAJ == 1 ? smem_desc_mn<TF>(b_addr + k * UKE * 128, BKE * 128)
                                            : smem_desc_sw128(b_addr + k * 32, 16, 1024);
              umma_any<TF, PAIR>(tmem_d, da, db, idesc, ((ks - ks0) | k) != 0);
            }
          }
          if (PAIR == 2) {
            umma_commit_pair(empty_bar + s);              // frees this stage in BOTH CTAs
            if (ks == ks1 - 1) umma_commit_pair(tmem_full_bar + acc);
          } else {
            umma_commit(empty_bar + s);                   // frees this smem stage when the MMAs retire
            if (ks == ks1 - 1) umma_commit(tmem_full_bar + acc);
          }
        }
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> (smem transpose) -> global =====
    const int q = warp & 3;                               // TMEM lane quarter this warp may access
    uint32_t* stage = (uint32_t*)(epi_smem + (warp - 2) * (32 * 17 * 4));
    uint32_t local = 0;
    for (int t = first_tile; t < num_tiles; t += tile_step, ++local) {
      const int split = t / mn_tiles, tt = t - split * mn_tiles;
      const int m_tile = (tt % m_tiles) * PAIR + (int)cta_rank, n_tile = tt / m_tiles;
      const int acc = local & 1;
      mbar_wait(tmem_full_bar + acc, (local >> 1) & 1);
      tc_fence_after();
      const int n_base = n_tile * BN;
      constexpr int kChunks = SUBS * (BN / 32);            // 32-column chunks per tile, split over the two warps of a quarter
      constexpr int kPer = kChunks / 2;                    // chunks drained by this warp
      const int half = (warp - 2) >> 2;
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (SUBS * BN);
      // direct path: bf16 output whose rows are 32-byte aligned -> every thread stores its row's 64
      // bytes of the chunk as two full 32-byte sectors straight from registers (no shared-memory
      // staging: the operand ring already saturates shared-memory bandwidth, see profiles/)
      const bool direct = p.out_fp32 == 0 && (p.ldc & 15) == 0 && (((uintptr_t)p.C) & 31) == 0 &&
                          (p.mask == nullptr || (((uintptr_t)p.mask) & 31) == 0);
      auto chunk_addr = [&](int i) {
        const int sc = half * kPer + i;
        const int sub = sc / (BN / 32), c = sc - sub * (BN / 32);
        return tacc + sub * BN + c * 32;
      };
      auto process = [&](const uint32_t (&cur)[32], int i) {
        const int sc = half * kPer + i;
        const int sub = sc / (BN / 32), c = sc - sub * (BN / 32);
        if (PAIR == 2 && m_tile >= m_tiles_rt) return;     // the peer's tile lies past the (device-side) row limit
        const int row0 = m_tile * p.tile_pix + sub * BM + q * 32;   // first output row of this warp
        const int m = row0 + lane;
        // rows of this tile that exist: inside the matrix and inside the tile's tile_pix valid pixels
        const int m_end_tile = m_tile * p.tile_pix + p.tile_pix;
        const int M_lim = m_end_tile < p.M ? m_end_tile : p.M;
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(cur[j]);
        const int n0 = n_base + c * 32;
        if (n0 >= p.N) return;
        if (p.bias && split == 0) {
#pragma unroll
          for (int j = 0; j < 32; ++j) if (n0 + j < p.N) v[j] += __ldg(p.bias + n0 + j);
        }
        if (p.relu == 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        const bool full = (n0 + 32 <= p.N);
        if (TF && p.out_fp32 == 1) {
          // fp32 activations: the thread's 32 columns are 128 contiguous bytes of its output row
          if (m >= M_lim) return;
          const int64_t off = (int64_t)m * p.ldc + n0;
          float* crow = (float*)p.C + off;
          const float* mrow = p.mask ? (const float*)p.mask + off : nullptr;
          const float* rrow = p.resid ? (const float*)p.resid + off : nullptr;
          float* c2row = p.C2 ? (float*)p.C2 + off : nullptr;
          const bool vec = full && (p.ldc & 7) == 0 && ((((uintptr_t)p.C) | ((uintptr_t)p.mask) | ((uintptr_t)p.resid) |
                                                          ((uintptr_t)p.C2)) & 31) == 0;
          if (vec) {
#pragma unroll
            for (int g8 = 0; g8 < 4; ++g8) {
              uint32_t w[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) w[j] = __float_as_uint(v[g8 * 8 + j]);
              if (mrow) {
                uint32_t mk[8];
                ld_global_nc_v8(mrow + g8 * 8, mk);
#pragma unroll
                for (int j = 0; j < 8; ++j) if (!(__uint_as_float(mk[j]) > 0.f)) w[j] = 0u;
              }
              if (rrow) {
                uint32_t rs[8];
                ld_global_nc_v8(rrow + g8 * 8, rs);
                if (c2row) st_global_v8(crow + g8 * 8, w);
#pragma unroll
                for (int j = 0; j < 8; ++j) w[j] = __float_as_uint(__uint_as_float(w[j]) + __uint_as_float(rs[j]));
                st_global_v8((c2row ? c2row : crow) + g8 * 8, w);
              } else {
                st_global_v8(crow + g8 * 8, w);
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (n0 + j >= p.N) break;
              float o = v[j];
              if (mrow && !(mrow[j] > 0.f)) o = 0.f;
              if (rrow) {
                if (c2row) { crow[j] = o; c2row[j] = o + rrow[j]; } else crow[j] = o + rrow[j];
              } else {
                crow[j] = o;
              }
            }
          }
          return;
        }
        if (direct && full) {
          if (m < M_lim) {
            const int64_t off = (int64_t)m * p.ldc + n0;
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
              uint32_t w[8];
              if (p.relu == 3) {
                // GELU backward fused into the data-gradient GEMM: d_pre = d_act * GELU'(pre)
                uint32_t pr[8];
                ld_global_nc_v8((const __nv_bfloat16*)p.mask + off + h2 * 16, pr);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float2 x2 = __bfloat1622float2(*(const __nv_bfloat162*)&pr[j]);
                  v[h2 * 16 + 2 * j] *= ceb_gelu_tanh_grad(x2.x);
                  v[h2 * 16 + 2 * j + 1] *= ceb_gelu_tanh_grad(x2.y);
                }
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                __nv_bfloat162 hh = __floats2bfloat162_rn(v[h2 * 16 + 2 * j], v[h2 * 16 + 2 * j + 1]);
                w[j] = *(uint32_t*)&hh;
              }
              if (p.relu == 2) {
                // GELU forward fused into the producing GEMM: the pre-activation is kept for the backward and the
                // activation is computed from its ROUNDED value (what a separate GELU kernel would read)
                st_global_v8((__nv_bfloat16*)p.C + off + h2 * 16, w);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float2 x2 = __bfloat1622float2(*(const __nv_bfloat162*)&w[j]);
                  const __nv_bfloat162 a2 = __floats2bfloat162_rn(ceb_gelu_tanh(x2.x), ceb_gelu_tanh(x2.y));
                  w[j] = *(const uint32_t*)&a2;
                }
                st_global_v8((__nv_bfloat16*)p.C2 + off + h2 * 16, w);
                continue;
              }
              if (p.mask && p.relu != 3) {
                // ReLU backward: keep a gradient only where the forward activation was positive
                // (bf16 > 0  <=>  sign bit clear and magnitude bits non-zero)
                uint32_t mk[8];
                ld_global_nc_v8((const __nv_bfloat16*)p.mask + off + h2 * 16, mk);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const uint32_t lo = ((mk[j] & 0x8000u) == 0u && (mk[j] & 0x7fffu) != 0u) ? 0x0000ffffu : 0u;
                  const uint32_t hi = ((mk[j] & 0x80000000u) == 0u && (mk[j] & 0x7fff0000u) != 0u) ? 0xffff0000u : 0u;
                  w[j] &= (lo | hi);
                }
              }
              if (p.resid) {
                uint32_t rs[8];
                ld_global_nc_v8((const __nv_bfloat16*)p.resid + off + h2 * 16, rs);
                if (p.C2) st_global_v8((__nv_bfloat16*)p.C + off + h2 * 16, w);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const __nv_bfloat162 sum = __hadd2(*(const __nv_bfloat162*)&w[j], *(const __nv_bfloat162*)&rs[j]);
                  w[j] = *(const uint32_t*)&sum;
                }
                st_global_v8((__nv_bfloat16*)(p.C2 ? p.C2 : p.C) + off + h2 * 16, w);
              } else {
                st_global_v8((__nv_bfloat16*)p.C + off + h2 * 16, w);
              }
            }
          }
          return;
        }
        if (p.mask && m < M_lim && !(p.out_fp32 == 0 && full)) {   // (the bf16 staged path masks while storing)
          const __nv_bfloat16* mk = (const __nv_bfloat16*)p.mask + (int64_t)m * p.ldc + n0;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (n0 + j < p.N && !(__bfloat162float(mk[j]) > 0.f)) v[j] = 0.f;
        }
        if (p.out_fp32 == 0 && full) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
            stage[lane * 17 + j] = *(uint32_t*)&h;
          }
          __syncwarp();
          // 4 lanes cover one 64-byte row segment; 8 rows per pass
#pragma unroll
          for (int pass = 0; pass < 4; ++pass) {
            const int r = pass * 8 + (lane >> 2);
            const int w4 = (lane & 3) * 4;
            if (row0 + r < M_lim) {
              uint4 pk;
              pk.x = stage[r * 17 + w4]; pk.y = stage[r * 17 + w4 + 1];
              pk.z = stage[r * 17 + w4 + 2]; pk.w = stage[r * 17 + w4 + 3];
              const int64_t off = (int64_t)(row0 + r) * p.ldc + n0 + w4 * 2;
              if (p.mask) {
                const uint4 mk = __ldg((const uint4*)((const __nv_bfloat16*)p.mask + off));
                auto keep = [](uint32_t m2, uint32_t val) {
                  const uint32_t lo = ((m2 & 0x8000u) == 0u && (m2 & 0x7fffu) != 0u) ? 0x0000ffffu : 0u;
                  const uint32_t hi = ((m2 & 0x80000000u) == 0u && (m2 & 0x7fff0000u) != 0u) ? 0xffff0000u : 0u;
                  return val & (lo | hi);
                };
                pk.x = keep(mk.x, pk.x); pk.y = keep(mk.y, pk.y); pk.z = keep(mk.z, pk.z); pk.w = keep(mk.w, pk.w);
              }
              *(uint4*)((__nv_bfloat16*)p.C + off) = pk;
            }
          }
          __syncwarp();
        } else if (m < M_lim) {
          if (p.out_fp32 == 0) {
            __nv_bfloat16* dst = (__nv_bfloat16*)p.C + (int64_t)m * p.ldc + n0;
#pragma unroll
            for (int j = 0; j < 32; ++j) if (n0 + j < p.N) dst[j] = __float2bfloat16(v[j]);
          } else {
            float* dst = (float*)p.C + (int64_t)m * p.ldc + n0;
            if (p.out_fp32 == 1) {
              if (full) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) *(float4*)(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) if (n0 + j < p.N) dst[j] = v[j];
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (n0 + j < p.N) red_add_f32(dst + j, v[j]);
            }
          }
        }
      };
      // software pipeline: the next chunk's TMEM read is in flight while this chunk is stored
      uint32_t ra[32], rb[32];
      tmem_ld_32x32_issue(chunk_addr(0), ra);
#pragma unroll
      for (int i = 0; i < kPer; i += 2) {
        tmem_ld_wait(ra);
        if (i + 1 < kPer) tmem_ld_32x32_issue(chunk_addr(i + 1), rb);
        process(ra, i);
        if (i + 1 < kPer) {
          tmem_ld_wait(rb);
          if (i + 2 < kPer) tmem_ld_32x32_issue(chunk_addr(i + 2), ra);
          process(rb, i + 1);
        }
      }
      // this warp is done reading accumulator `acc`
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR == 2) mbar_arrive_cluster(mapa_u32(smem_u32(tmem_empty_bar + acc), 0));
        else mbar_arrive(tmem_empty_bar + acc);
      }
    }
  }
  tc_fence_before();
  if (PAIR == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR == 2) tmem_dealloc_pair<L::kTmemCols>(tmem_base);
    else tmem_dealloc<L::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

int ceb_make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                       const uint64_t* strides_bytes, const uint32_t* box) {
  PFN_encodeTiled enc = get_encode_fn();
  CEB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t gdims[5], gstr[5];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) { gdims[i] = dims[i]; gbox[i] = box[i]; estr[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, (void*)base, gdims, gstr, gbox, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(0);          // see ceb_make_tmap
  }
  CEB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu/%llu, box %u/%u)",
              (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
  return 0;
}

int ceb_make_tmap(CUtensorMap* out, int elem_bytes, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* estr_in, bool atom32) {
  PFN_encodeTiled enc = get_encode_fn();
  CEB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  CEB_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "ceb_make_tmap: element size %d", elem_bytes);
  cuuint64_t gdims[5], gstr[5];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) { gdims[i] = dims[i]; gbox[i] = box[i]; estr[i] = estr_in ? estr_in[i] : 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank,
            (void*)base, gdims, gstr, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    // a thread that has made no runtime call yet (e.g. an autograd worker thread entering a backward function)
    // has no current context for this driver entry point: bind the primary context and try again
    cudaFree(0);
  }
  CEB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d, elem %d B, dims %llu/%llu, box %u/%u)",
              (int)r, rank, elem_bytes, (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
  return 0;
}

template <int BN, int MODE, int BMAJ = 0, int TF = 0>
static int launch_tile(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t st) {
  using L = SmemLayout<BN, MODE>;
  static bool configured = false;
  if (!configured) {
    CEB_CHECK_CUDA(cudaFuncSetAttribute(tc_tile_kernel<BN, MODE, BMAJ, 1, TF>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
    configured = true;
  }
  TcParams q = p;
  if (q.tile_pix == 0) q.tile_pix = BM * L::kSubTiles;
  if (q.a_bytes == 0) q.a_bytes = BM * 128;
  if (q.ks == 0) { q.ks = 3; q.stride = 1; q.pad = 1; }
  q.m_tiles = (int)ceb_cdiv(p.M, q.tile_pix);
  q.n_tiles = (int)ceb_cdiv(p.N, BN);
  if (q.k_split < 1) { q.k_split = 1; q.k_total = q.k_steps; }
  const int tiles = q.m_tiles * q.n_tiles * q.k_split;
  const int grid = tiles < ceb_num_sms() ? tiles : ceb_num_sms();
  tc_tile_kernel<BN, MODE, BMAJ, 1, TF><<<grid, kTcThreads, L::kTotal, st>>>(ta, tb, q);
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// CTA-pair launch: 2-CTA clusters, one pair per TPC, persistent over pair tiles (2 CTA tiles tall).
template <int BN, int MODE, int TF = 0>
static int launch_tile_pair(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t st) {
  using L = SmemLayout<BN, MODE, 2>;
  static_assert(L::kTotal <= 232448, "CTA-pair stage ring exceeds the 227 KB shared-memory limit");
  static bool configured = false;
  if (!configured) {
    CEB_CHECK_CUDA(cudaFuncSetAttribute(tc_tile_kernel<BN, MODE, 0, 2, TF>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
    configured = true;
  }
  TcParams q = p;
  if (q.tile_pix == 0) q.tile_pix = BM * L::kSubTiles;
  if (q.a_bytes == 0) q.a_bytes = BM * 128;
  if (q.ks == 0) { q.ks = 3; q.stride = 1; q.pad = 1; }
  q.m_tiles = (int)ceb_cdiv(p.M, q.tile_pix);
  q.n_tiles = (int)ceb_cdiv(p.N, BN);
  if (q.k_split < 1) { q.k_split = 1; q.k_total = q.k_steps; }
  const int pair_tiles = (int)ceb_cdiv(q.m_tiles, 2) * q.n_tiles * q.k_split;
  const int max_pairs = ceb_num_sms() / 2;
  const int pairs = pair_tiles < max_pairs ? pair_tiles : max_pairs;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * pairs), 1, 1);
  cfg.blockDim = dim3(kTcThreads, 1, 1);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, tc_tile_kernel<BN, MODE, 0, 2, TF>, ta, tb, q));
  CEB_COUNT_LAUNCH(1);
  return 0;
}

// CEB_TC_PAIR=0 disables / =1 forces the cta_group::2 variant wherever it exists.
static int pair_env() {
  static const int v = [] { const char* e = getenv("CEB_TC_PAIR"); return e ? atoi(e) : -1; }();
  return v;
}
// Default policy (measured, profiles/bench_tc.json): CTA pairs win once there are enough 2-tile
// units to occupy all 74 TPCs and the K loop is long enough to amortise the cluster hand-shakes;
// the halo convolution with a single 64-channel chunk (3 k-steps per tile) stays on one CTA.
static bool use_pair(int mode, int M, int N, int bn, int rows_per_cta_tile, int k_steps) {
  if (pair_env() >= 0) return pair_env() != 0;
  const int64_t pair_tiles = ceb_cdiv(ceb_cdiv(M, rows_per_cta_tile), 2) * ceb_cdiv(N, bn);
  if (pair_tiles < ceb_num_sms() / 2) return false;
  if (mode == 1) return k_steps >= 6;
  return k_steps >= 4;
}

// Widest tile that covers N -- unless the problem has so few row tiles that wide tiles would leave most SMs idle
// (one client's 8 images through a 512-channel layer: 1 row tile; every CTA then streams its whole slice of the
// weight matrix at one SM's L2 bandwidth, ~40 us with two 256-wide tiles): narrower tiles spread that read.
static int pick_bn(int N, int64_t m_tiles = 1 << 20) {
  int bn = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  while (bn > 64 && m_tiles * ceb_cdiv(N, bn) * 2 <= ceb_num_sms()) bn >>= 1;
  return bn;
}

template <int TF>
static int dispatch_tile_t(int bn, int mode, const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t st,
                           bool pair) {
  if (pair) {
    if (mode == 1) {
      if (bn == 64) return launch_tile_pair<64, 1, TF>(ta, tb, p, st);
      if (bn == 128) return launch_tile_pair<128, 1, TF>(ta, tb, p, st);
    } else if (mode == 0) {
      switch (bn) {
        case 64: return launch_tile_pair<64, 0, TF>(ta, tb, p, st);
        case 128: return launch_tile_pair<128, 0, TF>(ta, tb, p, st);
        case 256: return launch_tile_pair<256, 0, TF>(ta, tb, p, st);
      }
    }
    CEB_REQUIRE(false, "unsupported CTA-pair tile configuration (N tile %d, mode %d)", bn, mode);
  }
  if (mode == 1) {
    if (bn == 64) return launch_tile<64, 1, 0, TF>(ta, tb, p, st);
    if (bn == 128) return launch_tile<128, 1, 0, TF>(ta, tb, p, st);
  } else if (mode == 2) {      // GEMM with MN-major B
    switch (bn) {
      case 64: return launch_tile<64, 0, 1, TF>(ta, tb, p, st);
      case 128: return launch_tile<128, 0, 1, TF>(ta, tb, p, st);
      case 256: return launch_tile<256, 0, 1, TF>(ta, tb, p, st);
    }
  } else {
    switch (bn) {
      case 64: return launch_tile<64, 0, 0, TF>(ta, tb, p, st);
      case 128: return launch_tile<128, 0, 0, TF>(ta, tb, p, st);
      case 256: return launch_tile<256, 0, 0, TF>(ta, tb, p, st);
    }
  }
  CEB_REQUIRE(false, "unsupported tile configuration (N tile %d, mode %d)", bn, mode);
}

static int dispatch_tile(int bn, int mode, const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t st,
                         bool pair = false, int tf = 0) {
  return tf ? dispatch_tile_t<1>(bn, mode, ta, tb, p, st, pair) : dispatch_tile_t<0>(bn, mode, ta, tb, p, st, pair);
}

// C[M,N] (row stride ldc) = act(A[M,K] (row stride lda) * B[N,K]^T (row stride ldb) + bias).
// A, B bf16, K contiguous; lda, ldb multiples of 8 elements (16-byte rows for TMA), ldc a multiple
// of 8 (bf16 C) / 4 (fp32 C); M, N, K themselves are arbitrary: TMA zero-fills the K / N tails and
// the epilogue bounds its stores -- this is how the 50 262-row LM head runs without padding the
// embedding matrix (logits live in a buffer with ldc = 50 264).
// out_mode 0: bf16 C, 1: fp32 C, 2: fp32 C += (atomic).
static int gemm_tn_impl(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N, int K,
                        const float* bias, int relu, int out_mode, const int* m_limit, void* stream, int tf = 0,
                        void* c2 = nullptr, const void* aux = nullptr);

CEB_API int ceb_gemm_bf16_tn_ld(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M,
                                int N, int K, const float* bias, int relu, int out_mode, void* stream) {
  return gemm_tn_impl(A, lda, B, ldb, C, ldc, M, N, K, bias, relu, out_mode, nullptr, stream);
}

// Same, but only rows [0, *m_limit_dev) are computed (rounded up to whole 128-row tiles, never past
// M); rows beyond are left untouched.  m_limit_dev is a device int read by the kernel, so the launch
// can sit in a CUDA graph while the row count changes from replay to replay.
CEB_API int ceb_gemm_bf16_tn_rows(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M,
                                  int N, int K, const float* bias, int relu, int out_mode, const int* m_limit_dev,
                                  void* stream) {
  CEB_REQUIRE(m_limit_dev != nullptr, "ceb_gemm_bf16_tn_rows: m_limit_dev is NULL");
  return gemm_tn_impl(A, lda, B, ldb, C, ldc, M, N, K, bias, relu, out_mode, m_limit_dev, stream);
}

static int gemm_tn_impl(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N, int K,
                        const float* bias, int relu, int out_mode, const int* m_limit, void* stream, int tf, void* c2,
                        const void* aux) {
  const int esz = tf ? 4 : 2, bke = 128 / esz, al = 16 / esz;     // element bytes, elements per k-step, 16-byte granule
  CEB_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "ceb_gemm_tn: bad arguments");
  CEB_REQUIRE(lda % al == 0 && ldb % al == 0 && lda >= K && ldb >= K,
              "ceb_gemm_tn: lda, ldb must be multiples of %d and >= K (lda=%lld ldb=%lld K=%d)", al, (long long)lda,
              (long long)ldb, K);
  CEB_REQUIRE(!tf || out_mode != 0, "ceb_gemm_tf32_tn: the output is fp32 (out_mode 1 or 2)");
  CEB_REQUIRE(ldc >= N && ldc % (out_mode == 0 ? 8 : 4) == 0 && ldc < (1ll << 31),
              "ceb_gemm_tn: ldc=%lld must be >= N=%d and a multiple of %d", (long long)ldc, N, out_mode == 0 ? 8 : 4);
  CEB_REQUIRE((((uintptr_t)A | (uintptr_t)B | (uintptr_t)C) & 15) == 0, "ceb_gemm_tn: operands must be 16-byte aligned");
  const int bn = pick_bn(N, ceb_cdiv(M, BM));
  CUtensorMap ta, tb;
  uint64_t da[2] = {(uint64_t)K, (uint64_t)M}, sa[1] = {(uint64_t)lda * esz};
  uint32_t ba[2] = {(uint32_t)bke, BM};
  if (int e = ceb_make_tmap(&ta, esz, A, 2, da, sa, ba, nullptr)) return e;
  const bool pair = use_pair(0, M, N, bn, BM, (int)ceb_cdiv(K, bke));
  uint64_t db[2] = {(uint64_t)K, (uint64_t)N}, sb[1] = {(uint64_t)ldb * esz};
  uint32_t bb[2] = {(uint32_t)bke, (uint32_t)(pair ? bn / 2 : bn)};
  if (int e = ceb_make_tmap(&tb, esz, B, 2, db, sb, bb, nullptr)) return e;
  TcParams p{};
  p.C = C; p.bias = bias; p.M = M; p.N = N; p.ldc = (int)ldc; p.k_steps = (int)ceb_cdiv(K, bke);
  p.relu = relu; p.out_fp32 = out_mode; p.conv = 0; p.m_limit = m_limit; p.C2 = c2; p.mask = aux;
  if (out_mode == 2 && !relu) {
    // accumulate semantics allow split-K: pick the split that minimises (waves) x (k-steps per item + fixed cost)
    const int64_t units = pair ? ceb_cdiv(ceb_cdiv(M, BM), 2) * ceb_cdiv(N, bn) : ceb_cdiv(M, BM) * ceb_cdiv(N, bn);
    const int slots = pair ? ceb_num_sms() / 2 : ceb_num_sms();
    const int kt = p.k_steps;
    int best = 1;
    int64_t best_cost = ceb_cdiv(units, slots) * (kt + 8);
    for (int sp = 2; sp <= 16 && sp * 16 <= kt; ++sp) {
      const int per = (int)ceb_cdiv(kt, sp);
      const int eff = (int)ceb_cdiv(kt, per);
      const int64_t cost = ceb_cdiv(units * eff, slots) * (per + 8);
      if (cost < best_cost) { best_cost = cost; best = eff; }
    }
    if (best > 1) {
      p.k_total = kt;
      p.k_steps = (int)ceb_cdiv(kt, best);
      p.k_split = (int)ceb_cdiv(kt, p.k_steps);
    }
    if (m_limit) { p.auto_split = 1; p.k_total = kt; if (p.k_split < 1) p.k_split = 1; }
  }
  return dispatch_tile(bn, 0, ta, tb, p, (cudaStream_t)stream, pair, tf);
}

// GEMM with a tanh-GELU in the epilogue (bf16 [M][ldc] outputs, N % 32 == 0, ldc % 16 == 0, 32-byte aligned):
//   act = 2: C = A B^T + bias (the pre-activation, kept for the backward), C2 = GELU(C)        (aux unused)
//   act = 3: C = (A B^T) * GELU'(aux)        (aux = a saved pre-activation [M][ldc]; C2 unused)
// One pass over the [M, 4E] MLP activations each way instead of a GEMM plus an element-wise kernel.
CEB_API int ceb_gemm_bf16_tn_gelu(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, void* C2,
                                  const void* aux, int64_t ldc, int M, int N, int K, const float* bias, int act,
                                  void* stream) {
  CEB_REQUIRE(act == 2 || act == 3, "ceb_gemm_bf16_tn_gelu: act=%d (2: forward, 3: backward)", act);
  CEB_REQUIRE(A && B && C && (act == 2 ? C2 != nullptr : aux != nullptr), "ceb_gemm_bf16_tn_gelu: null pointer");
  CEB_REQUIRE(N % 32 == 0 && ldc % 16 == 0 && ((((uintptr_t)C) | ((uintptr_t)C2) | ((uintptr_t)aux)) & 31) == 0,
              "ceb_gemm_bf16_tn_gelu: need N %% 32 == 0, ldc %% 16 == 0 and 32-byte aligned outputs (N=%d ldc=%lld)", N,
              (long long)ldc);
  return gemm_tn_impl(A, lda, B, ldb, C, ldc, M, N, K, bias, act, 0, nullptr, stream, 0, act == 2 ? C2 : nullptr,
                      act == 3 ? aux : nullptr);
}

// fp32 operands on the TF32 tensor-core path: C[M,N] (fp32) = act(A[M,K] * B[N,K]^T + bias).  lda, ldb
// multiples of 4 elements; out_mode 1: store, 2: atomic add.
CEB_API int ceb_gemm_tf32_tn_ld(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int M,
                                int N, int K, const float* bias, int relu, int out_mode, void* stream) {
  return gemm_tn_impl(A, lda, B, ldb, C, ldc, M, N, K, bias, relu, out_mode, nullptr, stream, 1);
}

// Dense form of the above (lda = ldb = K, ldc = N): K % 8 == 0, N % 8 == 0.
CEB_API int ceb_gemm_bf16_tn(const void* A, const void* B, void* C, int M, int N, int K, const float* bias,
                             int relu, int out_mode, void* stream) {
  CEB_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "ceb_gemm_bf16_tn: bad arguments");
  CEB_REQUIRE(K % 8 == 0 && N % 8 == 0, "ceb_gemm_bf16_tn: K and N must be multiples of 8 (K=%d N=%d)", K, N);
  return ceb_gemm_bf16_tn_ld(A, K, B, K, C, N, M, N, K, bias, relu, out_mode, stream);
}

// C[M,N] = act(A[M,K] * B[K,N] + bias): like ceb_gemm_bf16_tn but B is row-major [K][N]
// (GPT-2 Conv1D weights; data gradients dX = dY * W).  K % 8 == 0, N % 8 == 0.
CEB_API int ceb_gemm_bf16_nn(const void* A, const void* B, void* C, int M, int N, int K, const float* bias,
                             int relu, int out_mode, void* stream) {
  CEB_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "ceb_gemm_bf16_nn: bad arguments");
  CEB_REQUIRE(K % 8 == 0 && N % 8 == 0, "ceb_gemm_bf16_nn: K and N must be multiples of 8 (K=%d N=%d)", K, N);
  const int bn = pick_bn(N);
  CUtensorMap ta, tb;
  uint64_t da[2] = {(uint64_t)K, (uint64_t)M}, sa[1] = {(uint64_t)K * 2};
  uint32_t ba[2] = {BK, BM};
  if (int e = ceb_make_tmap_bf16(&ta, A, 2, da, sa, ba)) return e;
  uint64_t db[2] = {(uint64_t)N, (uint64_t)K}, sb[1] = {(uint64_t)N * 2};
  uint32_t bb[2] = {64, BK};
  if (int e = ceb_make_tmap_bf16(&tb, B, 2, db, sb, bb)) return e;
  TcParams p{};
  p.C = C; p.bias = bias; p.M = M; p.N = N; p.ldc = N; p.k_steps = (int)ceb_cdiv(K, BK);
  p.relu = relu; p.out_fp32 = out_mode; p.conv = 0;
  return dispatch_tile(bn, 2, ta, tb, p, (cudaStream_t)stream);
}

// y[N,H,W,Cout] = act(conv3x3_s1_p1(x[N,H,W,Cin], w[Cout,3,3,Cin])), all bf16 NHWC / KRSC.
// Cin % 64 == 0, Cout % 8 == 0, W <= 128 and 128 % W == 0, H % (128/W) == 0 or (128/W) % H == 0.
static int conv3x3_impl(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout, const float* bias,
                        int relu, int out_mode, const void* relu_mask, const void* resid, void* y2, void* stream, int tf = 0);

// fp32 NHWC activations / KRSC weights on the TF32 tensor-core path (fp32 output, mask, residual).
// Cin % 32 == 0; resid (nullable) / y2 (nullable) as in ceb_conv3x3_bf16_nhwc_res.
CEB_API int ceb_conv3x3_tf32_nhwc(const float* x, const float* w, float* y, int Nimg, int H, int W, int Cin, int Cout,
                                  const float* bias, int relu, const float* relu_mask, const float* resid, float* y2,
                                  void* stream) {
  return conv3x3_impl(x, w, y, Nimg, H, W, Cin, Cout, bias, relu, 1, relu_mask, resid, y2, stream, 1);
}

CEB_API int ceb_conv3x3_bf16_nhwc(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin,
                                  int Cout, const float* bias, int relu, int out_mode, const void* relu_mask,
                                  void* stream) {
  return conv3x3_impl(x, w, y, Nimg, H, W, Cin, Cout, bias, relu, out_mode, relu_mask, nullptr, nullptr, stream);
}

// Same with a fused residual add (bf16 output only; Cout % 32 == 0, 32-byte aligned tensors):
// y2 == NULL: y = act(conv) + resid;  y2 != NULL: y = act(conv) and y2 = y + resid (the pre-add
// activation stays available as the ReLU mask of the backward pass).
CEB_API int ceb_conv3x3_bf16_nhwc_res(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout,
                                      const float* bias, int relu, const void* relu_mask, const void* resid, void* y2,
                                      void* stream) {
  CEB_REQUIRE(resid != nullptr, "ceb_conv3x3_bf16_nhwc_res: resid is NULL");
  CEB_REQUIRE(Cout % 32 == 0 && ((((uintptr_t)y) | ((uintptr_t)resid) | ((uintptr_t)y2) | ((uintptr_t)relu_mask)) & 31) == 0,
              "ceb_conv3x3_bf16_nhwc_res: need Cout %% 32 == 0 and 32-byte aligned tensors (Cout=%d)", Cout);
  return conv3x3_impl(x, w, y, Nimg, H, W, Cin, Cout, bias, relu, 0, relu_mask, resid, y2, stream);
}

// General NHWC convolution: filter ks x ks (1 or 3), stride 1 or 2, padding pad; H, W are the INPUT map dims.
static int conv_impl(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout, int ks, int stride,
                     int pad, const float* bias, int relu, int out_mode, const void* relu_mask, const void* resid, void* y2,
                     void* stream, int tf) {
  const int esz = tf ? 4 : 2, bke = 128 / esz;
  CEB_REQUIRE(x && w && y && Nimg > 0, "ceb_conv_nhwc: bad arguments");
  CEB_REQUIRE((ks == 1 || ks == 3) && (stride == 1 || stride == 2) && pad >= 0 && pad <= 1,
              "ceb_conv_nhwc: filter %d stride %d pad %d not supported", ks, stride, pad);
  CEB_REQUIRE(Cin % bke == 0 && Cout % 8 == 0, "ceb_conv_nhwc: need Cin %% %d == 0, Cout %% 8 == 0 (Cin=%d Cout=%d)", bke, Cin, Cout);
  const int Ho = (H + 2 * pad - ks) / stride + 1, Wo = (W + 2 * pad - ks) / stride + 1;
  CEB_REQUIRE(Ho > 0 && Wo > 0 && Wo <= BM, "ceb_conv_nhwc: output map %d x %d", Ho, Wo);
  // An M tile = whole rows of the OUTPUT map: tile_h rows of one image, or tile_n whole images.
  int tile_h = 1, tile_n = 1;
  for (int th = 1; th <= Ho; ++th)
    if (Ho % th == 0 && th * Wo <= BM) tile_h = th;
  if (tile_h == Ho) tile_n = BM / (Ho * Wo);
  if (tile_n < 1) tile_n = 1;
  if (tile_n > Nimg) tile_n = Nimg;
  const int tile_pix = tile_h * Wo * tile_n;
  const int bn = pick_bn(Cout, ceb_cdiv((int64_t)Nimg * Ho * Wo, tile_pix));
  // halo mode: 3x3 / stride 1, 256-pixel tiles that stay inside one image, weights tile <= 128 wide
  static const bool force_v1 = getenv("CEB_CONV_V1") != nullptr;
  const int rows256 = (Wo > 0 && 256 % Wo == 0) ? 256 / Wo : 0;
  const bool halo = !force_v1 && ks == 3 && stride == 1 && pad == 1 && bn <= 128 && (Wo == 16 || Wo == 32) && rows256 > 0 &&
                    rows256 <= Ho && Ho % rows256 == 0;
  CUtensorMap ta, tb;
  uint64_t da[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)Nimg};
  uint64_t sa[3] = {(uint64_t)Cin * esz, (uint64_t)W * Cin * esz, (uint64_t)H * W * Cin * esz};
  // with a traversal stride the box spans count * stride input elements (TMA keeps every stride-th one)
  uint32_t ba[4] = {(uint32_t)bke, (uint32_t)(Wo * stride), (uint32_t)((halo ? rows256 + 2 : tile_h) * stride),
                    (uint32_t)(halo ? 1 : tile_n)};
  uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
  if (int e = ceb_make_tmap(&ta, esz, x, 4, da, sa, ba, es)) return e;
  const int K = ks * ks * Cin;
  const int k_steps = (halo ? 3 : ks * ks) * (Cin / bke);
  const bool pair = use_pair(halo ? 1 : 0, Nimg * Ho * Wo, Cout, bn, halo ? 256 : tile_pix, k_steps);
  uint64_t db[2] = {(uint64_t)K, (uint64_t)Cout}, sb[1] = {(uint64_t)K * esz};
  uint32_t bb[2] = {(uint32_t)bke, (uint32_t)(pair ? bn / 2 : bn)};
  if (int e = ceb_make_tmap(&tb, esz, w, 2, db, sb, bb, nullptr)) return e;
  TcParams p{};
  p.C = y; p.bias = bias; p.mask = relu_mask; p.M = Nimg * Ho * Wo; p.N = Cout; p.ldc = Cout; p.resid = resid; p.C2 = y2;
  p.k_steps = k_steps;
  p.relu = relu; p.out_fp32 = out_mode; p.conv = 1; p.H = Ho; p.W = Wo; p.cin_steps = Cin / bke; p.tile_h = tile_h;
  p.halo_bytes = halo ? (rows256 + 2) * Wo * 128 : 0;
  p.ks = ks; p.stride = stride; p.pad = pad;
  p.tile_pix = halo ? 256 : tile_pix;
  p.a_bytes = tile_pix * 128;
  return dispatch_tile(bn, halo ? 1 : 0, ta, tb, p, (cudaStream_t)stream, pair, tf);
}

static int conv3x3_impl(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout, const float* bias,
                        int relu, int out_mode, const void* relu_mask, const void* resid, void* y2, void* stream, int tf) {
  return conv_impl(x, w, y, Nimg, H, W, Cin, Cout, 3, 1, 1, bias, relu, out_mode, relu_mask, resid, y2, stream, tf);
}

// y[N,Ho,Wo,Cout] = act(conv_{ks x ks, stride, pad}(x[N,H,W,Cin], w[Cout,ks,ks,Cin])) (+ mask / residual as in
// ceb_conv3x3_*): ks in {1, 3}, stride in {1, 2}, any map size whose output width is <= 128 (rows of the
// output map are tiled whole, so widths such as 28, 14, 7, 3 work).  f32 = 0: bf16 tensors (out_mode as in
// ceb_conv3x3_bf16_nhwc, mask / resid / y2 bf16); f32 = 1: fp32 tensors on the TF32 path.
CEB_API int ceb_conv2d_nhwc(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout, int ks,
                            int stride, int pad, int relu, int f32, const void* relu_mask, const void* resid, void* y2,
                            void* stream) {
  return conv_impl(x, w, y, Nimg, H, W, Cin, Cout, ks, stride, pad, nullptr, relu, f32 ? 1 : 0, relu_mask, resid, y2,
                   stream, f32);
}
