// Peer-memory collectives over NVLink 5 / NVSwitch, fused with the server update (sm_100a).
//
// The reference aggregates with one NCCL `reduce` to a parameter-server rank, updates the
// weights on the CPU and re-downloads them to every worker through host shared memory
// (/root/reference/CommEfficient/fed_worker.py:41,138, fed_aggregator.py:327-332,455).  Here
// every rank owns 1/W of the server state.  ONE kernel per round
//   1. waits until every rank's transmit buffer is complete (flag barrier in peer memory),
//   2. pulls its shard of every rank's buffer with 16-byte peer loads and sums them in rank
//      order (reduce-scatter; the sum of a shard is computed once, so replicas agree bitwise),
//   3. applies the server rule to the shard while it is in registers
//        sketch : Vv = rho*Vv + s/N ; Ve += Vv            -> broadcasts Ve
//        dense  : Vv = rho*Vv + s/N ; w -= lr*Vv           -> broadcasts w
//        plain  : broadcasts s                             (all-reduce),
//   4. pushes the result into every rank's destination buffer with 16-byte peer stores
//      (all-gather), and
//   5. signals completion and waits for the other ranks' pushes to land.
// No NCCL call is involved.  Buffers are cudaMalloc'd once and exchanged as CUDA IPC handles.
#include "peer.cuh"
#include <cstdlib>

struct OpSum {
  __device__ __forceinline__ float4 operator()(int64_t, float4 s) const { return s; }
};
struct OpSketch {   // Vv/Ve are this rank's shards, indexed from the shard start
  float* Vv; float* Ve; float inv_n, rho; int error_mode; int64_t lo;
  __device__ __forceinline__ float4 operator()(int64_t i, float4 s) const {
    float4 v = *reinterpret_cast<float4*>(Vv + (i - lo));
    v.x = fmaf(rho, v.x, s.x * inv_n); v.y = fmaf(rho, v.y, s.y * inv_n);
    v.z = fmaf(rho, v.z, s.z * inv_n); v.w = fmaf(rho, v.w, s.w * inv_n);
    *reinterpret_cast<float4*>(Vv + (i - lo)) = v;
    if (error_mode == 0) return v;
    float4 e = v;
    if (error_mode == 2) {
      e = *reinterpret_cast<float4*>(Ve + (i - lo));
      e.x += v.x; e.y += v.y; e.z += v.z; e.w += v.w;
    }
    *reinterpret_cast<float4*>(Ve + (i - lo)) = e;
    return e;
  }
};
struct OpDense {    // Vv is a shard; w is read from this rank's own replica (dst[rank])
  float* Vv; const float* w; const float* lr_vec; float inv_n, rho, lr; int64_t lo;
  __device__ __forceinline__ float4 operator()(int64_t i, float4 s) const {
    float4 v = *reinterpret_cast<float4*>(Vv + (i - lo));
    v.x = fmaf(rho, v.x, s.x * inv_n); v.y = fmaf(rho, v.y, s.y * inv_n);
    v.z = fmaf(rho, v.z, s.z * inv_n); v.w = fmaf(rho, v.w, s.w * inv_n);
    *reinterpret_cast<float4*>(Vv + (i - lo)) = v;
    float4 x = *reinterpret_cast<const float4*>(w + i);
    float4 l = lr_vec ? *reinterpret_cast<const float4*>(lr_vec + i) : make_float4(lr, lr, lr, lr);
    x.x -= l.x * v.x; x.y -= l.y * v.y; x.z -= l.z * v.z; x.w -= l.w * v.w;
    return x;
  }
};

// n is a multiple of 4 (callers pad); shard bounds are multiples of 4.  W = world size
// (compile time, so the per-peer loads unroll), U = float4s per thread per step chosen so
// that U*W ~ 8 sixteen-byte peer loads are in flight per thread: NVLink needs ~2-3 us x
// 770 GB/s ~ 2 MB outstanding per GPU to run at link speed.
struct OpTopkState {   // true_topk: Vv = rho*Vv + s/N ; Ve += Vv on this rank's coordinate shard; nothing is gathered
  float* Vv; float* Ve; float inv_n, rho; int64_t lo;
  __device__ __forceinline__ float4 operator()(int64_t i, float4 s) const {
    float4 v = *reinterpret_cast<float4*>(Vv + (i - lo));
    v.x = fmaf(rho, v.x, s.x * inv_n); v.y = fmaf(rho, v.y, s.y * inv_n);
    v.z = fmaf(rho, v.z, s.z * inv_n); v.w = fmaf(rho, v.w, s.w * inv_n);
    *reinterpret_cast<float4*>(Vv + (i - lo)) = v;
    float4 e = *reinterpret_cast<float4*>(Ve + (i - lo));
    e.x += v.x; e.y += v.y; e.z += v.z; e.w += v.w;
    *reinterpret_cast<float4*>(Ve + (i - lo)) = e;
    return e;
  }
};

// GATHER = false: reduce-scatter + update only (the shard's result stays with its owner; the second barrier
// then only means "every rank has finished reading every transmit buffer").
template <class Op, int W, int U, bool GATHER = true>
__global__ void __launch_bounds__(kPeerThreads)
peer_rs_op_ag_kernel(CebPeerCtx c, int64_t lo, int64_t hi, Op op) {
  __shared__ bool is_last;
  if (blockIdx.x == 0) peer_signal_all(c, 0);     // "my transmit buffer is complete"
  peer_wait_all(c, 0);
  const int64_t lane_step = (int64_t)blockDim.x * 4;
  const int64_t stride = (int64_t)gridDim.x * lane_step * U;
  for (int64_t base = lo + (int64_t)blockIdx.x * lane_step * U + (int64_t)threadIdx.x * 4; base < hi;
       base += stride) {
    float4 part[W][U];
#pragma unroll
    for (int q = 0; q < W; ++q)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = base + u * lane_step;
        part[q][u] = i < hi ? ld_peer_f4(c.src[q] + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + u * lane_step;
      if (i >= hi) break;
      float4 s = part[0][u];
#pragma unroll
      for (int q = 1; q < W; ++q) {
        s.x += part[q][u].x; s.y += part[q][u].y; s.z += part[q][u].z; s.w += part[q][u].w;
      }
      const float4 out = op(i, s);
      if (GATHER) {
#pragma unroll
        for (int q = 0; q < W; ++q) st_peer_f4(c.dst[q] + i, out);
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = atomicAdd(c.ticket, 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) *c.ticket = 0;
  }
  __syncthreads();
  if (is_last) {                                  // "my pushes have landed everywhere"
    __threadfence_system();
    peer_signal_all(c, 1);
    peer_wait_all(c, 1);
  }
}

// NVLS variant: the reduce-scatter is ONE multimem.ld_reduce per 16 bytes (the switch adds the W copies and
// returns the sum: this GPU receives 1/W of the bytes the peer-load version pulls), the all-gather ONE
// multimem.st (the switch replicates the store into every rank's copy: 1/W of the bytes leave this GPU).
// Same barriers, same Op functors, no dependence on W at compile time.
template <class Op, int U, bool GATHER = true>
__global__ void __launch_bounds__(kPeerThreads)
peer_rs_op_ag_nvls_kernel(CebPeerCtx c, int64_t lo, int64_t hi, Op op) {
  __shared__ bool is_last;
  if (blockIdx.x == 0) peer_signal_all(c, 0);
  peer_wait_all(c, 0);
  const int64_t lane_step = (int64_t)blockDim.x * 4;
  const int64_t stride = (int64_t)gridDim.x * lane_step * U;
  for (int64_t base = lo + (int64_t)blockIdx.x * lane_step * U + (int64_t)threadIdx.x * 4; base < hi;
       base += stride) {
    float4 s[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + u * lane_step;
      s[u] = i < hi ? multimem_ld_reduce_f4(c.mc_src + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + u * lane_step;
      if (i >= hi) break;
      const float4 out = op(i, s[u]);
      if (GATHER) multimem_st_f4(c.mc_dst + i, out);
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = atomicAdd(c.ticket, 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) *c.ticket = 0;
  }
  __syncthreads();
  if (is_last) {
    __threadfence_system();
    peer_signal_all(c, 1);
    peer_wait_all(c, 1);
  }
}

template <class Op, bool GATHER = true>
static void launch_peer(const CebPeerCtx& c, int64_t lo, int64_t hi, Op op, cudaStream_t st) {
  auto grid_for = [&](int U) {
    int64_t need = ceb_cdiv((hi - lo) / 4 + 1, (int64_t)kPeerThreads * U);
    int64_t cap = (int64_t)ceb_num_sms() * 4;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
  };
  if (c.mc_src && (c.mc_dst || !GATHER)) {
    peer_rs_op_ag_nvls_kernel<Op, 4, GATHER><<<grid_for(4), kPeerThreads, 0, st>>>(c, lo, hi, op);
    return;
  }
  // CEB_PEER_U=1|2|4 overrides the float4s in flight per thread and peer for the 8-rank kernels (A/B timing)
  int u8 = 1;
  if (c.world == 8) {
    const char* e = getenv("CEB_PEER_U");
    if (e) u8 = atoi(e);
  }
  switch (c.world) {
#define CEB_PEER_CASE(W_, U_) \
    case W_: peer_rs_op_ag_kernel<Op, W_, U_, GATHER><<<grid_for(U_), kPeerThreads, 0, st>>>(c, lo, hi, op); break;
    CEB_PEER_CASE(1, 4) CEB_PEER_CASE(2, 4) CEB_PEER_CASE(3, 2) CEB_PEER_CASE(4, 2)
    CEB_PEER_CASE(5, 1) CEB_PEER_CASE(6, 1) CEB_PEER_CASE(7, 1)
    case 8:
      if (u8 == 4) peer_rs_op_ag_kernel<Op, 8, 4, GATHER><<<grid_for(4), kPeerThreads, 0, st>>>(c, lo, hi, op);
      else if (u8 == 2) peer_rs_op_ag_kernel<Op, 8, 2, GATHER><<<grid_for(2), kPeerThreads, 0, st>>>(c, lo, hi, op);
      else peer_rs_op_ag_kernel<Op, 8, 1, GATHER><<<grid_for(1), kPeerThreads, 0, st>>>(c, lo, hi, op);
      break;
#undef CEB_PEER_CASE
    default: break;
  }
}

static int fill_ctx(CebPeerCtx& c, const void* const* src, const void* const* dst, const void* const* flags,
                    void* ticket, int rank, int world, uint32_t epoch) {
  CEB_REQUIRE(world >= 1 && world <= kMaxPeers, "peer collective: world=%d outside [1,%d]", world, kMaxPeers);
  CEB_REQUIRE(rank >= 0 && rank < world, "peer collective: rank=%d outside [0,%d)", rank, world);
  CEB_REQUIRE(src && dst && flags && ticket, "peer collective: null pointer table");
  for (int q = 0; q < kMaxPeers; ++q) {
    c.src[q] = q < world ? (float*)src[q] : nullptr;
    c.dst[q] = q < world ? (float*)dst[q] : nullptr;
    c.flags[q] = q < world ? (uint32_t*)flags[q] : nullptr;
  }
  c.ticket = (uint32_t*)ticket;
  c.rank = rank;
  c.world = world;
  c.epoch = epoch;
  c.mc_src = c.mc_dst = nullptr;
  return 0;
}

static void shard_bounds(int64_t n, int rank, int world, int64_t* lo, int64_t* hi) {
  int64_t per = ceb_cdiv(ceb_cdiv(n, world), 4) * 4;
  *lo = per * rank < n ? per * rank : n;
  *hi = per * (rank + 1) < n ? per * (rank + 1) : n;
}

CEB_API int64_t ceb_peer_shard_elems(int64_t n, int world) {
  return ceb_cdiv(ceb_cdiv(n, world), 4) * 4;
}

// src/dst/flags: host arrays of `world` device pointers (peer-mapped).  n % 4 == 0.
// mode 0: dst = sum(src)                                      (all-reduce; dst may equal src)
// mode 1: sketch -- Vv,Ve shards updated, dst = Ve             (error_mode 0: dst = Vv)
// mode 2: dense  -- Vv shard updated, dst (= weights) -= lr*Vv
// mode 3: true_topk state -- Vv, Ve shards updated (Ve += Vv), nothing gathered (dst unused; pass src)
static int fused_reduce_impl(int mode, const void* const* src, const void* const* dst, const void* const* flags,
                             void* ticket, int rank, int world, unsigned int epoch, int64_t n, float* Vv_shard,
                             float* Ve_shard, float inv_n, float rho, int error_mode, float lr, const float* lr_vec,
                             void* mc_src, void* mc_dst, void* stream);

CEB_API int ceb_peer_fused_reduce(int mode, const void* const* src, const void* const* dst,
                                  const void* const* flags, void* ticket, int rank, int world,
                                  unsigned int epoch, int64_t n, float* Vv_shard, float* Ve_shard,
                                  float inv_n, float rho, int error_mode, float lr, const float* lr_vec,
                                  void* stream) {
  return fused_reduce_impl(mode, src, dst, flags, ticket, rank, world, epoch, n, Vv_shard, Ve_shard, inv_n, rho,
                           error_mode, lr, lr_vec, nullptr, nullptr, stream);
}

// The same collective through the NVSwitch's multicast / in-switch reduction (NVLS): mc_src / mc_dst are the
// multicast addresses of the buffers whose per-rank addresses are src[] / dst[] (both bound to a multicast
// object spanning all `world` GPUs).  The flag barriers still run over the per-rank mappings.
CEB_API int ceb_peer_fused_reduce_nvls(int mode, const void* const* src, const void* const* dst,
                                       const void* const* flags, void* ticket, int rank, int world,
                                       unsigned int epoch, int64_t n, float* Vv_shard, float* Ve_shard,
                                       float inv_n, float rho, int error_mode, float lr, const float* lr_vec,
                                       void* mc_src, void* mc_dst, void* stream) {
  CEB_REQUIRE(mc_src && (mc_dst || mode == 3), "ceb_peer_fused_reduce_nvls: null multicast address");
  CEB_REQUIRE(((((uintptr_t)mc_src) | ((uintptr_t)mc_dst)) & 15) == 0, "ceb_peer_fused_reduce_nvls: multicast addresses must be 16-byte aligned");
  return fused_reduce_impl(mode, src, dst, flags, ticket, rank, world, epoch, n, Vv_shard, Ve_shard, inv_n, rho,
                           error_mode, lr, lr_vec, mc_src, mc_dst, stream);
}

static int fused_reduce_impl(int mode, const void* const* src, const void* const* dst, const void* const* flags,
                             void* ticket, int rank, int world, unsigned int epoch, int64_t n, float* Vv_shard,
                             float* Ve_shard, float inv_n, float rho, int error_mode, float lr, const float* lr_vec,
                             void* mc_src, void* mc_dst, void* stream) {
  CEB_REQUIRE(n > 0 && n % 4 == 0, "ceb_peer_fused_reduce: n=%lld must be a positive multiple of 4", (long long)n);
  CebPeerCtx c;
  if (int e = fill_ctx(c, src, dst, flags, ticket, rank, world, epoch)) return e;
  c.mc_src = (float*)mc_src;
  c.mc_dst = (float*)mc_dst;
  int64_t lo, hi;
  shard_bounds(n, rank, world, &lo, &hi);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0) {
    launch_peer(c, lo, hi, OpSum{}, st);
  } else if (mode == 1) {
    CEB_REQUIRE(Vv_shard && (error_mode == 0 || Ve_shard), "ceb_peer_fused_reduce: sketch mode needs state shards");
    launch_peer(c, lo, hi, OpSketch{Vv_shard, Ve_shard, inv_n, rho, error_mode, lo}, st);
  } else if (mode == 2) {
    CEB_REQUIRE(Vv_shard, "ceb_peer_fused_reduce: dense mode needs the momentum shard");
    launch_peer(c, lo, hi, OpDense{Vv_shard, (const float*)dst[rank], lr_vec, inv_n, rho, lr, lo}, st);
  } else if (mode == 3) {
    CEB_REQUIRE(Vv_shard && Ve_shard, "ceb_peer_fused_reduce: true_topk mode needs both state shards");
    launch_peer<OpTopkState, false>(c, lo, hi, OpTopkState{Vv_shard, Ve_shard, inv_n, rho, lo}, st);
  } else {
    CEB_REQUIRE(false, "ceb_peer_fused_reduce: unknown mode %d", mode);
  }
  CEB_COUNT_LAUNCH(1);
  CEB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ symmetric allocation (CUDA IPC)
// These three are the only entry points of the library that allocate: symmetric buffers must be
// whole cudaMalloc allocations to be exportable.  handle_out/handle: 64 bytes (cudaIpcMemHandle_t).
CEB_API int ceb_peer_alloc(int64_t bytes, void** ptr_out, void* handle_out) {
  CEB_REQUIRE(bytes > 0 && ptr_out && handle_out, "ceb_peer_alloc: bad arguments");
  void* p = nullptr;
  CEB_CHECK_CUDA(cudaMalloc(&p, (size_t)bytes));
  CEB_CHECK_CUDA(cudaMemset(p, 0, (size_t)bytes));
  cudaIpcMemHandle_t h;
  CEB_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = p;
  return 0;
}

CEB_API int ceb_peer_open(const void* handle, void** ptr_out) {
  CEB_REQUIRE(handle && ptr_out, "ceb_peer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  CEB_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = p;
  return 0;
}

CEB_API int ceb_peer_release(void* ptr, int is_owner) {
  if (!ptr) return 0;
  if (is_owner) CEB_CHECK_CUDA(cudaFree(ptr));
  else CEB_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

CEB_API int ceb_peer_handle_bytes() { return (int)sizeof(cudaIpcMemHandle_t); }
