// Device helpers shared by the peer-memory kernels (peer.cu, dselect.cu): system-scope flag barriers over
// CUDA-IPC-mapped memory with a BOUNDED wait -- a rank that died before signalling must not leave the other
// GPUs spinning until the host watchdog fires: after kPeerSpinNs of wall time (%globaltimer) the waiter records
// the phase in an error word and gives up, which the host reports (parallel/symm.py:PeerGroup.check).
#pragma once
#include "common.cuh"

constexpr int kMaxPeers = 8;
constexpr int kPeerThreads = 256;
constexpr int kFlagStride = 32;   // uint32 slots per phase (>= kMaxPeers, keeps phases on separate lines)

constexpr unsigned long long kPeerSpinNs = 120ull * 1000 * 1000 * 1000;   // two minutes: ranks legitimately arrive seconds apart (graph captures, first-touch allocations)

__device__ __forceinline__ unsigned long long peer_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// true once a wait that began at t0 has used up its budget (the clock is only read every 256 polls)
__device__ __forceinline__ bool peer_spin_expired(unsigned long long& polls, unsigned long long t0) {
  return ((++polls & 255u) == 0) && (peer_now_ns() - t0 > kPeerSpinNs);
}

struct CebPeerCtx {
  float* src[kMaxPeers];        // every rank's transmit (staging) buffer, peer-mapped
  float* dst[kMaxPeers];        // every rank's destination buffer, peer-mapped
  uint32_t* flags[kMaxPeers];   // every rank's flag block: [phases][kFlagStride]
  uint32_t* ticket;             // local: finished-block counter; ticket[2] = error word (0 = fine)
  int rank, world;
  uint32_t epoch;
  // NVLS: multicast addresses of the same two buffers (NULL when the buffers are not bound to a multicast
  // object).  A multimem load through mc_src returns the SUM over all ranks' copies, reduced inside the
  // NVSwitch; a multimem store through mc_dst lands in every rank's copy.
  float* mc_src;
  float* mc_dst;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// serial-number comparison so the 32-bit epoch may wrap
__device__ __forceinline__ bool epoch_reached(uint32_t seen, uint32_t want) { return (int32_t)(seen - want) >= 0; }

__device__ __forceinline__ void peer_wait_all(const CebPeerCtx& c, int phase) {
  if ((int)threadIdx.x < c.world) {
    const uint32_t* slot = c.flags[c.rank] + phase * kFlagStride + threadIdx.x;
    unsigned long long polls = 0;
    const unsigned long long t0 = peer_now_ns();
    while (!epoch_reached(ld_acquire_sys(slot), c.epoch)) {
      __nanosleep(20);
      if (peer_spin_expired(polls, t0)) {       // the peer never arrived: say so and stop waiting
        atomicExch(c.ticket + 2, 0x80000000u | ((uint32_t)phase << 8) | (uint32_t)threadIdx.x);
        break;
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void peer_signal_all(const CebPeerCtx& c, int phase) {
  if ((int)threadIdx.x < c.world)
    st_release_sys(c.flags[threadIdx.x] + phase * kFlagStride + c.rank, c.epoch);
}

