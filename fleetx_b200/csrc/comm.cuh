// comm.cuh — what the cross-GPU kernels (comm.cu, comm_rows.cu) share: system-scope flag / counter
// primitives on the ranks' signal pads, the two rank barriers, and the NVSwitch multimem accessors.
//
// Flag-page protocol (one page per rank, zero-initialised by the host, mapped into every peer). Every call
// on a set of buffers has an epoch 1, 2, 3, ... identical on all ranks, and EVERY kernel of this library
// advances every slot exactly once per call, so different kernels can be mixed on one set of pages:
//   SLOT_READY[r]   rank r stores the epoch into every rank's page when it reaches the call (= everything
//                   before the call in its stream has finished)
//   SLOT_RS_DONE    +1 per rank per call: "my reduced / published rows are stored"          (barrier A)
//   SLOT_AG_DONE    +1 per rank per call: "I have finished reading every copy"              (barrier B)
// A rank's contribution to a counter is made by the LAST of its blocks to get there (a ticket counter in the
// rank's own page elects it), so the targets are epoch * world whatever grid each call used: calls with
// different num_blocks can share a page.
#pragma once
#include "common.cuh"
#include "tc5.cuh"

namespace fx {
namespace comm {

using tc5::bf16_hi;
using tc5::bf16_lo;
using tc5::pack_bf16;

constexpr int MAX_RANKS = 8;
// u32 slots inside each rank's signal pad (byte offset 4096 keeps clear of torch's own use)
constexpr int SLOT_BASE = 1024;            // in u32 units
constexpr int SLOT_READY = SLOT_BASE;      // [MAX_RANKS] epoch flag written by each source rank
constexpr int SLOT_RS_DONE = SLOT_BASE + 16;
constexpr int SLOT_AG_DONE = SLOT_BASE + 17;
constexpr int SLOT_TICKET_A = SLOT_BASE + 24;  // local: blocks of THIS rank that reached barrier A / B
constexpr int SLOT_TICKET_B = SLOT_BASE + 25;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atom_add_acq_rel_gpu(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
// Wait until *p has reached `target` (wrap-safe). A peer that never arrives would hang the GPU, so
// the spin carries a 5 s watchdog.
__device__ __forceinline__ void spin_until(const uint32_t* p, uint32_t target, int tag) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (static_cast<int32_t>(ld_acquire_sys(p) - target) < 0) {
    if ((++spins & 1023u) == 0) {
      const uint64_t now = tc5::global_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > 5000000000ull) {
        printf("fleetx_b200: tp collective watchdog: block %d tag %d target %u have %u\n", blockIdx.x, tag, target,
               ld_acquire_sys(p));
        __trap();
      }
    }
  }
}

// The helpers below take the kernel's parameter struct itself (any type with members
//   uint64_t sig[MAX_RANKS]  signal pad of every rank, as mapped on this GPU;  int rank, world;  uint32_t epoch)
// so the flag addresses are read straight from kernel-parameter space.

// Entry handshake: every rank has reached this call, i.e. (stream order) finished everything before it.
template <class P>
__device__ __forceinline__ void ranks_ready(const P& c) {
  const int tid = threadIdx.x;
  uint32_t* my_sig = reinterpret_cast<uint32_t*>(c.sig[c.rank]);
  if (blockIdx.x == 0 && tid < c.world)
    st_release_sys(reinterpret_cast<uint32_t*>(c.sig[tid]) + SLOT_READY + c.rank, c.epoch);
  if (tid < c.world) spin_until(my_sig + SLOT_READY + tid, c.epoch, 0);
  __syncthreads();
}

// Rank barrier over `slot` (SLOT_RS_DONE with SLOT_TICKET_A, or SLOT_AG_DONE with SLOT_TICKET_B): whatever
// this block wrote before the call is visible system-wide to whoever passes the barrier afterwards. The last
// block of this rank to arrive adds the rank's one count to every rank's counter. wait = false only counts
// in (the epoch bookkeeping of a kernel that does not need this barrier); wait_all_blocks = false lets
// only block 0 wait (a barrier that protects the buffer against the NEXT kernel, which stream order puts
// behind the whole grid anyway).
template <class P>
__device__ __forceinline__ void rank_barrier(const P& c, int slot, int ticket_slot, bool wait, bool wait_all_blocks,
                                             int tag) {
  uint32_t* my_sig = reinterpret_cast<uint32_t*>(c.sig[c.rank]);
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t t = atom_add_acq_rel_gpu(my_sig + ticket_slot, 1u);
    if (t == gridDim.x - 1) {          // every block of this rank has arrived
      atomicExch(my_sig + ticket_slot, 0u);   // ready for the next call (stream order puts it behind this one)
      __threadfence_system();
      for (int i = 0; i < c.world; ++i) red_release_sys_add(reinterpret_cast<uint32_t*>(c.sig[i]) + slot, 1u);
    }
    if (wait && (wait_all_blocks || blockIdx.x == 0)) spin_until(my_sig + slot, c.epoch * static_cast<uint32_t>(c.world), tag);
  }
  __syncthreads();
}

__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const uint4* ptr) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(ptr)
               : "memory");
  return v;
}
// plain (weak) 16-byte load that does not allocate in L1: the data was published by the barrier
__device__ __forceinline__ uint4 ld_weak_v4(const uint4* ptr) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(ptr)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_v4(uint64_t mc_addr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
// 8 bf16 summed over all replicas by the switch, fp32 accumulation
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(uint64_t mc_addr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}
// 4 fp32 summed over all replicas by the switch
__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(uint64_t mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

}  // namespace comm
}  // namespace fx
