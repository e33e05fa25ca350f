// Peer-memory (NVLink 5 / NVSwitch) communicator shared by the stand-alone one-shot all-reduce
// (comm.cu) and the row-parallel GEMM whose epilogue reduces across the tensor-parallel group
// (fp8_gemm_smallm.cuh).  Replaces the reference's custom all-reduce
// (sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:203-225 one-shot kernel, :127-176 barriers;
// host side python/sglang/srt/distributed/device_communicators/custom_all_reduce.py:146-510).
//
// Every rank owns ONE symmetric buffer (cudaMalloc + cudaIpc handle, opened by every peer):
//
//   CommHeader   epoch  (local)   number of collective calls this rank has completed
//                done   (local)   CTAs of the running call that have finished
//                flags[2][8][kCommMaxBlocks]  written by PEERS: flags[e&1][src][cta] = e once CTA
//                                 `cta` of rank `src` has pushed its part of call e
//   ll[2][world][slot_bytes]      low-latency area: 8-byte words {32-bit payload, 32-bit flag = e}
//   data[2][world][slot_bytes]    data[e&1][src] = partial of rank `src` for call e
//
// Two protocols share the epoch counter:
//  * LL (messages whose doubled size fits a slot): the flag travels WITH the data — every 8-byte
//    word a rank pushes carries the call's epoch in its upper half (8-byte stores are single-copy
//    atomic, also over NVLink), the receiver polls the word itself until the flag matches.  One
//    one-way NVLink trip, no fence, no barrier; costs 2x the wire bytes.  Flags always sit in the
//    odd 32-bit words of the LL area, whatever the layout of the call, and the area is never used
//    for raw data, so a stale word can never look valid.
//  * flag protocol (larger messages), below.
//
// Protocol of call e (= epoch + 1), identical grid on every rank, CTA c:
//   push    the CTA writes its part of the partial into data[e&1][my rank] of EVERY peer
//           (fire-and-forget stores over NVLink; no round trip);
//   signal  bar.sync, then one thread per peer: st.release.sys flags[e&1][my rank][c] = e;
//   wait    one thread per peer spins (ld.acquire.sys) until its local flags[e&1][peer][c] == e;
//   reduce  the CTA sums the `world` slots of its part in RANK ORDER (fp32), so every rank
//           produces bit-identical results, and writes the output;
//   end     the last CTA to finish publishes epoch = e (device memory, so CUDA-graph replays
//           need no host-side sequence number).
// Parity double-buffering is sufficient: a peer can only start call e+2 (which reuses the slots of
// call e) after it has completed call e+1, i.e. after it has seen MY flags of call e+1, which I
// send only after my call e has finished.  No deadlock without co-residency: CTAs are dispatched
// in index order, and the lowest unfinished CTA index of every rank never waits before it pushes.
#pragma once

#include "common.cuh"

namespace b200 {

constexpr int kCommMaxRanks = 8;
constexpr int kCommMaxBlocks = 1024;
constexpr uint64_t kCommTimeoutNs = 10ull * 1000 * 1000 * 1000;  // 10 s

struct CommHeader {
  uint32_t epoch;
  uint32_t done;
  uint32_t pad[62];
  uint32_t flags[2][kCommMaxRanks][kCommMaxBlocks];
};
constexpr size_t kCommDataOffset = (sizeof(CommHeader) + 1023) / 1024 * 1024;

// Passed BY VALUE to kernels: base of every rank's buffer mapped into this process.
struct CommDev {
  uint8_t* peer[kCommMaxRanks];
  int rank, world;
  uint64_t slot_bytes;  // bytes of one data[parity][src] slot
};

__device__ __forceinline__ CommHeader* comm_header(const CommDev& c, int r) {
  return reinterpret_cast<CommHeader*>(c.peer[r]);
}
// data[e&1][src] in the buffer of rank `dst` (raw area comes after the 2*world LL slots)
__device__ __forceinline__ uint8_t* comm_slot(const CommDev& c, int dst, int src, uint32_t e) {
  return c.peer[dst] + kCommDataOffset + ((2u + (e & 1u)) * c.world + src) * c.slot_bytes;
}
// ll[e&1][src] in the buffer of rank `dst`
__device__ __forceinline__ uint2* comm_ll_slot(const CommDev& c, int dst, int src, uint32_t e) {
  return reinterpret_cast<uint2*>(c.peer[dst] + kCommDataOffset +
                                  ((e & 1u) * c.world + src) * c.slot_bytes);
}
// Push two LL words {d0,e},{d1,e} with one 16-byte store (either 8-byte half is self-validating).
__device__ __forceinline__ void ll_store2(uint2* dst, uint32_t d0, uint32_t d1, uint32_t e) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};\n" ::"l"(dst), "r"(d0), "r"(e),
               "r"(d1), "r"(e)
               : "memory");
}
__device__ __forceinline__ void ll_store1(uint2* dst, uint32_t d0, uint32_t e) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};\n" ::"l"(dst), "r"(d0), "r"(e) : "memory");
}
__device__ __forceinline__ void ll_timeout_check(uint64_t& t0) {
  uint64_t t1;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t1));
  if (t0 == 0) t0 = t1;
  if (t1 - t0 > kCommTimeoutNs) __trap();
}
// Poll two adjacent LL words until both carry epoch e; returns their payloads.
__device__ __forceinline__ uint2 ll_load2(const uint2* src, uint32_t e) {
  uint32_t d0, f0, d1, f1;
  uint64_t t0 = 0;
  while (true) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];\n"
                 : "=r"(d0), "=r"(f0), "=r"(d1), "=r"(f1)
                 : "l"(src)
                 : "memory");
    if (f0 == e && f1 == e) break;
    ll_timeout_check(t0);
  }
  return make_uint2(d0, d1);
}
__device__ __forceinline__ uint32_t ll_load1(const uint2* src, uint32_t e) {
  uint32_t d0, f0;
  uint64_t t0 = 0;
  while (true) {
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];\n" : "=r"(d0), "=r"(f0) : "l"(src) : "memory");
    if (f0 == e) break;
    ll_timeout_check(t0);
  }
  return d0;
}
__device__ __forceinline__ uint32_t comm_begin(const CommDev& c) {
  return *reinterpret_cast<volatile uint32_t*>(&comm_header(c, c.rank)->epoch) + 1u;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Called by thread `t` (0 <= t < world) AFTER a barrier that orders the CTA's pushes before it.
__device__ __forceinline__ void comm_signal(const CommDev& c, uint32_t e, int cta, int t) {
  if (t != c.rank) st_release_sys(&comm_header(c, t)->flags[e & 1u][c.rank][cta], e);
}
// Called by thread `t` (0 <= t < world); a barrier must follow before the CTA reads the slots.
__device__ __forceinline__ void comm_wait(const CommDev& c, uint32_t e, int cta, int t) {
  if (t == c.rank) return;
  const uint32_t* f = &comm_header(c, c.rank)->flags[e & 1u][t][cta];
  if (ld_acquire_sys(f) == e) return;
  // Bounded spin: a peer that never arrives (crashed rank, mismatched call sequence) traps this
  // kernel after kCommTimeoutNs instead of hanging the GPU.
  uint64_t t0;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t0));
  while (ld_acquire_sys(f) != e) {
    __nanosleep(32);
    uint64_t t1;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t1));
    if (t1 - t0 > kCommTimeoutNs) __trap();
  }
}
// Called by ONE thread per CTA once the CTA has written its outputs.
__device__ __forceinline__ void comm_end(const CommDev& c, uint32_t e, uint32_t num_ctas) {
  CommHeader* h = comm_header(c, c.rank);
  __threadfence();
  if (atomicAdd(&h->done, 1u) == num_ctas - 1u) {
    h->done = 0u;
    __threadfence();
    *reinterpret_cast<volatile uint32_t*>(&h->epoch) = e;
  }
}

// Host-side communicator object behind the opaque handle of the C-ABI.
struct Comm {
  CommDev dev;
  size_t buf_bytes;
  // NVLS (comm_nvls.cu): this rank's slice of a multicast-bound allocation, mapped twice — uc = plain
  // local view, mc = the multicast address whose loads are reduced inside the NVSwitch
  uint8_t* nvls_uc = nullptr;
  uint8_t* nvls_mc = nullptr;
  size_t nvls_slot = 0;  // bytes of one of the two parity slots
};

// comm_nvls.cu: in-switch reduction path of b200_comm_all_reduce (returns -1 when it does not apply)
int comm_all_reduce_nvls(const Comm* c, const void* in, void* out, int64_t bytes, int dtype,
                         cudaStream_t st);

}  // namespace b200
