// Device-side fabric primitives shared by comm.cu and bn.cu: flag pads, the
// block-level cross-GPU barrier and the scratch slots (layout in pdt_b200.h).
#pragma once
#include "common.cuh"
#include "../../include/pdt_b200.h"

namespace pdt {

constexpr int kMaxBlocks = PDT_COMM_MAX_BLOCKS;   // flag rows per channel
constexpr int kChannels = PDT_COMM_CHANNELS;

struct Comm {
  int rank, world;
  uint32_t* pad[PDT_MAX_RANKS];   // each rank's pad base (flags, then inf slots, then scratch)
};

__device__ __forceinline__ uint32_t* flag_addr(uint32_t* pad, int channel, int block, int src) {
  return pad + ((size_t)(channel * kMaxBlocks + block) * PDT_MAX_RANKS + src);
}
__device__ __forceinline__ uint32_t* inf_addr(uint32_t* pad, int block, int src) {
  return pad + PDT_PAD_INF_OFFSET_WORDS + (size_t)block * PDT_MAX_RANKS + src;
}
__device__ __forceinline__ float* scratch_addr(uint32_t* pad, int channel, int src) {
  return reinterpret_cast<float*>(pad + PDT_PAD_SCRATCH_OFFSET_WORDS) +
         ((size_t)channel * PDT_MAX_RANKS + src) * PDT_SCRATCH_FLOATS;
}

__device__ __forceinline__ void signal_flag(uint32_t* addr) {
  uint32_t old;
  const uint64_t t0 = globaltimer_ns();
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 0 && globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
  } while (old != 0);
}
__device__ __forceinline__ void wait_flag(uint32_t* addr) {
  uint32_t old;
  const uint64_t t0 = globaltimer_ns();
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 1 && globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
  } while (old != 1);
}

// All threads of the block call this.  Everything the block wrote before is
// visible to the peers' same-numbered block after they return, and vice versa.
// `row` selects the flag row; it must be the same number on every rank.
__device__ __forceinline__ void block_barrier(const Comm& c, int channel, int row) {
  __syncthreads();
  if ((int)threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    signal_flag(flag_addr(c.pad[peer], channel, row, c.rank));
    wait_flag(flag_addr(c.pad[c.rank], channel, row, peer));
  }
  __syncthreads();
}

__host__ inline Comm make_comm(const PdtCommDesc* d) {
  Comm c;
  c.rank = d ? d->rank : 0;
  c.world = d ? d->world : 1;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) c.pad[i] = d ? reinterpret_cast<uint32_t*>(d->pad[i]) : nullptr;
  return c;
}


}  // namespace pdt
