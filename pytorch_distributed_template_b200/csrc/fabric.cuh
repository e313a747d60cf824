// Device-side fabric primitives shared by comm.cu and bn.cu: flag pads, the
// block-level cross-GPU barrier and the scratch slots (layout in pdt_b200.h).
#pragma once
#include <cstdlib>
#include "common.cuh"
#include "../../include/pdt_b200.h"

namespace pdt {

constexpr int kMaxBlocks = PDT_COMM_MAX_BLOCKS;   // flag rows per channel
constexpr int kChannels = PDT_COMM_CHANNELS;

struct Comm {
  int rank, world;
  uint64_t timeout_ns;            // spin budget of every cross-GPU wait (PDT_SPIN_TIMEOUT_S, default 600 s)
  uint32_t* pad[PDT_MAX_RANKS];   // each rank's pad base (flags, then inf slots, then scratch)
  uint32_t* pad_mc;               // the same pad through the NVLS multicast mapping (one store reaches every rank), or null
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

// A rank may legitimately arrive late (rank 0 writing a checkpoint to a slow filesystem, a cold first batch),
// so the budget is minutes, not the few seconds that would do for an intra-step skew; a dead peer still turns
// into a trap (an error the host sees) instead of a hung box.
__device__ __forceinline__ void signal_flag(uint32_t* addr, uint64_t timeout_ns) {
  uint32_t old;
  const uint64_t t0 = globaltimer_ns();
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 0 && globaltimer_ns() - t0 > timeout_ns) __trap();
  } while (old != 0);
}
__device__ __forceinline__ void wait_flag(uint32_t* addr, uint64_t timeout_ns) {
  uint32_t old;
  const uint64_t t0 = globaltimer_ns();
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 1 && globaltimer_ns() - t0 > timeout_ns) __trap();
  } while (old != 1);
}

// ---- NVLS: one instruction on the multicast mapping of a symmetric buffer ------------------------------------
// ld_reduce: the switch reads the same offset on every rank and returns the sum (one NVLink transfer instead of W).
// st:        one store that the switch replicates into every rank's copy (W-fold less egress).
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void mc_st_b64(void* mc, uint2 v) {   // 8 bytes of any type (four 16-bit shadow weights)
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(mc), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)) : "memory");
}

// All threads of the block call this.  Everything the block wrote before is
// visible to the peers' same-numbered block after they return, and vice versa.
// `row` selects the flag row; it must be the same number on every rank.
__device__ __forceinline__ void block_barrier(const Comm& c, int channel, int row) {
  __syncthreads();
  if ((int)threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    signal_flag(flag_addr(c.pad[peer], channel, row, c.rank), c.timeout_ns);
    wait_flag(flag_addr(c.pad[c.rank], channel, row, peer), c.timeout_ns);
  }
  __syncthreads();
}

// The same for a SUBSET of the block's threads (the epilogue warps of a convolution kernel): `tid` numbers the
// participating threads from 0, `sync` synchronises exactly them (a named barrier).
template <class Sync>
__device__ __forceinline__ void block_barrier_t(const Comm& c, int channel, int row, int tid, Sync sync) {
  sync();
  if (tid < c.world) {
    signal_flag(flag_addr(c.pad[tid], channel, row, c.rank), c.timeout_ns);
    wait_flag(flag_addr(c.pad[c.rank], channel, row, tid), c.timeout_ns);
  }
  sync();
}

// Cross-rank sum of tot[0..n) (+ the count) through the peers' scratch slots; called by ONE block per rank (the
// one that finalises a statistics reduction) with `nthr` threads numbered `tid`.  With an NVLS mapping of the pad,
// this rank's vector reaches every peer's slot through ONE 16-byte multimem.st per four values (the switch
// replicates it) instead of `world` scalar peer stores per value.  Every rank then sums the slots in rank order,
// so all ranks compute bit-identical statistics on either data plane.  SyncBatchNorm's all_gather / all_reduce
// (torch/nn/modules/_functions.py:74-76, 159-164) as an in-kernel exchange.
template <class Sync>
__device__ __forceinline__ void exchange_sum_t(const Comm& c, double* tot, int n, double* count, int tid, int nthr, Sync sync) {
  // Channel 2 belongs to these exchanges alone (flags and scratch).  Two alternating slot sets make ONE cross-GPU
  // barrier per exchange enough: a rank can write exchange k+1 into the other set as soon as it has passed barrier
  // k, and it can only reach exchange k+2 (the same set again) after every peer has signalled barrier k+1, i.e.
  // after every peer has finished reading set k.  The parity comes from a per-rank call counter in this rank's own
  // pad (last word of its slot), which survives graph replays; all ranks run the same sequence of exchanges.
  constexpr int ch = 2;
  constexpr int kHalf = PDT_SCRATCH_FLOATS / 2;
  __shared__ int s_parity;
  if (tid == 0) {
    uint32_t* ctr = reinterpret_cast<uint32_t*>(scratch_addr(c.pad[c.rank], ch, c.rank)) + (PDT_SCRATCH_FLOATS - 1);
    const uint32_t k = *ctr;
    *ctr = k + 1;
    s_parity = (int)(k & 1u);
  }
  sync();
  const int off = s_parity * kHalf;
  if (c.pad_mc != nullptr) {
    float* slot = scratch_addr(c.pad_mc, ch, c.rank) + off;
    for (int i4 = tid * 4; i4 <= n; i4 += nthr * 4) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = i4 + j;
        v[j] = (i < n) ? (float)tot[i] : (i == n ? (float)*count : 0.f);
      }
      mc_st_f4(slot + i4, make_float4(v[0], v[1], v[2], v[3]));
    }
  } else {
    for (int i = tid; i <= n; i += nthr) {
      const float v = (i < n) ? (float)tot[i] : (float)*count;
      for (int q = 0; q < c.world; ++q) (scratch_addr(c.pad[q], ch, c.rank) + off)[i] = v;
    }
  }
  __threadfence_system();
  block_barrier_t(c, ch, 0, tid, sync);
  // (the per-rank values are fp32 already: summed in fp32 in rank order; only the count, which may pass 2^24, in fp64)
  for (int i = tid; i <= n; i += nthr) {
    if (i < n) {
      float a = 0.f;
      for (int p = 0; p < c.world; ++p) a += *reinterpret_cast<volatile float*>(scratch_addr(c.pad[c.rank], ch, p) + off + i);
      tot[i] = (double)a;
    } else {
      double a = 0.0;
      for (int p = 0; p < c.world; ++p) a += (double)*reinterpret_cast<volatile float*>(scratch_addr(c.pad[c.rank], ch, p) + off + i);
      *count = a;
    }
  }
  sync();
}

__host__ inline uint64_t spin_timeout_ns() {
  static const uint64_t ns = [] {
    const char* e = getenv("PDT_SPIN_TIMEOUT_S");
    double s = e ? atof(e) : 600.0;
    if (!(s > 0.0)) s = 600.0;
    return (uint64_t)(s * 1e9);
  }();
  return ns;
}

__host__ inline Comm make_comm(const PdtCommDesc* d) {
  Comm c;
  c.rank = d ? d->rank : 0;
  c.world = d ? d->world : 1;
  c.timeout_ns = spin_timeout_ns();
  for (int i = 0; i < PDT_MAX_RANKS; ++i) c.pad[i] = d ? reinterpret_cast<uint32_t*>(d->pad[i]) : nullptr;
  c.pad_mc = d ? reinterpret_cast<uint32_t*>(d->pad_mc) : nullptr;
  return c;
}


}  // namespace pdt
