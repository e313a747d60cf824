// Shared device helpers for the pdt_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <utility>

#ifndef PDT_MAX_RANKS
#define PDT_MAX_RANKS 8
#endif

#define PDT_CHECK(expr)                                                       \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return (int)_e;                                    \
  } while (0)

// dtype codes used across the C ABI
#ifndef PDT_BF16
#define PDT_BF16 0
#define PDT_FP16 1
#define PDT_FP32 2
#endif

#ifndef PDT_PEERS_DEFINED
#define PDT_PEERS_DEFINED
typedef struct PdtPeers { void* p[PDT_MAX_RANKS]; } PdtPeers;   // by-value kernel argument: one pointer per rank
#endif

namespace pdt {

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
// memory hung the GPU.  The launchers ask this before they pick a cluster size > 1 for a kernel whose tensor
// map covers `p`; anything that is not known to live on the current device runs without multicast (each CTA
// fetches the tile itself).  Answers are cached per (device, pointer); a pointer first seen during stream
// capture is not queried and counts as "not local".
inline bool memory_on_current_device(const void* p, cudaStream_t stream) {
  static std::mutex mu;
  static std::unordered_map<uint64_t, bool> cache;
  int dev = -1;
  if (p == nullptr || cudaGetDevice(&dev) != cudaSuccess) return false;
  const uint64_t key = reinterpret_cast<uint64_t>(p) ^ ((uint64_t)(dev + 1) << 56);
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess || st != cudaStreamCaptureStatusNone) { cudaGetLastError(); return false; }
  cudaPointerAttributes attr;
  bool local = false;
  if (cudaPointerGetAttributes(&attr, p) == cudaSuccess) local = attr.type == cudaMemoryTypeDevice && attr.device == dev;
  else cudaGetLastError();
  cache[key] = local;
  return local;
}

// ---- programmatic dependent launch (PDL) -------------------------------------------------------------------
// A kernel launched through launch_pdl() may START while its predecessor in the stream is still draining: its
// CTAs are scheduled as the predecessor's retire and run their prologue (shared-memory carve-up, mbarrier init,
// TMEM allocation, tensor-map prefetch).  pdl_wait() then blocks until the predecessor has COMPLETED and its
// memory is visible, so every kernel calls it before its first access to global memory; pdl_trigger() lets the
// NEXT kernel's CTAs start the same way.  At 150 images per GPU a step is ~170 kernels of 10-40 us each, and
// the launch gap + tail + prologue between two of them is what this can hide.  Kernels launched without the
// attribute (torch's own, <<<>>>) see both calls as no-ops.  Measured on B200 inside the captured step: 4.094 ->
// 4.062 ms at 150 images (-0.8 %), within noise at 1200 images (the kernels are persistent one-wave grids, so
// there is little tail to overlap and graph replay already has ~1 us gaps).  Hence opt-in: PDT_PDL=1.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_wait(); pdl_trigger(); }

inline bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("PDT_PDL"); v = (e != nullptr && e[0] == '1') ? 1 : 0; }
  return v == 1;
}
inline int pdl_attr(cudaLaunchAttribute* a) {      // fills a[0]; returns the number of attributes to use (0 or 1)
  a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  a[0].val.programmaticStreamSerializationAllowed = 1;
  return pdl_enabled() ? 1 : 0;
}
template <typename... P, typename... A>
inline cudaError_t launch_pdl(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  cfg.numAttrs = pdl_attr(attr);
  cfg.attrs = attr;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<A>(args)...);
}

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Every cross-GPU or mbarrier wait is bounded: a dead peer or a protocol bug becomes a trap (an error the
// host sees) instead of a hung box.  Cross-GPU waits take their budget from Comm::timeout_ns (fabric.cuh,
// PDT_SPIN_TIMEOUT_S, default 10 minutes); intra-GPU mbarrier waits use a fixed 4 s (tc.cuh).

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <typename T> struct Cvt;
template <> struct Cvt<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static __device__ __forceinline__ float2 unpack2(uint32_t u) {
    __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&u);
    return __bfloat1622float2(h);
  }
};
template <> struct Cvt<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static __device__ __forceinline__ float2 unpack2(uint32_t u) {
    __half2 h = *reinterpret_cast<__half2*>(&u);
    return __half22float2(h);
  }
};
template <> struct Cvt<float> {
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float from_f(float v) { return v; }
};

// 8 low-precision values <-> 8 floats through one 16-byte access
template <typename T>
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = Cvt<T>::unpack2(u.x), b = Cvt<T>::unpack2(u.y), c = Cvt<T>::unpack2(u.z), d = Cvt<T>::unpack2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = Cvt<T>::pack2(f[0], f[1]); u.y = Cvt<T>::pack2(f[2], f[3]);
  u.z = Cvt<T>::pack2(f[4], f[5]); u.w = Cvt<T>::pack2(f[6], f[7]);
  return u;
}

__device__ __forceinline__ uint4 ldg_stream(const void* p) {   // read-once data: do not pollute L1
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// Last-block fold of per-block partial rows in fp64: tot[i] = sum_b partial[b * n2 + i], i < n2 (n2 % 4 == 0,
// rows 16-byte aligned).  Run by threads tid = 0..nthr-1 that meet at sync(); `tot` needs n2 + 4*nthr doubles.
// One thread owns FOUR adjacent columns (16-byte loads) and keeps 16 rows in flight: the rows were just
// written by other SMs, so every load is an L2 round trip, and the scalar one-row-at-a-time version of this
// loop cost 40-60 us per launch for 1024-column rows (more than the streaming part of small layers).
// Sum rows b = b0, b0 + step, ... < nb of column quad c4 into a[4].  The loads of a batch are issued into an
// array before any is consumed: left to itself ptxas reuses one register per row and the loop runs at one L2
// round trip per row (measured: 0.25 us per row, 93 us for 296 rows).
__device__ __forceinline__ void fold_quad(const float* __restrict__ partial, int nb, int n2, int c4, int b0, int step, double (&a)[4]) {
  // 16 rows (16-byte loads) in flight per thread: with 128 folding threads that is 32 KB outstanding, about what one
  // SM needs to cover the L2 round trip.  (At 8 in flight the fold of 116-148 rows x 1024 columns took 15-38 serial
  // round trips: ~20 us of a ~33 us layer-4 convolution at 150 images per GPU.)
  constexpr int kB = 16;
  for (int b = b0; b < nb; b += kB * step) {
    float4 v[kB];
#pragma unroll
    for (int u = 0; u < kB; ++u) {
      const int bb = b + u * step;
      v[u] = bb < nb ? __ldcg(reinterpret_cast<const float4*>(partial + (size_t)bb * n2) + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // the batch is summed in fp32 (fixed order) and enters the fp64 total once: an F2F + DADD per element ran at
    // ~21 cycles per warp instruction on B200 and dominated the fold
    float4 f = v[0];
#pragma unroll
    for (int u = 1; u < kB; ++u) { f.x += v[u].x; f.y += v[u].y; f.z += v[u].z; f.w += v[u].w; }
    a[0] += (double)f.x; a[1] += (double)f.y; a[2] += (double)f.z; a[3] += (double)f.w;
  }
}

template <class Sync>
__device__ __forceinline__ void fold_rows_f64(double* tot, const float* __restrict__ partial, int nb, int n2, int tid, int nthr,
                                              Sync sync) {
  const int nc4 = n2 >> 2;
  sync();
  if (nc4 >= nthr) {
    for (int c4 = tid; c4 < nc4; c4 += nthr) {
      double a[4] = {0.0, 0.0, 0.0, 0.0};
      fold_quad(partial, nb, n2, c4, 0, 1, a);
      tot[4 * c4 + 0] = a[0]; tot[4 * c4 + 1] = a[1]; tot[4 * c4 + 2] = a[2]; tot[4 * c4 + 3] = a[3];
    }
    sync();
  } else {
    const int rg = nthr / nc4, c4 = tid % nc4, g = tid / nc4;
    double a[4] = {0.0, 0.0, 0.0, 0.0};
    if (g < rg) fold_quad(partial, nb, n2, c4, g, rg, a);
    double* stage = tot + n2;                 // [rg][n2], rg * n2 <= 4 * nthr
    if (g < rg) {
      double* sp = stage + (size_t)g * n2 + 4 * c4;
      sp[0] = a[0]; sp[1] = a[1]; sp[2] = a[2]; sp[3] = a[3];
    }
    sync();
    for (int i = tid; i < n2; i += nthr) {
      double t = 0.0;
      for (int gg = 0; gg < rg; ++gg) t += stage[(size_t)gg * n2 + i];
      tot[i] = t;
    }
    sync();
  }
}

// ---- two-level fold ------------------------------------------------------------------------------------
// One SM reads other SMs' rows out of L2 at ~50 GB/s (measured: 148 rows x 2 KB in 6.5 us, 116 x 4 KB in 9.1 us,
// at the tail of EVERY statistics-producing kernel), so the last CTA must not read them all.  CTAs form groups of
// ~sqrt(nb) by block index; the last CTA to arrive in a group adds the group's rows (fp32, row order) into a group
// row, and the last group to finish adds the ~sqrt(nb) group rows in fp64: ~2 sqrt(nb) rows on the critical path
// instead of nb, still a fixed summation order (deterministic, independent of arrival order).
//   counters: word 0 = top-level ticket (the CALLER resets it when it is done), words 1.. = group tickets (reset
//             here); all zero between launches.  The scratch header has 64 words.
//   partial:  [nb] CTA rows, then room for the group rows: nb + groups <= max_rows, else one level.
// Every CTA calls this with all `nthr` threads after writing its row; true in exactly one CTA, whose `tot` then
// holds the column sums (fp64, needs n2 + 4 * nthr doubles like fold_rows_f64).
template <class Sync>
__device__ __forceinline__ bool tree_fold(double* tot, float* partial, unsigned* counters, int nb, int n2, int max_rows, int tid,
                                          int nthr, Sync sync, int* s_flag) {
  int G = 8;
  while (G * G < nb) ++G;
  const int groups = (nb + G - 1) / G;
  const bool two = nb > 24 && groups <= 62 && nb + groups <= max_rows;
  __threadfence();                                           // this CTA's row
  sync();
  if (!two) {
    if (tid == 0) *s_flag = (atomicAdd(counters, 1u) == (unsigned)nb - 1u);
    sync();
    if (!*s_flag) return false;
    __threadfence();
    fold_rows_f64(tot, partial, nb, n2, tid, nthr, sync);
    return true;
  }
  const int grp = (int)blockIdx.x / G, r0 = grp * G, members = min(G, nb - r0);
  if (tid == 0) {
    const bool last = atomicAdd(counters + 1 + grp, 1u) == (unsigned)members - 1u;
    if (last) counters[1 + grp] = 0;                         // nobody else touches it in this launch
    *s_flag = last;
  }
  sync();
  if (!*s_flag) return false;
  __threadfence();
  {
    const int nc4 = n2 >> 2;
    float* grow = partial + (size_t)(nb + grp) * n2;
    for (int c4 = tid; c4 < nc4; c4 += nthr) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int b = 0; b < members; b += 16) {
        float4 v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u)
          v[u] = b + u < members ? __ldcg(reinterpret_cast<const float4*>(partial + (size_t)(r0 + b + u) * n2) + c4)
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 16; ++u) { a.x += v[u].x; a.y += v[u].y; a.z += v[u].z; a.w += v[u].w; }
      }
      reinterpret_cast<float4*>(grow)[c4] = a;
    }
  }
  __threadfence();
  sync();
  if (tid == 0) *s_flag = (atomicAdd(counters, 1u) == (unsigned)groups - 1u);
  sync();
  if (!*s_flag) return false;
  __threadfence();
  fold_rows_f64(tot, partial + (size_t)nb * n2, groups, n2, tid, nthr, sync);
  return true;
}

// mean / invstd / running statistics of channels [0, K) from tot = {sums[K], sums of squares[K]} over `cnt` values,
// by threads tid < nthr of the finalising CTA.  The cancellation E[x^2] - mean^2 stays in fp64 (two multiplies and
// one FMA per channel; 1 / cnt is an fp32 reciprocal refined by one fp64 Newton step); square root, reciprocal
// and the running-statistics update are fp32, and the running statistics of a thread's channels are loaded before
// any is stored (fp64 divide / sqrt sequences and four dependent L2 round trips cost this CTA 2-4 us per launch).
__device__ __forceinline__ void bn_stats_finalize(const double* tot, int K, double cnt, float eps, float momentum,
                                                  float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                                  float* running_mean, float* running_var, int tid, int nthr) {
  double inv_n = (double)(1.0f / (float)cnt);
  inv_n = inv_n * (2.0 - cnt * inv_n);
  const float unbias = cnt > 1.0 ? (float)cnt / ((float)cnt - 1.f) : 1.f;
  for (int c0 = tid; c0 < K; c0 += 4 * nthr) {
    float rm[4], rv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = c0 + u * nthr;
      rm[u] = rv[u] = 0.f;
      if (running_mean != nullptr && c < K) { rm[u] = running_mean[c]; rv[u] = running_var[c]; }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = c0 + u * nthr;
      if (c < K) {
        const double mean = tot[c] * inv_n;
        const float var = fmaxf((float)fma(-mean, mean, tot[K + c] * inv_n), 0.f);
        save_mean[c] = (float)mean;
        save_invstd[c] = 1.0f / sqrtf(var + eps);
        if (running_mean != nullptr) {
          running_mean[c] = fmaf(momentum, (float)mean - rm[u], rm[u]);
          running_var[c] = fmaf(momentum, var * unbias - rv[u], rv[u]);
        }
      }
    }
  }
}

}  // namespace pdt
