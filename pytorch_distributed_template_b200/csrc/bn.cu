// BatchNorm / SyncBatchNorm training kernels on NHWC activations (SURVEY §2.2 N5, N7;
// §2.4 X7, X8).
//
// Forward  = k_bn_fwd_stats  (per-channel sum, sum of squares -> mean, invstd,
//                             running-stat update; with world > 1 the raw moments
//                             are exchanged with every peer INSIDE the kernel, over
//                             NVLink peer stores into the peers' scratch slots)
//          + k_bn_fwd_apply  (normalise, affine, optional residual add, optional ReLU).
// Backward = k_bn_bwd_reduce (ReLU mask, sum dy, sum dy*(x-mean); local dgamma/dbeta;
//                             cross-rank sum of the two vectors in-kernel)
//          + k_bn_bwd_apply  (dx and the residual branch's gradient).
//
// This replaces, for the reference's SyncBN mode (distributed_syncBN_amp.py:143-145),
// ATen batch_norm_stats + all_gather_into_tensor + batch_norm_gather_stats_with_counts
// + batch_norm_elemt and batch_norm_backward_reduce + all_reduce +
// batch_norm_backward_elemt (torch/nn/modules/_functions.py:39-209): 2 collectives, 1
// host sync and ~8 launches per layer become 0 / 0 / 2.  Equal per-rank counts make the
// raw-moment sum algebraically identical to the reference's (mean, invstd, count) merge.
//
// Reduction scheme: every block reduces a slab of rows to one row of `partial`
// (scratch, [blocks][2C] floats); the block that takes the last ticket of `counter`
// (scratch, one word, self-resetting) folds the rows in a fixed order in fp64, so the
// result is deterministic for a given grid.
#include <atomic>
#include <cstdlib>
#include <initializer_list>
#include "fabric.cuh"
#include "tc.cuh"

namespace {

using namespace pdt;

constexpr int kThreads = 512;
constexpr int kSmemFloats = 10240;  // 40 KB staging, reused as double[5120] in the finalise step

template <typename T> struct Vec { static constexpr int N = 16 / sizeof(T); };

template <typename T>
__device__ __forceinline__ void load_vec(const T* p, float (&f)[Vec<T>::N]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  if constexpr (sizeof(T) == 4) {
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
  } else {
    unpack8<T>(u, f);
  }
}
template <typename T>
__device__ __forceinline__ void decode_vec(const uint4& u, float (&f)[Vec<T>::N]) {
  if constexpr (sizeof(T) == 4) {
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
  } else {
    unpack8<T>(u, f);
  }
}
template <typename T>
__device__ __forceinline__ void store_vec(T* p, const float (&f)[Vec<T>::N]) {
  uint4 u;
  if constexpr (sizeof(T) == 4) {
    u.x = __float_as_uint(f[0]); u.y = __float_as_uint(f[1]); u.z = __float_as_uint(f[2]); u.w = __float_as_uint(f[3]);
  } else {
    u = pack8<T>(f);
  }
  *reinterpret_cast<uint4*>(p) = u;
}

// Block-level: fold per-thread accumulators (V channels x 2 quantities) into partial[blockIdx][2C].
template <int V>
__device__ __forceinline__ void block_fold(float* smem, const float (&a)[V], const float (&b)[V], int C, int cg, int rpi,
                                           float* __restrict__ partial) {
  const int tid = threadIdx.x, cgi = tid % cg, r = tid / cg, n2 = 2 * C;
  if (r < rpi) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      smem[r * n2 + cgi * V + j] = a[j];
      smem[r * n2 + C + cgi * V + j] = b[j];
    }
  }
  __syncthreads();
  for (int i = tid; i < n2; i += blockDim.x) {
    float s = 0.f;
    for (int rr = 0; rr < rpi; ++rr) s += smem[rr * n2 + i];
    partial[(size_t)blockIdx.x * n2 + i] = s;
  }
}

// Every block calls this after block_fold; true in exactly one block per launch, whose tot[0..n2) (fp64, shared
// memory) then holds the column sums of all blocks' rows (common.cuh: tree_fold).
__device__ __forceinline__ bool fold_partials(double* tot, float* partial, unsigned* counter, int n2) {
  __shared__ int s_last;
  return tree_fold(tot, partial, counter, (int)gridDim.x, n2, PDT_BN_MAX_BLOCKS, (int)threadIdx.x, (int)blockDim.x,
                   [] { __syncthreads(); }, &s_last);
}

// Cross-rank sum of tot[0..n) (+ the count): the whole block takes part (fabric.cuh: exchange_sum_t).
__device__ __forceinline__ void exchange_sum(const Comm& c, double* tot, int n, double* count) {
  exchange_sum_t(c, tot, n, count, (int)threadIdx.x, (int)blockDim.x, [] { __syncthreads(); });
}

// After block_fold: the last block folds every block's row, exchanges with the peers (SyncBN) and writes
// mean / invstd / running statistics.  `smem` >= 2C doubles + blockDim doubles.
__device__ __forceinline__ void finish_fwd_stats(float* smem, float* partial, unsigned* counter, const Comm& comm,
                                                 int C, long count, float* running_mean, float* running_var, float momentum,
                                                 float eps, float* __restrict__ save_mean, float* __restrict__ save_invstd) {
  __shared__ double s_count;
  const int tid = threadIdx.x;
  double* tot = reinterpret_cast<double*>(smem);
  if (!fold_partials(tot, partial, counter, 2 * C)) return;
  if (tid == 0) { s_count = (double)count; *counter = 0; }
  __syncthreads();
  if (comm.world > 1) exchange_sum(comm, tot, 2 * C, &s_count);
  __syncthreads();
  bn_stats_finalize(tot, C, s_count, eps, momentum, save_mean, save_invstd, running_mean, running_var, tid, (int)blockDim.x);
}

// Same for the backward reduction: local dgamma / dbeta, then the (optionally cross-rank) k1, k2 vectors.
// out_k: [2C] floats consumed by the dx pass: k1 = sum_dy / N, k2 = sum_dy_xmu * invstd^2 / N (N global)
__device__ __forceinline__ void finish_bwd_reduce(float* smem, float* partial, unsigned* counter, const Comm& comm,
                                                  int C, long count, const float* __restrict__ invstd, float* __restrict__ dgamma,
                                                  float* __restrict__ dbeta, float* __restrict__ out_k) {
  __shared__ double s_count;
  const int tid = threadIdx.x;
  double* tot = reinterpret_cast<double*>(smem);
  if (!fold_partials(tot, partial, counter, 2 * C)) return;
  if (tid == 0) { s_count = (double)count; *counter = 0; }
  __syncthreads();
  for (int c = tid; c < C; c += blockDim.x) {   // parameter grads stay local: DDP averages them
    dbeta[c] = (float)tot[c];
    dgamma[c] = (float)tot[C + c] * invstd[c];
  }
  __syncthreads();
  if (comm.world > 1) exchange_sum(comm, tot, 2 * C, &s_count);
  __syncthreads();
  double inv_n = (double)(1.0f / (float)s_count);
  inv_n = inv_n * (2.0 - s_count * inv_n);                  // fp32 reciprocal + one fp64 Newton step (no fp64 divide sequence)
  for (int c = tid; c < C; c += blockDim.x) {
    const float is = invstd[c];
    out_k[c] = (float)(tot[c] * inv_n);
    out_k[C + c] = (float)(tot[C + c] * inv_n) * is * is;
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) k_bn_fwd_stats(const T* __restrict__ x, long M, int C, float* __restrict__ partial,
                                                          unsigned* counter, Comm comm, float* running_mean,
                                                          float* running_var, float momentum, float eps,
                                                          float* __restrict__ save_mean, float* __restrict__ save_invstd) {
  constexpr int V = Vec<T>::N;
  __shared__ __align__(16) float smem[kSmemFloats];
  const int tid = threadIdx.x, cg = C / V, rpi = kThreads / cg, cgi = tid % cg, r = tid / cg;
  const long chunk = (M + gridDim.x - 1) / gridDim.x;
  const long row0 = (long)blockIdx.x * chunk, row1 = min(M, row0 + chunk);
  float s[V], q[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = 0.f;
  if (r < rpi) {
    // all loads of a batch of kBatch rows are issued before any is consumed: the kernel runs at one
    // 512-thread block per SM, so memory-level parallelism has to come from each thread
    constexpr int kBatch = 8;
    for (long row = row0 + r; row < row1; row += (long)kBatch * rpi) {
      uint4 u[kBatch];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const long rr = row + (long)b * rpi;
        u[b] = rr < row1 ? ldg_stream(x + rr * C + cgi * V) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        float f[V];
        decode_vec<T>(u[b], f);
#pragma unroll
        for (int j = 0; j < V; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
      }
    }
  }
  block_fold<V>(smem, s, q, C, cg, rpi, partial);
  finish_fwd_stats(smem, partial, counter, comm, C, M, running_mean, running_var, momentum, eps, save_mean, save_invstd);
}

// stat2 is invstd (training) or running_var (eval, eps applied here).
template <typename T>
__global__ void __launch_bounds__(kThreads) k_bn_fwd_apply(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                                                          long nvec, int C, const float* __restrict__ mean,
                                                          const float* __restrict__ stat2, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, int relu, int eval, float eps) {
  constexpr int V = Vec<T>::N;
  extern __shared__ float sm[];
  float* scale = sm;
  float* shift = sm + C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float inv = eval ? rsqrtf(stat2[c] + eps) : stat2[c];
    const float sc = gamma[c] * inv;
    scale[c] = sc;
    shift[c] = beta[c] - mean[c] * sc;
  }
  __syncthreads();
  const int cg = C / V;
  const bool pow2 = (cg & (cg - 1)) == 0;     // always true for the shapes the statistics kernels serve
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    const int c0 = (pow2 ? (int)(i & (cg - 1)) : (int)(i % cg)) * V;
    float f[V];
    load_vec<T>(x + i * V, f);
#pragma unroll
    for (int j = 0; j < V; ++j) f[j] = fmaf(f[j], scale[c0 + j], shift[c0 + j]);
    if (res) {
      float g[V];
      load_vec<T>(res + i * V, g);
#pragma unroll
      for (int j = 0; j < V; ++j) f[j] += g[j];
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < V; ++j) f[j] = fmaxf(f[j], 0.f);
    }
    store_vec<T>(y + i * V, f);
  }
}

// out_k: [2C] floats consumed by k_bn_bwd_apply: k1 = sum_dy / N, k2 = sum_dy_xmu * invstd^2 / N (N global)
template <typename T>
__global__ void __launch_bounds__(kThreads) k_bn_bwd_reduce(const T* __restrict__ dy, const T* __restrict__ y,
                                                           const T* __restrict__ x, long M, int C,
                                                           const float* __restrict__ mean, const float* __restrict__ invstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           int relu, float* __restrict__ partial, unsigned* counter, Comm comm,
                                                           float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                           float* __restrict__ out_k, long count) {
  constexpr int V = Vec<T>::N;
  __shared__ __align__(16) float smem[kSmemFloats];
  const int tid = threadIdx.x, cg = C / V, rpi = kThreads / cg, cgi = tid % cg, r = tid / cg;
  const long chunk = (M + gridDim.x - 1) / gridDim.x;
  const long row0 = (long)blockIdx.x * chunk, row1 = min(M, row0 + chunk);
  // `count` = elements per channel the BatchNorm normalised over on this rank; it differs from the M rows
  // summed here when the caller feeds a compacted gradient (pdt_bn_pool_bwd_reduce).
  // y == nullptr with relu: no residual entered the ReLU, so its mask is recomputed from x with the
  // forward's own scale/shift instead of reading the output tensor (one pass of traffic less)
  const bool remask = relu && y == nullptr;
  float s[V], q[V], mu[V], sc[V], sh[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    s[j] = q[j] = 0.f;
    const int c = cgi * V + j;
    mu[j] = (r < rpi) ? mean[c] : 0.f;
    sc[j] = (r < rpi && remask) ? gamma[c] * invstd[c] : 0.f;
    sh[j] = (r < rpi && remask) ? beta[c] - mean[c] * sc[j] : 0.f;
  }
  if (r < rpi) {
    constexpr int kBatch = 4;              // loads of 4 rows in flight per thread (see k_bn_fwd_stats)
    const bool need_y = relu && !remask;
    for (long row = row0 + r; row < row1; row += (long)kBatch * rpi) {
      uint4 ug[kBatch], ux[kBatch], uy[kBatch];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const long rr = row + (long)b * rpi;
        const bool ok = rr < row1;
        const long o = rr * C + cgi * V;
        ug[b] = ok ? ldg_stream(dy + o) : make_uint4(0, 0, 0, 0);
        ux[b] = ok ? ldg_stream(x + o) : make_uint4(0, 0, 0, 0);
        uy[b] = (ok && need_y) ? ldg_stream(y + o) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        float g[V], xv[V];
        decode_vec<T>(ug[b], g);
        decode_vec<T>(ux[b], xv);
        if (remask) {
#pragma unroll
          for (int j = 0; j < V; ++j) g[j] = fmaf(xv[j], sc[j], sh[j]) > 0.f ? g[j] : 0.f;
        } else if (relu) {
          float yv[V];
          decode_vec<T>(uy[b], yv);
#pragma unroll
          for (int j = 0; j < V; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < V; ++j) { s[j] += g[j]; q[j] = fmaf(g[j], xv[j] - mu[j], q[j]); }
      }
    }
  }
  block_fold<V>(smem, s, q, C, cg, rpi, partial);
  finish_bwd_reduce(smem, partial, counter, comm, C, count, invstd, dgamma, dbeta, out_k);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) k_bn_bwd_apply(const T* __restrict__ dy, const T* __restrict__ y,
                                                          const T* __restrict__ x, T* __restrict__ dx, T* __restrict__ dres,
                                                          long nvec, int C, const float* __restrict__ mean,
                                                          const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, const float* __restrict__ k, int relu) {
  constexpr int V = Vec<T>::N;
  extern __shared__ float sm[];
  float* a = sm;            // invstd*gamma (also the forward scale)
  float* k1 = sm + C;
  float* k2 = sm + 2 * C;
  float* mu = sm + 3 * C;
  float* sh = sm + 4 * C;   // forward shift
  const bool remask = relu && y == nullptr;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    a[c] = gamma[c] * invstd[c];
    k1[c] = k[c];
    k2[c] = k[C + c];
    mu[c] = mean[c];
    sh[c] = beta[c] - mean[c] * a[c];
  }
  __syncthreads();
  const int cg = C / V;
  const bool pow2 = (cg & (cg - 1)) == 0;     // always true for the shapes the statistics kernels serve
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    const int c0 = (pow2 ? (int)(i & (cg - 1)) : (int)(i % cg)) * V;
    float g[V], xv[V];
    load_vec<T>(dy + i * V, g);
    load_vec<T>(x + i * V, xv);
    if (remask) {
#pragma unroll
      for (int j = 0; j < V; ++j) g[j] = fmaf(xv[j], a[c0 + j], sh[c0 + j]) > 0.f ? g[j] : 0.f;
    } else if (relu) {
      float yv[V];
      load_vec<T>(y + i * V, yv);
#pragma unroll
      for (int j = 0; j < V; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    }
    if (dres) store_vec<T>(dres + i * V, g);
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] = (g[j] - k1[c0 + j] - (xv[j] - mu[c0 + j]) * k2[c0 + j]) * a[c0 + j];
    store_vec<T>(dx + i * V, g);
  }
}

// ---- bulk-copy streaming variants ------------------------------------------------------------------------
// The kernels above keep their loads in registers, so bytes in flight per SM = threads x 16-32 B, and
// register pressure caps the threads: ncu showed 3.3-4.3 TB/s of the 6.5 TB/s HBM3e delivers.  Here one
// producer warp streams 8 KB chunks of every input into a shared-memory ring with cp.async.bulk (the TMA
// engine's 1-D mode) and 8 consumer warps read them back with LDS.128: 96 KB in flight per CTA, two CTAs
// per SM, independent of what the arithmetic needs in registers.  A chunk is 512 sixteen-byte vectors;
// the channel group of vector i is i mod (C/V) and C/V divides 256, so each consumer thread keeps ONE
// channel group for the whole kernel and its per-channel constants / accumulators live in registers.
// Grid = min(2 x SMs, chunks), chunk c goes to CTA c mod grid: fixed order, deterministic sums.
namespace sp {
using namespace pdt::tc;
constexpr int kCons = 256;                       // consumer threads
constexpr int kThr = kCons + 32;                 // + the producer warp
constexpr int kU = 2;                            // vectors per consumer thread and chunk
constexpr int kChunkVecs = kCons * kU;
constexpr int kChunkBytes = kChunkVecs * 16;
constexpr int kRingBytes = 96 * 1024;
constexpr int kSmemBytes = kRingBytes + 256;
__host__ __device__ constexpr int stages_for(int nin) { return kRingBytes / (nin * kChunkBytes) > 8 ? 8 : kRingBytes / (nin * kChunkBytes); }

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr) : "memory");
  return r;
}

template <int NIN>
struct Pipe {
  static constexpr int S = stages_for(NIN);
  uint32_t ring;
  uint64_t* full;
  uint64_t* empty;
  int s;
  uint32_t ph;
  __device__ __forceinline__ void init(uint8_t* dsm) {
    ring = smem_u32(dsm);
    full = reinterpret_cast<uint64_t*>(dsm + kRingBytes);
    empty = full + 8;
    s = 0; ph = 0;
    if (threadIdx.x == 0) {
      for (int i = 0; i < S; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], kCons / 32); }
      fence_barrier_init();
    }
    __syncthreads();
    pdl_enter();       // the ring is set up; nothing of the previous kernel's output has been touched yet
  }
  __device__ __forceinline__ void advance() { if (++s == S) { s = 0; ph ^= 1; } }
  // whole producer warp
  // rev: walk the chunks from the END of the tensor.  The kernel that wrote (or last read) these tensors went
  // through them front to back, so what is still in the 126 MB L2 is their tail: start there.
  __device__ __forceinline__ void produce(const void* const (&src)[NIN], long nvec, bool rev = false) {
    const long nchunks = (nvec + kChunkVecs - 1) / kChunkVecs;
    for (long ci = blockIdx.x; ci < nchunks; ci += gridDim.x) {
      const long ch = rev ? nchunks - 1 - ci : ci;
      mbar_wait(&empty[s], ph ^ 1);
      if (elect_one_sync()) {
        const long v0 = ch * kChunkVecs;
        const long left = nvec - v0;
        const uint32_t bytes = (uint32_t)(left < kChunkVecs ? left : kChunkVecs) * 16u;
        mbar_expect_tx(&full[s], bytes * NIN);
#pragma unroll
        for (int j = 0; j < NIN; ++j)
          bulk_g2s(ring + (uint32_t)((s * NIN + j) * kChunkBytes), reinterpret_cast<const char*>(src[j]) + v0 * 16, bytes, &full[s]);
      }
      __syncwarp();
      advance();
    }
  }
  // consumer side: address of vector v of input j in the current stage
  __device__ __forceinline__ uint32_t at(int j, int v) const { return ring + (uint32_t)((s * NIN + j) * kChunkBytes + v * 16); }
  __device__ __forceinline__ void wait_full() { mbar_wait(&full[s], ph); }
  __device__ __forceinline__ void release() {     // whole consumer warp
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[s]);
    advance();
  }
};

template <typename T> __device__ __forceinline__ void dec(const uint4& u, float (&f)[Vec<T>::N]) { decode_vec<T>(u, f); }
template <typename T>
__device__ __forceinline__ uint4 enc(const float (&f)[Vec<T>::N]) {
  if constexpr (sizeof(T) == 4) return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  else return pack8<T>(f);
}

template <typename T>
__global__ void __launch_bounds__(kThr, 2) k_fwd_stats(const T* __restrict__ x, long M, int C, float* __restrict__ partial,
                                                      unsigned* counter, Comm comm, float* running_mean, float* running_var,
                                                      float momentum, float eps, float* __restrict__ save_mean,
                                                      float* __restrict__ save_invstd) {
  constexpr int V = Vec<T>::N;
  extern __shared__ __align__(128) uint8_t dsm[];
  Pipe<1> pipe;
  pipe.init(dsm);
  const long nvec = M * C / V;
  const int tid = threadIdx.x, cg = C / V;
  float s[V], q[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = 0.f;
  if (tid >= kCons) {
    const void* src[1] = {x};
    pipe.produce(src, nvec);
  } else {
    const long nchunks = (nvec + kChunkVecs - 1) / kChunkVecs;
    for (long ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
      pipe.wait_full();
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int v = tid + kCons * u;
        if (ch * kChunkVecs + v < nvec) {
          float f[V];
          dec<T>(lds128(pipe.at(0, v)), f);
#pragma unroll
          for (int j = 0; j < V; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
        }
      }
      pipe.release();
    }
  }
  __syncthreads();                                   // every chunk consumed: the ring is free for the fold
  float* smem = reinterpret_cast<float*>(dsm);
  block_fold<V>(smem, s, q, C, cg, kCons / cg, partial);
  finish_fwd_stats(smem, partial, counter, comm, C, M, running_mean, running_var, momentum, eps, save_mean, save_invstd);
}

template <typename T, bool HAS_RES>
__global__ void __launch_bounds__(kThr, 2) k_fwd_apply(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                                                      long nvec, int C, const float* __restrict__ mean,
                                                      const float* __restrict__ stat2, const float* __restrict__ gamma,
                                                      const float* __restrict__ beta, int relu, int eval, float eps) {
  constexpr int V = Vec<T>::N;
  constexpr int NIN = HAS_RES ? 2 : 1;
  extern __shared__ __align__(128) uint8_t dsm[];
  Pipe<NIN> pipe;
  pipe.init(dsm);
  const int tid = threadIdx.x, cg = C / V;
  if (tid >= kCons) {
    if constexpr (HAS_RES) { const void* src[2] = {x, res}; pipe.produce(src, nvec, true); }
    else { const void* src[1] = {x}; pipe.produce(src, nvec, true); }
    return;
  }
  float sc[V], sh[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const int c = (tid % cg) * V + j;
    const float inv = eval ? rsqrtf(stat2[c] + eps) : stat2[c];
    sc[j] = gamma[c] * inv;
    sh[j] = beta[c] - mean[c] * sc[j];
  }
  uint4* out = reinterpret_cast<uint4*>(y);
  const long nchunks = (nvec + kChunkVecs - 1) / kChunkVecs;
  for (long ci = blockIdx.x; ci < nchunks; ci += gridDim.x) {
    const long ch = nchunks - 1 - ci;            // the producer's order (see Pipe::produce)
    pipe.wait_full();
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int v = tid + kCons * u;
      const long gi = ch * kChunkVecs + v;
      if (gi < nvec) {
        float f[V];
        dec<T>(lds128(pipe.at(0, v)), f);
#pragma unroll
        for (int j = 0; j < V; ++j) f[j] = fmaf(f[j], sc[j], sh[j]);
        if constexpr (HAS_RES) {
          float r[V];
          dec<T>(lds128(pipe.at(1, v)), r);
#pragma unroll
          for (int j = 0; j < V; ++j) f[j] += r[j];
        }
        if (relu) {
#pragma unroll
          for (int j = 0; j < V; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        out[gi] = enc<T>(f);
      }
    }
    pipe.release();
  }
}

// HAS_Y: the ReLU mask comes from the forward output y (a residual entered the ReLU); otherwise, with relu,
// it is recomputed from x.
template <typename T, bool HAS_Y>
__global__ void __launch_bounds__(kThr, 2) k_bwd_reduce(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x,
                                                       long M, int C, const float* __restrict__ mean,
                                                       const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, int relu, float* __restrict__ partial,
                                                       unsigned* counter, Comm comm, float* __restrict__ dgamma,
                                                       float* __restrict__ dbeta, float* __restrict__ out_k, long count) {
  constexpr int V = Vec<T>::N;
  constexpr int NIN = HAS_Y ? 3 : 2;
  extern __shared__ __align__(128) uint8_t dsm[];
  Pipe<NIN> pipe;
  pipe.init(dsm);
  const long nvec = M * C / V;
  const int tid = threadIdx.x, cg = C / V;
  float s[V], q[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = 0.f;
  if (tid >= kCons) {
    if constexpr (HAS_Y) { const void* src[3] = {dy, x, y}; pipe.produce(src, nvec, true); }
    else { const void* src[2] = {dy, x}; pipe.produce(src, nvec, true); }
  } else {
    const bool remask = relu && !HAS_Y;
    float mu[V], sc[V], sh[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = (tid % cg) * V + j;
      mu[j] = mean[c];
      sc[j] = remask ? gamma[c] * invstd[c] : 0.f;
      sh[j] = remask ? beta[c] - mean[c] * sc[j] : 0.f;
    }
    const long nchunks = (nvec + kChunkVecs - 1) / kChunkVecs;
    for (long ci = blockIdx.x; ci < nchunks; ci += gridDim.x) {
      const long ch = nchunks - 1 - ci;          // the producer's order (see Pipe::produce)
      pipe.wait_full();
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int v = tid + kCons * u;
        if (ch * kChunkVecs + v < nvec) {
          float g[V], xv[V];
          dec<T>(lds128(pipe.at(0, v)), g);
          dec<T>(lds128(pipe.at(1, v)), xv);
          if constexpr (HAS_Y) {
            float yv[V];
            dec<T>(lds128(pipe.at(2, v)), yv);
#pragma unroll
            for (int j = 0; j < V; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
          } else if (remask) {
#pragma unroll
            for (int j = 0; j < V; ++j) g[j] = fmaf(xv[j], sc[j], sh[j]) > 0.f ? g[j] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < V; ++j) { s[j] += g[j]; q[j] = fmaf(g[j], xv[j] - mu[j], q[j]); }
        }
      }
      pipe.release();
    }
  }
  __syncthreads();
  float* smem = reinterpret_cast<float*>(dsm);
  block_fold<V>(smem, s, q, C, cg, kCons / cg, partial);
  finish_bwd_reduce(smem, partial, counter, comm, C, count, invstd, dgamma, dbeta, out_k);
}

// dx = a*g + b*x + c per channel with a = invstd*gamma, b = -k2*a, c = (mean*k2 - k1)*a (the expanded form of
// (g - k1 - (x - mean)*k2) * a: three constants and two FMAs per element).
template <typename T, bool HAS_Y, bool HAS_DRES>
__global__ void __launch_bounds__(kThr, 2) k_bwd_apply(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x,
                                                      T* __restrict__ dx, T* __restrict__ dres, long nvec, int C,
                                                      const float* __restrict__ mean, const float* __restrict__ invstd,
                                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                                      const float* __restrict__ k, int relu) {
  constexpr int V = Vec<T>::N;
  constexpr int NIN = HAS_Y ? 3 : 2;
  extern __shared__ __align__(128) uint8_t dsm[];
  Pipe<NIN> pipe;
  pipe.init(dsm);
  const int tid = threadIdx.x, cg = C / V;
  if (tid >= kCons) {
    if constexpr (HAS_Y) { const void* src[3] = {dy, x, y}; pipe.produce(src, nvec); }
    else { const void* src[2] = {dy, x}; pipe.produce(src, nvec); }
    return;
  }
  const bool remask = relu && !HAS_Y;
  float fa[V], fb[V], fc[V], fs[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const int c = (tid % cg) * V + j;
    const float a = gamma[c] * invstd[c];
    fa[j] = a;
    fb[j] = -k[C + c] * a;
    fc[j] = (mean[c] * k[C + c] - k[c]) * a;
    fs[j] = beta[c] - mean[c] * a;
  }
  uint4* odx = reinterpret_cast<uint4*>(dx);
  uint4* odr = reinterpret_cast<uint4*>(dres);
  const long nchunks = (nvec + kChunkVecs - 1) / kChunkVecs;
  for (long ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    pipe.wait_full();
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int v = tid + kCons * u;
      const long gi = ch * kChunkVecs + v;
      if (gi < nvec) {
        float g[V], xv[V];
        dec<T>(lds128(pipe.at(0, v)), g);
        dec<T>(lds128(pipe.at(1, v)), xv);
        if constexpr (HAS_Y) {
          float yv[V];
          dec<T>(lds128(pipe.at(2, v)), yv);
#pragma unroll
          for (int j = 0; j < V; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        } else if (remask) {
#pragma unroll
          for (int j = 0; j < V; ++j) g[j] = fmaf(xv[j], fa[j], fs[j]) > 0.f ? g[j] : 0.f;
        }
        if constexpr (HAS_DRES) odr[gi] = enc<T>(g);
#pragma unroll
        for (int j = 0; j < V; ++j) g[j] = fmaf(g[j], fa[j], fmaf(xv[j], fb[j], fc[j]));
        odx[gi] = enc<T>(g);
      }
    }
    pipe.release();
  }
}

// Shapes the streaming kernels take: every thread must keep one channel group (C/V divides 256) and the
// bulk copies need 16-byte aligned sources.
template <typename T> bool ok(int C, std::initializer_list<const void*> ptrs) {
  const int V = Vec<T>::N;
  if (C % V) return false;
  const int cg = C / V;
  if (cg < 1 || cg > kCons || (cg & (cg - 1)) != 0) return false;
  if (2 * C * (kCons / cg) * (int)sizeof(float) > kRingBytes || 2 * C * 8 + kThr * 8 > kRingBytes) return false;
  for (const void* p : ptrs) if (reinterpret_cast<uintptr_t>(p) & 15) return false;
  return true;
}
// per_sm: CTAs per SM.  The apply kernels run 2 (more warps to hide the LDS -> math -> STG chain); the reduction
// kernels run 1: every CTA costs them a partial row that the last CTA folds serially, and one 96 KB ring per
// SM already saturates HBM (measured 6.2 TB/s at 148 CTAs vs 5.9 at 296 on the 56x56 layers).
int grid_for(long nvec, int per_sm = 2) {
  static int sms = 0;
  if (sms == 0) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
  const long nchunks = (nvec + kChunkVecs - 1) / kChunkVecs;
  static int cap = -1;                       // PDT_BN_GRID_CAP: upper bound on the grid (experiments)
  if (cap < 0) { const char* e = getenv("PDT_BN_GRID_CAP"); cap = e ? atoi(e) : 0; }
  long g = (long)per_sm * sms;
  if (cap > 0 && g > cap) g = cap;
  // every CTA costs the reduction kernels one partial row to fold at the end: small tensors get fewer, fatter CTAs
  if (nchunks / 8 < g) g = nchunks / 8 > sms / 4 ? nchunks / 8 : sms / 4;
  if (g > PDT_BN_MAX_BLOCKS) g = PDT_BN_MAX_BLOCKS;
  return (int)(nchunks < g ? (nchunks < 1 ? 1 : nchunks) : g);
}
template <typename K> cudaError_t optin(K kern) {
  return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
}
}  // namespace sp

// ---- BatchNorm + ReLU + MaxPool fused (the ResNet stem tail) -----------------------------------------
// The stem's BN output is the largest activation of the network (112 x 112 x 64 per image) and exists
// only to be max-pooled.  These kernels never materialise it:
//   forward : pooled = maxpool(relu(bn(x))) straight from the conv output x, arg-max tap saved as int8;
//   backward: the gradient w.r.t. the (virtual) BN output at (h, w) is gathered from the <= 4 pooled
//             outputs whose arg-max is (h, w), masked by relu'(bn(x)) recomputed from x; the BN
//             reduction and the dx pass consume it on the fly.
// Replaces batch_norm + relu_ + max_pool2d_with_indices and their three backward kernels.
struct PoolGeom { int N, H, W, C, Ho, Wo, ks, st, pd; };

template <typename T> struct Pair2;
template <> struct Pair2<__nv_bfloat16> { using type = __nv_bfloat162; };
template <> struct Pair2<__half> { using type = __half2; };
template <typename T>
__device__ __forceinline__ uint32_t add_packed(uint32_t a, uint32_t b) {
  using P = typename Pair2<T>::type;
  P r = __hadd2(*reinterpret_cast<P*>(&a), *reinterpret_cast<P*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// Gradient of the (virtual) BN+ReLU output at input pixel (n, h, w): sum of the pooled gradients of the <= 4
// windows whose arg-max is this pixel.  Accumulated in the activation dtype with packed 2-wide adds, like the
// reference's max_pool2d backward (atomic adds in the tensor dtype).
template <typename T>
__device__ __forceinline__ void gather_pool_grad(const T* __restrict__ dpool, const int8_t* __restrict__ idx, const PoolGeom& g,
                                                 int n, int h, int w, int c0, float (&acc)[8]) {
  uint4 sum = make_uint4(0, 0, 0, 0);
  const int ho_lo = max(0, (h + g.pd - g.ks + g.st) / g.st), ho_hi = min(g.Ho - 1, (h + g.pd) / g.st);
  const int wo_lo = max(0, (w + g.pd - g.ks + g.st) / g.st), wo_hi = min(g.Wo - 1, (w + g.pd) / g.st);
  for (int ho = ho_lo; ho <= ho_hi; ++ho) {
    const int kh = h + g.pd - ho * g.st;
    for (int wo = wo_lo; wo <= wo_hi; ++wo) {
      const int tap = kh * g.ks + (w + g.pd - wo * g.st);
      const size_t o = (((size_t)n * g.Ho + ho) * g.Wo + wo) * g.C + c0;
      const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
      uint4 dv = *reinterpret_cast<const uint4*>(dpool + o);
      // byte-wise compare of the 8 saved taps, widened to 16-bit lane masks; a masked-out 16-bit
      // float becomes +0.0, so no per-channel branch is needed
      const unsigned tap4 = (unsigned)tap * 0x01010101u;
      const unsigned mlo = __vcmpeq4(pk.x, tap4), mhi = __vcmpeq4(pk.y, tap4);
      dv.x &= __byte_perm(mlo, 0, 0x1100);
      dv.y &= __byte_perm(mlo, 0, 0x3322);
      dv.z &= __byte_perm(mhi, 0, 0x1100);
      dv.w &= __byte_perm(mhi, 0, 0x3322);
      sum.x = add_packed<T>(sum.x, dv.x);
      sum.y = add_packed<T>(sum.y, dv.y);
      sum.z = add_packed<T>(sum.z, dv.z);
      sum.w = add_packed<T>(sum.w, dv.w);
    }
  }
  unpack8<T>(sum, acc);
}

template <typename T>
__global__ void __launch_bounds__(256) k_bn_pool_fwd(const T* __restrict__ x, T* __restrict__ y, int8_t* __restrict__ idx,
                                                    T* __restrict__ xmax, PoolGeom g,
                                                    const float* __restrict__ mean, const float* __restrict__ invstd,
                                                    const float* __restrict__ gamma, const float* __restrict__ beta) {
  extern __shared__ float sm[];
  float* scale = sm;
  float* shift = sm + g.C;
  for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
    const float sc = gamma[c] * invstd[c];
    scale[c] = sc;
    shift[c] = beta[c] - mean[c] * sc;
  }
  __syncthreads();
  const int cg = g.C / 8;
  const unsigned total = (unsigned)g.N * g.Ho * g.Wo * cg;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c0 = (int)(i % cg) * 8;
    unsigned r = i / cg;
    const int wo = (int)(r % g.Wo); r /= g.Wo;
    const int ho = (int)(r % g.Ho);
    const int n = (int)(r / g.Ho);
    float best[8], bx[8];
    int bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; bx[j] = 0.f; }
    for (int kh = 0; kh < g.ks; ++kh) {
      const int h = ho * g.st - g.pd + kh;
      if (h < 0 || h >= g.H) continue;
      for (int kw = 0; kw < g.ks; ++kw) {
        const int w = wo * g.st - g.pd + kw;
        if (w < 0 || w >= g.W) continue;
        float f[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(x + (((size_t)n * g.H + h) * g.W + w) * g.C + c0), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = fmaxf(fmaf(f[j], scale[c0 + j], shift[c0 + j]), 0.f);
          if (v > best[j]) { best[j] = v; bi[j] = kh * g.ks + kw; bx[j] = f[j]; }
        }
      }
    }
    const size_t o = (((size_t)n * g.Ho + ho) * g.Wo + wo) * g.C + c0;
    *reinterpret_cast<uint4*>(y + o) = pack8<T>(best);
    *reinterpret_cast<uint4*>(xmax + o) = pack8<T>(bx);
    uint2 pk;
    pk.x = (bi[0] & 0xff) | ((bi[1] & 0xff) << 8) | ((bi[2] & 0xff) << 16) | ((bi[3] & 0xff) << 24);
    pk.y = (bi[4] & 0xff) | ((bi[5] & 0xff) << 8) | ((bi[6] & 0xff) << 16) | ((bi[7] & 0xff) << 24);
    *reinterpret_cast<uint2*>(idx + o) = pk;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) k_bn_pool_bwd_apply(const T* __restrict__ dpool, const int8_t* __restrict__ idx,
                                                          const T* __restrict__ x, T* __restrict__ dx, PoolGeom g,
                                                          const float* __restrict__ mean, const float* __restrict__ invstd,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          const float* __restrict__ k) {
  extern __shared__ float sm[];
  const int C = g.C;
  float* a = sm;            // invstd*gamma  (also the forward scale)
  float* sh = sm + C;       // forward shift
  float* k1 = sm + 2 * C;
  float* k2 = sm + 3 * C;
  float* mu = sm + 4 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    a[c] = invstd[c] * gamma[c];
    sh[c] = beta[c] - mean[c] * a[c];
    k1[c] = k[c];
    k2[c] = k[C + c];
    mu[c] = mean[c];
  }
  __syncthreads();
  const int cg = C / 8;                       // power of two (checked by the launcher)
  const int cg_shift = 31 - __clz(cg);
  const int n = blockIdx.x / g.H, h = blockIdx.x - n * g.H;          // one image row per block
  const size_t row_base = ((size_t)n * g.H + h) * g.W * cg;
  for (int j = threadIdx.x; j < g.W * cg; j += blockDim.x) {
    const int w = j >> cg_shift, c0 = (j & (cg - 1)) * 8;
    const size_t i = row_base + j;
    float gy[8], xv[8];
    gather_pool_grad<T>(dpool, idx, g, n, h, w, c0, gy);
    unpack8<T>(*reinterpret_cast<const uint4*>(x + i * 8), xv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gg = fmaf(xv[j], a[c0 + j], sh[c0 + j]) > 0.f ? gy[j] : 0.f;
      gy[j] = (gg - k1[c0 + j] - (xv[j] - mu[c0 + j]) * k2[c0 + j]) * a[c0 + j];
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8<T>(gy);
  }
}

// ---- 3x3 / stride 2 / pad 1 specialisations (the torchvision stem; the generic kernels above keep every
// other geometry working).  Both are bandwidth kernels whose generic form was latency bound: runtime loop
// bounds kept ptxas from hoisting the loads, so each thread had ONE 16-byte load in flight (ncu: 1.8 TB/s).
// Here every load of a work item is issued before the first is consumed.

// One pooled output x 8 channels per thread: nine predicated loads in flight, then the scan.
template <typename T>
__global__ void __launch_bounds__(256) k_bn_pool_fwd_321(const T* __restrict__ x, T* __restrict__ y, int8_t* __restrict__ idx,
                                                        T* __restrict__ xmax, PoolGeom g, const float* __restrict__ mean,
                                                        const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta) {
  pdl_enter();
  extern __shared__ float sm[];
  float* scale = sm;
  float* shift = sm + g.C;
  for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
    const float sc = gamma[c] * invstd[c];
    scale[c] = sc;
    shift[c] = beta[c] - mean[c] * sc;
  }
  __syncthreads();
  const int cg = g.C / 8;
  const unsigned total = (unsigned)g.N * g.Ho * g.Wo * cg;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c0 = (int)(i % cg) * 8;
    unsigned r = i / cg;
    const int wo = (int)(r % g.Wo); r /= g.Wo;
    const int ho = (int)(r % g.Ho);
    const int n = (int)(r / g.Ho);
    const int h0 = 2 * ho - 1, w0 = 2 * wo - 1;
    uint4 u[9];
    bool ok[9];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int h = h0 + kh, w = w0 + kw;
        ok[kh * 3 + kw] = (unsigned)h < (unsigned)g.H && (unsigned)w < (unsigned)g.W;
        u[kh * 3 + kw] = ok[kh * 3 + kw] ? __ldg(reinterpret_cast<const uint4*>(x + (((size_t)n * g.H + h) * g.W + w) * g.C + c0))
                                         : make_uint4(0, 0, 0, 0);   // neighbours share pixels: keep them in L1
      }
    float sc[8], sh[8];
    bool any_zero = false;
#pragma unroll
    for (int j = 0; j < 8; ++j) { sc[j] = scale[c0 + j]; sh[j] = shift[c0 + j]; any_zero |= (sc[j] == 0.f); }
    const size_t o = (((size_t)n * g.Ho + ho) * g.Wo + wo) * g.C + c0;
    if (!any_zero) {
      // relu(sc*x + sh) is monotonic in x, so the window's arg-max is the arg-max of x (sc > 0) or of -x
      // (sc < 0): the scan runs on the raw 16-bit pairs with packed compare / max / select (4 instructions per
      // two elements and tap instead of 12 in fp32; the fp32 scan made this kernel issue-bound at 2.9 TB/s),
      // and BN + ReLU are applied once, to the winner.  Ties and all-negative windows may pick a different tap
      // than the fp32 scan only where the pooled value is identical.
      using P = typename Pair2<T>::type;
      P sgn[4], best[4];
      uint32_t bi[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t s2 = Cvt<T>::pack2(sc[2 * q] > 0.f ? 1.f : -1.f, sc[2 * q + 1] > 0.f ? 1.f : -1.f);
        const uint32_t ninf = Cvt<T>::pack2(-INFINITY, -INFINITY);
        sgn[q] = *reinterpret_cast<const P*>(&s2);
        best[q] = *reinterpret_cast<const P*>(&ninf);
        bi[q] = 0u;
      }
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        if (!ok[t]) continue;
        const uint32_t w4[4] = {u[t].x, u[t].y, u[t].z, u[t].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const P key = __hmul2(*reinterpret_cast<const P*>(&w4[q]), sgn[q]);
          const uint32_t gt = __hgt2_mask(key, best[q]);            // 0xffff per winning 16-bit lane
          best[q] = __hmax2(key, best[q]);
          bi[q] = (bi[q] & ~gt) | ((uint32_t)(t * 0x00010001u) & gt);
        }
      }
      float bx[8], by[8];
      uint4 xm;
      uint32_t* xmw = reinterpret_cast<uint32_t*>(&xm);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const P xw = __hmul2(best[q], sgn[q]);                      // the winner's raw value back
        xmw[q] = *reinterpret_cast<const uint32_t*>(&xw);
        const float2 f2 = Cvt<T>::unpack2(xmw[q]);
        bx[2 * q] = f2.x; bx[2 * q + 1] = f2.y;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) by[j] = fmaxf(fmaf(bx[j], sc[j], sh[j]), 0.f);
      *reinterpret_cast<uint4*>(y + o) = pack8<T>(by);
      *reinterpret_cast<uint4*>(xmax + o) = xm;
      uint2 pk;
      pk.x = __byte_perm(bi[0], bi[1], 0x6420);
      pk.y = __byte_perm(bi[2], bi[3], 0x6420);
      *reinterpret_cast<uint2*>(idx + o) = pk;
    } else {
      // a zero scale makes every tap tie: keep the element-wise scan with first-tap-wins semantics
      float best[8], bx[8];
      int bi[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; bx[j] = 0.f; }
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        if (!ok[t]) continue;
        float f[8];
        unpack8<T>(u[t], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = fmaxf(fmaf(f[j], sc[j], sh[j]), 0.f);
          if (v > best[j]) { best[j] = v; bi[j] = t; bx[j] = f[j]; }
        }
      }
      *reinterpret_cast<uint4*>(y + o) = pack8<T>(best);
      *reinterpret_cast<uint4*>(xmax + o) = pack8<T>(bx);
      uint2 pk;
      pk.x = (bi[0] & 0xff) | ((bi[1] & 0xff) << 8) | ((bi[2] & 0xff) << 16) | ((bi[3] & 0xff) << 24);
      pk.y = (bi[4] & 0xff) | ((bi[5] & 0xff) << 8) | ((bi[6] & 0xff) << 16) | ((bi[7] & 0xff) << 24);
      *reinterpret_cast<uint2*>(idx + o) = pk;
    }
  }
}

// dv with the channels whose saved arg-max is not `tap` zeroed (see gather_pool_grad).
__device__ __forceinline__ uint4 keep_tap(uint4 dv, uint2 pk, unsigned tap) {
  const unsigned tap4 = tap * 0x01010101u;
  const unsigned mlo = __vcmpeq4(pk.x, tap4), mhi = __vcmpeq4(pk.y, tap4);
  dv.x &= __byte_perm(mlo, 0, 0x1100);
  dv.y &= __byte_perm(mlo, 0, 0x3322);
  dv.z &= __byte_perm(mhi, 0, 0x1100);
  dv.w &= __byte_perm(mhi, 0, 0x3322);
  return dv;
}
template <typename T>
__device__ __forceinline__ uint4 add_packed4(uint4 a, uint4 b) {
  return make_uint4(add_packed<T>(a.x, b.x), add_packed<T>(a.y, b.y), add_packed<T>(a.z, b.z), add_packed<T>(a.w, b.w));
}

// A 2x2 block of input pixels (rows 2a, 2a+1; columns 2b, 2b+1) x 8 channels per thread.  Only the pooled
// outputs (a..a+1, b..b+1) can have their arg-max inside the block, so four (tap, gradient) loads serve four
// pixels (the per-pixel gather needed 2.25 on average, each behind a runtime loop), and with the four x
// loads that is 160 bytes in flight per thread.  Accumulation order per pixel is the generic kernel's.
template <typename T>
__global__ void __launch_bounds__(256) k_bn_pool_bwd_apply_321(const T* __restrict__ dpool, const int8_t* __restrict__ idx,
                                                              const T* __restrict__ x, T* __restrict__ dx, PoolGeom g,
                                                              const float* __restrict__ mean, const float* __restrict__ invstd,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              const float* __restrict__ k) {
  pdl_enter();
  extern __shared__ float sm[];
  const int C = g.C;
  // dx = a*gg + b*x + c per channel: a = invstd*gamma, b = -k2*a, c = (mu*k2 - k1)*a; sh = forward shift
  float* ca = sm;
  float* cb = sm + C;
  float* cc = sm + 2 * C;
  float* sh = sm + 3 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float a = invstd[c] * gamma[c];
    ca[c] = a;
    cb[c] = -k[C + c] * a;
    cc[c] = (mean[c] * k[C + c] - k[c]) * a;
    sh[c] = beta[c] - mean[c] * a;
  }
  __syncthreads();
  const int cg = C / 8, Hh = g.H / 2, Wh = g.W / 2;
  const unsigned total = (unsigned)g.N * Hh * Wh * cg;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c0 = (int)(i % cg) * 8;
    unsigned r = i / cg;
    const int b = (int)(r % Wh); r /= Wh;
    const int a = (int)(r % Hh);
    const int n = (int)(r / Hh);
    const bool a1 = a + 1 < g.Ho, b1 = b + 1 < g.Wo;
    const size_t p00 = (((size_t)n * g.Ho + a) * g.Wo + b) * C + c0;
    const size_t p01 = p00 + C, p10 = p00 + (size_t)g.Wo * C, p11 = p10 + C;
    const size_t x00 = (((size_t)n * g.H + 2 * a) * g.W + 2 * b) * C + c0;
    const size_t x01 = x00 + C, x10 = x00 + (size_t)g.W * C, x11 = x10 + C;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    const uint2 z2 = make_uint2(0xffffffffu, 0xffffffffu);           // tap 255 never matches
    const uint2 k00 = __ldg(reinterpret_cast<const uint2*>(idx + p00));
    const uint2 k01 = b1 ? __ldg(reinterpret_cast<const uint2*>(idx + p01)) : z2;
    const uint2 k10 = a1 ? __ldg(reinterpret_cast<const uint2*>(idx + p10)) : z2;
    const uint2 k11 = (a1 && b1) ? __ldg(reinterpret_cast<const uint2*>(idx + p11)) : z2;
    const uint4 d00 = __ldg(reinterpret_cast<const uint4*>(dpool + p00));   // each pooled output serves 4 blocks: L1
    const uint4 d01 = b1 ? __ldg(reinterpret_cast<const uint4*>(dpool + p01)) : z4;
    const uint4 d10 = a1 ? __ldg(reinterpret_cast<const uint4*>(dpool + p10)) : z4;
    const uint4 d11 = (a1 && b1) ? __ldg(reinterpret_cast<const uint4*>(dpool + p11)) : z4;
    uint4 xv[4];
    xv[0] = ldg_stream(x + x00); xv[1] = ldg_stream(x + x01); xv[2] = ldg_stream(x + x10); xv[3] = ldg_stream(x + x11);
    uint4 gsum[4];
    gsum[0] = keep_tap(d00, k00, 4);
    gsum[1] = add_packed4<T>(keep_tap(d00, k00, 5), keep_tap(d01, k01, 3));
    gsum[2] = add_packed4<T>(keep_tap(d00, k00, 7), keep_tap(d10, k10, 1));
    gsum[3] = add_packed4<T>(add_packed4<T>(add_packed4<T>(keep_tap(d00, k00, 8), keep_tap(d01, k01, 6)), keep_tap(d10, k10, 2)),
                             keep_tap(d11, k11, 0));
    float fa[8], fb[8], fc[8], fs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { fa[j] = ca[c0 + j]; fb[j] = cb[c0 + j]; fc[j] = cc[c0 + j]; fs[j] = sh[c0 + j]; }
    const size_t xo[4] = {x00, x01, x10, x11};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float gy[8], xf[8];
      unpack8<T>(gsum[q], gy);
      unpack8<T>(xv[q], xf);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gg = fmaf(xf[j], fa[j], fs[j]) > 0.f ? gy[j] : 0.f;
        gy[j] = fmaf(gg, fa[j], fmaf(xf[j], fb[j], fc[j]));
      }
      *reinterpret_cast<uint4*>(dx + xo[q]) = pack8<T>(gy);
    }
  }
}

// ---- the same two kernels for fp32 activations (the fp32 training modes): FOUR channels per thread, so every access
// is still one 16-byte vector and the same number of loads is in flight.  Without them the fp32 stem tail ran as
// BN-apply + max-pool forward (2.7 ms at 1200 images) and max-pool backward + BN reduce + BN apply (6.9 ms),
// i.e. it wrote and re-read the 112 x 112 x 64 normalised tensor and its gradient (3.85 GB each).
__global__ void __launch_bounds__(256) k_bn_pool_fwd_321_f32(const float* __restrict__ x, float* __restrict__ y, int8_t* __restrict__ idx,
                                                            float* __restrict__ xmax, PoolGeom g, const float* __restrict__ mean,
                                                            const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta) {
  pdl_enter();
  extern __shared__ float sm[];
  float* scale = sm;
  float* shift = sm + g.C;
  for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
    const float sc = gamma[c] * invstd[c];
    scale[c] = sc;
    shift[c] = beta[c] - mean[c] * sc;
  }
  __syncthreads();
  const int cg = g.C / 4;
  const unsigned total = (unsigned)g.N * g.Ho * g.Wo * cg;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c0 = (int)(i % cg) * 4;
    unsigned r = i / cg;
    const int wo = (int)(r % g.Wo); r /= g.Wo;
    const int ho = (int)(r % g.Ho);
    const int n = (int)(r / g.Ho);
    const int h0 = 2 * ho - 1, w0 = 2 * wo - 1;
    float4 u[9];
    bool ok[9];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int h = h0 + kh, w = w0 + kw;
        ok[kh * 3 + kw] = (unsigned)h < (unsigned)g.H && (unsigned)w < (unsigned)g.W;
        u[kh * 3 + kw] = ok[kh * 3 + kw] ? __ldg(reinterpret_cast<const float4*>(x + (((size_t)n * g.H + h) * g.W + w) * g.C + c0))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);   // neighbours share pixels: keep them in L1
      }
    float sc[4], sh[4];
    bool any_zero = false;
#pragma unroll
    for (int j = 0; j < 4; ++j) { sc[j] = scale[c0 + j]; sh[j] = shift[c0 + j]; any_zero |= (sc[j] == 0.f); }
    float best[4], bx[4], by[4];
    int bi[4];
    if (!any_zero) {
      // relu(sc*x + sh) is monotonic in x: the window's arg-max is the arg-max of x (sc > 0) or of -x (sc < 0)
      float sg[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) { sg[j] = sc[j] > 0.f ? 1.f : -1.f; best[j] = -INFINITY; bi[j] = 0; }
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        if (!ok[t]) continue;
        const float f[4] = {u[t].x, u[t].y, u[t].z, u[t].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float key = f[j] * sg[j];
          if (key > best[j]) { best[j] = key; bi[j] = t; }
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) { bx[j] = best[j] * sg[j]; by[j] = fmaxf(fmaf(bx[j], sc[j], sh[j]), 0.f); }
    } else {
      // a zero scale makes every tap tie: element-wise scan with first-tap-wins semantics
#pragma unroll
      for (int j = 0; j < 4; ++j) { by[j] = -INFINITY; bi[j] = 0; bx[j] = 0.f; }
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        if (!ok[t]) continue;
        const float f[4] = {u[t].x, u[t].y, u[t].z, u[t].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float v = fmaxf(fmaf(f[j], sc[j], sh[j]), 0.f);
          if (v > by[j]) { by[j] = v; bi[j] = t; bx[j] = f[j]; }
        }
      }
    }
    const size_t o = (((size_t)n * g.Ho + ho) * g.Wo + wo) * g.C + c0;
    *reinterpret_cast<float4*>(y + o) = make_float4(by[0], by[1], by[2], by[3]);
    *reinterpret_cast<float4*>(xmax + o) = make_float4(bx[0], bx[1], bx[2], bx[3]);
    *reinterpret_cast<uint32_t*>(idx + o) = (uint32_t)(bi[0] & 0xff) | ((uint32_t)(bi[1] & 0xff) << 8) | ((uint32_t)(bi[2] & 0xff) << 16) |
                                            ((uint32_t)(bi[3] & 0xff) << 24);
  }
}

// d with the channels whose saved arg-max (one byte each in k) is not `tap` zeroed.
__device__ __forceinline__ float4 keep_tap_f32(const float4& d, uint32_t k, uint32_t tap) {
  return make_float4((k & 0xffu) == tap ? d.x : 0.f, ((k >> 8) & 0xffu) == tap ? d.y : 0.f, ((k >> 16) & 0xffu) == tap ? d.z : 0.f,
                     (k >> 24) == tap ? d.w : 0.f);
}
__device__ __forceinline__ float4 add4(const float4& a, const float4& b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

__global__ void __launch_bounds__(256) k_bn_pool_bwd_apply_321_f32(const float* __restrict__ dpool, const int8_t* __restrict__ idx,
                                                                  const float* __restrict__ x, float* __restrict__ dx, PoolGeom g,
                                                                  const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                  const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                  const float* __restrict__ k) {
  pdl_enter();
  extern __shared__ float sm[];
  const int C = g.C;
  float* ca = sm;            // dx = a*gg + b*x + c per channel (see k_bn_pool_bwd_apply_321); sh = forward shift
  float* cb = sm + C;
  float* cc = sm + 2 * C;
  float* sh = sm + 3 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float a = invstd[c] * gamma[c];
    ca[c] = a;
    cb[c] = -k[C + c] * a;
    cc[c] = (mean[c] * k[C + c] - k[c]) * a;
    sh[c] = beta[c] - mean[c] * a;
  }
  __syncthreads();
  const int cg = C / 4, Hh = g.H / 2, Wh = g.W / 2;
  const unsigned total = (unsigned)g.N * Hh * Wh * cg;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c0 = (int)(i % cg) * 4;
    unsigned r = i / cg;
    const int b = (int)(r % Wh); r /= Wh;
    const int a = (int)(r % Hh);
    const int n = (int)(r / Hh);
    const bool a1 = a + 1 < g.Ho, b1 = b + 1 < g.Wo;
    const size_t p00 = (((size_t)n * g.Ho + a) * g.Wo + b) * C + c0;
    const size_t p01 = p00 + C, p10 = p00 + (size_t)g.Wo * C, p11 = p10 + C;
    const size_t x00 = (((size_t)n * g.H + 2 * a) * g.W + 2 * b) * C + c0;
    const size_t x01 = x00 + C, x10 = x00 + (size_t)g.W * C, x11 = x10 + C;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const uint32_t zk = 0xffffffffu;                                       // tap 255 never matches
    const uint32_t k00 = __ldg(reinterpret_cast<const uint32_t*>(idx + p00));
    const uint32_t k01 = b1 ? __ldg(reinterpret_cast<const uint32_t*>(idx + p01)) : zk;
    const uint32_t k10 = a1 ? __ldg(reinterpret_cast<const uint32_t*>(idx + p10)) : zk;
    const uint32_t k11 = (a1 && b1) ? __ldg(reinterpret_cast<const uint32_t*>(idx + p11)) : zk;
    const float4 d00 = __ldg(reinterpret_cast<const float4*>(dpool + p00));   // each pooled output serves 4 blocks: L1
    const float4 d01 = b1 ? __ldg(reinterpret_cast<const float4*>(dpool + p01)) : z4;
    const float4 d10 = a1 ? __ldg(reinterpret_cast<const float4*>(dpool + p10)) : z4;
    const float4 d11 = (a1 && b1) ? __ldg(reinterpret_cast<const float4*>(dpool + p11)) : z4;
    uint4 xr[4];
    xr[0] = ldg_stream(x + x00); xr[1] = ldg_stream(x + x01); xr[2] = ldg_stream(x + x10); xr[3] = ldg_stream(x + x11);
    float4 gsum[4];
    gsum[0] = keep_tap_f32(d00, k00, 4);
    gsum[1] = add4(keep_tap_f32(d00, k00, 5), keep_tap_f32(d01, k01, 3));
    gsum[2] = add4(keep_tap_f32(d00, k00, 7), keep_tap_f32(d10, k10, 1));
    gsum[3] = add4(add4(add4(keep_tap_f32(d00, k00, 8), keep_tap_f32(d01, k01, 6)), keep_tap_f32(d10, k10, 2)), keep_tap_f32(d11, k11, 0));
    float fa[4], fb[4], fc[4], fs[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { fa[j] = ca[c0 + j]; fb[j] = cb[c0 + j]; fc[j] = cc[c0 + j]; fs[j] = sh[c0 + j]; }
    const size_t xo[4] = {x00, x01, x10, x11};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float gy[4] = {gsum[q].x, gsum[q].y, gsum[q].z, gsum[q].w};
      const float xf[4] = {__uint_as_float(xr[q].x), __uint_as_float(xr[q].y), __uint_as_float(xr[q].z), __uint_as_float(xr[q].w)};
      float o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float gg = fmaf(xf[j], fa[j], fs[j]) > 0.f ? gy[j] : 0.f;
        o[j] = fmaf(gg, fa[j], fmaf(xf[j], fb[j], fc[j]));
      }
      *reinterpret_cast<float4*>(dx + xo[q]) = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
}

template <typename T> bool shape_ok(int C) {
  const int V = Vec<T>::N;
  if (C % V) return false;
  const int cg = C / V;
  return cg >= 1 && cg <= kThreads && (cg & (cg - 1)) == 0 && 2 * C <= PDT_SCRATCH_FLOATS - 1 && 2 * C <= kSmemFloats / 2 - kThreads;
}

int stats_blocks(long M, int C, int V) {
  const int rpi = kThreads / (C / V);
  long nb = M / ((long)rpi * 8);             // at least 8 rows per thread
  if (nb < 1) nb = 1;
  if (nb > PDT_BN_MAX_BLOCKS) nb = PDT_BN_MAX_BLOCKS;
  return (int)nb;
}

int apply_blocks(long nvec) {
  long nb = (nvec + kThreads * 4 - 1) / (kThreads * 4);
  if (nb < 1) nb = 1;
  if (nb > 148 * 8) nb = 148 * 8;
  return (int)nb;
}

template <typename T>
int bwd_reduce_stream(const void* dy, const void* y, const void* x, long M, int C, const float* mean, const float* invstd,
                      const float* gamma, const float* beta, int relu, float* partial, unsigned* counter, const PdtCommDesc* comm,
                      float* dgamma, float* dbeta, float* out_k, long count, cudaStream_t stream) {
  const int grid = sp::grid_for(M * C / Vec<T>::N, 1);
  if (relu && y != nullptr) {
    PDT_CHECK(sp::optin(sp::k_bwd_reduce<T, true>));
    launch_pdl(sp::k_bwd_reduce<T, true>, dim3(grid), dim3(sp::kThr), sp::kSmemBytes, stream,
        reinterpret_cast<const T*>(dy), reinterpret_cast<const T*>(y), reinterpret_cast<const T*>(x), M, C, mean, invstd, gamma,
        beta, relu, partial, counter, make_comm(comm), dgamma, dbeta, out_k, count);
  } else {
    PDT_CHECK(sp::optin(sp::k_bwd_reduce<T, false>));
    launch_pdl(sp::k_bwd_reduce<T, false>, dim3(grid), dim3(sp::kThr), sp::kSmemBytes, stream,
        reinterpret_cast<const T*>(dy), nullptr, reinterpret_cast<const T*>(x), M, C, mean, invstd, gamma, beta, relu, partial,
        counter, make_comm(comm), dgamma, dbeta, out_k, count);
  }
  return (int)cudaGetLastError();
}

template <typename T, bool HY, bool HD>
int bwd_apply_launch(const void* dy, const void* y, const void* x, void* dx, void* dres, long nvec, int C, const float* mean,
                     const float* invstd, const float* gamma, const float* beta, const float* k, int relu, cudaStream_t stream) {
  PDT_CHECK(sp::optin(sp::k_bwd_apply<T, HY, HD>));
  launch_pdl(sp::k_bwd_apply<T, HY, HD>, dim3(sp::grid_for(nvec)), dim3(sp::kThr), sp::kSmemBytes, stream,
      reinterpret_cast<const T*>(dy), reinterpret_cast<const T*>(y), reinterpret_cast<const T*>(x), reinterpret_cast<T*>(dx),
      reinterpret_cast<T*>(dres), nvec, C, mean, invstd, gamma, beta, k, relu);
  return (int)cudaGetLastError();
}
template <typename T>
int bwd_apply_stream(const void* dy, const void* y, const void* x, void* dx, void* dres, long nvec, int C, const float* mean,
                     const float* invstd, const float* gamma, const float* beta, const float* k, int relu, cudaStream_t stream) {
  const bool hy = relu && y != nullptr, hd = dres != nullptr;
  if (hy && hd) return bwd_apply_launch<T, true, true>(dy, y, x, dx, dres, nvec, C, mean, invstd, gamma, beta, k, relu, stream);
  if (hy) return bwd_apply_launch<T, true, false>(dy, y, x, dx, dres, nvec, C, mean, invstd, gamma, beta, k, relu, stream);
  if (hd) return bwd_apply_launch<T, false, true>(dy, y, x, dx, dres, nvec, C, mean, invstd, gamma, beta, k, relu, stream);
  return bwd_apply_launch<T, false, false>(dy, y, x, dx, dres, nvec, C, mean, invstd, gamma, beta, k, relu, stream);
}

// PDT_BN_STREAM=0 falls back to the register-staged kernels (A/B measurements, debugging).
bool stream_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("PDT_BN_STREAM"); v = (e != nullptr && e[0] == '0') ? 0 : 1; }
  return v == 1;
}

}  // namespace

extern "C" {

size_t pdt_bn_scratch_bytes(int C) { return (size_t)PDT_BN_MAX_BLOCKS * 2 * C * sizeof(float) + 256; }

#define BN_DISPATCH(dtype, ...)                                                      \
  if (dtype == PDT_BF16) { using T = __nv_bfloat16; __VA_ARGS__; }                   \
  else if (dtype == PDT_FP16) { using T = __half; __VA_ARGS__; }                     \
  else if (dtype == PDT_FP32) { using T = float; __VA_ARGS__; }                      \
  else return PDT_E_UNSUPPORTED;

#define BNP_DISPATCH(dtype, ...)                                      \
  if (dtype == PDT_BF16) { using T = __nv_bfloat16; __VA_ARGS__; }    \
  else if (dtype == PDT_FP16) { using T = __half; __VA_ARGS__; }      \
  else return PDT_E_UNSUPPORTED;

static bool pool_geom(PoolGeom& g, int N, int H, int W, int C, int ks, int st, int pd) {
  g.N = N; g.H = H; g.W = W; g.C = C; g.ks = ks; g.st = st; g.pd = pd;
  g.Ho = (H + 2 * pd - ks) / st + 1;
  g.Wo = (W + 2 * pd - ks) / st + 1;
  return C % 8 == 0 && ks * ks <= 127 && g.Ho > 0 && g.Wo > 0 && (double)N * H * W * (C / 8) < 4.0e9;
}

int pdt_bn_pool_fwd(const void* x, void* y, int8_t* idx, void* xmax, int dtype, int N, int H, int W, int C, int ks, int st,
                    int pd, const float* mean, const float* invstd, const float* gamma, const float* beta, cudaStream_t stream) {
  PoolGeom g;
  if (!pool_geom(g, N, H, W, C, ks, st, pd)) return PDT_E_UNSUPPORTED;
  if (dtype == PDT_FP32) {         // fp32: the 3x3 / 2 / 1 window (the ResNet stem) only, four channels per thread
    if (!(ks == 3 && st == 2 && pd == 1) || (reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) |
                                             reinterpret_cast<uintptr_t>(xmax)) % 16 != 0) return PDT_E_UNSUPPORTED;
    long nb4 = ((long)N * g.Ho * g.Wo * (C / 4) + 255) / 256;
    if (nb4 > 148 * 16) nb4 = 148 * 16;
    launch_pdl(k_bn_pool_fwd_321_f32, dim3((unsigned)nb4), dim3(256), 2 * C * sizeof(float), stream, reinterpret_cast<const float*>(x),
               reinterpret_cast<float*>(y), idx, reinterpret_cast<float*>(xmax), g, mean, invstd, gamma, beta);
    return (int)cudaGetLastError();
  }
  const long total = (long)N * g.Ho * g.Wo * (C / 8);
  long nb = (total + 255) / 256;
  if (nb > 148 * 16) nb = 148 * 16;
  if (ks == 3 && st == 2 && pd == 1) {
    BNP_DISPATCH(dtype, launch_pdl(k_bn_pool_fwd_321<T>, dim3((unsigned)nb), dim3(256), 2 * C * sizeof(float), stream,
        reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), idx, reinterpret_cast<T*>(xmax), g, mean, invstd, gamma, beta));
  } else {
    BNP_DISPATCH(dtype, k_bn_pool_fwd<T><<<(int)nb, 256, 2 * C * sizeof(float), stream>>>(
        reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), idx, reinterpret_cast<T*>(xmax), g, mean, invstd, gamma, beta));
  }
  return (int)cudaGetLastError();
}

// Reduction half of the backward, taken from the POOLED side: with g = dpool where ypool > 0 (the ReLU mask
// of the arg-max element is exactly "the pooled value is positive"),
//   sum over input pixels of dy            = sum over pooled outputs of g
//   sum over input pixels of dy * (x - mu) = sum over pooled outputs of g * (x[arg-max] - mu)
// because every pooled gradient lands on exactly one input pixel.  x[arg-max] is what the forward saved as
// `xmax`, so this pass reads two pooled-size tensors + xmax instead of gathering over the 4x larger input.
int pdt_bn_pool_bwd_reduce(const void* dpool, const void* ypool, const void* xmax, int dtype, int N, int H, int W, int C,
                           int ks, int st, int pd, const float* mean, const float* invstd, const float* gamma,
                           const float* beta, void* scratch, const PdtCommDesc* comm, float* dgamma, float* dbeta,
                           float* out_k, cudaStream_t stream) {
  PoolGeom g;
  if (!pool_geom(g, N, H, W, C, ks, st, pd)) return PDT_E_UNSUPPORTED;
  unsigned* counter = reinterpret_cast<unsigned*>(scratch);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  const long Mp = (long)N * g.Ho * g.Wo;
  BN_DISPATCH(dtype, {
    if (!shape_ok<T>(C)) return PDT_E_UNSUPPORTED;
    if (stream_enabled() && sp::ok<T>(C, {dpool, ypool, xmax})) return bwd_reduce_stream<T>(dpool, ypool, xmax, Mp, C, mean, invstd, gamma, beta, 1,
                                                                                            partial, counter, comm, dgamma, dbeta, out_k,
                                                                                            (long)N * H * W, stream);
    k_bn_bwd_reduce<T><<<stats_blocks(Mp, C, Vec<T>::N), kThreads, 0, stream>>>(
        reinterpret_cast<const T*>(dpool), reinterpret_cast<const T*>(ypool), reinterpret_cast<const T*>(xmax), Mp, C, mean,
        invstd, gamma, beta, 1, partial, counter, make_comm(comm), dgamma, dbeta, out_k, (long)N * H * W);
  });
  return (int)cudaGetLastError();
}

int pdt_bn_pool_bwd_apply(const void* dpool, const int8_t* idx, const void* x, void* dx, int dtype, int N, int H, int W,
                          int C, int ks, int st, int pd, const float* mean, const float* invstd, const float* gamma,
                          const float* beta, const float* k, cudaStream_t stream) {
  PoolGeom g;
  if (!pool_geom(g, N, H, W, C, ks, st, pd)) return PDT_E_UNSUPPORTED;
  if (((C / 8) & (C / 8 - 1)) != 0) return PDT_E_UNSUPPORTED;
  if (dtype == PDT_FP32) {
    if (!(ks == 3 && st == 2 && pd == 1 && H % 2 == 0 && W % 2 == 0) ||
        (reinterpret_cast<uintptr_t>(dpool) | reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx)) % 16 != 0) return PDT_E_UNSUPPORTED;
    long nb4 = ((long)N * (H / 2) * (W / 2) * (C / 4) + 255) / 256;
    if (nb4 > 148 * 32) nb4 = 148 * 32;
    launch_pdl(k_bn_pool_bwd_apply_321_f32, dim3((unsigned)nb4), dim3(256), 4 * C * sizeof(float), stream,
               reinterpret_cast<const float*>(dpool), idx, reinterpret_cast<const float*>(x), reinterpret_cast<float*>(dx), g, mean, invstd,
               gamma, beta, k);
    return (int)cudaGetLastError();
  }
  if (ks == 3 && st == 2 && pd == 1 && H % 2 == 0 && W % 2 == 0) {
    const long total = (long)N * (H / 2) * (W / 2) * (C / 8);
    long nb = (total + 255) / 256;
    if (nb > 148 * 32) nb = 148 * 32;
    BNP_DISPATCH(dtype, launch_pdl(k_bn_pool_bwd_apply_321<T>, dim3((unsigned)nb), dim3(256), 4 * C * sizeof(float), stream,
        reinterpret_cast<const T*>(dpool), idx, reinterpret_cast<const T*>(x), reinterpret_cast<T*>(dx), g, mean, invstd,
        gamma, beta, k));
    return (int)cudaGetLastError();
  }
  const long nb = (long)N * H;                 // one block per image row
  BNP_DISPATCH(dtype, k_bn_pool_bwd_apply<T><<<(unsigned)nb, 256, 5 * C * sizeof(float), stream>>>(
      reinterpret_cast<const T*>(dpool), idx, reinterpret_cast<const T*>(x), reinterpret_cast<T*>(dx), g, mean, invstd,
      gamma, beta, k));
  return (int)cudaGetLastError();
}


int pdt_bn_fwd_stats(const void* x, long M, int C, int dtype, void* scratch, const PdtCommDesc* comm,
                     float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                     float* save_invstd, cudaStream_t stream) {
  unsigned* counter = reinterpret_cast<unsigned*>(scratch);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  BN_DISPATCH(dtype, {
    if (!shape_ok<T>(C)) return PDT_E_UNSUPPORTED;
    if (stream_enabled() && sp::ok<T>(C, {x})) {
      PDT_CHECK(sp::optin(sp::k_fwd_stats<T>));
      launch_pdl(sp::k_fwd_stats<T>, dim3(sp::grid_for(M * C / Vec<T>::N, 1)), dim3(sp::kThr), sp::kSmemBytes, stream,
          reinterpret_cast<const T*>(x), M, C, partial, counter, make_comm(comm), running_mean, running_var, momentum, eps,
          save_mean, save_invstd);
      return (int)cudaGetLastError();
    }
    k_bn_fwd_stats<T><<<stats_blocks(M, C, Vec<T>::N), kThreads, 0, stream>>>(
        reinterpret_cast<const T*>(x), M, C, partial, counter, make_comm(comm), running_mean, running_var, momentum, eps,
        save_mean, save_invstd);
  });
  return (int)cudaGetLastError();
}

// ---- allocation-free variant of pdt_bn_fwd_stats ---------------------------------------------------
// Some callers cannot hand us a scratch buffer because they cannot allocate one at the point of the
// call: code running inside a CUDA-graph-replayed region or a real-time inference/training loop that
// forbids cudaMalloc (it synchronises the device), plugins living under a foreign framework's memory
// pool that must not be grown, and plain C callers with no allocator of their own.  For them the
// scratch (ticket counter + per-block partial sums) lives in this file-scope device array instead.
//   * fixed size: channels <= PDT_BN_STATIC_MAX_C; larger C is refused with PDT_E_TOO_LARGE;
//   * every GPU that loads the library holds these ~2.3 MiB until the process exits;
//   * the array is SHARED by all calls on a device and keeps the last call's contents: the words that
//     must start at zero (the ticket counter block) are cleared on the caller's stream before every
//     launch, and two calls whose kernels can overlap on one GPU (different streams) race;
//   * the first call on a device resolves the array's device address (cudaGetSymbolAddress) — make
//     that first call outside any stream capture.
// Nothing here calls cudaMalloc, cudaFree or any allocating routine.
static __device__ __align__(256) unsigned char g_bn_static_scratch[PDT_BN_STATIC_SCRATCH_BYTES];

// Per device; filled once, with that device current.  The pointer is published with a release store and read
// with an acquire load, so two host threads making the first call on one device at the same time both end up
// with the same, fully written value (each may do the look-up; the result is identical).
static std::atomic<void*> g_bn_static_ptr[64];

static int bn_static_scratch_for_current_device(void** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  void* p = g_bn_static_ptr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    // In host code the array's name is only a handle for the runtime, not a device address.
    e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return (int)cudaErrorInvalidSymbol;
    g_bn_static_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

int pdt_bn_fwd_stats_noalloc(const void* x, long M, int C, int dtype, float* running_mean, float* running_var,
                             float momentum, float eps, float* save_mean, float* save_invstd, cudaStream_t stream) {
  if (C > PDT_BN_STATIC_MAX_C) return PDT_E_TOO_LARGE;
  if (pdt_bn_scratch_bytes(C) > (size_t)PDT_BN_STATIC_SCRATCH_BYTES) return PDT_E_TOO_LARGE;
  void* scratch = nullptr;
  int rc = bn_static_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  // the ticket counter must read zero when the kernel starts, whatever an earlier call left behind
  cudaError_t e = cudaMemsetAsync(scratch, 0, 256, stream);
  if (e != cudaSuccess) return (int)e;
  return pdt_bn_fwd_stats(x, M, C, dtype, scratch, nullptr, running_mean, running_var, momentum, eps, save_mean,
                          save_invstd, stream);
}

int pdt_bn_fwd_apply(const void* x, const void* res, void* y, long M, int C, int dtype, const float* mean,
                     const float* stat2, const float* gamma, const float* beta, int relu, int eval, float eps,
                     cudaStream_t stream) {
  BN_DISPATCH(dtype, {
    if (C % Vec<T>::N) return PDT_E_UNSUPPORTED;
    const long nvec = M * C / Vec<T>::N;
    if (stream_enabled() && sp::ok<T>(C, {x, res, y})) {
      if (res != nullptr) {
        PDT_CHECK(sp::optin(sp::k_fwd_apply<T, true>));
        launch_pdl(sp::k_fwd_apply<T, true>, dim3(sp::grid_for(nvec)), dim3(sp::kThr), sp::kSmemBytes, stream,
            reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(res), reinterpret_cast<T*>(y), nvec, C, mean, stat2,
            gamma, beta, relu, eval, eps);
      } else {
        PDT_CHECK(sp::optin(sp::k_fwd_apply<T, false>));
        launch_pdl(sp::k_fwd_apply<T, false>, dim3(sp::grid_for(nvec)), dim3(sp::kThr), sp::kSmemBytes, stream,
            reinterpret_cast<const T*>(x), nullptr, reinterpret_cast<T*>(y), nvec, C, mean, stat2, gamma, beta, relu, eval, eps);
      }
      return (int)cudaGetLastError();
    }
    k_bn_fwd_apply<T><<<apply_blocks(nvec), kThreads, 2 * C * sizeof(float), stream>>>(
        reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(res), reinterpret_cast<T*>(y), nvec, C, mean, stat2,
        gamma, beta, relu, eval, eps);
  });
  return (int)cudaGetLastError();
}

int pdt_bn_bwd_reduce(const void* dy, const void* y, const void* x, long M, int C, int dtype, const float* mean,
                      const float* invstd, const float* gamma, const float* beta, int relu, void* scratch,
                      const PdtCommDesc* comm, float* dgamma, float* dbeta, float* out_k, cudaStream_t stream) {
  unsigned* counter = reinterpret_cast<unsigned*>(scratch);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  BN_DISPATCH(dtype, {
    if (!shape_ok<T>(C)) return PDT_E_UNSUPPORTED;
    if (stream_enabled() && sp::ok<T>(C, {dy, y, x})) return bwd_reduce_stream<T>(dy, y, x, M, C, mean, invstd, gamma, beta, relu,
                                                                                  partial, counter, comm, dgamma, dbeta, out_k, M, stream);
    k_bn_bwd_reduce<T><<<stats_blocks(M, C, Vec<T>::N), kThreads, 0, stream>>>(
        reinterpret_cast<const T*>(dy), reinterpret_cast<const T*>(y), reinterpret_cast<const T*>(x), M, C, mean, invstd,
        gamma, beta, relu, partial, counter, make_comm(comm), dgamma, dbeta, out_k, M);
  });
  return (int)cudaGetLastError();
}

int pdt_bn_bwd_apply(const void* dy, const void* y, const void* x, void* dx, void* dres, long M, int C, int dtype,
                     const float* mean, const float* invstd, const float* gamma, const float* beta, const float* k,
                     int relu, cudaStream_t stream) {
  BN_DISPATCH(dtype, {
    if (C % Vec<T>::N) return PDT_E_UNSUPPORTED;
    const long nvec = M * C / Vec<T>::N;
    if (stream_enabled() && sp::ok<T>(C, {dy, y, x, dx, dres}))
      return bwd_apply_stream<T>(dy, y, x, dx, dres, nvec, C, mean, invstd, gamma, beta, k, relu, stream);
    k_bn_bwd_apply<T><<<apply_blocks(nvec), kThreads, 5 * C * sizeof(float), stream>>>(
        reinterpret_cast<const T*>(dy), reinterpret_cast<const T*>(y), reinterpret_cast<const T*>(x),
        reinterpret_cast<T*>(dx), reinterpret_cast<T*>(dres), nvec, C, mean, invstd, gamma, beta, k, relu);
  });
  return (int)cudaGetLastError();
}

}  // extern "C"
