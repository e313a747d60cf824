// Cross-GPU kernels over NVLink 5 / NVSwitch peer memory.
//
// Every rank maps every peer's communication arena (CUDA VMM / torch symmetric
// memory) and these kernels move the data themselves; there is no NCCL call,
// proxy thread or host round-trip on this path, so the whole training step stays
// CUDA-graph capturable.  Two data planes, chosen per launch:
//   NVLS  multimem.ld_reduce / multimem.st on the multicast mapping of the arena:
//         the switch sums the W copies (one NVLink transfer per 16 bytes instead of
//         W) or replicates one store into all W copies;
//   P2P   plain ld/st on the W mapped peer pointers (any world size 1..8, single-
//         process DataParallel, and the fallback when no multicast object exists).
//
//   k_reduce_scatter   gradient segment: each rank sums ITS shard over all
//                      peers' gradient arenas, folds in 1/world and the AMP
//                      1/loss_scale, and raises found_inf on a non-finite sum
//                      (replaces c10d Reducer all-reduce + _amp_foreach_non_
//                      finite_check_and_unscale_; reference call sites
//                      distributed.py:262, distributed_syncBN_amp.py:275-276).
//   k_sgd_allgather    inf consensus across ranks, momentum-SGD on the owned
//                      shard of EVERY segment in one launch, then the updated fp32
//                      master weights and their bf16/fp16 shadow are stored into
//                      every peer's parameter arena (replaces foreach SGD + the
//                      all-gather half of the all-reduce + autocast weight casts;
//                      optimizer.step() at distributed.py:263).
//   k_allreduce        in-place two-shot all-reduce for callers that keep
//                      their own optimizer.
//   k_allreduce_small  one-shot all-reduce of <= 8192 floats (loss/acc scalars,
//                      distributed.py:254-255; no separate barrier, the exchange
//                      is one), optionally accumulating the result into a device
//                      meter (AverageMeter.update, distributed.py:257-258).
//   k_bcast_pull       every rank pulls a range from the root's arena (DDP
//                      constructor sync and per-forward BN-buffer broadcast,
//                      torch/nn/parallel/distributed.py:879-890, 2214-2257).
//   k_scale_update     GradScaler.update() on device state.
//
// Synchronisation: per-(channel, block, peer) 32-bit flags in each rank's pad.
// signal = CAS 0->1 on the peer's flag (release.sys), wait = CAS 1->0 on the
// local flag (acquire.sys): self-resetting, so pads never need clearing and a
// replayed CUDA graph needs no epoch argument.  Block b of rank r only ever
// talks to block b of the peers, hence all ranks launch identical grids.
//
// Rank emulation (tests): the *_emu entry points run the SAME device
// bodies for all W ranks inside ONE cooperative launch on one GPU (rank =
// blockIdx.y, every rank's arena a separate buffer of that GPU), so shard
// ownership, reduction order and the flag protocol are testable without W GPUs.
#include <cooperative_groups.h>
#include "fabric.cuh"

namespace {

using namespace pdt;
__device__ __forceinline__ float4 ld_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_f4(float* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ bool finite4(const float4& v) {
  return isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w);
}

struct Segs {
  int n;
  size_t off[PDT_MAX_SEGMENTS], len[PDT_MAX_SEGMENTS];
};

// ---------------------------------------------------------------------------
// grads_mc != nullptr (and MC): the shard is summed by the switch.
template <int W, bool MC>
__device__ __forceinline__ void reduce_scatter_body(const Comm& c, const PdtPeers& grads, const float* grads_mc, size_t seg_off,
                                                    size_t seg_len, float* amp_state, int use_flags) {
  if (use_flags) block_barrier(c, 0, blockIdx.x);
  const size_t shard = seg_len / W;
  const size_t base = seg_off + (size_t)c.rank * shard;
  const float inv = 1.0f / ((float)W * amp_state[0]);
  bool bad = false;
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  float* mine = reinterpret_cast<float*>(grads.p[c.rank]);
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < shard; i += stride) {
    float4 a;
    if constexpr (MC) {
      a = mc_ld_reduce_f4(grads_mc + base + i);
    } else {
      float4 v[W];
#pragma unroll
      for (int p = 0; p < W; ++p) v[p] = ld_f4(reinterpret_cast<const float*>(grads.p[p]) + base + i);
      a = v[0];
#pragma unroll
      for (int p = 1; p < W; ++p) { a.x += v[p].x; a.y += v[p].y; a.z += v[p].z; a.w += v[p].w; }
    }
    a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
    bad |= !finite4(a);
    st_f4(mine + base + i, a);
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(reinterpret_cast<int*>(amp_state) + 2, 1);
  if (use_flags) block_barrier(c, 0, blockIdx.x);   // peers are done reading my arena
}

template <int W, bool MC>
__global__ void __launch_bounds__(512) k_reduce_scatter(Comm c, PdtPeers grads, const float* grads_mc, size_t seg_off,
                                                        size_t seg_len, float* amp_state, int use_flags) {
  reduce_scatter_body<W, MC>(c, grads, grads_mc, seg_off, seg_len, amp_state, use_flags);
}

// hyper = {lr, momentum, weight_decay, amp_enabled}
template <int W, typename T, bool MC>
__device__ __forceinline__ void sgd_allgather_body(const Comm& c, const PdtPeers& grads, const PdtPeers& params,
                                                   const PdtPeers& shadow, float* params_mc, T* shadow_mc,
                                                   float* __restrict__ momentum, const Segs& segs,
                                                   const float* __restrict__ hyper, float* amp_state, int use_flags) {
  __shared__ int s_skip;
  const bool amp = hyper[3] != 0.f;
  int skip = 0;
  if (use_flags) {
    // publish my found_inf to block b of every peer, then meet them
    if ((int)threadIdx.x < W)
      *inf_addr(c.pad[threadIdx.x], blockIdx.x, c.rank) = (uint32_t)reinterpret_cast<int*>(amp_state)[2];
    block_barrier(c, 0, blockIdx.x);
    if (threadIdx.x == 0) {
      int any = 0;
      for (int p = 0; p < W; ++p) any |= (int)*reinterpret_cast<volatile uint32_t*>(inf_addr(c.pad[c.rank], blockIdx.x, p));
      s_skip = any;
    }
    __syncthreads();
    skip = s_skip;
  } else {
    skip = reinterpret_cast<int*>(amp_state)[2];
  }
  skip = amp ? skip : 0;
  if (!skip) {
    const float lr = hyper[0], mu = hyper[1], wd = hyper[2];
    const float* g = reinterpret_cast<const float*>(grads.p[c.rank]);
    const float* pw = reinterpret_cast<const float*>(params.p[c.rank]);
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (int s = 0; s < segs.n; ++s) {
      const size_t shard = segs.len[s] / W;
      const size_t base = segs.off[s] + (size_t)c.rank * shard;
      for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < shard; i += stride) {
        float4 gv = ld_f4(g + base + i), pv = ld_f4(pw + base + i), mv = ld_f4(momentum + base + i);
        gv.x = fmaf(wd, pv.x, gv.x); gv.y = fmaf(wd, pv.y, gv.y); gv.z = fmaf(wd, pv.z, gv.z); gv.w = fmaf(wd, pv.w, gv.w);
        mv.x = fmaf(mu, mv.x, gv.x); mv.y = fmaf(mu, mv.y, gv.y); mv.z = fmaf(mu, mv.z, gv.z); mv.w = fmaf(mu, mv.w, gv.w);
        pv.x = fmaf(-lr, mv.x, pv.x); pv.y = fmaf(-lr, mv.y, pv.y); pv.z = fmaf(-lr, mv.z, pv.z); pv.w = fmaf(-lr, mv.w, pv.w);
        st_f4(momentum + base + i, mv);
        if constexpr (MC) {
          mc_st_f4(params_mc + base + i, pv);
        } else {
#pragma unroll
          for (int q = 0; q < W; ++q) st_f4(reinterpret_cast<float*>(params.p[q]) + base + i, pv);
        }
        if (shadow.p[0] != nullptr) {
          uint2 lo;
          lo.x = Cvt<T>::pack2(pv.x, pv.y);
          lo.y = Cvt<T>::pack2(pv.z, pv.w);
          if constexpr (MC) {
            mc_st_b64(shadow_mc + base + i, lo);
          } else {
#pragma unroll
            for (int q = 0; q < W; ++q) *reinterpret_cast<uint2*>(reinterpret_cast<T*>(shadow.p[q]) + base + i) = lo;
          }
        }
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<int*>(amp_state)[4] = skip;   // consensus, read by k_scale_update
  if (use_flags) {
    __threadfence_system();
    block_barrier(c, 0, blockIdx.x);   // every push has landed before anybody's next forward
  }
}

template <int W, typename T, bool MC>
__global__ void __launch_bounds__(512) k_sgd_allgather(Comm c, PdtPeers grads, PdtPeers params, PdtPeers shadow, float* params_mc,
                                                       T* shadow_mc, float* __restrict__ momentum, Segs segs,
                                                       const float* __restrict__ hyper, float* amp_state, int use_flags) {
  sgd_allgather_body<W, T, MC>(c, grads, params, shadow, params_mc, shadow_mc, momentum, segs, hyper, amp_state, use_flags);
}

template <int W>
__global__ void __launch_bounds__(512) k_allreduce(Comm c, PdtPeers bufs, size_t off, size_t len, float post_scale) {
  block_barrier(c, 0, blockIdx.x);
  const size_t shard = len / W;
  const size_t base = off + (size_t)c.rank * shard;
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < shard; i += stride) {
    float4 v[W];
#pragma unroll
    for (int p = 0; p < W; ++p) v[p] = ld_f4(reinterpret_cast<const float*>(bufs.p[p]) + base + i);
    float4 a = v[0];
#pragma unroll
    for (int p = 1; p < W; ++p) { a.x += v[p].x; a.y += v[p].y; a.z += v[p].z; a.w += v[p].w; }
    a.x *= post_scale; a.y *= post_scale; a.z *= post_scale; a.w *= post_scale;
#pragma unroll
    for (int q = 0; q < W; ++q) st_f4(reinterpret_cast<float*>(bufs.p[q]) + base + i, a);
  }
  __threadfence_system();
  block_barrier(c, 0, blockIdx.x);
}

// One block.  in/out are local; n <= PDT_SCRATCH_FLOATS.  meter (optional, local): [0..n) += out * weight,
// [n] += weight, [n+1..2n+1) = out  — AverageMeter's sum / count / val for n metrics, kept on the device.
__device__ __forceinline__ void allreduce_small_body(const Comm& c, const float* __restrict__ in, float* __restrict__ out, int n,
                                                     float post_scale, int channel, float* meter, float weight) {
  if (c.world > 1) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float v = in[i];
      for (int q = 0; q < c.world; ++q) scratch_addr(c.pad[q], channel, c.rank)[i] = v;
    }
    __threadfence_system();
    block_barrier(c, channel, blockIdx.x);
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float a = 0.f;
    if (c.world > 1) {
      for (int p = 0; p < c.world; ++p) a += *reinterpret_cast<volatile float*>(scratch_addr(c.pad[c.rank], channel, p) + i);
    } else {
      a = in[i];
    }
    a *= post_scale;
    out[i] = a;
    if (meter != nullptr) {
      meter[i] += a * weight;
      meter[n + 1 + i] = a;
      if (i == 0) meter[n] += weight;
    }
  }
  if (c.world > 1) block_barrier(c, channel, blockIdx.x);   // slots may be overwritten by the next call
}

__global__ void __launch_bounds__(256) k_allreduce_small(Comm c, const float* __restrict__ in, float* __restrict__ out,
                                                         int n, float post_scale, int channel, float* meter, float weight) {
  allreduce_small_body(c, in, out, n, post_scale, channel, meter, weight);
}

__device__ __forceinline__ void bcast_pull_body(const Comm& c, const PdtPeers& bufs, size_t off_bytes, size_t len_bytes, int root,
                                                int channel) {
  block_barrier(c, channel, blockIdx.x);
  if (c.rank != root) {
    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(bufs.p[root]) + off_bytes);
    uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(bufs.p[c.rank]) + off_bytes);
    const size_t n = len_bytes / 16;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
      uint4 v;
      asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
      dst[i] = v;
    }
  }
  block_barrier(c, channel, blockIdx.x);   // root may change the range again
}

__global__ void __launch_bounds__(512) k_bcast_pull(Comm c, PdtPeers bufs, size_t off_bytes, size_t len_bytes, int root,
                                                    int channel) {
  bcast_pull_body(c, bufs, off_bytes, len_bytes, root, channel);
}

__global__ void k_barrier(Comm c, int channel) { block_barrier(c, channel, blockIdx.x); }

// amp_state words: [0] scale f32, [1] growth_tracker i32, [2] found_inf (local) i32,
//                  [3] skipped-step counter i32, [4] found_inf consensus i32
__global__ void k_scale_update(float* amp_state, float growth, float backoff, int interval) {
  int* s = reinterpret_cast<int*>(amp_state);
  if (s[4]) {
    amp_state[0] *= backoff;
    s[1] = 0;
    s[3] += 1;
  } else if (++s[1] == interval) {
    amp_state[0] *= growth;
    s[1] = 0;
  }
  s[2] = 0;
  s[4] = 0;
}

// plain momentum-SGD on a flat arena (world 1 without a fabric; also the oracle for the fused path)
template <typename T>
__global__ void __launch_bounds__(512) k_sgd_flat(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                  T* __restrict__ shadow, size_t n, const float* __restrict__ hyper,
                                                  float* amp_state) {
  const int skip = (hyper[3] != 0.f) ? reinterpret_cast<const int*>(amp_state)[2] : 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<int*>(amp_state)[4] = skip;
  if (skip) return;
  const float lr = hyper[0], mu = hyper[1], wd = hyper[2];
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float pv = p[i];
    const float gv = fmaf(wd, pv, g[i]);
    const float mv = fmaf(mu, m[i], gv);
    const float nv = fmaf(-lr, mv, pv);
    m[i] = mv;
    p[i] = nv;
    if (shadow) shadow[i] = Cvt<T>::from_f(nv);
  }
}

// ---- rank emulation: all W ranks in one cooperative launch, rank = blockIdx.y -------------------------------
struct EmuRanks {
  uint32_t* pad[PDT_MAX_RANKS];
  float* amp_state[PDT_MAX_RANKS];
  float* momentum[PDT_MAX_RANKS];
  float* io[PDT_MAX_RANKS];      // allreduce_small: per-rank in/out vector
};
__device__ __forceinline__ Comm emu_comm(const EmuRanks& e, int world, uint64_t timeout_ns) {
  Comm c;
  c.rank = blockIdx.y;
  c.world = world;
  c.timeout_ns = timeout_ns;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) c.pad[i] = e.pad[i];
  c.pad_mc = nullptr;
  return c;
}
template <int W>
__global__ void __launch_bounds__(512) k_emu_step(EmuRanks e, PdtPeers grads, PdtPeers params, PdtPeers shadow, Segs segs,
                                                  const float* hyper, uint64_t timeout_ns) {
  const Comm c = emu_comm(e, W, timeout_ns);
  for (int s = 0; s < segs.n; ++s)
    reduce_scatter_body<W, false>(c, grads, nullptr, segs.off[s], segs.len[s], e.amp_state[c.rank], 1);
  sgd_allgather_body<W, __nv_bfloat16, false>(c, grads, params, shadow, nullptr, nullptr, e.momentum[c.rank], segs, hyper,
                                              e.amp_state[c.rank], 1);
}
__global__ void __launch_bounds__(256) k_emu_small(EmuRanks e, int world, int n, float post_scale, PdtPeers bufs, size_t len_bytes,
                                                   int root, uint64_t timeout_ns) {
  const Comm c = emu_comm(e, world, timeout_ns);
  allreduce_small_body(c, e.io[c.rank], e.io[c.rank], n, post_scale, 1, nullptr, 0.f);
  bcast_pull_body(c, bufs, 0, len_bytes, root, 1);
}

int clamp_blocks(int blocks) { return blocks < 1 ? 1 : (blocks > kMaxBlocks ? kMaxBlocks : blocks); }

#define DISPATCH_W(W, ...)                       \
  switch (W) {                                   \
    case 1: { constexpr int kW = 1; __VA_ARGS__; break; } \
    case 2: { constexpr int kW = 2; __VA_ARGS__; break; } \
    case 3: { constexpr int kW = 3; __VA_ARGS__; break; } \
    case 4: { constexpr int kW = 4; __VA_ARGS__; break; } \
    case 5: { constexpr int kW = 5; __VA_ARGS__; break; } \
    case 6: { constexpr int kW = 6; __VA_ARGS__; break; } \
    case 7: { constexpr int kW = 7; __VA_ARGS__; break; } \
    case 8: { constexpr int kW = 8; __VA_ARGS__; break; } \
    default: return (int)cudaErrorInvalidValue;  \
  }

bool segs_from(const PdtSegments* in, int world, Segs* out) {
  if (in == nullptr || in->n < 1 || in->n > PDT_MAX_SEGMENTS) return false;
  out->n = in->n;
  for (int i = 0; i < PDT_MAX_SEGMENTS; ++i) {
    out->off[i] = i < in->n ? in->off[i] : 0;
    out->len[i] = i < in->n ? in->len[i] : 0;
    if (i < in->n && (in->len[i] % (4 * (size_t)world) || in->off[i] % 4)) return false;
  }
  return true;
}

}  // namespace

extern "C" {

size_t pdt_comm_pad_bytes(void) { return PDT_PAD_BYTES; }

int pdt_reduce_scatter_f32(const PdtCommDesc* d, const PdtPeers* grads, const void* grads_mc, size_t seg_off, size_t seg_len,
                           float* amp_state, int use_flags, int blocks, cudaStream_t stream) {
  if (seg_len % (4 * (size_t)d->world) || seg_off % 4) return (int)cudaErrorInvalidValue;
  Comm c = make_comm(d);
  blocks = clamp_blocks(blocks);
  const float* mc = reinterpret_cast<const float*>(grads_mc);
  if (mc != nullptr && d->world > 1 && use_flags) {
    DISPATCH_W(d->world, k_reduce_scatter<kW, true><<<blocks, 512, 0, stream>>>(c, *grads, mc, seg_off, seg_len, amp_state, use_flags));
  } else {
    DISPATCH_W(d->world, k_reduce_scatter<kW, false><<<blocks, 512, 0, stream>>>(c, *grads, nullptr, seg_off, seg_len, amp_state, use_flags));
  }
  return (int)cudaGetLastError();
}

int pdt_sgd_allgather(const PdtCommDesc* d, const PdtPeers* grads, const PdtPeers* params, const PdtPeers* shadow,
                      void* params_mc, void* shadow_mc, int shadow_dtype, float* momentum, const PdtSegments* segments,
                      const float* hyper, float* amp_state, int use_flags, int blocks, cudaStream_t stream) {
  Segs segs;
  if (!segs_from(segments, d->world, &segs)) return (int)cudaErrorInvalidValue;
  Comm c = make_comm(d);
  blocks = clamp_blocks(blocks);
  PdtPeers sh;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) sh.p[i] = shadow ? shadow->p[i] : nullptr;
  float* pmc = reinterpret_cast<float*>(params_mc);
  const bool mc = pmc != nullptr && d->world > 1 && use_flags && (sh.p[0] == nullptr || shadow_mc != nullptr);
#define PDT_SGD_LAUNCH(T)                                                                                                   \
  if (mc) {                                                                                                                 \
    DISPATCH_W(d->world, k_sgd_allgather<kW, T, true><<<blocks, 512, 0, stream>>>(                                          \
        c, *grads, *params, sh, pmc, reinterpret_cast<T*>(shadow_mc), momentum, segs, hyper, amp_state, use_flags));        \
  } else {                                                                                                                  \
    DISPATCH_W(d->world, k_sgd_allgather<kW, T, false><<<blocks, 512, 0, stream>>>(                                         \
        c, *grads, *params, sh, nullptr, nullptr, momentum, segs, hyper, amp_state, use_flags));                            \
  }
  if (shadow_dtype == PDT_FP16) {
    PDT_SGD_LAUNCH(__half)
  } else {
    PDT_SGD_LAUNCH(__nv_bfloat16)
  }
#undef PDT_SGD_LAUNCH
  return (int)cudaGetLastError();
}

int pdt_allreduce_f32(const PdtCommDesc* d, const PdtPeers* bufs, size_t off, size_t len, float post_scale, int blocks,
                      cudaStream_t stream) {
  if (len % (4 * (size_t)d->world) || off % 4) return (int)cudaErrorInvalidValue;
  Comm c = make_comm(d);
  blocks = clamp_blocks(blocks);
  DISPATCH_W(d->world, k_allreduce<kW><<<blocks, 512, 0, stream>>>(c, *bufs, off, len, post_scale));
  return (int)cudaGetLastError();
}

int pdt_allreduce_small_f32(const PdtCommDesc* d, const float* in, float* out, int n, float post_scale, int channel,
                            cudaStream_t stream) {
  if (n < 0 || n > PDT_SCRATCH_FLOATS || channel < 0 || channel >= kChannels) return (int)cudaErrorInvalidValue;
  k_allreduce_small<<<1, 256, 0, stream>>>(make_comm(d), in, out, n, post_scale, channel, nullptr, 0.f);
  return (int)cudaGetLastError();
}

int pdt_metrics_update(const PdtCommDesc* d, const float* in, float* out, int n, float post_scale, int channel, float* meter,
                       float weight, cudaStream_t stream) {
  if (n < 1 || n > 1024 || channel < 0 || channel >= kChannels) return (int)cudaErrorInvalidValue;
  k_allreduce_small<<<1, 256, 0, stream>>>(make_comm(d), in, out, n, post_scale, channel, meter, weight);
  return (int)cudaGetLastError();
}

int pdt_bcast_pull(const PdtCommDesc* d, const PdtPeers* bufs, size_t off_bytes, size_t len_bytes, int root,
                   int channel, int blocks, cudaStream_t stream) {
  if ((off_bytes | len_bytes) % 16 || channel < 0 || channel >= kChannels) return (int)cudaErrorInvalidValue;
  k_bcast_pull<<<clamp_blocks(blocks), 512, 0, stream>>>(make_comm(d), *bufs, off_bytes, len_bytes, root, channel);
  return (int)cudaGetLastError();
}

int pdt_barrier(const PdtCommDesc* d, int channel, cudaStream_t stream) {
  if (channel < 0 || channel >= kChannels) return (int)cudaErrorInvalidValue;
  k_barrier<<<1, 32, 0, stream>>>(make_comm(d), channel);
  return (int)cudaGetLastError();
}

int pdt_scale_update(float* amp_state, float growth, float backoff, int interval, cudaStream_t stream) {
  k_scale_update<<<1, 1, 0, stream>>>(amp_state, growth, backoff, interval);
  return (int)cudaGetLastError();
}

int pdt_sgd_flat(float* p, const float* g, float* m, void* shadow, int shadow_dtype, size_t n, const float* hyper,
                 float* amp_state, cudaStream_t stream) {
  const int blocks = 148 * 2;
  if (shadow_dtype == PDT_FP16)
    k_sgd_flat<__half><<<blocks, 512, 0, stream>>>(p, g, m, reinterpret_cast<__half*>(shadow), n, hyper, amp_state);
  else
    k_sgd_flat<__nv_bfloat16><<<blocks, 512, 0, stream>>>(p, g, m, reinterpret_cast<__nv_bfloat16*>(shadow), n, hyper, amp_state);
  return (int)cudaGetLastError();
}

// ---- rank emulation on one GPU (tests only) ----------------------------------------------------------------------
int pdt_emu_step(const PdtEmuRanks* ranks, int world, const PdtPeers* grads, const PdtPeers* params, const PdtPeers* shadow,
                 const PdtSegments* segments, const float* hyper, int blocks, cudaStream_t stream) {
  Segs segs;
  if (world < 1 || world > PDT_MAX_RANKS || !segs_from(segments, world, &segs)) return (int)cudaErrorInvalidValue;
  EmuRanks e;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) {
    e.pad[i] = reinterpret_cast<uint32_t*>(ranks->pad[i]);
    e.amp_state[i] = ranks->amp_state[i];
    e.momentum[i] = ranks->momentum[i];
    e.io[i] = ranks->io[i];
  }
  PdtPeers g = *grads, p = *params, sh;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) sh.p[i] = shadow ? shadow->p[i] : nullptr;
  uint64_t timeout = 5ull * 1000000000ull;       // co-resident blocks of one GPU: seconds mean a protocol bug
  blocks = clamp_blocks(blocks);
  dim3 grid(blocks, world);
  void* args[] = {&e, &g, &p, &sh, &segs, (void*)&hyper, &timeout};
  const void* fn = nullptr;
  DISPATCH_W(world, fn = (const void*)k_emu_step<kW>);
  return (int)cudaLaunchCooperativeKernel(fn, grid, dim3(512), args, 0, stream);
}

int pdt_emu_small(const PdtEmuRanks* ranks, int world, int n, float post_scale, const PdtPeers* bufs, size_t len_bytes, int root,
                  cudaStream_t stream) {
  if (world < 1 || world > PDT_MAX_RANKS || n < 0 || n > PDT_SCRATCH_FLOATS || len_bytes % 16) return (int)cudaErrorInvalidValue;
  EmuRanks e;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) {
    e.pad[i] = reinterpret_cast<uint32_t*>(ranks->pad[i]);
    e.amp_state[i] = ranks->amp_state[i];
    e.momentum[i] = ranks->momentum[i];
    e.io[i] = ranks->io[i];
  }
  PdtPeers b = *bufs;
  uint64_t timeout = 5ull * 1000000000ull;
  dim3 grid(1, world);
  void* args[] = {&e, &world, &n, &post_scale, &b, &len_bytes, &root, &timeout};
  return (int)cudaLaunchCooperativeKernel((const void*)k_emu_small, grid, dim3(256), args, 0, stream);
}

}  // extern "C"
