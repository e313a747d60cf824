// Cross-GPU kernels over NVLink 5 / NVSwitch peer memory.
//
// Every rank maps every peer's communication arena (CUDA VMM / torch symmetric
// memory) and these kernels move the data themselves with ld/st on peer
// pointers; there is no NCCL call, proxy thread or host round-trip on this
// path, so the whole training step stays CUDA-graph capturable.
//
//   k_reduce_scatter   gradient segment: each rank sums ITS shard over all
//                      peers' gradient arenas, folds in 1/world and the AMP
//                      1/loss_scale, and raises found_inf on a non-finite sum
//                      (replaces c10d Reducer all-reduce + _amp_foreach_non_
//                      finite_check_and_unscale_; reference call sites
//                      distributed.py:262, distributed_syncBN_amp.py:275-276).
//   k_sgd_allgather    inf consensus across ranks, momentum-SGD on the owned
//                      shard, then the updated fp32 master weights and their
//                      bf16/fp16 shadow are stored into EVERY peer's parameter
//                      arena (replaces foreach SGD + the all-gather half of the
//                      all-reduce + autocast weight casts; optimizer.step() at
//                      distributed.py:263).
//   k_allreduce        in-place two-shot all-reduce for callers that keep
//                      their own optimizer.
//   k_allreduce_small  one-shot all-reduce of <= 2048 floats (loss/acc
//                      scalars, distributed.py:254-255; no separate barrier,
//                      the exchange is one).
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

// ---------------------------------------------------------------------------
template <int W>
__global__ void __launch_bounds__(512) k_reduce_scatter(Comm c, PdtPeers grads, size_t seg_off, size_t seg_len,
                                                        float* amp_state, int use_flags) {
  if (use_flags) block_barrier(c, 0, blockIdx.x);
  const size_t shard = seg_len / W;
  const size_t base = seg_off + (size_t)c.rank * shard;
  const float inv = 1.0f / ((float)W * amp_state[0]);
  bool bad = false;
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < shard; i += stride) {
    float4 v[W];
#pragma unroll
    for (int p = 0; p < W; ++p) v[p] = ld_f4(reinterpret_cast<const float*>(grads.p[p]) + base + i);
    float4 a = v[0];
#pragma unroll
    for (int p = 1; p < W; ++p) { a.x += v[p].x; a.y += v[p].y; a.z += v[p].z; a.w += v[p].w; }
    a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
    bad |= !finite4(a);
    st_f4(reinterpret_cast<float*>(grads.p[c.rank]) + base + i, a);
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(reinterpret_cast<int*>(amp_state) + 2, 1);
  if (use_flags) block_barrier(c, 0, blockIdx.x);   // peers are done reading my arena
}

// hyper = {lr, momentum, weight_decay, amp_enabled}
template <int W, typename T>
__global__ void __launch_bounds__(512) k_sgd_allgather(Comm c, PdtPeers grads, PdtPeers params, PdtPeers shadow,
                                                       float* __restrict__ momentum, size_t seg_off, size_t seg_len,
                                                       const float* __restrict__ hyper, float* amp_state,
                                                       int use_flags) {
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
    const size_t shard = seg_len / W;
    const size_t base = seg_off + (size_t)c.rank * shard;
    const float* g = reinterpret_cast<const float*>(grads.p[c.rank]);
    const float* pw = reinterpret_cast<const float*>(params.p[c.rank]);
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < shard; i += stride) {
      float4 gv = ld_f4(g + base + i), pv = ld_f4(pw + base + i), mv = ld_f4(momentum + base + i);
      gv.x = fmaf(wd, pv.x, gv.x); gv.y = fmaf(wd, pv.y, gv.y); gv.z = fmaf(wd, pv.z, gv.z); gv.w = fmaf(wd, pv.w, gv.w);
      mv.x = fmaf(mu, mv.x, gv.x); mv.y = fmaf(mu, mv.y, gv.y); mv.z = fmaf(mu, mv.z, gv.z); mv.w = fmaf(mu, mv.w, gv.w);
      pv.x = fmaf(-lr, mv.x, pv.x); pv.y = fmaf(-lr, mv.y, pv.y); pv.z = fmaf(-lr, mv.z, pv.z); pv.w = fmaf(-lr, mv.w, pv.w);
      st_f4(momentum + base + i, mv);
#pragma unroll
      for (int q = 0; q < W; ++q) st_f4(reinterpret_cast<float*>(params.p[q]) + base + i, pv);
      if (shadow.p[0] != nullptr) {
        uint2 lo;
        lo.x = Cvt<T>::pack2(pv.x, pv.y);
        lo.y = Cvt<T>::pack2(pv.z, pv.w);
#pragma unroll
        for (int q = 0; q < W; ++q) *reinterpret_cast<uint2*>(reinterpret_cast<T*>(shadow.p[q]) + base + i) = lo;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<int*>(amp_state)[4] = skip;   // consensus, read by k_scale_update
  if (use_flags) {
    __threadfence_system();
    block_barrier(c, 0, blockIdx.x);   // every push has landed before anybody's next forward
  }
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

// One block.  in/out are local; n <= PDT_SCRATCH_FLOATS.
__global__ void __launch_bounds__(256) k_allreduce_small(Comm c, const float* __restrict__ in, float* __restrict__ out,
                                                         int n, float post_scale, int channel) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = in[i];
    for (int q = 0; q < c.world; ++q) scratch_addr(c.pad[q], channel, c.rank)[i] = v;
  }
  __threadfence_system();
  block_barrier(c, channel, blockIdx.x);
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float a = 0.f;
    for (int p = 0; p < c.world; ++p) a += *reinterpret_cast<volatile float*>(scratch_addr(c.pad[c.rank], channel, p) + i);
    out[i] = a * post_scale;
  }
  block_barrier(c, channel, blockIdx.x);   // slots may be overwritten by the next call
}

__global__ void __launch_bounds__(512) k_bcast_pull(Comm c, PdtPeers bufs, size_t off_bytes, size_t len_bytes, int root,
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

int clamp_blocks(int blocks) { return blocks < 1 ? 1 : (blocks > kMaxBlocks ? kMaxBlocks : blocks); }

#define DISPATCH_W(W, ...)                       \
  switch (W) {                                   \
    case 1: { constexpr int kW = 1; __VA_ARGS__; break; } \
    case 2: { constexpr int kW = 2; __VA_ARGS__; break; } \
    case 4: { constexpr int kW = 4; __VA_ARGS__; break; } \
    case 8: { constexpr int kW = 8; __VA_ARGS__; break; } \
    default: return (int)cudaErrorInvalidValue;  \
  }

}  // namespace

extern "C" {

size_t pdt_comm_pad_bytes(void) { return PDT_PAD_BYTES; }

int pdt_reduce_scatter_f32(const PdtCommDesc* d, const PdtPeers* grads, size_t seg_off, size_t seg_len,
                           float* amp_state, int use_flags, int blocks, cudaStream_t stream) {
  if (seg_len % (4 * (size_t)d->world)) return (int)cudaErrorInvalidValue;
  Comm c = make_comm(d);
  blocks = clamp_blocks(blocks);
  DISPATCH_W(d->world, k_reduce_scatter<kW><<<blocks, 512, 0, stream>>>(c, *grads, seg_off, seg_len, amp_state, use_flags));
  return (int)cudaGetLastError();
}

int pdt_sgd_allgather(const PdtCommDesc* d, const PdtPeers* grads, const PdtPeers* params, const PdtPeers* shadow,
                      int shadow_dtype, float* momentum, size_t seg_off, size_t seg_len, const float* hyper,
                      float* amp_state, int use_flags, int blocks, cudaStream_t stream) {
  if (seg_len % (4 * (size_t)d->world)) return (int)cudaErrorInvalidValue;
  Comm c = make_comm(d);
  blocks = clamp_blocks(blocks);
  PdtPeers sh;
  for (int i = 0; i < PDT_MAX_RANKS; ++i) sh.p[i] = shadow ? shadow->p[i] : nullptr;
  if (shadow_dtype == PDT_FP16) {
    DISPATCH_W(d->world, k_sgd_allgather<kW, __half><<<blocks, 512, 0, stream>>>(c, *grads, *params, sh, momentum, seg_off, seg_len, hyper, amp_state, use_flags));
  } else {
    DISPATCH_W(d->world, k_sgd_allgather<kW, __nv_bfloat16><<<blocks, 512, 0, stream>>>(c, *grads, *params, sh, momentum, seg_off, seg_len, hyper, amp_state, use_flags));
  }
  return (int)cudaGetLastError();
}

int pdt_allreduce_f32(const PdtCommDesc* d, const PdtPeers* bufs, size_t off, size_t len, float post_scale, int blocks,
                      cudaStream_t stream) {
  if (len % (4 * (size_t)d->world)) return (int)cudaErrorInvalidValue;
  Comm c = make_comm(d);
  blocks = clamp_blocks(blocks);
  DISPATCH_W(d->world, k_allreduce<kW><<<blocks, 512, 0, stream>>>(c, *bufs, off, len, post_scale));
  return (int)cudaGetLastError();
}

int pdt_allreduce_small_f32(const PdtCommDesc* d, const float* in, float* out, int n, float post_scale, int channel,
                            cudaStream_t stream) {
  if (n < 0 || n > PDT_SCRATCH_FLOATS || channel < 0 || channel >= kChannels) return (int)cudaErrorInvalidValue;
  k_allreduce_small<<<1, 256, 0, stream>>>(make_comm(d), in, out, n, post_scale, channel);
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

}  // extern "C"
