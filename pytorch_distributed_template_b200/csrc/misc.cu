// Loss, pooling and layout/cast kernels on NHWC activations (§2.4(a)).
//
//   k_ce_fwd        CrossEntropyLoss(mean) + unit dlogits + top-1 hit fraction in ONE pass
//                   over the logits (replaces _log_softmax + nll_loss_forward + their
//                   backward pair + topk/eq/sum of utils.accuracy; reference call sites
//                   distributed.py:247,250 and utils.py:105-111).
//   k_scale_cast    dlogits = unit_dlogits * grad_output (a device scalar, e.g. the AMP
//                   loss scale) cast to the activation dtype.
//   k_maxpool_*     MaxPool2d(3,2,1) forward (saves the arg-max tap as int8) / backward
//                   (gather form, no atomics).
//   k_avgpool_*     AdaptiveAvgPool2d(1) forward / backward.
//   k_nchw_to_nhwc  fp32 NCHW image batch -> low-precision NHWC with the channel dimension
//                   zero-padded (the autocast input cast + channels_last conversion).
#include "common.cuh"
#include "../../include/pdt_b200.h"

namespace {

using namespace pdt;

// ---- fused cross entropy ---------------------------------------------------------------
// One block (128 threads) per row.  row_out[b] = {loss_b, hit_b}; the block that takes the
// last ticket folds the rows in index order (deterministic) into out = {mean loss, hit fraction}.
template <typename T>
__global__ void __launch_bounds__(128) k_ce_fwd(const T* __restrict__ logits, const long long* __restrict__ target, int B, int K,
                                               float* __restrict__ dlogits_unit, float2* __restrict__ row_out,
                                               unsigned* counter, float* __restrict__ out) {
  pdl_enter();
  __shared__ float s_red[4];
  __shared__ int s_idx[4];
  __shared__ int s_last;
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const T* row = logits + (size_t)b * K;
  // Labels: nn.CrossEntropyLoss semantics.  ignore_index = -100 rows contribute neither loss nor gradient and
  // are left out of the mean; any other value outside [0, K) is a bug in the caller: trap (torch raises a
  // device-side assert) instead of reading out of bounds.
  const long long tgt_ll = target[b];
  const bool ignored = tgt_ll == -100;
  if (!ignored && (tgt_ll < 0 || tgt_ll >= K)) __trap();
  const int tgt = ignored ? -1 : (int)tgt_ll;
  int nvalid = 0;
  for (int i = tid; i < B; i += 128) nvalid += target[i] != -100;
  for (int o = 16; o > 0; o >>= 1) nvalid += __shfl_xor_sync(0xffffffffu, nvalid, o);
  __shared__ int s_nv[4];
  if (lane == 0) s_nv[wid] = nvalid;
  __syncthreads();
  nvalid = s_nv[0] + s_nv[1] + s_nv[2] + s_nv[3];

  float mx = -INFINITY;
  int arg = 0x7fffffff;
  for (int i = tid; i < K; i += 128) {
    const float v = Cvt<T>::to_f(row[i]);
    if (v > mx) { mx = v; arg = i; }
  }
  // arg-max with lowest-index tie break
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
  }
  if (lane == 0) { s_red[wid] = mx; s_idx[wid] = arg; }
  __syncthreads();
  mx = s_red[0]; arg = s_idx[0];
  for (int w = 1; w < 4; ++w)
    if (s_red[w] > mx || (s_red[w] == mx && s_idx[w] < arg)) { mx = s_red[w]; arg = s_idx[w]; }
  __syncthreads();

  float se = 0.f;
  for (int i = tid; i < K; i += 128) se += __expf(Cvt<T>::to_f(row[i]) - mx);
  se = warp_sum(se);
  if (lane == 0) s_red[wid] = se;
  __syncthreads();
  se = s_red[0] + s_red[1] + s_red[2] + s_red[3];
  const float lse = mx + __logf(se);
  const float inv = 1.f / se, invB = ignored ? 0.f : 1.f / (float)max(nvalid, 1);
  float* drow = dlogits_unit + (size_t)b * K;
  for (int i = tid; i < K; i += 128) {
    const float p = __expf(Cvt<T>::to_f(row[i]) - mx) * inv;
    drow[i] = (p - (i == tgt ? 1.f : 0.f)) * invB;
  }
  if (tid == 0) row_out[b] = ignored ? make_float2(0.f, 0.f) : make_float2(lse - Cvt<T>::to_f(row[tgt]), arg == tgt ? 1.f : 0.f);

  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // fixed-order fold: thread t sums rows t, t+128, ... then a fixed tree
  double l = 0.0, h = 0.0;
  for (int i = tid; i < B; i += 128) {
    const float2 v = __ldcg(row_out + i);
    l += v.x; h += v.y;
  }
  __shared__ double s_l[128], s_h[128];
  s_l[tid] = l; s_h[tid] = h;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if (tid < o) { s_l[tid] += s_l[tid + o]; s_h[tid] += s_h[tid + o]; }
    __syncthreads();
  }
  if (tid == 0) {
    out[0] = (float)(s_l[0] / max(nvalid, 1));     // mean over the rows that are not ignored (torch: reduction='mean')
    out[1] = (float)(s_h[0] / B);                  // utils.accuracy divides by the batch size
    *counter = 0;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) k_scale_cast(const float* __restrict__ in, const float* __restrict__ scale, T* __restrict__ out,
                                                   size_t n) {
  pdl_enter();
  const float s = *scale;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = Cvt<T>::from_f(in[i] * s);
}

// ---- pooling ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_maxpool_fwd(const T* __restrict__ x, T* __restrict__ y, int8_t* __restrict__ idx, int N, int H,
                                                    int W, int C, int Ho, int Wo, int ks, int st, int pd) {
  const int cg = C / 8;
  const long total = (long)N * Ho * Wo * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    long r = i / cg;
    const int wo = (int)(r % Wo); r /= Wo;
    const int ho = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float best[8];
    int bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
    for (int kh = 0; kh < ks; ++kh) {
      const int hi = ho * st - pd + kh;
      if (hi < 0 || hi >= H) continue;
      for (int kw = 0; kw < ks; ++kw) {
        const int wi = wo * st - pd + kw;
        if (wi < 0 || wi >= W) continue;
        float f[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(x + (((long)n * H + hi) * W + wi) * C + c8 * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (f[j] > best[j]) { best[j] = f[j]; bi[j] = kh * ks + kw; }
      }
    }
    const long o = (((long)n * Ho + ho) * Wo + wo) * C + c8 * 8;
    *reinterpret_cast<uint4*>(y + o) = pack8<T>(best);
    uint2 pk;
    pk.x = (bi[0] & 0xff) | ((bi[1] & 0xff) << 8) | ((bi[2] & 0xff) << 16) | ((bi[3] & 0xff) << 24);
    pk.y = (bi[4] & 0xff) | ((bi[5] & 0xff) << 8) | ((bi[6] & 0xff) << 16) | ((bi[7] & 0xff) << 24);
    *reinterpret_cast<uint2*>(idx + o) = pk;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) k_maxpool_bwd(const T* __restrict__ dy, const int8_t* __restrict__ idx, T* __restrict__ dx, int N,
                                                    int H, int W, int C, int Ho, int Wo, int ks, int st, int pd) {
  const int cg = C / 8;
  const long total = (long)N * H * W * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    long r = i / cg;
    const int wi = (int)(r % W); r /= W;
    const int hi = (int)(r % H);
    const int n = (int)(r / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    // outputs (ho, wo) whose window covers (hi, wi): ho*st - pd + kh == hi with 0 <= kh < ks,
    // i.e. ho in [ceil((hi + pd - ks + 1) / st), floor((hi + pd) / st)] — at most 2 x 2 for (3, 2, 1)
    const int ho_lo = max(0, (hi + pd - ks + st) / st), ho_hi = min(Ho - 1, (hi + pd) / st);
    const int wo_lo = max(0, (wi + pd - ks + st) / st), wo_hi = min(Wo - 1, (wi + pd) / st);
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
      const int kh = hi + pd - ho * st;
      for (int wo = wo_lo; wo <= wo_hi; ++wo) {
        const int kw = wi + pd - wo * st;
        const long o = (((long)n * Ho + ho) * Wo + wo) * C + c8 * 8;
        const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
        float g[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(dy + o), g);
        const int tap = kh * ks + kw;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int sel = ((j < 4 ? pk.x : pk.y) >> ((j & 3) * 8)) & 0xff;
          if (sel == tap) acc[j] += g[j];
        }
      }
    }
    *reinterpret_cast<uint4*>(dx + (((long)n * H + hi) * W + wi) * C + c8 * 8) = pack8<T>(acc);
  }
}

// fp32 variants (the reference's fp32 modes, distributed.py / dataparallel.py): 4 channels per 16-byte vector.
__global__ void __launch_bounds__(256) k_maxpool_fwd_f32(const float* __restrict__ x, float* __restrict__ y, int8_t* __restrict__ idx,
                                                        int N, int H, int W, int C, int Ho, int Wo, int ks, int st, int pd) {
  pdl_enter();
  const int cg = C / 4;
  const long total = (long)N * Ho * Wo * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % cg);
    long r = i / cg;
    const int wo = (int)(r % Wo); r /= Wo;
    const int ho = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    int bi[4] = {0, 0, 0, 0};
    for (int kh = 0; kh < ks; ++kh) {
      const int hi = ho * st - pd + kh;
      if (hi < 0 || hi >= H) continue;
      for (int kw = 0; kw < ks; ++kw) {
        const int wi = wo * st - pd + kw;
        if (wi < 0 || wi >= W) continue;
        const float4 v = *reinterpret_cast<const float4*>(x + (((long)n * H + hi) * W + wi) * C + c4 * 4);
        const float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (f[j] > best[j]) { best[j] = f[j]; bi[j] = kh * ks + kw; }
      }
    }
    const long o = (((long)n * Ho + ho) * Wo + wo) * C + c4 * 4;
    *reinterpret_cast<float4*>(y + o) = make_float4(best[0], best[1], best[2], best[3]);
    *reinterpret_cast<uint32_t*>(idx + o) = (bi[0] & 0xff) | ((bi[1] & 0xff) << 8) | ((bi[2] & 0xff) << 16) | ((bi[3] & 0xff) << 24);
  }
}

__global__ void __launch_bounds__(256) k_maxpool_bwd_f32(const float* __restrict__ dy, const int8_t* __restrict__ idx, float* __restrict__ dx,
                                                        int N, int H, int W, int C, int Ho, int Wo, int ks, int st, int pd) {
  pdl_enter();
  const int cg = C / 4;
  const long total = (long)N * H * W * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % cg);
    long r = i / cg;
    const int wi = (int)(r % W); r /= W;
    const int hi = (int)(r % H);
    const int n = (int)(r / H);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int ho_lo = max(0, (hi + pd - ks + st) / st), ho_hi = min(Ho - 1, (hi + pd) / st);
    const int wo_lo = max(0, (wi + pd - ks + st) / st), wo_hi = min(Wo - 1, (wi + pd) / st);
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
      const int kh = hi + pd - ho * st;
      for (int wo = wo_lo; wo <= wo_hi; ++wo) {
        const int kw = wi + pd - wo * st;
        const long o = (((long)n * Ho + ho) * Wo + wo) * C + c4 * 4;
        const uint32_t pk = *reinterpret_cast<const uint32_t*>(idx + o);
        const float4 g = *reinterpret_cast<const float4*>(dy + o);
        const float gf[4] = {g.x, g.y, g.z, g.w};
        const int tap = kh * ks + kw;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if ((int)((pk >> (j * 8)) & 0xff) == tap) acc[j] += gf[j];
      }
    }
    *reinterpret_cast<float4*>(dx + (((long)n * H + hi) * W + wi) * C + c4 * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  }
}

__global__ void __launch_bounds__(256) k_avgpool_fwd_f32(const float* __restrict__ x, float* __restrict__ y, int N, int HW, int C) {
  pdl_enter();
  const int cg = C / 4;
  const long total = (long)N * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % cg);
    const long n = i / cg;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < HW; ++p) {
      const float4 v = *reinterpret_cast<const float4*>(x + (n * HW + p) * C + c4 * 4);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    const float inv = 1.f / (float)HW;
    *reinterpret_cast<float4*>(y + n * C + c4 * 4) = make_float4(a.x * inv, a.y * inv, a.z * inv, a.w * inv);
  }
}

__global__ void __launch_bounds__(256) k_avgpool_bwd_f32(const float* __restrict__ dy, float* __restrict__ dx, int N, int HW, int C) {
  pdl_enter();
  const int cg = C / 4;
  const long total = (long)N * HW * cg;
  const float inv = 1.f / (float)HW;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % cg);
    const long n = i / cg / HW;
    const float4 v = *reinterpret_cast<const float4*>(dy + n * C + c4 * 4);
    *reinterpret_cast<float4*>(dx + i * 4) = make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) k_avgpool_fwd(const T* __restrict__ x, T* __restrict__ y, int N, int HW, int C) {
  pdl_enter();
  const int cg = C / 8;
  const long total = (long)N * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    const long n = i / cg;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int p = 0; p < HW; ++p) {
      float f[8];
      unpack8<T>(*reinterpret_cast<const uint4*>(x + (n * HW + p) * C + c8 * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    *reinterpret_cast<uint4*>(y + n * C + c8 * 8) = pack8<T>(acc);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) k_avgpool_bwd(const T* __restrict__ dy, T* __restrict__ dx, int N, int HW, int C) {
  pdl_enter();
  const int cg = C / 8;
  const long total = (long)N * HW * cg;
  const float inv = 1.f / (float)HW;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    const long n = i / cg / HW;
    float f[8];
    unpack8<T>(*reinterpret_cast<const uint4*>(dy + n * C + c8 * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= inv;
    *reinterpret_cast<uint4*>(dx + i * 8) = pack8<T>(f);
  }
}

// x: [N][Cin][H][W] fp32 -> y: [N][H][W][Cpad] T, channels >= Cin zero.  Cpad in {4, 8}.
template <typename T, int CP>
__global__ void __launch_bounds__(256) k_nchw_to_nhwc(const float* __restrict__ x, T* __restrict__ y, int N, int Cin, long HW) {
  const long total = (long)N * HW;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long n = i / HW, p = i % HW;
    float f[CP];
#pragma unroll
    for (int c = 0; c < CP; ++c) f[c] = (c < Cin) ? x[(n * Cin + c) * HW + p] : 0.f;
    if constexpr (CP == 8) {
      float g[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) g[c] = f[c];
      *reinterpret_cast<uint4*>(y + i * 8) = pack8<T>(g);
    } else {
      uint2 u;
      u.x = Cvt<T>::pack2(f[0], f[1]);
      u.y = Cvt<T>::pack2(f[2], f[3]);
      *reinterpret_cast<uint2*>(y + i * 4) = u;
    }
  }
}

// Stem input transform: x [N][3][H][W] fp32 (NCHW) -> y [N][HP][WP][16] low precision with
//   y[n][i][j][(dy*2+dx)*3 + c] = xpad[n][c][2i+dy][2j+dx],  xpad = x zero-padded by `pad` on every side,
// channels 12..15 zero (2x2 space-to-depth).  A 7x7 stride-2 pad-3 convolution over x becomes a
// 4x4 stride-1 convolution over y whose four horizontal taps x 16 channels are 64 CONTIGUOUS
// elements — exactly one 128-byte tensor-core K row (see ops/stem.py).
// TI = float: the loader already normalised the image.  TI = uint8_t: raw 0..255 pixels; the per-channel
// (x / 255 - mean) / std of torchvision's ToTensor + Normalize is applied here as x * sc[c] + bs[c], so the
// host ships one byte per value instead of four (the H2D copy bounds the end-to-end step at 8 GPUs).
struct StemNorm { float sc[3], bs[3]; };
template <typename T, typename TI>
__global__ void __launch_bounds__(256) k_stem_s2d(const TI* __restrict__ x, T* __restrict__ y, int N, int H, int W, int HP, int WP,
                                                 int pad, StemNorm nm) {
  pdl_enter();
  const long total = (long)N * HP * WP;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int j = (int)(i % WP);
    long r = i / WP;
    const int ii = (int)(r % HP);
    const long n = r / HP;
    float f[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) f[e] = 0.f;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int yy = 2 * ii + dy - pad;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int xx = 2 * j + dx - pad;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const TI v = __ldg(x + ((n * 3 + c) * H + yy) * (long)W + xx);
            if constexpr (sizeof(TI) == 1) f[(dy * 2 + dx) * 3 + c] = fmaf((float)v, nm.sc[c], nm.bs[c]);
            else f[(dy * 2 + dx) * 3 + c] = (float)v;
          }
        }
      }
    }
    if constexpr (sizeof(T) == 4) {
      float4* dst = reinterpret_cast<float4*>(y + i * 16);
#pragma unroll
      for (int e = 0; e < 4; ++e) dst[e] = make_float4(f[4 * e], f[4 * e + 1], f[4 * e + 2], f[4 * e + 3]);
    } else {
      float a[8], b[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { a[e] = f[e]; b[e] = f[8 + e]; }
      uint4* dst = reinterpret_cast<uint4*>(y + i * 16);
      dst[0] = pack8<T>(a);
      dst[1] = pack8<T>(b);
    }
  }
}

int grid_for(long total, int threads) {
  long nb = (total + threads - 1) / threads;
  if (nb < 1) nb = 1;
  if (nb > 148 * 16) nb = 148 * 16;
  return (int)nb;
}

}  // namespace

extern "C" {

#define LP_DISPATCH(dtype, ...)                                      \
  if (dtype == PDT_BF16) { using T = __nv_bfloat16; __VA_ARGS__; }   \
  else if (dtype == PDT_FP16) { using T = __half; __VA_ARGS__; }     \
  else return PDT_E_UNSUPPORTED;

size_t pdt_ce_scratch_bytes(int B) { return 256 + (size_t)B * sizeof(float2); }

int pdt_ce_fwd(const void* logits, int dtype, const long long* target, int B, int K, float* dlogits_unit, void* scratch,
               float* out, cudaStream_t stream) {
  unsigned* counter = reinterpret_cast<unsigned*>(scratch);
  float2* rows = reinterpret_cast<float2*>(reinterpret_cast<char*>(scratch) + 256);
  if (dtype == PDT_FP32)
    launch_pdl(k_ce_fwd<float>, dim3(B), dim3(128), 0, stream, reinterpret_cast<const float*>(logits), target, B, K, dlogits_unit, rows, counter, out);
  else
    LP_DISPATCH(dtype, launch_pdl(k_ce_fwd<T>, dim3(B), dim3(128), 0, stream, reinterpret_cast<const T*>(logits), target, B, K, dlogits_unit, rows, counter, out));
  return (int)cudaGetLastError();
}

int pdt_scale_cast(const float* in, const float* scale, void* out, int dtype, size_t n, cudaStream_t stream) {
  const int nb = grid_for((long)n, 256 * 4);
  if (dtype == PDT_FP32)
    launch_pdl(k_scale_cast<float>, dim3(nb), dim3(256), 0, stream, in, scale, reinterpret_cast<float*>(out), n);
  else
    LP_DISPATCH(dtype, launch_pdl(k_scale_cast<T>, dim3(nb), dim3(256), 0, stream, in, scale, reinterpret_cast<T*>(out), n));
  return (int)cudaGetLastError();
}

int pdt_maxpool_fwd(const void* x, void* y, int8_t* idx, int dtype, int N, int H, int W, int C, int ks, int st, int pd,
                    cudaStream_t stream) {
  if (C % 8 || ks * ks > 127) return PDT_E_UNSUPPORTED;
  const int Ho = (H + 2 * pd - ks) / st + 1, Wo = (W + 2 * pd - ks) / st + 1;
  if (dtype == PDT_FP32) {
    launch_pdl(k_maxpool_fwd_f32, dim3(grid_for((long)N * Ho * Wo * (C / 4), 256)), dim3(256), 0, stream,
               reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), idx, N, H, W, C, Ho, Wo, ks, st, pd);
    return (int)cudaGetLastError();
  }
  LP_DISPATCH(dtype, k_maxpool_fwd<T><<<grid_for((long)N * Ho * Wo * (C / 8), 256), 256, 0, stream>>>(
      reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), idx, N, H, W, C, Ho, Wo, ks, st, pd));
  return (int)cudaGetLastError();
}

int pdt_maxpool_bwd(const void* dy, const int8_t* idx, void* dx, int dtype, int N, int H, int W, int C, int ks, int st,
                    int pd, cudaStream_t stream) {
  if (C % 8) return PDT_E_UNSUPPORTED;
  const int Ho = (H + 2 * pd - ks) / st + 1, Wo = (W + 2 * pd - ks) / st + 1;
  if (dtype == PDT_FP32) {
    launch_pdl(k_maxpool_bwd_f32, dim3(grid_for((long)N * H * W * (C / 4), 256)), dim3(256), 0, stream,
               reinterpret_cast<const float*>(dy), idx, reinterpret_cast<float*>(dx), N, H, W, C, Ho, Wo, ks, st, pd);
    return (int)cudaGetLastError();
  }
  LP_DISPATCH(dtype, k_maxpool_bwd<T><<<grid_for((long)N * H * W * (C / 8), 256), 256, 0, stream>>>(
      reinterpret_cast<const T*>(dy), idx, reinterpret_cast<T*>(dx), N, H, W, C, Ho, Wo, ks, st, pd));
  return (int)cudaGetLastError();
}

int pdt_avgpool_fwd(const void* x, void* y, int dtype, int N, int HW, int C, cudaStream_t stream) {
  if (C % 8) return PDT_E_UNSUPPORTED;
  if (dtype == PDT_FP32) {
    launch_pdl(k_avgpool_fwd_f32, dim3(grid_for((long)N * (C / 4), 256)), dim3(256), 0, stream,
               reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), N, HW, C);
    return (int)cudaGetLastError();
  }
  LP_DISPATCH(dtype, launch_pdl(k_avgpool_fwd<T>, dim3(grid_for((long)N * (C / 8), 256)), dim3(256), 0, stream, 
      reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), N, HW, C));
  return (int)cudaGetLastError();
}

int pdt_avgpool_bwd(const void* dy, void* dx, int dtype, int N, int HW, int C, cudaStream_t stream) {
  if (C % 8) return PDT_E_UNSUPPORTED;
  if (dtype == PDT_FP32) {
    launch_pdl(k_avgpool_bwd_f32, dim3(grid_for((long)N * HW * (C / 4), 256)), dim3(256), 0, stream,
               reinterpret_cast<const float*>(dy), reinterpret_cast<float*>(dx), N, HW, C);
    return (int)cudaGetLastError();
  }
  LP_DISPATCH(dtype, launch_pdl(k_avgpool_bwd<T>, dim3(grid_for((long)N * HW * (C / 8), 256)), dim3(256), 0, stream, 
      reinterpret_cast<const T*>(dy), reinterpret_cast<T*>(dx), N, HW, C));
  return (int)cudaGetLastError();
}

int pdt_stem_s2d(const float* x, void* y, int dtype, int N, int H, int W, int pad, cudaStream_t stream) {
  const int HP = (H + 2 * pad + 1) / 2, WP = (W + 2 * pad + 1) / 2;
  const int nb = grid_for((long)N * HP * WP, 256);
  StemNorm nm = {{1.f, 1.f, 1.f}, {0.f, 0.f, 0.f}};
  if (dtype == PDT_FP32) {
    launch_pdl(k_stem_s2d<float, float>, dim3(nb), dim3(256), 0, stream, x, reinterpret_cast<float*>(y), N, H, W, HP, WP, pad, nm);
    return (int)cudaGetLastError();
  }
  LP_DISPATCH(dtype, launch_pdl(k_stem_s2d<T, float>, dim3(nb), dim3(256), 0, stream, x, reinterpret_cast<T*>(y), N, H, W, HP, WP, pad, nm));
  return (int)cudaGetLastError();
}

// Same transform from raw uint8 NCHW pixels with the normalisation fused: value = x * scale[c] + bias[c]
// (scale = 1 / (255 std), bias = -mean / std; host arrays of 3 floats).
int pdt_stem_s2d_u8(const unsigned char* x, void* y, int dtype, int N, int H, int W, int pad, const float* scale, const float* bias,
                    cudaStream_t stream) {
  if (pad < 0) return PDT_E_UNSUPPORTED;
  const int HP = (H + 2 * pad + 1) / 2, WP = (W + 2 * pad + 1) / 2;
  const int nb = grid_for((long)N * HP * WP, 256);
  StemNorm nm;
  for (int c = 0; c < 3; ++c) { nm.sc[c] = scale[c]; nm.bs[c] = bias[c]; }
  if (dtype == PDT_FP32) {
    launch_pdl(k_stem_s2d<float, unsigned char>, dim3(nb), dim3(256), 0, stream, x, reinterpret_cast<float*>(y), N, H, W, HP, WP, pad, nm);
    return (int)cudaGetLastError();
  }
  LP_DISPATCH(dtype, launch_pdl(k_stem_s2d<T, unsigned char>, dim3(nb), dim3(256), 0, stream, x, reinterpret_cast<T*>(y), N, H, W, HP, WP, pad, nm));
  return (int)cudaGetLastError();
}

int pdt_nchw_to_nhwc(const float* x, void* y, int dtype, int N, int Cin, int H, int W, int Cpad, cudaStream_t stream) {
  if (Cin > Cpad || (Cpad != 4 && Cpad != 8)) return PDT_E_UNSUPPORTED;
  const long HW = (long)H * W;
  const int nb = grid_for((long)N * HW, 256);
  if (Cpad == 8) {
    LP_DISPATCH(dtype, k_nchw_to_nhwc<T, 8><<<nb, 256, 0, stream>>>(x, reinterpret_cast<T*>(y), N, Cin, HW));
  } else {
    LP_DISPATCH(dtype, k_nchw_to_nhwc<T, 4><<<nb, 256, 0, stream>>>(x, reinterpret_cast<T*>(y), N, Cin, HW));
  }
  return (int)cudaGetLastError();
}

}  // extern "C"

// ---- peer copy (DataParallel scatter / gather / buffer broadcast) ------------------------------
// Runs on the DESTINATION device and pulls from a peer-mapped source pointer over NVLink
// (replaces torch.cuda.comm.scatter / gather / broadcast_coalesced, torch/nn/parallel/comm.py).
namespace {
__global__ void __launch_bounds__(512) k_peer_copy(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16,
                                                  const unsigned char* __restrict__ src_tail, unsigned char* __restrict__ dst_tail,
                                                  int tail) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
    uint4 v;
    asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
    dst[i] = v;
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}
}  // namespace

extern "C" int pdt_peer_copy(const void* src, void* dst, size_t nbytes, cudaStream_t stream) {
  if ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) return PDT_E_UNSUPPORTED;
  const size_t n16 = nbytes / 16;
  const int tail = (int)(nbytes % 16);
  long nb = (long)((n16 + 512 * 4 - 1) / (512 * 4));
  if (nb < 1) nb = 1;
  if (nb > 148 * 4) nb = 148 * 4;
  k_peer_copy<<<(int)nb, 512, 0, stream>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), n16,
                                           reinterpret_cast<const unsigned char*>(src) + n16 * 16,
                                           reinterpret_cast<unsigned char*>(dst) + n16 * 16, tail);
  return (int)cudaGetLastError();
}

// ---- filter re-layout for the data gradient --------------------------------------------------------
// dx of a convolution is itself a convolution of dy with a filter whose channel roles are swapped and whose
// taps are a (reversed) subset of the forward taps: out[c][i][j][k] = w[k][rs[i]][ss[j]][c] for 16-bit
// elements, w = [K][R][S][C].  Stride 1 uses all taps reversed; one parity class of a stride-2 gradient uses
// every second tap.  This replaces the flip / index_select x2 / transpose / contiguous chain (3-5 ATen launches
// per convolution and step) by one tiled transpose through shared memory.
namespace {
struct TapSel { int rs[8], ss[8]; };
__global__ void __launch_bounds__(256) k_dgrad_weight(const uint16_t* __restrict__ w, uint16_t* __restrict__ out, int K, int C, int R,
                                                      int S, int Th, int Tw, TapSel sel) {
  __shared__ uint16_t tile[32][34];
  const int ij = blockIdx.z, i = ij / Tw, j = ij - i * Tw;
  const int r = sel.rs[i], s = sel.ss[j];
  const int k0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
  for (int kk = ty; kk < 32; kk += 8) {
    const int k = k0 + kk, c = c0 + tx;
    tile[kk][tx] = (k < K && c < C) ? w[(((size_t)k * R + r) * S + s) * C + c] : (uint16_t)0;
  }
  __syncthreads();
  for (int cc = ty; cc < 32; cc += 8) {
    const int c = c0 + cc, k = k0 + tx;
    if (c < C && k < K) out[(((size_t)c * Th + i) * Tw + j) * K + k] = tile[tx][cc];
  }
}
}  // namespace

extern "C" int pdt_dgrad_weight(const void* w, void* out, int K, int C, int R, int S, int Th, int Tw, const int* rs, const int* ss,
                                cudaStream_t stream) {
  if (Th < 1 || Tw < 1 || Th > 8 || Tw > 8 || K < 1 || C < 1) return PDT_E_UNSUPPORTED;
  TapSel sel;
  for (int i = 0; i < 8; ++i) { sel.rs[i] = i < Th ? rs[i] : 0; sel.ss[i] = i < Tw ? ss[i] : 0; }
  for (int i = 0; i < Th; ++i) if (sel.rs[i] < 0 || sel.rs[i] >= R) return PDT_E_UNSUPPORTED;
  for (int j = 0; j < Tw; ++j) if (sel.ss[j] < 0 || sel.ss[j] >= S) return PDT_E_UNSUPPORTED;
  dim3 grid((K + 31) / 32, (C + 31) / 32, Th * Tw);
  k_dgrad_weight<<<grid, 256, 0, stream>>>(reinterpret_cast<const uint16_t*>(w), reinterpret_cast<uint16_t*>(out), K, C, R, S, Th, Tw, sel);
  return (int)cudaGetLastError();
}

// All filter re-layouts of a model in ONE launch: `jobs` is a device array (built once, the pointers are arena
// slots and persistent buffers), a block finds its job from the running tile count.
namespace {
constexpr int kDT = 64;      // tile edge: 64 output channels x 64 input channels (128-byte rows on both sides)
__global__ void __launch_bounds__(256) k_dgrad_weight_batched(const PdtDgradJob* __restrict__ jobs, int njobs) {
  pdl_enter();
  __shared__ __align__(16) uint16_t tile[kDT][kDT + 2];
  __shared__ PdtDgradJob jb;
  if (threadIdx.x == 0) {
    int lo = 0, hi = njobs - 1;                       // last job whose tile_begin <= blockIdx.x
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (jobs[mid].tile_begin <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
    }
    jb = jobs[lo];
  }
  __syncthreads();
  const int K = jb.K, C = jb.C, R = jb.R, S = jb.S, Th = jb.Th, Tw = jb.Tw;
  const int kt = (K + kDT - 1) / kDT, ct = (C + kDT - 1) / kDT;
  int t = (int)blockIdx.x - jb.tile_begin;
  const int bx = t % kt; t /= kt;
  const int by = t % ct; t /= ct;
  const int i = t / Tw, j = t - i * Tw;
  const int r = jb.rs[i], s = jb.ss[j];
  const uint16_t* __restrict__ w = reinterpret_cast<const uint16_t*>(jb.src);
  uint16_t* __restrict__ out = reinterpret_cast<uint16_t*>(jb.dst);
  const int k0 = bx * kDT, c0 = by * kDT;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 lanes x 8 rows; a lane moves two adjacent elements
  // The 32 x 32 version of this kernel moved 64-byte row pieces with one load in flight per thread (53 us for the 22 MB
  // of ResNet-18 filters, 0.8 TB/s).  Here a warp reads / writes whole 128-byte rows and has eight loads in flight.
  const bool vec = ((C | K) & 1) == 0 && ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(out)) & 3) == 0;
  uint32_t v[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int k = k0 + ty + 8 * m, c = c0 + 2 * tx;
    const size_t at = (((size_t)k * R + r) * S + s) * C + c;
    uint32_t x = 0;
    if (k < K && c < C) {
      if (vec) x = *reinterpret_cast<const uint32_t*>(w + at);        // c + 1 < C because C is even
      else x = (uint32_t)w[at] | (c + 1 < C ? (uint32_t)w[at + 1] << 16 : 0u);
    }
    v[m] = x;
  }
#pragma unroll
  for (int m = 0; m < 8; ++m) *reinterpret_cast<uint32_t*>(&tile[ty + 8 * m][2 * tx]) = v[m];
  __syncthreads();
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int cc = ty + 8 * m, c = c0 + cc, k = k0 + 2 * tx;
    if (c < C && k < K) {
      const size_t at = (((size_t)c * Th + i) * Tw + j) * K + k;
      const uint32_t lo = tile[2 * tx][cc], hi = tile[2 * tx + 1][cc];
      if (vec) *reinterpret_cast<uint32_t*>(out + at) = lo | (hi << 16);
      else { out[at] = (uint16_t)lo; if (k + 1 < K) out[at + 1] = (uint16_t)hi; }
    }
  }
}
}  // namespace

extern "C" int pdt_dgrad_weight_batched(const PdtDgradJob* jobs_dev, int njobs, int total_tiles, cudaStream_t stream) {
  if (njobs < 1 || total_tiles < 1) return 0;
  launch_pdl(k_dgrad_weight_batched, dim3(total_tiles), dim3(256), 0, stream, jobs_dev, njobs);
  return (int)cudaGetLastError();
}

// ---- stem filter in space-to-depth form ---------------------------------------------------------------------
// wp[k][r'][s'][(dy, dx, c) padded to 16] = w[k][c][2r' + dy][2s' + dx] (zero where 2r' + dy or 2s' + dx is 7): the filter
// of the 4x4 / 1 convolution over the space-to-depth image that equals torchvision's 7x7 / 2 conv1.  One launch and
// its inverse below instead of the pad / reshape / permute / contiguous / pad chain (five ATen launches per forward,
// three per backward, ~30 us of a 3.8 ms step at 150 images per GPU).  sk, sc, sr, ss: element strides of w.
namespace {
template <typename T>
__global__ void __launch_bounds__(256) k_stem_weight_s2d(const T* __restrict__ w, T* __restrict__ wp, int K, long sk, long sc, long sr, long ss) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // one element of wp: [K][4][4][16]
  if (i >= K * 256) return;
  const int c16 = i & 15, sp = (i >> 4) & 3, rp = (i >> 6) & 3, k = i >> 8;
  T v = T(0.f);
  if (c16 < 12) {
    const int dy = c16 / 6, dx = (c16 / 3) & 1, c = c16 % 3;
    const int r = 2 * rp + dy, q = 2 * sp + dx;
    if (r < 7 && q < 7) v = w[k * sk + c * sc + r * sr + q * ss];
  }
  wp[i] = v;
}
// dw[k][c][r][s] (element strides sk, sc, sr, ss) = dwp[k][r / 2][s / 2][(r % 2, s % 2, c)], fp32
__global__ void __launch_bounds__(256) k_stem_weight_grad_unpack(const float* __restrict__ dwp, float* __restrict__ dw, int K, long sk, long sc,
                                                                 long sr, long ss) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // one element of dw: K * 3 * 7 * 7
  if (i >= K * 147) return;
  const int s7 = i % 7, r = (i / 7) % 7, c = (i / 49) % 3, k = i / 147;
  dw[k * sk + c * sc + r * sr + s7 * ss] = dwp[k * 256 + (r >> 1) * 64 + (s7 >> 1) * 16 + (r & 1) * 6 + (s7 & 1) * 3 + c];
}
}  // namespace

extern "C" int pdt_stem_weight_s2d(const void* w, void* wp, int dtype, int K, long long sk, long long sc, long long sr, long long ss,
                                   cudaStream_t stream) {
  if (K < 1) return PDT_E_UNSUPPORTED;
  const int nb = (K * 256 + 255) / 256;
  if (dtype == PDT_FP32) launch_pdl(k_stem_weight_s2d<float>, dim3(nb), dim3(256), 0, stream, reinterpret_cast<const float*>(w),
                                    reinterpret_cast<float*>(wp), K, (long)sk, (long)sc, (long)sr, (long)ss);
  else if (dtype == PDT_FP16) launch_pdl(k_stem_weight_s2d<__half>, dim3(nb), dim3(256), 0, stream, reinterpret_cast<const __half*>(w),
                                         reinterpret_cast<__half*>(wp), K, (long)sk, (long)sc, (long)sr, (long)ss);
  else if (dtype == PDT_BF16) launch_pdl(k_stem_weight_s2d<__nv_bfloat16>, dim3(nb), dim3(256), 0, stream,
                                         reinterpret_cast<const __nv_bfloat16*>(w), reinterpret_cast<__nv_bfloat16*>(wp), K, (long)sk,
                                         (long)sc, (long)sr, (long)ss);
  else return PDT_E_UNSUPPORTED;
  return (int)cudaGetLastError();
}

extern "C" int pdt_stem_weight_grad_unpack(const float* dwp, float* dw, int K, long long sk, long long sc, long long sr, long long ss,
                                           cudaStream_t stream) {
  if (K < 1) return PDT_E_UNSUPPORTED;
  launch_pdl(k_stem_weight_grad_unpack, dim3((K * 147 + 255) / 256), dim3(256), 0, stream, dwp, dw, K, (long)sk, (long)sc, (long)sr, (long)ss);
  return (int)cudaGetLastError();
}

// Tile edge of pdt_dgrad_weight_batched: a job covers ceil(K / t) * ceil(C / t) * Th * Tw tiles (PdtDgradJob.tile_begin).
extern "C" int pdt_dgrad_tile(void) { return kDT; }

extern "C" int pdt_abi_version(void) { return PDT_ABI_VERSION; }
