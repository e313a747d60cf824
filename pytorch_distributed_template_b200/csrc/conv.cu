// Implicit-GEMM convolution on 5th-gen tensor cores (SURVEY K-CONV-F / K-CONV-D, N6).
//
// y[n,p,q,k] = sum_{r,s,c} x[n, p*stride - pad + r, q*stride - pad + s, c] * w[k,r,s,c]
// as a GEMM  D[M = N*P*Q, Cout] = A[M, R*S*C] * B[Cout, R*S*C]^T  with
//   A tiles (128 output pixels x 64 channels of one filter tap) fetched by TMA in IM2COL mode
//     straight from the NHWC activation: the hardware walks the output-pixel order across
//     row / image boundaries and zero-fills padding, no im2col buffer exists;
//   B tiles (BN filters x 64 channels) fetched by tiled TMA from the KRSC weights;
//   both land in 128B-swizzled shared memory, K-major, and feed tcgen05.mma (kind::f16,
//     M=128, N=BN, K=16) issued by ONE thread, accumulating in TMEM (fp32);
//   the accumulator is double-buffered in TMEM (2 x BN columns): 4 epilogue warps drain
//     tile i with tcgen05.ld -> bf16/fp16 -> global while the MMA thread already runs tile i+1.
// Persistent grid (one CTA per SM), warp-specialised: warp 0 = TMA producer, warp 1 = MMA
// issuer, warp 2 = TMEM allocator, warps 4-7 = epilogue.
//
// The same kernel computes the data gradient of stride-1 convolutions (feed dy as the
// activation, the flipped/transposed filter as the weight, pad' = R-1-pad).
// Replaces cudnnConvolutionForward / BackwardData (reference: every nn.Conv2d of
// torchvision resnet18, /root/reference/distributed.py:137,246,262).
#include <type_traits>
#include "tc.cuh"
#include "../../include/pdt_b200.h"

namespace {

using namespace pdt;
using namespace pdt::tc;

constexpr int BM = 128;        // output pixels per tile (= TMEM lanes)
constexpr int BK = 64;         // channels per k-block: 64 x 16 bit = one 128 B swizzle row
constexpr int kThreads = 256;

struct ConvGeom {
  int N, H, W, C;              // input  (NHWC)
  int K, R, S;                 // filter (KRSC)
  int P, Q;                    // output spatial
  int stride, pad;
  int M;                       // N*P*Q
  int m_tiles, n_tiles;
};

template <int BN> struct Cfg {
  static constexpr int kStageBytesA = BM * BK * 2;
  static constexpr int kStageBytesB = BN * BK * 2;
  static constexpr int kStageBytes = kStageBytesA + kStageBytesB;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kTmemCols = 2 * BN;     // double-buffered accumulator (power of two >= 32 for BN in {64,128,256})
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/;
};

template <int BN, typename T>
__global__ void __launch_bounds__(kThreads, 1)
k_conv_igemm(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, T* __restrict__ out,
             const ConvGeom g) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + C::kStages * C::kStageBytesA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* tfull = bars + 2 * C::kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = g.m_tiles * g.n_tiles;
  const int c_chunks = g.C / BK;
  const int k_blocks = g.R * g.S * c_chunks;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, C::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mt = tile / g.n_tiles, nt = tile % g.n_tiles;
        const int m0 = mt * BM;
        const int img = m0 / (g.P * g.Q);
        const int rem = m0 - img * (g.P * g.Q);
        const int p0 = rem / g.Q, q0 = rem - p0 * g.Q;
        const int cw = q0 * g.stride - g.pad, ch = p0 * g.stride - g.pad;
        for (int r = 0; r < g.R; ++r)
          for (int s = 0; s < g.S; ++s)
            for (int cc = 0; cc < c_chunks; ++cc) {
              mbar_wait(&empty[stage], phase ^ 1);
              mbar_expect_tx(&full[stage], C::kStageBytes);
              tma_load_im2col_4d(smA + stage * C::kStageBytesA, &tmA, &full[stage], cc * BK, cw, ch, img, (uint16_t)s, (uint16_t)r);
              tma_load_2d(smB + stage * C::kStageBytesB, &tmB, &full[stage], (r * g.S + s) * g.C + cc * BK, nt * BN);
              if (++stage == C::kStages) { stage = 0; phase ^= 1; }
            }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BM, BN, sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_sw128_desc(smem_u32(smA + stage * C::kStageBytesA), 1024, 16);
          const uint64_t db = make_sw128_desc(smem_u32(smB + stage * C::kStageBytesB), 1024, 16);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)   // +32 bytes along K per UMMA_K=16 step: +2 in the (addr >> 4) field
            umma_f16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          umma_commit(&empty[stage]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue: TMEM -> registers -> global =================
    const int quarter = warp & 3;                 // TMEM lanes [32*quarter, +32) are the ones this warp may read
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int mt = tile / g.n_tiles, nt = tile % g.n_tiles;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const long row = (long)mt * BM + quarter * 32 + lane;
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);
      T* orow = out + row * g.K + (long)nt * BN;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
        if (row < g.M) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            if (nt * BN + c0 + j < g.K) {
              uint4 o;
              o.x = Cvt<T>::pack2(__uint_as_float(v[j + 0]), __uint_as_float(v[j + 1]));
              o.y = Cvt<T>::pack2(__uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
              o.z = Cvt<T>::pack2(__uint_as_float(v[j + 4]), __uint_as_float(v[j + 5]));
              o.w = Cvt<T>::pack2(__uint_as_float(v[j + 6]), __uint_as_float(v[j + 7]));
              *reinterpret_cast<uint4*>(orow + c0 + j) = o;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ---- host side: tensor maps ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn g_encode_tiled = nullptr;
EncodeIm2colFn g_encode_im2col = nullptr;

int load_driver_fns() {
  if (g_encode_tiled && g_encode_im2col) return 0;
  cudaDriverEntryPointQueryResult q;
  void* fn = nullptr;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
  g_encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
  g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
  return 0;
}

CUtensorMapDataType tm_dtype(int dtype) { return dtype == PDT_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; }

template <int BN>
int launch_fprop(const void* x, const void* w, void* y, int dtype, const ConvGeom& g, cudaStream_t stream) {
  using C = Cfg<BN>;
  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[4] = {(cuuint64_t)g.C, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.N};
    cuuint64_t strides[3] = {(cuuint64_t)g.C * 2, (cuuint64_t)g.W * g.C * 2, (cuuint64_t)g.H * g.W * g.C * 2};
    int lower[2] = {-g.pad, -g.pad};
    int upper[2] = {g.pad - (g.S - 1), g.pad - (g.R - 1)};
    cuuint32_t estr[4] = {1, (cuuint32_t)g.stride, (cuuint32_t)g.stride, 1};
    CUresult r = g_encode_im2col(&tmA, tm_dtype(dtype), 4, const_cast<void*>(x), dims, strides, lower, upper, BK, BM, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 10000 + (int)r;
    // Driver <= 13.1 mis-encodes im2col maps of tensors smaller than 128 KiB: clear bit 21 of the
    // second descriptor word (the same fix-up CUTLASS applies in copy_traits_sm90_im2col.hpp).
    int drv = 0;
    cudaDriverGetVersion(&drv);
    if (drv <= 13010 && (size_t)g.N * g.H * g.W * g.C * 2 < 131072)
      reinterpret_cast<uint64_t*>(&tmA)[1] &= ~(1ull << 21);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)g.R * g.S * g.C, (cuuint64_t)g.K};
    cuuint64_t strides[1] = {(cuuint64_t)g.R * g.S * g.C * 2};
    cuuint32_t box[2] = {BK, (cuuint32_t)BN};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode_tiled(&tmB, tm_dtype(dtype), 2, const_cast<void*>(w), dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 20000 + (int)r;
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int tiles = g.m_tiles * g.n_tiles;
  const int grid = tiles < sms ? tiles : sms;
  if (dtype == PDT_FP16) {
    auto kern = k_conv_igemm<BN, __half>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    kern<<<grid, kThreads, C::kSmemBytes, stream>>>(tmA, tmB, reinterpret_cast<__half*>(y), g);
  } else {
    auto kern = k_conv_igemm<BN, __nv_bfloat16>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    kern<<<grid, kThreads, C::kSmemBytes, stream>>>(tmA, tmB, reinterpret_cast<__nv_bfloat16*>(y), g);
  }
  return (int)cudaGetLastError();
}

}  // namespace

extern "C" {

// x: [N][H][W][C], w: [K][R][S][C], y: [N][P][Q][K]; C % 64 == 0, K % 8 == 0, square filter/stride/pad.
int pdt_conv_fprop(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, int R, int S,
                   int stride, int pad, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16) return PDT_E_UNSUPPORTED;
  if (C % BK || K % 8 || R < 1 || S < 1 || stride < 1 || pad < 0) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  ConvGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  if (g.P < 1 || g.Q < 1) return PDT_E_UNSUPPORTED;
  g.M = N * g.P * g.Q;
  g.m_tiles = (g.M + BM - 1) / BM;
  if (K <= 64) { g.n_tiles = 1; return launch_fprop<64>(x, w, y, dtype, g, stream); }
  if (K <= 128) { g.n_tiles = 1; return launch_fprop<128>(x, w, y, dtype, g, stream); }
  g.n_tiles = (K + 255) / 256;
  return launch_fprop<256>(x, w, y, dtype, g, stream);
}

}  // extern "C"
