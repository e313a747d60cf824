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
#include <cstdlib>
#include <type_traits>
#include "tc.cuh"
#include "fabric.cuh"

namespace {

using namespace pdt;
using namespace pdt::tc;

constexpr int BM = 128;        // output pixels per tile (= TMEM lanes)
// Element-type traits: one k-block is always one 128-byte swizzle row, i.e. 64 bf16/fp16 or 32 tf32
// channels; one tcgen05.mma consumes 32 bytes of it along K (UMMA_K = 16 or 8).
template <typename T> struct Elem {
  static constexpr int kRow = 128 / (int)sizeof(T);
  static constexpr int kUK = 32 / (int)sizeof(T);
  static constexpr int kFmt = std::is_same<T, __nv_bfloat16>::value ? 1 : (std::is_same<T, __half>::value ? 0 : 2);
};
template <typename T>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if constexpr (sizeof(T) == 4) umma_tf32(tmem_d, da, db, idesc, acc);
  else umma_f16(tmem_d, da, db, idesc, acc);
}
template <typename T>
__device__ __forceinline__ void umma_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if constexpr (sizeof(T) == 4) umma_tf32_pair(tmem_d, da, db, idesc, acc);
  else umma_f16_pair(tmem_d, da, db, idesc, acc);
}
constexpr int kThreads = 256;

struct ConvGeom {
  int N, H, W, C;              // input  (NHWC)
  int K, R, S;                 // filter (KRSC)
  int P, Q;                    // output spatial
  int stride;
  int dil_w;                   // horizontal distance (in TMA pixels) between filter taps, normally 1
  int pad_h, pad_w;            // lower (top/left) padding
  int pad_hi_h, pad_hi_w;      // upper (bottom/right) padding
  int M;                       // N*P*Q
  int m_tiles, n_tiles;
  // where output pixel (n,p,q) lands: ((n*OH + p*os + oa)*OW + q*os + ob)*K.  Dense NHWC: OH=P, OW=Q, os=1.
  int OH, OW, os, oa, ob;
  // activation tensor as TMA sees it: Cx channels per pixel, byte strides of w / h / n
  int Cx;
  long long x_sw, x_sh, x_sn;
  // 1: the whole filter (k_blocks tiles) is loaded ONCE into the B ring, slot kb = k-block kb, and stays there
  // (one N tile, k_blocks <= ring depth, no cluster).  The stem (4 k-blocks) otherwise re-fetched 32 KB of
  // filter per 128-pixel tile next to 64 KB of activations and ran at the L2 delivery limit.
  int resident_b;
  // optional epilogue addend, same layout / dtype as the output: out = conv(...) + addend.  The data gradient
  // of a residual block's first convolution adds the block's shortcut gradient here instead of leaving the sum
  // to a separate elementwise pass (3 tensor passes saved per block).
  const void* addend;
};

// Optional epilogue fusion: per-channel sum / sum of squares of the convolution output (the BatchNorm
// batch statistics), so the following BatchNorm needs no pass of its own over the activation.
// Every CTA accumulates its tiles' column sums in shared memory, writes one row of `partial`, and the
// CTA that takes the last ticket folds the rows in fp64 and finalises mean / invstd / running stats.
constexpr int kStatMaxK = 512;
struct BnStats {
  int enabled;
  float* partial;          // [gridDim][2*K] scratch
  unsigned* counter;       // ticket, zero before the launch, reset by the kernel
  float* running_mean;     // may be null
  float* running_var;
  float* save_mean;
  float* save_invstd;
  float momentum, eps;
  Comm comm;               // world > 1: SyncBatchNorm — the last CTA exchanges the raw moments with its peers (fabric.cuh)
};

template <int BN, bool PAIR = false> struct Cfg {
  static constexpr int kStageBytesA = BM * 128;
  static constexpr int kStageBytesB = (PAIR ? BN / 2 : BN) * 128;       // a CTA of a pair holds half of the filter tile
  static constexpr int kStageBytes = kStageBytesA + kStageBytesB;
  static constexpr int kStages = PAIR ? (BN >= 256 ? 6 : 8) : ((BN >= 256) ? 4 : (BN >= 128 ? 6 : 8));
  static constexpr int kTmemCols = 2 * BN;     // double-buffered accumulator (power of two >= 32 for BN in {64,128,256})
  // per epilogue warp: [sum, sumsq][K] accumulators + a 32 x 33 float transpose tile
  static constexpr int kStatBytes = 4 * (2 * kStatMaxK + 32 * 33) * 4;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/ + kStatBytes;
};

// Column sums over the 32 lanes of a warp for 32 values per lane: after the call lane L holds
// sum over lanes of a[L] in a[0].  31 shuffles (butterfly that halves the live values each step).
__device__ __forceinline__ void warp_transpose_reduce(float (&a)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float keep = up ? a[i + s] : a[i];
      const float send = up ? a[i] : a[i + s];
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// CL > 1: launched as clusters of CL CTAs that work on CL consecutive M tiles of the SAME N tile in lock
// step.  The filter tile of a k-block is then fetched once per cluster: CTA r loads rows [r*BN/CL, (r+1)*BN/CL)
// and TMA multicasts them into every CTA's ring slot.  The 256-wide layers were bound by the L2's total
// delivery rate (~42 B/clk per SM when all SMs stream): per k-block a CTA pulled 16 KB of activations and
// 32 KB of filter for 512 clk of MMA; with CL = 4 it is 16 + 8 KB.  A slot may be refilled only when every
// CTA of the cluster has consumed it, so the MMA warp's commit arrives on the `empty` barrier of all CL CTAs.
// Measured on B200 (fp16, 1200 images, in isolation): 256->256 3x3 at 14x14 209.2 -> 196.1 us (1.33 -> 1.42 PFLOP/s),
// 512->512 3x3 at 7x7 197.0 -> 189.1 us (1.47 PFLOP/s), 128->256 3x3/2 125.7 -> 121.4 us; slower for 2-4 k-block 1x1
// projections, which keep the multicast form.  Whole step 25.00 -> 24.82 ms.
// PAIR (requires CL == 2): the two CTAs of a cluster run ONE M = 256 tcgen05.mma.cta_group::2 per k-step instead of two
// M = 128 ones.  Each CTA still loads its own 128-pixel activation tile, but only HALF of the filter tile (rows
// [rank*BN/2, +BN/2), no multicast); both CTAs' loads complete on the leader's `full` barrier, the leader's MMA
// thread issues for the pair, its commits arrive on the `empty` / `tfull` barriers of both CTAs, and every
// epilogue warp of either CTA hands the accumulator back on the leader's `tempty`.
template <int BN, typename T, int CL, bool PAIR = false>
__global__ void __launch_bounds__(kThreads, 1)
k_conv_igemm(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, T* __restrict__ out,
             const float* __restrict__ bias, const ConvGeom g, const BnStats bs) {
  using C = Cfg<BN, PAIR>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + C::kStages * C::kStageBytesA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* tfull = bars + 2 * C::kStages;
  uint64_t* tempty = tfull + 2;
  uint64_t* bres = tempty + 2;                  // resident filter has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bres + 1);
  float* s_stat = reinterpret_cast<float*>(smem + C::kStages * C::kStageBytes + 256);   // [4][2][kStatMaxK]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // work unit = CL consecutive M tiles x one N tile; this CTA takes M tile (unit / n_tiles) * CL + rank.
  // M tiles past the end (last group) run on zero-filled activations and store nothing.
  const int crank = CL > 1 ? (int)cluster_ctarank() : 0;
  const int cluster_id = (int)blockIdx.x / CL, num_clusters = (int)gridDim.x / CL;
  const int num_units = ((g.m_tiles + CL - 1) / CL) * g.n_tiles;
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1);
  constexpr int BK = Elem<T>::kRow;
  const int c_chunks = (g.C + BK - 1) / BK;
  const int k_blocks = g.R * g.S * c_chunks;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  static_assert(!PAIR || CL == 2, "a CTA pair is a cluster of two");
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], PAIR ? 1 : CL); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], PAIR ? 8 : 4); }
    mbar_init(bres, 1);
    fence_barrier_init();
  }
  if (warp == 2) { if constexpr (PAIR) tmem_alloc_pair(tmem_slot, C::kTmemCols); else tmem_alloc(tmem_slot, C::kTmemCols); }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();        // peers' barriers are initialised before anything is multicast at them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();                           // everything above overlapped the previous kernel's tail

  if (warp == 0) {
    // ================= TMA producer (whole warp converged, one elected lane issues) =================
    {
      int stage = 0;
      uint32_t phase = 0;
      const bool resb = CL == 1 && g.resident_b != 0;
      if (resb) {
        if (elect_one_sync()) {
          mbar_expect_tx(bres, (uint32_t)k_blocks * C::kStageBytesB);
          for (int kb = 0; kb < k_blocks; ++kb) {
            const int tap = kb / c_chunks, cc = kb - tap * c_chunks;
            tma_load_2d(smB + kb * C::kStageBytesB, &tmB, bres, tap * g.C + cc * BK, 0);
          }
        }
        __syncwarp();
      }
      for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
        const int mt = (unit / g.n_tiles) * CL + crank, nt = unit % g.n_tiles;
        const int m0 = mt * BM;
        const int img = m0 / (g.P * g.Q);
        const int rem = m0 - img * (g.P * g.Q);
        const int p0 = rem / g.Q, q0 = rem - p0 * g.Q;
        const int cw = q0 * g.stride - g.pad_w, ch = p0 * g.stride - g.pad_h;
        for (int r = 0; r < g.R; ++r)
          for (int s = 0; s < g.S; ++s)
            for (int cc = 0; cc < c_chunks; ++cc) {
              mbar_wait(&empty[stage], phase ^ 1);
              if constexpr (PAIR) {
                if (elect_one_sync()) {
                  // both CTAs' bytes (2 x A tile + the two halves of the B tile) complete on the LEADER's barrier
                  const uint32_t lead_full = mapa_cluster(smem_u32(&full[stage]), 0);
                  if (crank == 0) mbar_expect_tx(&full[stage], 2 * C::kStageBytes);
                  tma_load_im2col_4d_pair(smA + stage * C::kStageBytesA, &tmA, lead_full, cc * BK, cw, ch, img,
                                          (uint16_t)(s * g.dil_w), (uint16_t)r);
                  tma_load_2d_pair(smB + stage * C::kStageBytesB, &tmB, lead_full, (r * g.S + s) * g.C + cc * BK,
                                   nt * BN + crank * (BN / 2));
                }
              } else if (elect_one_sync()) {
                mbar_expect_tx(&full[stage], resb ? C::kStageBytesA : C::kStageBytes);
                tma_load_im2col_4d(smA + stage * C::kStageBytesA, &tmA, &full[stage], cc * BK, cw, ch, img, (uint16_t)(s * g.dil_w), (uint16_t)r);
                if (resb) {
                } else if (CL == 1) {
                  tma_load_2d(smB + stage * C::kStageBytesB, &tmB, &full[stage], (r * g.S + s) * g.C + cc * BK, nt * BN);
                } else {
                  tma_load_2d_mc(smB + stage * C::kStageBytesB + crank * (C::kStageBytesB / CL), &tmB, &full[stage],
                                 (r * g.S + s) * g.C + cc * BK, nt * BN + crank * (BN / CL), kMask);
                }
              }
              __syncwarp();
              if (++stage == C::kStages) { stage = 0; phase ^= 1; }
            }
      }
    }
  } else if (warp == 1 && (!PAIR || crank == 0)) {
    // ================= MMA issuer (whole warp converged, one elected lane issues; the leader CTA of a pair) =================
    {
      constexpr uint32_t idesc = make_idesc(PAIR ? 2 * BM : BM, BN, Elem<T>::kFmt, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      const bool resb = CL == 1 && g.resident_b != 0;
      if (resb) { mbar_wait(bres, 0); tc_fence_after(); }
      for (int unit = cluster_id; unit < num_units; unit += num_clusters, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_sw128_desc(smem_u32(smA + stage * C::kStageBytesA), 1024, 16);
          const uint64_t db = make_sw128_desc(smem_u32(smB + (resb ? kb : stage) * C::kStageBytesB), 1024, 16);
          if (elect_one_sync()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {  // +32 bytes along K per UMMA_K step: +2 in the (addr >> 4) field
              if constexpr (PAIR) umma_pair<T>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
              else umma<T>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            }
            if constexpr (PAIR) umma_commit_pair(&empty[stage], kMask);
            else if (CL == 1) umma_commit(&empty[stage]);
            else umma_commit_mc(&empty[stage], kMask);
          }
          __syncwarp();
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        if (elect_one_sync()) { if constexpr (PAIR) umma_commit_pair(&tfull[acc], kMask); else umma_commit(&tfull[acc]); }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue: TMEM -> registers -> global =================
    const int quarter = warp & 3;                 // TMEM lanes [32*quarter, +32) are the ones this warp may read
    float* my_stat = s_stat + quarter * (2 * kStatMaxK + 32 * 33);
    const uint32_t stat_a = smem_u32(my_stat);
    const uint32_t tp_a = stat_a + 2 * kStatMaxK * 4u;   // 32 x 33 transpose tile
    if (bs.enabled) {
      for (int i = lane; i < 2 * kStatMaxK; i += 32) my_stat[i] = 0.f;
      __syncwarp();
    }
    // N = 64 (the stem, 1x1 projections): each lane keeps its own per-column partial sums for the whole
    // kernel (2 instructions per element, no shared-memory transposes; 128 registers, affordable at 256
    // threads per CTA) and the lanes are combined once at the end.  Wider tiles keep the transpose scheme.
    constexpr bool kRegStats = (BN == 64);
    float rcs[kRegStats ? 64 : 1], rcq[kRegStats ? 64 : 1];
#pragma unroll
    for (int j = 0; j < (kRegStats ? 64 : 1); ++j) rcs[j] = rcq[j] = 0.f;
    int it = 0;
    for (int unit = cluster_id; unit < num_units; unit += num_clusters, ++it) {
      const int mt = (unit / g.n_tiles) * CL + crank, nt = unit % g.n_tiles;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const long row = (long)mt * BM + quarter * 32 + lane;
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);
      long opix = row;
      if (g.os != 1 || g.OH != g.P || g.OW != g.Q) {       // strided / offset placement (parity classes of a strided dgrad)
        const int pq = g.P * g.Q;
        const long n = row / pq;
        const int rem = (int)(row - n * pq);
        const int p = rem / g.Q, q = rem - p * g.Q;
        opix = (n * g.OH + (long)p * g.os + g.oa) * g.OW + (long)q * g.os + g.ob;
      }
      T* orow = out + opix * g.K + (long)nt * BN;
      const T* arow = g.addend != nullptr ? reinterpret_cast<const T*>(g.addend) + opix * g.K + (long)nt * BN : nullptr;
#pragma unroll (kRegStats ? 2 : 1)
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
        if (kRegStats) {
          if (bs.enabled) {       // rows past M are exactly 0 (zero-filled pixels): no mask needed
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float t = __uint_as_float(v[j]);
              rcs[(kRegStats ? c0 : 0) + j] += t;
              rcq[(kRegStats ? c0 : 0) + j] = fmaf(t, t, rcq[(kRegStats ? c0 : 0) + j]);
            }
          }
        } else if (bs.enabled) {
          // rows past M come from zero-filled (out-of-bounds) pixels: their accumulators are exactly 0.
          // Column sums over the warp's 32 rows via a padded shared-memory transpose (conflict-free both ways).
#pragma unroll
          for (int j = 0; j < 32; ++j) sts_f32(tp_a + (uint32_t)(lane * 33 + j) * 4u, __uint_as_float(v[j]));
          __syncwarp();
          float cs = 0.f, cq = 0.f;
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) { const float t = lds_f32(tp_a + (uint32_t)(rr * 33 + lane) * 4u); cs += t; cq = fmaf(t, t, cq); }
          __syncwarp();
          const int col = nt * BN + c0 + lane;
          if (col < g.K) {
            const uint32_t a = stat_a + (uint32_t)col * 4u;
            sts_f32(a, lds_f32(a) + cs);
            sts_f32(a + kStatMaxK * 4u, lds_f32(a + kStatMaxK * 4u) + cq);
          }
        }
        if (row < g.M) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            if (nt * BN + c0 + j < g.K) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j + e]);
              if (bias != nullptr) {
                const float4 b0 = *reinterpret_cast<const float4*>(bias + nt * BN + c0 + j);
                const float4 b1 = *reinterpret_cast<const float4*>(bias + nt * BN + c0 + j + 4);
                f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
                f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
              }
              if (arow != nullptr) {
                if constexpr (sizeof(T) == 4) {
                  const float4 a0 = *reinterpret_cast<const float4*>(arow + c0 + j);
                  const float4 a1 = *reinterpret_cast<const float4*>(arow + c0 + j + 4);
                  f[0] += a0.x; f[1] += a0.y; f[2] += a0.z; f[3] += a0.w;
                  f[4] += a1.x; f[5] += a1.y; f[6] += a1.z; f[7] += a1.w;
                } else {
                  float a[8];
                  unpack8<T>(ldg_stream(arow + c0 + j), a);
#pragma unroll
                  for (int e = 0; e < 8; ++e) f[e] += a[e];
                }
              }
              if constexpr (sizeof(T) == 4) {
                *reinterpret_cast<float4*>(orow + c0 + j) = make_float4(f[0], f[1], f[2], f[3]);
                *reinterpret_cast<float4*>(orow + c0 + j + 4) = make_float4(f[4], f[5], f[6], f[7]);
              } else {
                uint4 o;
                o.x = Cvt<T>::pack2(f[0], f[1]);
                o.y = Cvt<T>::pack2(f[2], f[3]);
                o.z = Cvt<T>::pack2(f[4], f[5]);
                o.w = Cvt<T>::pack2(f[6], f[7]);
                *reinterpret_cast<uint4*>(orow + c0 + j) = o;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR) mbar_arrive_cluster(mapa_cluster(smem_u32(&tempty[acc]), 0));   // the leader's MMA thread waits for both CTAs
        else mbar_arrive(&tempty[acc]);
      }
    }
    if (kRegStats && bs.enabled) {
      float k0 = 0.f, k1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
      for (int j = 0; j < (kRegStats ? 64 : 1); ++j) {
        float a = rcs[j], b = rcq[j];
#pragma unroll
        for (int sft = 16; sft >= 1; sft >>= 1) {
          a += __shfl_xor_sync(0xffffffffu, a, sft);
          b += __shfl_xor_sync(0xffffffffu, b, sft);
        }
        if (lane == (j & 31)) { if (j < 32) { k0 = a; q0 = b; } else { k1 = a; q1 = b; } }
      }
      my_stat[lane] = k0; my_stat[32 + lane] = k1;
      my_stat[kStatMaxK + lane] = q0; my_stat[kStatMaxK + 32 + lane] = q1;
    }
    if (bs.enabled) {
      // ---- fold the 4 warps' column sums, publish this CTA's row, last CTA finalises ----
      __shared__ int s_last;
      const int te = threadIdx.x - 128;           // 0..127 among the epilogue threads
      const int K = g.K;
      epi_bar_sync();
      for (int i = te; i < 2 * K; i += 128) {
        const int which = i / K, c = i - which * K;
        float t = 0.f;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) t += s_stat[q4 * (2 * kStatMaxK + 32 * 33) + which * kStatMaxK + c];
        bs.partial[(size_t)blockIdx.x * 2 * K + i] = t;
      }
      double* tot = reinterpret_cast<double*>(s_stat);        // 2K + 512 doubles <= 12 KB of the 16 KB region
      if (tree_fold(tot, bs.partial, bs.counter, (int)gridDim.x, 2 * K, PDT_BN_MAX_BLOCKS, te, 128, [] { epi_bar_sync(); }, &s_last)) {
        __shared__ double s_cnt;
        if (te == 0) s_cnt = (double)g.M;
        epi_bar_sync();
        if (bs.comm.world > 1) exchange_sum_t(bs.comm, tot, 2 * K, &s_cnt, te, 128, [] { epi_bar_sync(); });
        bn_stats_finalize(tot, K, s_cnt, bs.eps, bs.momentum, bs.save_mean, bs.save_invstd, bs.running_mean, bs.running_var, te, 128);
        if (te == 0) *bs.counter = 0;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();        // no CTA leaves while a peer may still multicast into it or arrive on its barriers
  if (warp == 2) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, C::kTmemCols); else tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ---- weight gradient --------------------------------------------------------------------------------------
// dW[k][r][s][c] = sum_{n,p,q} dy[n,p,q,k] * x[n, p*stride - pad + r, q*stride - pad + s, c]
// as a GEMM over the PIXEL dimension:  D[(r,s,c), k] = sum_pix X_im2col[pix, (r,s,c)] * dY[pix, k].
// Both operands are "MN-major" for the tensor core (the reduction index is the row index in
// memory), which tcgen05 reads directly from 128B-swizzled [pixel][64 channel] tiles:
//   A: two 64-channel blocks (each one filter tap x 64 input channels) via IM2COL TMA, 64 pixels deep;
//   B: BN/64 blocks of 64 output channels via tiled TMA on dy viewed as [pixels][Cout].
// The pixel range is split across CTAs (split-K); partial tiles are accumulated into the fp32
// gradient with red.global.add (the caller zeroes dW first).  TMEM lane = (tap, c) row, so a warp's
// 32 lanes hit 32 consecutive floats of dW for a fixed k: coalesced reductions.
// Replaces cudnnConvolutionBackwardFilter (distributed.py:262).
struct WgradGeom {
  ConvGeom c;
  int rsc;                     // R*S*C
  int m_tiles, n_tiles;        // over (r,s,c) in 128s, over Cout in BN
  int k_steps;                 // ceil(M / 64)
  int splits, steps_per_split;
};

template <int BN, typename T, bool PAIR = false> struct WCfg {
  static constexpr int kRow = Elem<T>::kRow;             // channels per 128-byte block row: 64 (16 bit) / 32 (tf32)
  static constexpr int kPix = 4 * Elem<T>::kUK;          // pixels per k-step (4 MMAs): 64 / 32
  static constexpr int kBlkBytes = kPix * 128;           // one [kPix pixel][128 byte] block
  static constexpr int kABlocks = 128 / kRow;            // blocks forming the M = 128 side
  static constexpr int kBBlocks = BN / kRow;
  static constexpr int kBLocal = PAIR ? kBBlocks / 2 : kBBlocks;     // dy blocks held by ONE CTA (a pair splits the N side)
  static constexpr int kStageBytesA = kABlocks * kBlkBytes;
  static constexpr int kStageBytesB = kBLocal * kBlkBytes;
  static constexpr int kStageBytes = kStageBytesA + kStageBytesB;
  static constexpr int kStages = PAIR ? (BN >= 256 ? 6 : 8) : ((BN >= 256) ? 4 : (BN >= 128 ? 6 : 8));
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

// PAIR: clusters of two CTAs run ONE M = 256 tcgen05.mma.cta_group::2 per k-step: the pair takes two adjacent
// (r,s,c) tiles of the same output-channel tile and pixel split; each CTA loads its own x blocks and HALF of the
// dy blocks (the N side), completing on the leader's barrier — the same scheme as k_conv_igemm<.., PAIR>.
template <int BN, typename T, bool PAIR = false>
__global__ void __launch_bounds__(kThreads, 1)
k_conv_wgrad(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ dw,
             float* __restrict__ part, const WgradGeom wg) {
  using C = WCfg<BN, T, PAIR>;
  static_assert(!PAIR || WCfg<BN, T, false>::kBBlocks % 2 == 0, "a pair splits the dy blocks in halves");
  constexpr int kRow = C::kRow, BKP = C::kPix, kBlkBytes = C::kBlkBytes, kAB = C::kABlocks;
  const ConvGeom& g = wg.c;
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
  // PAIR: a work item is (pair of adjacent M tiles, N tile, pixel split) and belongs to a cluster of two CTAs;
  // CTA `crank` of the cluster owns M tile 2 * pair + crank (past the end for the last odd one: computed, never stored).
  const int crank = PAIR ? (int)cluster_ctarank() : 0;
  const int first_item = PAIR ? (int)blockIdx.x / 2 : (int)blockIdx.x;
  const int item_step = PAIR ? (int)gridDim.x / 2 : (int)gridDim.x;
  const int m_groups = PAIR ? (wg.m_tiles + 1) / 2 : wg.m_tiles;
  const int items = m_groups * wg.n_tiles * wg.splits;
  const int c_chunks = g.C / kRow;
  const int m_blocks = wg.rsc / kRow;   // wgrad needs C % kRow == 0: an (r,s,c) block never straddles two taps

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmX); prefetch_tmap(&tmDY); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], PAIR ? 8 : 4); }
    fence_barrier_init();
  }
  if (warp == 2) { if constexpr (PAIR) tmem_alloc_pair(tmem_slot, C::kTmemCols); else tmem_alloc(tmem_slot, C::kTmemCols); }
  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) cluster_sync_all();      // the peer's barriers exist before anything is signalled at them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();

  if (warp == 0) {
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = first_item; item < items; item += item_step) {
        const int sp = item % wg.splits;
        const int tile = item / wg.splits;
        const int mt = PAIR ? (tile / wg.n_tiles) * 2 + crank : tile / wg.n_tiles, nt = tile % wg.n_tiles;
        const int ks0 = sp * wg.steps_per_split;
        const int ks1 = min(wg.k_steps, ks0 + wg.steps_per_split);
        const int nA = max(0, min(kAB, m_blocks - mt * kAB));
        const int nB = min(C::kBBlocks, (g.K - nt * BN + kRow - 1) / kRow);
        // PAIR: this CTA loads dy blocks [jb0, jb1) of the N tile into its own ring slot; the leader expects the
        // bytes of both CTAs (its peer's M tile may be short or absent)
        const int jb0 = PAIR ? min(nB, crank * C::kBLocal) : 0;
        const int jb1 = PAIR ? min(nB, (crank + 1) * C::kBLocal) : nB;
        const int nA_peer = PAIR ? max(0, min(kAB, m_blocks - (mt ^ 1) * kAB)) : 0;
        int tapr[kAB], taps[kAB], cc[kAB];
        for (int b = 0; b < kAB; ++b) {
          const int gb = min(mt * kAB + b, m_blocks - 1);
          const int tap = gb / c_chunks;
          cc[b] = gb - tap * c_chunks;
          tapr[b] = tap / g.S;
          taps[b] = tap - tapr[b] * g.S;
        }
        for (int ks = ks0; ks < ks1; ++ks) {
          const int m0 = ks * BKP;
          const int img = m0 / (g.P * g.Q);
          const int rem = m0 - img * (g.P * g.Q);
          const int p0 = rem / g.Q, q0 = rem - p0 * g.Q;
          const int cw = q0 * g.stride - g.pad_w, ch = p0 * g.stride - g.pad_h;
          mbar_wait(&empty[stage], phase ^ 1);
          if constexpr (PAIR) {
            if (elect_one_sync()) {
              const uint32_t lead_full = mapa_cluster(smem_u32(&full[stage]), 0);
              if (crank == 0) mbar_expect_tx(&full[stage], (uint32_t)(nA + nA_peer + nB) * kBlkBytes);
#pragma unroll
              for (int b = 0; b < kAB; ++b)
                if (b < nA)
                  tma_load_im2col_4d_pair(smA + stage * C::kStageBytesA + b * kBlkBytes, &tmX, lead_full, cc[b] * kRow, cw, ch, img,
                                          (uint16_t)(taps[b] * g.dil_w), (uint16_t)tapr[b]);
              for (int j = jb0; j < jb1; ++j)
                tma_load_2d_pair(smB + stage * C::kStageBytesB + (j - jb0) * kBlkBytes, &tmDY, lead_full, nt * BN + j * kRow, m0);
            }
          } else if (elect_one_sync()) {
            mbar_expect_tx(&full[stage], (uint32_t)(nA + nB) * kBlkBytes);
#pragma unroll
            for (int b = 0; b < kAB; ++b)
              if (b < nA)
                tma_load_im2col_4d(smA + stage * C::kStageBytesA + b * kBlkBytes, &tmX, &full[stage], cc[b] * kRow, cw, ch, img,
                                   (uint16_t)(taps[b] * g.dil_w), (uint16_t)tapr[b]);
            for (int j = 0; j < nB; ++j)
              tma_load_2d(smB + stage * C::kStageBytesB + j * kBlkBytes, &tmDY, &full[stage], nt * BN + j * kRow, m0);
          }
          __syncwarp();
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1 && (!PAIR || crank == 0)) {
    {
      constexpr uint32_t idesc = make_idesc(PAIR ? 256 : 128, BN, Elem<T>::kFmt, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int item = first_item; item < items; item += item_step, ++it) {
        const int sp = item % wg.splits;
        const int ks0 = sp * wg.steps_per_split;
        const int ks1 = min(wg.k_steps, ks0 + wg.steps_per_split);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int ks = ks0; ks < ks1; ++ks) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          // MN-major, 128B swizzle: 8-pixel groups 1024 B apart (SBO), 64-channel blocks kBlkBytes apart (LBO)
          // 32-bit MN-major operands use the 32-byte-atom swizzle: 4-pixel groups 512 B apart
          constexpr uint32_t kSbo = sizeof(T) == 4 ? 512 : 1024, kLayout = sizeof(T) == 4 ? 1 : 2;
          const uint64_t da = make_sw128_desc(smem_u32(smA + stage * C::kStageBytesA), kSbo, kBlkBytes, kLayout);
          const uint64_t db = make_sw128_desc(smem_u32(smB + stage * C::kStageBytesB), kSbo, kBlkBytes, kLayout);
          if (elect_one_sync()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {  // UMMA_K pixels = kUK * 128 bytes per step, in (addr >> 4) units
              if constexpr (PAIR)
                umma_pair<T>(tmem_d, da + (Elem<T>::kUK * 8) * k, db + (Elem<T>::kUK * 8) * k, idesc, (ks > ks0 || k > 0) ? 1u : 0u);
              else
                umma<T>(tmem_d, da + (Elem<T>::kUK * 8) * k, db + (Elem<T>::kUK * 8) * k, idesc, (ks > ks0 || k > 0) ? 1u : 0u);
            }
            if constexpr (PAIR) umma_commit_pair(&empty[stage], (uint16_t)3); else umma_commit(&empty[stage]);
          }
          __syncwarp();
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        if (elect_one_sync()) { if constexpr (PAIR) umma_commit_pair(&tfull[acc], (uint16_t)3); else umma_commit(&tfull[acc]); }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    const int quarter = warp & 3;
    int it = 0;
    for (int item = first_item; item < items; item += item_step, ++it) {
      const int tile = item / wg.splits;
      const int mt = PAIR ? (tile / wg.n_tiles) * 2 + crank : tile / wg.n_tiles, nt = tile % wg.n_tiles;
      const int sp = item % wg.splits;
      const bool nonempty = sp * wg.steps_per_split < wg.k_steps;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const int m = mt * 128 + quarter * 32 + lane;           // (tap, c) index == column of dW's [Cout][R*S*C] view
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
        if (m < wg.rsc && nonempty) {
          if (part != nullptr) {
            // deterministic mode: every split stores its partial tile into its own slab; k_wgrad_fold adds the
            // slabs in split order afterwards (no floating-point atomics anywhere)
            float* slab = part + (size_t)sp * ((size_t)g.K * wg.rsc);
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int k = nt * BN + c0 + j;
              if (k < g.K) slab[(size_t)k * wg.rsc + m] = __uint_as_float(v[j]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int k = nt * BN + c0 + j;
              if (k < g.K) atomicAdd(dw + (size_t)k * wg.rsc + m, __uint_as_float(v[j]));
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR) mbar_arrive_cluster(mapa_cluster(smem_u32(&tempty[acc]), 0));
        else mbar_arrive(&tempty[acc]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) cluster_sync_all();      // no CTA leaves while its peer may still signal its barriers / write its TMEM
  if (warp == 2) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, C::kTmemCols); else tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// Ordered fold of the split-K slabs of the deterministic weight gradient: dw[i] += part[0][i] + part[1][i] + ...
__global__ void __launch_bounds__(256) k_wgrad_fold(float* __restrict__ dw, const float* __restrict__ part, size_t total, int splits) {
  pdl_enter();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float a = 0.f;
    for (int s = 0; s < splits; ++s) a += part[(size_t)s * total + i];
    dw[i] += a;
  }
}

// Per-device workspace for the deterministic weight gradient (pdt_set_wgrad_workspace); null = atomics.
struct WgradWs { float* ptr; size_t floats; };
WgradWs g_wgrad_ws[64];

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

CUtensorMapDataType tm_dtype(int dtype) {
  return dtype == PDT_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                           : (dtype == PDT_FP32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
}
int esize_of(int dtype) { return dtype == PDT_FP32 ? 4 : 2; }

// Cluster size for the forward / data-gradient kernel: multicast pays when the filter tile is at least as
// large as the activation tile (BN >= 128) and there are enough M tiles to keep every cluster busy.
// PDT_CONV_CLUSTER=1|2|4 overrides.
int cluster_size_for(int BN, const ConvGeom& g, const void* filter, cudaStream_t stream) {
  static int env = -1;
  if (env < 0) { const char* e = getenv("PDT_CONV_CLUSTER"); env = e ? atoi(e) : 0; }
  int cl = env > 0 ? env : 2;     // 4 measured the same per launch but only 33 clusters of 4 are co-resident (132 of 148 SMs)
  if (BN < 128 || (cl != 2 && cl != 4)) return 1;
  // fewer M tiles than SMs: a multicast cluster would leave SMs idle, but a CTA PAIR (cta_group::2, no multicast load) keeps
  // the same number of SMs busy as single CTAs and runs the wider instruction: -2 = "pairs only" (layer 4 at 150 images
  // per GPU: 58 M tiles)
  if (g.m_tiles < 148) return (cl == 2 && g.m_tiles >= 2) ? -2 : 1;
  if (!memory_on_current_device(filter, stream)) return cl == 2 ? -2 : 1;  // a multicast tile (the filter) must be this GPU's memory
  return cl;
}

int conv_pair_enabled() {
  static int pair = -1;
  if (pair < 0) { const char* e = getenv("PDT_CONV_PAIR"); pair = (e != nullptr && e[0] == '0') ? 0 : 1; }
  return pair;
}

template <int BN, typename T, int CL, bool PAIR = false>
int launch_igemm_cl(const CUtensorMap& tmA, const CUtensorMap& tmB, void* y, const float* bias, const ConvGeom& g, const BnStats& bs,
                    int sms, cudaStream_t stream) {
  using C = Cfg<BN, PAIR>;
  auto kern = k_conv_igemm<BN, T, CL, PAIR>;
  PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  const int units = ((g.m_tiles + CL - 1) / CL) * g.n_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int nattr = 0;
  int clusters = sms / CL;
  if (CL > 1) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    nattr = 1;
    // a persistent kernel must fit in ONE wave: ask how many clusters can be resident (GPC boundaries make
    // this less than SMs / CL on some parts) instead of assuming
    static int cached[3] = {-1, -1, -1};      // (static per template instance: PAIR and multicast variants each keep their own)
    int& mc = cached[CL == 2 ? 1 : 2];
    if (mc < 0) {
      cfg.gridDim = dim3(sms / CL * CL);
      cfg.attrs = attr; cfg.numAttrs = nattr;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n < 1) { cudaGetLastError(); n = sms / CL / 2; }
      mc = n;
    }
    if (clusters > mc) clusters = mc;
  }
  if (clusters > units) clusters = units;
  cfg.gridDim = dim3(clusters * CL);
  nattr += pdl_attr(attr + nattr);
  cfg.attrs = attr; cfg.numAttrs = nattr;
  ConvGeom gg = g;
  {
    static int env = -1;                       // PDT_CONV_RESIDENT_B=0 keeps streaming the filter
    if (env < 0) { const char* e = getenv("PDT_CONV_RESIDENT_B"); env = (e != nullptr && e[0] == '0') ? 0 : 1; }
    const int bk = 128 / (int)sizeof(T);
    const int k_blocks = g.R * g.S * ((g.C + bk - 1) / bk);
    gg.resident_b = (env == 1 && CL == 1 && g.n_tiles == 1 && k_blocks <= C::kStages) ? 1 : 0;
  }
  PDT_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, reinterpret_cast<T*>(y), bias, gg, bs));
  return (int)cudaGetLastError();
}

template <int BN, typename T>
int launch_igemm(const CUtensorMap& tmA, const CUtensorMap& tmB, void* y, const float* bias, const ConvGeom& g, const BnStats& bs, int cl,
                 int sms, cudaStream_t stream) {
  if constexpr (BN >= 128) {
    if (cl == 4) return launch_igemm_cl<BN, T, 4>(tmA, tmB, y, bias, g, bs, sms, stream);
    if (cl == 2) {
      const int pair = conv_pair_enabled();  // CTA pairs run one M = 256 cta_group::2 MMA (see k_conv_igemm); PDT_CONV_PAIR=0: multicast pairs
      // pairs pay when the k loop is long enough to amortise the pair's hand-shakes (3x3 layers: 18-72 k-blocks;
      // measured slower on 1x1 projections with 2-4 k-blocks)
      const int kb = g.R * g.S * ((g.C + (128 / (int)sizeof(T)) - 1) / (128 / (int)sizeof(T)));
      if (pair == 1 && kb >= 9) return launch_igemm_cl<BN, T, 2, true>(tmA, tmB, y, bias, g, bs, sms, stream);
      return launch_igemm_cl<BN, T, 2>(tmA, tmB, y, bias, g, bs, sms, stream);
    }
  }
  return launch_igemm_cl<BN, T, 1>(tmA, tmB, y, bias, g, bs, sms, stream);
}

template <int BN>
int launch_fprop(const void* x, const void* w, void* y, const float* bias, int dtype, const ConvGeom& g, const BnStats& bs,
                 cudaStream_t stream) {
  CUtensorMap tmA, tmB;
  int cl = cluster_size_for(BN, g, w, stream);
  if (cl == -2) {                        // "pairs only": the same condition launch_igemm uses to pick the pair form
    const int bk = 128 / esize_of(dtype);
    const int kb = g.R * g.S * ((g.C + bk - 1) / bk);
    cl = (conv_pair_enabled() == 1 && kb >= 9) ? 2 : 1;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)g.Cx, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.N};
    cuuint64_t strides[3] = {(cuuint64_t)g.x_sw, (cuuint64_t)g.x_sh, (cuuint64_t)g.x_sn};
    int lower[2] = {-g.pad_w, -g.pad_h};
    int upper[2] = {g.pad_hi_w - (g.S - 1) * g.dil_w, g.pad_hi_h - (g.R - 1)};
    cuuint32_t estr[4] = {1, (cuuint32_t)g.stride, (cuuint32_t)g.stride, 1};
    const int BK = 128 / esize_of(dtype);
    CUresult r = g_encode_im2col(&tmA, tm_dtype(dtype), 4, const_cast<void*>(x), dims, strides, lower, upper, BK, BM, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 10000 + (int)r;
    // Driver <= 13.1 mis-encodes im2col maps of tensors smaller than 128 KiB: clear bit 21 of the
    // second descriptor word (the same fix-up CUTLASS applies in copy_traits_sm90_im2col.hpp).
    int drv = 0;
    cudaDriverGetVersion(&drv);
    if (drv <= 13010 && (size_t)g.N * (size_t)g.x_sn < 131072)
      reinterpret_cast<uint64_t*>(&tmA)[1] &= ~(1ull << 21);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)g.R * g.S * g.C, (cuuint64_t)g.K};
    cuuint64_t strides[1] = {(cuuint64_t)g.R * g.S * g.C * esize_of(dtype)};
    cuuint32_t box[2] = {(cuuint32_t)(128 / esize_of(dtype)), (cuuint32_t)(BN / cl)};   // one CTA's slice of the filter tile
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode_tiled(&tmB, tm_dtype(dtype), 2, const_cast<void*>(w), dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 20000 + (int)r;
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (dtype == PDT_FP16) return launch_igemm<BN, __half>(tmA, tmB, y, bias, g, bs, cl, sms, stream);
  if (dtype == PDT_FP32) return launch_igemm<BN, float>(tmA, tmB, y, bias, g, bs, cl, sms, stream);
  return launch_igemm<BN, __nv_bfloat16>(tmA, tmB, y, bias, g, bs, cl, sms, stream);
}

template <int BN>
int launch_wgrad(const void* x, const void* dy, float* dw, int dtype, WgradGeom& wg, cudaStream_t stream) {
  const ConvGeom& g = wg.c;
  const int es = esize_of(dtype);
  const int kRow = 128 / es, BKP = 4 * (32 / es);
  const int smem_bytes = (dtype == PDT_FP32) ? WCfg<BN, float>::kSmemBytes : WCfg<BN, __half>::kSmemBytes;
  // MN-major 32-bit operands need the 32-byte-atom flavour of the 128-byte swizzle (see tc.cuh)
  const CUtensorMapSwizzle swz = (dtype == PDT_FP32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUtensorMap tmX, tmDY;
  {
    cuuint64_t dims[4] = {(cuuint64_t)g.Cx, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.N};
    cuuint64_t strides[3] = {(cuuint64_t)g.x_sw, (cuuint64_t)g.x_sh, (cuuint64_t)g.x_sn};
    int lower[2] = {-g.pad_w, -g.pad_h};
    int upper[2] = {g.pad_hi_w - (g.S - 1) * g.dil_w, g.pad_hi_h - (g.R - 1)};
    cuuint32_t estr[4] = {1, (cuuint32_t)g.stride, (cuuint32_t)g.stride, 1};
    CUresult r = g_encode_im2col(&tmX, tm_dtype(dtype), 4, const_cast<void*>(x), dims, strides, lower, upper, kRow, BKP, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 10000 + (int)r;
    int drv = 0;
    cudaDriverGetVersion(&drv);
    if (drv <= 13010 && (size_t)g.N * (size_t)g.x_sn < 131072)
      reinterpret_cast<uint64_t*>(&tmX)[1] &= ~(1ull << 21);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)g.K, (cuuint64_t)g.M};
    cuuint64_t strides[1] = {(cuuint64_t)g.K * es};
    cuuint32_t box[2] = {(cuuint32_t)kRow, (cuuint32_t)BKP};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode_tiled(&tmDY, tm_dtype(dtype), 2, const_cast<void*>(dy), dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 20000 + (int)r;
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // CTA pairs (one M = 256 cta_group::2 MMA over two adjacent (r,s,c) tiles): PDT_WGRAD_PAIR=0 keeps single CTAs
  // (measured per launch at 1200 images: BN = 256 tiles +12 %, BN = 128 tiles -2.5 %: only the wide tiles pair up)
  bool pair = false;
  if constexpr (BN >= 256) {
    static int env = -1;
    if (env < 0) { const char* e = getenv("PDT_WGRAD_PAIR"); env = (e != nullptr && e[0] == '0') ? 0 : 1; }
    pair = env == 1 && wg.m_tiles >= 2;
  }
  int units = sms;                                            // co-resident work units: CTAs, or clusters of two
  if (pair) {
    static int cached = -1;                                   // clusters of 2 that fit at once (GPC boundaries decide)
    if (cached < 0) {
      int n = 0;
      if constexpr (BN >= 256) {
        auto kern = k_conv_wgrad<BN, __half, true>;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, WCfg<BN, __half, true>::kSmemBytes);
        cudaLaunchConfig_t q = {};
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
        q.gridDim = dim3(sms / 2 * 2); q.blockDim = dim3(kThreads);
        q.dynamicSmemBytes = WCfg<BN, __half, true>::kSmemBytes; q.attrs = qa; q.numAttrs = 1;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &q) != cudaSuccess || n < 1) { cudaGetLastError(); n = sms / 4; }
      }
      cached = n < sms / 2 ? n : sms / 2;
    }
    units = cached;
  }
  const int tiles = (pair ? (wg.m_tiles + 1) / 2 : wg.m_tiles) * wg.n_tiles;
  int splits = units / tiles;                                 // one wave: never more work items than resident units
  if (splits > wg.k_steps) splits = wg.k_steps;
  if (splits < 1) splits = 1;
  wg.steps_per_split = (wg.k_steps + splits - 1) / splits;
  wg.splits = (wg.k_steps + wg.steps_per_split - 1) / wg.steps_per_split;
  // deterministic mode (a workspace is registered for this device): split-K partials go to slabs of the
  // workspace and are folded in order; the split count shrinks to what the workspace holds
  float* part = nullptr;
  const size_t total = (size_t)wg.c.K * wg.rsc;
  if (dev >= 0 && dev < 64 && g_wgrad_ws[dev].ptr != nullptr && wg.splits > 1) {
    const size_t fit = g_wgrad_ws[dev].floats / total;
    if (fit < 2) {
      wg.steps_per_split = wg.k_steps;           // one writer per element: deterministic without slabs
      wg.splits = 1;
    } else {
      if ((size_t)wg.splits > fit) {
        wg.steps_per_split = (wg.k_steps + (int)fit - 1) / (int)fit;
        wg.splits = (wg.k_steps + wg.steps_per_split - 1) / wg.steps_per_split;
      }
      if (wg.splits > 1) part = g_wgrad_ws[dev].ptr;
    }
  }
  const int items = tiles * wg.splits;
  const int grid = (items < units ? items : units) * (pair ? 2 : 1);
  auto go = [&](auto kern, int smem) -> int {
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (!pair) {
      PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(kThreads), smem, stream, tmX, tmDY, dw, part, wg));
      return 0;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    int nattr = 1 + pdl_attr(attr + 1);
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cfg.attrs = attr; cfg.numAttrs = nattr;
    PDT_CHECK(cudaLaunchKernelEx(&cfg, kern, tmX, tmDY, dw, part, wg));
    return 0;
  };
  int rc = 0;
  if constexpr (BN >= 256) {
    if (pair) {
      if (dtype == PDT_FP16) rc = go(k_conv_wgrad<BN, __half, true>, WCfg<BN, __half, true>::kSmemBytes);
      else if (dtype == PDT_FP32) rc = go(k_conv_wgrad<BN, float, true>, WCfg<BN, float, true>::kSmemBytes);
      else rc = go(k_conv_wgrad<BN, __nv_bfloat16, true>, WCfg<BN, __nv_bfloat16, true>::kSmemBytes);
    }
  }
  if (!pair) {
    if (dtype == PDT_FP16) rc = go(k_conv_wgrad<BN, __half>, smem_bytes);
    else if (dtype == PDT_FP32) rc = go(k_conv_wgrad<BN, float>, smem_bytes);
    else rc = go(k_conv_wgrad<BN, __nv_bfloat16>, smem_bytes);
  }
  if (rc != 0) return rc;
  if (part != nullptr) {
    PDT_CHECK(cudaGetLastError());
    const int fb = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
    k_wgrad_fold<<<fb, 256, 0, stream>>>(dw, part, total, wg.splits);
  }
  return (int)cudaGetLastError();
}

void fill_dense(ConvGeom& g, int N, int H, int W, int C, int K, int R, int S, int stride, int pad, int es) {
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.dil_w = 1;
  g.pad_h = g.pad_w = g.pad_hi_h = g.pad_hi_w = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  g.M = N * g.P * g.Q;
  g.m_tiles = (g.M + BM - 1) / BM;
  g.n_tiles = 0;
  g.OH = g.P; g.OW = g.Q; g.os = 1; g.oa = g.ob = 0;
  g.Cx = C;
  g.x_sw = (long long)C * es; g.x_sh = (long long)W * C * es; g.x_sn = (long long)H * W * C * es;
  g.addend = nullptr;
}

int dispatch_fprop(const void* x, const void* w, void* y, const float* bias, int dtype, ConvGeom& g, cudaStream_t stream,
                   const BnStats* stats = nullptr) {
  BnStats bs;
  if (stats != nullptr) bs = *stats; else { bs = BnStats(); bs.enabled = 0; }
  if (g.K <= 64) { g.n_tiles = 1; return launch_fprop<64>(x, w, y, bias, dtype, g, bs, stream); }
  if (g.K <= 128) { g.n_tiles = 1; return launch_fprop<128>(x, w, y, bias, dtype, g, bs, stream); }
  g.n_tiles = (g.K + 255) / 256;
  return launch_fprop<256>(x, w, y, bias, dtype, g, bs, stream);
}

bool wgrad_halo_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("PDT_WGRAD_HALO"); v = (e != nullptr && e[0] == '0') ? 0 : 1; }
  if (v != 1) return false;
  int dev = 0;
  cudaGetDevice(&dev);      // the patch-reuse kernel accumulates with red.add: not used in deterministic mode
  return !(dev >= 0 && dev < 64 && g_wgrad_ws[dev].ptr != nullptr);
}

}  // namespace

extern "C" {

// Deterministic weight gradients for the CURRENT device: with a workspace registered, split-K partial tiles
// are stored into slabs of it and folded in a fixed order (and the atomics-based patch-reuse kernel is not
// used), so two runs produce bit-identical gradients (--seed, distributed.py:91-99: cudnn.deterministic).
// ws == NULL switches back to the faster red.global.add accumulation.  Launches that share the workspace
// must be stream-ordered.
int pdt_set_wgrad_workspace(void* ws, size_t bytes) {
  int dev = 0;
  PDT_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  if ((reinterpret_cast<uintptr_t>(ws) & 15) != 0) return (int)cudaErrorInvalidValue;
  g_wgrad_ws[dev].ptr = reinterpret_cast<float*>(ws);
  g_wgrad_ws[dev].floats = ws ? bytes / sizeof(float) : 0;
  return 0;
}

// x: [N][H][W][C], dy: [N][P][Q][K] (bf16/fp16), dw: [K][R][S][C] fp32, ACCUMULATED into (zero it first).
int pdt_conv_wgrad(const void* x, const void* dy, float* dw, int dtype, int N, int H, int W, int C, int K, int R, int S,
                   int stride, int pad, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  const int es = esize_of(dtype), BKP = 4 * (32 / es);
  if (C % (128 / es) || K % 8 || R < 1 || S < 1 || stride < 1 || pad < 0) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dw)) & 15) return PDT_E_UNSUPPORTED;
  if (R == 3 && S == 3 && stride == 1 && pad == 1 && wgrad_halo_enabled() && pdt_conv3x3_wgrad_halo_supported(dtype, N, H, W, C, K))
    return pdt_conv3x3_wgrad_halo(x, dy, dw, dtype, N, H, W, C, K, stream);
  int e = load_driver_fns();
  if (e) return e;
  WgradGeom wg;
  fill_dense(wg.c, N, H, W, C, K, R, S, stride, pad, es);
  if (wg.c.P < 1 || wg.c.Q < 1) return PDT_E_UNSUPPORTED;
  wg.rsc = R * S * C;
  wg.m_tiles = (wg.rsc + 127) / 128;
  wg.k_steps = (wg.c.M + BKP - 1) / BKP;
  if (K <= 64) { wg.n_tiles = 1; return launch_wgrad<64>(x, dy, dw, dtype, wg, stream); }
  if (K <= 128) { wg.n_tiles = 1; return launch_wgrad<128>(x, dy, dw, dtype, wg, stream); }
  wg.n_tiles = (K + 255) / 256;
  return launch_wgrad<256>(x, dy, dw, dtype, wg, stream);
}

// General form of the weight gradient: explicit activation strides / per-pixel channel count, separate
// lower/upper padding and output size (the space-to-depth stem, see pdt_conv_fprop_ex).
int pdt_conv_wgrad_ex(const void* x, const void* dy, float* dw, int dtype, int N, int H, int W, int C, int Cx,
                      long long x_sw, long long x_sh, long long x_sn, int K, int R, int S, int stride, int pad_lo_h,
                      int pad_lo_w, int pad_hi_h, int pad_hi_w, int P, int Q, int dil_w, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  const int es = esize_of(dtype), BKP = 4 * (32 / es);
  if (C % (128 / es) || K % 8 || R < 1 || S < 1 || stride < 1 || P < 1 || Q < 1 || dil_w < 1 || (x_sw | x_sh | x_sn) % 16) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dw)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  WgradGeom wg;
  ConvGeom& g = wg.c;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.dil_w = dil_w;
  g.pad_h = pad_lo_h; g.pad_w = pad_lo_w; g.pad_hi_h = pad_hi_h; g.pad_hi_w = pad_hi_w;
  g.P = P; g.Q = Q;
  g.M = N * P * Q;
  g.m_tiles = g.n_tiles = 0;
  g.OH = P; g.OW = Q; g.os = 1; g.oa = g.ob = 0;
  g.Cx = Cx; g.x_sw = x_sw; g.x_sh = x_sh; g.x_sn = x_sn;
  g.addend = nullptr;
  wg.rsc = R * S * C;
  wg.m_tiles = (wg.rsc + 127) / 128;
  wg.k_steps = (g.M + BKP - 1) / BKP;
  if (K <= 64) { wg.n_tiles = 1; return launch_wgrad<64>(x, dy, dw, dtype, wg, stream); }
  if (K <= 128) { wg.n_tiles = 1; return launch_wgrad<128>(x, dy, dw, dtype, wg, stream); }
  wg.n_tiles = (K + 255) / 256;
  return launch_wgrad<256>(x, dy, dw, dtype, wg, stream);
}

// x: [N][H][W][C], w: [K][R][S][C], y: [N][P][Q][K]; C % 8 == 0, K % 8 == 0, square stride/pad.
int pdt_conv_fprop(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, int R, int S,
                   int stride, int pad, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  const int es = esize_of(dtype);
  if (C % (16 / es) || K % 8 || R < 1 || S < 1 || stride < 1 || pad < 0) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  ConvGeom g;
  fill_dense(g, N, H, W, C, K, R, S, stride, pad, es);
  if (g.P < 1 || g.Q < 1) return PDT_E_UNSUPPORTED;
  return dispatch_fprop(x, w, y, nullptr, dtype, g, stream);
}

// y = conv(x, w) + addend (addend: [N][P][Q][K] like y).  Used as the data gradient of the first convolution of
// a residual block with the block's shortcut gradient as addend (autograd's AccumulateGrad add, an ATen
// elementwise kernel per block in the reference's backward, distributed.py:262).
int pdt_conv_fprop_add(const void* x, const void* w, void* y, const void* addend, int dtype, int N, int H, int W, int C, int K,
                       int R, int S, int stride, int pad, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  const int es = esize_of(dtype);
  if (C % (16 / es) || K % 8 || R < 1 || S < 1 || stride < 1 || pad < 0) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y) |
       reinterpret_cast<uintptr_t>(addend)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  ConvGeom g;
  fill_dense(g, N, H, W, C, K, R, S, stride, pad, es);
  if (g.P < 1 || g.Q < 1) return PDT_E_UNSUPPORTED;
  g.addend = addend;
  return dispatch_fprop(x, w, y, nullptr, dtype, g, stream);
}

// Convolution + BatchNorm batch statistics in one kernel (dense NHWC output, K <= 512).  scratch:
// pdt_bn_scratch_bytes(K) bytes (CTA rows + group rows of the two-level fold), first 256 bytes zero before the first call.
int pdt_conv_fprop_bnstats(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, int R, int S,
                           int stride, int pad, void* scratch, const PdtCommDesc* comm, float* running_mean, float* running_var,
                           float momentum, float eps, float* save_mean, float* save_invstd, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  const int es = esize_of(dtype);
  if (C % (16 / es) || K % 8 || K > kStatMaxK || R < 1 || S < 1 || stride < 1 || pad < 0) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  ConvGeom g;
  fill_dense(g, N, H, W, C, K, R, S, stride, pad, es);
  if (g.P < 1 || g.Q < 1) return PDT_E_UNSUPPORTED;
  BnStats bs;
  bs.enabled = 1;
  bs.counter = reinterpret_cast<unsigned*>(scratch);
  bs.partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  bs.running_mean = running_mean; bs.running_var = running_var;
  bs.save_mean = save_mean; bs.save_invstd = save_invstd;
  bs.momentum = momentum; bs.eps = eps;
  bs.comm = make_comm(comm);
  return dispatch_fprop(x, w, y, nullptr, dtype, g, stream, &bs);
}

// General form.  Adds: an fp32 bias [K] in the epilogue (nn.Linear as a 1x1 convolution); separate
// lower/upper padding and an explicit output size P x Q; placement of output pixel (p,q) at
// (p*os + oa, q*os + ob) of an OH x OW image (one parity class of a strided data gradient);
// explicit byte strides and per-pixel channel count of the activation as TMA should see it
// (overlapping "pixels" let a space-to-depth stem read 4 neighbouring pixels as one 64-channel row).
int pdt_conv_fprop_ex(const void* x, const void* w, void* y, const float* bias, int dtype, int N, int H, int W, int C,
                      int Cx, long long x_sw, long long x_sh, long long x_sn, int K, int R, int S, int stride,
                      int pad_lo_h, int pad_lo_w, int pad_hi_h, int pad_hi_w, int P, int Q, int OH, int OW, int os,
                      int oa, int ob, int dil_w, const void* addend, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  if (reinterpret_cast<uintptr_t>(addend) & 15) return PDT_E_UNSUPPORTED;
  if (K % 8 || R < 1 || S < 1 || stride < 1 || P < 1 || Q < 1 || dil_w < 1 || (x_sw | x_sh | x_sn) % 16) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  ConvGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.dil_w = dil_w;
  g.pad_h = pad_lo_h; g.pad_w = pad_lo_w; g.pad_hi_h = pad_hi_h; g.pad_hi_w = pad_hi_w;
  g.P = P; g.Q = Q;
  g.M = N * P * Q;
  g.m_tiles = (g.M + BM - 1) / BM;
  g.OH = OH; g.OW = OW; g.os = os; g.oa = oa; g.ob = ob;
  g.Cx = Cx; g.x_sw = x_sw; g.x_sh = x_sh; g.x_sn = x_sn;
  g.addend = addend;
  return dispatch_fprop(x, w, y, bias, dtype, g, stream);
}

// The general form with the BatchNorm-statistics epilogue (dense output only: os == 1, OH x OW == P x Q, no addend, no bias;
// K <= 512).  The fp32 stem (a 4x4 convolution over a strided view of the space-to-depth image) uses it so that its
// BatchNorm needs no statistics pass over the 112 x 112 x 64 output.  scratch / statistics as in pdt_conv_fprop_bnstats.
int pdt_conv_fprop_ex_bnstats(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int Cx, long long x_sw,
                              long long x_sh, long long x_sn, int K, int R, int S, int stride, int pad_lo_h, int pad_lo_w,
                              int pad_hi_h, int pad_hi_w, int P, int Q, int dil_w, void* scratch, const PdtCommDesc* comm,
                              float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                              float* save_invstd, cudaStream_t stream) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return PDT_E_UNSUPPORTED;
  if (scratch == nullptr || save_mean == nullptr || save_invstd == nullptr) return PDT_E_UNSUPPORTED;
  if (K % 8 || K > kStatMaxK || R < 1 || S < 1 || stride < 1 || P < 1 || Q < 1 || dil_w < 1 || (x_sw | x_sh | x_sn) % 16) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return PDT_E_UNSUPPORTED;
  int e = load_driver_fns();
  if (e) return e;
  ConvGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.dil_w = dil_w;
  g.pad_h = pad_lo_h; g.pad_w = pad_lo_w; g.pad_hi_h = pad_hi_h; g.pad_hi_w = pad_hi_w;
  g.P = P; g.Q = Q;
  g.M = N * P * Q;
  g.m_tiles = (g.M + BM - 1) / BM;
  g.OH = P; g.OW = Q; g.os = 1; g.oa = g.ob = 0;
  g.Cx = Cx; g.x_sw = x_sw; g.x_sh = x_sh; g.x_sn = x_sn;
  g.addend = nullptr;
  BnStats bs;
  bs.enabled = 1;
  bs.counter = reinterpret_cast<unsigned*>(scratch);
  bs.partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  bs.running_mean = running_mean; bs.running_var = running_var;
  bs.save_mean = save_mean; bs.save_invstd = save_invstd;
  bs.momentum = momentum; bs.eps = eps;
  bs.comm = make_comm(comm);
  return dispatch_fprop(x, w, y, nullptr, dtype, g, stream, &bs);
}

}  // extern "C"
