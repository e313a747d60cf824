// Patch-reuse convolution kernels (SURVEY K-CONV-F / -D / -W for the large feature maps).  Three kernels share
// one idea — load a contiguous input patch ONCE and form every filter tap as a row-shifted shared-memory
// descriptor on it:
//   k_conv3x3_halo   3x3 / 1 / 1 forward and data gradient, 64- and 128-channel layers   (pdt_conv3x3_halo)
//   k_wgrad3x3_halo  3x3 / 1 / 1 weight gradient, 64 -> 64 channels, two taps per MMA     (pdt_conv3x3_wgrad_halo)
//   k_stem_conv      the 7x7 / 2 stem as 4x4 / 1 over the space-to-depth image            (pdt_stem_conv)
//
// k_conv3x3_halo:
// The generic implicit-GEMM kernel (conv.cu) fetches the activation once PER FILTER TAP (9 TMA im2col loads
// per 64-channel block).  On the 56x56 and 28x28 layers that makes the L2 -> SM fabric (about 42 B/clk per
// SM), not the tensor core, the limit: ncu shows 22 % tensor-pipe utilisation at 128x64 tiles.
// Here a tile is TR whole rows of one image in PADDED coordinates (W + 2 wide, so TR*(W+2) <= 128 outputs,
// the two border columns are junk).  ONE tiled TMA load brings the (TR+2) x (W+2) x 64-channel patch, with
// the halo zero-filled by TMA's out-of-bounds handling, and all nine taps read it through shared-memory
// descriptors whose start address is shifted by (r*(W+2) + s - 1) rows: tcgen05 derives the 128B swizzle
// from the absolute address, so a row-shifted view of a swizzled tile is still a valid K-major operand
// (verified by csrc/probe.cu).  The 64x64 filter bank of the first stage (72 KB for all taps) stays resident
// in shared memory for the whole kernel; wider layers stream their taps through a second ring.
// Activation traffic per output drops ~5x; the filter-tap loop issues 36 tcgen05.mma per patch.
// Same warp roles, TMEM double buffering and BatchNorm-statistics epilogue as conv.cu.  The data gradient
// of these layers is the same kernel on dy with the flipped/transposed filter.
#include <cstdlib>
#include <type_traits>
#include "tc.cuh"
#include "fabric.cuh"

namespace {

using namespace pdt;
using namespace pdt::tc;

constexpr int kThreads = 384;       // warps: 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4..11 epilogue
constexpr int kEpiWarps = 8;        // two warps per TMEM lane quarter, each draining half of the columns
constexpr int kStatMaxK = 128;
#ifndef PDT_HALO64_CHAINS
#define PDT_HALO64_CHAINS 3      // (4 measured 3 % slower: the chains are not the limit)
#endif

struct HaloGeom {
  int N, H, W, C, K;
  int Wp;                 // W + 2
  int TR;                 // image rows per tile
  int tiles_per_image;
  int num_tiles;
  int c_chunks;
  uint32_t patch_bytes;   // (TR + 2) * Wp * 128
  const void* addend;     // optional [N][H][W][K] tensor added to the output in the epilogue (residual-gradient sum), or null
  int dw_ct;              // weight-gradient kernels: input channels of the WHOLE layer (the row length of dw); a launch
                          // handles one 64 x 64 (input, output) channel slice of it
};

struct BnStats {
  int enabled;
  float* partial;
  unsigned* counter;
  float* running_mean;
  float* running_var;
  float* save_mean;
  float* save_invstd;
  float momentum, eps;
  long long count;        // N*H*W
  Comm comm;              // world > 1: SyncBatchNorm — the last CTA exchanges the raw moments with its peers (fabric.cuh)
};

// PAIR: two CTAs of a cluster run ONE M = 256 tcgen05.mma.cta_group::2 per tap and k-step over their two patches; each
// holds HALF of every filter tile (BN / 2 output channels), so the 64 -> 64 filter banks stay resident in fp32 too
// (72 KB per CTA) and a CTA reads A + B/2 from shared memory per instruction (see tc.cuh, k_conv_igemm<.., PAIR>).
template <int BN, bool RES, bool F32 = false, bool PAIR = false> struct HCfg {
  static constexpr int kAStage = 32768;                       // 1 KB front slack + patch + tail slack (Wp <= 58)
  static constexpr int kAStages = 3;
  static constexpr int kBTile = (PAIR ? BN / 2 : BN) * 128;   // rows of one (tap, channel chunk) filter tile held by this CTA
  static constexpr int kChunks = F32 ? 2 : 1;                 // resident form: C = 64 is one 16-bit chunk or two fp32 chunks
  static constexpr int kBStages = RES ? 9 * kChunks : (PAIR ? 8 : 4);   // resident: every tap of every chunk
  // Several accumulator chains per tile, each double-buffered: consecutive tcgen05.mma that accumulate into
  // the SAME TMEM columns serialise on the accumulate latency (~96 clk measured, i.e. 3x the 32-clk rate of
  // an N=64 MMA and 1.5x the 64-clk rate of an N=128 one); rotating the 36 MMAs of a patch over 3 (N=64) or
  // 2 (N=128) independent chains hides it.  The epilogue adds the chains.
  static constexpr int kChains = BN == 64 ? PDT_HALO64_CHAINS : 2;
  static constexpr int kAccCols = BN == 64 ? 256 : 2 * BN;      // columns reserved per accumulator buffer (chains * BN, padded)
  static constexpr int kTmemCols = 2 * kAccCols;                // 512
  // per epilogue warp: column accumulators [2][kStatMaxK] + a 32 x 33 float transpose tile
  // N=64: the output tile is staged in shared memory (dense rows, 128B-swizzled, double-buffered) and leaves
  // through one TMA store, and the statistics stay in registers, so the transpose tiles are not used.
  static constexpr bool kTmaStore = BN == 64 && !F32;           // (fp32 rows are 256 B: they leave through direct 16-byte stores)
  static constexpr int kOutBuf = 128 * 128;
  static constexpr int kOutBytes = kTmaStore ? 2 * kOutBuf : 0;
  static constexpr int kStatStride = kTmaStore ? 2 * kStatMaxK : 2 * kStatMaxK + 32 * 33;   // floats per epilogue warp
  static constexpr int kStatBytes = kTmaStore ? 16384 : kEpiWarps * kStatStride * 4;   // also the fp64 fold scratch of the last CTA
  static constexpr int kSmemBytes = kAStages * kAStage + kBStages * kBTile + kOutBytes + 1024 + 256 + kStatBytes;
  static_assert(kSmemBytes <= 232448, "over the 227 KB shared-memory limit of sm_100");
};

__device__ __forceinline__ void tma_load_4d(void* smem, const CUtensorMap* m, uint64_t* bar, int c, int w, int h, int n) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n) : "memory");
}

__device__ __forceinline__ void tma_load_3d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// The same for a CTA pair: the data lands in THIS CTA's shared memory, the bytes are signalled on the leader's barrier.
__device__ __forceinline__ void tma_load_4d_pair(void* smem, const CUtensorMap* m, uint32_t bar_cluster_addr, int c, int w, int h, int n) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c), "r"(w), "r"(h), "r"(n) : "memory");
}

// Pull a future patch into L2 (no shared-memory destination, no barrier): the 3-deep shared-memory ring alone
// cannot hide DRAM latency for 30 KB patches, an L2 hit can be.
__device__ __forceinline__ void tma_prefetch_4d(const CUtensorMap* m, int c, int w, int h, int n) {
  asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global [%0, {%1, %2, %3, %4}];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c), "r"(w), "r"(h), "r"(n) : "memory");
}

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, uint32_t smem_src, int c, int w, int h, int n) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_src), "r"(c), "r"(w), "r"(h), "r"(n) : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, uint32_t smem_src, int c, int w, int h) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_src), "r"(c), "r"(w), "r"(h) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

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
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// CL = 2: CTA pairs (a thread-block cluster) stream ONE copy of the filter taps between them: each CTA loads
// half of a tap's rows and TMA multicasts it into both rings, which halves the L2 traffic of the layers whose
// filter bank does not fit in shared memory (they were bound by it: 9 x 16 KB of taps per 23 KB patch).  A ring
// slot is refilled only when both CTAs have consumed it, so the MMA commit arrives on both `bempty` barriers;
// CTAs of a pair run the same number of tiles (the last one may be a phantom tile that stores nothing).
template <int BN, bool RES, typename T, int CL, bool PAIR = false>
__global__ void __launch_bounds__(kThreads, 1)
k_conv3x3_halo(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmY, T* __restrict__ out, const HaloGeom g, const BnStats bs) {
  using C = HCfg<BN, RES, sizeof(T) == 4, PAIR>;
  static_assert(!PAIR || CL == 2, "a CTA pair is a cluster of two");
  // one 128-byte K row of a patch pixel / filter row: 64 channels in 16 bit, 32 as fp32 (kind::tf32); a layer's
  // channels are g.c_chunks such chunks, each with its own patch load and nine tap tiles
  constexpr int kCh = 128 / (int)sizeof(T);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + C::kAStages * C::kAStage;
  uint8_t* smO = smB + C::kBStages * C::kBTile;          // output staging (1024-byte aligned: every ring above is n KB)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smO + C::kOutBytes);
  uint64_t* afull = bars;                        // [kAStages]
  uint64_t* aempty = afull + C::kAStages;
  uint64_t* bfull = aempty + C::kAStages;        // [kBStages] (RES: only bfull[0] is used, once)
  uint64_t* bempty = bfull + C::kBStages;
  uint64_t* tfull = bempty + C::kBStages;
  uint64_t* tempty = tfull + 2;
  uint64_t* sfree = tempty + 2;                  // [2] output staging buffer has been read by its TMA store
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sfree + 2);
  float* s_stat = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  static_assert(CL == 1 || PAIR || (!RES && !C::kTmaStore), "the multicast ring is for the streamed, direct-store variant");
  static_assert(sizeof(T) == 2 || !RES || PAIR, "the fp32 filter bank (144 KB for 64 x 64 x 9) is resident only as the halves of a CTA pair");
  const int crank = CL > 1 ? (int)cluster_ctarank() : 0;
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1);
  // every CTA of a cluster runs the same number of iterations
  const int tile_end = CL > 1 ? ((g.num_tiles + (int)gridDim.x - 1) / (int)gridDim.x) * (int)gridDim.x : g.num_tiles;

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmX); prefetch_tmap(&tmW); if (C::kTmaStore) prefetch_tmap(&tmY); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kAStages; ++i) { mbar_init(&afull[i], 1); mbar_init(&aempty[i], 1); }
    // PAIR: the `full` barriers that count are the leader's (both CTAs' loads complete on them); one MMA commit
    // reaches both CTAs' `empty` / tfull barriers; both CTAs' epilogue warps arrive on the leader's tempty
    for (int i = 0; i < C::kBStages; ++i) { mbar_init(&bfull[i], 1); mbar_init(&bempty[i], PAIR ? 1 : CL); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], PAIR ? 2 * kEpiWarps : kEpiWarps); mbar_init(&sfree[i], 1); }
    fence_barrier_init();
  }
  if (warp == 2) { if constexpr (PAIR) tmem_alloc_pair(tmem_slot, C::kTmemCols); else tmem_alloc(tmem_slot, C::kTmemCols); }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();

  if (warp == 0) {
    // ================= TMA producer (whole warp converged, one elected lane issues) =================
    {
      if (RES) {     // the whole 3x3x64xBN filter bank, once (PAIR: this CTA's BN / 2 rows of every tile)
        if (elect_one_sync()) {
          if constexpr (PAIR) {
            const uint32_t lead = mapa_cluster(smem_u32(&bfull[0]), 0);
            if (crank == 0) mbar_expect_tx(&bfull[0], 2 * C::kBStages * C::kBTile);
            for (int cc = 0; cc < C::kChunks; ++cc)
              for (int tap = 0; tap < 9; ++tap)
                tma_load_2d_pair(smB + (cc * 9 + tap) * C::kBTile, &tmW, lead, tap * g.C + cc * kCh, crank * (BN / 2));
          } else {
            mbar_expect_tx(&bfull[0], 9 * C::kBTile);
            for (int tap = 0; tap < 9; ++tap) tma_load_2d(smB + tap * C::kBTile, &tmW, &bfull[0], tap * g.C, 0);
          }
        }
        __syncwarp();
      }
      int as = 0, bsg = 0;
      uint32_t aph = 0, bph = 0;
      for (int tile = blockIdx.x; tile < tile_end; tile += gridDim.x) {
        const int n = tile / g.tiles_per_image;
        const int h0 = (tile - n * g.tiles_per_image) * g.TR;
        {   // L2 prefetch of the patch this CTA will want kPrefetchAhead tiles from now
          constexpr int kPrefetchAhead = 6;
          const int ft = tile + kPrefetchAhead * (int)gridDim.x;
          if (ft < g.num_tiles && elect_one_sync()) {
            const int fn = ft / g.tiles_per_image;
            const int fh = (ft - fn * g.tiles_per_image) * g.TR;
            for (int cc = 0; cc < g.c_chunks; ++cc) tma_prefetch_4d(&tmX, cc * kCh, -1, fh - 1, fn);
          }
          __syncwarp();
        }
        for (int cc = 0; cc < g.c_chunks; ++cc) {
          mbar_wait(&aempty[as], aph ^ 1);
          if (elect_one_sync()) {
            if constexpr (PAIR) {
              if (crank == 0) mbar_expect_tx(&afull[as], 2 * g.patch_bytes);
              tma_load_4d_pair(smA + as * C::kAStage + 1024, &tmX, mapa_cluster(smem_u32(&afull[as]), 0), cc * kCh, -1, h0 - 1, n);
            } else {
              mbar_expect_tx(&afull[as], g.patch_bytes);
              tma_load_4d(smA + as * C::kAStage + 1024, &tmX, &afull[as], cc * kCh, -1, h0 - 1, n);
            }
          }
          __syncwarp();
          if (++as == C::kAStages) { as = 0; aph ^= 1; }
          if (!RES) {
            for (int tap = 0; tap < 9; ++tap) {
              mbar_wait(&bempty[bsg], bph ^ 1);
              if (elect_one_sync()) {
                if constexpr (PAIR) {
                  if (crank == 0) mbar_expect_tx(&bfull[bsg], 2 * C::kBTile);
                  tma_load_2d_pair(smB + bsg * C::kBTile, &tmW, mapa_cluster(smem_u32(&bfull[bsg]), 0), tap * g.C + cc * kCh,
                                   crank * (BN / 2));
                } else {
                mbar_expect_tx(&bfull[bsg], C::kBTile);
                if (CL == 1) tma_load_2d(smB + bsg * C::kBTile, &tmW, &bfull[bsg], tap * g.C + cc * kCh, 0);
                else tma_load_2d_mc(smB + bsg * C::kBTile + crank * (C::kBTile / CL), &tmW, &bfull[bsg], tap * g.C + cc * kCh,
                                    crank * (BN / CL), kMask);
                }
              }
              __syncwarp();
              if (++bsg == C::kBStages) { bsg = 0; bph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1 && (!PAIR || crank == 0)) {
    // ================= MMA issuer (whole warp converged, one elected lane issues; the leader CTA of a pair) =================
    {
      constexpr uint32_t idesc = make_idesc(PAIR ? 256 : 128, BN, sizeof(T) == 4 ? 2 : (std::is_same<T, __nv_bfloat16>::value ? 1 : 0), 0, 0);
      constexpr int NCH = C::kChains;
      auto mma = [](uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {      // 32 bytes of K per instruction either way
        if constexpr (PAIR) { if constexpr (sizeof(T) == 4) umma_tf32_pair(d, a, b, id, acc); else umma_f16_pair(d, a, b, id, acc); }
        else { if constexpr (sizeof(T) == 4) umma_tf32(d, a, b, id, acc); else umma_f16(d, a, b, id, acc); }
      };
      auto commit = [](uint64_t* bar) { if constexpr (PAIR) umma_commit_pair(bar, (uint16_t)3); else umma_commit(bar); };
      // Everything the 36 MMAs of a patch need is loop-invariant except the ring slot: the tap's operand is
      // the patch shifted by whole 128-byte pixel rows, i.e. a constant added to the descriptor's
      // (address >> 4) field.  Keeping those constants in registers and unrolling the tap loop matters: the
      // single issuing warp was the bottleneck when it recomputed them (ncu: 70 % of its samples in
      // descriptor arithmetic, 264 clk per tap for 4 MMAs worth 128-192 clk of tensor-pipe time).
      int64_t tapoff[9];
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) tapoff[tap] = (int64_t)(((tap / 3) * g.Wp + (tap % 3) - 1) * 8);
      const uint64_t db_base = make_sw128_desc(smem_u32(smB), 1024, 16);
      if (RES) { mbar_wait(&bfull[0], 0); tc_fence_after(); }
      int as = 0, bsg = 0, it = 0;
      uint32_t aph = 0, bph = 0;
      for (int tile = blockIdx.x; tile < tile_end; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * C::kAccCols;
        for (int cc = 0; cc < g.c_chunks; ++cc) {
          mbar_wait(&afull[as], aph);
          tc_fence_after();
          const uint64_t da_base = make_sw128_desc(smem_u32(smA + as * C::kAStage + 1024), 1024, 16);
          const uint32_t later = cc != 0;
          if (RES) {
            if (elect_one_sync()) {
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                const uint64_t da = da_base + (uint64_t)tapoff[tap];
                const uint64_t db = db_base + (uint64_t)((cc * 9 + tap) * (C::kBTile >> 4));
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const int i = tap * 4 + k;
                  mma(tmem_d + (i % NCH) * BN, da + 2 * k, db + 2 * k, idesc, i >= NCH ? 1u : later);
                }
              }
              commit(&aempty[as]);
            }
            __syncwarp();
          } else {
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
              mbar_wait(&bfull[bsg], bph);
              tc_fence_after();
              const uint64_t da = da_base + (uint64_t)tapoff[tap];
              const uint64_t db = db_base + (uint64_t)(bsg * (C::kBTile >> 4));
              if (elect_one_sync()) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const int i = tap * 4 + k;
                  mma(tmem_d + (i % NCH) * BN, da + 2 * k, db + 2 * k, idesc, i >= NCH ? 1u : later);
                }
                if constexpr (PAIR) commit(&bempty[bsg]);
                else if (CL == 1) umma_commit(&bempty[bsg]); else umma_commit_mc(&bempty[bsg], kMask);
                if (tap == 8) commit(&aempty[as]);
              }
              __syncwarp();
              if (++bsg == C::kBStages) { bsg = 0; bph ^= 1; }
            }
          }
          if (++as == C::kAStages) { as = 0; aph ^= 1; }
        }
        if (elect_one_sync()) commit(&tfull[acc]);
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    auto release_acc = [&](int acc) {        // one arrival per epilogue warp; a pair's accumulators are released at the leader
      if constexpr (PAIR) mbar_arrive_cluster(mapa_cluster(smem_u32(&tempty[acc]), 0)); else mbar_arrive(&tempty[acc]);
    };
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may read
    const int half = (warp - 4) >> 2;             // which half of the BN columns it drains
    const int ew = warp - 4;
    constexpr int NCH = C::kChains;
    float* my_stat = s_stat + ew * C::kStatStride;
    const int m = quarter * 32 + lane;            // output slot inside the tile (padded coordinates)
    const int ti = m / g.Wp, tj = m - ti * g.Wp;
    // tile -> (image, first row) advanced incrementally: no division inside the loop
    int n = (int)blockIdx.x / g.tiles_per_image, tr = (int)blockIdx.x - n * g.tiles_per_image;
    const int dn = (int)gridDim.x / g.tiles_per_image, dr = (int)gridDim.x - dn * g.tiles_per_image;
    if constexpr (C::kTmaStore) {
      // ---- N = 64: statistics in registers, output through a swizzled staging tile + one TMA store ----
      // Each lane owns one accumulator row, so it keeps its own partial sums for its 32 columns across all
      // tiles (2 instructions per element, no transposes); lanes are combined once at the end of the kernel.
      float cs[32], cq[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) cs[j] = cq[j] = 0.f;
      const bool slot_ok = ti < g.TR && tj >= 1 && tj <= g.W;
      const int drow = ti * g.W + (tj - 1);                      // row of this slot in the dense staging tile
      const uint32_t so = smem_u32(smO);
      const int c0 = half * 32;
      int it = 0;
      for (int tile = blockIdx.x; tile < tile_end; tile += gridDim.x, ++it) {
        const int h0 = tr * g.TR;
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        const bool valid = slot_ok && (h0 + ti) < g.H;
        const uint32_t taddr = tmem_base + acc * C::kAccCols + ((uint32_t)(quarter * 32) << 16);
        uint4 ad[4];
        const bool has_add = g.addend != nullptr && valid;
        if (has_add) {            // issued before the accumulator wait: the loads fly while the MMAs finish
          const T* arow = reinterpret_cast<const T*>(g.addend) + (((size_t)n * g.H + h0 + ti) * g.W + (tj - 1)) * g.K + c0;
#pragma unroll
          for (int j = 0; j < 4; ++j) ad[j] = ldg_stream(arow + 8 * j);
        }
        mbar_wait(&tfull[acc], acc_phase);
        tc_fence_after();
        uint32_t v[32];
        {
          uint32_t v2[32], v3[32], v4[NCH >= 4 ? 32 : 1];
          tmem_ld32(taddr + c0, v);
          tmem_ld32(taddr + BN + c0, v2);
          if (NCH >= 3) tmem_ld32(taddr + 2 * BN + c0, v3);
          if constexpr (NCH >= 4) tmem_ld32(taddr + 3 * BN + c0, v4);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float t = __uint_as_float(v[j]) + __uint_as_float(v2[j]);
            if (NCH >= 3) t += __uint_as_float(v3[j]);
            if constexpr (NCH >= 4) t += __uint_as_float(v4[j]);
            v[j] = __float_as_uint(t);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);                // accumulators are in registers: hand the buffer back
        if (has_add) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float a[8];
            unpack8<T>(ad[j], a);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[8 * j + e] = __float_as_uint(__uint_as_float(v[8 * j + e]) + a[e]);
          }
        }
        uint4 o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[j].x = Cvt<T>::pack2(__uint_as_float(v[8 * j + 0]), __uint_as_float(v[8 * j + 1]));
          o[j].y = Cvt<T>::pack2(__uint_as_float(v[8 * j + 2]), __uint_as_float(v[8 * j + 3]));
          o[j].z = Cvt<T>::pack2(__uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]));
          o[j].w = Cvt<T>::pack2(__uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
        }
        const int ob = it & 1;
        mbar_wait(&sfree[ob], ((it >> 1) & 1) ^ 1);              // the store that last used this buffer has read it
        if (slot_ok) {
          const uint32_t row = so + (uint32_t)(ob * C::kOutBuf + drow * 128);
#pragma unroll
          for (int j = 0; j < 4; ++j) sts128(row + (uint32_t)((((half * 4 + j) ^ (drow & 7))) << 4), o[j]);
        }
        fence_proxy_async();
        epi_bar_sync();
        if (threadIdx.x == 128) {
          tma_store_4d(&tmY, so + (uint32_t)(ob * C::kOutBuf), 0, 0, h0, n);   // rows past H are clipped by TMA
          bulk_commit();
          bulk_wait_read<1>();                                   // the previous tile's store no longer needs its buffer
          if (it >= 1) mbar_arrive(&sfree[ob ^ 1]);
        }
        if (bs.enabled && valid) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float t = __uint_as_float(v[j]);
            cs[j] += t;
            cq[j] = fmaf(t, t, cq[j]);
          }
        }
        tr += dr; n += dn;
        if (tr >= g.tiles_per_image) { tr -= g.tiles_per_image; ++n; }
      }
      if (threadIdx.x == 128) bulk_wait_all<0>();
      if (bs.enabled) {
        for (int i = lane; i < 2 * kStatMaxK; i += 32) my_stat[i] = 0.f;
        __syncwarp();
        float ks = 0.f, kq = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float a = cs[j], b = cq[j];
#pragma unroll
          for (int sft = 16; sft >= 1; sft >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, sft);
            b += __shfl_xor_sync(0xffffffffu, b, sft);
          }
          if (lane == j) { ks = a; kq = b; }
        }
        my_stat[c0 + lane] = ks;
        my_stat[kStatMaxK + c0 + lane] = kq;
      }
    } else {
    constexpr int NIT = BN / 64;                  // 32-column groups per warp
    const uint32_t tp_a = smem_u32(my_stat + 2 * kStatMaxK);   // 32 x 33 transpose tile
    if (bs.enabled) {
      for (int i = lane; i < 2 * kStatMaxK; i += 32) my_stat[i] = 0.f;
      __syncwarp();
    }
    float cs[NIT], cq[NIT];                       // this lane's column (c0 + lane): running sum / sum of squares
#pragma unroll
    for (int q = 0; q < NIT; ++q) cs[q] = cq[q] = 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < tile_end; tile += gridDim.x, ++it) {
      const int h0 = tr * g.TR;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const bool valid = ti < g.TR && tj >= 1 && tj <= g.W && (h0 + ti) < g.H && tile < g.num_tiles;
      T* orow = out + (((size_t)n * g.H + h0 + ti) * g.W + (tj - 1)) * g.K;
      const uint32_t taddr = tmem_base + acc * C::kAccCols + ((uint32_t)(quarter * 32) << 16);
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
#pragma unroll
      for (int q = 0; q < NIT; ++q) {
        const int c0 = half * (BN / 2) + q * 32;
        uint32_t v[32];
        {
          uint32_t v2[32];
          tmem_ld32(taddr + c0, v);
          tmem_ld32(taddr + BN + c0, v2);
          if (NCH == 3) {
            uint32_t v3[32];
            tmem_ld32(taddr + 2 * BN + c0, v3);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]) + __uint_as_float(v3[j]));
          } else {
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]));
          }
        }
        if (q == NIT - 1) {       // accumulators are in registers: hand the buffer back before the long tail
          tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(acc);
        }
        if (valid && g.addend != nullptr) {
          const T* arow = reinterpret_cast<const T*>(g.addend) + (orow - out) + c0;
          if constexpr (sizeof(T) == 4) {
            uint4 ad[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) ad[j] = ldg_stream(arow + 4 * j);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[4 * j + 0] = __float_as_uint(__uint_as_float(v[4 * j + 0]) + __uint_as_float(ad[j].x));
              v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + __uint_as_float(ad[j].y));
              v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + __uint_as_float(ad[j].z));
              v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + __uint_as_float(ad[j].w));
            }
          } else {
            uint4 ad[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) ad[j] = ldg_stream(arow + 8 * j);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float a[8];
              unpack8<T>(ad[j], a);
#pragma unroll
              for (int e = 0; e < 8; ++e) v[8 * j + e] = __float_as_uint(__uint_as_float(v[8 * j + e]) + a[e]);
            }
          }
        }
        if (valid) {
          if constexpr (sizeof(T) == 4) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<uint4*>(orow + c0 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 o;
              o.x = Cvt<T>::pack2(__uint_as_float(v[j + 0]), __uint_as_float(v[j + 1]));
              o.y = Cvt<T>::pack2(__uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
              o.z = Cvt<T>::pack2(__uint_as_float(v[j + 4]), __uint_as_float(v[j + 5]));
              o.w = Cvt<T>::pack2(__uint_as_float(v[j + 6]), __uint_as_float(v[j + 7]));
              *reinterpret_cast<uint4*>(orow + c0 + j) = o;
            }
          }
        }
        if (bs.enabled) {
          // column sums over this warp's 32 rows through a padded shared-memory transpose (conflict-free
          // both ways): 32 STS + 32 LDS + 64 FMA per lane, cheaper than a shuffle butterfly
#pragma unroll
          for (int j = 0; j < 32; ++j)
            sts_f32(tp_a + (uint32_t)(lane * 33 + j) * 4u, valid ? __uint_as_float(v[j]) : 0.f);   // junk slots excluded
          __syncwarp();
          float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
          for (int rr = 0; rr < 32; rr += 2) {
            const float t0 = lds_f32(tp_a + (uint32_t)(rr * 33 + lane) * 4u);
            const float t1 = lds_f32(tp_a + (uint32_t)((rr + 1) * 33 + lane) * 4u);
            s0 += t0; q0 = fmaf(t0, t0, q0);
            s1 += t1; q1 = fmaf(t1, t1, q1);
          }
          __syncwarp();
          cs[q] += s0 + s1;
          cq[q] += q0 + q1;
        }
      }
      tr += dr; n += dn;
      if (tr >= g.tiles_per_image) { tr -= g.tiles_per_image; ++n; }
    }
    if (bs.enabled) {
#pragma unroll
      for (int q = 0; q < NIT; ++q) {
        my_stat[half * (BN / 2) + q * 32 + lane] = cs[q];
        my_stat[kStatMaxK + half * (BN / 2) + q * 32 + lane] = cq[q];
      }
    }
    }   // !kTmaStore
    if (bs.enabled) {
      __shared__ int s_last;
      const int te = threadIdx.x - 128;            // 0..255 among the epilogue threads
      const int K = g.K;
      epi_bar_sync();
      for (int i = te; i < 2 * K; i += 256) {
        const int which = i / K, c = i - which * K;
        float t = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < kEpiWarps; ++w8) t += s_stat[w8 * C::kStatStride + which * kStatMaxK + c];
        bs.partial[(size_t)blockIdx.x * 2 * K + i] = t;
      }
      double* tot = reinterpret_cast<double*>(s_stat);
      if (tree_fold(tot, bs.partial, bs.counter, (int)gridDim.x, 2 * K, PDT_BN_MAX_BLOCKS, te, 256, [] { epi_bar_sync(); }, &s_last)) {
        __shared__ double s_cnt;
        if (te == 0) s_cnt = (double)bs.count;
        epi_bar_sync();
        if (bs.comm.world > 1) exchange_sum_t(bs.comm, tot, 2 * K, &s_cnt, te, 256, [] { epi_bar_sync(); });
        bn_stats_finalize(tot, K, s_cnt, bs.eps, bs.momentum, bs.save_mean, bs.save_invstd, bs.running_mean, bs.running_var, te, 256);
        if (te == 0) *bs.counter = 0;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, C::kTmemCols); else tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ---- 3x3 / stride 1 / pad 1 weight gradient, 64 -> 64 channels, with the same patch reuse -------------
// dW[k][r][s][c] = sum over pixels of dy[pixel][k] * x[pixel + (r-1, s-1)][c].  The generic split-K kernel
// (conv.cu) fetches one im2col tile of x per tap and 64-pixel step, which makes it L2-bound by ~3x on the
// 56x56 layers.  Here a work item is again TR image rows in padded coordinates: ONE x patch (with halo) and
// ONE dy tile (its two border columns and out-of-image rows zero-filled by TMA) are loaded per item, and the
// nine taps are nine row-shifted MN-major views of the patch.  Two taps share one M=128 instruction: the
// descriptor's leading-dimension offset (distance between the two 64-channel halves of A) is set to the
// distance between the two taps' views (one pixel = 128 B, or W+2-2 pixels across a filter-row boundary), so
// D rows 0-63 / 64-127 are dW of tap a / tap b.  Five accumulators (taps 01 23 45 67 8x) live in TMEM for the
// whole kernel; each CTA adds them into dw once at the end (red.global.add.f32, dw must start zeroed).
struct WHCfg {
  static constexpr int kStages = 4;
  static constexpr int kAStage = 32768;          // 1 KB front slack + patch + tail (rows read past the patch are zero)
  static constexpr int kBStage = 16384;          // 128 pixel slots x 64 channels
  static constexpr int kSmemBytes = kStages * (kAStage + kBStage) + 1024 + 128;
};

template <typename T>
__global__ void __launch_bounds__(128, 1)
k_wgrad3x3_halo(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ dw,
                const HaloGeom g) {
  using W = WHCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + W::kStages * W::kAStage;
  uint64_t* full = reinterpret_cast<uint64_t*>(smB + W::kStages * W::kBStage);
  uint64_t* empty = full + W::kStages;
  uint64_t* done = empty + W::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmX); prefetch_tmap(&tmDY); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < W::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  // Slots the TMA boxes never write (tail of every stage) are read by the last k-step / the largest shifts:
  // they must hold zeros, not whatever the previous kernel left in shared memory (0 * NaN = NaN).
  {
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = W::kStages * (W::kAStage + W::kBStage) / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();
  const int slots = g.TR * g.Wp;
  const int ksteps = (slots + 15) >> 4;
  const uint32_t dy_bytes = (uint32_t)slots * 128u;

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    int n = (int)blockIdx.x / g.tiles_per_image, tr = (int)blockIdx.x - n * g.tiles_per_image;
    const int dn = (int)gridDim.x / g.tiles_per_image, dr = (int)gridDim.x - dn * g.tiles_per_image;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      const int h0 = tr * g.TR;
      mbar_wait(&empty[stage], phase ^ 1);
      if (elect_one_sync()) {
        mbar_expect_tx(&full[stage], g.patch_bytes + dy_bytes);
        tma_load_4d(smA + stage * W::kAStage + 1024, &tmX, &full[stage], 0, -1, h0 - 1, n);
        tma_load_4d(smB + stage * W::kBStage, &tmDY, &full[stage], 0, -1, h0, n);
      }
      __syncwarp();
      if (++stage == W::kStages) { stage = 0; phase ^= 1; }
      tr += dr; n += dn;
      if (tr >= g.tiles_per_image) { tr -= g.tiles_per_image; ++n; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(128, 64, std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 1, 1);
    // pair j = taps (2j, 2j+1); view of tap t starts (t/3)*Wp + t%3 - 1 pixel rows into the patch
    uint64_t tmpl[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int ta = 2 * j, tb = 2 * j + 1;
      const int sa = (ta / 3) * g.Wp + ta % 3 - 1, sb = (tb / 3) * g.Wp + tb % 3 - 1;
      tmpl[j] = make_sw128_desc(0, 1024, (uint32_t)(sb - sa) * 128u) + (uint64_t)(int64_t)(sa * 8);
    }
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      const uint64_t a_addr = (uint64_t)((smem_u32(smA + stage * W::kAStage + 1024) & 0x3FFFF) >> 4);
      const uint64_t db0 = make_sw128_desc(smem_u32(smB + stage * W::kBStage), 1024, 1024);
      if (elect_one_sync()) {
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint32_t accum = (first && ks == 0) ? 0u : 1u;
#pragma unroll
          for (int j = 0; j < 5; ++j)
            umma_f16(tmem_base + j * 64, tmpl[j] + a_addr + (uint64_t)(ks * 128), db0 + (uint64_t)(ks * 128), idesc, accum);
        }
        umma_commit(&empty[stage]);
      }
      __syncwarp();
      first = false;
      if (++stage == W::kStages) { stage = 0; phase ^= 1; }
    }
    if (elect_one_sync()) umma_commit(done);
    __syncwarp();
  }
  __syncwarp();
  // ---- drain: every warp adds its 32 accumulator rows (lane = tap-half * 64 + input channel) into dw ----
  if (blockIdx.x < g.num_tiles) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int m = warp * 32 + lane;
    const int hi = m >> 6, c = m & 63;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
      const int tap = 2 * j + hi;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + j * 64 + c0, v);
        tmem_ld_wait();
        if (tap < 9) {
#pragma unroll
          for (int i = 0; i < 32; ++i)       // dw[k][tap][c], k = c0 + i: the 32 lanes of a warp cover 32 consecutive c
            atomicAdd(dw + ((size_t)(c0 + i) * 9 + tap) * g.dw_ct + c, __uint_as_float(v[i]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// fp32 (kind::tf32) form of k_wgrad3x3_halo.  A pixel's 64 fp32 channels are two 128-byte rows, so the patch and
// the dy tile are kept as two PLANES of 32 channels each (two TMA loads, 32-byte-atom swizzle: the MN-major layout
// tcgen05 reads for 32-bit operands, as in k_conv_wgrad<float>).  One M = 128 instruction takes the FOUR views
// (r, s = 0..3) of one plane — leading-dimension offset = one pixel slot, the fourth view is a phantom tap whose rows
// are never stored — against all 64 output channels (two planes of dy, N = 64): six accumulator chains
// (3 filter rows x 2 planes, 384 TMEM columns) per 8-pixel k-step, 192 tensor clocks per 8 pixels, and every
// activation byte leaves L2 once instead of once per tap (k_conv_wgrad<64, float>: 1.73 ms per layer-1
// convolution at 1200 images, bound by L2 -> SM traffic).
struct WHCfg32 {
  static constexpr int kStages = 2;
  static constexpr int kAPlane = 32768;          // 1 KB front slack + patch plane + tail
  static constexpr int kBPlane = 16384;          // 128 pixel slots x 32 channels
  static constexpr int kAStage = 2 * kAPlane, kBStage = 2 * kBPlane;
  static constexpr int kSmemBytes = kStages * (kAStage + kBStage) + 1024 + 128;
  static_assert(kSmemBytes <= 232448, "over the 227 KB shared-memory limit of sm_100");
};

__global__ void __launch_bounds__(128, 1)
k_wgrad3x3_halo_tf32(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ dw,
                     const HaloGeom g) {
  using W = WHCfg32;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + W::kStages * W::kAStage;
  uint64_t* full = reinterpret_cast<uint64_t*>(smB + W::kStages * W::kBStage);
  uint64_t* empty = full + W::kStages;
  uint64_t* done = empty + W::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmX); prefetch_tmap(&tmDY); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < W::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  {   // slots no TMA box writes are read by the last k-step / the largest shifts: zeros, not stale shared memory
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = W::kStages * (W::kAStage + W::kBStage) / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();
  const int slots = g.TR * g.Wp;
  const int ksteps = (slots + 7) >> 3;
  const uint32_t dy_bytes = (uint32_t)slots * 128u;          // per plane

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    int n = (int)blockIdx.x / g.tiles_per_image, tr = (int)blockIdx.x - n * g.tiles_per_image;
    const int dn = (int)gridDim.x / g.tiles_per_image, dr = (int)gridDim.x - dn * g.tiles_per_image;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      const int h0 = tr * g.TR;
      mbar_wait(&empty[stage], phase ^ 1);
      if (elect_one_sync()) {
        mbar_expect_tx(&full[stage], 2 * (g.patch_bytes + dy_bytes));
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          tma_load_4d(smA + stage * W::kAStage + p * W::kAPlane + 1024, &tmX, &full[stage], 32 * p, -1, h0 - 1, n);
          tma_load_4d(smB + stage * W::kBStage + p * W::kBPlane, &tmDY, &full[stage], 32 * p, -1, h0, n);
        }
      }
      __syncwarp();
      if (++stage == W::kStages) { stage = 0; phase ^= 1; }
      tr += dr; n += dn;
      if (tr >= g.tiles_per_image) { tr -= g.tiles_per_image; ++n; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(128, 64, 2, 1, 1);            // kind::tf32, both operands MN-major
    // filter row r: the views of taps (r, 0..3) start (r * Wp - 1), +1, +2, +3 slots into the plane
    uint64_t tmpl[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) tmpl[r] = make_sw128_desc(0, 512, 128, 1) + (uint64_t)(int64_t)((r * g.Wp - 1) * 8);
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      const uint64_t a_addr = (uint64_t)((smem_u32(smA + stage * W::kAStage + 1024) & 0x3FFFF) >> 4);
      const uint64_t db0 = make_sw128_desc(smem_u32(smB + stage * W::kBStage), 512, W::kBPlane, 1);
      if (elect_one_sync()) {
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint32_t accum = (first && ks == 0) ? 0u : 1u;
#pragma unroll
          for (int j = 0; j < 6; ++j)      // chain j = (filter row j / 2, channel plane j % 2)
            umma_tf32(tmem_base + j * 64, tmpl[j >> 1] + a_addr + (uint64_t)((j & 1) * (W::kAPlane >> 4) + ks * 64),
                      db0 + (uint64_t)(ks * 64), idesc, accum);
        }
        umma_commit(&empty[stage]);
      }
      __syncwarp();
      first = false;
      if (++stage == W::kStages) { stage = 0; phase ^= 1; }
    }
    if (elect_one_sync()) umma_commit(done);
    __syncwarp();
  }
  __syncwarp();
  // ---- drain: accumulator row = (tap column s) * 32 + input channel of the plane; warp 3 holds the phantom tap ----
  if (blockIdx.x < g.num_tiles) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int sx = warp, c = lane;
#pragma unroll 1
    for (int j = 0; j < 6; ++j) {
      const int tap = (j >> 1) * 3 + sx, ch = (j & 1) * 32 + c;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + j * 64 + c0, v);
        tmem_ld_wait();
        if (sx < 3) {
#pragma unroll
          for (int i = 0; i < 32; ++i)       // dw[k][tap][ch], k = c0 + i: a warp's lanes cover 32 consecutive channels
            atomicAdd(dw + ((size_t)(c0 + i) * 9 + tap) * g.dw_ct + ch, __uint_as_float(v[i]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---- stem weight gradient: dW'[k][r][s][c] = sum over output pixels of dy[n,p,q,k] * xs[n, p+r, q+s, c] (4x4 taps over the
// space-to-depth image, 16 channels per pixel).  The generic split-K kernel builds each 128-byte operand row (four
// neighbouring pixels) with an im2col load per filter row and M tile, pulling every pixel 16 times and every dy
// element twice out of L2 (1.0 ms at 1200 images, bound by L2 -> SM traffic).  Here a tile is TR output rows of one
// image: the TR + 3 image rows it touches arrive ONCE as linear 32-byte pixel rows (32-byte swizzle), and the A
// operand of filter row r is that patch read MN-major: 16 channels of a pixel are one 32-byte atom along M, the
// NEXT atom (leading-dimension offset 32 B) is the next pixel, i.e. the next horizontal tap, and the reduction index
// (the output pixel q) also advances one pixel per row: M = 128 rows are taps s = 0..7 x 16 channels of which
// s < 4 are real (an M = 64 instruction costs the same tensor time), against dy's 64 channels (N = 64).  Four
// accumulators (r = 0..3) stay in TMEM for the whole kernel and are added into dW' once per CTA.
struct SWCfg {
  static constexpr int kStages = 2;
  static constexpr int kTR = 4;                        // output rows per tile
  static constexpr int kAStage = 1024 * 32;            // 1024 linear pixel slots (>= (kTR + 3) * WP + 8)
  static constexpr int kBStage = kTR * 128 * 128;      // kTR rows of up to 128 pixels x 64 channels
  static constexpr int kSmemBytes = kStages * (kAStage + kBStage) + 1024 + 128;
  static_assert(kSmemBytes <= 232448, "over the 227 KB shared-memory limit of sm_100");
};
struct StemWGeom {
  int N, HP, WP, P, Q;
  int tiles_per_image, num_tiles;
};

template <typename T>
__global__ void __launch_bounds__(128, 1)
k_stem_wgrad(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ dw, const StemWGeom g) {
  using W = SWCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + W::kStages * W::kAStage;
  uint64_t* full = reinterpret_cast<uint64_t*>(smB + W::kStages * W::kBStage);
  uint64_t* empty = full + W::kStages;
  uint64_t* done = empty + W::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmX); prefetch_tmap(&tmDY); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < W::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 256);
  {   // dy rows shorter than 128 pixels leave gaps no TMA box writes: zeros, not stale shared memory (0 * NaN = NaN)
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = W::kStages * (W::kAStage + W::kBStage) / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();
  const int qsteps = g.Q >> 4;                               // 16-pixel k-steps per output row
  const uint32_t dy_half_bytes = (uint32_t)(2 * g.Q) * 128u; // one dy load: two output rows

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      const int n = tile / g.tiles_per_image;
      const int p0 = (tile - n * g.tiles_per_image) * W::kTR;
      mbar_wait(&empty[stage], phase ^ 1);
      if (elect_one_sync()) {
        mbar_expect_tx(&full[stage], (uint32_t)W::kAStage + 2 * dy_half_bytes);
        const long long slot0 = ((long long)n * g.HP + p0) * g.WP;
#pragma unroll
        for (int i = 0; i < 4; ++i)          // 4 x 256 pixel slots; slots past the image array are zero-filled
          tma_load_2d(smA + stage * W::kAStage + i * 8192, &tmX, &full[stage], 0, (int)(slot0 + 256 * i));
#pragma unroll
        for (int i = 0; i < 2; ++i)          // rows p0 + 2i, p0 + 2i + 1 of dy; rows past P are zero-filled
          tma_load_3d(smB + stage * W::kBStage + i * dy_half_bytes, &tmDY, &full[stage], 0, (p0 + 2 * i) * g.Q, n);
      }
      __syncwarp();
      if (++stage == W::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(128, 64, std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 1, 1);
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      // A: 32-byte swizzle (layout type 6), 8-pixel groups 256 B apart along the reduction, atoms 32 B apart along M
      const uint64_t da0 = make_sw128_desc(smem_u32(smA + stage * W::kAStage), 256, 32, 6);
      const uint64_t db0 = make_sw128_desc(smem_u32(smB + stage * W::kBStage), 1024, 1024);
      if (elect_one_sync()) {
        for (int j = 0; j < W::kTR; ++j)
          for (int ks = 0; ks < qsteps; ++ks) {
            const uint32_t accum = (first && j == 0 && ks == 0) ? 0u : 1u;
            const uint64_t db = db0 + (uint64_t)((j * g.Q + ks * 16) * 8);
#pragma unroll
            for (int r = 0; r < 4; ++r)
              umma_f16(tmem_base + r * 64, da0 + (uint64_t)(((j + r) * g.WP + ks * 16) * 2), db, idesc, accum);
          }
        umma_commit(&empty[stage]);
      }
      __syncwarp();
      first = false;
      if (++stage == W::kStages) { stage = 0; phase ^= 1; }
    }
    if (elect_one_sync()) umma_commit(done);
    __syncwarp();
  }
  __syncwarp();
  // ---- drain: accumulator row m = s * 16 + c; rows 64..127 (warps 2, 3) are the phantom taps s = 4..7 ----
  if (blockIdx.x < g.num_tiles) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int m = warp * 32 + lane;
#pragma unroll 1
    for (int r = 0; r < 4; ++r) {
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + r * 64 + c0, v);
        tmem_ld_wait();
        if (warp < 2) {
#pragma unroll
          for (int i = 0; i < 32; ++i)       // dw'[k][r][s][c], k = c0 + i: a warp's lanes cover 32 consecutive (s, c)
            atomicAdd(dw + (size_t)(c0 + i) * 256 + r * 64 + m, __uint_as_float(v[i]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// ---- stem: 4x4 / stride 1 convolution over the space-to-depth image, 16 channels per pixel -----------
// (torchvision's 7x7 / stride 2 conv1 after pdt_stem_s2d).  The generic kernel builds each K row from four
// neighbouring 32-byte pixels with an im2col load per filter row, i.e. it pulls every pixel 16 times from
// L2 (ncu: 71 % L2 throughput, 21 % tensor pipe).  Here a tile is 128 consecutive pixel slots of the
// row-major image; ONE linear TMA load brings the 512 pixels a tile can touch, stored as 32-byte rows with
// the 32-byte swizzle, and each of the 16 taps is a K = 16 MMA whose A operand is that patch shifted by
// r * WP + s rows (same absolute-address swizzle argument as the 3x3 kernels).  A tile is ONE output row
// (Q <= 128; the slots past Q are junk and are neither stored nor counted), so the finished tile is a dense
// Q x 64 block that leaves through one TMA store from a 128B-swizzled staging buffer, as in k_conv3x3_halo.
// The 16 x 64 x 16 filter (32 KB) stays in shared memory; three TMEM accumulator chains hide the accumulate
// latency; 8 epilogue warps keep the BatchNorm statistics in registers.
struct StemGeom {
  int N, HP, WP, P, Q, K;       // K = 64 output channels
  long long slots;              // N * HP * WP
  int num_tiles;                // N * P: one output row per tile (Q <= 128 slots of it are real)
};
struct SCfg {
  static constexpr int kStages = 4;
  static constexpr int kAStage = 512 * 32;                 // 16 KB patch
  static constexpr int kBBytes = 16 * 64 * 32;             // resident filter
  static constexpr int kAccCols = 256;                     // 3 chains x 64 columns, padded
  static constexpr int kStatBytes = 16384;
  static constexpr int kOutBuf = 128 * 128;                // staging tile, double-buffered
  static constexpr int kSmemBytes = kStages * kAStage + kBBytes + 2 * kOutBuf + 1024 + 256 + kStatBytes;
};

template <typename T>
__global__ void __launch_bounds__(kThreads, 1)
k_stem_conv(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
            const __grid_constant__ CUtensorMap tmY, const StemGeom g, const BnStats bs) {
  using C = SCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + C::kStages * C::kAStage;
  uint8_t* smO = smB + C::kBBytes;               // 1024-byte aligned (everything above is a multiple of 16 KB)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smO + 2 * C::kOutBuf);
  uint64_t* afull = bars;
  uint64_t* aempty = afull + C::kStages;
  uint64_t* bfull = aempty + C::kStages;
  uint64_t* tfull = bfull + 1;
  uint64_t* tempty = tfull + 2;
  uint64_t* sfree = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sfree + 2);
  float* s_stat = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmX); prefetch_tmap(&tmW); prefetch_tmap(&tmY); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) { mbar_init(&afull[i], 1); mbar_init(&aempty[i], 1); }
    mbar_init(bfull, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], kEpiWarps); mbar_init(&sfree[i], 1); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_enter();

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one_sync()) {
      mbar_expect_tx(bfull, C::kBBytes);
      for (int tap = 0; tap < 16; ++tap) tma_load_2d(smB + tap * 2048, &tmW, bfull, tap * 16, 0);
    }
    __syncwarp();
    int as = 0;
    uint32_t aph = 0;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      const int n = tile / g.P, p = tile - n * g.P;
      const int base = (n * g.HP + p) * g.WP;              // first pixel slot of output row (n, p)
      mbar_wait(&aempty[as], aph ^ 1);
      if (elect_one_sync()) {
        mbar_expect_tx(&afull[as], C::kAStage);
        tma_load_2d(smA + as * C::kAStage, &tmX, &afull[as], 0, base);
        tma_load_2d(smA + as * C::kAStage + 8192, &tmX, &afull[as], 0, base + 256);
      }
      __syncwarp();
      if (++as == C::kStages) { as = 0; aph ^= 1; }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc = make_idesc(128, 64, std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 0, 0);
    int64_t tapoff[16];
#pragma unroll
    for (int tap = 0; tap < 16; ++tap) tapoff[tap] = (int64_t)(((tap >> 2) * g.WP + (tap & 3)) * 2);   // 32-byte rows, >> 4
    // 32-byte swizzle (layout type 6): 8-row groups 256 B apart
    const uint64_t db_base = make_sw128_desc(smem_u32(smB), 256, 16, 6);
    mbar_wait(bfull, 0);
    tc_fence_after();
    int as = 0, it = 0;
    uint32_t aph = 0;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * C::kAccCols;
      mbar_wait(&afull[as], aph);
      tc_fence_after();
      const uint64_t da_base = make_sw128_desc(smem_u32(smA + as * C::kAStage), 256, 16, 6);
      if (elect_one_sync()) {
#pragma unroll
        for (int tap = 0; tap < 16; ++tap)
          umma_f16(tmem_d + (tap % 3) * 64, da_base + (uint64_t)tapoff[tap], db_base + (uint64_t)(tap * 128), idesc, tap >= 3 ? 1u : 0u);
        umma_commit(&aempty[as]);
        umma_commit(&tfull[acc]);
      }
      __syncwarp();
      if (++as == C::kStages) { as = 0; aph ^= 1; }
    }
  } else if (warp >= 4) {
    // ================= epilogue: 8 warps, (lane quarter) x (column half) =================
    const int quarter = warp & 3, half = (warp - 4) >> 2, ew = warp - 4;
    const int c0 = half * 32;
    float* my_stat = s_stat + ew * (2 * kStatMaxK);
    float cs[32], cq[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) cs[j] = cq[j] = 0.f;
    const int m = quarter * 32 + lane;                     // slot = output column q
    const bool valid = m < g.Q;
    const uint32_t so = smem_u32(smO);
    int it = 0;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const uint32_t taddr = tmem_base + acc * C::kAccCols + ((uint32_t)(quarter * 32) << 16);
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      uint32_t v[32];
      {
        uint32_t v2[32], v3[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld32(taddr + 64 + c0, v2);
        tmem_ld32(taddr + 128 + c0, v3);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j)
          v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]) + __uint_as_float(v3[j]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      uint4 o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        o[j].x = Cvt<T>::pack2(__uint_as_float(v[8 * j + 0]), __uint_as_float(v[8 * j + 1]));
        o[j].y = Cvt<T>::pack2(__uint_as_float(v[8 * j + 2]), __uint_as_float(v[8 * j + 3]));
        o[j].z = Cvt<T>::pack2(__uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]));
        o[j].w = Cvt<T>::pack2(__uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
      }
      const int ob = it & 1;
      mbar_wait(&sfree[ob], ((it >> 1) & 1) ^ 1);            // the store that last used this buffer has read it
      if (valid) {
        const uint32_t row = so + (uint32_t)(ob * C::kOutBuf + m * 128);
#pragma unroll
        for (int j = 0; j < 4; ++j) sts128(row + (uint32_t)((((half * 4 + j) ^ (m & 7))) << 4), o[j]);
      }
      fence_proxy_async();
      epi_bar_sync();
      if (threadIdx.x == 128) {
        tma_store_3d(&tmY, so + (uint32_t)(ob * C::kOutBuf), 0, 0, tile);     // (channel, q, n * P + p)
        bulk_commit();
        bulk_wait_read<1>();
        if (it >= 1) mbar_arrive(&sfree[ob ^ 1]);
      }
      if (valid) {
        if (bs.enabled) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float t = __uint_as_float(v[j]);
            cs[j] += t;
            cq[j] = fmaf(t, t, cq[j]);
          }
        }
      }
    }
    if (threadIdx.x == 128) bulk_wait_all<0>();
    if (bs.enabled) {
      for (int i = lane; i < 2 * kStatMaxK; i += 32) my_stat[i] = 0.f;
      __syncwarp();
      float ks = 0.f, kq = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float a = cs[j], b = cq[j];
#pragma unroll
        for (int sft = 16; sft >= 1; sft >>= 1) {
          a += __shfl_xor_sync(0xffffffffu, a, sft);
          b += __shfl_xor_sync(0xffffffffu, b, sft);
        }
        if (lane == j) { ks = a; kq = b; }
      }
      my_stat[c0 + lane] = ks;
      my_stat[kStatMaxK + c0 + lane] = kq;
      // ---- fold the 8 warps, publish this CTA's row, last CTA finalises (as in k_conv3x3_halo) ----
      __shared__ int s_last;
      const int te = threadIdx.x - 128;
      const int K = 64;
      epi_bar_sync();
      for (int i = te; i < 2 * K; i += 256) {
        const int which = i / K, c = i - which * K;
        float t = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < kEpiWarps; ++w8) t += s_stat[w8 * (2 * kStatMaxK) + which * kStatMaxK + c];
        bs.partial[(size_t)blockIdx.x * 2 * K + i] = t;
      }
      double* tot = reinterpret_cast<double*>(s_stat);
      if (tree_fold(tot, bs.partial, bs.counter, (int)gridDim.x, 2 * K, PDT_BN_MAX_BLOCKS, te, 256, [] { epi_bar_sync(); }, &s_last)) {
        __shared__ double s_cnt;
        if (te == 0) s_cnt = (double)bs.count;
        epi_bar_sync();
        if (bs.comm.world > 1) exchange_sum_t(bs.comm, tot, 2 * K, &s_cnt, te, 256, [] { epi_bar_sync(); });
        bn_stats_finalize(tot, K, s_cnt, bs.eps, bs.momentum, bs.save_mean, bs.save_invstd, bs.running_mean, bs.running_var, te, 256);
        if (te == 0) *bs.counter = 0;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_enc = nullptr;

template <int BN, bool RES, typename T, int CL, bool PAIR = false>
int launch_t(void* y, const HaloGeom& g, const BnStats& bs, const CUtensorMap& tmX, const CUtensorMap& tmW, const CUtensorMap& tmY,
             cudaStream_t stream) {
  using C = HCfg<BN, RES, sizeof(T) == 4, PAIR>;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  auto kern = k_conv3x3_halo<BN, RES, T, CL, PAIR>;
  PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int nattr = 0;
  int grid = g.num_tiles < sms ? g.num_tiles : sms;
  if (CL > 1) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    nattr = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    static int max_clusters = -1;       // resident clusters of this kernel: the persistent grid must be one wave
    if (max_clusters < 0) {
      cfg.gridDim = dim3(sms / CL * CL);
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n < 1) { cudaGetLastError(); n = sms / CL / 2; }
      max_clusters = n;
    }
    int clusters = sms / CL < max_clusters ? sms / CL : max_clusters;
    grid = clusters * CL;
  }
  cfg.gridDim = dim3(grid);
  nattr += pdl_attr(attr + nattr);
  cfg.attrs = attr; cfg.numAttrs = nattr;
  PDT_CHECK(cudaLaunchKernelEx(&cfg, kern, tmX, tmW, tmY, reinterpret_cast<T*>(y), g, bs));
  return (int)cudaGetLastError();
}

template <int BN, bool RES, int CL = 1, bool PAIR = false>
int launch(const void* x, const void* w, void* y, int dtype, const HaloGeom& g, const BnStats& bs, const CUtensorMap& tmX,
           const CUtensorMap& tmW, const CUtensorMap& tmY, cudaStream_t stream) {
  (void)x; (void)w;
  if constexpr (!RES || PAIR) {
    if (dtype == PDT_FP32) return launch_t<BN, RES, float, CL, PAIR>(y, g, bs, tmX, tmW, tmY, stream);
  }
  if constexpr (BN == 64 && CL > 1 && !PAIR) {
    return PDT_E_UNSUPPORTED;                   // the 16-bit N = 64 kernel stores through TMA: no multicast-cluster form
  } else {
    if (dtype == PDT_FP16) return launch_t<BN, RES, __half, CL, PAIR>(y, g, bs, tmX, tmW, tmY, stream);
    return launch_t<BN, RES, __nv_bfloat16, CL, PAIR>(y, g, bs, tmX, tmW, tmY, stream);
  }
}

// Filter-tap multicast between CTA pairs for the streamed 128-wide variant (PDT_CONV_CLUSTER=1 disables).
int halo_cluster(const HaloGeom& g, int sms, const void* filter, cudaStream_t stream) {
  static int env = -1;
  if (env < 0) { const char* e = getenv("PDT_CONV_CLUSTER"); env = e ? atoi(e) : 0; }
  if (env == 1) return 1;
  if (g.num_tiles < 2 * sms) return 1;
  return memory_on_current_device(filter, stream) ? 2 : 1;   // the multicast tile (the filter) must be this GPU's memory
}

}  // namespace

extern "C" {

// 1 if pdt_conv3x3_halo serves this shape (3x3, stride 1, pad 1 implied).
int pdt_conv3x3_halo_supported(int dtype, int N, int H, int W, int C, int K) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return 0;
  if (C % 64 || (K != 64 && K != 128) || W + 2 > 58 || W < 6 || H < 1 || N < 1) return 0;
  return 1;
}

// x: [N][H][W][C], w: [K][3][3][C], y: [N][H][W][K] (bf16 / fp16 / fp32 as tf32).  scratch/stat pointers as in
// pdt_conv_fprop_bnstats; pass scratch == NULL for a plain convolution.
int pdt_conv3x3_halo(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, void* scratch,
                     float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                     float* save_invstd, cudaStream_t stream) {
  return pdt_conv3x3_halo_add(x, w, y, nullptr, dtype, N, H, W, C, K, scratch, nullptr, running_mean, running_var, momentum, eps,
                              save_mean, save_invstd, stream);
}

// The same with an optional addend [N][H][W][K] summed into the output in the epilogue (y = conv + addend): the
// data gradient of a residual block's first 3x3 convolution takes the shortcut gradient here.  Not combined
// with the statistics epilogue (addend != NULL requires scratch == NULL).
int pdt_conv3x3_halo_add(const void* x, const void* w, void* y, const void* addend, int dtype, int N, int H, int W, int C, int K,
                         void* scratch, const PdtCommDesc* comm, float* running_mean, float* running_var, float momentum, float eps,
                         float* save_mean, float* save_invstd, cudaStream_t stream) {
  if (!pdt_conv3x3_halo_supported(dtype, N, H, W, C, K)) return PDT_E_UNSUPPORTED;
  if (addend != nullptr && scratch != nullptr) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y) |
       reinterpret_cast<uintptr_t>(addend)) & 15) return PDT_E_UNSUPPORTED;
  if (g_enc == nullptr) {
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
    g_enc = reinterpret_cast<EncodeTiledFn>(fn);
  }
  HaloGeom g;
  g.addend = addend;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K;
  g.Wp = W + 2;
  g.TR = 128 / g.Wp;
  if (g.TR > H) g.TR = H;
  g.tiles_per_image = (H + g.TR - 1) / g.TR;
  g.num_tiles = N * g.tiles_per_image;
  const bool f32 = dtype == PDT_FP32;
  const cuuint64_t es = f32 ? 4 : 2;
  const cuuint32_t kch = f32 ? 32 : 64;          // channels per 128-byte K row
  g.c_chunks = C / (int)kch;
  g.patch_bytes = (uint32_t)((g.TR + 2) * g.Wp * 128);
  if (1024 + (128 + 2 * g.Wp + 2) * 128 > 32768 || g.patch_bytes + 1024 > 32768) return PDT_E_UNSUPPORTED;
  int sms_q = 148;
  { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms_q, cudaDevAttrMultiProcessorCount, dev); }
  // CTA pairs (cta_group::2, no multicast load involved): when every pair has at least two tiles each; PDT_HALO_PAIR=0 disables
  static int pair_env = -1;
  if (pair_env < 0) { const char* e = getenv("PDT_HALO_PAIR"); pair_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
  // (measured per launch at 1200 images: 128-wide fp16 274 -> 245 us, fp32 550 -> 465 us; 64-wide fp32 846 -> 648 us with the
  // filter bank resident as two halves; the 16-bit 64-wide kernel got SLOWER, 322 -> 386 us: its M = 256 x N = 64
  // instructions are too short for one issuing thread, so it keeps single CTAs)
  const bool pair = pair_env == 1 && g.num_tiles >= 2 * sms_q && (K == 128 || f32);
  const bool resident = (C == 64 && K == 64 && (!f32 || pair));
  const int cl = pair ? 2 : ((K == 128 || f32) ? halo_cluster(g, sms_q, w, stream) : 1);
  const CUtensorMapDataType dt = f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                     : (dtype == PDT_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
  CUtensorMap tmX, tmW;
  {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * es, (cuuint64_t)W * C * es, (cuuint64_t)H * W * C * es};
    cuuint32_t box[4] = {kch, (cuuint32_t)g.Wp, (cuuint32_t)(g.TR + 2), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (g_enc(&tmX, dt, 4, const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10003;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)9 * C, (cuuint64_t)K};
    cuuint64_t strides[1] = {(cuuint64_t)9 * C * es};
    cuuint32_t box[2] = {kch, (cuuint32_t)(K / cl)};    // with CTA pairs each CTA loads (and multicasts) half of a tap's rows
    cuuint32_t estr[2] = {1, 1};
    if (g_enc(&tmW, dt, 2, const_cast<void*>(w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10004;
  }
  CUtensorMap tmY;
  {   // output as (K, W, H, N); one box = the TR x W x 64-channel tile the N=64 kernel stages in shared memory
    cuuint64_t dims[4] = {(cuuint64_t)K, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)K * es, (cuuint64_t)W * K * es, (cuuint64_t)H * W * K * es};
    cuuint32_t box[4] = {kch, (cuuint32_t)W, (cuuint32_t)g.TR, 1};      // (only the 16-bit N = 64 kernel stores through it)
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (g_enc(&tmY, dt, 4, y, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10005;
  }
  BnStats bs;
  bs.enabled = scratch != nullptr;
  bs.counter = reinterpret_cast<unsigned*>(scratch);
  bs.partial = scratch ? reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256) : nullptr;
  bs.running_mean = running_mean; bs.running_var = running_var;
  bs.save_mean = save_mean; bs.save_invstd = save_invstd;
  bs.momentum = momentum; bs.eps = eps;
  bs.count = (long long)N * H * W;
  bs.comm = make_comm(comm);
  if (K == 64) {
    if (resident && pair) return launch<64, true, 2, true>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
    if (pair) return launch<64, false, 2, true>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
    if (resident) return launch<64, true>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
    if (f32 && cl == 2) return launch<64, false, 2>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
    return launch<64, false>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
  }
  if (pair) return launch<128, false, 2, true>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
  if (cl == 2) return launch<128, false, 2>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
  return launch<128, false>(x, w, y, dtype, g, bs, tmX, tmW, tmY, stream);
}

// 1 if pdt_conv3x3_wgrad_halo serves this weight gradient (3x3, stride 1, pad 1 implied).
int pdt_conv3x3_wgrad_halo_supported(int dtype, int N, int H, int W, int C, int K) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16 && dtype != PDT_FP32) return 0;
  if (W + 2 > 58 || W < 6 || H < 1 || N < 1) return 0;
  if (C == 64 && K == 64) return 1;
  // 128 -> 128 as four 64 x 64 channel slices, one launch each: pays in fp32, where the generic kernel is bound by
  // its per-tap im2col traffic (993 us at 1200 images against 4 x 160); in 16 bit the generic kernel is as fast
  if (C == 128 && K == 128 && dtype == PDT_FP32) {
    static int env = -1;
    if (env < 0) { const char* e = getenv("PDT_WGRAD_HALO_SLICES"); env = (e != nullptr && e[0] == '0') ? 0 : 1; }
    return env;
  }
  return 0;
}

// x: [N][H][W][64], dy: [N][H][W][64] (bf16 / fp16 / fp32 as tf32), dw: [64][3][3][64] fp32, ACCUMULATED into (zero it first).
int pdt_conv3x3_wgrad_halo(const void* x, const void* dy, float* dw, int dtype, int N, int H, int W, int C, int K,
                           cudaStream_t stream) {
  if (!pdt_conv3x3_wgrad_halo_supported(dtype, N, H, W, C, K)) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dw)) & 15) return PDT_E_UNSUPPORTED;
  if (g_enc == nullptr) {
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
    g_enc = reinterpret_cast<EncodeTiledFn>(fn);
  }
  HaloGeom g;
  g.addend = nullptr;
  g.N = N; g.H = H; g.W = W; g.C = 64; g.K = 64;
  g.dw_ct = C;
  g.Wp = W + 2;
  g.TR = 128 / g.Wp;
  if (g.TR > H) g.TR = H;
  g.tiles_per_image = (H + g.TR - 1) / g.TR;
  g.num_tiles = N * g.tiles_per_image;
  g.c_chunks = 1;
  g.patch_bytes = (uint32_t)((g.TR + 2) * g.Wp * 128);
  const bool f32 = dtype == PDT_FP32;
  // the largest view starts 2*Wp + 2 rows in and is read for up to 128 rows (16 bit: phantom tap 9, paired with tap 8;
  // fp32: phantom tap (2, 3) of the four-view instruction of the last filter row)
  if (1024 + (2 * g.Wp + 2 + 128) * 128 > WHCfg::kAStage || g.patch_bytes + 1024 > (uint32_t)WHCfg::kAStage) return PDT_E_UNSUPPORTED;
  const CUtensorMapDataType dt = f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                     : (dtype == PDT_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
  // fp32: a pixel is two 128-byte rows; the box takes 32 channels (one plane), MN-major 32-bit operands use the
  // 32-byte-atom flavour of the 128-byte swizzle
  const CUtensorMapSwizzle swz = f32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  const cuuint64_t es = f32 ? 4 : 2;
  const cuuint32_t bc = f32 ? 32 : 64;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = g.num_tiles < sms ? g.num_tiles : sms;
  const void* x_all = x;
  const void* dy_all = dy;
  float* dw_all = dw;
  for (int k0 = 0; k0 < K; k0 += 64)
  for (int c0 = 0; c0 < C; c0 += 64) {
  // one 64 x 64 channel slice: the tensor maps see 64 of the layer's channels (pixel stride = the whole pixel)
  x = reinterpret_cast<const char*>(x_all) + (size_t)c0 * es;
  dy = reinterpret_cast<const char*>(dy_all) + (size_t)k0 * es;
  dw = dw_all + (size_t)k0 * 9 * C + c0;
  CUtensorMap tmX, tmDY;
  cuuint32_t estr[4] = {1, 1, 1, 1};
  {
    cuuint64_t dims[4] = {64, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * es, (cuuint64_t)W * C * es, (cuuint64_t)H * W * C * es};
    cuuint32_t box[4] = {bc, (cuuint32_t)g.Wp, (cuuint32_t)(g.TR + 2), 1};
    if (g_enc(&tmX, dt, 4, const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10006;
  }
  {
    cuuint64_t dims[4] = {64, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)K * es, (cuuint64_t)W * K * es, (cuuint64_t)H * W * K * es};
    cuuint32_t box[4] = {bc, (cuuint32_t)g.Wp, (cuuint32_t)g.TR, 1};
    if (g_enc(&tmDY, dt, 4, const_cast<void*>(dy), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10007;
  }
  if (f32) {
    auto kern = k_wgrad3x3_halo_tf32;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, WHCfg32::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(128), WHCfg32::kSmemBytes, stream, tmX, tmDY, dw, g));
  } else if (dtype == PDT_FP16) {
    auto kern = k_wgrad3x3_halo<__half>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, WHCfg::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(128), WHCfg::kSmemBytes, stream, tmX, tmDY, dw, g));
  } else {
    auto kern = k_wgrad3x3_halo<__nv_bfloat16>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, WHCfg::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(128), WHCfg::kSmemBytes, stream, tmX, tmDY, dw, g));
  }
  }   // channel slices
  return (int)cudaGetLastError();
}

// 1 if pdt_stem_wgrad serves this stem weight gradient (16-bit, 64 output channels, output rows of 16..128 pixels in
// whole 16-pixel steps).  Not used in deterministic mode (it accumulates with red.add).
int pdt_stem_wgrad_supported(int dtype, int N, int HP, int WP, int K) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16) return 0;
  const int P = HP - 3, Q = WP - 3;
  if (K != 64 || N < 1 || P < 1 || Q < 16 || Q > 128 || Q % 16) return 0;
  if ((SWCfg::kTR + 3) * WP + 8 > 1024) return 0;
  if ((long long)N * HP * WP > 2000000000ll || (long long)N * HP * WP < 256) return 0;
  return 1;
}

// xs: [N][HP][WP][16] space-to-depth image, dy: [N][HP-3][WP-3][64], dwp: [64][4][4][16] fp32, ACCUMULATED into.
int pdt_stem_wgrad(const void* xs, const void* dy, float* dwp, int dtype, int N, int HP, int WP, int K, cudaStream_t stream) {
  if (!pdt_stem_wgrad_supported(dtype, N, HP, WP, K)) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(xs) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dwp)) & 15) return PDT_E_UNSUPPORTED;
  if (g_enc == nullptr) {
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
    g_enc = reinterpret_cast<EncodeTiledFn>(fn);
  }
  StemWGeom g;
  g.N = N; g.HP = HP; g.WP = WP; g.P = HP - 3; g.Q = WP - 3;
  g.tiles_per_image = (g.P + SWCfg::kTR - 1) / SWCfg::kTR;
  g.num_tiles = N * g.tiles_per_image;
  const CUtensorMapDataType dt = dtype == PDT_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMap tmX, tmDY;
  {
    cuuint64_t dims[2] = {16, (cuuint64_t)N * HP * WP};
    cuuint64_t strides[1] = {32};
    cuuint32_t box[2] = {16, 256};
    cuuint32_t estr[2] = {1, 1};
    if (g_enc(&tmX, dt, 2, const_cast<void*>(xs), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10011;
  }
  {
    cuuint64_t dims[3] = {64, (cuuint64_t)g.P * g.Q, (cuuint64_t)N};
    cuuint64_t strides[2] = {128, (cuuint64_t)g.P * g.Q * 128};
    cuuint32_t box[3] = {64, (cuuint32_t)(2 * g.Q), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    if (g_enc(&tmDY, dt, 3, const_cast<void*>(dy), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10012;
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = g.num_tiles < sms ? g.num_tiles : sms;
  if (dtype == PDT_FP16) {
    auto kern = k_stem_wgrad<__half>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SWCfg::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(128), SWCfg::kSmemBytes, stream, tmX, tmDY, dwp, g));
  } else {
    auto kern = k_stem_wgrad<__nv_bfloat16>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SWCfg::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(128), SWCfg::kSmemBytes, stream, tmX, tmDY, dwp, g));
  }
  return (int)cudaGetLastError();
}

// 1 if pdt_stem_conv serves this stem (space-to-depth image of WP pixels per row, K output channels).
int pdt_stem_conv_supported(int dtype, int N, int HP, int WP, int K) {
  if (dtype != PDT_BF16 && dtype != PDT_FP16) return 0;
  if (K != 64 || WP < 4 || HP < 4 || WP > 127 || N < 1) return 0;      // WP <= 127: row fits a 128-slot tile and the 512-pixel patch
  if ((long long)N * HP * WP > 2000000000ll) return 0;
  if ((long long)N * HP * WP < 512) return 0;      // the patch loads use 256-pixel boxes: tiny inputs take the generic kernel
  return 1;
}

// xs: [N][HP][WP][16] space-to-depth image (pdt_stem_s2d), wp: [64][4][4][16] filter, y: [N][HP-3][WP-3][64].
// scratch / statistics pointers as in pdt_conv_fprop_bnstats; scratch == NULL for a plain convolution.
int pdt_stem_conv(const void* xs, const void* wp, void* y, int dtype, int N, int HP, int WP, int K, void* scratch,
                  const PdtCommDesc* comm, float* running_mean, float* running_var, float momentum, float eps, float* save_mean, float* save_invstd,
                  cudaStream_t stream) {
  if (!pdt_stem_conv_supported(dtype, N, HP, WP, K)) return PDT_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(xs) | reinterpret_cast<uintptr_t>(wp) | reinterpret_cast<uintptr_t>(y)) & 15) return PDT_E_UNSUPPORTED;
  if (g_enc == nullptr) {
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
    g_enc = reinterpret_cast<EncodeTiledFn>(fn);
  }
  StemGeom g;
  g.N = N; g.HP = HP; g.WP = WP; g.P = HP - 3; g.Q = WP - 3; g.K = K;
  g.slots = (long long)N * HP * WP;
  g.num_tiles = N * g.P;
  const CUtensorMapDataType dt = dtype == PDT_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMap tmX, tmW, tmY;
  cuuint32_t estr[2] = {1, 1};
  {   // output rows: (channel, q, n * P + p); one box = one output row
    cuuint64_t dims[3] = {64, (cuuint64_t)g.Q, (cuuint64_t)N * g.P};
    cuuint64_t strides[2] = {128, (cuuint64_t)g.Q * 128};
    cuuint32_t box[3] = {64, (cuuint32_t)g.Q, 1};
    cuuint32_t estr3[3] = {1, 1, 1};
    if (g_enc(&tmY, dt, 3, y, dims, strides, box, estr3, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10010;
  }
  {
    cuuint64_t dims[2] = {16, (cuuint64_t)g.slots};
    cuuint64_t strides[1] = {32};
    cuuint32_t box[2] = {16, 256};
    if (g_enc(&tmX, dt, 2, const_cast<void*>(xs), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10008;
  }
  {
    cuuint64_t dims[2] = {256, 64};
    cuuint64_t strides[1] = {512};
    cuuint32_t box[2] = {16, 64};
    if (g_enc(&tmW, dt, 2, const_cast<void*>(wp), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10009;
  }
  BnStats bs;
  bs.enabled = scratch != nullptr;
  bs.counter = reinterpret_cast<unsigned*>(scratch);
  bs.partial = scratch ? reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256) : nullptr;
  bs.running_mean = running_mean; bs.running_var = running_var;
  bs.save_mean = save_mean; bs.save_invstd = save_invstd;
  bs.momentum = momentum; bs.eps = eps;
  bs.count = (long long)N * g.P * g.Q;
  bs.comm = make_comm(comm);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = g.num_tiles < sms ? g.num_tiles : sms;
  if (dtype == PDT_FP16) {
    auto kern = k_stem_conv<__half>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SCfg::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(kThreads), SCfg::kSmemBytes, stream, tmX, tmW, tmY, g, bs));
  } else {
    auto kern = k_stem_conv<__nv_bfloat16>;
    PDT_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SCfg::kSmemBytes));
    PDT_CHECK(launch_pdl(kern, dim3(grid), dim3(kThreads), SCfg::kSmemBytes, stream, tmX, tmW, tmY, g, bs));
  }
  return (int)cudaGetLastError();
}

}  // extern "C"
