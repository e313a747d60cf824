// Hardware probe (test/diagnostic only): can a K-major, 128B-swizzled A operand be read by tcgen05.mma
// from a start address that is shifted by whole 128-byte rows inside a larger tile written by ONE TMA
// load?  That is what a convolution needs to reuse one input patch for all filter taps (each tap is the
// same patch shifted by a few pixels = rows).  D[128 x 64] = X[shift : shift + 128, 0:64] * B[64 x 64]^T.
// `base_offset_mode`: 0 = descriptor base_offset field left 0, 1 = base_offset = (start_address >> 7) & 7.
#include <type_traits>
#include "tc.cuh"
#include "../../include/pdt_b200.h"

namespace {
using namespace pdt;
using namespace pdt::tc;

__global__ void __launch_bounds__(128, 1)
k_probe_shift(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmB, float* __restrict__ out,
              int shift, int base_offset_mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smX = smem;                    // 256 rows x 128 B
  uint8_t* smB = smem + 256 * 128;        // 64 rows x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 320 * 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bars[0], 320 * 128);
    tma_load_2d(smX, &tmX, &bars[0], 0, 0);
    tma_load_2d(smB, &tmB, &bars[0], 0, 0);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t a_addr = smem_u32(smX) + (uint32_t)shift * 128u;
    uint64_t da = make_sw128_desc(a_addr, 1024, 16);
    if (base_offset_mode == 1) da |= (uint64_t)((a_addr >> 7) & 7) << 49;
    const uint64_t db = make_sw128_desc(smem_u32(smB), 1024, 16);
    constexpr uint32_t idesc = make_idesc(128, 64, 1, 0, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_f16(tmem_base, da + 2 * k, db + 2 * k, idesc, k != 0);
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();
  uint32_t v[32];
  for (int c0 = 0; c0 < 64; c0 += 32) {
    tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
}  // namespace

extern "C" int pdt_probe_shifted_desc(const void* x /*[256][64] bf16*/, const void* b /*[64][64] bf16*/, float* out /*[128][64]*/,
                                      int shift, int base_offset_mode, cudaStream_t stream) {
  cudaDriverEntryPointQueryResult q;
  void* fn = nullptr;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || fn == nullptr) return (int)(e != cudaSuccess ? e : cudaErrorUnknown);
  EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(fn);
  if (shift < 0 || shift > 128) return PDT_E_UNSUPPORTED;
  CUtensorMap tmX, tmB;
  cuuint32_t estr[2] = {1, 1};
  {
    cuuint64_t dims[2] = {64, 256};
    cuuint64_t strides[1] = {128};
    cuuint32_t box[2] = {64, 256};
    if (enc(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10001;
  }
  {
    cuuint64_t dims[2] = {64, 64};
    cuuint64_t strides[1] = {128};
    cuuint32_t box[2] = {64, 64};
    if (enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(b), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 10002;
  }
  const int smem = 320 * 128 + 1024 + 64;
  PDT_CHECK(cudaFuncSetAttribute(k_probe_shift, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  k_probe_shift<<<1, 128, smem, stream>>>(tmX, tmB, out, shift, base_offset_mode);
  return (int)cudaGetLastError();
}
