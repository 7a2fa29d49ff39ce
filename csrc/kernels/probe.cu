// Hardware probes used to pin down descriptor semantics that the kernels in cnn.cu rely on.
// They are tiny single-CTA kernels driven by tests/test_probe_gpu.py:
//   tfdl_probe_shifted_umma   A tile of `rows` x (64|32) bf16 is TMA-loaded with 128B/64B swizzle; the MMA
//                             reads 128 rows starting at row `shift` by offsetting the descriptor start
//                             address (optionally setting the descriptor base-offset field). This is the
//                             primitive behind im2col-free implicit GEMM: one raw tile in smem, one MMA per
//                             filter tap on a shifted window.
//   tfdl_probe_tma_coremajor  a 4-D tensor map whose strides are a permutation of a row-major matrix
//                             writes the no-swizzle UMMA core-matrix layout directly (weights via TMA).
#include <cuda.h>
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {


typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn2 encode_fn() {
  static EncodeTiledFn2 fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn2>(sym);
  }
  return fn;
}

// Generic N-d bf16 tensor map. dims/strides innermost first; strides in BYTES for dims 1..rank-1.
int make_tmap_bf16_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle /*0 none, 1 32B, 2 64B, 3 128B*/) {
  EncodeTiledFn2 fn = encode_fn();
  TFDL_REQUIRE(fn, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = swizzle == 3 ? CU_TENSOR_MAP_SWIZZLE_128B : swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFDL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(rank %d) failed with CUresult %d", rank, (int)r);
  return 0;
}

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, int c0, int c1, int c2, int c3, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar)) : "memory");
}

// D[128][N] = A[shift : shift+128][0:KC] * B[N][KC]^T ; KC = 64 (swizzle 128B) or 32 (swizzle 64B)
__global__ void __launch_bounds__(128, 1)
probe_shifted_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                          float* D, int rows, int N, int kc, int shift, int base_offset_mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_tma, bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t row_bytes = kc * 2;
  const uint32_t a_smem = smem_u32(smem);
  const uint32_t b_smem = a_smem + ((rows * row_bytes + 1023) & ~1023u);
  if (warp == 0) tmem_alloc(&slot, tmem_cols_for(N));
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, rows * row_bytes + N * row_bytes);
    for (int r0 = 0; r0 < rows; r0 += 128) tma_load_2d(a_smem + r0 * row_bytes, &tmA, 0, r0, &bar_tma);
    tma_load_2d(b_smem, &tmB, 0, 0, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    const uint32_t swz = kc == 64 ? SWZ_128B : SWZ_64B;
    const uint32_t sbo = 8 * row_bytes;  // 8 rows per swizzle atom
    const uint32_t idesc = umma_idesc_bf16(128, N, false, false);
    for (int k = 0; k < kc / 16; ++k) {
      const uint32_t a_addr = a_smem + shift * row_bytes + k * 32;
      uint64_t ad = umma_smem_desc(a_addr, 16, sbo, swz);
      if (base_offset_mode == 1) ad |= (uint64_t)((a_addr >> 7) & 7) << 49;
      uint64_t bd = umma_smem_desc(b_smem + k * 32, 16, sbo, swz);
      umma_bf16(tmem, ad, bd, idesc, k > 0);
    }
    umma_commit(&bar_mma);
  }
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  const int row = warp * 32 + (int)lane_id();
  for (int c = 0; c < N; c += 16) {
    uint32_t v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    tmem_ld_wait();
    for (int i = 0; i < 16; ++i) D[row * N + c + i] = __uint_as_float(v[i]);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, tmem_cols_for(N));
}

// Copies W[rows][K] (row-major) to smem through ONE 4-d TMA whose dims are (8 elems, 8 rows, K/8 chunks,
// rows/8 groups) - i.e. the no-swizzle K-major core-matrix order - then dumps smem to `out` linearly.
__global__ void probe_tma_coremajor_kernel(const __grid_constant__ CUtensorMap tm, uint16_t* out, int bytes) {
  extern __shared__ __align__(1024) uint8_t smem2[];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    mbar_expect_tx(&bar, bytes);
    tma_load_4d(smem_u32(smem2), &tm, 0, 0, 0, 0, &bar);
  }
  __syncthreads();
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = reinterpret_cast<uint16_t*>(smem2)[i];
}

// MN-major swizzled operands with a shift along K (the pixel index of a weight-gradient GEMM):
//   D[128][32] = sum_k A[k][m] * B[k + shift][n],  A global [Krows][64 or 128] (MN-major, 128B swizzle),
//   B global [Krows + shift][32] (MN-major, 64B swizzle). lbo_mode 0: A has 128 columns, two 64-wide blocks
//   LBO apart; lbo_mode 1: A has 64 columns and LBO = 0, so MMA rows 64..127 should replicate rows 0..63.
__global__ void __launch_bounds__(128, 1)
probe_mn_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, float* D,
                     int krows, int shift, int lbo_mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw3[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw3) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_tma, bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t a_block = krows * 128u;
  const uint32_t a_smem = smem_u32(smem);
  const uint32_t b_smem = a_smem + 2 * ((a_block + 1023) & ~1023u);
  const int brows = krows + 64;
  if (warp == 0) tmem_alloc(&slot, 32);
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    const int nblk = lbo_mode == 0 ? 2 : 1;
    mbar_expect_tx(&bar_tma, nblk * a_block + brows * 64);
    for (int b = 0; b < nblk; ++b) tma_load_2d(a_smem + b * a_block, &tmA, b * 64, 0, &bar_tma);
    tma_load_2d(b_smem, &tmB, 0, 0, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    const uint32_t idesc = umma_idesc_bf16(128, 32, true, true);
    for (int k = 0; k < krows / 16; ++k) {
      uint64_t ad = umma_smem_desc(a_smem + k * 2048u, lbo_mode == 0 ? a_block : 0u, 1024u, SWZ_128B);
      uint64_t bd = umma_smem_desc(b_smem + (shift + k * 16) * 64u, brows * 64u, 512u, SWZ_64B);
      umma_bf16(tmem, ad, bd, idesc, k > 0);
    }
    umma_commit(&bar_mma);
  }
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  const int row = warp * 32 + (int)lane_id();
  uint32_t v[32];
  tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
  tmem_ld_wait();
  for (int i = 0; i < 32; ++i) D[row * 32 + i] = __uint_as_float(v[i]);
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 32);
}

}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_probe_mn_umma(const void* A, const void* B, float* D, int krows, int shift, int lbo_mode, void* stream) {
  TFDL_REQUIRE(krows % 16 == 0 && krows <= 192 && shift >= 0 && shift <= 64, "probe_mn_umma: bad arguments");
  CUtensorMap tmA, tmB;
  const int acols = lbo_mode == 0 ? 128 : 64, brows = krows + 64;
  uint64_t dA[2] = {(uint64_t)acols, (uint64_t)krows}, sA[1] = {(uint64_t)acols * 2};
  uint32_t bA[2] = {64, (uint32_t)krows};
  if (int rc = make_tmap_bf16_nd(&tmA, A, 2, dA, sA, bA, 3)) return rc;
  uint64_t dB[2] = {32, (uint64_t)brows}, sB[1] = {64};
  uint32_t bB[2] = {32, (uint32_t)brows};
  if (int rc = make_tmap_bf16_nd(&tmB, B, 2, dB, sB, bB, 2)) return rc;
  size_t smem = 2 * ((size_t)krows * 128 + 1024) + (size_t)brows * 64 + 4096;
  TFDL_CUDA_OK(cudaFuncSetAttribute(probe_mn_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  probe_mn_umma_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(tmA, tmB, D, krows, shift, lbo_mode);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}


extern "C" int tfdl_probe_shifted_umma(const void* A, const void* B, float* D, int rows, int N, int kc, int shift,
                                       int base_offset_mode, void* stream) {
  TFDL_REQUIRE((kc == 64 || kc == 32) && rows % 128 == 0 && shift + 128 <= rows && N % 16 == 0 && N <= 256,
               "probe_shifted_umma: bad arguments");
  CUtensorMap tmA, tmB;
  uint64_t dA[2] = {(uint64_t)kc, (uint64_t)rows}, sA[1] = {(uint64_t)kc * 2};
  uint32_t bA[2] = {(uint32_t)kc, 128};
  if (int rc = make_tmap_bf16_nd(&tmA, A, 2, dA, sA, bA, kc == 64 ? 3 : 2)) return rc;
  uint64_t dB[2] = {(uint64_t)kc, (uint64_t)N};
  uint32_t bB[2] = {(uint32_t)kc, (uint32_t)N};
  if (int rc = make_tmap_bf16_nd(&tmB, B, 2, dB, sA, bB, kc == 64 ? 3 : 2)) return rc;
  size_t smem = (size_t)rows * kc * 2 + (size_t)N * kc * 2 + 4096;
  TFDL_CUDA_OK(cudaFuncSetAttribute(probe_shifted_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  probe_shifted_umma_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(tmA, tmB, D, rows, N, kc, shift, base_offset_mode);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_probe_tma_coremajor(const void* W, void* out, int rows, int K, void* stream) {
  TFDL_REQUIRE(rows % 8 == 0 && K % 8 == 0 && K / 8 <= 256 && rows / 8 <= 256, "probe_tma_coremajor: bad arguments");
  CUtensorMap tm;
  uint64_t dims[4] = {8, 8, (uint64_t)K / 8, (uint64_t)rows / 8};
  uint64_t strides[3] = {(uint64_t)K * 2, 16, (uint64_t)K * 16};
  uint32_t box[4] = {8, 8, (uint32_t)K / 8, (uint32_t)rows / 8};
  if (int rc = make_tmap_bf16_nd(&tm, W, 4, dims, strides, box, 0)) return rc;
  int bytes = rows * K * 2;
  TFDL_CUDA_OK(cudaFuncSetAttribute(probe_tma_coremajor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes + 1024));
  probe_tma_coremajor_kernel<<<1, 128, bytes, (cudaStream_t)stream>>>(tm, (uint16_t*)out, bytes);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
