// tcgen05 GEMM whose operands are *gathered* into shared memory by the CTA's threads
// (cp.async 16-byte chunks into the no-swizzle UMMA core-matrix layout) instead of tiled
// by TMA. This is the building block behind every layer-sized GEMM of the reference CNN
// (Dense 1600->128 fwd/dgrad/wgrad, /root/reference/tf_dist_example.py:46-47), where the
// operand is either tiny, transposed, or an im2col view, and it accepts all four
// operand-major combinations (NT / NN / TN / TT) so backward passes need no transposes.
//
//   D[M,N] (fp32) (+)= A[M,K] * B[N,K]^T        A, B bf16, fp32 accumulation in TMEM
//     a_mn = 0: A stored row-major [M][K] (K-major)      a_mn = 1: stored [K][M] (MN-major)
//     b_mn = 0: B stored row-major [N][K] (K-major)      b_mn = 1: stored [K][N] (MN-major)
//
// One CTA = one 128 x n_tile accumulator tile; 128 threads. Thread 0 issues the MMAs, all
// four warps drain TMEM in the epilogue (bias / ReLU / split-K atomics optional).
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

struct GatherGemmParams {
  const __nv_bfloat16* A;
  const __nv_bfloat16* B;
  float* D;               // fp32 output (may be null if Dh given)
  __nv_bfloat16* Dh;      // optional bf16 output
  const float* bias;      // optional, length N
  int M, N, K;
  int lda, ldb, ldd;
  int a_mn, b_mn;
  int n_tile;             // multiple of 16, <= 256
  int k_block;            // multiple of 16; K is processed in blocks of this many columns
  int k_split;            // gridDim.z; each z handles a contiguous range of k-blocks
  int relu;
  int atomic_out;         // accumulate into D with red.add (split-K)
  int trans_out;          // write D transposed: element (row, col) goes to D[col*ldd + row]
};

__global__ void __launch_bounds__(128, 1) gather_gemm_kernel(GatherGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int m0 = blockIdx.x * 128;
  const int n0 = blockIdx.y * p.n_tile;
  const int kb_total = (p.K + p.k_block - 1) / p.k_block;
  const int kb_per = (kb_total + p.k_split - 1) / p.k_split;
  const int kb_begin = blockIdx.z * kb_per;
  const int kb_end = min(kb_total, kb_begin + kb_per);

  const uint32_t a_smem = smem_u32(smem);
  const uint32_t b_smem = a_smem + 128u * p.k_block * 2u;
  const uint32_t ncols = tmem_cols_for(p.n_tile);

  if (warp == 0) tmem_alloc(&tmem_slot, ncols);
  if (threadIdx.x == 32) {
    mbar_init(&mma_bar, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_d = tmem_slot;
  const uint32_t idesc = umma_idesc_bf16(128, p.n_tile, p.a_mn != 0, p.b_mn != 0);

  uint32_t phase = 0;
  bool any = false;
  for (int kb = kb_begin; kb < kb_end; ++kb) {
    const int k0 = kb * p.k_block;
    const int kcols = min(p.k_block, p.K - k0);  // multiple of 16 (host-checked)
    if (any) {
      // the previous block's MMAs must have finished reading smem before we overwrite it
      mbar_wait(&mma_bar, phase);
      phase ^= 1;
    }
    if (p.a_mn) stage_mnmajor(a_smem, p.A, p.lda, m0, p.M, k0, 128, kcols);
    else        stage_kmajor(a_smem, p.A, p.lda, m0, p.M, k0, 128, kcols);
    if (p.b_mn) stage_mnmajor(b_smem, p.B, p.ldb, n0, p.N, k0, p.n_tile, kcols);
    else        stage_kmajor(b_smem, p.B, p.ldb, n0, p.N, k0, p.n_tile, kcols);
    cp_async_wait_all();
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after_sync();
      const uint32_t sbo = (uint32_t)(kcols >> 3) * 128u;
      for (int ks = 0; ks < (kcols >> 4); ++ks) {
        uint64_t ad = umma_smem_desc(a_smem + ks * 256u, 128u, sbo, SWZ_NONE);
        uint64_t bd = umma_smem_desc(b_smem + ks * 256u, 128u, sbo, SWZ_NONE);
        umma_bf16(tmem_d, ad, bd, idesc, any || ks > 0);
      }
      umma_commit(&mma_bar);
    }
    any = true;
  }
  if (any) {
    mbar_wait(&mma_bar, phase);
    tc_fence_after_sync();
    const int row = m0 + warp * 32 + (int)lane_id();
    const uint32_t t_row = tmem_d + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < p.n_tile; c += 32) {
      uint32_t v[32];
      if (p.n_tile - c >= 32) {
        tmem_ld32(t_row + c, v);
      } else {
        uint32_t v16[16];
        tmem_ld16(t_row + c, v16);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = v16[i];
      }
      tmem_ld_wait();
      const int ncol = min(32, p.n_tile - c);
      if (row < p.M) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          if (i >= ncol) break;
          const int col = n0 + c + i;
          if (col >= p.N) break;
          float x = __uint_as_float(v[i]);
          const size_t o = p.trans_out ? (size_t)col * p.ldd + row : (size_t)row * p.ldd + col;
          if (p.atomic_out) {
            atomicAdd(p.D + o, x);
          } else {
            if (p.bias) x += p.bias[p.trans_out ? row : col];
            if (p.relu) x = fmaxf(x, 0.f);
            if (p.D) p.D[o] = x;
            if (p.Dh) p.Dh[o] = __float2bfloat16(x);
          }
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, ncols);
}

}  // namespace tfdl

extern "C" int tfdl_gemm_bf16_gather_ex(const void* A, const void* B, float* D, void* Dh, const float* bias, int M,
                                        int N, int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int relu,
                                        int k_split, int n_tile, int k_block, int trans_out, void* stream);

extern "C" int tfdl_gemm_bf16_gather(const void* A, const void* B, float* D, void* Dh, const float* bias, int M,
                                     int N, int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int relu,
                                     int k_split, void* stream) {
  return tfdl_gemm_bf16_gather_ex(A, B, D, Dh, bias, M, N, K, lda, ldb, ldd, a_mn, b_mn, relu, k_split, 0, 0, 0,
                                  stream);
}

extern "C" int tfdl_gemm_bf16_gather_ex(const void* A, const void* B, float* D, void* Dh, const float* bias, int M,
                                        int N, int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int relu,
                                        int k_split, int n_tile, int k_block, int trans_out, void* stream) {
  using namespace tfdl;
  TFDL_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_gather: empty problem");
  TFDL_REQUIRE(K % 16 == 0, "gemm_gather: K=%d must be a multiple of 16", K);
  TFDL_REQUIRE(!a_mn || (M % 8 == 0 && lda % 8 == 0), "gemm_gather: MN-major A needs M,lda %% 8 == 0");
  TFDL_REQUIRE(!b_mn || (N % 8 == 0 && ldb % 8 == 0), "gemm_gather: MN-major B needs N,ldb %% 8 == 0");
  TFDL_REQUIRE(a_mn || lda % 8 == 0, "gemm_gather: lda must be a multiple of 8");
  TFDL_REQUIRE(b_mn || ldb % 8 == 0, "gemm_gather: ldb must be a multiple of 8");
  GatherGemmParams p{};
  p.A = (const __nv_bfloat16*)A; p.B = (const __nv_bfloat16*)B; p.D = D; p.Dh = (__nv_bfloat16*)Dh;
  p.bias = bias; p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldd = ldd;
  p.a_mn = a_mn; p.b_mn = b_mn; p.relu = relu;
  int n_pad = (N + 15) / 16 * 16;
  p.n_tile = n_tile > 0 ? n_tile : (n_pad <= 256 ? n_pad : 128);
  TFDL_REQUIRE(p.n_tile % 16 == 0 && p.n_tile <= 256, "gemm_gather: n_tile=%d must be a multiple of 16, <= 256",
               p.n_tile);
  // K block: as large as fits in ~192 KB with both operands resident
  int kb = (192 * 1024) / (2 * (128 + p.n_tile));
  kb = kb / 16 * 16;
  if (k_block > 0) kb = k_block;
  if (kb > K) kb = K;
  if (kb > 512) kb = 512;
  TFDL_REQUIRE(kb % 16 == 0 && (size_t)(128 + p.n_tile) * kb * 2 <= 200 * 1024, "gemm_gather: bad k_block %d", kb);
  p.k_block = kb;
  p.k_split = k_split < 1 ? 1 : k_split;
  p.atomic_out = p.k_split > 1;
  p.trans_out = trans_out;
  TFDL_REQUIRE(!p.atomic_out || (D && !Dh && !bias && !relu),
               "gemm_gather: split-K accumulates into fp32 D only (caller zeroes D, applies bias later)");
  size_t smem = (size_t)(128 + p.n_tile) * p.k_block * 2;
  TFDL_CUDA_OK(cudaFuncSetAttribute(gather_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)(200 * 1024)));
  dim3 grid((M + 127) / 128, (N + p.n_tile - 1) / p.n_tile, p.k_split);
  gather_gemm_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>(p);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
