// Fused self-attention for S = 128 tokens, head dim 64 (BERT-base at the sequence length of BASELINE.json config 4;
// "a plain fused QK^T-softmax-V for 12 heads x 64 dims at S = 128"). The reference repo has no attention.
//
// One CTA per (sequence, head). Q, K and V are 128 x 64 bf16 tiles cut by TMA straight out of the fused QKV
// projection [B*S][3H] (no head split / transpose copies), both GEMMs of the forward and all five of the backward run
// on tcgen05 with accumulators in TMEM, and the softmax lives in registers: with M = 128 every thread owns one full
// row of the 128 x 128 score tile (TMEM lane = query row), so max / sum need no cross-thread reduction.
//
// The same shared-memory bytes serve as different operands (cf. tests/test_probe_gpu.py):
//   a [128 rows][64 cols] 128B-swizzled tile is a K-major operand (rows = M or N, cols = K) AND an MN-major operand
//   (rows = K, cols = 64 contiguous M or N);  P / dS tiles are two such 64-column blocks 16 KB apart.
//   forward : S = Q K^T (Q, K K-major)            O  = P V      (P K-major A, V MN-major B)
//   backward: S = Q K^T, dP = dO V^T (V K-major)  dV = P^T dO   (P MN-major A, dO MN-major B)
//             dQ = dS K (dS K-major A, K MN-major B)            dK = dS^T Q (dS MN-major A, Q MN-major B)
#include <cuda.h>
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

int make_tmap_bf16_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle);

constexpr int AT_S = 128, AT_D = 64, AT_THREADS = 128;
constexpr uint32_t AT_TILE = AT_S * AT_D * 2;  // 16 KB: one [128][64] bf16 tile

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2f(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Store 32 consecutive columns [c0, c0 + 32) of row `row` of a [128][128] bf16 tile kept as two 128B-swizzled
// [128][64] blocks (16 KB apart): 16-byte chunk j of a row sits at chunk (j ^ (row & 7)).
__device__ __forceinline__ void st_row32_sw128(uint32_t tile, int row, int c0, const float (&f)[32]) {
  const uint32_t base = tile + (uint32_t)(c0 >> 6) * AT_TILE + (uint32_t)row * 128;
  const int j0 = (c0 & 63) >> 3;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    st_shared_v4(base + ((uint32_t)((j0 + j) ^ (row & 7)) << 4), pack_bf16x2(f[8 * j], f[8 * j + 1]),
                 pack_bf16x2(f[8 * j + 2], f[8 * j + 3]), pack_bf16x2(f[8 * j + 4], f[8 * j + 5]),
                 pack_bf16x2(f[8 * j + 6], f[8 * j + 7]));
}
// descriptors of the two operand views of a [128][64] swizzled tile (k = K-step of 16)
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile, int k) { return umma_smem_desc(tile + k * 32, 16, 1024, SWZ_128B); }
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile, int k, uint32_t lbo) {
  return umma_smem_desc(tile + k * 2048, lbo, 1024, SWZ_128B);
}

// -------------------------------------------------------------------------------------------------------------
// forward
// -------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(AT_THREADS)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out, float* __restrict__ lse2,
                     int H, int heads, float scale_log2e) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_tma, bar_s, bar_o;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, row = threadIdx.x;
  const int b = blockIdx.x / heads, h = blockIdx.x % heads;
  const uint32_t sQ = smem_u32(smem), sK = sQ + AT_TILE, sV = sK + AT_TILE, sP = sV + AT_TILE;  // sP: 2 tiles
  if (warp == 0) tmem_alloc(&slot, 256);
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_s, 1);
    mbar_init(&bar_o, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, 3 * AT_TILE);
    tma_load_2d(sQ, &tmQKV, h * AT_D, b * AT_S, &bar_tma);
    tma_load_2d(sK, &tmQKV, H + h * AT_D, b * AT_S, &bar_tma);
    tma_load_2d(sV, &tmQKV, 2 * H + h * AT_D, b * AT_S, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, 128, false, false);
#pragma unroll
    for (int k = 0; k < AT_D / 16; ++k) umma_bf16(tmem, desc_kmajor(sQ, k), desc_kmajor(sK, k), idesc, k != 0);
    umma_commit(&bar_s);
  }
  mbar_wait(&bar_s, 0);
  tc_fence_after_sync();
  const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);
  // pass 1: row maximum (the row is re-read from TMEM in pass 2 instead of being held in 128 registers)
  float mx = -INFINITY;
#pragma unroll 1
  for (int c = 0; c < AT_S; c += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + c, v);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
  }
  mx *= scale_log2e;
  float sum = 0.f;
#pragma unroll 1
  for (int c = 0; c < AT_S; c += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + c, v);
    tmem_ld_wait();
    float p[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      p[i] = ex2f(fmaf(__uint_as_float(v[i]), scale_log2e, -mx));
      // the sum uses the bf16-rounded value the tensor core will see, so that P / sum is normalised exactly
      p[i] = __bfloat162float(__float2bfloat16(p[i]));
      sum += p[i];
    }
    st_row32_sw128(sP, row, c, p);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, AT_D, false, true);
#pragma unroll
    for (int k = 0; k < AT_S / 16; ++k)
      umma_bf16(tmem + 128, desc_kmajor(sP + (k >> 2) * AT_TILE, k & 3), desc_mnmajor(sV, k, AT_TILE), idesc, k != 0);
    umma_commit(&bar_o);
  }
  mbar_wait(&bar_o, 0);
  tc_fence_after_sync();
  const float inv = 1.f / sum;
  __nv_bfloat16* dst = out + ((size_t)(b * AT_S + row)) * H + h * AT_D;
#pragma unroll
  for (int c = 0; c < AT_D; c += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + 128 + c, v);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      uint4 o = make_uint4(pack_bf16x2(__uint_as_float(v[i]) * inv, __uint_as_float(v[i + 1]) * inv),
                           pack_bf16x2(__uint_as_float(v[i + 2]) * inv, __uint_as_float(v[i + 3]) * inv),
                           pack_bf16x2(__uint_as_float(v[i + 4]) * inv, __uint_as_float(v[i + 5]) * inv),
                           pack_bf16x2(__uint_as_float(v[i + 6]) * inv, __uint_as_float(v[i + 7]) * inv));
      *reinterpret_cast<uint4*>(dst + c + i) = o;
    }
  }
  if (lse2) lse2[(size_t)blockIdx.x * AT_S + row] = mx + lg2f(sum);
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// -------------------------------------------------------------------------------------------------------------
// backward: dQKV [B*S][3H] from dO [B*S][H], recomputing P from Q, K and the saved log-sum-exp
// -------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_out64(__nv_bfloat16* dst, uint32_t t_addr) {
#pragma unroll
  for (int c = 0; c < AT_D; c += 32) {
    uint32_t v[32];
    tmem_ld32(t_addr + c, v);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; i += 8)
      *reinterpret_cast<uint4*>(dst + c + i) =
          make_uint4(pack_bf16x2(__uint_as_float(v[i]), __uint_as_float(v[i + 1])),
                     pack_bf16x2(__uint_as_float(v[i + 2]), __uint_as_float(v[i + 3])),
                     pack_bf16x2(__uint_as_float(v[i + 4]), __uint_as_float(v[i + 5])),
                     pack_bf16x2(__uint_as_float(v[i + 6]), __uint_as_float(v[i + 7])));
  }
}

__global__ void __launch_bounds__(AT_THREADS)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                     const float* __restrict__ lse2, __nv_bfloat16* __restrict__ dqkv, int H, int heads, float scale,
                     float scale_log2e) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_tma, bar_s, bar_g;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, row = threadIdx.x;
  const int b = blockIdx.x / heads, h = blockIdx.x % heads;
  const uint32_t sQ = smem_u32(smem), sK = sQ + AT_TILE, sV = sK + AT_TILE, sDO = sV + AT_TILE, sP = sDO + AT_TILE,
                 sDS = sP + 2 * AT_TILE;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DQ = 320, T_DK = 384;
  if (warp == 0) tmem_alloc(&slot, 512);
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_s, 1);
    mbar_init(&bar_g, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, 4 * AT_TILE);
    tma_load_2d(sQ, &tmQKV, h * AT_D, b * AT_S, &bar_tma);
    tma_load_2d(sK, &tmQKV, H + h * AT_D, b * AT_S, &bar_tma);
    tma_load_2d(sV, &tmQKV, 2 * H + h * AT_D, b * AT_S, &bar_tma);
    tma_load_2d(sDO, &tmDO, h * AT_D, b * AT_S, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, 128, false, false);
#pragma unroll
    for (int k = 0; k < AT_D / 16; ++k) umma_bf16(tmem + T_S, desc_kmajor(sQ, k), desc_kmajor(sK, k), idesc, k != 0);
#pragma unroll
    for (int k = 0; k < AT_D / 16; ++k) umma_bf16(tmem + T_DP, desc_kmajor(sDO, k), desc_kmajor(sV, k), idesc, k != 0);
    umma_commit(&bar_s);
  }
  const float l2 = lse2[(size_t)blockIdx.x * AT_S + row];
  mbar_wait(&bar_s, 0);
  tc_fence_after_sync();
  const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);
  // pass 1: D = sum_j P_qj dP_qj   (== rowsum(dO * O))
  float D = 0.f;
#pragma unroll 1
  for (int c = 0; c < AT_S; c += 32) {
    uint32_t s[32], g[32];
    tmem_ld32(t_row + T_S + c, s);
    tmem_ld32(t_row + T_DP + c, g);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) D = fmaf(ex2f(fmaf(__uint_as_float(s[i]), scale_log2e, -l2)), __uint_as_float(g[i]), D);
  }
  // pass 2: P and dS = scale * P * (dP - D) as bf16 operand tiles
#pragma unroll 1
  for (int c = 0; c < AT_S; c += 32) {
    uint32_t s[32], g[32];
    tmem_ld32(t_row + T_S + c, s);
    tmem_ld32(t_row + T_DP + c, g);
    tmem_ld_wait();
    float p[32], ds[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      p[i] = ex2f(fmaf(__uint_as_float(s[i]), scale_log2e, -l2));
      ds[i] = scale * p[i] * (__uint_as_float(g[i]) - D);
    }
    st_row32_sw128(sP, row, c, p);
    st_row32_sw128(sDS, row, c, ds);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (threadIdx.x == 0) {
    constexpr uint32_t i_tt = umma_idesc_bf16(128, AT_D, true, true), i_nt = umma_idesc_bf16(128, AT_D, false, true);
#pragma unroll
    for (int k = 0; k < AT_S / 16; ++k)  // dV[kv][d] = sum_q P[q][kv] dO[q][d]
      umma_bf16(tmem + T_DV, desc_mnmajor(sP, k, AT_TILE), desc_mnmajor(sDO, k, AT_TILE), i_tt, k != 0);
#pragma unroll
    for (int k = 0; k < AT_S / 16; ++k)  // dQ[q][d] = sum_kv dS[q][kv] K[kv][d]
      umma_bf16(tmem + T_DQ, desc_kmajor(sDS + (k >> 2) * AT_TILE, k & 3), desc_mnmajor(sK, k, AT_TILE), i_nt, k != 0);
#pragma unroll
    for (int k = 0; k < AT_S / 16; ++k)  // dK[kv][d] = sum_q dS[q][kv] Q[q][d]
      umma_bf16(tmem + T_DK, desc_mnmajor(sDS, k, AT_TILE), desc_mnmajor(sQ, k, AT_TILE), i_tt, k != 0);
    umma_commit(&bar_g);
  }
  mbar_wait(&bar_g, 0);
  tc_fence_after_sync();
  __nv_bfloat16* dst = dqkv + ((size_t)(b * AT_S + row)) * (3 * H) + h * AT_D;
  st_out64(dst, t_row + T_DQ);
  st_out64(dst + H, t_row + T_DK);
  st_out64(dst + 2 * H, t_row + T_DV);
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace tfdl

using namespace tfdl;

static int attention_tmap(CUtensorMap* tm, const void* base, long long rows, int cols) {
  uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows}, strides[1] = {(uint64_t)cols * 2};
  uint32_t box[2] = {AT_D, AT_S};
  return make_tmap_bf16_nd(tm, base, 2, dims, strides, box, 3 /* 128B swizzle */);
}

extern "C" int tfdl_attention_supported(int S, int H, int heads) {
  return S == AT_S && heads > 0 && H == heads * AT_D;
}

// qkv: bf16 [B*S][3H] (query | key | value column blocks, heads of 64 inside each), out: bf16 [B*S][H],
// lse2: fp32 [B*heads][S] (log2-domain log-sum-exp of the scaled scores; what the backward pass needs).
extern "C" int tfdl_attention_fwd(const void* qkv, void* out, float* lse2, int B, int S, int H, int heads, void* stream) {
  TFDL_REQUIRE(tfdl_attention_supported(S, H, heads) && B > 0, "attention_fwd: needs S = 128 and H = heads * 64 (got S=%d H=%d heads=%d)", S, H, heads);
  TFDL_REQUIRE((((uintptr_t)qkv | (uintptr_t)out) & 15) == 0, "attention_fwd: tensors must be 16-byte aligned");
  CUtensorMap tm;
  if (int rc = attention_tmap(&tm, qkv, (long long)B * S, 3 * H)) return rc;
  const size_t smem = 5 * AT_TILE + 1024;
  static bool attr = false;
  if (!attr) {
    TFDL_CUDA_OK(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = true;
  }
  const float scale_log2e = 1.4426950408889634f / sqrtf((float)AT_D);
  attention_fwd_kernel<<<B * heads, AT_THREADS, smem, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, lse2, H, heads,
                                                                            scale_log2e);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// dout: bf16 [B*S][H]; dqkv: bf16 [B*S][3H] (every element is written).
extern "C" int tfdl_attention_bwd(const void* qkv, const void* dout, const float* lse2, void* dqkv, int B, int S, int H,
                                  int heads, void* stream) {
  TFDL_REQUIRE(tfdl_attention_supported(S, H, heads) && B > 0, "attention_bwd: needs S = 128 and H = heads * 64");
  TFDL_REQUIRE((((uintptr_t)qkv | (uintptr_t)dout | (uintptr_t)dqkv) & 15) == 0, "attention_bwd: tensors must be 16-byte aligned");
  CUtensorMap tmQ, tmD;
  if (int rc = attention_tmap(&tmQ, qkv, (long long)B * S, 3 * H)) return rc;
  if (int rc = attention_tmap(&tmD, dout, (long long)B * S, H)) return rc;
  const size_t smem = 8 * AT_TILE + 1024;
  static bool attr = false;
  if (!attr) {
    TFDL_CUDA_OK(cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = true;
  }
  const float scale = 1.f / sqrtf((float)AT_D);
  attention_bwd_kernel<<<B * heads, AT_THREADS, smem, (cudaStream_t)stream>>>(tmQ, tmD, lse2, (__nv_bfloat16*)dqkv, H, heads,
                                                                            scale, scale * 1.4426950408889634f);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
