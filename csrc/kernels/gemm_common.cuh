// Pieces shared by the persistent TMA + tcgen05 GEMM (gemm_tma.cu) and the implicit-GEMM convolution
// (conv_igemm.cu): problem description, the TMEM -> registers -> smem -> TMA-store epilogue, tile rasterisation.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace tfdl {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row

struct TmaGemmParams {
  float* Df;               // fp32 out (optional)
  __nv_bfloat16* Dh;       // bf16 out (optional)
  const float* bias;       // optional
  int M, N, K, ldd;
  int act;                 // 0 none, 1 relu, 2 gelu(tanh)
  int m_tiles, n_tiles;
  int a_mn, b_mn;          // operand stored [K][rows] (MN-major) instead of [rows][K]
  int ksplit;              // split-K factor (> 1: fp32 output accumulated with TMA reduce-add into a zeroed buffer)
  int tma_store;           // epilogue through smem + TMA store (needs a 16-byte-multiple output pitch)
};

constexpr uint32_t EPI_STAGE_BYTES = 4096;                  // 32 rows x 128 B, per warp per buffer
constexpr uint32_t EPI_SMEM_BYTES = 4 * 2 * EPI_STAGE_BYTES;  // 4 epilogue warps, double buffered

// Activation as a compile-time choice: with a run-time switch per element the compiler evaluates the
// tanh path unconditionally and the four epilogue warps, not the tensor core, set the tile time.
template <int ACT>
__device__ __forceinline__ float act_fn(float x) {
  if (ACT == 1) return fmaxf(x, 0.f);
  if (ACT == 2) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float t, u = k0 * (x + k1 * x * x * x);
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
    return 0.5f * x * (1.f + t);
  }
  return x;
}
__device__ __forceinline__ uint32_t pack2(uint32_t lo, uint32_t hi) {
  return pack_bf16x2(__uint_as_float(lo), __uint_as_float(hi));
}
// v[i] <- act(v[i] + bias[col0 + i]) for 32 accumulator columns held as fp32 bit patterns.
template <int ACT, bool BIAS>
__device__ __forceinline__ void epi_apply_t(uint32_t (&v)[32], const float* bias, int col0, int N) {
  const bool vec = BIAS && col0 + 32 <= N && ((reinterpret_cast<uintptr_t>(bias + col0) & 15) == 0);
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (BIAS) {
      if (vec) {
        b = __ldg(reinterpret_cast<const float4*>(bias + col0 + i));
      } else {
        b.x = col0 + i < N ? __ldg(bias + col0 + i) : 0.f;
        b.y = col0 + i + 1 < N ? __ldg(bias + col0 + i + 1) : 0.f;
        b.z = col0 + i + 2 < N ? __ldg(bias + col0 + i + 2) : 0.f;
        b.w = col0 + i + 3 < N ? __ldg(bias + col0 + i + 3) : 0.f;
      }
    }
    v[i] = __float_as_uint(act_fn<ACT>(__uint_as_float(v[i]) + b.x));
    v[i + 1] = __float_as_uint(act_fn<ACT>(__uint_as_float(v[i + 1]) + b.y));
    v[i + 2] = __float_as_uint(act_fn<ACT>(__uint_as_float(v[i + 2]) + b.z));
    v[i + 3] = __float_as_uint(act_fn<ACT>(__uint_as_float(v[i + 3]) + b.w));
  }
}
__device__ __forceinline__ void epi_apply(uint32_t (&v)[32], const float* bias, int col0, int N, int act) {
  if (bias) {
    if (act == 1) epi_apply_t<1, true>(v, bias, col0, N);
    else if (act == 2) epi_apply_t<2, true>(v, bias, col0, N);
    else epi_apply_t<0, true>(v, bias, col0, N);
  } else {
    if (act == 1) epi_apply_t<1, false>(v, bias, col0, N);
    else if (act == 2) epi_apply_t<2, false>(v, bias, col0, N);
  }
}

// Epilogue of one accumulator tile for one warp: 32 rows (TMEM lanes of this warp) x BN_T columns starting at
// global element (row0, n0). Shared by the 1-CTA and the CTA-pair kernels.
template <int BN_T>
__device__ __forceinline__ void epilogue_tile(const TmaGemmParams& p, const CUtensorMap& tmD, uint32_t t_row,
                                              uint32_t stg_base, uint32_t& sb, int row0, int n0, bool first_split) {
  constexpr int BN = BN_T;
  const int row = row0 + (int)lane_id();
  const float* bias = first_split ? p.bias : nullptr;  // split-K: only one split adds the bias
  if (p.tma_store) {
    const int CH = p.Dh ? 64 : 32;  // output columns per 128-byte staged row
    const uint32_t lane = lane_id();
#pragma unroll 1
    for (int c = 0; c < BN && row0 < p.M; c += CH) {
      const int col0 = n0 + c;
      if (col0 >= p.N) break;
      uint32_t v[32], v2[32];
      tmem_ld32(t_row + c, v);
      if (p.Dh) tmem_ld32(t_row + c + 32, v2);
      tmem_ld_wait();
      epi_apply(v, bias, col0, p.N, p.act);
      if (p.Dh) epi_apply(v2, bias, col0 + 32, p.N, p.act);
      // the bulk store that last read this staging buffer (two chunks ago) must have drained it
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncwarp();
      const uint32_t sbuf = stg_base + sb * EPI_STAGE_BYTES;
      const uint32_t srow = sbuf + lane * 128;
      if (p.Dh) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          st_shared_v4(srow + ((uint32_t)(j ^ (lane & 7)) << 4), pack2(v[j * 8], v[j * 8 + 1]),
                       pack2(v[j * 8 + 2], v[j * 8 + 3]), pack2(v[j * 8 + 4], v[j * 8 + 5]),
                       pack2(v[j * 8 + 6], v[j * 8 + 7]));
          st_shared_v4(srow + ((uint32_t)((j + 4) ^ (lane & 7)) << 4), pack2(v2[j * 8], v2[j * 8 + 1]),
                       pack2(v2[j * 8 + 2], v2[j * 8 + 3]), pack2(v2[j * 8 + 4], v2[j * 8 + 5]),
                       pack2(v2[j * 8 + 6], v2[j * 8 + 7]));
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          st_shared_v4(srow + ((uint32_t)(j ^ (lane & 7)) << 4), v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (p.ksplit > 1) tma_reduce_add_2d(&tmD, sbuf, col0, row0);
        else tma_store_2d(&tmD, sbuf, col0, row0);
        tma_store_commit();
      }
      sb ^= 1;
    }
  } else {
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t v[32];
      tmem_ld32(t_row + c, v);
      tmem_ld_wait();
      const int col0 = n0 + c;
      if (row < p.M && col0 < p.N) {
        epi_apply(v, bias, col0, p.N, p.act);
        const bool full = (col0 + 32 <= p.N);
        if (p.Dh) {
          __nv_bfloat16* dst = p.Dh + (size_t)row * p.ldd + col0;
          if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
            for (int i = 0; i < 32; i += 8)
              *reinterpret_cast<uint4*>(dst + i) = make_uint4(pack2(v[i], v[i + 1]), pack2(v[i + 2], v[i + 3]),
                                                              pack2(v[i + 4], v[i + 5]), pack2(v[i + 6], v[i + 7]));
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col0 + i < p.N) dst[i] = __float2bfloat16(__uint_as_float(v[i]));
          }
        }
        if (p.Df) {
          float* dst = p.Df + (size_t)row * p.ldd + col0;
          if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) *reinterpret_cast<uint4*>(dst + i) = make_uint4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col0 + i < p.N) dst[i] = __uint_as_float(v[i]);
          }
        }
      }
    }
  }
}

// Tile order: GROUP_M row-tiles are swept before moving to the next column-tile, so one wave of CTAs works on a
// roughly square patch of the output and re-reads far fewer operand bytes from HBM than a column sweep would.
__device__ __forceinline__ void tile_coords(int t, int m_tiles, int n_tiles, int group_m, int& mt, int& nt) {
  const int group = group_m * n_tiles;
  const int g = t / group, r = t - g * group;
  const int first = g * group_m;
  const int gm = min(group_m, m_tiles - first);
  mt = first + r % gm;
  nt = r / gm;
}


// host helpers defined in gemm_tma.cu
int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols);
int make_tmap_out(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, bool f32);

}  // namespace tfdl
