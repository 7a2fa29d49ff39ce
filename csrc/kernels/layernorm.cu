// LayerNorm forward / backward for transformer blocks (BASELINE.json config 4, BERT-base MLM). The reference repo has
// no transformer; this is the op behind tf.keras.layers.LayerNormalization for the driver's larger config.
//
//   forward : z = x (+ residual);  y = (z - mean) * rstd * gamma + beta     bf16 in / bf16 out, fp32 statistics
//   backward: dz = rstd * (g*dy - mean_h(g*dy) - xhat * mean_h(g*dy*xhat)),  dgamma = sum_t dy*xhat,  dbeta = sum_t dy
//
// Both are pure streaming kernels (one pass over the rows): one warp per row, 16-byte loads, the row lives in
// registers between the statistics and the normalisation, so HBM traffic is the minimum (forward: read x [+ r],
// write y [+ z]; backward: read dy, z, write dz). The column sums for dgamma / dbeta are kept in registers across all
// rows a warp handles, reduced across the CTA's warps in shared memory and added to the fp32 gradients with one
// atomic per column per CTA.
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

constexpr int LN_WARPS = 8;

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
// eight consecutive fp32 parameters; vector loads when the pointer allows (arena views are only 4-byte aligned)
__device__ __forceinline__ void load8(const float* p, float (&o)[8]) {
  if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p + 4));
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = __ldg(p + e);
  }
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// VPL = 16-byte vectors per lane; H = VPL * 256
template <int VPL>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                     const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     __nv_bfloat16* __restrict__ z_out, float* __restrict__ mean_out, float* __restrict__ rstd_out,
                     int T, float eps) {
  constexpr int H = VPL * 256;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int row = blockIdx.x * LN_WARPS + warp; row < T; row += gridDim.x * LN_WARPS) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * H);
    const uint4* rr = res ? reinterpret_cast<const uint4*>(res + (size_t)row * H) : nullptr;
    float v[VPL][8];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      unpack8(xr[j * 32 + lane], v[j]);
      if (rr) {
        float r[8];
        unpack8(rr[j * 32 + lane], r);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[j][e] += r[e];
        if (z_out) {
          // the sum is re-rounded to bf16: that is the tensor the backward pass (and the next residual) sees
          const uint4 zp = pack8(v[j]);
          reinterpret_cast<uint4*>(z_out + (size_t)row * H)[j * 32 + lane] = zp;
          unpack8(zp, v[j]);
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) s += v[j][e];
    }
    const float mean = warp_sum(s) * (1.f / H);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float dlt = v[j][e] - mean;
        q += dlt * dlt;
      }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / H) + eps);
    if (lane == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      const int c0 = (j * 32 + lane) * 8;
      float g[8], b[8];
      load8(gamma + c0, g);
      load8(beta + c0, b);
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (v[j][e] - mean) * rstd * g[e] + b[e];
      reinterpret_cast<uint4*>(y + (size_t)row * H)[j * 32 + lane] = pack8(o);
    }
  }
}

template <int VPL>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z,
                     const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                     __nv_bfloat16* __restrict__ dz, float* __restrict__ dgamma, float* __restrict__ dbeta, int T) {
  constexpr int H = VPL * 256;
  __shared__ float red[LN_WARPS][256];  // one 256-column slab at a time
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float dg[VPL][8], db[VPL][8];
#pragma unroll
  for (int j = 0; j < VPL; ++j)
#pragma unroll
    for (int e = 0; e < 8; ++e) dg[j][e] = db[j][e] = 0.f;
  for (int row = blockIdx.x * LN_WARPS + warp; row < T; row += gridDim.x * LN_WARPS) {
    const uint4* dr = reinterpret_cast<const uint4*>(dy + (size_t)row * H);
    const uint4* zr = reinterpret_cast<const uint4*>(z + (size_t)row * H);
    const float mu = mean[row], rs = rstd[row];
    float gd[VPL][8], xh[VPL][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      float d[8], zz[8];
      unpack8(dr[j * 32 + lane], d);
      unpack8(zr[j * 32 + lane], zz);
      const int c0 = (j * 32 + lane) * 8;
      float g[8];
      load8(gamma + c0, g);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        xh[j][e] = (zz[e] - mu) * rs;
        gd[j][e] = g[e] * d[e];
        s1 += gd[j][e];
        s2 += gd[j][e] * xh[j][e];
        dg[j][e] += d[e] * xh[j][e];
        db[j][e] += d[e];
      }
    }
    const float m1 = warp_sum(s1) * (1.f / H), m2 = warp_sum(s2) * (1.f / H);
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = rs * (gd[j][e] - m1 - xh[j][e] * m2);
      reinterpret_cast<uint4*>(dz + (size_t)row * H)[j * 32 + lane] = pack8(o);
    }
  }
  // column sums: warps -> shared -> one atomic per column per CTA
#pragma unroll
  for (int j = 0; j < VPL; ++j) {
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
      __syncthreads();
#pragma unroll
      for (int e = 0; e < 8; ++e) red[warp][lane * 8 + e] = pass == 0 ? dg[j][e] : db[j][e];
      __syncthreads();
      for (int c = threadIdx.x; c < 256; c += LN_WARPS * 32) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) t += red[w][c];
        atomicAdd((pass == 0 ? dgamma : dbeta) + j * 256 + c, t);
      }
    }
  }
}

}  // namespace tfdl

using namespace tfdl;

#define TFDL_LN_DISPATCH(V, ...)                     \
  switch (V) {                                       \
    case 1: { constexpr int VPL = 1; __VA_ARGS__; break; }  \
    case 2: { constexpr int VPL = 2; __VA_ARGS__; break; }  \
    case 3: { constexpr int VPL = 3; __VA_ARGS__; break; }  \
    case 4: { constexpr int VPL = 4; __VA_ARGS__; break; }  \
    default: TFDL_REQUIRE(false, "layernorm: hidden size %d unsupported", (V) * 256); \
  }

extern "C" int tfdl_layernorm_supported(int H) {
  if (H <= 0 || H % 256) return 0;
  const int v = H / 256;
  return v >= 1 && v <= 4;  // the backward kernel keeps four row-sized register arrays: 1024 columns is its limit
}

static int ln_grid(int T) {
  int blocks = (T + LN_WARPS - 1) / LN_WARPS;
  return blocks < 148 * 4 ? blocks : 148 * 4;
}

extern "C" int tfdl_layernorm_fwd(const void* x, const void* residual, const float* gamma, const float* beta, void* y,
                                  void* z_out, float* mean, float* rstd, int T, int H, float eps, void* stream) {
  TFDL_REQUIRE(tfdl_layernorm_supported(H) && T > 0, "layernorm_fwd: unsupported shape [%d, %d]", T, H);
  TFDL_REQUIRE((((uintptr_t)x | (uintptr_t)y | (uintptr_t)residual | (uintptr_t)z_out) & 15) == 0,
               "layernorm_fwd: pointers must be 16-byte aligned");
  TFDL_LN_DISPATCH(H / 256, (layernorm_fwd_kernel<VPL><<<ln_grid(T), LN_WARPS * 32, 0, (cudaStream_t)stream>>>(
                                (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, gamma, beta, (__nv_bfloat16*)y,
                                (__nv_bfloat16*)z_out, mean, rstd, T, eps)));
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// dgamma / dbeta must be zero on entry (they are accumulated with atomics).
extern "C" int tfdl_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                                  void* dz, float* dgamma, float* dbeta, int T, int H, void* stream) {
  TFDL_REQUIRE(tfdl_layernorm_supported(H) && T > 0, "layernorm_bwd: unsupported shape [%d, %d]", T, H);
  TFDL_REQUIRE((((uintptr_t)dy | (uintptr_t)z | (uintptr_t)dz) & 15) == 0,
               "layernorm_bwd: pointers must be 16-byte aligned");
  int grid = ln_grid(T);
  if (grid > 148) grid = 148;  // fewer CTAs = fewer atomics per column; each still streams many rows
  TFDL_LN_DISPATCH(H / 256, (layernorm_bwd_kernel<VPL><<<grid, LN_WARPS * 32, 0, (cudaStream_t)stream>>>(
                                (const __nv_bfloat16*)dy, (const __nv_bfloat16*)z, mean, rstd, gamma, (__nv_bfloat16*)dz,
                                dgamma, dbeta, T)));
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Column sums of a bf16 matrix [T, N] into fp32 [N] (bias gradients of Dense layers: db = sum over tokens of dy).
// One pass at streaming speed: a CTA owns a slab of rows, each thread 8 consecutive columns (16-byte loads, several
// rows in flight), partial sums stay in registers, one atomic per column per CTA at the end. `out` must be zeroed.
// ---------------------------------------------------------------------------------------------------------
namespace tfdl {
constexpr int CS_THREADS = 256;
constexpr int CS_MAX_GROUPS = 4;  // column groups of 8 per thread -> N <= 8 * 256 * 4 = 8192

constexpr int CS_UNROLL = 8;  // rows in flight per row-lane: a CTA keeps 8-16 KB of loads outstanding

__global__ void __launch_bounds__(CS_THREADS)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, long long T, int N, int ld) {
  extern __shared__ float sh[];  // [N]
  const int groups = N >> 3;
  const int gpt = (groups + CS_THREADS - 1) / CS_THREADS;  // column groups per thread (<= CS_MAX_GROUPS)
  const int tpr = (groups + gpt - 1) / gpt;                // threads per row
  const int rpb = CS_THREADS / tpr;                        // rows per CTA iteration
  const int ci = threadIdx.x % tpr, rl = threadIdx.x / tpr;
  for (int i = threadIdx.x; i < N; i += CS_THREADS) sh[i] = 0.f;
  __syncthreads();
  float acc[CS_MAX_GROUPS][8];
#pragma unroll
  for (int g = 0; g < CS_MAX_GROUPS; ++g)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
  if (rl < rpb) {
    const long long stride = (long long)gridDim.x * rpb;
    for (long long r = (long long)blockIdx.x * rpb + rl; r < T; r += CS_UNROLL * stride) {
      uint4 v[CS_UNROLL][CS_MAX_GROUPS];
#pragma unroll
      for (int u = 0; u < CS_UNROLL; ++u) {
        const long long rr = r + u * stride;
#pragma unroll
        for (int g = 0; g < CS_MAX_GROUPS; ++g) {
          const int cg = ci + g * tpr;
          v[u][g] = (rr < T && g < gpt && cg < groups) ? __ldg(reinterpret_cast<const uint4*>(x + rr * ld) + cg)
                                                       : make_uint4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int u = 0; u < CS_UNROLL; ++u)
#pragma unroll
        for (int g = 0; g < CS_MAX_GROUPS; ++g) {
          float f[8];
          unpack8(v[u][g], f);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[g][e] += f[e];
        }
    }
#pragma unroll
    for (int g = 0; g < CS_MAX_GROUPS; ++g) {
      const int cg = ci + g * tpr;
      if (g < gpt && cg < groups) {
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(&sh[cg * 8 + e], acc[g][e]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < N; i += CS_THREADS) atomicAdd(out + i, sh[i]);
}
}  // namespace tfdl

// out[n] += sum_t x[t][n]; x bf16 [T, N] with row pitch ld (elements, multiple of 8), N % 8 == 0, N <= 8192.
extern "C" int tfdl_colsum_bf16(const void* x, float* out, long long T, int N, int ld, void* stream) {
  TFDL_REQUIRE(T > 0 && N > 0 && N % 8 == 0 && N <= 8 * CS_THREADS * CS_MAX_GROUPS && ld % 8 == 0 && ld >= N,
               "colsum_bf16: unsupported shape [%lld, %d] (pitch %d)", T, N, ld);
  TFDL_REQUIRE(((uintptr_t)x & 15) == 0, "colsum_bf16: x must be 16-byte aligned");
  long long blocks = (T + 15) / 16;  // at least a few rows per row-lane, so the atomics stay a small fraction of the work
  if (blocks > 148 * 4) blocks = 148 * 4;
  colsum_bf16_kernel<<<(unsigned)blocks, CS_THREADS, (size_t)N * sizeof(float), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, out, T, N, ld);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
