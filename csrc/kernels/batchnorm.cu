// BatchNorm (training and inference) for NHWC bf16 activations, with the ReLU and the residual add of a ResNet
// bottleneck fused in: y = relu(BN(x) + shortcut). This is tf.keras.layers.BatchNormalization (+ Add + ReLU) of the
// driver's ResNet-50 config (BASELINE.json config 3); the reference repo itself has no normalisation layer.
//
// BatchNorm is a streaming problem: the activations of ResNet-50 at batch 256 are ~5.4 GB per pass, far beyond
// L2, so the cost is the number of passes over HBM. Here:
//   forward : 1 read of x for the statistics, then 1 read of x (+ 1 of the shortcut) and 1 write of y
//   backward: 1 read of dy, x, y for the two channel sums, then 1 read of dy, x, y and 1 write of dx (+ dshortcut)
// with ReLU / ReLU-backward / the residual add costing no extra pass. Each thread owns 8 consecutive channels
// (one 16-byte load) and strides over rows with 4 loads in flight; channel sums go registers -> shared-memory
// atomics -> one global atomic per channel per CTA.
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

constexpr int BN_THREADS = 256;
constexpr int BN_MAX_C = 2048;
constexpr int BN_UNROLL = 4;

__device__ __forceinline__ void bn_unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
__device__ __forceinline__ uint4 bn_pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// sums[0][c] += sum_rows x, sums[1][c] += sum_rows x^2
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums, long long M, int C) {
  __shared__ float sh[2][BN_MAX_C];
  const int tpr = C >> 3, rpb = BN_THREADS / tpr, cg = threadIdx.x % tpr, r0 = threadIdx.x / tpr;
  for (int i = threadIdx.x; i < 2 * C; i += BN_THREADS) (&sh[0][0])[(i / C) * BN_MAX_C + i % C] = 0.f;
  __syncthreads();
  float s[8], q[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s[e] = q[e] = 0.f;
  if (r0 < rpb) {
    const long long stride = (long long)gridDim.x * rpb;
    long long row = (long long)blockIdx.x * rpb + r0;
    for (; row + (BN_UNROLL - 1) * stride < M; row += BN_UNROLL * stride) {
      uint4 v[BN_UNROLL];
#pragma unroll
      for (int u = 0; u < BN_UNROLL; ++u) v[u] = ld_stream(reinterpret_cast<const uint4*>(x + (row + u * stride) * C) + cg);
#pragma unroll
      for (int u = 0; u < BN_UNROLL; ++u) {
        float f[8];
        bn_unpack8(v[u], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) { s[e] += f[e]; q[e] = fmaf(f[e], f[e], q[e]); }
      }
    }
    for (; row < M; row += stride) {
      float f[8];
      bn_unpack8(ld_stream(reinterpret_cast<const uint4*>(x + row * C) + cg), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) { s[e] += f[e]; q[e] = fmaf(f[e], f[e], q[e]); }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      atomicAdd(&sh[0][cg * 8 + e], s[e]);
      atomicAdd(&sh[1][cg * 8 + e], q[e]);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += BN_THREADS) {
    atomicAdd(sums + c, sh[0][c]);
    atomicAdd(sums + C + c, sh[1][c]);
  }
}

// y = [relu]( (x - mean) * rstd * gamma + beta [+ res] ). TRAIN: mean / var from `sums` (and CTA 0 publishes
// mean / rstd and updates the running statistics); otherwise from the running statistics.
template <bool TRAIN>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res, const float* __restrict__ sums,
                const float* __restrict__ in_bias, const float* __restrict__ gamma, const float* __restrict__ beta,
                float* __restrict__ running_mean,
                float* __restrict__ running_var, float* __restrict__ mean_out, float* __restrict__ rstd_out,
                __nv_bfloat16* __restrict__ y, long long M, int C, float eps, float momentum, int relu) {
  __shared__ float sc[BN_MAX_C], sf[BN_MAX_C];
  const float inv_m = 1.f / (float)M;
  for (int c = threadIdx.x; c < C; c += BN_THREADS) {
    float mean, var;
    // in_bias: a per-channel bias the producer (a convolution) left out. Batch statistics absorb it exactly
    // (BN(x + b) == BN(x)), so in training it only shifts the running mean; in inference it shifts the input.
    const float ib = in_bias ? in_bias[c] : 0.f;
    if (TRAIN) {
      mean = sums[c] * inv_m;
      var = fmaxf(sums[C + c] * inv_m - mean * mean, 0.f);
    } else {
      mean = running_mean[c] - ib;
      var = running_var[c];
    }
    const float rstd = rsqrtf(var + eps);
    const float g = gamma[c] * rstd;
    sc[c] = g;
    sf[c] = beta[c] - mean * g;
    if (TRAIN && blockIdx.x == 0) {
      mean_out[c] = mean;
      rstd_out[c] = rstd;
      if (running_mean) {
        const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (mean + ib);
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
      }
    }
  }
  __syncthreads();
  const int tpr = C >> 3;
  const long long nvec = M * tpr;
  // two vectors per thread per trip, all loads issued before the arithmetic; C/8 is a power of two, so the channel
  // group is a mask, not a 64-bit division
  const long long stride = (long long)gridDim.x * BN_THREADS;
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  for (long long i = (long long)blockIdx.x * BN_THREADS + threadIdx.x; i < nvec; i += 2 * stride) {
    const long long i1 = i + stride;
    const bool two = i1 < nvec;
    const uint4 x0 = ld_stream(reinterpret_cast<const uint4*>(x) + i);
    const uint4 x1 = two ? ld_stream(reinterpret_cast<const uint4*>(x) + i1) : zero4;
    const uint4 r0 = res ? ld_stream(reinterpret_cast<const uint4*>(res) + i) : zero4;
    const uint4 r1 = (res && two) ? ld_stream(reinterpret_cast<const uint4*>(res) + i1) : zero4;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !two) break;
      const long long ii = h ? i1 : i;
      const int cg = (int)(ii & (tpr - 1));
      float f[8], r[8];
      bn_unpack8(h ? x1 : x0, f);
      bn_unpack8(h ? r1 : r0, r);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        f[e] = fmaf(f[e], sc[cg * 8 + e], sf[cg * 8 + e]) + r[e];
        if (relu) f[e] = fmaxf(f[e], 0.f);
      }
      reinterpret_cast<uint4*>(y)[ii] = bn_pack8(f);
    }
  }
}

// sums[0][c] += sum dy_eff, sums[1][c] += sum dy_eff * xhat,  dy_eff = relu ? dy * (y > 0) : dy
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ rstd,
                     float* __restrict__ sums, long long M, int C, int relu) {
  __shared__ float sh[2][BN_MAX_C];
  const int tpr = C >> 3, rpb = BN_THREADS / tpr, cg = threadIdx.x % tpr, r0 = threadIdx.x / tpr;
  for (int i = threadIdx.x; i < 2 * C; i += BN_THREADS) (&sh[0][0])[(i / C) * BN_MAX_C + i % C] = 0.f;
  __syncthreads();
  if (r0 < rpb) {
    float mu[8], rs[8], s[8], q[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      mu[e] = mean[cg * 8 + e];
      rs[e] = rstd[cg * 8 + e];
      s[e] = q[e] = 0.f;
    }
    const long long stride = (long long)gridDim.x * rpb;
    for (long long row = (long long)blockIdx.x * rpb + r0; row < M; row += 2 * stride) {
      const bool two = row + stride < M;
      const long long o0 = row * tpr + cg, o1 = (row + stride) * tpr + cg;
      const uint4 d0 = ld_stream(reinterpret_cast<const uint4*>(dy) + o0), x0 = ld_stream(reinterpret_cast<const uint4*>(x) + o0);
      uint4 y0 = make_uint4(0, 0, 0, 0), d1 = y0, x1 = y0, y1 = y0;
      if (relu) y0 = ld_stream(reinterpret_cast<const uint4*>(y) + o0);
      if (two) {
        d1 = ld_stream(reinterpret_cast<const uint4*>(dy) + o1);
        x1 = ld_stream(reinterpret_cast<const uint4*>(x) + o1);
        if (relu) y1 = ld_stream(reinterpret_cast<const uint4*>(y) + o1);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if (h == 1 && !two) break;
        float fd[8], fx[8], fy[8];
        bn_unpack8(h ? d1 : d0, fd);
        bn_unpack8(h ? x1 : x0, fx);
        bn_unpack8(h ? y1 : y0, fy);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float g = (relu && !(fy[e] > 0.f)) ? 0.f : fd[e];
          s[e] += g;
          q[e] = fmaf(g, (fx[e] - mu[e]) * rs[e], q[e]);
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      atomicAdd(&sh[0][cg * 8 + e], s[e]);
      atomicAdd(&sh[1][cg * 8 + e], q[e]);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += BN_THREADS) {
    atomicAdd(sums + c, sh[0][c]);
    atomicAdd(sums + C + c, sh[1][c]);
  }
}

// dx = gamma * rstd * (dy_eff - mean(dy_eff) - xhat * mean(dy_eff * xhat));  dres = dy_eff (optional)
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                    const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ gamma, const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx,
                    __nv_bfloat16* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta, long long M,
                    int C, int relu) {
  __shared__ float a[BN_MAX_C], b[BN_MAX_C], c0[BN_MAX_C];  // dx = a*dy_eff + b*x + c0
  const float inv_m = 1.f / (float)M;
  for (int c = threadIdx.x; c < C; c += BN_THREADS) {
    const float rs = rstd[c], mu = mean[c], g = gamma[c];
    const float m1 = sums[c] * inv_m, m2 = sums[C + c] * inv_m;
    // xhat = (x - mu) * rs  ->  dx = g*rs*dy - g*rs*m1 - g*rs*m2*rs*(x - mu)
    a[c] = g * rs;
    b[c] = -g * rs * rs * m2;
    c0[c] = -g * rs * m1 + g * rs * rs * m2 * mu;
    if (blockIdx.x == 0) {
      dgamma[c] = sums[C + c];
      dbeta[c] = sums[c];
    }
  }
  __syncthreads();
  const int tpr = C >> 3;
  const long long nvec = M * tpr;
  const long long stride = (long long)gridDim.x * BN_THREADS;
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  for (long long i = (long long)blockIdx.x * BN_THREADS + threadIdx.x; i < nvec; i += 2 * stride) {
    const long long i1 = i + stride;
    const bool two = i1 < nvec;
    const uint4 d0 = ld_stream(reinterpret_cast<const uint4*>(dy) + i), x0 = ld_stream(reinterpret_cast<const uint4*>(x) + i);
    const uint4 d1 = two ? ld_stream(reinterpret_cast<const uint4*>(dy) + i1) : zero4;
    const uint4 x1 = two ? ld_stream(reinterpret_cast<const uint4*>(x) + i1) : zero4;
    const uint4 y0 = relu ? ld_stream(reinterpret_cast<const uint4*>(y) + i) : zero4;
    const uint4 y1 = (relu && two) ? ld_stream(reinterpret_cast<const uint4*>(y) + i1) : zero4;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !two) break;
      const long long ii = h ? i1 : i;
      const int cg = (int)(ii & (tpr - 1));
      float fd[8], fx[8], fy[8];
      bn_unpack8(h ? d1 : d0, fd);
      bn_unpack8(h ? x1 : x0, fx);
      bn_unpack8(h ? y1 : y0, fy);
      if (relu) {
#pragma unroll
        for (int e = 0; e < 8; ++e) fd[e] = fy[e] > 0.f ? fd[e] : 0.f;
      }
      if (dres) reinterpret_cast<uint4*>(dres)[ii] = bn_pack8(fd);
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = fmaf(a[cg * 8 + e], fd[e], fmaf(b[cg * 8 + e], fx[e], c0[cg * 8 + e]));
      reinterpret_cast<uint4*>(dx)[ii] = bn_pack8(o);
    }
  }
}

}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_batchnorm_supported(int C) {
  // threads-per-row = C/8 must divide the CTA: C in {8, 16, ..., 2048} with C/8 a power of two
  if (C < 8 || C > BN_MAX_C || C % 8) return 0;
  const int t = C / 8;
  return (t & (t - 1)) == 0;
}

static int bn_grid_rows(long long M, int C) {
  const int rpb = BN_THREADS / (C / 8);
  long long blocks = (M + (long long)rpb * BN_UNROLL - 1) / ((long long)rpb * BN_UNROLL);
  const long long cap = 148 * 6;
  return (int)(blocks < 1 ? 1 : blocks < cap ? blocks : cap);
}
static int bn_grid_vec(long long M, int C) {
  long long blocks = (M * (C / 8) + BN_THREADS * 4 - 1) / (BN_THREADS * 4);
  const long long cap = 148 * 8;
  return (int)(blocks < 1 ? 1 : blocks < cap ? blocks : cap);
}

// sums: 2*C floats of scratch (any contents; cleared here). mean_out / rstd_out [C] are what the backward pass needs.
extern "C" int tfdl_batchnorm_fwd(const void* x, const void* residual, const float* in_bias, const float* gamma,
                                  const float* beta,
                                  float* running_mean, float* running_var, float* sums, float* mean_out, float* rstd_out,
                                  void* y, long long M, int C, float eps, float momentum, int relu, int training,
                                  void* stream) {
  TFDL_REQUIRE(tfdl_batchnorm_supported(C) && M > 0, "batchnorm_fwd: unsupported shape [%lld, %d]", M, C);
  TFDL_REQUIRE((((uintptr_t)x | (uintptr_t)y | (uintptr_t)residual) & 15) == 0, "batchnorm_fwd: activations must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  if (training) {
    TFDL_REQUIRE(sums && mean_out && rstd_out, "batchnorm_fwd: training needs sums / mean_out / rstd_out");
    TFDL_CUDA_OK(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * C, s));
    bn_stats_kernel<<<bn_grid_rows(M, C), BN_THREADS, 0, s>>>((const __nv_bfloat16*)x, sums, M, C);
    bn_apply_kernel<true><<<bn_grid_vec(M, C), BN_THREADS, 0, s>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, sums, in_bias, gamma, beta, running_mean, running_var, mean_out,
        rstd_out, (__nv_bfloat16*)y, M, C, eps, momentum, relu);
  } else {
    TFDL_REQUIRE(running_mean && running_var, "batchnorm_fwd: inference needs the running statistics");
    bn_apply_kernel<false><<<bn_grid_vec(M, C), BN_THREADS, 0, s>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, nullptr, in_bias, gamma, beta, running_mean, running_var, nullptr,
        nullptr, (__nv_bfloat16*)y, M, C, eps, momentum, relu);
  }
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// y is only read when relu != 0 (ReLU mask); dres (optional) receives the gradient of the fused shortcut input.
extern "C" int tfdl_batchnorm_bwd(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                                  const float* gamma, float* sums, void* dx, void* dres, float* dgamma, float* dbeta,
                                  long long M, int C, int relu, void* stream) {
  TFDL_REQUIRE(tfdl_batchnorm_supported(C) && M > 0, "batchnorm_bwd: unsupported shape [%lld, %d]", M, C);
  TFDL_REQUIRE((((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dx | (uintptr_t)dres | (uintptr_t)(relu ? y : nullptr)) & 15) == 0,
               "batchnorm_bwd: activations must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  TFDL_CUDA_OK(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * C, s));
  bn_bwd_reduce_kernel<<<bn_grid_rows(M, C), BN_THREADS, 0, s>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                                                 (const __nv_bfloat16*)y, mean, rstd, sums, M, C, relu);
  bn_bwd_apply_kernel<<<bn_grid_vec(M, C), BN_THREADS, 0, s>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                                               (const __nv_bfloat16*)y, mean, rstd, gamma, sums,
                                                               (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, dgamma, dbeta, M, C,
                                                               relu);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
