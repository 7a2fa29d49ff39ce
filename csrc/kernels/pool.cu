// Pooling for NHWC bf16 activations (tf.keras.layers.MaxPooling2D of /root/reference/tf_dist_example.py:42,44 for
// general windows / strides / paddings, and the GlobalAveragePooling2D of the ResNet-50 config), plus the strided
// scatter that turns the input gradient of a strided 1x1 convolution back into the full-resolution tensor.
//
// All kernels are streaming: 8 channels (16 bytes) per thread, no shared memory, no atomics. Max-pool forward stores
// the winning window position (r*S + s, one byte per output element); backward is a GATHER - every input element
// adds up the gradients of the (at most ceil(R/stride)*ceil(S/stride)) windows that cover it and elected it - so it is
// deterministic and needs no zero-filled output.
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

struct PoolGeom {
  int N, H, W, C, P, Q, R, S, sh, sw, pt, pl;
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}

__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y,
                                                          uint2* __restrict__ idx, PoolGeom g) {
  const int cg = g.C >> 3;
  const long total = (long)g.N * g.P * g.Q * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    long pix = i / cg;
    const int q = (int)(pix % g.Q);
    pix /= g.Q;
    const int p = (int)(pix % g.P), n = (int)(pix / g.P);
    float best[8];
    unsigned char arg[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      best[k] = -INFINITY;
      arg[k] = 0;
    }
    const int h0 = p * g.sh - g.pt, w0 = q * g.sw - g.pl;
    for (int r = 0; r < g.R; ++r) {
      const int h = h0 + r;
      if (h < 0 || h >= g.H) continue;
      for (int s = 0; s < g.S; ++s) {
        const int w = w0 + s;
        if (w < 0 || w >= g.W) continue;
        float f[8];
        unpack8(__ldg(x + ((long)(n * g.H + h) * g.W + w) * cg + c8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (f[k] > best[k]) {  // first maximum wins (the order TensorFlow's and cuDNN's max-pool gradients use)
            best[k] = f[k];
            arg[k] = (unsigned char)(r * g.S + s);
          }
      }
    }
    y[i] = pack8(best);
    if (idx) {
      uint2 a;
      a.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | ((unsigned)arg[3] << 24);
      a.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | ((unsigned)arg[7] << 24);
      idx[i] = a;
    }
  }
}

__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const uint4* __restrict__ dy, const uint2* __restrict__ idx,
                                                          uint4* __restrict__ dx, PoolGeom g) {
  const int cg = g.C >> 3;
  const long total = (long)g.N * g.H * g.W * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    long pix = i / cg;
    const int w = (int)(pix % g.W);
    pix /= g.W;
    const int h = (int)(pix % g.H), n = (int)(pix / g.H);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // windows p with p*sh - pt <= h <= p*sh - pt + R - 1
    const int hp = h + g.pt, wp = w + g.pl;
    const int p_lo = max(0, (hp - g.R + g.sh) / g.sh), p_hi = min(g.P - 1, hp / g.sh);
    const int q_lo = max(0, (wp - g.S + g.sw) / g.sw), q_hi = min(g.Q - 1, wp / g.sw);
    for (int p = p_lo; p <= p_hi; ++p) {
      const int r = hp - p * g.sh;
      for (int q = q_lo; q <= q_hi; ++q) {
        const int s = wp - q * g.sw;
        const unsigned want = (unsigned)(r * g.S + s);
        const long o = ((long)(n * g.P + p) * g.Q + q) * cg + c8;
        const uint2 a = __ldg(idx + o);
        float f[8];
        unpack8(__ldg(dy + o), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const unsigned ak = ((k < 4 ? a.x : a.y) >> (8 * (k & 3))) & 0xFFu;
          if (ak == want) acc[k] += f[k];
        }
      }
    }
    dx[i] = pack8(acc);
  }
}

// out[n][c] = mean over the H*W positions. One block per (image, slab of channel groups); the block's threads split
// the positions, partial sums meet in shared memory.
__global__ void __launch_bounds__(256) gap_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int HW, int C,
                                                      float inv) {
  extern __shared__ float part[];  // [lanes][groups*8]
  const int cg = C >> 3;
  const int groups = min(cg - (int)blockIdx.y * 256, 256);
  const int lanes = blockDim.x / groups;
  const int grp = threadIdx.x % groups, lane = threadIdx.x / groups;
  const int n = blockIdx.x, c8 = blockIdx.y * 256 + grp;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (lane < lanes) {
    for (int i = lane; i < HW; i += lanes) {
      float f[8];
      unpack8(__ldg(x + ((long)n * HW + i) * cg + c8), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) part[(lane * groups + grp) * 8 + k] = acc[k];
  }
  __syncthreads();
  if (lane == 0) {
    for (int l = 1; l < lanes; ++l)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += part[(l * groups + grp) * 8 + k];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] *= inv;
    y[(long)n * cg + c8] = pack8(acc);
  }
}

__global__ void __launch_bounds__(256) gap_bwd_kernel(const uint4* __restrict__ dy, uint4* __restrict__ dx, int HW, int C,
                                                      float inv, long total) {
  const int cg = C >> 3;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    const long n = i / ((long)HW * cg);
    float f[8];
    unpack8(__ldg(dy + n * cg + c8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] *= inv;
    dx[i] = pack8(f);
  }
}

// dst[N][H][W][C]: dst[n][p*sh][q*sw] = src[n][p][q], zero elsewhere (input gradient of a strided 1x1 convolution).
__global__ void __launch_bounds__(256) scatter_strided_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                                              PoolGeom g) {
  const int cg = g.C >> 3;
  const long total = (long)g.N * g.H * g.W * cg;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cg);
    long pix = i / cg;
    const int w = (int)(pix % g.W);
    pix /= g.W;
    const int h = (int)(pix % g.H), n = (int)(pix / g.H);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (h % g.sh == 0 && w % g.sw == 0 && h / g.sh < g.P && w / g.sw < g.Q)
      v = __ldg(src + ((long)(n * g.P + h / g.sh) * g.Q + w / g.sw) * cg + c8);
    dst[i] = v;
  }
}

// Patch matrix of a convolution whose input has few channels (the 1- or 3-channel first layer of an image model):
// cols[m][k], m = output pixel, k = (r*S + s)*C + c, zero-padded to kp columns. One thread per 16-byte chunk of the
// output; the input window is re-read through L1/L2. IN_F32: the input is still fp32 (cast fused into the gather).
template <bool IN_F32>
__global__ void __launch_bounds__(256) im2col_small_kernel(const void* __restrict__ xin, uint4* __restrict__ cols, PoolGeom g,
                                                           int kp) {
  const int chunks = kp >> 3;
  const long total = (long)g.N * g.P * g.Q * chunks;
  const int K = g.R * g.S * g.C, SC = g.S * g.C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int j = (int)(i % chunks);
    long pix = i / chunks;
    const int q = (int)(pix % g.Q);
    pix /= g.Q;
    const int p = (int)(pix % g.P), n = (int)(pix / g.P);
    const int h0 = p * g.sh - g.pt, w0 = q * g.sw - g.pl;
    float f[8];
    // decode the first column of the chunk once, then walk (c, s, r) incrementally: no division per element
    int k = j * 8;
    int r = k / SC, rem = k - r * SC;
    int sx = rem / g.C, c = rem - sx * g.C;
    const long img = (long)n * g.H;
#pragma unroll
    for (int e = 0; e < 8; ++e, ++k) {
      float v = 0.f;
      if (k < K) {
        const int h = h0 + r, w = w0 + sx;
        if (h >= 0 && h < g.H && w >= 0 && w < g.W) {
          const long o = ((img + h) * g.W + w) * g.C + c;
          v = IN_F32 ? __ldg(static_cast<const float*>(xin) + o)
                     : __bfloat162float(static_cast<const __nv_bfloat16*>(xin)[o]);
        }
      }
      f[e] = v;
      if (++c == g.C) {
        c = 0;
        if (++sx == g.S) {
          sx = 0;
          ++r;
        }
      }
    }
    cols[i] = pack8(f);
  }
}

static int grid_for(long total) {
  long b = (total + 255) / 256;
  return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, int R, int S, int stride_h,
                                int stride_w, int pad_t, int pad_l, int P, int Q, void* stream) {
  TFDL_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && R > 0 && S > 0 && R * S <= 255 && P > 0 && Q > 0,
               "maxpool_fwd: bad shape (C must be a multiple of 8, window at most 255 positions)");
  TFDL_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0 && ((uintptr_t)idx & 7) == 0, "maxpool_fwd: alignment");
  PoolGeom g{N, H, W, C, P, Q, R, S, stride_h, stride_w, pad_t, pad_l};
  const long total = (long)N * P * Q * (C / 8);
  maxpool_fwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)x, (uint4*)y, (uint2*)idx, g);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int N, int H, int W, int C, int R, int S,
                                int stride_h, int stride_w, int pad_t, int pad_l, int P, int Q, void* stream) {
  TFDL_REQUIRE(N > 0 && C % 8 == 0 && idx && dy && dx, "maxpool_bwd: bad arguments");
  PoolGeom g{N, H, W, C, P, Q, R, S, stride_h, stride_w, pad_t, pad_l};
  const long total = (long)N * H * W * (C / 8);
  maxpool_bwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)dy, (const uint2*)idx, (uint4*)dx, g);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_global_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
  TFDL_REQUIRE(N > 0 && HW > 0 && C > 0 && C % 8 == 0, "global_avgpool_fwd: C must be a multiple of 8");
  const int cg = C / 8;
  const int gy = (cg + 255) / 256;
  TFDL_REQUIRE(gy == 1 || cg % 256 == 0, "global_avgpool_fwd: C above 2048 must be a multiple of 2048");
  const int groups = cg < 256 ? cg : 256;
  const int lanes = 256 / groups;
  const size_t smem = (size_t)lanes * groups * 8 * sizeof(float);
  gap_fwd_kernel<<<dim3(N, gy), lanes * groups, smem, (cudaStream_t)stream>>>((const uint4*)x, (uint4*)y, HW, C, 1.f / HW);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_global_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
  TFDL_REQUIRE(N > 0 && HW > 0 && C > 0 && C % 8 == 0, "global_avgpool_bwd: C must be a multiple of 8");
  const long total = (long)N * HW * (C / 8);
  gap_bwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)dy, (uint4*)dx, HW, C, 1.f / HW, total);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_scatter_strided_nhwc(const void* src, void* dst, int N, int H, int W, int C, int P, int Q, int stride_h,
                                         int stride_w, void* stream) {
  TFDL_REQUIRE(N > 0 && C % 8 == 0 && stride_h > 0 && stride_w > 0, "scatter_strided_nhwc: bad arguments");
  PoolGeom g{N, H, W, C, P, Q, 1, 1, stride_h, stride_w, 0, 0};
  const long total = (long)N * H * W * (C / 8);
  scatter_strided_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)src, (uint4*)dst, g);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_im2col_small(const void* x, int x_is_f32, void* cols, int N, int H, int W, int C, int R, int S,
                                 int stride_h, int stride_w, int pad_t, int pad_l, int P, int Q, int kp, void* stream) {
  TFDL_REQUIRE(N > 0 && C > 0 && R > 0 && S > 0 && P > 0 && Q > 0 && kp % 8 == 0 && kp >= R * S * C,
               "im2col_small: kp must be a multiple of 8 and at least R*S*C");
  TFDL_REQUIRE(((uintptr_t)cols & 15) == 0, "im2col_small: cols must be 16-byte aligned");
  PoolGeom g{N, H, W, C, P, Q, R, S, stride_h, stride_w, pad_t, pad_l};
  const long total = (long)N * P * Q * (kp / 8);
  if (x_is_f32) im2col_small_kernel<true><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(x, (uint4*)cols, g, kp);
  else im2col_small_kernel<false><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(x, (uint4*)cols, g, kp);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
