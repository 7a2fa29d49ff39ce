// Sparse softmax cross-entropy over a wide vocabulary (the MLM head of BERT: 8192 tokens x 30522 classes per step),
// forward and backward, on bf16 logits. tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True) of the
// reference (tf_dist_example.py:50) in the shape the driver's transformer config needs.
//
// The eager formulation costs ~9 passes over a 1 GB fp32 logits tensor (cast, log-softmax, gather, their backward).
// Here the logits stay bf16: the forward reads a row once into shared memory, reduces max and sum in the CTA (one exp
// per logit), writes
// the loss and the softmax probabilities (bf16, in the same padded layout as the logits); the backward turns the
// probabilities into dlogits = (p - onehot) * dloss in place. Two read+write passes over 0.5 GB in total.
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

constexpr int XE_THREADS = 256;

__device__ __forceinline__ float xe_block_reduce(float v, bool is_max, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float w = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, w) : v + w;
  }
  __syncthreads();  // red may still be read from the previous reduction
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int i = 1; i < XE_THREADS / 32; ++i) r = is_max ? fmaxf(r, red[i]) : r + red[i];
  return r;
}

// logits: bf16 [T][ld], columns [V, ld) are padding (ignored on input, written as 0 in probs). ld % 8 == 0.
__global__ void __launch_bounds__(XE_THREADS)
softmax_xent_fwd_kernel(const __nv_bfloat16* __restrict__ logits, const long long* __restrict__ labels,
                        float* __restrict__ loss, __nv_bfloat16* __restrict__ probs, int V, int ld, long long ignore) {
  extern __shared__ uint4 row_s[];  // ld / 8 vectors
  __shared__ float red[XE_THREADS / 32];
  const long long t = blockIdx.x;
  const uint4* src = reinterpret_cast<const uint4*>(logits + t * ld);
  const int nvec = ld >> 3;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < nvec; i += XE_THREADS) {
    const uint4 v = __ldg(src + i);
    row_s[i] = v;
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float f = __uint_as_float((e & 1) ? (w[e >> 1] & 0xFFFF0000u) : (w[e >> 1] << 16));
      if (i * 8 + e < V) mx = fmaxf(mx, f);
    }
  }
  mx = xe_block_reduce(mx, true, red);
  const float L2E = 1.4426950408889634f;
  // exp once per logit: the unnormalised exponentials replace the logits in shared memory (bf16), the sum is fp32
  float sum = 0.f;
  for (int i = threadIdx.x; i < nvec; i += XE_THREADS) {
    const uint4 v = row_s[i];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float p[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float f = __uint_as_float((e & 1) ? (w[e >> 1] & 0xFFFF0000u) : (w[e >> 1] << 16));
      p[e] = (i * 8 + e < V) ? exp2f((f - mx) * L2E) : 0.f;
      sum += p[e];
    }
    row_s[i] = make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]), pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
  }
  sum = xe_block_reduce(sum, false, red);
  const float inv = 1.f / sum;
  uint4* dst = reinterpret_cast<uint4*>(probs + t * ld);
  for (int i = threadIdx.x; i < nvec; i += XE_THREADS) {  // each thread re-reads only what it wrote itself
    const uint4 v = row_s[i];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float p[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) p[e] = __uint_as_float((e & 1) ? (w[e >> 1] & 0xFFFF0000u) : (w[e >> 1] << 16)) * inv;
    dst[i] = make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]), pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
  }
  if (threadIdx.x == 0) {
    const long long y = labels[t];
    float l = 0.f;
    if (y != ignore && y >= 0 && y < V) {
      const float xy = __bfloat162float(logits[t * ld + y]);  // shared memory holds exponentials by now
      l = (mx - xy) + logf(sum);
    }
    loss[t] = l;
  }
}

// probs -> dlogits in place: (p - onehot(label)) * gloss[t]; rows with an ignored label become zero.
__global__ void __launch_bounds__(XE_THREADS)
softmax_xent_bwd_kernel(__nv_bfloat16* __restrict__ probs, const long long* __restrict__ labels,
                        const float* __restrict__ gloss, int V, int ld, long long ignore) {
  const long long t = blockIdx.x;
  const long long y = labels[t];
  const bool live = y != ignore && y >= 0 && y < V;
  const float g = live ? gloss[t] : 0.f;
  uint4* row = reinterpret_cast<uint4*>(probs + t * ld);
  const int nvec = ld >> 3;
  for (int i = threadIdx.x; i < nvec; i += XE_THREADS) {
    const uint4 v = row[i];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float p[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float f = __uint_as_float((e & 1) ? (w[e >> 1] & 0xFFFF0000u) : (w[e >> 1] << 16));
      p[e] = (f - ((live && (long long)(i * 8 + e) == y) ? 1.f : 0.f)) * g;
    }
    row[i] = make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]), pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
  }
}

}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_softmax_xent_supported(int V, int ld) { return V > 0 && ld >= V && ld % 8 == 0 && ld * 2 <= 200 * 1024; }

// logits bf16 [T][ld] (first V columns valid), labels int64 [T], loss fp32 [T], probs bf16 [T][ld] (may not alias logits).
extern "C" int tfdl_softmax_xent_fwd(const void* logits, const long long* labels, float* loss, void* probs, long long T,
                                     int V, int ld, long long ignore_index, void* stream) {
  TFDL_REQUIRE(tfdl_softmax_xent_supported(V, ld) && T > 0, "softmax_xent_fwd: unsupported shape [%lld, %d] (pitch %d)", T, V, ld);
  TFDL_REQUIRE((((uintptr_t)logits | (uintptr_t)probs) & 15) == 0, "softmax_xent_fwd: tensors must be 16-byte aligned");
  const size_t smem = (size_t)ld * 2;
  static size_t have = 0;
  if (smem > have) {
    TFDL_CUDA_OK(cudaFuncSetAttribute(softmax_xent_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    have = smem;
  }
  softmax_xent_fwd_kernel<<<(unsigned)T, XE_THREADS, smem, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits, labels, loss,
                                                                                 (__nv_bfloat16*)probs, V, ld, ignore_index);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_softmax_xent_bwd(void* probs_inout, const long long* labels, const float* gloss, long long T, int V,
                                     int ld, long long ignore_index, void* stream) {
  TFDL_REQUIRE(tfdl_softmax_xent_supported(V, ld) && T > 0, "softmax_xent_bwd: unsupported shape");
  softmax_xent_bwd_kernel<<<(unsigned)T, XE_THREADS, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)probs_inout, labels, gloss, V, ld,
                                                                              ignore_index);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
