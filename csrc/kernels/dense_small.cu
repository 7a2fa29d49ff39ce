// Small-batch Dense forward (tf.keras.layers.Dense, /root/reference/tf_dist_example.py:46-47) as a
// standalone op: out[B][N] = act(x[B][K] * W[N][K]^T + bias[N]).
//
// At training/inference batch sizes like 64 the GEMM has a single 128-row tile, so the K dimension is
// split across CTAs (tcgen05 gather GEMM, swap-AB so the units are the 128 MMA rows) and the partial
// products are added into an fp32 accumulator with RED; a second tiny kernel applies bias/activation,
// writes bf16, and hands the accumulator back zeroed. That accumulator is scratch memory: it is neither
// an input nor an output of the op; the caller passes it (B*N floats, zero on entry, zero on exit).
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

__global__ void dense_finish_kernel(float* __restrict__ acc, const float* __restrict__ bias,
                                    __nv_bfloat16* __restrict__ out, int total, int N, int relu) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  float v = acc[i] + (bias ? bias[i % N] : 0.f);
  acc[i] = 0.f;
  if (relu) v = fmaxf(v, 0.f);
  out[i] = __float2bfloat16(v);
}

static int dense_fwd_impl(const void* x, const void* w, const float* bias, void* out, float* scratch, int B, int N,
                          int K, int relu, int k_split, cudaStream_t st) {
  TFDL_REQUIRE(B > 0 && N > 0 && K > 0, "dense_fwd_splitk: empty problem");
  TFDL_REQUIRE(B % 16 == 0 && B <= 256, "dense_fwd_splitk: batch %d must be a multiple of 16 and <= 256", B);
  TFDL_REQUIRE(K % 16 == 0, "dense_fwd_splitk: K=%d must be a multiple of 16", K);
  const int kb = 64;
  int splits = k_split > 0 ? k_split : (K + kb - 1) / kb;
  // swap-AB: rows = units (W is [N][K], K-major), columns = samples; transposed RED into scratch[b][n]
  int rc = tfdl_gemm_bf16_gather_ex(w, x, scratch, nullptr, nullptr, N, B, K, K, K, N, 0, 0, 0, splits, 0, kb, 1, st);
  if (rc != 0) return rc;
  const int total = B * N;
  dense_finish_kernel<<<(total + 255) / 256, 256, 0, st>>>(scratch, bias, (__nv_bfloat16*)out, total, N, relu);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_dense_fwd_splitk(const void* x, const void* w, const float* bias, void* out, float* scratch, int B,
                                     int N, int K, int relu, int k_split, void* stream) {
  TFDL_REQUIRE(scratch, "dense_fwd_splitk: scratch buffer is required (B*N floats, zero on entry)");
  return dense_fwd_impl(x, w, bias, out, scratch, B, N, K, relu, k_split, (cudaStream_t)stream);
}
