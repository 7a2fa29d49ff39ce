// Small-batch Dense forward (tf.keras.layers.Dense, /root/reference/tf_dist_example.py:46-47) as a
// standalone op: out[B][N] = act(x[B][K] * W[N][K]^T + bias[N]).
//
// At training/inference batch sizes like 64 the GEMM has a single 128-row tile, so the K dimension is
// split across CTAs (tcgen05 gather GEMM, swap-AB so the units are the 128 MMA rows) and the partial
// products are added into an fp32 accumulator with RED; a second tiny kernel applies bias/activation,
// writes bf16, and hands the accumulator back zeroed. That accumulator is scratch memory: it is neither
// an input nor an output of the op. tfdl_dense_fwd_splitk takes it from the caller (B*N floats, zero on
// entry, zero on exit); tfdl_dense_fwd_splitk_static keeps it in a fixed array inside this library.
#include <atomic>
#include <string.h>
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

// ---------------------------------------------------------------------------------------------
// Library-owned scratch for callers that cannot allocate.
//
// Who that is: a process that hands all GPU memory to its own pool at start-up (a later cudaMalloc
// simply fails there), a latency-critical loop that cannot tolerate the device-wide synchronisation a
// cudaMalloc/cudaFree implies, or a thin C / ctypes caller with no tensor library at hand to make a
// workspace. For them the accumulator lives in the array below: 256 KiB that every GPU which loads
// this library keeps until the process exits.
//
// The name `g_static_scratch` is only meaningful in device code. On the host it is a stand-in
// object, not a device address, so host code asks the runtime for the real address (once per device,
// with that device current) and passes that pointer to the kernels.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_static_scratch[TFDL_STATIC_SCRATCH_FLOATS];

static float* static_scratch_for_current_device(int* rc_out) {
  constexpr int kMaxDevices = 64;
  // one slot per device, filled on first use; two threads racing on the same device store the same address
  static std::atomic<float*> ptrs[kMaxDevices];
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_last_error("dense_fwd_splitk_static: cannot identify the current device (%s)",
                   e == cudaSuccess ? "index out of range" : cudaGetErrorString(e));
    *rc_out = e == cudaSuccess ? -1 : (int)e;
    return nullptr;
  }
  float* cached = ptrs[dev].load(std::memory_order_acquire);
  if (!cached) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      set_last_error("dense_fwd_splitk_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                     cudaGetErrorString(e));
      *rc_out = e == cudaSuccess ? -1 : (int)e;
      return nullptr;
    }
    ptrs[dev].store(p, std::memory_order_release);
    cached = p;
  }
  *rc_out = 0;
  return cached;
}

}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_dense_fwd_splitk(const void* x, const void* w, const float* bias, void* out, float* scratch, int B,
                                     int N, int K, int relu, int k_split, void* stream) {
  TFDL_REQUIRE(scratch, "dense_fwd_splitk: scratch buffer is required (B*N floats, zero on entry)");
  return dense_fwd_impl(x, w, bias, out, scratch, B, N, K, relu, k_split, (cudaStream_t)stream);
}

extern "C" int tfdl_dense_fwd_splitk_static(const void* x, const void* w, const float* bias, void* out, int B, int N,
                                            int K, int relu, int k_split, void* stream) {
  TFDL_REQUIRE(B > 0 && N > 0, "dense_fwd_splitk_static: empty problem");
  TFDL_REQUIRE((long long)B * N <= (long long)TFDL_STATIC_SCRATCH_FLOATS,
               "dense_fwd_splitk_static: B*N = %lld exceeds the built-in scratch of %d floats; use "
               "tfdl_dense_fwd_splitk with your own buffer",
               (long long)B * N, (int)TFDL_STATIC_SCRATCH_FLOATS);
  int rc = 0;
  float* scratch = static_scratch_for_current_device(&rc);
  if (!scratch) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // The array is shared by every call and keeps whatever the last one left (an aborted call may have
  // left partial sums), so the part used here is cleared first.
  TFDL_CUDA_OK(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)B * N, st));
  return dense_fwd_impl(x, w, bias, out, scratch, B, N, K, relu, k_split, st);
}

// Diagnostic hook: fills the first n floats of the built-in array with `value` on the current device (stream-ordered).
// Tests use it to leave the array dirty - NaNs - before a call, to prove the entry point clears what it uses.
extern "C" int tfdl_static_scratch_fill(float value, int n, void* stream) {
  TFDL_REQUIRE(n >= 0 && n <= TFDL_STATIC_SCRATCH_FLOATS, "static_scratch_fill: n out of range");
  int rc = 0;
  float* scratch = static_scratch_for_current_device(&rc);
  if (!scratch) return rc;
  if (n == 0) return 0;
  if (value == 0.f) {
    TFDL_CUDA_OK(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)n, (cudaStream_t)stream));
    return 0;
  }
  unsigned bits;
  memcpy(&bits, &value, 4);
  // every float gets the same 32-bit pattern; a byte memset only works for patterns of four equal bytes (0xFF.. = NaN)
  TFDL_REQUIRE(((bits >> 24) & 0xFF) == (bits & 0xFF) && ((bits >> 16) & 0xFF) == (bits & 0xFF) &&
               ((bits >> 8) & 0xFF) == (bits & 0xFF), "static_scratch_fill: only 0 and byte-uniform patterns (e.g. NaN 0xFFFFFFFF)");
  TFDL_CUDA_OK(cudaMemsetAsync(scratch, (int)(bits & 0xFF), sizeof(float) * (size_t)n, (cudaStream_t)stream));
  return 0;
}
