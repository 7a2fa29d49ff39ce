/* C ABI of libtfdl_kernels.so — the sm_100a CUDA kernels of the framework.
 *
 * Conventions
 *   - every entry point returns 0 on success, a cudaError_t (>0) or -1 (argument error) on
 *     failure; tfdl_last_error() then returns a human-readable message (thread-local).
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream).
 *   - pointers are device pointers unless the name says host; bf16 buffers are passed as void*.
 *   - nothing here allocates device memory unless its comment says so; the caller owns all
 *     buffers.
 *
 * The "ref:" notes point at the line of /root/reference that the op stands in for.
 */
#ifndef TFDL_KERNELS_H_
#define TFDL_KERNELS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* tfdl_last_error(void);
int tfdl_kernels_abi_version(void);

/* ---- GEMM (Dense layers; ref: tf_dist_example.py:46-47 tf.keras.layers.Dense) ------------ */

/* D[M,N] (+)= A[M,K] * B[N,K]^T with tcgen05; operands gathered with cp.async into the
 * no-swizzle UMMA layout, any operand may be MN-major (a_mn/b_mn = 1: stored [K][M] / [K][N]).
 * D fp32 and/or Dh bf16 (same ldd). k_split > 1: partial products are atomically added into a
 * caller-zeroed fp32 D (bias/relu/Dh then not allowed). K % 16 == 0. */
int tfdl_gemm_bf16_gather(const void* A, const void* B, float* D, void* Dh, const float* bias, int M, int N,
                          int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int relu, int k_split,
                          void* stream);

/* D[M,N] = act(A[M,K] * B[N,K]^T + bias): persistent warp-specialised TMA + tcgen05 kernel
 * (128B-swizzled tiles, 4-8 stage mbarrier ring, double-buffered TMEM accumulators).
 * act: 0 none, 1 relu, 2 gelu(tanh). K, lda, ldb % 8 == 0; A, B 16-byte aligned. */
int tfdl_gemm_bf16_tma(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M, int N,
                       int K, int lda, int ldb, int ldd, int act, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TFDL_KERNELS_H_ */
