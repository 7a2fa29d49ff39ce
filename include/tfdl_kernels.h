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

/* Same with explicit tiling: n_tile (multiple of 16, 0 = auto), k_block (0 = auto), and trans_out = 1 to
 * write element (row, col) to D[col*ldd + row] (coalesced split-K atomics when rows are the fast index). */
int tfdl_gemm_bf16_gather_ex(const void* A, const void* B, float* D, void* Dh, const float* bias, int M, int N,
                             int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int relu, int k_split,
                             int n_tile, int k_block, int trans_out, void* stream);

/* Small-batch Dense forward (ref: tf_dist_example.py:46-47): out[B][N] (bf16) = act(x[B][K] * W[N][K]^T + bias).
 * Split-K tcgen05 GEMM (k_split CTAs along K, 0 = one per 64 columns) accumulating with RED into
 * `scratch` (B*N floats, all zero on entry; handed back all zero), then bias + optional ReLU + cast.
 * B multiple of 16 and <= 256, K multiple of 16. */
int tfdl_dense_fwd_splitk(const void* x_bf16, const void* w_bf16, const float* bias, void* out_bf16, float* scratch,
                          int B, int N, int K, int relu, int k_split, void* stream);

/* The same Dense forward for callers that cannot allocate device memory. (No counterpart in the
 * reference: this entry point exists for embedders of this library, not for parity.)
 *
 * Some hosts have no way to give tfdl_dense_fwd_splitk its scratch buffer: a serving process that handed
 * the whole GPU to its own memory pool at start-up (any later cudaMalloc fails), a latency-critical
 * loop that must not take the device-wide synchronisation cudaMalloc/cudaFree imply, or a bare C /
 * ctypes caller without a tensor library. Here the accumulator is a fixed array inside the library:
 * TFDL_STATIC_SCRATCH_FLOATS floats (256 KiB) per GPU, held by every GPU that loads the library until the
 * process exits. The call allocates nothing - no cudaMalloc/cudaFree, no allocating library routine -
 * and the caller still passes x, w, bias and out.
 *   - Limit: B*N <= TFDL_STATIC_SCRATCH_FLOATS, otherwise the call fails with a message and does nothing.
 *   - The first call on each device looks up the array's device address (with that device current);
 *     make that first call before you begin any stream capture on that device.
 *   - All calls on a device share the one array, which keeps what the previous call left: each call
 *     first clears the B*N floats it uses, but two calls that overlap on the GPU (different streams)
 *     race on it and corrupt each other. Serialise them, or use tfdl_dense_fwd_splitk with
 *     separate buffers. */
#define TFDL_STATIC_SCRATCH_FLOATS 65536
int tfdl_dense_fwd_splitk_static(const void* x_bf16, const void* w_bf16, const float* bias, void* out_bf16, int B,
                                 int N, int K, int relu, int k_split, void* stream);
/* Diagnostic: fill the first n floats of that built-in array with 0 or with a byte-uniform pattern such as the NaN
 * 0xFFFFFFFF (stream-ordered, current device). Tests dirty the array with it before a call. */
int tfdl_static_scratch_fill(float value, int n, void* stream);

/* D[M,N] = act(A[M,K] * B[N,K]^T + bias): persistent warp-specialised TMA + tcgen05 kernel
 * (128B-swizzled tiles, 4-8 stage mbarrier ring, double-buffered TMEM accumulators).
 * act: 0 none, 1 relu, 2 gelu(tanh). K, lda, ldb % 8 == 0; A, B 16-byte aligned. */
int tfdl_gemm_bf16_tma(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M, int N,
                       int K, int lda, int ldb, int ldd, int act, void* stream);
/* Same kernel with either operand MN-major (a_mn: A stored [K][M]; b_mn: B stored [K][N]), loaded as 64-wide
 * 128B-swizzled column blocks: Dense dgrad (dy W) and wgrad (dy^T x) without transposing anything. */
int tfdl_gemm_bf16_tma_ex(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M, int N,
                          int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int act, void* stream);
/* same, with the reduction split `ksplit` ways across CTAs (long-K, few-tile problems such as weight gradients):
 * fp32 output only, no activation, and the caller zero-initialises Df - splits accumulate with TMA reduce-add. */
int tfdl_gemm_bf16_tma_splitk(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M, int N, int K,
                              int lda, int ldb, int ldd, int a_mn, int b_mn, int act, int ksplit, void* stream);

/* ---- LayerNorm (tf.keras.layers.LayerNormalization of the transformer config) -------------------------------------
 * bf16 rows of H = 256, 512, 768 or 1024 elements, fp32 gamma/beta/statistics. forward: z = x (+ residual, optional;
 * z_out receives the bf16 sum), y = LN(z); mean/rstd [T] are saved for the backward pass. backward: dgamma/dbeta are
 * ACCUMULATED (zero them first). */
int tfdl_layernorm_supported(int H);
int tfdl_layernorm_fwd(const void* x, const void* residual, const float* gamma, const float* beta, void* y, void* z_out,
                       float* mean, float* rstd, int T, int H, float eps, void* stream);
int tfdl_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma, void* dz,
                       float* dgamma, float* dbeta, int T, int H, void* stream);
/* out[n] += sum_t x[t][n]: bias gradient of a Dense layer. x bf16 [T, N], row pitch ld; N % 8 == 0, N <= 8192; `out`
 * (fp32 [N]) is accumulated into - zero it first. */
int tfdl_colsum_bf16(const void* x, float* out, long long T, int N, int ld, void* stream);

/* ---- Fused self-attention, S = 128 tokens, head dim 64 (BERT-base at sequence length 128) ----------------------------
 * qkv: bf16 [B*S][3H] = the fused query|key|value projection (H = heads*64); out: bf16 [B*S][H]; lse2: fp32 [B*heads][S]
 * saved for the backward pass. backward: dout bf16 [B*S][H] -> dqkv bf16 [B*S][3H]. No mask, softmax scale 1/8. */
int tfdl_attention_supported(int S, int H, int heads);
int tfdl_attention_fwd(const void* qkv, void* out, float* lse2, int B, int S, int H, int heads, void* stream);
int tfdl_attention_bwd(const void* qkv, const void* dout, const float* lse2, void* dqkv, int B, int S, int H, int heads,
                       void* stream);

/* ---- Sparse softmax cross-entropy on bf16 logits [T][ld] (first V columns valid; ld % 8 == 0, ld <= 100 K) --------------
 * forward: loss fp32 [T] (0 where label == ignore_index) and the softmax probabilities bf16 [T][ld] (pad columns 0);
 * backward: probs -> dlogits = (p - onehot) * gloss[t], in place. labels are int64. */
int tfdl_softmax_xent_supported(int V, int ld);
int tfdl_softmax_xent_fwd(const void* logits, const long long* labels, float* loss, void* probs, long long T, int V, int ld,
                          long long ignore_index, void* stream);
int tfdl_softmax_xent_bwd(void* probs_inout, const long long* labels, const float* gloss, long long T, int V, int ld,
                          long long ignore_index, void* stream);

/* ---- BatchNorm for NHWC bf16 activations [M = N*H*W rows][C channels] (tf.keras.layers.BatchNormalization of the
 * ResNet-50 config), with the bottleneck's ReLU and residual add fused: y = relu(BN(x) + residual).
 * C/8 must be a power of two, C <= 2048. `sums` is 2*C floats of scratch. Training: batch statistics, mean/rstd [C]
 * saved for the backward pass, running statistics updated in place (momentum = weight of the NEW value, running_var
 * gets the unbiased variance); inference: running statistics. backward: y is read only for the ReLU mask.
 * in_bias (optional, [C]): a bias the producing convolution skipped - BN(x + b) == BN(x) under batch statistics, so it
 * only shifts the running mean (training) or the input (inference); its gradient is exactly zero. */
int tfdl_batchnorm_supported(int C);
int tfdl_batchnorm_fwd(const void* x, const void* residual, const float* in_bias, const float* gamma, const float* beta,
                       float* running_mean, float* running_var, float* sums, float* mean_out, float* rstd_out, void* y, long long M, int C,
                       float eps, float momentum, int relu, int training, void* stream);
int tfdl_batchnorm_bwd(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                       const float* gamma, float* sums, void* dx, void* dres, float* dgamma, float* dbeta, long long M,
                       int C, int relu, void* stream);


/* ---- Conv2D as an implicit GEMM on tcgen05 (ref: tf.keras.layers.Conv2D, tf_dist_example.py:41,43) ---------------
 * NHWC bf16 activations, OHWI bf16 weights [K][R][S][C] (the storage layout of a Conv2D kernel variable), fp32
 * accumulation. The activation operand is fetched by TMA in im2col mode (one [pixels x 64 channels] tile per filter
 * tap, stride and zero padding applied by the TMA unit), so there is no im2col buffer. C and K multiples of 8.
 * Output size: P = (H + pad_t + pad_b - R) / stride_h + 1, Q likewise. act: 0 none, 1 relu.
 *   fwd    y[N][P][Q][K] = act(conv(x, w) + bias)
 *   dgrad  dx[N][H][W][C] from dy[N][P][Q][K]             (stride 1)
 *   wgrad  dw[K][R][S][C] fp32 += dy^T im2col(x)          (ACCUMULATES: zero dw first; split over pixels, TMA reduce-add) */
int tfdl_conv2d_supported(int C, int K);
int tfdl_conv2d_fwd(const void* x, const void* w, const float* bias, void* y, int N, int H, int W, int C, int K, int R,
                    int S, int stride_h, int stride_w, int pad_t, int pad_l, int pad_b, int pad_r, int act, void* stream);
int tfdl_conv2d_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R, int S, int pad_t,
                      int pad_l, int pad_b, int pad_r, void* stream);
int tfdl_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int K, int R, int S,
                      int stride_h, int stride_w, int pad_t, int pad_l, int pad_b, int pad_r, void* stream);

/* ---- pooling for NHWC bf16 activations (ref: tf.keras.layers.MaxPooling2D, tf_dist_example.py:42,44; C % 8 == 0) --------
 * maxpool_fwd stores the winning window position r*S+s per output element in idx (uint8 [N][P][Q][C]); maxpool_bwd is a
 * deterministic gather over the windows covering each input element (dx is fully written). Padding positions never win.
 * global_avgpool: y[N][C] = mean over HW. scatter_strided: dst[n][p*sh][q*sw] = src[n][p][q], zero elsewhere - the input
 * gradient of a strided 1x1 convolution from its GEMM result. */
int tfdl_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, int R, int S, int stride_h,
                     int stride_w, int pad_t, int pad_l, int P, int Q, void* stream);
int tfdl_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int N, int H, int W, int C, int R, int S, int stride_h,
                     int stride_w, int pad_t, int pad_l, int P, int Q, void* stream);
int tfdl_global_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int tfdl_global_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
int tfdl_scatter_strided_nhwc(const void* src, void* dst, int N, int H, int W, int C, int P, int Q, int stride_h,
                              int stride_w, void* stream);
/* Patch matrix for a convolution over few input channels (the first layer of an image model; ref: tf_dist_example.py:41,
 * Conv2D over a 1-channel image): cols[N*P*Q][kp] bf16, column (r*S+s)*C+c, zero-padded to kp (multiple of 8). The input is
 * NHWC fp32 (x_is_f32) or bf16. The convolution is then a plain GEMM of cols with the [K][kp] weight matrix. */
int tfdl_im2col_small(const void* x, int x_is_f32, void* cols, int N, int H, int W, int C, int R, int S, int stride_h,
                      int stride_w, int pad_t, int pad_l, int P, int Q, int kp, void* stream);

/* ---- fused gradient all-reduce + SGD over NVLink peer memory ------------------------------
 * ref: README.md:17,21,23,67 (NcclAllReduce / CollectiveOps after every batch) and
 * tf_dist_example.py:51 (SGD). One kernel per step: cross-GPU flag barrier, peer loads of the
 * gradient arenas, 1/N scale, SGD/momentum update of the fp32 master weights, bf16 shadow
 * refresh, clear of the other-parity gradient buffer. See csrc/kernels/allreduce.cu. */
#define TFDL_MAX_RANKS 8
enum { TFDL_AR_ONESHOT = 0, TFDL_AR_TWOSHOT = 1, TFDL_AR_MULTIMEM = 2, TFDL_AR_PUSH = 3, TFDL_AR_PUSH2 = 4 };

typedef struct {
  float lr, momentum, nesterov, scale; /* read on the device each step: LR schedules need no re-capture */
  float beta1, beta2, eps, t_offset;   /* Adam when beta2 > 0 (slots mom = m, mom2 = v); step number = epoch counter + t_offset */
} TfdlHyper;

typedef struct {
  int world, rank;
  int64_t n;                            /* fp32 elements in the arena, multiple of 4 */
  float* params[TFDL_MAX_RANKS];        /* every rank's master-weight arena (peer-mapped); one-shot uses [rank] only */
  void* shadow[TFDL_MAX_RANKS];         /* every rank's bf16 shadow arena (optional, may be NULL) */
  float* grads[2][TFDL_MAX_RANKS];      /* [parity][rank] gradient arenas (peer-mapped) */
  float* mom;                           /* local momentum slot arena or NULL */
  uint32_t* flags[TFDL_MAX_RANKS];      /* every rank's flag pad (peer-mapped), 3 pads of pad_stride u32 */
  uint32_t pad_stride;                  /* >= max_blocks * world */
  int max_blocks;                       /* grid cap (also bounds flag slots) */
  float* mc_params;                     /* multicast (NVLS) addresses, NULL when unsupported */
  void* mc_shadow;
  float* mc_grads[2];
  uint32_t* epoch;                      /* device-local step counter (parity source for the backward kernels) */
  uint32_t* done;                       /* device-local block ticket */
  const TfdlHyper* hyper;               /* device memory */
  const uint32_t* abort_word;           /* host-mapped or device word; non-zero releases every spin */
  int parity;                           /* gradient buffer of this step: 0/1 fixed by the host, -1 = (epoch & 1) */
  void* inbox[TFDL_MAX_RANKS];          /* push / push2 only: every rank's packet inbox (peer-mapped, zeroed once) */
  int64_t inbox_bytes;                  /* size of each inbox, >= tfdl_fused_ar_inbox_bytes(algo, world, n) */
  float* mom2;                          /* second slot arena (Adam's v) or NULL */
} TfdlFusedArParams;

/* TFDL_AR_PUSH / TFDL_AR_PUSH2 move 16-byte {value, epoch, value, epoch} packets into the peers' inboxes instead of
 * raising flags and loading from peers: the receiver polls its own memory and needs no barrier (allreduce.cu).
 * n must be a multiple of 4 except on the one-shot path, which finishes a ragged tail with scalar accesses. */
int64_t tfdl_fused_ar_inbox_bytes(int algo, int world, int64_t n);
int tfdl_fused_allreduce_sgd(const TfdlFusedArParams* p, int algo, void* stream);
/* In-place sum of grads[0][*] over ranks (no update); the bandwidth-sweep primitive. */
int tfdl_allreduce_sum(const TfdlFusedArParams* p, int algo, void* stream);
/* Copy grads[0][root] into every other rank's grads[0][rank] over NVLink (initial variable broadcast from the chief). */
int tfdl_broadcast_f32(const TfdlFusedArParams* p, int root, void* stream);
/* Sum n <= 256 doubles across ranks (metric totals): grads[0][r] = rank r's slot buffer of 2 x 256 doubles, the caller
 * puts its values into slot [epoch & 1] of its own buffer before the call; every rank receives the same totals. */
int tfdl_allreduce_small_f64(const TfdlFusedArParams* p, double* out, int n, void* stream);
/* dst[i] = (double)metrics[i] for i < n <= 256, then metrics[i] = 0 if `zero`: closes an epoch's device-side counters
 * (loss sum, correct, examples) in stream order. dst: device memory or mapped pinned host memory. */
int tfdl_metrics_snapshot(float* metrics, int n, double* dst, int zero, void* stream);
int tfdl_cast_f32_to_bf16(const float* src, void* dst_bf16, int64_t n, void* stream);
int tfdl_enable_peer_access(int device, int peer);


/* ---- reference CNN training step (ref: tf_dist_example.py:40-52) ---------------------------
 * Conv2D(C1,3,relu)-MaxPool-Conv2D(C2,3,relu)-MaxPool-Flatten-Dense(D1,relu)-Dense(NC) with
 * sparse softmax cross-entropy; see csrc/kernels/cnn.cu for the kernel list and fusion plan. */
typedef struct {
  int B;          /* per-replica batch, multiple of 16 */
  int H0, W0;     /* input height/width (single channel) */
  int C1;         /* conv1 filters (32 or 64) */
  int P1h, P1w;   /* pooled conv1 output = (H0-2)/2, (W0-2)/2 */
  int C2;         /* conv2 filters (multiple of 32, <= 128) */
  int P2h, P2w;   /* pooled conv2 output */
  int F;          /* P2h*P2w*C2 (flatten width) */
  int D1;         /* dense units (multiple of 16, <= 256) */
  int NC;         /* classes */
  int nrep;       /* replicas in sync: the loss is mean over the replica batch / nrep */
} TfdlCnnDims;

/* clear2: optional pointer to two floats that the kernel zeroes (the step's loss/accuracy slot). */
int tfdl_cnn_conv1_fwd(const float* x, const float* w /*[C1][9]*/, const float* b, void* a1_bf16, uint8_t* idx1,
                       float* clear2, const TfdlCnnDims* d, void* stream);
int tfdl_cnn_conv2_fwd(const void* a1_bf16, const void* w2_bf16 /*[C2][9*C1]*/, const float* b2, void* a2_bf16,
                       uint8_t* idx2, const TfdlCnnDims* d, void* stream);
/* g_w2 / g_b2 / g_b1, metrics and step_out are ACCUMULATED (the caller clears them between steps). */
int tfdl_cnn_head(float* acc /*[B][D1], re-zeroed*/, const float* b1, const float* w2 /*[NC][D1]*/, const float* b2,
                  const int* labels, void* dh_bf16, float* g_w2, float* g_b2, float* g_b1, float* metrics /*[3]*/,
                  float* step_out /*[2]: mean loss, accuracy*/, const TfdlCnnDims* d, void* stream);
/* dy2 lives on the conv2 INPUT pixel grid, [B][P1h][P1w][C2]; only [0:2P2h, 0:2P2w] is ever written, the
 * rest must stay zero (allocate it zeroed). */
int tfdl_cnn_dense1_dgrad(const void* dh_bf16, const void* w1_bf16 /*[D1][F]*/, const void* a2_bf16,
                          const uint8_t* idx2, void* dy2_bf16 /*[B][P1h][P1w][C2]*/, const TfdlCnnDims* d,
                          void* stream);
int tfdl_cnn_conv2_wgrad(const void* a1_bf16, const void* dy2_bf16, float* g_w2c /*[C2][9*C1], accumulated*/,
                         float* g_b2c /*accumulated*/, const TfdlCnnDims* d, void* stream);
int tfdl_cnn_conv2_dgrad(const void* dy2_bf16, const void* w2_bf16, const void* a1_bf16, void* g1_bf16,
                         const TfdlCnnDims* d, void* stream);
int tfdl_cnn_conv1_wgrad(const float* x, const void* g1_bf16, const uint8_t* idx1, float* g_w1c /*accumulated*/,
                         float* g_b1c /*accumulated*/, const TfdlCnnDims* d, void* stream);

/* TMA-fed, im2col-free versions of the three conv2 kernels (csrc/kernels/cnn_tma.cu): one raw activation tile
 * per image in shared memory, the 3x3 filter as nine MMAs on shifted operand descriptors. Same arguments and
 * results as the gather versions; available when tfdl_cnn_tma_supported(d) (the reference geometry). */
int tfdl_cnn_tma_supported(const TfdlCnnDims* d);
int tfdl_cnn_conv2_fwd_tma(const void* a1_bf16, const void* w2_bf16, const float* b2, void* a2_bf16, uint8_t* idx2,
                           const TfdlCnnDims* d, void* stream);
int tfdl_cnn_conv2_dgrad_tma(const void* dy2_bf16, const void* w2_bf16, const void* a1_bf16, void* g1_bf16,
                             const TfdlCnnDims* d, void* stream);
int tfdl_cnn_conv2_wgrad_tma(const void* a1_bf16, const void* dy2_bf16, float* g_w2c, float* g_b2c,
                             const TfdlCnnDims* d, void* stream);

/* Fused dense block (csrc/kernels/dense_block.cu): Dense(128,relu) + Dense(NC) + softmax-xent forward AND
 * backward down to dy2, plus dW1/db1/dW2/db2, loss and accuracy, in one kernel of P2h*P2w CTAs with one
 * grid barrier. acc: [B][128] fp32, zero on entry; acc_next: the other parity's accumulator (cleared here);
 * sync: one uint32 arrival counter (zero-initialised once, never reset). Gradients are STORED (not
 * accumulated). dh_out optional (tests). Requires tfdl_cnn_dense_block_supported(d). */
int tfdl_cnn_dense_block_supported(const TfdlCnnDims* d);
int tfdl_cnn_dense_block(const void* a2_bf16, const uint8_t* idx2, const void* w1_bf16 /*[128][F]*/,
                         const void* w2_bf16 /*[NC][128]*/, const float* b1, const float* b2, const int* labels,
                         float* acc, float* acc_next, uint32_t* sync, void* dy2_bf16, float* g_w1, float* g_b1,
                         float* g_w2, float* g_b2, float* metrics, float* step_out, void* dh_out_bf16,
                         const TfdlCnnDims* d, void* stream);

/* conv1 forward fused into the TMA conv2 forward (a1/idx1 are still stored for the backward pass), and conv1
 * wgrad fused into the TMA conv2 dgrad (g1 never leaves the SM; g1_out optional, for tests). */
int tfdl_cnn_conv12_fwd_tma(const float* x, const float* w1, const float* b1, const void* w2_bf16, const float* b2,
                            void* a1_bf16, uint8_t* idx1, void* a2_bf16, uint8_t* idx2, float* clear2,
                            const TfdlCnnDims* d, void* stream);
int tfdl_cnn_conv21_bwd_tma(const void* dy2_bf16, const void* w2_bf16, const void* a1_bf16, const uint8_t* idx1,
                            const float* x, float* g_w1c, float* g_b1c, void* g1_out_bf16, const TfdlCnnDims* d,
                            void* stream);

/* One training step = the 9 kernels above + the fused all-reduce/SGD, described by pointers into the
 * caller's buffers. Two descriptors (gradient-buffer parity 0 and 1) make a program. */
typedef struct {
  TfdlCnnDims dims;
  const float* x;            /* [B][H0][W0] fp32 */
  const int* labels;         /* [B] int32 */
  const float* params;       /* fp32 master arena */
  const void* shadow;        /* bf16 shadow arena (same element offsets) */
  float* grads;              /* gradient arena of this parity */
  int64_t off_w1c, off_b1c, off_w2c, off_b2c, off_w1d, off_b1d, off_w2d, off_b2d; /* element offsets */
  void *a1, *a2, *dh, *dy2, *g1; /* bf16 activations / activation gradients */
  uint8_t *idx1, *idx2;          /* pooling argmax */
  float* acc;                /* [B][D1] fp32 split-K scratch, must be zero on entry, zero on exit */
  float* metrics;            /* [3] running loss sum / correct / examples */
  float* step_out;           /* [2] this step's mean loss and accuracy */
  int dense1_ksplit;         /* CTAs along K for dense1 forward */
  int use_tma_conv;          /* 1: conv2 fwd/dgrad/wgrad via cnn_tma.cu (requires tfdl_cnn_tma_supported) */
  int use_dense_block;       /* 1: dense1 fwd + head + dense1 dgrad/wgrad fused (tfdl_cnn_dense_block) */
  float* acc_next;           /* the other parity's split-K accumulator (dense block clears it) */
  uint32_t* sync;            /* grid-barrier arrival counter of the dense block */
  int use_fused_conv;        /* 1: conv1 fwd fused into conv2 fwd, conv1 wgrad fused into conv2 dgrad (needs use_tma_conv) */
} TfdlCnnStep;

/* Launch the forward+backward kernels of one step on `stream`, using `side` for the branches that are
 * off the critical path (dW1, conv2 wgrad); fork/join is done with the two events. Capturable. */
int tfdl_cnn_step_launch(const TfdlCnnStep* s, void* stream, void* side, void* ev_fork, void* ev_join);
/* Same, with bucketed overlap: `early` (may be NULL) is the fused all-reduce+SGD of the dense-layer gradient
 * bucket; it is launched on `ar_stream` as soon as the dense block has produced those gradients and runs
 * concurrently with the conv backward kernels. The caller launches the remaining bucket after this returns. */
int tfdl_cnn_step_launch_overlap(const TfdlCnnStep* s, const TfdlFusedArParams* early, int early_algo, void* stream,
                                 void* side, void* ar_stream, void* ev_fork, void* ev_join, void* ev_ar_fork,
                                 void* ev_ar_join);

/* Native step executor: owns two CUDA graphs (parity 0/1), each = step kernels + fused all-reduce/SGD.
 * tfdl_cnn_program_run drives n_steps steps from host memory without returning to Python: per step an
 * async H2D copy of the batch (x then labels, packed) from pinned memory on a copy stream, one graph
 * launch, and an async D2H copy of (loss, accuracy) into host_out. */
typedef struct TfdlCnnProgram TfdlCnnProgram;
int tfdl_cnn_program_create(const TfdlCnnStep steps[2], const TfdlFusedArParams ar[2], int algo, int use_graph,
                            TfdlCnnProgram** out);
/* Two gradient buckets: `ar_early` (dense layers; reduced while the conv backward still runs) and `ar` (the rest). */
int tfdl_cnn_program_create2(const TfdlCnnStep steps[2], const TfdlFusedArParams ar[2], int algo,
                             const TfdlFusedArParams ar_early[2], int algo_early, int use_graph, TfdlCnnProgram** out);
int tfdl_cnn_program_destroy(TfdlCnnProgram* p);
/* The run entry points launch up to 16 steps per CUDA graph (TFDL_STEPS_PER_GRAPH=1|2|4|8|16): a graph boundary after
 * kernels that wrote peer memory costs ~3-4 us. Each step's batch is fetched by a copy kernel on a parallel branch of
 * the step before it (source pointers come from a device table rewritten before each launch), and the same kernel
 * writes the previous step's (loss, accuracy) into host_out through mapped pinned memory.
 * tfdl_cnn_program_prepare captures and uploads every graph variant up front; nothing is executed. */
int tfdl_cnn_program_prepare(TfdlCnnProgram* p, void* stream);
/* One step on data already resident in steps[parity].x / .labels. */
int tfdl_cnn_program_step(TfdlCnnProgram* p, void* stream);
/* host_batches: n_steps records of (B*H0*W0 floats, B int32) back to back in PINNED memory (record i at
 * byte offset (i % n_records) * record_bytes). host_out: pinned, 2 floats per step. Returns after the
 * last step has been enqueued; the caller synchronises `stream`. */
int tfdl_cnn_program_run(TfdlCnnProgram* p, const void* host_batches, int64_t n_records, int64_t n_steps,
                         float* host_out, void* stream);
/* as tfdl_cnn_program_run, but step i takes its images from xs[i] and its int32 labels from ys[i] (page-locked
 * host memory, e.g. dataset arrays passed through tfdl_host_register): no staging copy on the host */
int tfdl_cnn_program_run_ptrs(TfdlCnnProgram* prog, const void* const* xs, const void* const* ys, int64_t n_steps,
                              float* host_out, void* stream);
int tfdl_host_register(void* ptr, uint64_t bytes);   /* cudaHostRegister; ok if already page-locked */
int tfdl_host_unregister(void* ptr);
/* As tfdl_cnn_program_run with the records already on the device (one D2D copy + one graph launch per step). */
int tfdl_cnn_program_run_device(TfdlCnnProgram* p, const void* dev_batches, int64_t n_records, int64_t n_steps,
                                void* stream);
/* Step timeline for diagnostics: after this call every step launched or captured drops one-thread marker kernels
 * between its stages (0 step start, 1 conv forward done, 2 dense block done, 3 conv2 wgrad done, 4 conv backward done,
 * 5 early-bucket all-reduce done, 6 last all-reduce done). Marker id stores %globaltimer (ns) into
 * buf[id*TFDL_TRACE_CAP + k % TFDL_TRACE_CAP], k = counters[id]++ (device memory, 8 ids). NULL switches it off. */
#define TFDL_TRACE_CAP 256
int tfdl_trace_enable(unsigned long long* buf, unsigned* counters);
int64_t tfdl_cnn_program_steps_done(const TfdlCnnProgram* p);
int64_t tfdl_cnn_program_launches_per_step(const TfdlCnnProgram* p);


/* ---- hardware probes (csrc/kernels/probe.cu): pin down descriptor semantics the conv kernels rely on */
int tfdl_probe_shifted_umma(const void* A_bf16 /*[rows][kc]*/, const void* B_bf16 /*[N][kc]*/, float* D /*[128][N]*/,
                            int rows, int N, int kc /*64: 128B swizzle, 32: 64B swizzle*/, int shift,
                            int base_offset_mode, void* stream);
int tfdl_probe_mn_umma(const void* A_bf16 /*[krows][128|64]*/, const void* B_bf16 /*[krows+64][32]*/, float* D /*[128][32]*/,
                       int krows, int shift, int lbo_mode, void* stream);
int tfdl_probe_tma_coremajor(const void* W_bf16 /*[rows][K]*/, void* out_bf16 /*rows*K*/, int rows, int K, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TFDL_KERNELS_H_ */
