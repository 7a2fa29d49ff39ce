// Training-step kernels of the reference CNN (/root/reference/tf_dist_example.py:40-52):
//   Conv2D(32,3,relu) -> MaxPool -> Conv2D(64,3,relu) -> MaxPool -> Flatten -> Dense(128,relu) -> Dense(10)
//   -> SparseCategoricalCrossentropy(from_logits) -> SGD, forward and backward.
//
// TensorFlow runs this as ~40 library/elementwise launches per step (cuDNN conv + BiasAdd + Relu +
// MaxPool + ... and their gradients). Here the step is 9 launches: the GEMM-shaped layers run on
// tcgen05 with fp32 accumulation in TMEM and everything elementwise (bias, ReLU, 2x2 max-pool with
// arg-max, ReLU/pool gradients, bias gradients, loss, metrics) is fused into a prologue/epilogue.
//
//   conv1_fwd          SIMT   x[B,28,28] -> a1 = pool(relu(conv+b)) [B,13,13,32] bf16 + argmax      (K=9: bandwidth bound)
//   conv2_fwd          UMMA   implicit GEMM M=B*5*5*4 (pool windows contiguous), N=64, K=288;
//                             epilogue: +bias, 2x2 max/argmax by warp shuffles, ReLU -> a2 [B,1600] bf16
//   dense1_fwd         UMMA   gather GEMM (gemm_gather.cu) split-K into a zeroed fp32 scratch `acc`
//   head               SIMT   h = relu(acc+b); logits; softmax-xent; metrics; dW2, db2, db1; dh (ReLU-masked); re-zero acc
//   dense1_wgrad       UMMA   gather GEMM, both operands MN-major: dW1 = dh^T a2
//   dense1_dgrad       UMMA   da2 = dh W1 ; epilogue: ReLU mask + un-pool through argmax -> dy2 [B,10,10,64] bf16
//   conv2_wgrad        UMMA   dW2 += im2col(a1)^T dy2, reduction over pixels split across CTAs (fp32 RED), + db2
//   conv2_dgrad        UMMA   da1 = dy2 (*) rot(W2): implicit GEMM M=B*13*13, N=32, K=576; epilogue ReLU mask -> g1
//   conv1_wgrad        SIMT   dW1, db1 of the first conv through the argmax of pool 1
//
// Layouts: activations NHWC; conv kernels OHWI, dense kernels [out][in] (the arena's storage layout,
// see parallel/values.py), so every forward operand is K-major and every backward operand is either
// K-major or MN-major - both are legal tcgen05 smem-descriptor majors, so nothing is ever transposed.
// Gathered operands are written to shared memory in the no-swizzle core-matrix layout (common.cuh).
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

__device__ __forceinline__ float bf16_bits_to_float(uint16_t b) { return __uint_as_float((uint32_t)b << 16); }

// Optional in-kernel phase timer (build with TFDL_NVCC_EXTRA=-DTFDL_PROFILE_PHASES): thread 0 of every
// CTA stamps clock64() at named points into g_phase[cta][slot]; tools/phase_times.py reads it back.
#ifdef TFDL_PROFILE_PHASES
__device__ long long g_phase[256][8];
#define PHASE(slot) do { if (threadIdx.x == 0 && blockIdx.x < 256 && blockIdx.y == 0) g_phase[blockIdx.x][slot] = clock64(); } while (0)
#else
#define PHASE(slot) do { } while (0)
#endif

// Issue all K-steps of one accumulator tile. A/B staged in the no-swizzle layout with `kchunks`
// 16-byte chunks along K (SBO = kchunks*128, LBO = 128; one K-step of 16 = +256 bytes).
__device__ __forceinline__ void issue_tile(uint32_t tmem_d, uint32_t a_smem, uint32_t b_smem, int kchunks,
                                           uint32_t idesc) {
  const uint32_t sbo = (uint32_t)kchunks * 128u;
  for (int ks = 0; ks < (kchunks >> 1); ++ks) {
    uint64_t ad = umma_smem_desc(a_smem + ks * 256u, 128u, sbo, SWZ_NONE);
    uint64_t bd = umma_smem_desc(b_smem + ks * 256u, 128u, sbo, SWZ_NONE);
    umma_bf16(tmem_d, ad, bd, idesc, ks > 0);
  }
}

struct TileCtx {
  uint32_t tmem;
  int warp;
};

// Common prologue: TMEM alloc (warp 0), mbarrier init, block sync. TC_THREADS threads: warps w and w+4
// share TMEM lane quarter (w & 3) and split the accumulator columns between them in the epilogue.
constexpr int TC_THREADS = 256;
__device__ __forceinline__ TileCtx tile_begin(uint64_t* bar, uint32_t* slot, uint32_t ncols) {
  TileCtx c;
  c.warp = threadIdx.x >> 5;
  if (c.warp == 0) tmem_alloc(slot, ncols);
  if (threadIdx.x == 32) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  c.tmem = *slot;
  return c;
}
__device__ __forceinline__ void tile_end(const TileCtx& c, uint32_t ncols) {
  tc_fence_before_sync();
  __syncthreads();
  if (c.warp == 0) tmem_dealloc(c.tmem, ncols);
}
// All staging cp.asyncs of this thread are complete and visible to the tensor core.
__device__ __forceinline__ void staging_done() {
  cp_async_wait_all();
  fence_proxy_async_smem();
  __syncthreads();
}

// =============================================================================================
// conv1 forward: direct 3x3 conv (Cin = 1) + bias + 2x2 max-pool (argmax) + ReLU
// grid = B * P1h blocks (one pooled row each), block = P1w * C1 threads
// =============================================================================================
__global__ void conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                 __nv_bfloat16* __restrict__ a1, uint8_t* __restrict__ idx1, float* __restrict__ clear2,
                                 TfdlCnnDims d) {
  extern __shared__ float rows[];  // 4 input rows
  if (clear2 && blockIdx.x == 0 && threadIdx.x < 2) clear2[threadIdx.x] = 0.f;  // this step's (loss, accuracy) slot
  const int n = blockIdx.x / d.P1h, ph = blockIdx.x % d.P1h;
  const float* src = x + ((size_t)n * d.H0 + 2 * ph) * d.W0;
  for (int i = threadIdx.x; i < 4 * d.W0; i += blockDim.x) rows[i] = src[i];
  const int pw = threadIdx.x / d.C1, c = threadIdx.x % d.C1;
  float wr[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wr[t] = w[c * 9 + t];
  const float bias = b[c];
  __syncthreads();
  if (pw >= d.P1w) return;
  float best = -INFINITY;
  int arg = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int dy = q >> 1, dx = q & 1;
    float acc = bias;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) acc = fmaf(rows[(dy + kh) * d.W0 + 2 * pw + dx + kw], wr[kh * 3 + kw], acc);
    if (acc > best) { best = acc; arg = q; }
  }
  const size_t o = (((size_t)n * d.P1h + ph) * d.P1w + pw) * d.C1 + c;
  a1[o] = __float2bfloat16(fmaxf(best, 0.f));
  idx1[o] = (uint8_t)arg;
}

// =============================================================================================
// conv2 forward: implicit GEMM on tcgen05 with fused bias + 2x2 max-pool (argmax) + ReLU epilogue.
//   D[cout][pixel] = sum_k W[cout][k] * im2col(a1)[pixel][k]        (weights are the M operand)
// The 128 pixels of a tile are ordered (n, ph, pw, dy, dx), so the four pixels of a pooling window
// are four CONSECUTIVE accumulator columns of one TMEM lane: the pool is a register-local max, the
// bias is one scalar per thread, and a warp stores 32 consecutive channels of a window (64 B).
// The C2 filter rows are replicated to fill the 128 MMA rows so that all four lane quarters (and
// both warps of each quarter) own a distinct column range.
// grid = ceil(B*P2h*P2w*4 / 128), block = 256
// =============================================================================================
template <int C1, int C2>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv2_fwd_kernel(const __nv_bfloat16* __restrict__ a1, const __nv_bfloat16* __restrict__ w2h,
                 const float* __restrict__ b2, __nv_bfloat16* __restrict__ a2, uint8_t* __restrict__ idx2,
                 TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  constexpr int K = 9 * C1, KCH = K / 8, CPT = C1 / 8, REP = 128 / C2, NCOL = 128 / REP / 2;
  static_assert(128 % C2 == 0 && NCOL % 4 == 0 && CPT % 2 == 0, "unsupported channel counts");
  const uint32_t w_smem = smem_u32(smem), p_smem = w_smem + 128u * K * 2u;
  PHASE(0);
  TileCtx tc = tile_begin(&bar, &slot, 128);
  PHASE(1);
  const int Mtot = d.B * d.P2h * d.P2w * 4;
  // pixels (N operand). A warp-wide cp.async must touch few 128-byte lines: lanes = 8 consecutive tile
  // rows x 4 consecutive 16-byte channel chunks. Rows are ordered (.., pw, dy, dx), so 8 rows are two
  // pooling windows = two runs of 4 adjacent pixels in a1 (NHWC), i.e. 4 lines instead of 32.
  __shared__ int rowoff[128];
  if (threadIdx.x < 128) {
    const int m = blockIdx.x * 128 + threadIdx.x;
    const int q = m & 3, win = m >> 2;
    const int pw = win % d.P2w, t2 = win / d.P2w, ph = t2 % d.P2h, n = t2 / d.P2h;
    rowoff[threadIdx.x] = m < Mtot ? ((n * d.P1h + 2 * ph + (q >> 1)) * d.P1w + 2 * pw + (q & 1)) * C1 : -1;
  }
  __syncthreads();
  {
    const int r_in = threadIdx.x & 7;
#pragma unroll 3
    for (int i = threadIdx.x >> 3; i < 16 * 9 * CPT; i += TC_THREADS / 8) {
      const int c8 = i % CPT, rest = i / CPT, rg = rest & 15, tap = rest >> 4;
      const int r = rg * 8 + r_in, off = rowoff[r];
      const bool ok = off >= 0;
      const __nv_bfloat16* src = a1 + (ok ? off + ((tap / 3) * d.P1w + tap % 3) * C1 + c8 * 8 : 0);
      cp_async16(p_smem + core_off_kmajor(r, tap * CPT + c8, KCH), src, ok);
    }
    // weights (M operand): 128 rows = the C2 filters repeated REP times
#pragma unroll 2
    for (int i = threadIdx.x; i < 128 * KCH; i += TC_THREADS) {
      const int r_in2 = i & 7, rest = i >> 3, j = rest % KCH, row = (rest / KCH) * 8 + r_in2;
      cp_async16(w_smem + core_off_kmajor(row, j, KCH), w2h + (size_t)(row % C2) * K + j * 8, true);
    }
  }
  PHASE(2);
  staging_done();
  PHASE(3);
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    issue_tile(tc.tmem, w_smem, p_smem, KCH, umma_idesc_bf16(128, 128, false, false));
    umma_commit(&bar);
  }
  PHASE(4);
  const int lq = tc.warp & 3, row = lq * 32 + (int)lane_id(), cout = row % C2;
  const float bias = __ldg(b2 + cout);
  const int col0 = (row / C2) * (128 / REP) + (tc.warp >> 2) * NCOL;
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
  PHASE(5);
  const uint32_t t_row = tc.tmem + ((uint32_t)(lq * 32) << 16);
#pragma unroll 1
  for (int c = 0; c < NCOL; c += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + col0 + c, v);
    tmem_ld_wait();
    constexpr int WPL = NCOL >= 32 ? 8 : NCOL / 4;  // pooling windows in this load that belong to this thread
#pragma unroll
    for (int wdw = 0; wdw < WPL; ++wdw) {
      const int m = blockIdx.x * 128 + col0 + c + wdw * 4;
      float best = __uint_as_float(v[wdw * 4]);
      int arg = 0;
#pragma unroll
      for (int q = 1; q < 4; ++q) {
        const float x = __uint_as_float(v[wdw * 4 + q]);
        if (x > best) { best = x; arg = q; }
      }
      if (m < Mtot) {
        const size_t o = (size_t)(m >> 2) * C2 + cout;
        a2[o] = __float2bfloat16(fmaxf(best + bias, 0.f));
        idx2[o] = (uint8_t)arg;
      }
    }
  }
  PHASE(6);
  tile_end(tc, 128);
  PHASE(7);
}

// =============================================================================================
// head: Dense(D1->NC) + softmax cross-entropy + metrics + all tiny gradients.
// grid = B / 8 blocks, 256 threads: warp w of a block owns sample 8*blockIdx.x + w.
//   acc  [B][D1] fp32 : dense1 pre-activation WITHOUT bias (split-K scratch), re-zeroed here
//   dh   [B][D1] bf16 : dL/d(dense1 pre-activation) (ReLU-masked)
//   g_w2 / g_b2 / g_b1 are ACCUMULATED (fp32 RED): the arena is cleared by the all-reduce kernel
//   metrics[0..2] += (sum of per-example loss, correct predictions, examples)
//   step_out[0..1] += (mean loss, accuracy) of this step; the step's first kernel clears it
// =============================================================================================
constexpr int HEAD_S = 8;
__global__ void __launch_bounds__(HEAD_S * 32, 1)
head_kernel(float* __restrict__ acc, const float* __restrict__ b1, const float* __restrict__ w2 /*[NC][D1]*/,
            const float* __restrict__ b2, const int* __restrict__ labels, __nv_bfloat16* __restrict__ dh,
            float* __restrict__ g_w2, float* __restrict__ g_b2, float* __restrict__ g_b1,
            float* __restrict__ metrics, float* __restrict__ step_out, TfdlCnnDims d) {
  extern __shared__ float sm[];
  const int D1 = d.D1, NC = d.NC, t = threadIdx.x, w = t >> 5, lane = t & 31;
  float* h = sm;                     // [HEAD_S][D1]   relu(acc + b1)
  float* dhs = h + HEAD_S * D1;      // [HEAD_S][D1]   dL/dpre
  float* w2s = dhs + HEAD_S * D1;    // [NC][D1]
  float* dl = w2s + NC * D1;         // [HEAD_S][NC]   logits -> dlogits
  float* red = dl + HEAD_S * NC;     // [2] block loss, block correct
  const int s = blockIdx.x * HEAD_S + w;
  // phase 1: all loads issued before any use
  for (int i = t; i < NC * D1; i += HEAD_S * 32) w2s[i] = w2[i];
  for (int k = lane; k < D1; k += 32) {
    const float v = acc[(size_t)s * D1 + k] + b1[k];
    acc[(size_t)s * D1 + k] = 0.f;
    h[w * D1 + k] = fmaxf(v, 0.f);
  }
  if (t < 2) red[t] = 0.f;
  __syncthreads();
  // phase 2: logits, one warp per sample, lanes split D1
  for (int j = 0; j < NC; ++j) {
    float a = 0.f;
    for (int k = lane; k < D1; k += 32) a = fmaf(h[w * D1 + k], w2s[j * D1 + k], a);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (lane == 0) dl[w * NC + j] = a + b2[j];
  }
  __syncwarp();
  // phase 3: softmax cross-entropy (NC <= 64: two classes per lane at most)
  {
    const int y = labels[s];
    const float z0 = lane < NC ? dl[w * NC + lane] : -INFINITY;
    const float z1 = lane + 32 < NC ? dl[w * NC + lane + 32] : -INFINITY;
    float mx = fmaxf(z0, z1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float e0 = lane < NC ? __expf(z0 - mx) : 0.f, e1 = lane + 32 < NC ? __expf(z1 - mx) : 0.f;
    float sum = e0 + e1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    // first index attaining the maximum (argmax for the accuracy metric)
    int am = z0 == mx ? lane : (z1 == mx ? lane + 32 : 1 << 20);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) am = min(am, __shfl_xor_sync(0xffffffffu, am, o));
    const float inv = 1.f / ((float)d.B * (float)d.nrep);
    const float zy = dl[w * NC + y];
    __syncwarp();
    if (lane < NC) dl[w * NC + lane] = (e0 / sum - (lane == y ? 1.f : 0.f)) * inv;
    if (lane + 32 < NC) dl[w * NC + lane + 32] = (e1 / sum - (lane + 32 == y ? 1.f : 0.f)) * inv;
    if (lane == 0) {
      atomicAdd(&red[0], __logf(sum) + mx - zy);
      atomicAdd(&red[1], am == y ? 1.f : 0.f);
    }
  }
  __syncwarp();
  // phase 4: dh[s][k] = relu'(h) * sum_j dl[s][j] w2[j][k]
  for (int k = lane; k < D1; k += 32) {
    float a = 0.f;
    for (int j = 0; j < NC; ++j) a = fmaf(dl[w * NC + j], w2s[j * D1 + k], a);
    a = h[w * D1 + k] > 0.f ? a : 0.f;
    dhs[w * D1 + k] = a;
    dh[(size_t)s * D1 + k] = __float2bfloat16(a);
  }
  __syncthreads();
  // phase 5: this block's contribution to dW2, db2, db1
  for (int i = t; i < NC * D1; i += HEAD_S * 32) {
    const int j = i / D1, k = i - j * D1;
    float a = 0.f;
#pragma unroll
    for (int q = 0; q < HEAD_S; ++q) a = fmaf(dl[q * NC + j], h[q * D1 + k], a);
    atomicAdd(g_w2 + i, a);
  }
  for (int k = t; k < D1; k += HEAD_S * 32) {
    float a = 0.f;
#pragma unroll
    for (int q = 0; q < HEAD_S; ++q) a += dhs[q * D1 + k];
    atomicAdd(g_b1 + k, a);
  }
  if (t < NC) {
    float a = 0.f;
#pragma unroll
    for (int q = 0; q < HEAD_S; ++q) a += dl[q * NC + t];
    atomicAdd(g_b2 + t, a);
  }
  if (t == 0) {
    atomicAdd(metrics + 0, red[0]);
    atomicAdd(metrics + 1, red[1]);
    atomicAdd(metrics + 2, (float)HEAD_S);
    atomicAdd(step_out + 0, red[0] / (float)d.B);
    atomicAdd(step_out + 1, red[1] / (float)d.B);
  }
}

// =============================================================================================
// dense1 dgrad: da2 = dh * W1, epilogue = ReLU mask (a2 > 0) + max-pool backward through argmax
// M = samples (tile of 128), N tile = C2 channels of ONE pooled pixel, K = D1
// grid = (P2h*P2w, ceil(B/128)), block = 256 (warps w / w+4 take alternate 32-channel chunks)
// =============================================================================================
template <int C2, int D1>
__global__ void __launch_bounds__(TC_THREADS, 1)
dense1_dgrad_kernel(const __nv_bfloat16* __restrict__ dh, const __nv_bfloat16* __restrict__ w1h /*[D1][F]*/,
                    const __nv_bfloat16* __restrict__ a2, const uint8_t* __restrict__ idx2,
                    __nv_bfloat16* __restrict__ dy2 /*[B][P1h][P1w][C2], only [0:2P2h, 0:2P2w] written*/, TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  constexpr int K = D1, KCH = K / 8;
  const uint32_t a_smem = smem_u32(smem), b_smem = a_smem + 128u * K * 2u;
  constexpr uint32_t ncols = tmem_cols_for(C2);
  TileCtx tc = tile_begin(&bar, &slot, ncols);
  const int pix = blockIdx.x, s0 = blockIdx.y * 128;
  stage_kmajor(a_smem, dh, D1, s0, d.B, 0, 128, K);                 // A: [B][D1] K-major
  stage_mnmajor(b_smem, w1h, d.F, pix * C2, d.F, 0, C2, K);         // B: [K=D1][N=F] MN-major
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    issue_tile(tc.tmem, a_smem, b_smem, KCH, umma_idesc_bf16(128, C2, false, true));
    umma_commit(&bar);
  }
  const int lq = tc.warp & 3;
  const int s = s0 + lq * 32 + (int)lane_id();
  const uint32_t t_row = tc.tmem + ((uint32_t)(lq * 32) << 16);
  const int ph = pix / d.P2w, pw = pix % d.P2w;
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
#pragma unroll 1
  for (int c0 = (tc.warp >> 2) * 32; c0 < C2; c0 += 64) {
    uint32_t v[32];
    tmem_ld32(t_row + c0, v);
    tmem_ld_wait();
    if (s < d.B) {
      const size_t io = (size_t)s * d.F + (size_t)pix * C2 + c0;
      uint4 av[4];
      uint2 iv[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        av[k] = *reinterpret_cast<const uint4*>(a2 + io + k * 8);
        iv[k] = *reinterpret_cast<const uint2*>(idx2 + io + k * 8);
      }
      const uint32_t* aw = reinterpret_cast<const uint32_t*>(av);  // 16 words = 32 bf16
      const uint32_t* iw = reinterpret_cast<const uint32_t*>(iv);  // 8 words = 32 bytes
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t outw[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const uint32_t a_pair = aw[i >> 1];
          const uint32_t i0 = (iw[i >> 2] >> ((i & 3) * 8)) & 0xFF, i1 = (iw[i >> 2] >> (((i + 1) & 3) * 8)) & 0xFF;
          const bool on0 = (a_pair & 0x7FFFu) != 0 && !(a_pair & 0x8000u) && i0 == (uint32_t)q;
          const bool on1 = ((a_pair >> 16) & 0x7FFFu) != 0 && !((a_pair >> 16) & 0x8000u) && i1 == (uint32_t)q;
          outw[i >> 1] = pack_bf16x2(on0 ? __uint_as_float(v[i]) : 0.f, on1 ? __uint_as_float(v[i + 1]) : 0.f);
        }
        __nv_bfloat16* dst = dy2 + (((size_t)s * d.P1h + 2 * ph + (q >> 1)) * d.P1w + 2 * pw + (q & 1)) * C2 + c0;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          reinterpret_cast<uint4*>(dst)[k] = make_uint4(outw[4 * k], outw[4 * k + 1], outw[4 * k + 2], outw[4 * k + 3]);
      }
    }
  }
  tile_end(tc, ncols);
}

// =============================================================================================
// conv2 wgrad: dW2[co][(tap,ci)] += sum_pix a1[pix+tap][ci] * dy2[pix][co]   (+ db2[co] += sum_pix dy2)
// GEMM view: M = (tap,ci) = 9*C1 rows in ceil(9*C1/128) accumulator tiles, N = C2, reduction = pixels.
// Each CTA reduces KP = 128 pixels and adds its partial result to the gradient arena (fp32 RED).
// Both operands are MN-major (the reduction index is the slow index of a1 and of dy2).
// dy2 is stored on the a1 pixel grid [B][P1h][P1w][C2] (non-zero only in [0:2P2h, 0:2P2w]).
// grid = ceil(B*R*R2 / 128), block = 256
// =============================================================================================
template <int C1, int C2>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv2_wgrad_kernel(const __nv_bfloat16* __restrict__ a1, const __nv_bfloat16* __restrict__ dy2,
                   float* __restrict__ g_w2c /*[C2][9*C1]*/, float* __restrict__ g_b2c, TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  constexpr int KP = 128, KCH = KP / 8, KF = 9 * C1, MT = (KF + 127) / 128, CPT = C1 / 8, G2 = C2 / 8;
  constexpr uint32_t A_TILE = 128u * KP * 2u;
  const uint32_t a_smem = smem_u32(smem), b_smem = a_smem + MT * A_TILE;
  constexpr uint32_t ncols = tmem_cols_for(MT * C2);
  TileCtx tc = tile_begin(&bar, &slot, ncols);
  const int R = 2 * d.P2h, R2 = 2 * d.P2w, Ptot = d.B * R * R2;
  // lanes = 8 consecutive pixels x 4 consecutive channel chunks: a few 128-byte lines per cp.async
  __shared__ int rowoff[128];
  if (threadIdx.x < 128) {
    const int pix = blockIdx.x * KP + threadIdx.x;
    const int ow = pix % R2, t2 = pix / R2, oh = t2 % R, n = t2 / R;
    rowoff[threadIdx.x] = pix < Ptot ? ((n * d.P1h + oh) * d.P1w + ow) * C1 : -1;
  }
  __syncthreads();
  {
    const int k_in = threadIdx.x & 7;
#pragma unroll 3
    for (int i = threadIdx.x >> 3; i < 16 * MT * 16; i += TC_THREADS / 8) {
      // i -> (channel chunk c8, pixel group kg, tap); feature chunk g = tap*CPT + c8 (zero-filled past 9 taps)
      const int c8 = i % CPT, rest = i / CPT, kg = rest & 15, tap = rest >> 4;
      const int g = tap * CPT + c8;
      if (g >= MT * 16) break;
      const int k = kg * 8 + k_in, off = rowoff[k];
      const bool ok = off >= 0 && tap < 9;
      const __nv_bfloat16* src = a1 + (ok ? off + ((tap / 3) * d.P1w + tap % 3) * C1 + c8 * 8 : 0);
      cp_async16(a_smem + (g >> 4) * A_TILE + (g & 15) * (KCH * 128u) + kg * 128u + k_in * 16u, src, ok);
    }
#pragma unroll 2
    for (int i = threadIdx.x >> 3; i < 16 * G2; i += TC_THREADS / 8) {
      const int g2 = i % G2, kg = i / G2, k = kg * 8 + k_in;
      const int off = rowoff[k];  // a1 and dy2 share the pixel grid: a1 offset / C1 = pixel index
      const bool ok = off >= 0;
      cp_async16(b_smem + g2 * (KCH * 128u) + kg * 128u + k_in * 16u,
                 dy2 + (ok ? (size_t)(off / C1) * C2 + g2 * 8 : 0), ok);
    }
  }
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, C2, true, true);
    for (int t = 0; t < MT; ++t) issue_tile(tc.tmem + t * C2, a_smem + t * A_TILE, b_smem, KCH, idesc);
    umma_commit(&bar);
  }
  // bias gradient from the staged dy2 tile while the tensor core works (the MMA only reads this smem)
  if (threadIdx.x >= 128 && threadIdx.x < 128 + C2) {
    const int co = threadIdx.x - 128;
    float sacc = 0.f;
    const uint8_t* bs = smem + (b_smem - a_smem) + (co >> 3) * (KCH * 128) + (co & 7) * 2;
#pragma unroll 8
    for (int kk = 0; kk < KP; ++kk)
      sacc += bf16_bits_to_float(*reinterpret_cast<const uint16_t*>(bs + (kk >> 3) * 128 + (kk & 7) * 16));
    atomicAdd(g_b2c + co, sacc);
  }
  const int lq = tc.warp & 3, row = lq * 32 + (int)lane_id();
  const uint32_t t_row = tc.tmem + ((uint32_t)(lq * 32) << 16);
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
  // (tile, 32-column chunk) pairs are dealt alternately to the two warps of a lane quarter
#pragma unroll 1
  for (int job = tc.warp >> 2; job < MT * (C2 / 32); job += 2) {
    const int t = job / (C2 / 32), c0 = (job % (C2 / 32)) * 32;
    const int feat = t * 128 + row;
    uint32_t v[32];
    tmem_ld32(t_row + t * C2 + c0, v);
    tmem_ld_wait();
    if (feat < KF) {
#pragma unroll
      for (int i = 0; i < 32; ++i) atomicAdd(g_w2c + (size_t)(c0 + i) * KF + feat, __uint_as_float(v[i]));
    }
  }
  tile_end(tc, ncols);
}

// =============================================================================================
// conv2 dgrad: g1[n,ih,iw,ci] = relu'(a1) * sum_{tap,co} dy2[n,ih-kh,iw-kw,co] * W2[co][tap][ci]
// implicit GEMM: M = input pixels (tile 128), N = C1, K = 9*C2; A K-major (zero-filled halo),
// B = W2 viewed [K=(tap,co)][N=ci] MN-major.  grid = ceil(B*P1h*P1w / 128), block = 256
// =============================================================================================
template <int C1, int C2>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv2_dgrad_kernel(const __nv_bfloat16* __restrict__ dy2, const __nv_bfloat16* __restrict__ w2h /*[C2][9][C1]*/,
                   const __nv_bfloat16* __restrict__ a1, __nv_bfloat16* __restrict__ g1, TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  constexpr int K = 9 * C2, KCH = K / 8, CPT = C2 / 8, NG = C1 / 8, KF = 9 * C1;
  const uint32_t a_smem = smem_u32(smem), b_smem = a_smem + 128u * K * 2u;
  constexpr uint32_t ncols = tmem_cols_for(C1);
  TileCtx tc = tile_begin(&bar, &slot, ncols);
  const int R = 2 * d.P2h, R2 = 2 * d.P2w, Mtot = d.B * d.P1h * d.P1w;
  // lanes = 8 consecutive input pixels x 4 consecutive channel chunks of dy2 (coalesced cp.async)
  __shared__ int rowpos[128];  // (n*R*R2 + ih*R2 + iw) with ih, iw packed alongside: [31:16] base pixel, [15:8] ih, [7:0] iw
  __shared__ int rowbase[128];
  if (threadIdx.x < 128) {
    const int m = blockIdx.x * 128 + threadIdx.x;
    const int iw = m % d.P1w, t2 = m / d.P1w, ih = t2 % d.P1h, n = t2 / d.P1h;
    rowbase[threadIdx.x] = m < Mtot ? m : -(1 << 28);  // dy2 lives on the same P1h x P1w grid as a1
    rowpos[threadIdx.x] = (ih << 8) | iw;
  }
  __syncthreads();
  {
    const int r_in = threadIdx.x & 7;
#pragma unroll 3
    for (int i = threadIdx.x >> 3; i < 16 * 9 * CPT; i += TC_THREADS / 8) {
      const int c8 = i % CPT, rest = i / CPT, rg = rest & 15, tap = rest >> 4;
      const int r = rg * 8 + r_in, kh = tap / 3, kw = tap % 3;
      const int pos = rowpos[r], base = rowbase[r];
      const int oh = (pos >> 8) - kh, ow = (pos & 0xFF) - kw;
      const bool ok = base >= 0 && oh >= 0 && oh < R && ow >= 0 && ow < R2;
      const __nv_bfloat16* src = dy2 + (ok ? (size_t)(base - kh * d.P1w - kw) * C2 + c8 * 8 : 0);
      cp_async16(a_smem + core_off_kmajor(r, tap * CPT + c8, KCH), src, ok);
    }
    // B: chunk (k=(tap,co), g=ci8) <- w2h[co][tap][g*8..]; a quarter-warp fills one core matrix
#pragma unroll 2
    for (int i = threadIdx.x; i < K * NG; i += TC_THREADS) {
      const int k_in = i & 7, rest = i >> 3, g = rest % NG, kk = (rest / NG) * 8 + k_in;
      const int tap = kk / C2, co = kk % C2;
      cp_async16(b_smem + core_off_mnmajor(kk, g, KCH), w2h + (size_t)co * KF + tap * C1 + g * 8, true);
    }
  }
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    issue_tile(tc.tmem, a_smem, b_smem, KCH, umma_idesc_bf16(128, C1, false, true));
    umma_commit(&bar);
  }
  const int lq = tc.warp & 3;
  const int row_m = blockIdx.x * 128 + lq * 32 + (int)lane_id();
  const uint32_t t_row = tc.tmem + ((uint32_t)(lq * 32) << 16);
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
  // 16-channel chunks dealt alternately to the two warps of a lane quarter
#pragma unroll 1
  for (int c0 = (tc.warp >> 2) * 16; c0 < C1; c0 += 32) {
    uint32_t v[16];
    tmem_ld16(t_row + c0, v);
    tmem_ld_wait();
    if (row_m < Mtot) {
      const size_t o = (size_t)row_m * C1 + c0;
      uint4 av[2];
#pragma unroll
      for (int k4 = 0; k4 < 2; ++k4) av[k4] = *reinterpret_cast<const uint4*>(a1 + o + k4 * 8);
      const uint32_t* aw = reinterpret_cast<const uint32_t*>(av);
      uint32_t outw[8];
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        const uint32_t a_pair = aw[i >> 1];
        const bool on0 = (a_pair & 0x7FFFu) != 0 && !(a_pair & 0x8000u);
        const bool on1 = ((a_pair >> 16) & 0x7FFFu) != 0 && !((a_pair >> 16) & 0x8000u);
        outw[i >> 1] = pack_bf16x2(on0 ? __uint_as_float(v[i]) : 0.f, on1 ? __uint_as_float(v[i + 1]) : 0.f);
      }
#pragma unroll
      for (int k4 = 0; k4 < 2; ++k4)
        reinterpret_cast<uint4*>(g1 + o)[k4] = make_uint4(outw[4 * k4], outw[4 * k4 + 1], outw[4 * k4 + 2], outw[4 * k4 + 3]);
    }
  }
  tile_end(tc, ncols);
}

// =============================================================================================
// conv1 wgrad: dW1[c][tap] += sum_{n,ph,pw} g1[n,ph,pw,c] * x[n, 2ph+dy+kh, 2pw+dx+kw], (dy,dx) = argmax
// grid = B blocks (one image each), 1024 threads: warp = channel, lanes stride over the pooled pixels
// with 9 independent accumulators each; warp-shuffle reduction, one RED per (channel, tap) per image.
// =============================================================================================
__global__ void __launch_bounds__(1024, 1)
conv1_wgrad_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ g1, const uint8_t* __restrict__ idx1,
                   float* __restrict__ g_w1c, float* __restrict__ g_b1c, TfdlCnnDims d) {
  extern __shared__ float smf[];
  const int n = blockIdx.x, npix = d.P1h * d.P1w, C1 = d.C1;
  const int ld = npix | 1;                            // odd pitch: conflict-free transposed stores
  float* xs = smf;                                    // [H0*W0]
  float* gs = xs + d.H0 * d.W0;                       // [C1][ld]
  uint8_t* qs = reinterpret_cast<uint8_t*>(gs + (size_t)C1 * ld);  // [C1][ld]
  for (int i = threadIdx.x; i < d.H0 * d.W0; i += blockDim.x) xs[i] = x[(size_t)n * d.H0 * d.W0 + i];
  {
    // 8 elements per thread per pass: one 16-byte load of g1 and one 8-byte load of idx1
    const uint4* gsrc = reinterpret_cast<const uint4*>(g1 + (size_t)n * npix * C1);
    const uint2* isrc = reinterpret_cast<const uint2*>(idx1 + (size_t)n * npix * C1);
    const int nvec = npix * C1 / 8;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
      const uint4 gv = gsrc[v];
      const uint2 iv = isrc[v];
      const int e = v * 8, p = e / C1, c0 = e - p * C1;
      const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w};
      const uint32_t iw[2] = {iv.x, iv.y};
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t bits = (gw[k >> 1] >> ((k & 1) * 16)) & 0xFFFFu;
        gs[(c0 + k) * ld + p] = __uint_as_float(bits << 16);
        qs[(c0 + k) * ld + p] = (uint8_t)((iw[k >> 2] >> ((k & 3) * 8)) & 0xFF);
      }
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int c = warp; c < C1; c += nwarps) {
    float acc[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float bacc = 0.f;
    for (int p = lane; p < npix; p += 32) {
      const float g = gs[c * ld + p];
      const int q = qs[c * ld + p];
      const int ph = p / d.P1w, pw = p - ph * d.P1w;
      const float* xp = xs + (2 * ph + (q >> 1)) * d.W0 + 2 * pw + (q & 1);
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) acc[kh * 3 + kw] = fmaf(g, xp[kh * d.W0 + kw], acc[kh * 3 + kw]);
      bacc += g;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], o);
      bacc += __shfl_xor_sync(0xffffffffu, bacc, o);
    }
    if (lane < 9) {
      float v = acc[0];
#pragma unroll
      for (int t = 1; t < 9; ++t) v = lane == t ? acc[t] : v;
      atomicAdd(g_w1c + c * 9 + lane, v);
    } else if (lane == 9) {
      atomicAdd(g_b1c + c, bacc);
    }
  }
}

}  // namespace tfdl

using namespace tfdl;

// Copies the phase stamps out (64 CTAs x 8 slots); returns -1 when the library was built without them.
extern "C" int tfdl_debug_read_phases(long long* host_out /*[256*8]*/) {
#ifdef TFDL_PROFILE_PHASES
  TFDL_CUDA_OK(cudaMemcpyFromSymbol(host_out, g_phase, sizeof(long long) * 256 * 8));
  return 0;
#else
  (void)host_out;
  return -1;
#endif
}

static int check_dims(const TfdlCnnDims& d) {
  TFDL_REQUIRE(d.B > 0 && d.B % 16 == 0, "cnn: per-replica batch %d must be a positive multiple of 16", d.B);
  TFDL_REQUIRE(d.C1 % 32 == 0 && d.C1 <= 64 && d.P1w * d.C1 <= 1024, "cnn: conv1 filters %d unsupported", d.C1);
  TFDL_REQUIRE(d.C2 % 32 == 0 && d.C2 <= 128, "cnn: conv2 filters %d unsupported (multiple of 32, <= 128)", d.C2);
  TFDL_REQUIRE(d.F == d.P2h * d.P2w * d.C2 && d.F % 16 == 0, "cnn: flatten size mismatch");
  TFDL_REQUIRE(d.D1 % 16 == 0 && d.D1 <= 256, "cnn: dense units %d unsupported (multiple of 16, <= 256)", d.D1);
  TFDL_REQUIRE(d.NC >= 2 && d.NC <= 64, "cnn: %d classes unsupported", d.NC);
  TFDL_REQUIRE(d.P1h == (d.H0 - 2) / 2 && d.P1w == (d.W0 - 2) / 2 && d.P2h == (d.P1h - 2) / 2 &&
                   d.P2w == (d.P1w - 2) / 2, "cnn: inconsistent spatial dims");
  TFDL_REQUIRE(d.nrep >= 1, "cnn: nrep must be >= 1");
  return 0;
}

template <class K>
static int set_smem(K kern, size_t bytes) {
  TFDL_REQUIRE(bytes <= 227 * 1024, "cnn: kernel needs %zu bytes of shared memory (> 227 KB)", bytes);
  TFDL_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

extern "C" int tfdl_cnn_conv1_fwd(const float* x, const float* w, const float* b, void* a1, uint8_t* idx1,
                                  float* clear2, const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  conv1_fwd_kernel<<<d->B * d->P1h, d->P1w * d->C1, 4 * d->W0 * sizeof(float), (cudaStream_t)stream>>>(
      x, w, b, (__nv_bfloat16*)a1, idx1, clear2, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// Channel counts are template parameters (address math folds to immediates, loops unroll); the
// supported combinations are instantiated here.
#define TFDL_DISPATCH_C1C2(C1v, C2v, CALL)                                                   \
  if (C1v == 32 && C2v == 64) { CALL(32, 64); }                                               \
  else if (C1v == 32 && C2v == 32) { CALL(32, 32); }                                          \
  else if (C1v == 32 && C2v == 128) { CALL(32, 128); }                                        \
  else if (C1v == 64 && C2v == 64) { CALL(64, 64); }                                          \
  else if (C1v == 64 && C2v == 128) { CALL(64, 128); }                                        \
  else { TFDL_REQUIRE(false, "cnn: no kernel instance for conv filters (%d, %d)", C1v, C2v); }

extern "C" int tfdl_cnn_conv2_fwd(const void* a1, const void* w2h, const float* b2, void* a2, uint8_t* idx2,
                                  const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)(128 + 128) * 9 * d->C1 * 2;
  int M = d->B * d->P2h * d->P2w * 4;
#define CALL(A, Bc)                                                                                            \
  {                                                                                                            \
    if (int rc = set_smem(conv2_fwd_kernel<A, Bc>, smem)) return rc;                                           \
    conv2_fwd_kernel<A, Bc><<<(M + 127) / 128, TC_THREADS, smem, (cudaStream_t)stream>>>(                      \
        (const __nv_bfloat16*)a1, (const __nv_bfloat16*)w2h, b2, (__nv_bfloat16*)a2, idx2, *d);                \
  }
  TFDL_DISPATCH_C1C2(d->C1, d->C2, CALL)
#undef CALL
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_head(float* acc, const float* b1, const float* w2, const float* b2, const int* labels, void* dh,
                             float* g_w2, float* g_b2, float* g_b1, float* metrics, float* step_out,
                             const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  TFDL_REQUIRE(d->B % HEAD_S == 0, "cnn head: batch must be a multiple of %d", HEAD_S);
  size_t smem = ((size_t)2 * HEAD_S * d->D1 + (size_t)d->NC * d->D1 + (size_t)HEAD_S * d->NC + 2) * sizeof(float);
  if (int rc = set_smem(head_kernel, smem)) return rc;
  head_kernel<<<d->B / HEAD_S, HEAD_S * 32, smem, (cudaStream_t)stream>>>(
      acc, b1, w2, b2, labels, (__nv_bfloat16*)dh, g_w2, g_b2, g_b1, metrics, step_out, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_dense1_dgrad(const void* dh, const void* w1h, const void* a2, const uint8_t* idx2, void* dy2,
                                     const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)(128 + d->C2) * d->D1 * 2;
  dim3 grid(d->P2h * d->P2w, (d->B + 127) / 128);
#define CALL(C2v, D1v)                                                                                          \
  {                                                                                                             \
    if (int rc = set_smem(dense1_dgrad_kernel<C2v, D1v>, smem)) return rc;                                      \
    dense1_dgrad_kernel<C2v, D1v><<<grid, TC_THREADS, smem, (cudaStream_t)stream>>>(                            \
        (const __nv_bfloat16*)dh, (const __nv_bfloat16*)w1h, (const __nv_bfloat16*)a2, idx2, (__nv_bfloat16*)dy2, *d); \
  }
  if (d->C2 == 64 && d->D1 == 128) CALL(64, 128)
  else if (d->C2 == 64 && d->D1 == 256) CALL(64, 256)
  else if (d->C2 == 64 && d->D1 == 64) CALL(64, 64)
  else if (d->C2 == 32 && d->D1 == 128) CALL(32, 128)
  else if (d->C2 == 128 && d->D1 == 128) CALL(128, 128)
  else if (d->C2 == 128 && d->D1 == 256) CALL(128, 256)
  else TFDL_REQUIRE(false, "cnn: no dense1_dgrad instance for (C2=%d, D1=%d)", d->C2, d->D1);
#undef CALL
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_wgrad(const void* a1, const void* dy2, float* g_w2c, float* g_b2c, const TfdlCnnDims* d,
                                    void* stream) {
  if (int rc = check_dims(*d)) return rc;
  int mtiles = (9 * d->C1 + 127) / 128;
  TFDL_REQUIRE(mtiles * d->C2 <= 512, "cnn: conv2 wgrad accumulators exceed TMEM");
  size_t smem = (size_t)mtiles * 128 * 128 * 2 + (size_t)128 * d->C2 * 2;
  int P = d->B * 4 * d->P2h * d->P2w;
#define CALL(A, Bc)                                                                                  \
  {                                                                                                  \
    if (int rc = set_smem(conv2_wgrad_kernel<A, Bc>, smem)) return rc;                               \
    conv2_wgrad_kernel<A, Bc><<<(P + 127) / 128, TC_THREADS, smem, (cudaStream_t)stream>>>(          \
        (const __nv_bfloat16*)a1, (const __nv_bfloat16*)dy2, g_w2c, g_b2c, *d);                      \
  }
  TFDL_DISPATCH_C1C2(d->C1, d->C2, CALL)
#undef CALL
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_dgrad(const void* dy2, const void* w2h, const void* a1, void* g1, const TfdlCnnDims* d,
                                    void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)(128 + d->C1) * 9 * d->C2 * 2;
  int M = d->B * d->P1h * d->P1w;
#define CALL(A, Bc)                                                                                       \
  {                                                                                                       \
    if (int rc = set_smem(conv2_dgrad_kernel<A, Bc>, smem)) return rc;                                    \
    conv2_dgrad_kernel<A, Bc><<<(M + 127) / 128, TC_THREADS, smem, (cudaStream_t)stream>>>(               \
        (const __nv_bfloat16*)dy2, (const __nv_bfloat16*)w2h, (const __nv_bfloat16*)a1, (__nv_bfloat16*)g1, *d); \
  }
  TFDL_DISPATCH_C1C2(d->C1, d->C2, CALL)
#undef CALL
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv1_wgrad(const float* x, const void* g1, const uint8_t* idx1, float* g_w1c, float* g_b1c,
                                    const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  const int ld = (d->P1h * d->P1w) | 1;
  TFDL_REQUIRE((d->P1h * d->P1w * d->C1) % 8 == 0, "cnn conv1 wgrad: pooled map size must be a multiple of 8 elements");
  size_t smem = (size_t)d->H0 * d->W0 * 4 + (size_t)d->C1 * ld * 5 + 16;
  if (int rc = set_smem(conv1_wgrad_kernel, smem)) return rc;
  conv1_wgrad_kernel<<<d->B, 1024, smem, (cudaStream_t)stream>>>(x, (const __nv_bfloat16*)g1, idx1, g_w1c, g_b1c, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
