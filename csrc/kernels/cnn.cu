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

// Common prologue: TMEM alloc (warp 0), mbarrier init, block sync. 128 threads.
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
                                 __nv_bfloat16* __restrict__ a1, uint8_t* __restrict__ idx1, TfdlCnnDims d) {
  extern __shared__ float rows[];  // 4 input rows
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
// conv2 forward: implicit GEMM on tcgen05 with fused bias + pool + ReLU epilogue
// rows of the M tile are ordered (n, ph, pw, dy, dx): 4 consecutive rows = one pooling window
// grid = ceil(B*P2h*P2w*4 / 128), block = 128
// =============================================================================================
__global__ void __launch_bounds__(128, 1)
conv2_fwd_kernel(const __nv_bfloat16* __restrict__ a1, const __nv_bfloat16* __restrict__ w2h,
                 const float* __restrict__ b2, __nv_bfloat16* __restrict__ a2, uint8_t* __restrict__ idx2,
                 TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int K = 9 * d.C1, kchunks = K >> 3, cpt = d.C1 >> 3;
  const uint32_t a_smem = smem_u32(smem), b_smem = a_smem + 128u * K * 2u;
  const uint32_t ncols = tmem_cols_for(d.C2);
  TileCtx tc = tile_begin(&bar, &slot, ncols);

  const int Mtot = d.B * d.P2h * d.P2w * 4;
  const int r = threadIdx.x, m = blockIdx.x * 128 + r;
  {  // A: thread = row; 9 taps x cpt chunks
    const bool ok = m < Mtot;
    const int q = m & 3, win = m >> 2;
    const int pw = win % d.P2w, t2 = win / d.P2w, ph = t2 % d.P2h, n = t2 / d.P2h;
    const int oh = 2 * ph + (q >> 1), ow = 2 * pw + (q & 1);
    const __nv_bfloat16* base = ok ? a1 + (((size_t)n * d.P1h + oh) * d.P1w + ow) * d.C1 : a1;
    int j = 0;
    for (int kh = 0; kh < 3; ++kh)
      for (int kw = 0; kw < 3; ++kw) {
        const __nv_bfloat16* src = base + ((size_t)kh * d.P1w + kw) * d.C1;
        for (int c8 = 0; c8 < cpt; ++c8, ++j)
          cp_async16(a_smem + core_off_kmajor(r, j, kchunks), ok ? src + c8 * 8 : a1, ok);
      }
  }
  stage_kmajor(b_smem, w2h, K, 0, d.C2, 0, d.C2, K);  // B: [C2][K] K-major
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    issue_tile(tc.tmem, a_smem, b_smem, kchunks, umma_idesc_bf16(128, d.C2, false, false));
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after_sync();

  const int lane = lane_id(), q = lane & 3;
  const uint32_t t_row = tc.tmem + ((uint32_t)(tc.warp * 32) << 16);
  const int win = m >> 2;
  for (int c0 = 0; c0 < d.C2; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + c0, v);
    tmem_ld_wait();
    // per column: max / first-argmax over the 4 lanes of the window
    uint32_t packed_idx[8];  // 32 two-bit... kept as bytes below
    float outv[32];
    uint8_t outi[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float val = __uint_as_float(v[i]) + __ldg(b2 + c0 + i);
      int arg = q;
      float o1 = __shfl_xor_sync(0xffffffffu, val, 1);
      int a1i = q ^ 1;
      if (o1 > val || (o1 == val && a1i < arg)) { val = o1; arg = a1i; }
      float o2 = __shfl_xor_sync(0xffffffffu, val, 2);
      int a2i = __shfl_xor_sync(0xffffffffu, arg, 2);
      if (o2 > val || (o2 == val && a2i < arg)) { val = o2; arg = a2i; }
      outv[i] = fmaxf(val, 0.f);
      outi[i] = (uint8_t)arg;
    }
    (void)packed_idx;
    // lane q of each window writes 8 of the 32 columns (16 B bf16 + 8 B argmax)
    if (m < Mtot) {
      const size_t o = (size_t)win * d.C2 + c0 + q * 8;
      uint4 pk;
      uint32_t lo = 0, hi = 0;
      // select this lane's 8 columns without dynamic register indexing
      float s[8];
      uint8_t si[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i] = q == 0 ? outv[i] : q == 1 ? outv[8 + i] : q == 2 ? outv[16 + i] : outv[24 + i];
        si[i] = q == 0 ? outi[i] : q == 1 ? outi[8 + i] : q == 2 ? outi[16 + i] : outi[24 + i];
      }
      pk = make_uint4(pack_bf16x2(s[0], s[1]), pack_bf16x2(s[2], s[3]), pack_bf16x2(s[4], s[5]),
                      pack_bf16x2(s[6], s[7]));
      lo = si[0] | (si[1] << 8) | (si[2] << 16) | (si[3] << 24);
      hi = si[4] | (si[5] << 8) | (si[6] << 16) | (si[7] << 24);
      *reinterpret_cast<uint4*>(a2 + o) = pk;
      *reinterpret_cast<uint2*>(idx2 + o) = make_uint2(lo, hi);
    }
  }
  tile_end(tc, ncols);
}

// =============================================================================================
// head: Dense(D1->NC) + softmax cross-entropy + metrics + all tiny gradients. One block.
//   acc  [B][D1] fp32 : dense1 pre-activation WITHOUT bias (split-K scratch), re-zeroed here
//   dh   [B][D1] bf16 : dL/d(dense1 pre-activation) (ReLU-masked)
//   metrics[0..2] += (sum of per-example loss, correct predictions, examples); step_out[0] = mean loss
// =============================================================================================
__global__ void __launch_bounds__(512, 1)
head_kernel(float* __restrict__ acc, const float* __restrict__ b1, const float* __restrict__ w2 /*[NC][D1]*/,
            const float* __restrict__ b2, const int* __restrict__ labels, __nv_bfloat16* __restrict__ dh,
            float* __restrict__ g_w2, float* __restrict__ g_b2, float* __restrict__ g_b1,
            float* __restrict__ metrics, float* __restrict__ step_out, TfdlCnnDims d) {
  extern __shared__ float sm[];
  const int B = d.B, D1 = d.D1, NC = d.NC, T = blockDim.x, t = threadIdx.x;
  float* h = sm;                    // [B][D1]
  float* w2s = h + (size_t)B * D1;  // [NC][D1]
  float* dl = w2s + NC * D1;        // [B][NC] logits -> dlogits
  float* red = dl + B * NC;         // [2] loss, correct
  for (int i = t; i < B * D1; i += T) {
    h[i] = fmaxf(acc[i] + b1[i % D1], 0.f);
    acc[i] = 0.f;
  }
  for (int i = t; i < NC * D1; i += T) w2s[i] = w2[i];
  if (t < 2) red[t] = 0.f;
  __syncthreads();
  for (int i = t; i < B * NC; i += T) {
    const int s = i / NC, j = i % NC;
    float a = b2[j];
    const float* hs = h + (size_t)s * D1;
    const float* wj = w2s + j * D1;
    for (int k = 0; k < D1; ++k) a = fmaf(hs[k], wj[k], a);
    dl[i] = a;
  }
  __syncthreads();
  const float inv = 1.f / ((float)B * (float)d.nrep);
  for (int s = t; s < B; s += T) {
    float* z = dl + s * NC;
    float mx = z[0];
    int am = 0;
    for (int j = 1; j < NC; ++j)
      if (z[j] > mx) { mx = z[j]; am = j; }
    float sum = 0.f;
    for (int j = 0; j < NC; ++j) sum += __expf(z[j] - mx);
    const int y = labels[s];
    const float loss = __logf(sum) + mx - z[y];
    for (int j = 0; j < NC; ++j) z[j] = (__expf(z[j] - mx) / sum - (j == y ? 1.f : 0.f)) * inv;
    atomicAdd(&red[0], loss);
    atomicAdd(&red[1], am == y ? 1.f : 0.f);
  }
  __syncthreads();
  // dW2[j][k] = sum_s dl[s][j] h[s][k];  db2[j] = sum_s dl[s][j]
  for (int i = t; i < NC * D1; i += T) {
    const int j = i / D1, k = i % D1;
    float a = 0.f;
    for (int s = 0; s < B; ++s) a = fmaf(dl[s * NC + j], h[(size_t)s * D1 + k], a);
    g_w2[i] = a;
  }
  for (int j = t; j < NC; j += T) {
    float a = 0.f;
    for (int s = 0; s < B; ++s) a += dl[s * NC + j];
    g_b2[j] = a;
  }
  // dh[s][k] = relu'(h) * sum_j dl[s][j] w2[j][k]
  for (int i = t; i < B * D1; i += T) {
    const int s = i / D1, k = i % D1;
    float a = 0.f;
    for (int j = 0; j < NC; ++j) a = fmaf(dl[s * NC + j], w2s[j * D1 + k], a);
    a = h[i] > 0.f ? a : 0.f;
    h[i] = a;  // reuse as dh (fp32) for the bias gradient
    dh[i] = __float2bfloat16(a);
  }
  __syncthreads();
  for (int k = t; k < D1; k += T) {
    float a = 0.f;
    for (int s = 0; s < B; ++s) a += h[(size_t)s * D1 + k];
    g_b1[k] = a;
  }
  if (t == 0) {
    metrics[0] += red[0];
    metrics[1] += red[1];
    metrics[2] += (float)B;
    step_out[0] = red[0] / (float)B;
    step_out[1] = red[1] / (float)B;
  }
}

// =============================================================================================
// dense1 dgrad: da2 = dh * W1, epilogue = ReLU mask (a2 > 0) + max-pool backward through argmax
// M = samples (tile of 128), N tile = C2 channels of ONE pooled pixel, K = D1
// grid = (P2h*P2w, ceil(B/128)), block = 128
// =============================================================================================
__global__ void __launch_bounds__(128, 1)
dense1_dgrad_kernel(const __nv_bfloat16* __restrict__ dh, const __nv_bfloat16* __restrict__ w1h /*[D1][F]*/,
                    const __nv_bfloat16* __restrict__ a2, const uint8_t* __restrict__ idx2,
                    __nv_bfloat16* __restrict__ dy2 /*[B][2P2h][2P2w][C2]*/, TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int K = d.D1, kchunks = K >> 3;
  const uint32_t a_smem = smem_u32(smem), b_smem = a_smem + 128u * K * 2u;
  const uint32_t ncols = tmem_cols_for(d.C2);
  TileCtx tc = tile_begin(&bar, &slot, ncols);
  const int pix = blockIdx.x, s0 = blockIdx.y * 128;
  stage_kmajor(a_smem, dh, d.D1, s0, d.B, 0, 128, K);                 // A: [B][D1] K-major
  stage_mnmajor(b_smem, w1h, d.F, pix * d.C2, d.F, 0, d.C2, K);       // B: [K=D1][N=F] MN-major
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    issue_tile(tc.tmem, a_smem, b_smem, kchunks, umma_idesc_bf16(128, d.C2, false, true));
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
  const int s = s0 + tc.warp * 32 + (int)lane_id();
  const uint32_t t_row = tc.tmem + ((uint32_t)(tc.warp * 32) << 16);
  const int ph = pix / d.P2w, pw = pix % d.P2w, R = 2 * d.P2h, R2 = 2 * d.P2w;
  for (int c0 = 0; c0 < d.C2; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + c0, v);
    tmem_ld_wait();
    if (s < d.B) {
      const size_t io = (size_t)s * d.F + (size_t)pix * d.C2 + c0;
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
        __nv_bfloat16* dst = dy2 + (((size_t)s * R + 2 * ph + (q >> 1)) * R2 + 2 * pw + (q & 1)) * d.C2 + c0;
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
// grid = ceil(B*R*R2 / 128), block = 128
// =============================================================================================
__global__ void __launch_bounds__(128, 1)
conv2_wgrad_kernel(const __nv_bfloat16* __restrict__ a1, const __nv_bfloat16* __restrict__ dy2,
                   float* __restrict__ g_w2c /*[C2][9*C1]*/, float* __restrict__ g_b2c, TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  constexpr int KP = 128, kchunks = KP >> 3;
  const int KF = 9 * d.C1, mtiles = (KF + 127) / 128, cpt = d.C1 >> 3;
  const uint32_t a_smem = smem_u32(smem);                       // mtiles tiles of [128 feat][KP] MN-major
  const uint32_t a_tile_bytes = 128u * KP * 2u;
  const uint32_t b_smem = a_smem + mtiles * a_tile_bytes;       // [KP][C2] MN-major
  const uint32_t ncols = tmem_cols_for(mtiles * d.C2);
  TileCtx tc = tile_begin(&bar, &slot, ncols);
  const int R = 2 * d.P2h, R2 = 2 * d.P2w, Ptot = d.B * R * R2;
  const int k = threadIdx.x, pix = blockIdx.x * KP + k;
  const bool ok = pix < Ptot;
  const int ow = pix % R2, t2 = pix / R2, oh = t2 % R, n = t2 / R;
  {
    const __nv_bfloat16* base = ok ? a1 + (((size_t)n * d.P1h + oh) * d.P1w + ow) * d.C1 : a1;
    int g = 0;  // global feature chunk index = tap*cpt + c8
    for (int kh = 0; kh < 3; ++kh)
      for (int kw = 0; kw < 3; ++kw) {
        const __nv_bfloat16* src = base + ((size_t)kh * d.P1w + kw) * d.C1;
        for (int c8 = 0; c8 < cpt; ++c8, ++g)
          cp_async16(a_smem + (g >> 4) * a_tile_bytes + core_off_mnmajor(k, g & 15, kchunks), ok ? src + c8 * 8 : a1, ok);
      }
    // zero the unused feature chunks of the last tile: garbage * 0 would still be NaN-prone
    for (; g < mtiles * 16; ++g)
      cp_async16(a_smem + (g >> 4) * a_tile_bytes + core_off_mnmajor(k, g & 15, kchunks), a1, false);
    const __nv_bfloat16* bsrc = ok ? dy2 + (size_t)pix * d.C2 : dy2;
    for (int g2 = 0; g2 < (d.C2 >> 3); ++g2)
      cp_async16(b_smem + core_off_mnmajor(k, g2, kchunks), ok ? bsrc + g2 * 8 : dy2, ok);
  }
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    const uint32_t idesc = umma_idesc_bf16(128, d.C2, true, true);
    for (int t = 0; t < mtiles; ++t) issue_tile(tc.tmem + t * d.C2, a_smem + t * a_tile_bytes, b_smem, kchunks, idesc);
    umma_commit(&bar);
  }
  // bias gradient from the staged dy2 tile while the tensor core works (generic-proxy reads of
  // smem that the MMA only reads are safe)
  if (threadIdx.x < d.C2) {
    const int co = threadIdx.x;
    float sacc = 0.f;
    const uint8_t* bs = smem + (b_smem - a_smem);
    for (int kk = 0; kk < KP; ++kk) {
      const uint16_t bits = *reinterpret_cast<const uint16_t*>(bs + (co >> 3) * (kchunks * 128) + (kk >> 3) * 128 +
                                                               (kk & 7) * 16 + (co & 7) * 2);
      sacc += bf16_bits_to_float(bits);
    }
    atomicAdd(g_b2c + co, sacc);
  }
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
  const int row = tc.warp * 32 + (int)lane_id();
  const uint32_t t_row = tc.tmem + ((uint32_t)(tc.warp * 32) << 16);
  for (int t = 0; t < mtiles; ++t) {
    const int feat = t * 128 + row;
    for (int c0 = 0; c0 < d.C2; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(t_row + t * d.C2 + c0, v);
      tmem_ld_wait();
      if (feat < KF) {
#pragma unroll
        for (int i = 0; i < 32; ++i) atomicAdd(g_w2c + (size_t)(c0 + i) * KF + feat, __uint_as_float(v[i]));
      }
    }
  }
  tile_end(tc, ncols);
}

// =============================================================================================
// conv2 dgrad: g1[n,ih,iw,ci] = relu'(a1) * sum_{tap,co} dy2[n,ih-kh,iw-kw,co] * W2[co][tap][ci]
// implicit GEMM: M = input pixels (tile 128), N = C1, K = 9*C2; A K-major (zero-filled halo),
// B = W2 viewed [K=(tap,co)][N=ci] MN-major.  grid = ceil(B*P1h*P1w / 128), block = 128
// =============================================================================================
__global__ void __launch_bounds__(128, 1)
conv2_dgrad_kernel(const __nv_bfloat16* __restrict__ dy2, const __nv_bfloat16* __restrict__ w2h /*[C2][9][C1]*/,
                   const __nv_bfloat16* __restrict__ a1, __nv_bfloat16* __restrict__ g1, TfdlCnnDims d) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int K = 9 * d.C2, kchunks = K >> 3, cpt = d.C2 >> 3;
  const uint32_t a_smem = smem_u32(smem), b_smem = a_smem + 128u * K * 2u;
  const uint32_t ncols = tmem_cols_for(d.C1);
  TileCtx tc = tile_begin(&bar, &slot, ncols);
  const int R = 2 * d.P2h, R2 = 2 * d.P2w, Mtot = d.B * d.P1h * d.P1w;
  const int r = threadIdx.x, m = blockIdx.x * 128 + r;
  {
    const bool okm = m < Mtot;
    const int iw = m % d.P1w, t2 = m / d.P1w, ih = t2 % d.P1h, n = t2 / d.P1h;
    int j = 0;
    for (int kh = 0; kh < 3; ++kh)
      for (int kw = 0; kw < 3; ++kw) {
        const int oh = ih - kh, ow = iw - kw;
        const bool ok = okm && oh >= 0 && oh < R && ow >= 0 && ow < R2;
        const __nv_bfloat16* src = ok ? dy2 + (((size_t)n * R + oh) * R2 + ow) * d.C2 : dy2;
        for (int c8 = 0; c8 < cpt; ++c8, ++j) cp_async16(a_smem + core_off_kmajor(r, j, kchunks), src + (ok ? c8 * 8 : 0), ok);
      }
    // B: chunk (k=(tap,co), g=ci8) <- w2h[co][tap][g*8..]
    const int ng = d.C1 >> 3, total = K * ng, KF = 9 * d.C1;
    for (int i = threadIdx.x; i < total; i += 128) {
      const int g = i % ng, kk = i / ng, tap = kk / d.C2, co = kk % d.C2;
      cp_async16(b_smem + core_off_mnmajor(kk, g, kchunks), w2h + (size_t)co * KF + tap * d.C1 + g * 8, true);
    }
  }
  staging_done();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    issue_tile(tc.tmem, a_smem, b_smem, kchunks, umma_idesc_bf16(128, d.C1, false, true));
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after_sync();
  const int row_m = blockIdx.x * 128 + tc.warp * 32 + (int)lane_id();
  const uint32_t t_row = tc.tmem + ((uint32_t)(tc.warp * 32) << 16);
  for (int c0 = 0; c0 < d.C1; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(t_row + c0, v);
    tmem_ld_wait();
    if (row_m < Mtot) {
      const size_t o = (size_t)row_m * d.C1 + c0;
      uint4 av[4];
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) av[k4] = *reinterpret_cast<const uint4*>(a1 + o + k4 * 8);
      const uint32_t* aw = reinterpret_cast<const uint32_t*>(av);
      uint32_t outw[16];
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        const uint32_t a_pair = aw[i >> 1];
        const bool on0 = (a_pair & 0x7FFFu) != 0 && !(a_pair & 0x8000u);
        const bool on1 = ((a_pair >> 16) & 0x7FFFu) != 0 && !((a_pair >> 16) & 0x8000u);
        outw[i >> 1] = pack_bf16x2(on0 ? __uint_as_float(v[i]) : 0.f, on1 ? __uint_as_float(v[i + 1]) : 0.f);
      }
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4)
        reinterpret_cast<uint4*>(g1 + o)[k4] = make_uint4(outw[4 * k4], outw[4 * k4 + 1], outw[4 * k4 + 2], outw[4 * k4 + 3]);
    }
  }
  tile_end(tc, ncols);
}

// =============================================================================================
// conv1 wgrad: dW1[c][tap] += sum_{n,ph,pw} g1[n,ph,pw,c] * x[n, 2ph+dy+kh, 2pw+dx+kw], (dy,dx) = argmax
// grid = B blocks (one image each), block = 9*C1 threads
// =============================================================================================
__global__ void conv1_wgrad_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ g1,
                                   const uint8_t* __restrict__ idx1, float* __restrict__ g_w1c, float* __restrict__ g_b1c,
                                   TfdlCnnDims d) {
  extern __shared__ float smf[];
  const int n = blockIdx.x, npix = d.P1h * d.P1w;
  float* xs = smf;                                   // [H0*W0]
  float* gs = xs + d.H0 * d.W0;                      // [npix][C1]
  uint8_t* is = reinterpret_cast<uint8_t*>(gs + (size_t)npix * d.C1);  // [npix][C1]
  for (int i = threadIdx.x; i < d.H0 * d.W0; i += blockDim.x) xs[i] = x[(size_t)n * d.H0 * d.W0 + i];
  for (int i = threadIdx.x; i < npix * d.C1; i += blockDim.x) {
    gs[i] = __bfloat162float(g1[(size_t)n * npix * d.C1 + i]);
    is[i] = idx1[(size_t)n * npix * d.C1 + i];
  }
  __syncthreads();
  const int c = threadIdx.x % d.C1, tap = threadIdx.x / d.C1;
  if (tap >= 9) return;
  const int kh = tap / 3, kw = tap % 3;
  float acc = 0.f, bacc = 0.f;
  for (int p = 0; p < npix; ++p) {
    const float g = gs[p * d.C1 + c];
    const int q = is[p * d.C1 + c];
    const int ph = p / d.P1w, pw = p % d.P1w;
    acc = fmaf(g, xs[(2 * ph + (q >> 1) + kh) * d.W0 + 2 * pw + (q & 1) + kw], acc);
    bacc += g;
  }
  atomicAdd(g_w1c + c * 9 + tap, acc);
  if (tap == 0) atomicAdd(g_b1c + c, bacc);
}

}  // namespace tfdl

using namespace tfdl;

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
                                  const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  conv1_fwd_kernel<<<d->B * d->P1h, d->P1w * d->C1, 4 * d->W0 * sizeof(float), (cudaStream_t)stream>>>(
      x, w, b, (__nv_bfloat16*)a1, idx1, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_fwd(const void* a1, const void* w2h, const float* b2, void* a2, uint8_t* idx2,
                                  const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)(128 + d->C2) * 9 * d->C1 * 2;
  if (int rc = set_smem(conv2_fwd_kernel, smem)) return rc;
  int M = d->B * d->P2h * d->P2w * 4;
  conv2_fwd_kernel<<<(M + 127) / 128, 128, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)a1, (const __nv_bfloat16*)w2h, b2, (__nv_bfloat16*)a2, idx2, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_head(float* acc, const float* b1, const float* w2, const float* b2, const int* labels, void* dh,
                             float* g_w2, float* g_b2, float* g_b1, float* metrics, float* step_out,
                             const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = ((size_t)d->B * d->D1 + (size_t)d->NC * d->D1 + (size_t)d->B * d->NC + 2) * sizeof(float);
  if (int rc = set_smem(head_kernel, smem)) return rc;
  head_kernel<<<1, 512, smem, (cudaStream_t)stream>>>(acc, b1, w2, b2, labels, (__nv_bfloat16*)dh, g_w2, g_b2, g_b1,
                                                      metrics, step_out, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_dense1_dgrad(const void* dh, const void* w1h, const void* a2, const uint8_t* idx2, void* dy2,
                                     const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)(128 + d->C2) * d->D1 * 2;
  if (int rc = set_smem(dense1_dgrad_kernel, smem)) return rc;
  dim3 grid(d->P2h * d->P2w, (d->B + 127) / 128);
  dense1_dgrad_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dh, (const __nv_bfloat16*)w1h, (const __nv_bfloat16*)a2, idx2, (__nv_bfloat16*)dy2, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_wgrad(const void* a1, const void* dy2, float* g_w2c, float* g_b2c, const TfdlCnnDims* d,
                                    void* stream) {
  if (int rc = check_dims(*d)) return rc;
  int mtiles = (9 * d->C1 + 127) / 128;
  TFDL_REQUIRE(mtiles * d->C2 <= 512, "cnn: conv2 wgrad accumulators exceed TMEM");
  size_t smem = (size_t)mtiles * 128 * 128 * 2 + (size_t)128 * d->C2 * 2;
  if (int rc = set_smem(conv2_wgrad_kernel, smem)) return rc;
  int P = d->B * 4 * d->P2h * d->P2w;
  conv2_wgrad_kernel<<<(P + 127) / 128, 128, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)a1, (const __nv_bfloat16*)dy2, g_w2c, g_b2c, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_dgrad(const void* dy2, const void* w2h, const void* a1, void* g1, const TfdlCnnDims* d,
                                    void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)(128 + d->C1) * 9 * d->C2 * 2;
  if (int rc = set_smem(conv2_dgrad_kernel, smem)) return rc;
  int M = d->B * d->P1h * d->P1w;
  conv2_dgrad_kernel<<<(M + 127) / 128, 128, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy2, (const __nv_bfloat16*)w2h, (const __nv_bfloat16*)a1, (__nv_bfloat16*)g1, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv1_wgrad(const float* x, const void* g1, const uint8_t* idx1, float* g_w1c, float* g_b1c,
                                    const TfdlCnnDims* d, void* stream) {
  if (int rc = check_dims(*d)) return rc;
  size_t smem = (size_t)d->H0 * d->W0 * 4 + (size_t)d->P1h * d->P1w * d->C1 * 5;
  if (int rc = set_smem(conv1_wgrad_kernel, smem)) return rc;
  conv1_wgrad_kernel<<<d->B, 9 * d->C1, smem, (cudaStream_t)stream>>>(x, (const __nv_bfloat16*)g1, idx1, g_w1c, g_b1c,
                                                                       *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
