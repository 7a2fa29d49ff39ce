// TMA-fed, im2col-free tcgen05 convolution kernels for the reference CNN's second conv layer
// (tf.keras.layers.Conv2D(64, 3) forward / input-gradient / filter-gradient,
// /root/reference/tf_dist_example.py:43).
//
// The gather kernels in cnn.cu materialise an im2col tile in shared memory with cp.async: every input
// pixel is fetched from L2 nine times in 16-byte requests, and that request stream - not the math - was
// ~70% of their run time. These kernels use two facts measured on the hardware (tests/test_probe_gpu.py):
//   (1) a swizzled UMMA operand can be read at ANY row offset by moving the descriptor start address,
//   (2) LBO = 0 on an MN-major operand replicates its first 64 rows into MMA rows 64..127.
// So ONE TMA brings the raw activation tile of an image (rows = pixels, a row = one pixel's channels),
// and the 3x3 filter becomes nine MMAs whose operand descriptor is shifted by (kh*W + kw) rows:
//   conv2_fwd    D[cout][pix]      += W_tap[cout][ci]  * X[pix + kh*13 + kw][ci]        (shift along N)
//   conv2_dgrad  D[pix'][ci]       += DYpad[pix' + (2-kh)*15 + (2-kw)][co] * W_tap[co][ci] (shift along M;
//                                     the zero halo comes from TMA out-of-bounds fill of a 15x15 box)
//   conv2_wgrad  D_tap[co][ci]     += DY[pix][co] * X[pix + kh*13 + kw][ci]             (shift along K;
//                                     the bias gradient is a tenth "tap" against a tile of ones)
// Outputs are computed on the 13-wide (15-wide) input grid; the few junk columns are never stored.
// No thread issues a load for an operand: the SM's LSU is idle until the epilogue.
#include <cuda.h>
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

int make_tmap_bf16_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle);

__device__ __forceinline__ void tma_load_4d_(uint32_t dst, const void* tmap, int c0, int c1, int c2, int c3, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_5d_(uint32_t dst, const void* tmap, int c0, int c1, int c2, int c3, int c4,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
      ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

constexpr int TT = 256;  // threads: warps w and w+4 share TMEM lane quarter w&3

// =============================================================================================
// conv2 forward. CTA = one image. C1 = 32 (64-byte pixel rows, 64B swizzle), C2 = 64 (x2 replicas).
// =============================================================================================
template <int P1W, int P2W>
__global__ void __launch_bounds__(TT, 1)
conv2_fwd_tma_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                     const float* __restrict__ b2, __nv_bfloat16* __restrict__ a2, uint8_t* __restrict__ idx2,
                     TfdlCnnDims d) {
  constexpr int C1 = 32, C2 = 64, XROWS = 176, NPIX = 128;
  static_assert(2 * P1W <= 32 && 2 * P2W <= P1W, "pooled row must fit one 32-column TMEM load");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t bar_tma, bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, n = blockIdx.x;
  const uint32_t x_smem = smem_u32(smem), w_smem = x_smem + XROWS * 64;
  if (warp == 0) tmem_alloc(&slot, 256);  // 128 accumulator columns + slack for the 32-wide epilogue loads
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, XROWS * 64 + 9 * 128 * 64);
    tma_load_2d(x_smem, &tmX, 0, n * d.P1h * P1W, &bar_tma);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {  // filters twice: MMA rows 64..127 replicate rows 0..63
      tma_load_2d(w_smem + tap * 8192, &tmW, tap * C1, 0, &bar_tma);
      tma_load_2d(w_smem + tap * 8192 + 4096, &tmW, tap * C1, 0, &bar_tma);
    }
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, NPIX, false, false);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      const uint32_t shift = (tap / 3) * P1W + tap % 3;
#pragma unroll
      for (int s = 0; s < C1 / 16; ++s) {
        uint64_t ad = umma_smem_desc(w_smem + tap * 8192 + s * 32, 16, 512, SWZ_64B);
        uint64_t bd = umma_smem_desc(x_smem + shift * 64 + s * 32, 16, 512, SWZ_64B);
        umma_bf16(tmem, ad, bd, idesc, (tap | s) != 0);
      }
    }
    umma_commit(&bar_mma);
  }
  const int lq = warp & 3, row = lq * 32 + (int)lane_id(), cout = row & (C2 - 1);
  const float bias = __ldg(b2 + cout);
  const int worker = (row >> 6) * 2 + (warp >> 2);  // 4 workers share the pooled rows of the image
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  const uint32_t t_row = tmem + ((uint32_t)(lq * 32) << 16);
#pragma unroll 1
  for (int ph = worker; ph < d.P2h; ph += 4) {
    uint32_t v[32];
    tmem_ld32(t_row + 2 * ph * P1W, v);  // conv rows 2ph and 2ph+1 of this image (13-wide grid)
    tmem_ld_wait();
#pragma unroll
    for (int pw = 0; pw < P2W; ++pw) {
      const float x0 = __uint_as_float(v[2 * pw]), x1 = __uint_as_float(v[2 * pw + 1]);
      const float x2 = __uint_as_float(v[P1W + 2 * pw]), x3 = __uint_as_float(v[P1W + 2 * pw + 1]);
      float best = x0;
      int arg = 0;
      if (x1 > best) { best = x1; arg = 1; }
      if (x2 > best) { best = x2; arg = 2; }
      if (x3 > best) { best = x3; arg = 3; }
      const size_t o = ((size_t)(n * d.P2h + ph) * P2W + pw) * C2 + cout;
      a2[o] = __float2bfloat16(fmaxf(best + bias, 0.f));
      idx2[o] = (uint8_t)arg;
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// =============================================================================================
// conv2 dgrad. CTA = one image. dy2 is stored on the [B][13][13][64] grid (only [0:10,0:10] non-zero);
// one 4-D TMA box {64, 15, 15, 1} at (0,-2,-2,n) brings the zero-haloed 15x15 grid (OOB fill = 0).
// Weights: nine 2-D boxes W_tap[co][ci] (64-byte rows), used as MN-major 64B-swizzled operands.
// =============================================================================================
template <int P1W>
__global__ void __launch_bounds__(TT, 1)
conv2_dgrad_tma_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmW,
                       const __nv_bfloat16* __restrict__ a1, __nv_bfloat16* __restrict__ g1, TfdlCnnDims d) {
  constexpr int C1 = 32, C2 = 64, GW = P1W + 2, PROWS = 288;
  static_assert((P1W - 1) * GW + P1W - 1 < 256 - 32 + 32, "image does not fit two M tiles");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t bar_tma, bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, n = blockIdx.x;
  const uint32_t p_smem = smem_u32(smem), w_smem = p_smem + PROWS * 128;
  if (warp == 0) tmem_alloc(&slot, 64);
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  pdl_launch_dependents();
  pdl_wait();  // dy2 (dense block) is complete from here on; everything above overlapped its tail
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, GW * GW * 128 + 9 * C2 * C1 * 2);
    tma_load_4d_(p_smem, &tmDY, 0, -2, -2, n, &bar_tma);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) tma_load_2d(w_smem + tap * (C2 * 64), &tmW, tap * C1, 0, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, C1, false, true);
#pragma unroll 1
    for (int t = 0; t < 2; ++t)
#pragma unroll 1
      for (int tap = 0; tap < 9; ++tap) {
        const uint32_t shift = (2 - tap / 3) * GW + (2 - tap % 3);
#pragma unroll
        for (int s = 0; s < C2 / 16; ++s) {
          uint64_t ad = umma_smem_desc(p_smem + (t * 128 + shift) * 128 + s * 32, 16, 1024, SWZ_128B);
          // W_tap[co][ci]: rows = co = the K index, 64 bytes = 32 ci = the N index -> MN-major, 64B swizzle
          uint64_t bd = umma_smem_desc(w_smem + tap * (C2 * 64) + s * 1024, 16, 512, SWZ_64B);
          umma_bf16(tmem + t * C1, ad, bd, idesc, (tap | s) != 0);
        }
      }
    umma_commit(&bar_mma);
  }
  const int lq = warp & 3, t = warp >> 2;
  const int mp = t * 128 + lq * 32 + (int)lane_id();  // pixel index on the 15-wide grid
  const int ih = mp / GW, iw = mp - ih * GW;
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  uint32_t v[32];
  tmem_ld32(tmem + ((uint32_t)(lq * 32) << 16) + t * C1, v);
  tmem_ld_wait();
  if (ih < d.P1h && iw < P1W) {
    const size_t o = ((size_t)(n * d.P1h + ih) * P1W + iw) * C1;
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
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 64);
}

// =============================================================================================
// conv2 wgrad (+ bias grad). CTA = IMG images; K = the 144 pixels (13-wide grid rows 0..11) of an image.
// A = dy2 tile [144][64] (MN-major, 128B swizzle, LBO = 0: rows 64..127 replicate 0..63),
// B_tap = a1 tile shifted by kh*13+kw rows (MN-major, 64B swizzle); nine accumulators of 32 columns
// plus one of 16 columns against a ones tile (bias gradient). fp32 RED (v4) into the gradient arena.
// =============================================================================================
template <int P1W, int IMG>
__global__ void __launch_bounds__(TT, 1)
conv2_wgrad_tma_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                       float* __restrict__ g_w2c, float* __restrict__ g_b2c, TfdlCnnDims d) {
  constexpr int C1 = 32, C2 = 64, KROWS = 144, XROWS = 176, KF = 9 * C1;
  constexpr uint32_t A_BYTES = KROWS * 128, B_BYTES = XROWS * 64, IMG_BYTES = A_BYTES + B_BYTES;
  static_assert(IMG_BYTES % 1024 == 0, "per-image stage must keep 1024-byte alignment");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t bar_tma[IMG], bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, n0 = blockIdx.x * IMG;
  const int nimg = min(IMG, d.B - n0);
  const uint32_t base = smem_u32(smem), ones_smem = base + IMG * IMG_BYTES;
  if (warp == 0) tmem_alloc(&slot, 512);
  if (threadIdx.x == 32) {
    for (int i = 0; i < IMG; ++i) mbar_init(&bar_tma[i], 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  // 16x16 tile of bf16 ones (any layout: all elements equal)
  reinterpret_cast<uint32_t*>(smem + IMG * IMG_BYTES)[threadIdx.x & 127] = 0x3F803F80u;
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    const int img_rows = d.P1h * P1W;
    for (int i = 0; i < nimg; ++i) {
      mbar_expect_tx(&bar_tma[i], IMG_BYTES);
      tma_load_2d(base + i * IMG_BYTES, &tmDY, 0, (n0 + i) * img_rows, &bar_tma[i]);
      tma_load_2d(base + i * IMG_BYTES + A_BYTES, &tmX, 0, (n0 + i) * img_rows, &bar_tma[i]);
    }
    constexpr uint32_t idesc3 = umma_idesc_bf16(128, 3 * C1, true, true), idesc_b = umma_idesc_bf16(128, 16, true, true);
    for (int i = 0; i < nimg; ++i) {
      mbar_wait(&bar_tma[i], 0);
      tc_fence_after_sync();
      const uint32_t a_s = base + i * IMG_BYTES, b_s = a_s + A_BYTES;
      // the three taps of a filter row are the same a1 tile shifted by 0, 1, 2 pixels = three 32-wide MN
      // blocks exactly one 64-byte row apart: LBO = 64 makes them ONE N = 96 operand, so dy2 is read from
      // shared memory three times per K-step instead of nine
#pragma unroll 1
      for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
        for (int s = 0; s < KROWS / 16; ++s) {
          uint64_t ad = umma_smem_desc(a_s + s * 2048, 0, 1024, SWZ_128B);
          uint64_t bd = umma_smem_desc(b_s + (kh * P1W + 16 * s) * 64, 64, 512, SWZ_64B);
          umma_bf16(tmem + kh * 3 * C1, ad, bd, idesc3, (i | s) != 0);
        }
      }
#pragma unroll
      for (int s = 0; s < KROWS / 16; ++s) {
        uint64_t ad = umma_smem_desc(a_s + s * 2048, 0, 1024, SWZ_128B);
        uint64_t bd = umma_smem_desc(ones_smem, 128, 256, SWZ_NONE);
        umma_bf16(tmem + 9 * C1, ad, bd, idesc_b, (i | s) != 0);
      }
    }
    umma_commit(&bar_mma);
  }
  const int lq = warp & 3, row = lq * 32 + (int)lane_id(), co = row & (C2 - 1);
  const int worker = (row >> 6) * 2 + (warp >> 2);
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  const uint32_t t_row = tmem + ((uint32_t)(lq * 32) << 16);
#pragma unroll 1
  for (int tap = worker; tap < 9; tap += 4) {
    uint32_t v[32];
    tmem_ld32(t_row + tap * C1, v);
    tmem_ld_wait();
    float* dst = g_w2c + (size_t)co * KF + tap * C1;
#pragma unroll
    for (int i = 0; i < 32; i += 4)
      red_add_v4(dst + i, __uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]),
                 __uint_as_float(v[i + 3]));
  }
  if (worker == 3) {
    uint32_t vb[16];
    tmem_ld16(t_row + 9 * C1, vb);
    tmem_ld_wait();
    atomicAdd(g_b2c + co, __uint_as_float(vb[0]));
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}


// =============================================================================================
// conv1 forward FUSED into conv2 forward. CTA = one image. While the conv2 filters stream in by TMA the
// CTA computes conv1 + bias + 2x2 max-pool + ReLU for its image on the SIMT cores (Cin = 1, K = 9: not
// tensor-core shaped) and writes the result straight into the 64B-swizzled shared-memory tile that the
// conv2 MMAs read - a1 never makes the round trip through L2 for the forward pass (it is still stored,
// together with the pooling argmax, because the backward pass needs both).
// =============================================================================================
template <int P1W, int P2W>
__global__ void __launch_bounds__(TT, 1)
conv12_fwd_tma_kernel(const float* __restrict__ x, const float* __restrict__ w1, const float* __restrict__ b1,
                      const __grid_constant__ CUtensorMap tmW, const float* __restrict__ b2,
                      __nv_bfloat16* __restrict__ a1, uint8_t* __restrict__ idx1, __nv_bfloat16* __restrict__ a2,
                      uint8_t* __restrict__ idx2, float* __restrict__ clear2, TfdlCnnDims d) {
  constexpr int C1 = 32, C2 = 64, XROWS = 176, NPIX = 128;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t bar_tma, bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n = blockIdx.x;
  const uint32_t x_smem = smem_u32(smem), w_smem = x_smem + XROWS * 64;
  float* xs = reinterpret_cast<float*>(smem + XROWS * 64 + 9 * 128 * 64);  // the input image, fp32
  pdl_launch_dependents();  // the dense block may start its prologue (TMEM alloc, barrier init) now
  if (clear2 && n == 0 && threadIdx.x < 2) clear2[threadIdx.x] = 0.f;
  if (warp == 0) tmem_alloc(&slot, 256);
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  const int npx = d.H0 * d.W0;
  for (int i = threadIdx.x; i < npx / 4; i += TT)
    reinterpret_cast<float4*>(xs)[i] = reinterpret_cast<const float4*>(x + (size_t)n * npx)[i];
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, 9 * 128 * 64);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      tma_load_2d(w_smem + tap * 8192, &tmW, tap * C1, 0, &bar_tma);
      tma_load_2d(w_smem + tap * 8192 + 4096, &tmW, tap * C1, 0, &bar_tma);
    }
  }
  {  // conv1 + pool + relu: lane = channel, warps stride over the pooled pixels of the image
    float wr[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) wr[t] = __ldg(w1 + lane * 9 + t);
    const float bias1 = __ldg(b1 + lane);
    const int np1 = d.P1h * P1W;
    for (int p = warp; p < np1; p += TT / 32) {
      const int ph = p / P1W, pw = p - ph * P1W;
      const float* xp = xs + (2 * ph) * d.W0 + 2 * pw;
      float in[16];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) in[r * 4 + c] = xp[r * d.W0 + c];  // same address in all lanes: broadcast
      float best = -INFINITY;
      int arg = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float acc = bias1;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) acc = fmaf(in[((q >> 1) + kh) * 4 + (q & 1) + kw], wr[kh * 3 + kw], acc);
        if (acc > best) { best = acc; arg = q; }
      }
      const __nv_bfloat16 o = __float2bfloat16(fmaxf(best, 0.f));
      // 64B-swizzled row p of the conv2 operand tile: 16-byte chunk (lane/8) is XORed with (p/2)%4
      *reinterpret_cast<__nv_bfloat16*>(smem + p * 64 + ((((lane >> 3) ^ ((p >> 1) & 3))) << 4) + (lane & 7) * 2) = o;
      const size_t go = ((size_t)n * np1 + p) * C1 + lane;
      a1[go] = o;
      idx1[go] = (uint8_t)arg;
    }
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, NPIX, false, false);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      const uint32_t shift = (tap / 3) * P1W + tap % 3;
#pragma unroll
      for (int s = 0; s < C1 / 16; ++s) {
        uint64_t ad = umma_smem_desc(w_smem + tap * 8192 + s * 32, 16, 512, SWZ_64B);
        uint64_t bd = umma_smem_desc(x_smem + shift * 64 + s * 32, 16, 512, SWZ_64B);
        umma_bf16(tmem, ad, bd, idesc, (tap | s) != 0);
      }
    }
    umma_commit(&bar_mma);
  }
  const int lq = warp & 3, row = lq * 32 + lane, cout = row & (C2 - 1);
  const float bias = __ldg(b2 + cout);
  const int worker = (row >> 6) * 2 + (warp >> 2);
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  const uint32_t t_row = tmem + ((uint32_t)(lq * 32) << 16);
#pragma unroll 1
  for (int ph = worker; ph < d.P2h; ph += 4) {
    uint32_t v[32];
    tmem_ld32(t_row + 2 * ph * P1W, v);
    tmem_ld_wait();
#pragma unroll
    for (int pw = 0; pw < P2W; ++pw) {
      const float x0 = __uint_as_float(v[2 * pw]), x1 = __uint_as_float(v[2 * pw + 1]);
      const float x2 = __uint_as_float(v[P1W + 2 * pw]), x3 = __uint_as_float(v[P1W + 2 * pw + 1]);
      float best = x0;
      int arg = 0;
      if (x1 > best) { best = x1; arg = 1; }
      if (x2 > best) { best = x2; arg = 2; }
      if (x3 > best) { best = x3; arg = 3; }
      const size_t o = ((size_t)(n * d.P2h + ph) * P2W + pw) * C2 + cout;
      a2[o] = __float2bfloat16(fmaxf(best + bias, 0.f));
      idx2[o] = (uint8_t)arg;
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// =============================================================================================
// conv2 dgrad FUSED with conv1 wgrad. CTA = one image. The input gradient g1 of conv2 (after the ReLU mask)
// is never written to global memory: the epilogue drops it into shared memory transposed ([channel][pixel])
// and the same CTA contracts it with the 28x28 input through the pool-1 argmax to produce its image's
// contribution to dW1 / db1 of the first conv (fp32 RED into the gradient arena).
// =============================================================================================
template <int P1W>
__global__ void __launch_bounds__(TT, 1)
conv21_bwd_tma_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmW,
                      const __nv_bfloat16* __restrict__ a1, const uint8_t* __restrict__ idx1,
                      const float* __restrict__ x, float* __restrict__ g_w1c, float* __restrict__ g_b1c,
                      __nv_bfloat16* __restrict__ g1_out /*optional*/, TfdlCnnDims d) {
  constexpr int C1 = 32, C2 = 64, GW = P1W + 2, PROWS = 288;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  __shared__ uint64_t bar_tma, bar_mma;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n = blockIdx.x;
  const int np1 = d.P1h * P1W, ld = np1 | 1, npx = d.H0 * d.W0;
  const uint32_t p_smem = smem_u32(smem), w_smem = p_smem + PROWS * 128;
  float* xs = reinterpret_cast<float*>(smem + PROWS * 128 + 9 * C2 * 64);  // [H0*W0]
  float* gs = xs + npx;                                                       // [C1][ld] masked g1, fp32
  uint8_t* qs = reinterpret_cast<uint8_t*>(gs + C1 * ld);                      // [C1][ld] pool-1 argmax
  if (warp == 0) tmem_alloc(&slot, 64);
  if (threadIdx.x == 32) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  pdl_launch_dependents();
  pdl_wait();  // dy2 (dense block) is complete from here on; everything above overlapped its tail
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_tma, GW * GW * 128 + 9 * C2 * C1 * 2);
    tma_load_4d_(p_smem, &tmDY, 0, -2, -2, n, &bar_tma);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) tma_load_2d(w_smem + tap * (C2 * 64), &tmW, tap * C1, 0, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    constexpr uint32_t idesc = umma_idesc_bf16(128, C1, false, true);
#pragma unroll 1
    for (int t = 0; t < 2; ++t)
#pragma unroll 1
      for (int tap = 0; tap < 9; ++tap) {
        const uint32_t shift = (2 - tap / 3) * GW + (2 - tap % 3);
#pragma unroll
        for (int s = 0; s < C2 / 16; ++s) {
          uint64_t ad = umma_smem_desc(p_smem + (t * 128 + shift) * 128 + s * 32, 16, 1024, SWZ_128B);
          uint64_t bd = umma_smem_desc(w_smem + tap * (C2 * 64) + s * 1024, 16, 512, SWZ_64B);
          umma_bf16(tmem + t * C1, ad, bd, idesc, (tap | s) != 0);
        }
      }
    umma_commit(&bar_mma);
  }
  // while the tensor core works: stage the input image and the pool-1 argmax (transposed) for the conv1 wgrad
  for (int i = threadIdx.x; i < npx / 4; i += TT)
    reinterpret_cast<float4*>(xs)[i] = reinterpret_cast<const float4*>(x + (size_t)n * npx)[i];
  {
    const uint2* isrc = reinterpret_cast<const uint2*>(idx1 + (size_t)n * np1 * C1);
    for (int v8 = threadIdx.x; v8 < np1 * C1 / 8; v8 += TT) {
      const uint2 iv = isrc[v8];
      const int e = v8 * 8, p = e / C1, c0 = e - p * C1;
      const uint32_t iw[2] = {iv.x, iv.y};
#pragma unroll
      for (int k = 0; k < 8; ++k) qs[(c0 + k) * ld + p] = (uint8_t)((iw[k >> 2] >> ((k & 3) * 8)) & 0xFF);
    }
  }
  const int lq = warp & 3, t = warp >> 2;
  const int mp = t * 128 + lq * 32 + lane;  // pixel index on the 15-wide grid
  const int ih = mp / GW, iw = mp - ih * GW;
  const bool valid = ih < d.P1h && iw < P1W;
  const int p1 = ih * P1W + iw;
  uint4 av[4];
  if (valid) {
    const size_t o = ((size_t)n * np1 + p1) * C1;
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) av[k4] = *reinterpret_cast<const uint4*>(a1 + o + k4 * 8);
  }
  mbar_wait(&bar_mma, 0);
  tc_fence_after_sync();
  uint32_t v[32];
  tmem_ld32(tmem + ((uint32_t)(lq * 32) << 16) + t * C1, v);
  tmem_ld_wait();
  if (valid) {
    const uint32_t* aw = reinterpret_cast<const uint32_t*>(av);
    uint32_t outw[16];
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      const uint32_t a_pair = aw[i >> 1];
      const bool on0 = (a_pair & 0x7FFFu) != 0 && !(a_pair & 0x8000u);
      const bool on1 = ((a_pair >> 16) & 0x7FFFu) != 0 && !((a_pair >> 16) & 0x8000u);
      const float g0 = on0 ? __uint_as_float(v[i]) : 0.f, g1v = on1 ? __uint_as_float(v[i + 1]) : 0.f;
      // bf16-round like the unfused path does when it stores g1, then keep fp32 in shared memory
      const uint32_t pk = pack_bf16x2(g0, g1v);
      outw[i >> 1] = pk;
      gs[i * ld + p1] = __uint_as_float(pk << 16);
      gs[(i + 1) * ld + p1] = __uint_as_float(pk & 0xFFFF0000u);
    }
    if (g1_out) {
      const size_t o = ((size_t)n * np1 + p1) * C1;
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4)
        reinterpret_cast<uint4*>(g1_out + o)[k4] = make_uint4(outw[4 * k4], outw[4 * k4 + 1], outw[4 * k4 + 2], outw[4 * k4 + 3]);
    }
  }
  __syncthreads();
  // conv1 wgrad for this image: warp = channel (4 per warp), lanes stride over the pooled pixels
  for (int c = warp; c < C1; c += TT / 32) {
    float acc[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float bacc = 0.f;
    for (int p = lane; p < np1; p += 32) {
      const float g = gs[c * ld + p];
      const int q = qs[c * ld + p];
      const int ph = p / P1W, pw = p - ph * P1W;
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
      for (int k = 0; k < 9; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
      bacc += __shfl_xor_sync(0xffffffffu, bacc, o);
    }
    if (lane < 9) {
      float r = acc[0];
#pragma unroll
      for (int k = 1; k < 9; ++k) r = lane == k ? acc[k] : r;
      atomicAdd(g_w1c + c * 9 + lane, r);
    } else if (lane == 9) {
      atomicAdd(g_b1c + c, bacc);
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 64);
}

}  // namespace tfdl

using namespace tfdl;


// The fast path covers the reference geometry (13x13x32 -> 5x5x64); other sizes use the gather kernels.
extern "C" int tfdl_cnn_tma_supported(const TfdlCnnDims* d) {
  return d->C1 == 32 && d->C2 == 64 && d->P1w == 13 && d->P1h == 13 && d->P2w == 5 && d->P2h == 5;
}

template <class K>
static int set_smem_tma(K kern, size_t bytes) {
  TFDL_REQUIRE(bytes <= 227 * 1024, "cnn_tma: kernel needs %zu bytes of shared memory", bytes);
  TFDL_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

extern "C" int tfdl_cnn_conv2_fwd_tma(const void* a1, const void* w2h, const float* b2, void* a2, uint8_t* idx2,
                                      const TfdlCnnDims* d, void* stream) {
  TFDL_REQUIRE(tfdl_cnn_tma_supported(d), "conv2_fwd_tma: unsupported geometry");
  CUtensorMap tmX, tmW;
  const uint64_t rows = (uint64_t)d->B * d->P1h * d->P1w;
  uint64_t dx[2] = {32, rows}, sx[1] = {64};
  uint32_t bx[2] = {32, 176};
  if (int rc = make_tmap_bf16_nd(&tmX, a1, 2, dx, sx, bx, 2)) return rc;
  uint64_t dw[2] = {288, 64}, sw[1] = {576};
  uint32_t bw[2] = {32, 64};
  if (int rc = make_tmap_bf16_nd(&tmW, w2h, 2, dw, sw, bw, 2)) return rc;
  const size_t smem = 176 * 64 + 9 * 128 * 64 + 1024;
  auto kern = conv2_fwd_tma_kernel<13, 5>;
  if (int rc = set_smem_tma(kern, smem)) return rc;
  kern<<<d->B, TT, smem, (cudaStream_t)stream>>>(tmX, tmW, b2, (__nv_bfloat16*)a2, idx2, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_dgrad_tma(const void* dy2, const void* w2h, const void* a1, void* g1,
                                        const TfdlCnnDims* d, void* stream) {
  TFDL_REQUIRE(tfdl_cnn_tma_supported(d), "conv2_dgrad_tma: unsupported geometry");
  CUtensorMap tmDY, tmW;
  // dy2 on the [B][13][13][64] grid
  uint64_t dd[4] = {64, 13, 13, (uint64_t)d->B}, sd[3] = {128, 13 * 128, 169 * 128};
  uint32_t bd[4] = {64, 15, 15, 1};
  if (int rc = make_tmap_bf16_nd(&tmDY, dy2, 4, dd, sd, bd, 3)) return rc;
  uint64_t dw[2] = {288, 64}, sw[1] = {576};
  uint32_t bw[2] = {32, 64};
  if (int rc = make_tmap_bf16_nd(&tmW, w2h, 2, dw, sw, bw, 2)) return rc;
  const size_t smem = 288 * 128 + 9 * 64 * 32 * 2 + 1024;
  auto kern = conv2_dgrad_tma_kernel<13>;
  if (int rc = set_smem_tma(kern, smem)) return rc;
  kern<<<d->B, TT, smem, (cudaStream_t)stream>>>(tmDY, tmW, (const __nv_bfloat16*)a1, (__nv_bfloat16*)g1, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv2_wgrad_tma(const void* a1, const void* dy2, float* g_w2c, float* g_b2c,
                                        const TfdlCnnDims* d, void* stream) {
  TFDL_REQUIRE(tfdl_cnn_tma_supported(d), "conv2_wgrad_tma: unsupported geometry");
  CUtensorMap tmDY, tmX;
  const uint64_t rows = (uint64_t)d->B * 169;
  uint64_t dd[2] = {64, rows}, sd[1] = {128};
  uint32_t bd[2] = {64, 144};
  if (int rc = make_tmap_bf16_nd(&tmDY, dy2, 2, dd, sd, bd, 3)) return rc;
  uint64_t dx[2] = {32, rows}, sx[1] = {64};
  uint32_t bx[2] = {32, 176};
  if (int rc = make_tmap_bf16_nd(&tmX, a1, 2, dx, sx, bx, 2)) return rc;
  constexpr int IMG = 4;
  const size_t smem = (size_t)IMG * (144 * 128 + 176 * 64) + 512 + 1024;
  auto kern = conv2_wgrad_tma_kernel<13, IMG>;
  if (int rc = set_smem_tma(kern, smem)) return rc;
  kern<<<(d->B + IMG - 1) / IMG, TT, smem, (cudaStream_t)stream>>>(tmDY, tmX, g_w2c, g_b2c, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv12_fwd_tma(const float* x, const float* w1, const float* b1, const void* w2h, const float* b2,
                                       void* a1, uint8_t* idx1, void* a2, uint8_t* idx2, float* clear2,
                                       const TfdlCnnDims* d, void* stream) {
  TFDL_REQUIRE(tfdl_cnn_tma_supported(d) && (d->H0 * d->W0) % 4 == 0, "conv12_fwd_tma: unsupported geometry");
  CUtensorMap tmW;
  uint64_t dw[2] = {288, 64}, sw[1] = {576};
  uint32_t bw[2] = {32, 64};
  if (int rc = make_tmap_bf16_nd(&tmW, w2h, 2, dw, sw, bw, 2)) return rc;
  const size_t smem = 176 * 64 + 9 * 128 * 64 + (size_t)d->H0 * d->W0 * 4 + 1024;
  auto kern = conv12_fwd_tma_kernel<13, 5>;
  if (int rc = set_smem_tma(kern, smem)) return rc;
  kern<<<d->B, TT, smem, (cudaStream_t)stream>>>(x, w1, b1, tmW, b2, (__nv_bfloat16*)a1, idx1, (__nv_bfloat16*)a2, idx2,
                                                clear2, *d);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cnn_conv21_bwd_tma(const void* dy2, const void* w2h, const void* a1, const uint8_t* idx1,
                                       const float* x, float* g_w1c, float* g_b1c, void* g1_out, const TfdlCnnDims* d,
                                       void* stream) {
  TFDL_REQUIRE(tfdl_cnn_tma_supported(d) && (d->H0 * d->W0) % 4 == 0, "conv21_bwd_tma: unsupported geometry");
  CUtensorMap tmDY, tmW;
  uint64_t dd[4] = {64, 13, 13, (uint64_t)d->B}, sd[3] = {128, 13 * 128, 169 * 128};
  uint32_t bd[4] = {64, 15, 15, 1};
  if (int rc = make_tmap_bf16_nd(&tmDY, dy2, 4, dd, sd, bd, 3)) return rc;
  uint64_t dw[2] = {288, 64}, sw[1] = {576};
  uint32_t bw[2] = {32, 64};
  if (int rc = make_tmap_bf16_nd(&tmW, w2h, 2, dw, sw, bw, 2)) return rc;
  const int ld = (d->P1h * d->P1w) | 1;
  const size_t smem = 288 * 128 + 9 * 64 * 64 + (size_t)d->H0 * d->W0 * 4 + (size_t)32 * ld * 5 + 16 + 1024;
  auto kern = conv21_bwd_tma_kernel<13>;
  if (int rc = set_smem_tma(kern, smem)) return rc;
  TFDL_CUDA_OK(launch_kernel_pdl(2, kern, dim3(d->B), dim3(TT), smem, (cudaStream_t)stream, tmDY, tmW,
                                 (const __nv_bfloat16*)a1, idx1, x, g_w1c, g_b1c, (__nv_bfloat16*)g1_out, *d));
  return 0;
}
