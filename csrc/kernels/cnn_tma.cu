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
