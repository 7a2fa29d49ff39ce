// Implicit-GEMM Conv2D on tcgen05 for NHWC bf16 activations: forward, input gradient and weight gradient of
// tf.keras.layers.Conv2D (/root/reference/tf_dist_example.py:41,43) for general filter sizes, strides and
// paddings - the convolutions of the reference CNN with other filter counts / 'same' padding and every
// convolution of the Keras ResNet-50 config (BASELINE.json config 3) except its 3-channel stem.
//
// There is no im2col buffer. The activation operand is fetched by TMA in *im2col mode*
// (cp.async.bulk.tensor.4d...im2col): one instruction brings a [pixels x 64 channels] tile for ONE filter
// tap (r, s), walking the output pixels of the tile across rows and images, applying the stride, and
// zero-filling whatever falls into the padding halo. The tile lands in shared memory in the same
// 128B-swizzled layout a plain 2-D TMA tile has, so the tcgen05 mainloop is the GEMM's:
//
//   forward   y[m][k]   = sum_{r,s,c} x(m; r,s)[c] * W[k][r][s][c]       A = im2col(x) K-major, B = W K-major
//   dgrad     dx[m][c]  = sum_{r,s,k} dy(m; r',s')[k] * W[k][r][s][c]    A = im2col(dy), flipped taps, B = W MN-major
//   wgrad     dW[k][r][s][c] = sum_m dy[m][k] * x(m; r,s)[c]            A = dy MN-major, B = im2col(x) MN-major
//
// (m runs over output pixels n*P*Q + p*Q + q.) The K loop of forward / dgrad runs over (tap, 64-channel block); a
// channel count that is not a multiple of 64 is padded by the TMA unit's out-of-bounds zero fill on the activation
// side, which also neutralises whatever the weight tile reads past its tap. The weight gradient is split over the
// pixel dimension across CTAs and accumulated with TMA reduce-add into a zero-initialised fp32 buffer in the storage
// layout of the kernel variable (OHWI), i.e. straight into the gradient arena.
//
// Kernel shape: as gemm_tma.cu (warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue, STAGES-deep mbarrier
// ring, two TMEM accumulator stages, persistent over tiles).
#include <cuda.h>
#include <stdlib.h>
#include "common.cuh"
#include "gemm_common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

struct ConvGeom {
  int P, Q;            // output rows / columns of the im2col walk
  int PQ;
  int lower_h, lower_w;  // input coordinate of tap (0,0) for output pixel (0,0)
  int stride_h, stride_w;
  int R, S;            // filter taps
  int Cin_k;           // channels of the im2col tensor (K extent per tap for forward / dgrad)
  int cblocks;         // ceil(Cin_k / 64)
  int w_tap_stride;    // columns per tap in the weight matrix [Cout][R*S*Cin] (= Cin)
  int flip;            // dgrad: weight tap = (R-1-r, S-1-s)
  int mode;            // 0 forward, 1 dgrad, 2 wgrad
};

// TMA im2col load of a [pixels x 64 channels] tile for tap offsets (off_w, off_h) starting at base pixel (w, h, n).
__device__ __forceinline__ void tma_load_im2col(uint32_t dst_smem, const void* tmap, int32_t c, int32_t w, int32_t h,
                                                int32_t n, uint16_t off_w, uint16_t off_h, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], "
      "[%6], {%7, %8};" ::"r"(dst_smem),
      "l"(tmap), "r"(c), "r"(w), "r"(h), "r"(n), "r"(smem_u32(bar)), "h"(off_w), "h"(off_h)
      : "memory");
}

// linear output pixel -> TMA base coordinates
__device__ __forceinline__ void pixel_coords(const ConvGeom& g, int m, int32_t& w, int32_t& h, int32_t& n) {
  n = m / g.PQ;
  const int rem = m - n * g.PQ;
  const int p = rem / g.Q, q = rem - p * g.Q;
  w = g.lower_w + q * g.stride_w;
  h = g.lower_h + p * g.stride_h;
}

// 224 threads: warp 0 = activation (A) producer, warp 6 = weight / second-operand (B) producer, warp 1 = MMA issuer,
// warps 2-5 = epilogue. With 64-channel layers a K block is only 128 x BN x 64 of MMA work, and ONE thread issuing both
// operands' TMA loads (~70 dependent instructions per block) was the kernel's limiter, not the tensor core or memory.
template <int BN, int STAGES>
__global__ void __launch_bounds__(224, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmD, TmaGemmParams p, ConvGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = BM * BK * 2;
  constexpr uint32_t B_BYTES = BN * BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int UNITS = BN / 64;  // 64-column blocks of the B tile (wgrad: one (tap, channel block) each)
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int num_tiles = p.m_tiles * p.n_tiles;
  const bool wgrad = g.mode == 2;
  const int n_units = g.R * g.S * g.cblocks;                      // (tap, channel block) pairs
  const int k_blocks = wgrad ? (p.K + BK - 1) / BK : n_units;     // wgrad: K = output pixels
  const int k_per = (k_blocks + p.ksplit - 1) / p.ksplit, num_units = num_tiles * p.ksplit;
  constexpr uint32_t TMEM_COLS = tmem_cols_for(2 * BN);

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      tma_prefetch_desc(&tmD);
    }
    tmem_alloc(&tmem_slot, TMEM_COLS);
  }
  if (threadIdx.x == 32) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4);
    }
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t smem_base = smem_u32(smem);

  if (warp == 0 || warp == 6) {
    // ------------------------------ TMA producers ------------------------------
    // warp 0 posts the stage's byte count and loads operand A, warp 6 loads operand B; both wait for the slot to be free.
    const bool prodA = warp == 0;
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
        int mt, nt;
        tile_coords(t, p.m_tiles, p.n_tiles, 16, mt, nt);
        // No division inside the K loop: tap / channel-block / pixel counters are decoded once per work unit and then
        // advanced incrementally.
        int32_t w0 = 0, h0 = 0, n0 = 0;
        if (!wgrad) {
          if (prodA) pixel_coords(g, mt * BM, w0, h0, n0);
          int tap = kb0 / g.cblocks, cb = kb0 - tap * g.cblocks;
          int r = tap / g.S, sx = tap - r * g.S;
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + A_BYTES;
            if (prodA) {
              mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
              tma_load_im2col(sa, &tmA, cb * 64, w0, h0, n0, (uint16_t)sx, (uint16_t)r, &full_bar[stage]);
            } else if (g.mode == 0) {
              tma_load_2d(sb, &tmB, tap * g.w_tap_stride + cb * 64, nt * BN, &full_bar[stage]);
            } else {
              const int wt = (g.R - 1 - r) * g.S + (g.S - 1 - sx);
#pragma unroll
              for (int b = 0; b < UNITS; ++b)
                tma_load_2d(sb + b * 8192, &tmB, wt * g.w_tap_stride + nt * BN + b * 64, cb * 64, &full_bar[stage]);
            }
            if (++cb == g.cblocks) {
              cb = 0;
              ++tap;
              if (++sx == g.S) { sx = 0; ++r; }
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        } else {
          const int valid = min(UNITS, n_units - nt * UNITS);
          int u_r[UNITS], u_s[UNITS], u_c[UNITS];
#pragma unroll
          for (int b = 0; b < UNITS; ++b) {
            const int un = min(nt * UNITS + b, n_units - 1);
            const int tap = un / g.cblocks;
            u_c[b] = (un - tap * g.cblocks) * 64;
            u_r[b] = tap / g.S;
            u_s[b] = tap - u_r[b] * g.S;
          }
          // output pixel (n, p, q) of the first row of K block kb0, advanced by BK pixels per block
          int pn = (kb0 * BK) / g.PQ, rem = kb0 * BK - pn * g.PQ;
          int pp = rem / g.Q, pq = rem - pp * g.Q;
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + A_BYTES;
            if (prodA) {
              mbar_expect_tx(&full_bar[stage], A_BYTES + (uint32_t)valid * 8192u);
#pragma unroll
              for (int b = 0; b < BM / 64; ++b) tma_load_2d(sa + b * 8192, &tmA, mt * BM + b * 64, kb * BK, &full_bar[stage]);
            } else {
              w0 = g.lower_w + pq * g.stride_w;
              h0 = g.lower_h + pp * g.stride_h;
#pragma unroll
              for (int b = 0; b < UNITS; ++b)
                if (b < valid)
                  tma_load_im2col(sb + b * 8192, &tmB, u_c[b], w0, h0, pn, (uint16_t)u_s[b], (uint16_t)u_r[b], &full_bar[stage]);
              pq += BK;
              while (pq >= g.Q) {
                pq -= g.Q;
                if (++pp == g.P) { pp = 0; ++pn; }
              }
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (elect_one()) {
      const bool a_mn = wgrad, b_mn = g.mode != 0;
      const uint32_t idesc = umma_idesc_bf16(BM, BN, a_mn, b_mn);
      // Descriptors differ from block to block only in their 14-bit start-address field: build them once for stage 0 and
      // add (byte offset >> 4) per stage and per K step (the sum stays below 2^14: shared memory is < 256 KB).
      const uint64_t a_desc0 = a_mn ? umma_smem_desc(smem_base, 8192, 1024, SWZ_128B) : umma_smem_desc(smem_base, 16, 1024, SWZ_128B);
      const uint64_t b_desc0 = b_mn ? umma_smem_desc(smem_base + A_BYTES, 8192, 1024, SWZ_128B)
                                    : umma_smem_desc(smem_base + A_BYTES, 16, 1024, SWZ_128B);
      const uint32_t a_step = a_mn ? (2048u >> 4) : (32u >> 4), b_step = b_mn ? (2048u >> 4) : (32u >> 4);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
      for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
        const int kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
        mbar_wait(&acc_empty[as], aphase ^ 1);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t soff = (stage * STAGE_BYTES) >> 4;
          const uint64_t ad = a_desc0 + soff, bd = b_desc0 + soff;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) umma_bf16(d_tmem, ad + k * a_step, bd + k * b_step, idesc, (kb > kb0) || (k != 0));
          umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&acc_full[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    const int q = warp & 3;
    uint32_t as = 0, aphase = 0, sbuf_i = 0;
    const uint32_t stg_base = smem_base + STAGES * STAGE_BYTES + q * 2 * EPI_STAGE_BYTES;
    for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
      const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per;
      int mt, nt;
      tile_coords(t, p.m_tiles, p.n_tiles, 16, mt, nt);
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after_sync();
      const uint32_t t_row = tmem_base + as * BN + ((uint32_t)(q * 32) << 16);
      const int row0 = mt * BM + q * 32;
      if (!wgrad) {
        epilogue_tile<BN>(p, tmD, t_row, stg_base, sbuf_i, row0, nt * BN, kb0 == 0);
      } else {
        // fp32 partial sums of dW: 32-column chunks, each reduce-added at the column of its (tap, channel block)
        const uint32_t lane = lane_id();
#pragma unroll 1
        for (int c = 0; c < BN && row0 < p.M; c += 32) {
          const int un = nt * UNITS + c / 64;
          if (un >= n_units) break;
          const int tap = un / g.cblocks, cb = un - tap * g.cblocks;
          const int col0 = tap * g.w_tap_stride + cb * 64 + (c & 63);
          if (cb * 64 + (c & 63) >= g.w_tap_stride) continue;  // channel block tail beyond Cin: all zeros
          uint32_t v[32];
          tmem_ld32(t_row + c, v);
          tmem_ld_wait();
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          const uint32_t sbuf = stg_base + sbuf_i * EPI_STAGE_BYTES;
          const uint32_t srow = sbuf + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(srow + ((uint32_t)(j ^ (lane & 7)) << 4), v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_reduce_add_2d(&tmD, sbuf, col0, row0);
            tma_store_commit();
          }
          sbuf_i ^= 1;
        }
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&acc_empty[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane_id() == 0) tma_store_wait_all();
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);
static EncodeIm2colFn get_im2col_fn() {
  static EncodeIm2colFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeIm2colFn>(sym);
  }
  return fn;
}

// NHWC bf16 tensor [N][H][W][C] -> im2col tensor map: `pixels` output positions x 64 channels per load, 128B swizzle.
// lower/upper: offsets of the box of tap-(0,0) positions from the tensor's near / far corner, per (w, h).
static int make_tmap_im2col(CUtensorMap* out, const void* base, int N, int H, int W, int C, int lower_w, int lower_h,
                            int upper_w, int upper_h, int stride_w, int stride_h, int pixels) {
  EncodeIm2colFn fn = get_im2col_fn();
  TFDL_REQUIRE(fn, "cuTensorMapEncodeIm2col not available from the driver");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lower[2] = {lower_w, lower_h}, upper[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride_w, (cuuint32_t)stride_h, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper,
                  /*channelsPerPixel=*/64, /*pixelsPerColumn=*/(cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFDL_REQUIRE(r == CUDA_SUCCESS,
               "cuTensorMapEncodeIm2col failed with CUresult %d (N=%d H=%d W=%d C=%d lower=(%d,%d) upper=(%d,%d) "
               "stride=(%d,%d))", (int)r, N, H, W, C, lower_w, lower_h, upper_w, upper_h, stride_w, stride_h);
  // Drivers up to 13.1 mis-encode im2col maps of tensors smaller than 128 KiB (bit 21 of the second descriptor word);
  // CUTLASS clears the bit by hand (copy_traits_sm90_im2col.hpp) and so do we.
  int drv = 0;
  if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010 && (size_t)N * H * W * C * 2 < 131072)
    reinterpret_cast<uint64_t*>(out)[1] &= ~(1ull << 21);
  return 0;
}

struct ConvShape {
  int N, H, W, C, K, R, S, sh, sw, pt, pl, pb, pr, P, Q;
};
static int check_shape(const ConvShape& c, const char* who) {
  TFDL_REQUIRE(c.N > 0 && c.H > 0 && c.W > 0 && c.C > 0 && c.K > 0 && c.R > 0 && c.S > 0 && c.sh > 0 && c.sw > 0,
               "%s: empty problem", who);
  TFDL_REQUIRE(c.C % 8 == 0 && c.K % 8 == 0, "%s: channel counts must be multiples of 8 (C=%d, K=%d)", who, c.C, c.K);
  TFDL_REQUIRE(c.P == (c.H + c.pt + c.pb - c.R) / c.sh + 1 && c.Q == (c.W + c.pl + c.pr - c.S) / c.sw + 1 && c.P > 0 &&
               c.Q > 0, "%s: output size %dx%d does not match input %dx%d, filter %dx%d, stride (%d,%d), padding "
               "(%d,%d,%d,%d)", who, c.P, c.Q, c.H, c.W, c.R, c.S, c.sh, c.sw, c.pt, c.pl, c.pb, c.pr);
  TFDL_REQUIRE(c.R <= 16 && c.S <= 16 && c.pt < 128 && c.pl < 128, "%s: filter / padding too large", who);
  return 0;
}

template <int BN, int STAGES>
static int launch_conv(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD, TmaGemmParams p,
                       const ConvGeom& g, cudaStream_t s) {
  constexpr size_t smem = (size_t)STAGES * (BM * BK * 2 + BN * BK * 2) + EPI_SMEM_BYTES + 1024;
  auto kern = conv_igemm_kernel<BN, STAGES>;
  static int sms = 0;
  if (!sms) {
    TFDL_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, n = 148;
    TFDL_CUDA_OK(cudaGetDevice(&dev));
    TFDL_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    sms = n;
  }
  const int units = p.m_tiles * p.n_tiles * p.ksplit;
  const int grid = units < sms ? units : sms;
  kern<<<grid, 224, smem, s>>>(tmA, tmB, tmD, p, g);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace tfdl

using namespace tfdl;

static void fill_geom(ConvGeom& g, int P, int Q, int lower_h, int lower_w, int sh, int sw, int R, int S, int Ck, int wtap,
                      int mode) {
  g.P = P; g.Q = Q; g.PQ = P * Q; g.lower_h = lower_h; g.lower_w = lower_w; g.stride_h = sh; g.stride_w = sw;
  g.R = R; g.S = S; g.Cin_k = Ck; g.cblocks = (Ck + 63) / 64; g.w_tap_stride = wtap; g.flip = mode == 1; g.mode = mode;
}

extern "C" int tfdl_conv2d_supported(int C, int K) { return (C % 8 == 0 && K % 8 == 0 && C >= 8 && K >= 8) ? 1 : 0; }

extern "C" int tfdl_conv2d_fwd(const void* x, const void* w, const float* bias, void* y, int N, int H, int W, int C, int K,
                               int R, int S, int stride_h, int stride_w, int pad_t, int pad_l, int pad_b, int pad_r,
                               int act, void* stream) {
  ConvShape c{N, H, W, C, K, R, S, stride_h, stride_w, pad_t, pad_l, pad_b, pad_r,
              (H + pad_t + pad_b - R) / stride_h + 1, (W + pad_l + pad_r - S) / stride_w + 1};
  if (int rc = check_shape(c, "conv2d_fwd")) return rc;
  TFDL_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)y & 15) == 0,
               "conv2d_fwd: x, w, y must be 16-byte aligned");
  CUtensorMap tmA, tmB, tmD;
  if (int rc = make_tmap_im2col(&tmA, x, N, H, W, C, -pad_l, -pad_t, pad_r - (S - 1), pad_b - (R - 1), stride_w, stride_h, BM))
    return rc;
  const int BN = K <= 64 ? 64 : ((K >= 256 && (long)((N * c.P * c.Q + 127) / 128) * ((K + 255) / 256) >= 120) ? 256 : 128);
  if (int rc = make_tmap_bf16_2d(&tmB, w, K, (uint64_t)R * S * C, (uint64_t)R * S * C, BN, BK)) return rc;
  const long M = (long)N * c.P * c.Q;
  TFDL_REQUIRE(M < (1l << 31), "conv2d_fwd: too many output pixels");
  if (int rc = make_tmap_out(&tmD, y, M, K, K, false)) return rc;
  TmaGemmParams p{};
  p.Dh = (__nv_bfloat16*)y; p.bias = bias; p.M = (int)M; p.N = K; p.K = R * S * C; p.ldd = K; p.act = act;
  p.ksplit = 1; p.tma_store = 1;
  p.m_tiles = (p.M + BM - 1) / BM; p.n_tiles = (K + BN - 1) / BN;
  ConvGeom g;
  fill_geom(g, c.P, c.Q, -pad_t, -pad_l, stride_h, stride_w, R, S, C, C, 0);
  cudaStream_t s = (cudaStream_t)stream;
  if (BN == 64) return launch_conv<64, 8>(tmA, tmB, tmD, p, g, s);
  if (BN == 128) return launch_conv<128, 6>(tmA, tmB, tmD, p, g, s);
  return launch_conv<256, 4>(tmA, tmB, tmD, p, g, s);
}

// dx[N][H][W][C] = sum over taps and output channels of dy * W (stride 1 only: with a stride the input gradient is a
// strided scatter; ops/conv.py runs those layers as per-phase sub-convolutions or GEMM + scatter).
extern "C" int tfdl_conv2d_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R, int S,
                                 int pad_t, int pad_l, int pad_b, int pad_r, void* stream) {
  ConvShape c{N, H, W, C, K, R, S, 1, 1, pad_t, pad_l, pad_b, pad_r, H + pad_t + pad_b - R + 1, W + pad_l + pad_r - S + 1};
  if (int rc = check_shape(c, "conv2d_dgrad")) return rc;
  TFDL_REQUIRE(pad_t <= R - 1 && pad_l <= S - 1, "conv2d_dgrad: padding larger than the filter reach");
  TFDL_REQUIRE(((uintptr_t)dy & 15) == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)dx & 15) == 0,
               "conv2d_dgrad: dy, w, dx must be 16-byte aligned");
  const int lw = -(S - 1 - pad_l), lh = -(R - 1 - pad_t);
  CUtensorMap tmA, tmB, tmD;
  if (int rc = make_tmap_im2col(&tmA, dy, N, c.P, c.Q, K, lw, lh, W - c.Q + lw, H - c.P + lh, 1, 1, BM)) return rc;
  if (int rc = make_tmap_bf16_2d(&tmB, w, K, (uint64_t)R * S * C, (uint64_t)R * S * C, BK, 64)) return rc;
  const long M = (long)N * H * W;
  TFDL_REQUIRE(M < (1l << 31), "conv2d_dgrad: too many pixels");
  if (int rc = make_tmap_out(&tmD, dx, M, C, C, false)) return rc;
  const int BN = C <= 64 ? 64 : ((C >= 256 && (long)((M + 127) / 128) * ((C + 255) / 256) >= 120) ? 256 : 128);
  TmaGemmParams p{};
  p.Dh = (__nv_bfloat16*)dx; p.M = (int)M; p.N = C; p.K = R * S * K; p.ldd = C; p.act = 0; p.ksplit = 1; p.tma_store = 1;
  p.m_tiles = (p.M + BM - 1) / BM; p.n_tiles = (C + BN - 1) / BN;
  ConvGeom g;
  fill_geom(g, H, W, lh, lw, 1, 1, R, S, K, C, 1);
  cudaStream_t s = (cudaStream_t)stream;
  if (BN == 64) return launch_conv<64, 8>(tmA, tmB, tmD, p, g, s);
  if (BN == 128) return launch_conv<128, 6>(tmA, tmB, tmD, p, g, s);
  return launch_conv<256, 4>(tmA, tmB, tmD, p, g, s);
}

// dw[K][R][S][C] (fp32, ACCUMULATED: zero it first - or point it at a cleared gradient arena) += dy^T * im2col(x).
extern "C" int tfdl_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int K, int R, int S,
                                 int stride_h, int stride_w, int pad_t, int pad_l, int pad_b, int pad_r, void* stream) {
  ConvShape c{N, H, W, C, K, R, S, stride_h, stride_w, pad_t, pad_l, pad_b, pad_r,
              (H + pad_t + pad_b - R) / stride_h + 1, (W + pad_l + pad_r - S) / stride_w + 1};
  if (int rc = check_shape(c, "conv2d_wgrad")) return rc;
  TFDL_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)dy & 15) == 0 && ((uintptr_t)dw & 15) == 0 && (R * S * C) % 4 == 0,
               "conv2d_wgrad: x, dy, dw must be 16-byte aligned");
  const long M = (long)N * c.P * c.Q;
  TFDL_REQUIRE(M < (1l << 31), "conv2d_wgrad: too many output pixels");
  CUtensorMap tmA, tmB, tmD;
  if (int rc = make_tmap_bf16_2d(&tmA, dy, M, K, K, BK, 64)) return rc;
  if (int rc = make_tmap_im2col(&tmB, x, N, H, W, C, -pad_l, -pad_t, pad_r - (S - 1), pad_b - (R - 1), stride_w, stride_h, BK))
    return rc;
  if (int rc = make_tmap_out(&tmD, dw, K, (uint64_t)R * S * C, (uint64_t)R * S * C, true)) return rc;
  ConvGeom g;
  fill_geom(g, c.P, c.Q, -pad_t, -pad_l, stride_h, stride_w, R, S, C, C, 2);
  const int n_units = R * S * g.cblocks;
  constexpr int BN = 128;
  TmaGemmParams p{};
  p.Df = dw; p.M = K; p.N = n_units * 64; p.K = (int)M; p.ldd = R * S * C; p.act = 0; p.tma_store = 1;
  p.m_tiles = (K + BM - 1) / BM; p.n_tiles = (n_units + BN / 64 - 1) / (BN / 64);
  const int k_blocks = (int)((M + BK - 1) / BK), tiles = p.m_tiles * p.n_tiles;
  // Split factor: the kernel's time is (waves of work units over 148 SMs) x (K blocks per unit + a fixed per-unit cost of
  // about six blocks for the accumulator hand-over and the reduce-add epilogue); take the split that minimises it, with
  // at least four pixel blocks per unit.
  int ks = 1;
  long best = -1;
  for (int c = 1; c <= 64 && c <= (k_blocks + 3) / 4; ++c) {
    const long waves = ((long)tiles * c + 147) / 148;
    const long cost = waves * ((k_blocks + c - 1) / c + 6);
    if (best < 0 || cost < best) { best = cost; ks = c; }
  }
  const int k_per = (k_blocks + ks - 1) / ks;
  p.ksplit = (k_blocks + k_per - 1) / k_per;        // no empty split
  return launch_conv<BN, 6>(tmA, tmB, tmD, p, g, (cudaStream_t)stream);
}
