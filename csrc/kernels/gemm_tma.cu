// Persistent, warp-specialised, TMA-fed tcgen05 GEMM for layer-sized matrices (the Dense
// layers of the larger configs in BASELINE.json: ResNet-50's classifier, BERT-base's QKV /
// FFN projections; same op as tf.keras.layers.Dense in /root/reference/tf_dist_example.py:46).
//
//   D[M,N] = act(A[M,K] * B[N,K]^T + bias[N])      A, B bf16; each operand either K-major (row-major [rows][K])
//                                                  or MN-major (stored transposed, [K][rows]) - the two layouts a
//                                                  Dense layer's forward / dgrad / wgrad need, with no transposes:
//       forward  y  = x  W^T      A = x  [tok][in]  K-major    B = W  [out][in]  K-major
//       dgrad    dx = dy W        A = dy [tok][out] K-major    B = W  [out][in]  MN-major (K = out)
//       wgrad    dW = dy^T x      A = dy [tok][out] MN-major   B = x  [tok][in]  MN-major (K = tok)
//
// Shape of the kernel (one CTA per SM, 192 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor.2d tiles (128B swizzle) into a STAGES-deep ring
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma (128 x BN x 16) from smem descriptors,
//               accumulating in TMEM; tcgen05.commit releases ring slots / publishes accumulators
//   warps 2-5   epilogue: tcgen05.ld the accumulator (each warp its own 32-lane quarter), bias +
//               activation, then 128-byte rows staged in shared memory (128B swizzle, conflict-free) and
//               written with cp.async.bulk.tensor stores - fully coalesced, clipped at the matrix edge by
//               the tensor map; each warp double-buffers its own 32-row staging tile. (A row-per-thread
//               direct store is kept for outputs whose pitch is not a multiple of 16 bytes.) Two TMEM
//               accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
#include <cuda.h>
#include <stdlib.h>
#include "common.cuh"
#include "gemm_common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

template <int BN, int STAGES>
__global__ void __launch_bounds__(192, 1)
tma_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmD, TmaGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128B swizzle atom (8 rows x 128 B).
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = BM * BK * 2;
  constexpr uint32_t B_BYTES = BN * BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int num_tiles = p.m_tiles * p.n_tiles;
  const int k_blocks = (p.K + BK - 1) / BK;
  const int k_per = (k_blocks + p.ksplit - 1) / p.ksplit, num_units = num_tiles * p.ksplit;
  constexpr uint32_t TMEM_COLS = tmem_cols_for(2 * BN);

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      if (p.tma_store) tma_prefetch_desc(&tmD);
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
      mbar_init(&acc_empty[i], 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t smem_base = smem_u32(smem);

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
        int mt, nt;
        tile_coords(t, p.m_tiles, p.n_tiles, 16, mt, nt);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          if (!p.a_mn) {
            tma_load_2d(sa, &tmA, kb * BK, mt * BM, &full_bar[stage]);
          } else {  // [K rows][64 MN columns] blocks of 8 KB, one per 64 rows of the tile
#pragma unroll
            for (int b = 0; b < BM / 64; ++b) tma_load_2d(sa + b * 8192, &tmA, mt * BM + b * 64, kb * BK, &full_bar[stage]);
          }
          if (!p.b_mn) {
            tma_load_2d(sa + A_BYTES, &tmB, kb * BK, nt * BN, &full_bar[stage]);
          } else {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b)
              tma_load_2d(sa + A_BYTES + b * 8192, &tmB, nt * BN + b * 64, kb * BK, &full_bar[stage]);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(BM, BN, p.a_mn != 0, p.b_mn != 0);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
      for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
        mbar_wait(&acc_empty[as], aphase ^ 1);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major, 128B swizzle: 8-row atoms 1024 B apart; a K step of 16 bf16 = +32 B.
            // MN-major: rows are K, 64-element MN blocks 8 KB apart (LBO); a K step = 16 rows = +2 KB.
            uint64_t ad = p.a_mn ? umma_smem_desc(sa + k * 2048, 8192, 1024, SWZ_128B)
                                 : umma_smem_desc(sa + k * 32, 16, 1024, SWZ_128B);
            uint64_t bd = p.b_mn ? umma_smem_desc(sb + k * 2048, 8192, 1024, SWZ_128B)
                                 : umma_smem_desc(sb + k * 32, 16, 1024, SWZ_128B);
            umma_bf16(d_tmem, ad, bd, idesc, (kb > kb0) || (k != 0));
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&acc_full[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    uint32_t as = 0, aphase = 0, sb = 0;
    const uint32_t stg_base = smem_base + STAGES * STAGE_BYTES + q * 2 * EPI_STAGE_BYTES;
    for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
      int mt, nt;
      tile_coords(t, p.m_tiles, p.n_tiles, 16, mt, nt);
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after_sync();
      epilogue_tile<BN>(p, tmD, tmem_base + as * BN + ((uint32_t)(q * 32) << 16), stg_base, sb, mt * BM + q * 32, nt * BN, kb0 == 0);
      tc_fence_before_sync();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&acc_empty[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane_id() == 0) tma_store_wait_all();  // staging smem must outlive the bulk stores reading it
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------
// CTA-pair kernel (cta_group::2): a cluster of two CTAs on the two SMs of a TPC computes one 256 x 256 tile.
// Each CTA stages its own 128 rows of A and its own 128 rows of B (halving the shared-memory operand traffic
// per SM, which is what bounds the 1-CTA kernel), the leader CTA's elected thread issues
// tcgen05.mma.cta_group::2 (M = 256, N = 256), each CTA keeps its 128 accumulator rows in its own TMEM and runs
// its own epilogue. Barriers: `full` lives in the leader (both CTAs' TMA loads complete bytes on it), `empty`
// and `acc_full` exist in both CTAs and are signalled by multicast tcgen05.commit, the leader's `acc_empty`
// collects the eight epilogue warps of the pair.
// ---------------------------------------------------------------------------------------
template <int PAIR_BN, int STAGES>
__global__ void __launch_bounds__(192, 1)
tma_gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmD, TmaGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int PAIR_HALF = PAIR_BN / 2;  // B rows staged per CTA
  constexpr uint32_t A_BYTES = BM * BK * 2;
  constexpr uint32_t B_BYTES = PAIR_HALF * BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_tiles = p.m_tiles * p.n_tiles;  // 256 x 256 tiles
  const int k_blocks = (p.K + BK - 1) / BK;
  const int k_per = (k_blocks + p.ksplit - 1) / p.ksplit, num_units = num_tiles * p.ksplit;
  constexpr uint32_t TMEM_COLS = 2 * PAIR_BN;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      if (p.tma_store) tma_prefetch_desc(&tmD);
    }
    tmem_alloc_pair(&tmem_slot, TMEM_COLS);
  }
  if (threadIdx.x == 32) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 8);  // four epilogue warps of each CTA of the pair (used in the leader only)
    }
    fence_mbar_init();
  }
  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers are initialised before anything can arrive on them
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t smem_base = smem_u32(smem);

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------------------
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      for (int u = pair; u < num_units; u += num_pairs) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
        int mt, nt;
        tile_coords(t, p.m_tiles, p.n_tiles, 8, mt, nt);
        const int m0 = mt * 256 + (int)rank * BM, n0 = nt * PAIR_BN + (int)rank * PAIR_HALF;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);  // own smem slot free (multicast commit from the leader)
          const uint32_t full_leader = mapa_shared(smem_u32(&full_bar[stage]), 0);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          if (!p.a_mn) {
            tma_load_2d_pair(sa, &tmA, kb * BK, m0, full_leader);
          } else {
#pragma unroll
            for (int b = 0; b < BM / 64; ++b) tma_load_2d_pair(sa + b * 8192, &tmA, m0 + b * 64, kb * BK, full_leader);
          }
          if (!p.b_mn) {
            tma_load_2d_pair(sa + A_BYTES, &tmB, kb * BK, n0, full_leader);
          } else {
#pragma unroll
            for (int b = 0; b < PAIR_HALF / 64; ++b)
              tma_load_2d_pair(sa + A_BYTES + b * 8192, &tmB, n0 + b * 64, kb * BK, full_leader);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) --------------------------
    if (leader && elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(256, PAIR_BN, p.a_mn != 0, p.b_mn != 0);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
      for (int u = pair; u < num_units; u += num_pairs) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
        mbar_wait(&acc_empty[as], aphase ^ 1);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * PAIR_BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            uint64_t ad = p.a_mn ? umma_smem_desc(sa + k * 2048, 8192, 1024, SWZ_128B)
                                 : umma_smem_desc(sa + k * 32, 16, 1024, SWZ_128B);
            uint64_t bd = p.b_mn ? umma_smem_desc(sb + k * 2048, 8192, 1024, SWZ_128B)
                                 : umma_smem_desc(sb + k * 32, 16, 1024, SWZ_128B);
            umma_bf16_pair(d_tmem, ad, bd, idesc, (kb > kb0) || (k != 0));
          }
          umma_commit_pair(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_pair(&acc_full[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ------------------------------ epilogue (both CTAs, own 128 rows) --------------------
    const int q = warp & 3;
    uint32_t as = 0, aphase = 0, sb = 0;
    const uint32_t stg_base = smem_base + STAGES * STAGE_BYTES + q * 2 * EPI_STAGE_BYTES;
    for (int u = pair; u < num_units; u += num_pairs) {
        const int t = u / p.ksplit, kb0 = (u % p.ksplit) * k_per, kb1 = min(k_blocks, kb0 + k_per);
      int mt, nt;
      tile_coords(t, p.m_tiles, p.n_tiles, 8, mt, nt);
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after_sync();
      epilogue_tile<PAIR_BN>(p, tmD, tmem_base + as * PAIR_BN + ((uint32_t)(q * 32) << 16), stg_base, sb,
                             mt * 256 + (int)rank * BM + q * 32, nt * PAIR_BN, kb0 == 0);
      tc_fence_before_sync();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[as]), 0));
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane_id() == 0) tma_store_wait_all();
  }
  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();  // neither CTA may retire (or free TMEM) while the pair's MMAs / remote arrives are in flight
  if (warp == 0) tmem_dealloc_pair(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

// Row-major bf16 matrix [rows][cols] (pitch ld elements) -> tensor map with a (box_rows x 64)
// box and 128-byte swizzle.
int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn fn = get_encode_fn();
  TFDL_REQUIRE(fn, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFDL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// Output matrix [rows][cols] (pitch ld elements, 2- or 4-byte elements): 32-row x 128-byte boxes, 128B swizzle.
int make_tmap_out(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, bool f32) {
  EncodeTiledFn fn = get_encode_fn();
  TFDL_REQUIRE(fn, "cuTensorMapEncodeTiled not available from the driver");
  const uint64_t es = f32 ? 4 : 2;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * es};
  cuuint32_t box[2] = {(cuuint32_t)(128 / es), 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFDL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (output) failed with CUresult %d", (int)r);
  return 0;
}

template <int BN, int STAGES>
static int launch_tma_gemm(const void* A, const void* B, TmaGemmParams p, int lda, int ldb, cudaStream_t s) {
  CUtensorMap tmA, tmB, tmD;
  // K-major operand: matrix [rows][K], box (BK x tile rows). MN-major: matrix [K][rows], box (64 x BK).
  int rc = p.a_mn ? make_tmap_bf16_2d(&tmA, A, p.K, p.M, lda, BK, 64) : make_tmap_bf16_2d(&tmA, A, p.M, p.K, lda, BM, BK);
  if (rc) return rc;
  rc = p.b_mn ? make_tmap_bf16_2d(&tmB, B, p.K, p.N, ldb, BK, 64) : make_tmap_bf16_2d(&tmB, B, p.N, p.K, ldb, BN, BK);
  if (rc) return rc;
  const void* D = p.Dh ? (const void*)p.Dh : (const void*)p.Df;
  const size_t es = p.Dh ? 2 : 4;
  p.tma_store = (!(p.Dh && p.Df) && ((size_t)p.ldd * es) % 16 == 0 && ((uintptr_t)D & 15) == 0) ? 1 : 0;
  if (p.tma_store) {
    rc = make_tmap_out(&tmD, D, p.M, p.N, p.ldd, !p.Dh);
    if (rc) return rc;
  } else {
    tmD = tmA;  // unused by the kernel
  }
  p.m_tiles = (p.M + BM - 1) / BM;
  p.n_tiles = (p.N + BN - 1) / BN;
  constexpr size_t smem = (size_t)STAGES * (BM * BK * 2 + BN * BK * 2) + EPI_SMEM_BYTES + 1024;
  auto kern = tma_gemm_kernel<BN, STAGES>;
  static int sms = 0;  // per template instance: attribute + SM count are set once (all devices of a node are alike)
  if (!sms) {
    TFDL_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, n = 148;
    TFDL_CUDA_OK(cudaGetDevice(&dev));
    TFDL_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    sms = n;
  }
  const int units = p.m_tiles * p.n_tiles * p.ksplit;
  int grid = units < sms ? units : sms;
  kern<<<grid, 192, smem, s>>>(tmA, tmB, tmD, p);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int PAIR_BN, int STAGES>
static int launch_tma_gemm_pair(const void* A, const void* B, TmaGemmParams p, int lda, int ldb, cudaStream_t s) {
  CUtensorMap tmA, tmB, tmD;
  int rc = p.a_mn ? make_tmap_bf16_2d(&tmA, A, p.K, p.M, lda, BK, 64) : make_tmap_bf16_2d(&tmA, A, p.M, p.K, lda, BM, BK);
  if (rc) return rc;
  rc = p.b_mn ? make_tmap_bf16_2d(&tmB, B, p.K, p.N, ldb, BK, 64) : make_tmap_bf16_2d(&tmB, B, p.N, p.K, ldb, PAIR_BN / 2, BK);
  if (rc) return rc;
  const void* D = p.Dh ? (const void*)p.Dh : (const void*)p.Df;
  const size_t es = p.Dh ? 2 : 4;
  p.tma_store = (!(p.Dh && p.Df) && ((size_t)p.ldd * es) % 16 == 0 && ((uintptr_t)D & 15) == 0) ? 1 : 0;
  if (p.tma_store) {
    rc = make_tmap_out(&tmD, D, p.M, p.N, p.ldd, !p.Dh);
    if (rc) return rc;
  } else {
    tmD = tmA;
  }
  p.m_tiles = (p.M + 255) / 256;
  p.n_tiles = (p.N + PAIR_BN - 1) / PAIR_BN;
  constexpr size_t smem = (size_t)STAGES * (BM * BK * 2 + (PAIR_BN / 2) * BK * 2) + EPI_SMEM_BYTES + 1024;
  auto kern = tma_gemm_pair_kernel<PAIR_BN, STAGES>;
  static int sms = 0;
  if (!sms) {
    TFDL_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, n = 148;
    TFDL_CUDA_OK(cudaGetDevice(&dev));
    TFDL_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    sms = n;
  }
  const int units = p.m_tiles * p.n_tiles * p.ksplit;
  int pairs = units < sms / 2 ? units : sms / 2;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(192);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TFDL_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmD, p));
  return 0;
}

}  // namespace tfdl

extern "C" int tfdl_gemm_bf16_tma(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M,
                                  int N, int K, int lda, int ldb, int ldd, int act, void* stream) {
  return tfdl_gemm_bf16_tma_ex(A, B, Df, Dh, bias, M, N, K, lda, ldb, ldd, 0, 0, act, stream);
}

extern "C" int tfdl_gemm_bf16_tma_ex(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M,
                                     int N, int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int act,
                                     void* stream) {
  return tfdl_gemm_bf16_tma_splitk(A, B, Df, Dh, bias, M, N, K, lda, ldb, ldd, a_mn, b_mn, act, 1, stream);
}

extern "C" int tfdl_gemm_bf16_tma_splitk(const void* A, const void* B, float* Df, void* Dh, const float* bias, int M,
                                         int N, int K, int lda, int ldb, int ldd, int a_mn, int b_mn, int act,
                                         int ksplit, void* stream) {
  using namespace tfdl;
  TFDL_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_tma: empty problem");
  TFDL_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_tma: lda, ldb must be multiples of 8");
  TFDL_REQUIRE((a_mn ? M : K) % 8 == 0 && (b_mn ? N : K) % 8 == 0,
               "gemm_tma: the contiguous dimension of each operand must be a multiple of 8");
  TFDL_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm_tma: A/B must be 16-byte aligned");
  TmaGemmParams p{};
  p.Df = Df; p.Dh = (__nv_bfloat16*)Dh; p.bias = bias; p.M = M; p.N = N; p.K = K; p.ldd = ldd; p.act = act;
  p.a_mn = a_mn; p.b_mn = b_mn;
  const int k_blocks = (K + BK - 1) / BK;
  if (ksplit < 1) ksplit = 1;
  if (ksplit > k_blocks) ksplit = k_blocks;
  if (ksplit > 1) {
    TFDL_REQUIRE(Df && !Dh && act == 0 && (ldd % 4) == 0 && ((uintptr_t)Df & 15) == 0,
                 "gemm_tma split-K: needs a zero-initialised fp32 output with a 16-byte-multiple pitch and no activation");
    const int k_per = (k_blocks + ksplit - 1) / ksplit;
    ksplit = (k_blocks + k_per - 1) / k_per;  // no empty split
  }
  p.ksplit = ksplit;
  cudaStream_t s = (cudaStream_t)stream;
  // Tile choice: 128x256 when there are enough tiles to fill the chip, else 128x128 / 128x64
  long tiles256 = (long)((M + 127) / 128) * ((N + 255) / 256) * ksplit;
  static int pair_mode = -1;  // TFDL_GEMM_PAIR=0 keeps every problem on the 1-CTA kernel
  if (pair_mode < 0) {
    const char* e = getenv("TFDL_GEMM_PAIR");
    pair_mode = e ? atoi(e) : 1;
  }
  if (pair_mode && N >= 128 && M >= 256 && tiles256 >= 60) {
    // 256 x 256 or 256 x 128 tiles per CTA pair: pick the one with the cheaper last wave (74 pairs per wave)
    const long mt = (M + 255) / 256, pairs = 74;
    const long c256 = ((mt * ((N + 255) / 256) * ksplit + pairs - 1) / pairs) * 256;
    const long c128 = ((mt * ((N + 127) / 128) * ksplit + pairs - 1) / pairs) * 205;  // measured: a 128-wide tile costs ~0.8 of a 256-wide one
    if (c128 < c256) return launch_tma_gemm_pair<128, 8>(A, B, p, lda, ldb, s);
    return launch_tma_gemm_pair<256, 6>(A, B, p, lda, ldb, s);
  }
  if (N >= 256 && tiles256 >= 120) return launch_tma_gemm<256, 4>(A, B, p, lda, ldb, s);
  if (N > 64) return launch_tma_gemm<128, 6>(A, B, p, lda, ldb, s);
  return launch_tma_gemm<64, 8>(A, B, p, lda, ldb, s);
}
