// Fused "dense block" of the reference CNN: everything between the second pooling layer and the
// gradient w.r.t. the conv2 output, in ONE kernel (tf_dist_example.py:45-50 forward and backward):
//
//   Flatten -> Dense(128, relu) -> Dense(10) -> softmax cross-entropy -> dLogits -> dDense(10) -> dDense(128)
//   -> ReLU / max-pool backward, plus the weight gradients dW1, db1, dW2, db2, the loss and the accuracy.
//
// The unfused path runs this as four kernels (split-K GEMM, head, dgrad GEMM, wgrad GEMM) with three
// round trips through L2; at batch 64 each of them is latency-bound. Here CTA p (one per pooled pixel,
// 25 CTAs) owns the 64-feature slice p of the flatten axis, which is at the same time
//   * one split-K chunk of the forward GEMM,          * one N tile of the input-gradient GEMM,
//   * one N tile of the weight-gradient GEMM,
// so the W1 slice and the a2 slice are brought by TMA ONCE and reused by three MMAs under different
// descriptors (K-major for the forward, MN-major for the two backward GEMMs: same bytes). The 64x10
// "head" is also done on the tensor core (logits = H*W2^T and dH = dLogits*W2 are 128x16x128 / 128x128x16
// MMAs), every CTA recomputes it redundantly (it is ~1 MFLOP), and column sums (db1) are an MMA against
// a tile of ones. One grid-wide barrier (25 co-resident CTAs, arrival counter in global memory)
// separates the split-K reduction from its consumers.
//
// Shared-memory tiles (all 128-byte rows, 128B swizzle unless noted):
//   Wt [128 units][64 feats]   TMA   fwd: A K-major            dgrad: B MN-major (k = units)
//   At [128 smpls][64 feats]   TMA   fwd: B K-major            wgrad: B MN-major (k = samples)
//   H  2x[128 smpls][64 units] thr   logits: A K-major         dW2:   A MN-major (k = samples)
//   DH 2x[128 smpls][64 units] thr   dgrad: A K-major          wgrad/db1: A MN-major (k = samples)
//   W2 2x[16 cls][64 units]    TMA   logits: B K-major         dH: B MN-major (k = classes)
//   DL [128 smpls][16 cls]     thr   (no swizzle) dH: A K-major, dW2: B MN-major
#include <cuda.h>
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

int make_tmap_bf16_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle);

namespace {

#ifdef TFDL_PROFILE_PHASES
__device__ long long g_phase_db[64][16];
#define DBPHASE(slot) do { if (threadIdx.x == 0 && blockIdx.x < 64) g_phase_db[blockIdx.x][slot] = clock64(); } while (0)
#else
#define DBPHASE(slot) do { } while (0)
#endif

constexpr int DB_THREADS = 256;
constexpr int D1 = 128, C2 = 64;
// smem byte offsets (from a 1024-aligned base)
constexpr uint32_t OFF_WT = 0, OFF_AT = 16384, OFF_H = 32768, OFF_DH = 65536, OFF_W2 = 98304, OFF_DL = 102400,
                   OFF_ONES = 106496, OFF_RED = 107008, DB_SMEM = 107520;
// TMEM columns
constexpr uint32_t TC_FWD = 0, TC_LOG = 128, TC_DH = 144, TC_DG = 272, TC_WG = 336, TC_W2 = 400, TC_B1 = 416;

__device__ __forceinline__ uint32_t sw128_off(uint32_t row, uint32_t chunk) {  // 16-byte chunk in a 128B-swizzled row
  return row * 128u + ((chunk ^ (row & 7u)) << 4);
}
__device__ __forceinline__ float4 ld_cg_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void tma_store_4d(const void* tmap, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(tmap),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

struct DenseBlockArgs {
  float* acc;              // [B][D1] fp32 split-K accumulator of this parity (zero on entry)
  float* acc_next;         // the other parity's accumulator: cleared here for the next step
  uint32_t* sync;          // arrival counter of the grid barrier (monotonic)
  const float* b1;         // [D1]
  const float* b2;         // [NC]
  const int* labels;       // [B]
  const __nv_bfloat16* a2; // [B][F]
  const uint8_t* idx2;     // [B][F]
  __nv_bfloat16* dy2;      // [B][P1h][P1w][C2]
  float* g_w1;             // [D1][F]
  float* g_b1;             // [D1]
  float* g_w2;             // [NC][D1]
  float* g_b2;             // [NC]
  float* metrics;          // [3]
  float* step_out;         // [2]
  __nv_bfloat16* dh_out;   // optional [B][D1] (tests)
  TfdlCnnDims d;
};

__global__ void __launch_bounds__(DB_THREADS, 1)
dense_block_kernel(const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmA2,
                   const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmDY,
                   DenseBlockArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_tma, bar_m[5];
  __shared__ uint32_t slot;
  const TfdlCnnDims& d = a.d;
  const int B = d.B, NC = d.NC, F = d.F;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, lq = warp & 3, half = warp >> 2;
  const int p = blockIdx.x, NP = gridDim.x;
  const uint32_t sb = smem_u32(smem);
  float* red = reinterpret_cast<float*>(smem + OFF_RED);  // [0]: the one reduction this CTA reports

  DBPHASE(0);
  if (warp == 0) tmem_alloc(&slot, 512);
  if (tid == 32) {
    mbar_init(&bar_tma, 1);
    for (int i = 0; i < 5; ++i) mbar_init(&bar_m[i], 1);
    fence_mbar_init();
  }
  if (tid < 128) reinterpret_cast<uint32_t*>(smem + OFF_ONES)[tid] = 0x3F803F80u;  // bf16 ones
  if (tid < 32) red[tid] = 0.f;
  // clear the other parity's accumulator (nobody reads it during this step)
  for (int i = p * DB_THREADS + tid; i < B * D1 / 4; i += NP * DB_THREADS)
    reinterpret_cast<float4*>(a.acc_next)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  const uint32_t t_row = tmem + ((uint32_t)(lq * 32) << 16);
  const int ksteps_b = B / 16;  // K-steps when the reduction runs over samples
  // values the later phases need from global memory are fetched now, off the critical path
  const int my_s = lq * 32 + lane;
  const int my_y = (half == 0 && my_s < B) ? a.labels[my_s] : 0;
  float my_b2[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) my_b2[j] = (half == 0 && j < NC) ? __ldg(a.b2 + j) : 0.f;

  DBPHASE(1);
  pdl_launch_dependents();
  pdl_wait();  // a2 / idx2 (conv forward) are complete; the prologue above overlapped the previous kernel's tail
  // ---------------- phase 1: forward split-K chunk -> RED into acc -------------------------------------
  if (tid == 0) {
    mbar_expect_tx(&bar_tma, 16384 + 16384 + 2 * 2048);
    tma_load_2d(sb + OFF_WT, &tmW1, p * C2, 0, &bar_tma);   // W1[:, p*64 : p*64+64]
    tma_load_2d(sb + OFF_AT, &tmA2, p * C2, 0, &bar_tma);   // a2[:, p*64 : ...] (rows >= B zero-filled)
    tma_load_2d(sb + OFF_W2, &tmW2, 0, 0, &bar_tma);        // W2[:, 0:64]   (rows >= NC zero-filled)
    tma_load_2d(sb + OFF_W2 + 2048, &tmW2, 64, 0, &bar_tma);
    mbar_wait(&bar_tma, 0);
    tc_fence_after_sync();
    const uint32_t idesc = umma_idesc_bf16(128, B, false, false);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks)
      umma_bf16(tmem + TC_FWD, umma_smem_desc(sb + OFF_WT + ks * 32, 16, 1024, SWZ_128B),
                umma_smem_desc(sb + OFF_AT + ks * 32, 16, 1024, SWZ_128B), idesc, ks > 0);
    umma_commit(&bar_m[0]);
  }
  mbar_wait(&bar_m[0], 0);
  tc_fence_after_sync();
  DBPHASE(2);
  {
    const int unit = lq * 32 + lane;
    for (int c0 = half * 32; c0 < B; c0 += 64) {  // sample columns; lanes = units -> coalesced REDs
      uint32_t v[32];
      tmem_ld32(t_row + TC_FWD + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c0 + i < B) atomicAdd(a.acc + (size_t)(c0 + i) * D1 + unit, __uint_as_float(v[i]));
    }
  }
  DBPHASE(3);
  // ---------------- grid barrier: all split-K partials are in acc --------------------------------------
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const uint32_t ticket = atomicAdd(a.sync, 1u);
    const uint32_t target = (ticket / NP + 1u) * NP;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_gpu(a.sync) - target) < 0) {
      if (++spins > TFDL_SPIN_LIMIT) __trap();
    }
  }
  __syncthreads();

  DBPHASE(4);
  // ---------------- phase 2: H = relu(acc + b1) as a bf16 swizzled tile -----------------------------------
  {
    // 16-byte chunks (8 units) of H: chunk i -> sample i/16, chunk i%16. All loads are issued before any use.
    constexpr int MAXIT = 128 * 16 / DB_THREADS;  // B <= 128
    float4 x0[MAXIT], x1[MAXIT];
#pragma unroll
    for (int it = 0; it < MAXIT; ++it) {
      const int i = tid + it * DB_THREADS;
      if (i < B * 16) {
        x0[it] = ld_cg_f4(a.acc + (size_t)i * 8);
        x1[it] = ld_cg_f4(a.acc + (size_t)i * 8 + 4);
      }
    }
    const int c = tid & 15;  // DB_THREADS % 16 == 0: the chunk column is the same in every iteration
    const float4 b0 = *reinterpret_cast<const float4*>(a.b1 + c * 8), b1v = *reinterpret_cast<const float4*>(a.b1 + c * 8 + 4);
#pragma unroll
    for (int it = 0; it < MAXIT; ++it) {
      const int i = tid + it * DB_THREADS;
      if (i < B * 16) {
        const int r = i >> 4;
        uint4 o;
        o.x = pack_bf16x2(fmaxf(x0[it].x + b0.x, 0.f), fmaxf(x0[it].y + b0.y, 0.f));
        o.y = pack_bf16x2(fmaxf(x0[it].z + b0.z, 0.f), fmaxf(x0[it].w + b0.w, 0.f));
        o.z = pack_bf16x2(fmaxf(x1[it].x + b1v.x, 0.f), fmaxf(x1[it].y + b1v.y, 0.f));
        o.w = pack_bf16x2(fmaxf(x1[it].z + b1v.z, 0.f), fmaxf(x1[it].w + b1v.w, 0.f));
        *reinterpret_cast<uint4*>(smem + OFF_H + (c >> 3) * 16384 + sw128_off(r, c & 7)) = o;
      }
    }
  }
  fence_proxy_async_smem();
  __syncthreads();

  DBPHASE(5);
  // ---------------- phase 3: logits on the tensor core, softmax cross-entropy per thread ------------------
  if (tid == 0) {
    tc_fence_after_sync();
    const uint32_t idesc = umma_idesc_bf16(128, 16, false, false);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)
      umma_bf16(tmem + TC_LOG, umma_smem_desc(sb + OFF_H + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024, SWZ_128B),
                umma_smem_desc(sb + OFF_W2 + (ks >> 2) * 2048 + (ks & 3) * 32, 16, 1024, SWZ_128B), idesc, ks > 0);
    umma_commit(&bar_m[1]);
  }
  mbar_wait(&bar_m[1], 0);
  tc_fence_after_sync();
  DBPHASE(6);
  if (half == 0) {
    const int s = lq * 32 + lane;  // this thread's sample
    uint32_t v[16];
    tmem_ld16(t_row + TC_LOG, v);
    tmem_ld_wait();
    float z[16], mx = -INFINITY;
    int am = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      z[j] = j < NC ? __uint_as_float(v[j]) + my_b2[j] : -INFINITY;
      if (z[j] > mx) { mx = z[j]; am = j; }
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) { z[j] = j < NC ? __expf(z[j] - mx) : 0.f; sum += z[j]; }
    const bool valid = s < B;
    const int y = my_y;
    const float inv = valid ? 1.f / ((float)B * (float)d.nrep) : 0.f, rs = 1.f / sum;
    float loss = 0.f, dl[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float pj = z[j] * rs;
      if (j == y) loss = -__logf(fmaxf(pj, 1e-30f));
      dl[j] = (valid && j < NC) ? (pj - (j == y ? 1.f : 0.f)) * inv : 0.f;  // rows >= B hold junk: force 0
    }
    // DL tile, no swizzle: chunk c (8 classes) of row s at (s/8)*256 + c*128 + (s%8)*16
    uint4 o0 = make_uint4(pack_bf16x2(dl[0], dl[1]), pack_bf16x2(dl[2], dl[3]), pack_bf16x2(dl[4], dl[5]), pack_bf16x2(dl[6], dl[7]));
    uint4 o1 = make_uint4(pack_bf16x2(dl[8], dl[9]), pack_bf16x2(dl[10], dl[11]), pack_bf16x2(dl[12], dl[13]), pack_bf16x2(dl[14], dl[15]));
    uint8_t* row = smem + OFF_DL + (s >> 3) * 256 + (s & 7) * 16;
    *reinterpret_cast<uint4*>(row) = o0;
    *reinterpret_cast<uint4*>(row + 128) = o1;
    // every CTA holds all dLogits; the reductions that are reported once are dealt out: CTA j < NC sums
    // class j (db2), CTA NC the loss, CTA NC+1 the correct predictions
    if (p <= NC + 1) {
      float x = p == NC ? (valid ? loss : 0.f) : p == NC + 1 ? ((valid && am == y) ? 1.f : 0.f) : 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) x = (j == p && j < NC) ? dl[j] : x;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == 0) atomicAdd(&red[0], x);
    }
  }
  fence_proxy_async_smem();
  __syncthreads();

  DBPHASE(7);
  // ---------------- phase 4: dH = relu'(H) * (dLogits * W2) ------------------------------------------------
  if (tid == 0) {
    tc_fence_after_sync();
    // A = DL (K-major, no swizzle: K chunks 128 B apart, 8-row groups 256 B apart); B = W2 as [K=classes][N=units]
    umma_bf16(tmem + TC_DH, umma_smem_desc(sb + OFF_DL, 128, 256, SWZ_NONE),
              umma_smem_desc(sb + OFF_W2, 2048, 1024, SWZ_128B), umma_idesc_bf16(128, 128, false, true), false);
    umma_commit(&bar_m[2]);
  }
  mbar_wait(&bar_m[2], 0);
  tc_fence_after_sync();
  DBPHASE(8);
  {
    const int s = lq * 32 + lane;
    for (int c0 = half * 32; c0 < D1; c0 += 64) {
      uint32_t v[32];
      tmem_ld32(t_row + TC_DH + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {  // four 16-byte chunks of 8 units
        const int c = (c0 >> 3) + cc;
        const uint32_t off = (c >> 3) * 16384 + sw128_off(s, c & 7);
        const uint4 h = *reinterpret_cast<const uint4*>(smem + OFF_H + off);
        const uint32_t hw[4] = {h.x, h.y, h.z, h.w};
        uint32_t ow[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float g0 = (hw[k] & 0x7FFFu) ? __uint_as_float(v[cc * 8 + 2 * k]) : 0.f;   // h >= 0 always: nonzero <=> positive
          const float g1 = (hw[k] >> 16) ? __uint_as_float(v[cc * 8 + 2 * k + 1]) : 0.f;
          ow[k] = pack_bf16x2(g0, g1);
        }
        const uint4 o = make_uint4(ow[0], ow[1], ow[2], ow[3]);
        *reinterpret_cast<uint4*>(smem + OFF_DH + off) = o;
        if (a.dh_out && p == 0 && s < B) *reinterpret_cast<uint4*>(a.dh_out + (size_t)s * D1 + c * 8) = o;
      }
    }
  }
  fence_proxy_async_smem();
  __syncthreads();

  DBPHASE(9);
  // ---------------- phase 5: input gradient, weight gradients ---------------------------------------------
  if (tid == 0) {
    tc_fence_after_sync();
    const uint32_t i_dg = umma_idesc_bf16(128, C2, false, true), i_wg = umma_idesc_bf16(128, C2, true, true),
                   i_16 = umma_idesc_bf16(128, 16, true, true);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)  // da2[s][f] = sum_u DH[s][u] * W1[u][f]
      umma_bf16(tmem + TC_DG, umma_smem_desc(sb + OFF_DH + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024, SWZ_128B),
                umma_smem_desc(sb + OFF_WT + ks * 2048, 16, 1024, SWZ_128B), i_dg, ks > 0);
    for (int ks = 0; ks < ksteps_b; ++ks)  // dW1[u][f] = sum_s DH[s][u] * a2[s][f]
      umma_bf16(tmem + TC_WG, umma_smem_desc(sb + OFF_DH + ks * 2048, 16384, 1024, SWZ_128B),
                umma_smem_desc(sb + OFF_AT + ks * 2048, 16, 1024, SWZ_128B), i_wg, ks > 0);
    if (p == NP - 1) {
      for (int ks = 0; ks < ksteps_b; ++ks) {
        // dW2[u][j] = sum_s H[s][u] * dl[s][j]   (DL as MN-major: class chunks 128 B apart, sample groups 256 B apart)
        umma_bf16(tmem + TC_W2, umma_smem_desc(sb + OFF_H + ks * 2048, 16384, 1024, SWZ_128B),
                  umma_smem_desc(sb + OFF_DL + ks * 512, 256, 128, SWZ_NONE), i_16, ks > 0);
        // db1[u] = sum_s DH[s][u] * 1
        umma_bf16(tmem + TC_B1, umma_smem_desc(sb + OFF_DH + ks * 2048, 16384, 1024, SWZ_128B),
                  umma_smem_desc(sb + OFF_ONES, 128, 256, SWZ_NONE), i_16, ks > 0);
      }
    }
    umma_commit(&bar_m[3]);
  }
  // dgrad epilogue inputs (a2 for the ReLU mask, idx2 for the un-pool) do not depend on the MMAs: fetch first
  const int ph = p / d.P2w, pw = p % d.P2w;
  uint4 av[4];
  uint2 iv[4];
  {
    const int c0 = half * 32;
    if (my_s < B) {
      const size_t io = (size_t)my_s * F + (size_t)p * C2 + c0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        av[k] = *reinterpret_cast<const uint4*>(a.a2 + io + k * 8);
        iv[k] = *reinterpret_cast<const uint2*>(a.idx2 + io + k * 8);
      }
    }
  }
  mbar_wait(&bar_m[3], 0);
  tc_fence_after_sync();
  DBPHASE(10);
  // ReLU mask (a2 > 0) + max-pool backward through argmax. The four window positions are four [B][64] bf16
  // tiles (128B-swizzled, reusing the H/DH area: all MMAs have retired) that leave through TMA stores,
  // so the scattered per-sample rows of dy2 are written as full 128-byte lines by the copy engine.
  {
    const int c0 = half * 32;  // C2 = 64: one 32-channel chunk per warp of the pair
    uint32_t v[32];
    tmem_ld32(t_row + TC_DG + c0, v);
    tmem_ld_wait();
    if (my_s < B) {
      const uint32_t* aw = reinterpret_cast<const uint32_t*>(av);
      const uint32_t* iw = reinterpret_cast<const uint32_t*>(iv);
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
#pragma unroll
        for (int k = 0; k < 4; ++k)
          *reinterpret_cast<uint4*>(smem + OFF_H + q * 16384 + sw128_off(my_s, (c0 >> 3) + k)) =
              make_uint4(outw[4 * k], outw[4 * k + 1], outw[4 * k + 2], outw[4 * k + 3]);
      }
    }
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) tma_store_4d(&tmDY, sb + OFF_H + q * 16384, 0, 2 * pw + (q & 1), 2 * ph + (q >> 1), 0);
    tma_store_commit();
  }
  DBPHASE(11);
  // wgrad epilogue: dW1[:, p*64 : p*64+64] (lanes = units), plain stores
  {
    const int unit = lq * 32 + lane;
    for (int c0 = half * 32; c0 < C2; c0 += 64) {
      uint32_t v[32];
      tmem_ld32(t_row + TC_WG + c0, v);
      tmem_ld_wait();
      float4* dst = reinterpret_cast<float4*>(a.g_w1 + (size_t)unit * F + (size_t)p * C2 + c0);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        dst[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                             __uint_as_float(v[4 * i + 3]));
    }
  }
  DBPHASE(12);
  if (p == NP - 1) {
    const int unit = lq * 32 + lane;
    uint32_t v[16];
    tmem_ld16(t_row + (half == 0 ? TC_W2 : TC_B1), v);
    tmem_ld_wait();
    if (half == 0) {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (j < NC) a.g_w2[(size_t)j * D1 + unit] = __uint_as_float(v[j]);
    } else {
      a.g_b1[unit] = __uint_as_float(v[0]);
    }
  }
  if (tid == 0) {
    if (p < NC) a.g_b2[p] = red[0];
    if (p == NC) {
      atomicAdd(a.metrics + 0, red[0]);
      atomicAdd(a.metrics + 2, (float)B);
      a.step_out[0] = red[0] / (float)B;
    }
    if (p == NC + 1) {
      atomicAdd(a.metrics + 1, red[0]);
      a.step_out[1] = red[0] / (float)B;
    }
    tma_store_wait_all();  // the dy2 tiles must have left shared memory before the CTA exits
  }
  DBPHASE(13);
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
  DBPHASE(14);
}

}  // namespace
}  // namespace tfdl

using namespace tfdl;

extern "C" int tfdl_debug_read_phases_db(long long* host_out /*[64*16]*/) {
#ifdef TFDL_PROFILE_PHASES
  TFDL_CUDA_OK(cudaMemcpyFromSymbol(host_out, g_phase_db, sizeof(long long) * 64 * 16));
  return 0;
#else
  (void)host_out;
  return -1;
#endif
}

extern "C" int tfdl_cnn_dense_block_supported(const TfdlCnnDims* d) {
  return d->D1 == 128 && d->C2 == 64 && d->NC <= 16 && d->B % 16 == 0 && d->B <= 128 && d->P2h * d->P2w <= 148 &&
         d->P2h * d->P2w >= d->NC + 3;
}

extern "C" int tfdl_cnn_dense_block(const void* a2, const uint8_t* idx2, const void* w1h, const void* w2h, const float* b1,
                                    const float* b2, const int* labels, float* acc, float* acc_next, uint32_t* sync,
                                    void* dy2, float* g_w1, float* g_b1, float* g_w2, float* g_b2, float* metrics,
                                    float* step_out, void* dh_out, const TfdlCnnDims* d, void* stream) {
  TFDL_REQUIRE(tfdl_cnn_dense_block_supported(d), "dense_block: unsupported dims (needs D1=128, C2=64, NC<=16, B<=128)");
  CUtensorMap tmW1, tmA2, tmW2;
  uint64_t dw[2] = {(uint64_t)d->F, 128}, sw[1] = {(uint64_t)d->F * 2};
  uint32_t bw[2] = {64, 128};
  if (int rc = make_tmap_bf16_nd(&tmW1, w1h, 2, dw, sw, bw, 3)) return rc;
  uint64_t da[2] = {(uint64_t)d->F, (uint64_t)d->B};
  if (int rc = make_tmap_bf16_nd(&tmA2, a2, 2, da, sw, bw, 3)) return rc;
  uint64_t d2[2] = {128, (uint64_t)d->NC}, s2[1] = {256};
  uint32_t b2x[2] = {64, 16};
  if (int rc = make_tmap_bf16_nd(&tmW2, w2h, 2, d2, s2, b2x, 3)) return rc;
  CUtensorMap tmDY;
  uint64_t dd[4] = {64, (uint64_t)d->P1w, (uint64_t)d->P1h, (uint64_t)d->B};
  uint64_t sd[3] = {128, (uint64_t)d->P1w * 128, (uint64_t)d->P1h * d->P1w * 128};
  uint32_t bd[4] = {64, 1, 1, (uint32_t)d->B};
  if (int rc = make_tmap_bf16_nd(&tmDY, dy2, 4, dd, sd, bd, 3)) return rc;
  DenseBlockArgs a{};
  a.acc = acc; a.acc_next = acc_next; a.sync = sync; a.b1 = b1; a.b2 = b2; a.labels = labels;
  a.a2 = (const __nv_bfloat16*)a2; a.idx2 = idx2; a.dy2 = (__nv_bfloat16*)dy2; a.g_w1 = g_w1; a.g_b1 = g_b1;
  a.g_w2 = g_w2; a.g_b2 = g_b2; a.metrics = metrics; a.step_out = step_out; a.dh_out = (__nv_bfloat16*)dh_out; a.d = *d;
  const size_t smem = DB_SMEM + 1024;
  TFDL_CUDA_OK(cudaFuncSetAttribute(dense_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  TFDL_CUDA_OK(launch_kernel_pdl(1, dense_block_kernel, dim3(d->P2h * d->P2w), dim3(DB_THREADS), smem, (cudaStream_t)stream,
                                 tmW1, tmA2, tmW2, tmDY, a));
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
