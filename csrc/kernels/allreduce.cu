// Fused gradient all-reduce + optimizer update over NVLink peer memory.
//
// This replaces the reference's per-step gradient aggregation - `CollectiveAllReduce` /
// `NcclAllReduce` followed by one `ResourceApplyGradientDescent` per variable
// (/root/reference/README.md:17,21,23,67; tf_dist_example.py:51,59) - with ONE kernel per step that
//   * waits until every peer's backward pass has finished (flag barrier over NVLink),
//   * sums the peers' gradient arenas by loading them directly from peer memory,
//   * applies the 1/N scale and the SGD / momentum / Nesterov update to the fp32 master weights,
//   * refreshes the bf16 shadow copy of the weights that the tcgen05 kernels consume,
//   * clears the gradient buffer of the *other* parity for the next step,
// so no NCCL call, no pack/unpack copy and no separate optimizer launch is on the step path.
//
// Three data paths, chosen per message size by ops/allreduce.py (the reference's `AUTO`):
//   one-shot   every rank reads all W arenas and updates its own full copy        (latency-optimal)
//   two-shot   rank r reduces slice r, updates it, and stores the new weights
//              into every rank's arena (reduce-scatter + all-gather by peer stores) (bandwidth-optimal)
//   multimem   as two-shot, but the slice is reduced inside the NVSwitch with
//              multimem.ld_reduce and broadcast with multimem.st                    (NVLS)
//   push       every rank WRITES its gradients into every peer's inbox as 16-byte packets
//              {value, epoch, value, epoch}; the receiver polls its own memory and the arrival of
//              the data is the signal - no flag barrier, no peer load: one NVLink one-way trip   (lowest latency)
//   push2      the same packets in two hops: gradient slices to their owner, updated weight
//              slices from the owner to everyone (reduce-scatter + all-gather, no barrier)      (mid sizes)
//
// Synchronisation protocol (all variants):
//   * `epoch` is a device-local step counter. Step e (1-based) uses gradient buffer parity
//     (e-1)&1; backward kernels read the same counter to pick the buffer they write.
//   * start barrier: block b of rank r writes e into slot [b][r] of every peer's flag pad
//     (st.release.sys) and spins until its own pad holds >= e in slots [b][*] (ld.acquire.sys).
//     A peer's flag can only arrive after that peer's stream reached this kernel, i.e. after its
//     backward for step e is complete and its step e-1 reads are over. That is why gradient
//     buffers are double-buffered by parity instead of guarded by a second barrier: the buffer a
//     fast rank starts overwriting in step e+1 is never the one a slow rank still reads in step e.
//   * end barrier (two-shot / multimem only): same exchange on a second pad after the peer stores,
//     so the kernel does not retire before every peer's slice has landed in this rank's arena.
//   * every spin is bounded (trap after TFDL_SPIN_LIMIT polls) and polls a host-visible abort word.
// All ranks compute bit-identical weights: one-shot sums in rank order on every rank; in the
// sliced variants only the slice owner computes.
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

namespace tfdl {

__device__ __forceinline__ bool spin_until_ge(const uint32_t* flag, uint32_t want, const uint32_t* abort_word) {
  uint32_t spins = 0;
  while ((int32_t)(ld_acquire_sys(flag) - want) < 0) {  // wrap-safe: epochs of the ranks of a group differ by at most 1
    if ((++spins & 0x3FF) == 0) {
      if (abort_word && ld_relaxed_sys(abort_word) != 0) return false;
      if (spins > TFDL_SPIN_LIMIT) __trap();
    }
  }
  return true;
}

// Cross-rank barrier for block `b`: returns false if aborted.
__device__ __forceinline__ bool block_barrier(const TfdlFusedArParams& p, uint32_t pad_off, uint32_t value) {
  const int W = p.world;
  const uint32_t slot_base = pad_off + blockIdx.x * W;
  __shared__ int ok;
  if (threadIdx.x == 0) ok = 1;
  __syncthreads();  // all prior global writes of this block are ordered before the release below
  if ((int)threadIdx.x < W && (int)threadIdx.x != p.rank) {
    st_release_sys(p.flags[threadIdx.x] + slot_base + p.rank, value);  // release.sys: cumulative over the bar.sync above
    if (!spin_until_ge(p.flags[p.rank] + slot_base + threadIdx.x, value, p.abort_word)) ok = 0;
  }
  __syncthreads();
  return ok != 0;
}

// Sum of element i over the W ranks' buffers. All W loads are issued before the first add, so the cost is ONE
// NVLink round trip, not W of them (a rolled loop of volatile peer loads serialises: 7 x ~1.7 us at W = 8).
// Summation is in rank order on every rank -> bit-identical results everywhere.
__device__ __forceinline__ float4 gather_sum(float* const* bufs, int64_t i, int W, int rank) {
  float4 x[TFDL_MAX_RANKS];
#pragma unroll
  for (int r = 0; r < TFDL_MAX_RANKS; ++r) {
    if (r < W) {
      const float4* src = reinterpret_cast<const float4*>(bufs[r]) + i;
      x[r] = (r == rank) ? *src : ld_peer_f4(src);
    }
  }
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < TFDL_MAX_RANKS; ++r) {
    if (r < W) { g.x += x[r].x; g.y += x[r].y; g.z += x[r].z; g.w += x[r].w; }
  }
  return g;
}

__device__ __forceinline__ float4 sgd_update(float4 w, float4 g, float4* v, const TfdlHyper& h) {
  if (h.momentum == 0.f) {
    w.x -= h.lr * g.x; w.y -= h.lr * g.y; w.z -= h.lr * g.z; w.w -= h.lr * g.w;
    return w;
  }
  float4 m = *v;
  m.x = h.momentum * m.x - h.lr * g.x; m.y = h.momentum * m.y - h.lr * g.y;
  m.z = h.momentum * m.z - h.lr * g.z; m.w = h.momentum * m.w - h.lr * g.w;
  *v = m;
  if (h.nesterov != 0.f) {
    w.x += h.momentum * m.x - h.lr * g.x; w.y += h.momentum * m.y - h.lr * g.y;
    w.z += h.momentum * m.z - h.lr * g.z; w.w += h.momentum * m.w - h.lr * g.w;
  } else {
    w.x += m.x; w.y += m.y; w.z += m.z; w.w += m.w;
  }
  return w;
}

// Adam (tf.keras.optimizers.Adam: m = b1 m + (1-b1) g, v = b2 v + (1-b2) g^2, w -= lr_t m / (sqrt(v) + eps) with the bias
// correction folded into lr_t = lr sqrt(1 - b2^t) / (1 - b1^t)). Selected by hyper.beta2 > 0; `m` and `v` are the two slot
// arenas; the step number t is the kernel's own epoch counter plus hyper.t_offset (restored runs), so LR schedules and
// the bias correction need no re-capture of a step graph.
struct OptStep {
  bool adam;
  float lr_t;
};
__device__ __forceinline__ OptStep opt_step(const TfdlHyper& h, uint32_t e) {
  OptStep o;
  o.adam = h.beta2 > 0.f;
  o.lr_t = h.lr;
  if (o.adam) {
    const float t = (float)e + h.t_offset;
    o.lr_t = h.lr * sqrtf(1.f - powf(h.beta2, t)) / (1.f - powf(h.beta1, t));
  }
  return o;
}
__device__ __forceinline__ float adam1(float w, float g, float& m, float& v, const TfdlHyper& h, float lr_t) {
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  return w - lr_t * m / (sqrtf(v) + h.eps);
}
__device__ __forceinline__ float4 opt_update(float4 w, float4 g, float4* mp, float4* vp, const TfdlHyper& h, const OptStep& o) {
  if (!o.adam) return sgd_update(w, g, mp, h);
  float4 m = *mp, v = *vp;
  w.x = adam1(w.x, g.x, m.x, v.x, h, o.lr_t);
  w.y = adam1(w.y, g.y, m.y, v.y, h, o.lr_t);
  w.z = adam1(w.z, g.z, m.z, v.z, h, o.lr_t);
  w.w = adam1(w.w, g.w, m.w, v.w, h, o.lr_t);
  *mp = m;
  *vp = v;
  return w;
}

// Scalar tail for element counts that are not a multiple of 4 (the vector loops cover [0, n & ~3)).
__device__ __forceinline__ float ld_peer_f1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f1(float* p, float v) {
  asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ float gather_sum1(float* const* bufs, int64_t i, int W, int rank) {
  float g = 0.f;
  for (int r = 0; r < W; ++r) g += (r == rank) ? bufs[r][i] : ld_peer_f1(bufs[r] + i);
  return g;
}

__device__ __forceinline__ uint2 to_bf16x4(float4 w) {
  return make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
}

__device__ __forceinline__ void finish_step(const TfdlFusedArParams& p, uint32_t e) {
  // last block to leave publishes the new epoch for the next step's kernels
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    uint32_t t = atomicAdd(p.done, 1u);
    if (t == gridDim.x - 1) {
      *p.done = 0;
      __threadfence();
      *p.epoch = e;
    }
  }
}

// -------------------------------------------------------------------------------------------
// one-shot
// -------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) fused_ar_oneshot_kernel(TfdlFusedArParams p) {
  pdl_wait();  // launched as a programmatic dependent of the last backward kernel: its gradients are final from here
  const uint32_t e = *p.epoch + 1;
  const int par = p.parity >= 0 ? p.parity : (int)((e - 1) & 1);
  const TfdlHyper h = *p.hyper;
  if (p.world > 1 && !block_barrier(p, 0, e)) return;
  const int64_t n4 = p.n >> 2;
  const int W = p.world;
  float4* w4 = reinterpret_cast<float4*>(p.params[p.rank]);
  float4* v4 = reinterpret_cast<float4*>(p.mom);
  float4* u4 = reinterpret_cast<float4*>(p.mom2);
  const OptStep os = opt_step(h, e);
  uint2* s2 = reinterpret_cast<uint2*>(p.shadow[p.rank]);
  float4* z4 = reinterpret_cast<float4*>(p.grads[par ^ 1][p.rank]);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 g = gather_sum(p.grads[par], i, W, p.rank);
    g.x *= h.scale; g.y *= h.scale; g.z *= h.scale; g.w *= h.scale;
    float4 w = opt_update(w4[i], g, v4 ? v4 + i : nullptr, u4 ? u4 + i : nullptr, h, os);
    w4[i] = w;
    if (s2) s2[i] = to_bf16x4(w);
    if (z4) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if ((p.n & 3) && blockIdx.x == 0 && threadIdx.x == 0) {  // scalar tail (n not a multiple of 4)
    for (int64_t i = n4 * 4; i < p.n; ++i) {
      float4 g = make_float4(gather_sum1(p.grads[par], i, W, p.rank) * h.scale, 0.f, 0.f, 0.f);
      float4 m = p.mom ? make_float4(p.mom[i], 0.f, 0.f, 0.f) : g;
      float4 v2 = p.mom2 ? make_float4(p.mom2[i], 1.f, 1.f, 1.f) : g;
      float4 w = opt_update(make_float4(p.params[p.rank][i], 0.f, 0.f, 0.f), g, p.mom ? &m : nullptr, p.mom2 ? &v2 : nullptr, h, os);
      if (p.mom) p.mom[i] = m.x;
      if (p.mom2) p.mom2[i] = v2.x;
      p.params[p.rank][i] = w.x;
      if (p.shadow[p.rank]) reinterpret_cast<__nv_bfloat16*>(p.shadow[p.rank])[i] = __float2bfloat16(w.x);
      if (p.grads[par ^ 1][p.rank]) p.grads[par ^ 1][p.rank][i] = 0.f;
    }
  }
  finish_step(p, e);
}

// -------------------------------------------------------------------------------------------
// two-shot (MC = false) and multimem / NVLS (MC = true)
// -------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float4* src) {
  float4 g;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w)
               : "l"(src)
               : "memory");
  return g;
}
__device__ __forceinline__ void mc_st_f4(float4* dst, float4 w) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(w.x), "f"(w.y), "f"(w.z),
               "f"(w.w)
               : "memory");
}

// U float4 per thread per trip: the loop is bound by the NVLink round trip, so the bytes in flight per SM set the
// bandwidth (one float4 per thread was 296 GB/s at W = 2 by Little's law).
template <bool MC, int U>
__global__ void __launch_bounds__(512) fused_ar_sliced_kernel(TfdlFusedArParams p) {
  const uint32_t e = *p.epoch + 1;
  const int par = p.parity >= 0 ? p.parity : (int)((e - 1) & 1);
  const TfdlHyper h = *p.hyper;
  const int W = p.world;
  if (!block_barrier(p, 0, e)) return;
  const int64_t n4 = p.n >> 2;
  const int64_t per = (n4 + W - 1) / W;
  const int64_t lo = per * p.rank, hi = min(n4, lo + per);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  float4* v4 = reinterpret_cast<float4*>(p.mom);
  float4* u4 = reinterpret_cast<float4*>(p.mom2);
  const OptStep os = opt_step(h, e);
  const float4* w_own = reinterpret_cast<const float4*>(p.params[p.rank]);
  // The other-parity gradient buffer (whole local arena, n4 float4) is cleared INSIDE the reduction loop: the thread
  // that handles element j of its slice also zeroes elements [j*W, j*W + W) while its NVLink loads are in flight, so
  // this HBM work hides under the transfer instead of following it (per * W >= n4 covers the arena).
  float4* z4 = reinterpret_cast<float4*>(p.grads[par ^ 1][p.rank]);
  for (int64_t i0 = lo + tid; i0 < hi; i0 += nth * U) {
    float4 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * nth;
      if (i < hi) g[u] = MC ? mc_ld_reduce_f4(reinterpret_cast<const float4*>(p.mc_grads[par]) + i)
                            : gather_sum(p.grads[par], i, W, p.rank);
    }
    if (z4) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t j = i0 + u * nth - lo;
        if (i0 + u * nth < hi)
          for (int64_t z = j * W; z < min(n4, j * W + W); ++z) z4[z] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * nth;
      if (i >= hi) continue;
      float4 gg = g[u];
      gg.x *= h.scale; gg.y *= h.scale; gg.z *= h.scale; gg.w *= h.scale;
      const float4 w = opt_update(w_own[i], gg, v4 ? v4 + i : nullptr, u4 ? u4 + i : nullptr, h, os);
      if (MC) {
        mc_st_f4(reinterpret_cast<float4*>(p.mc_params) + i, w);
      } else {
#pragma unroll
        for (int r = 0; r < TFDL_MAX_RANKS; ++r)
          if (r < W) {
            float4* dst = reinterpret_cast<float4*>(p.params[r]) + i;
            if (r == p.rank) *dst = w; else st_peer_f4(dst, w);
          }
      }
    }
  }
  // (a last rank with a short slice leaves a tail of the arena to clear: elements >= (hi - lo) * W)
  if (z4)
    for (int64_t i = (hi - lo) * W + tid; i < n4; i += nth) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  // end barrier: block b of every peer has stored its part of every slice into my arena
  if (!block_barrier(p, p.pad_stride, e)) return;
  // bf16 shadow of the WHOLE arena, cast locally (the weights travel over NVLink once, as fp32; the shadow never
  // does). Block b converts exactly the elements that the peers' blocks b wrote - the ones its barrier covers.
  uint2* s2 = reinterpret_cast<uint2*>(p.shadow[p.rank]);
  if (s2) {
    for (int r = 0; r < W; ++r) {
      const int64_t rlo = per * r, rhi = min(n4, rlo + per);
      for (int64_t i = rlo + tid; i < rhi; i += nth) s2[i] = to_bf16x4(ld_peer_f4(w_own + i));
    }
  }
  finish_step(p, e);
}

// -------------------------------------------------------------------------------------------
// push / push2: flag-in-data packets (the data IS the signal)
// -------------------------------------------------------------------------------------------
// A packet is 16 bytes {v0, e, v1, e}: two fp32 values, each next to a copy of the step's epoch. NVLink
// delivers aligned 8-byte units whole, so a receiver that sees `e` in both flag words holds both values of
// step e; nothing orders packets against each other and nothing has to. A float4 is two packets. Inboxes are
// double-buffered by epoch parity: a sender can reach step e+2 (same slot) only after its step e+1 completed,
// which needed this rank's step e+1 packets, which this rank sent after it finished reading step e.
// Every rank adds the W contributions in rank order (its own from the local arena), so all replicas compute
// bit-identical weights. Peers' gradient arenas are never read, so the parity double buffer of the gradient
// arena is not needed for correctness here (it is kept: backward accumulates into a pre-cleared buffer).
#define TFDL_PUSH_SPIN_LIMIT (1u << 25)

__device__ __forceinline__ void st_packet(uint4* dst, float a, float b, uint32_t e) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(__float_as_uint(a)), "r"(e),
               "r"(__float_as_uint(b)), "r"(e)
               : "memory");
}
__device__ __forceinline__ uint4 ld_packet(const uint4* src) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(src)
               : "memory");
  return v;
}
// Packet layout of one inbox region: float4 #i occupies packets pkt_of(i) and pkt_of(i) + 32, so the 32 lanes of a
// warp (consecutive i) store 512 contiguous bytes per instruction instead of 16-byte pieces at a 32-byte stride.
__host__ __device__ __forceinline__ int64_t pkt_of(int64_t i) { return ((i >> 5) << 6) + (i & 31); }
__host__ __device__ __forceinline__ int64_t region_packets(int64_t n4) { return ((n4 + 31) >> 5) << 6; }
__device__ __forceinline__ void push_f4(uint4* region, int64_t i, float4 g, uint32_t e) {
  uint4* dst = region + pkt_of(i);
  st_packet(dst, g.x, g.y, e);
  st_packet(dst + 32, g.z, g.w, e);
}

// Sum over ranks of float4 #slot: peers' values from this rank's inbox regions (region r at base + r*stride),
// this rank's own value `own` in its rank-order position. All 2(W-1) packet loads are in flight together.
__device__ __forceinline__ bool poll_sum(const uint4* base, int64_t stride, int64_t slot, int W, int rank, uint32_t e,
                                         float4 own, const uint32_t* abort_word, float4& out) {
  uint4 a[TFDL_MAX_RANKS], b[TFDL_MAX_RANKS];
  uint32_t spins = 0;
  for (;;) {
#pragma unroll
    for (int r = 0; r < TFDL_MAX_RANKS; ++r)
      if (r < W && r != rank) {
        a[r] = ld_packet(base + r * stride + pkt_of(slot));
        b[r] = ld_packet(base + r * stride + pkt_of(slot) + 32);
      }
    bool ready = true;
#pragma unroll
    for (int r = 0; r < TFDL_MAX_RANKS; ++r)
      if (r < W && r != rank) ready = ready && a[r].y == e && a[r].w == e && b[r].y == e && b[r].w == e;
    if (ready) break;
    if ((++spins & 0x3FF) == 0) {
      if (abort_word && ld_relaxed_sys(abort_word) != 0) return false;
      if (spins > TFDL_PUSH_SPIN_LIMIT) __trap();
    }
  }
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < TFDL_MAX_RANKS; ++r)
    if (r < W) {
      if (r == rank) { g.x += own.x; g.y += own.y; g.z += own.z; g.w += own.w; }
      else {
        g.x += __uint_as_float(a[r].x); g.y += __uint_as_float(a[r].z);
        g.z += __uint_as_float(b[r].x); g.w += __uint_as_float(b[r].z);
      }
    }
  out = g;
  return true;
}
__device__ __forceinline__ bool poll_one(const uint4* region, int64_t slot, uint32_t e, const uint32_t* abort_word,
                                         float4& out) {
  const uint4* src = region + pkt_of(slot);
  uint32_t spins = 0;
  for (;;) {
    uint4 a = ld_packet(src), b = ld_packet(src + 32);
    if (a.y == e && a.w == e && b.y == e && b.w == e) {
      out = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
      return true;
    }
    if ((++spins & 0x3FF) == 0) {
      if (abort_word && ld_relaxed_sys(abort_word) != 0) return false;
      if (spins > TFDL_PUSH_SPIN_LIMIT) __trap();
    }
  }
}

__global__ void __launch_bounds__(256) fused_ar_push_kernel(TfdlFusedArParams p) {
  pdl_wait();
  const uint32_t e = *p.epoch + 1;
  const int par = p.parity >= 0 ? p.parity : (int)((e - 1) & 1);
  const int ipar = (int)((e - 1) & 1);  // inbox parity always follows the epoch (the gradient parity may be fixed)
  const TfdlHyper h = *p.hyper;
  const int W = p.world, rank = p.rank;
  const int64_t n4 = p.n >> 2;
  const int64_t region = region_packets(n4);  // packets per (parity, source)
  const float4* g4 = reinterpret_cast<const float4*>(p.grads[par][rank]);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  // ---- out: my gradients into every peer's inbox ----
  for (int64_t i = tid; i < n4; i += nth) {
    const float4 g = g4[i];
#pragma unroll
    for (int r = 0; r < TFDL_MAX_RANKS; ++r)
      if (r < W && r != rank)
        push_f4(reinterpret_cast<uint4*>(p.inbox[r]) + ((int64_t)ipar * W + rank) * region, i, g, e);
  }
  // ---- in: sum in rank order, update ----
  const uint4* mine = reinterpret_cast<const uint4*>(p.inbox[rank]) + (int64_t)ipar * W * region;
  float4* w4 = reinterpret_cast<float4*>(p.params[rank]);
  float4* v4 = reinterpret_cast<float4*>(p.mom);
  float4* u4 = reinterpret_cast<float4*>(p.mom2);
  const OptStep os = opt_step(h, e);
  uint2* s2 = reinterpret_cast<uint2*>(p.shadow[rank]);
  float4* z4 = reinterpret_cast<float4*>(p.grads[par ^ 1][rank]);
  for (int64_t i = tid; i < n4; i += nth) {
    float4 g;
    if (!poll_sum(mine, region, i, W, rank, e, g4[i], p.abort_word, g)) return;
    g.x *= h.scale; g.y *= h.scale; g.z *= h.scale; g.w *= h.scale;
    float4 w = opt_update(w4[i], g, v4 ? v4 + i : nullptr, u4 ? u4 + i : nullptr, h, os);
    w4[i] = w;
    if (s2) s2[i] = to_bf16x4(w);
    if (z4) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  finish_step(p, e);
}

// Two hops. Inbox layout per parity: [hop 0: W source regions][hop 1: W owner regions], each `per` float4 slots.
__global__ void __launch_bounds__(256) fused_ar_push2_kernel(TfdlFusedArParams p) {
  pdl_wait();
  const uint32_t e = *p.epoch + 1;
  const int par = p.parity >= 0 ? p.parity : (int)((e - 1) & 1);
  const int ipar = (int)((e - 1) & 1);
  const TfdlHyper h = *p.hyper;
  const int W = p.world, rank = p.rank;
  const int64_t n4 = p.n >> 2;
  const int64_t per = (n4 + W - 1) / W, region = region_packets(per);
  const int64_t lo = per * rank, hi = min(n4, lo + per);
  const float4* g4 = reinterpret_cast<const float4*>(p.grads[par][rank]);
  float4* w4 = reinterpret_cast<float4*>(p.params[rank]);
  float4* v4 = reinterpret_cast<float4*>(p.mom);
  float4* u4 = reinterpret_cast<float4*>(p.mom2);
  const OptStep os = opt_step(h, e);
  uint2* s2 = reinterpret_cast<uint2*>(p.shadow[rank]);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  const int64_t hop0 = ((int64_t)ipar * 2 + 0) * W * region, hop1 = ((int64_t)ipar * 2 + 1) * W * region;
  // ---- hop 0 out: slice j of my gradients to rank j ----
  for (int64_t i = tid; i < n4; i += nth) {
    const int owner = (int)(i / per);
    if (owner == rank) continue;
    push_f4(reinterpret_cast<uint4*>(p.inbox[owner]) + hop0 + (int64_t)rank * region, i - owner * per, g4[i], e);
  }
  // ---- my slice: sum, update, hop 1 out (the new weights) ----
  const uint4* in0 = reinterpret_cast<const uint4*>(p.inbox[rank]) + hop0;
  for (int64_t i = lo + tid; i < hi; i += nth) {
    float4 g;
    if (!poll_sum(in0, region, i - lo, W, rank, e, g4[i], p.abort_word, g)) return;
    g.x *= h.scale; g.y *= h.scale; g.z *= h.scale; g.w *= h.scale;
    float4 w = opt_update(w4[i], g, v4 ? v4 + i : nullptr, u4 ? u4 + i : nullptr, h, os);
    w4[i] = w;
    if (s2) s2[i] = to_bf16x4(w);
#pragma unroll
    for (int r = 0; r < TFDL_MAX_RANKS; ++r)
      if (r < W && r != rank)
        push_f4(reinterpret_cast<uint4*>(p.inbox[r]) + hop1 + (int64_t)rank * region, i - lo, w, e);
  }
  // ---- clear the other-parity gradients while the owners work ----
  float4* z4 = reinterpret_cast<float4*>(p.grads[par ^ 1][rank]);
  if (z4)
    for (int64_t i = tid; i < n4; i += nth) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  // ---- hop 1 in: everybody else's slices ----
  const uint4* in1 = reinterpret_cast<const uint4*>(p.inbox[rank]) + hop1;
  // Four elements per trip with all eight packet loads in flight together: a thread that polled its elements one after
  // the other paid one L2 round trip each.
  constexpr int U = 4;
  for (int64_t i0 = tid; i0 < n4; i0 += nth * U) {
    const uint4* src[U];
    bool need[U];
    float4 w[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * nth;
      need[u] = i < n4 && (int)(i / per) != rank;
      if (need[u]) {
        const int owner = (int)(i / per);
        src[u] = in1 + (int64_t)owner * region + pkt_of(i - owner * per);
      }
    }
    uint32_t spins = 0;
    for (;;) {
      uint4 a[U], b[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (need[u]) {
          a[u] = ld_packet(src[u]);
          b[u] = ld_packet(src[u] + 32);
        }
      bool pending = false;
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (need[u]) {
          if (a[u].y == e && a[u].w == e && b[u].y == e && b[u].w == e) {
            w[u] = make_float4(__uint_as_float(a[u].x), __uint_as_float(a[u].z), __uint_as_float(b[u].x), __uint_as_float(b[u].z));
            const int64_t i = i0 + u * nth;
            w4[i] = w[u];
            if (s2) s2[i] = to_bf16x4(w[u]);
            need[u] = false;
          } else {
            pending = true;
          }
        }
      if (!pending) break;
      if ((++spins & 0x3FF) == 0) {
        if (p.abort_word && ld_relaxed_sys(p.abort_word) != 0) return;
        if (spins > TFDL_PUSH_SPIN_LIMIT) __trap();
      }
    }
  }
  finish_step(p, e);
}

// Plain in-place all-reduce (sum) of a symmetric fp32 buffer - used by the bandwidth sweep and by
// callers that want the reduced gradients themselves (metrics, optimizers without a fused form).
// Two-shot over peer memory; `epoch`/flags as above but without parity (the caller separates calls).
template <bool MC>
__global__ void __launch_bounds__(512) allreduce_sum_kernel(TfdlFusedArParams p) {
  const uint32_t e = *p.epoch + 1;
  const int W = p.world;
  if (!block_barrier(p, 0, e)) return;
  const int64_t n4 = p.n >> 2;
  const int64_t per = (n4 + W - 1) / W;
  const int64_t lo = per * p.rank, hi = min(n4, lo + per);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  constexpr int U = MC ? 4 : 2;
  for (int64_t i0 = lo + tid; i0 < hi; i0 += nth * U) {
    float4 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * nth;
      if (i < hi) g[u] = MC ? mc_ld_reduce_f4(reinterpret_cast<const float4*>(p.mc_grads[0]) + i)
                            : gather_sum(p.grads[0], i, W, p.rank);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * nth;
      if (i >= hi) continue;
      if (MC) {
        mc_st_f4(reinterpret_cast<float4*>(p.mc_grads[0]) + i, g[u]);
      } else {
        // slice `rank` of every arena is read by this rank alone, so it can be overwritten right away
#pragma unroll
        for (int r = 0; r < TFDL_MAX_RANKS; ++r)
          if (r < W) {
            float4* dst = reinterpret_cast<float4*>(p.grads[0][r]) + i;
            if (r == p.rank) *dst = g[u]; else st_peer_f4(dst, g[u]);
          }
      }
    }
  }
  if ((p.n & 3) && p.rank == 0 && blockIdx.x == 0 && threadIdx.x == 0) {  // scalar tail, owned by rank 0
    for (int64_t i = n4 * 4; i < p.n; ++i) {
      const float g = gather_sum1(p.grads[0], i, W, 0);
      for (int r = 0; r < W; ++r)
        if (r == 0) p.grads[0][0][i] = g; else st_peer_f1(p.grads[0][r] + i, g);
    }
  }
  if (!block_barrier(p, p.pad_stride, e)) return;
  finish_step(p, e);
}

// Broadcast of a symmetric fp32 buffer from `root` (the chief's initial variable values, SURVEY K15 - TF sends them
// with CollectiveBcastSend/Recv): after a start barrier every other rank copies the root's buffer over NVLink with
// 16-byte peer loads; the end barrier keeps the root from touching its copy before everyone has read it.
// Buffers are p.grads[0][*], as for allreduce_sum_kernel.
__global__ void __launch_bounds__(512) broadcast_kernel(TfdlFusedArParams p, int root) {
  const uint32_t e = *p.epoch + 1;
  if (!block_barrier(p, 0, e)) return;
  if (p.rank != root) {
    const int64_t n4 = p.n >> 2;
    const float4* src = reinterpret_cast<const float4*>(p.grads[0][root]);
    float4* dst = reinterpret_cast<float4*>(p.grads[0][p.rank]);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
      dst[i] = ld_peer_f4(src + i);
    if ((p.n & 3) && blockIdx.x == 0 && threadIdx.x == 0)
      for (int64_t i = n4 * 4; i < p.n; ++i) p.grads[0][p.rank][i] = ld_peer_f1(p.grads[0][root] + i);
  }
  if (!block_barrier(p, p.pad_stride, e)) return;
  finish_step(p, e);
}

// Sum of up to 256 doubles across ranks (metric totals, SURVEY K16: sync-on-read variables are SUM-aggregated when
// read). Each rank's values sit in its own symmetric slot [parity][n]; every rank adds the W slots in rank order, so
// all ranks get bit-identical totals. Slots alternate by call parity, which makes a second barrier unnecessary.
__global__ void __launch_bounds__(256) allreduce_f64_kernel(TfdlFusedArParams p, double* __restrict__ out, int n) {
  const uint32_t e = *p.epoch + 1;
  const int par = (int)((e - 1) & 1);
  if (!block_barrier(p, 0, e)) return;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    double t = 0.0;
    for (int r = 0; r < p.world; ++r) {
      const double* slot = reinterpret_cast<const double*>(p.grads[0][r]) + (size_t)par * 256 + i;
      unsigned long long bits;
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(bits) : "l"(slot) : "memory");
      t += __longlong_as_double((long long)bits);
    }
    out[i] = t;
  }
  finish_step(p, e);
}

// fp32 -> bf16 shadow refresh (after set_weights / checkpoint load) and buffer clear
__global__ void cast_shadow_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    float4 w = *reinterpret_cast<const float4*>(src + i);
    *reinterpret_cast<uint2*>(dst + i) = to_bf16x4(w);
  } else {
    for (; i < n; ++i) dst[i] = __float2bfloat16(src[i]);
  }
}

}  // namespace tfdl

using namespace tfdl;

static int pick_grid(const TfdlFusedArParams* p, int64_t work4, int threads) {
  int64_t blocks = (work4 + threads - 1) / threads;
  int cap = p->max_blocks > 0 ? p->max_blocks : 148;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// Bytes of inbox every rank must provide for the push paths (peer-mapped, zero-initialised once).
extern "C" int64_t tfdl_fused_ar_inbox_bytes(int algo, int world, int64_t n) {
  const int64_t n4 = (n + 3) / 4;
  if (algo == TFDL_AR_PUSH) return 2 * (int64_t)world * tfdl::region_packets(n4) * 16;  // [parity][source][region]
  if (algo == TFDL_AR_PUSH2)                                                             // [parity][hop][rank][slice region]
    return 2 * 2 * (int64_t)world * tfdl::region_packets((n4 + world - 1) / world) * 16;
  return 0;
}

extern "C" int tfdl_fused_allreduce_sgd(const TfdlFusedArParams* p, int algo, void* stream) {
  TFDL_REQUIRE(p && p->world >= 1 && p->world <= TFDL_MAX_RANKS, "fused_allreduce_sgd: world %d out of range",
               p ? p->world : -1);
  TFDL_REQUIRE(p->n > 0, "fused_allreduce_sgd: n=%lld must be positive", (long long)p->n);
  TFDL_REQUIRE(p->n % 4 == 0 || algo == TFDL_AR_ONESHOT || p->world == 1,
               "fused_allreduce_sgd: n=%lld is not a multiple of 4: only the one-shot path takes a ragged tail "
               "(parameter arenas are padded to 16 bytes)", (long long)p->n);
  TFDL_REQUIRE(p->epoch && p->done && p->hyper, "fused_allreduce_sgd: epoch/done/hyper pointers are required");
  // (Adam, hyper.beta2 > 0, needs both slot arenas; the kernels cannot check device memory here, the Python side does)
  TFDL_REQUIRE(p->world == 1 || p->flags[p->rank], "fused_allreduce_sgd: flag pads missing");
  cudaStream_t s = (cudaStream_t)stream;
  const int threads = 256;
  const int64_t n4 = (p->n + 3) / 4;
  if (algo == TFDL_AR_ONESHOT || p->world == 1) {
    int grid = pick_grid(p, n4, threads);
    TFDL_CUDA_OK(launch_kernel_pdl(4, fused_ar_oneshot_kernel, dim3(grid), dim3(threads), 0, s, *p));
  } else if (algo == TFDL_AR_TWOSHOT) {
    int grid = pick_grid(p, (n4 + p->world - 1) / p->world, 512);
    fused_ar_sliced_kernel<false, 2><<<grid, 512, 0, s>>>(*p);
  } else if (algo == TFDL_AR_MULTIMEM) {
    TFDL_REQUIRE(p->mc_params && p->mc_grads[0] && p->mc_grads[1], "fused_allreduce_sgd: multicast pointers missing");
    int grid = pick_grid(p, (n4 + p->world - 1) / p->world, 512);
    fused_ar_sliced_kernel<true, 4><<<grid, 512, 0, s>>>(*p);
  } else if (algo == TFDL_AR_PUSH || algo == TFDL_AR_PUSH2) {
    TFDL_REQUIRE(p->inbox[p->rank] && p->inbox_bytes >= tfdl_fused_ar_inbox_bytes(algo, p->world, p->n),
                 "fused_allreduce_sgd: push needs an inbox of %lld bytes per rank (tfdl_fused_ar_inbox_bytes), got %lld",
                 (long long)tfdl_fused_ar_inbox_bytes(algo, p->world, p->n), (long long)p->inbox_bytes);
    if (algo == TFDL_AR_PUSH) {
      int grid = pick_grid(p, n4, threads);
      TFDL_CUDA_OK(launch_kernel_pdl(4, fused_ar_push_kernel, dim3(grid), dim3(threads), 0, s, *p));
    } else {
      int grid = pick_grid(p, n4, threads);  // hop 0 / hop 1-in walk the whole arena
      TFDL_CUDA_OK(launch_kernel_pdl(4, fused_ar_push2_kernel, dim3(grid), dim3(threads), 0, s, *p));
    }
  } else {
    TFDL_REQUIRE(false, "fused_allreduce_sgd: unknown algorithm %d", algo);
  }
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_allreduce_sum(const TfdlFusedArParams* p, int algo, void* stream) {
  TFDL_REQUIRE(p && p->world >= 2 && p->world <= TFDL_MAX_RANKS, "allreduce_sum: world %d out of range",
               p ? p->world : -1);
  TFDL_REQUIRE(p->n > 0, "allreduce_sum: n must be positive");
  cudaStream_t s = (cudaStream_t)stream;
  const int threads = 512;
  int grid = pick_grid(p, ((p->n + 3) / 4 + p->world - 1) / p->world, threads);
  if (algo == TFDL_AR_MULTIMEM) {
    TFDL_REQUIRE(p->mc_grads[0], "allreduce_sum: multicast pointer missing");
    allreduce_sum_kernel<true><<<grid, threads, 0, s>>>(*p);
  } else {
    allreduce_sum_kernel<false><<<grid, threads, 0, s>>>(*p);
  }
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream) {
  if (n <= 0) return 0;
  int threads = 256;
  int64_t blocks = ((n + 3) / 4 + threads - 1) / threads;
  cast_shadow_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int tfdl_enable_peer_access(int device, int peer) {
  int cur = 0;
  TFDL_CUDA_OK(cudaGetDevice(&cur));
  TFDL_CUDA_OK(cudaSetDevice(device));
  int can = 0;
  TFDL_CUDA_OK(cudaDeviceCanAccessPeer(&can, device, peer));
  TFDL_REQUIRE(can, "device %d cannot access peer %d", device, peer);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
  cudaSetDevice(cur);
  TFDL_CUDA_OK(e);
  return 0;
}

extern "C" int tfdl_broadcast_f32(const TfdlFusedArParams* p, int root, void* stream) {
  TFDL_REQUIRE(p && p->world >= 2 && p->world <= TFDL_MAX_RANKS && root >= 0 && root < p->world,
               "broadcast_f32: bad world / root");
  TFDL_REQUIRE(p->n > 0, "broadcast_f32: n must be positive");
  const int threads = 512;
  int grid = pick_grid(p, (p->n + 3) / 4, threads);
  broadcast_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(*p, root);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// p->grads[0][r] = rank r's slot buffer of 2 x 256 doubles (the caller writes its n values into slot [*p->epoch & 1]
// of its own buffer, stream-ordered before this call); out = n doubles of device memory.
extern "C" int tfdl_allreduce_small_f64(const TfdlFusedArParams* p, double* out, int n, void* stream) {
  TFDL_REQUIRE(p && p->world >= 2 && p->world <= TFDL_MAX_RANKS && out && n > 0 && n <= 256,
               "allreduce_small_f64: bad arguments (n = %d, at most 256 values)", n);
  allreduce_f64_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*p, out, n);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}

// Epoch totals without a host round trip: dst[i] = (double)metrics[i], optionally clearing the counters - dst is either a
// mapped pinned host slot (one GPU: the host reads it after the event) or this rank's slot of the symmetric buffer that
// tfdl_allreduce_small_f64 sums over the ranks (its `out` may itself be a mapped host pointer).
namespace tfdl {
__global__ void metrics_snapshot_kernel(float* metrics, int n, double* dst, int zero) {
  const int i = threadIdx.x;
  if (i < n) {
    dst[i] = (double)metrics[i];
    if (zero) metrics[i] = 0.f;
  }
}
}  // namespace tfdl
extern "C" int tfdl_metrics_snapshot(float* metrics, int n, double* dst, int zero, void* stream) {
  TFDL_REQUIRE(metrics && dst && n > 0 && n <= 256, "metrics_snapshot: bad arguments");
  tfdl::metrics_snapshot_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(metrics, n, dst, zero);
  TFDL_CUDA_OK(cudaGetLastError());
  return 0;
}
