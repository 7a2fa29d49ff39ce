// Native step executor for the reference CNN (what model.fit(...) at
// /root/reference/tf_dist_example.py:59 drives on a B200).
//
// TensorFlow's fit() dispatches a traced graph step by step from Python/C++ with an input iterator,
// an all-reduce and eight optimizer launches. Here one training step is a CUDA graph
// (9 compute kernels on two captured branches + the fused all-reduce/SGD kernel), there are two such
// graphs (gradient-buffer parity 0/1, see allreduce.cu), and tfdl_cnn_program_run() drives any number
// of steps from C++: per step one async H2D copy of the packed batch from pinned host memory on a
// copy stream, one graph launch, one async 8-byte D2H copy of (loss, accuracy). The host never
// waits for the device inside the loop.
#include <stdlib.h>
#include <vector>
#include "common.cuh"
#include "../../include/tfdl_kernels.h"

using namespace tfdl;

#define TFDL_RC(expr)        \
  do {                       \
    int _rc = (expr);        \
    if (_rc != 0) return _rc; \
  } while (0)

// ---- step timeline (diagnostics) -------------------------------------------------------------------------------
// tfdl_trace_enable(buf, counters) makes every step launched (or captured) afterwards drop one-thread marker
// kernels between its stages; marker `id` stores %globaltimer into buf[id*TFDL_TRACE_CAP + (counters[id]++ % CAP)].
// The markers serialise what programmatic dependent launch would overlap, so a traced step is a few microseconds
// slower than an untraced one: read the timeline for where the time goes, not for the step time.
static unsigned long long* g_trace_buf = nullptr;
static unsigned* g_trace_cnt = nullptr;
__global__ void trace_mark_kernel(unsigned long long* buf, unsigned* cnt, int id) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const unsigned k = cnt[id];
  cnt[id] = k + 1;
  buf[(size_t)id * TFDL_TRACE_CAP + (k % TFDL_TRACE_CAP)] = t;
}
static unsigned g_trace_mask = 0xFFu;
static inline void trace_mark(int id, cudaStream_t s) {
  if (g_trace_buf && ((g_trace_mask >> id) & 1u)) trace_mark_kernel<<<1, 1, 0, s>>>(g_trace_buf, g_trace_cnt, id);
}
extern "C" int tfdl_trace_enable(unsigned long long* buf, unsigned* counters) {
  g_trace_buf = buf;
  g_trace_cnt = counters;
  const char* m = getenv("TFDL_TRACE_MASK");  // bit i = marker i; fewer markers perturb the step less
  g_trace_mask = m ? (unsigned)strtoul(m, nullptr, 0) : 0xFFu;
  return 0;
}

extern "C" int tfdl_cnn_step_launch(const TfdlCnnStep* s, void* stream, void* side, void* ev_fork, void* ev_join) {
  return tfdl_cnn_step_launch_overlap(s, nullptr, 0, stream, side, nullptr, ev_fork, ev_join, nullptr, nullptr);
}

static int step_launch_impl(const TfdlCnnStep* s, const TfdlFusedArParams* early, int early_algo, void* stream, void* side,
                            void* ar_stream, void* ev_fork, void* ev_join, void* ev_ar_fork, void* ev_ar_join, bool join_ar);

extern "C" int tfdl_cnn_step_launch_overlap(const TfdlCnnStep* s, const TfdlFusedArParams* early, int early_algo,
                                            void* stream, void* side, void* ar_stream, void* ev_fork, void* ev_join,
                                            void* ev_ar_fork, void* ev_ar_join) {
  return step_launch_impl(s, early, early_algo, stream, side, ar_stream, ev_fork, ev_join, ev_ar_fork, ev_ar_join, true);
}

// join_ar = false leaves the early bucket's all-reduce un-joined: the caller enqueues the last bucket's all-reduce
// first (the two are independent and run side by side) and then waits on ev_ar_join itself.
static int step_launch_impl(const TfdlCnnStep* s, const TfdlFusedArParams* early, int early_algo, void* stream, void* side,
                            void* ar_stream, void* ev_fork, void* ev_join, void* ev_ar_fork, void* ev_ar_join, bool join_ar) {
  const TfdlCnnDims* d = &s->dims;
  cudaStream_t st = (cudaStream_t)stream, sd = (cudaStream_t)side;
  cudaEvent_t ef = (cudaEvent_t)ev_fork, ej = (cudaEvent_t)ev_join;
  const float* P = s->params;
  const __nv_bfloat16* H = (const __nv_bfloat16*)s->shadow;
  float* G = s->grads;
  const bool forked = sd != nullptr && sd != st && ef != nullptr && ej != nullptr;
  // ---- forward ----
  const bool tma = s->use_tma_conv != 0, fusedc = tma && s->use_fused_conv != 0;
  trace_mark(0, st);
  if (fusedc) {
    TFDL_RC(tfdl_cnn_conv12_fwd_tma(s->x, P + s->off_w1c, P + s->off_b1c, H + s->off_w2c, P + s->off_b2c, s->a1, s->idx1,
                                    s->a2, s->idx2, s->step_out, d, st));
  } else {
    TFDL_RC(tfdl_cnn_conv1_fwd(s->x, P + s->off_w1c, P + s->off_b1c, s->a1, s->idx1, s->step_out, d, st));
    if (tma) TFDL_RC(tfdl_cnn_conv2_fwd_tma(s->a1, H + s->off_w2c, P + s->off_b2c, s->a2, s->idx2, d, st));
    else TFDL_RC(tfdl_cnn_conv2_fwd(s->a1, H + s->off_w2c, P + s->off_b2c, s->a2, s->idx2, d, st));
  }
  cudaStream_t off = forked ? sd : st;
  trace_mark(1, st);
  if (s->use_dense_block) {
    // Dense(128)+Dense(NC)+loss forward and backward down to dy2, and all four dense gradients, in one kernel
    TFDL_RC(tfdl_cnn_dense_block(s->a2, s->idx2, H + s->off_w1d, H + s->off_w2d, P + s->off_b1d, P + s->off_b2d,
                                 s->labels, s->acc, s->acc_next, s->sync, s->dy2, G + s->off_w1d, G + s->off_b1d,
                                 G + s->off_w2d, G + s->off_b2d, s->metrics, s->step_out, nullptr, d, st));
  } else {
    // dense1: acc[s][u] += sum_k W1[u][k] a2[s][k]   (swap-AB: M = units, N = samples, split-K, transposed RED)
    TFDL_RC(tfdl_gemm_bf16_gather_ex(H + s->off_w1d, s->a2, s->acc, nullptr, nullptr, d->D1, d->B, d->F, d->F, d->F,
                                     d->D1, 0, 0, 0, s->dense1_ksplit > 0 ? s->dense1_ksplit : 1, 0, 64, 1, st));
    TFDL_RC(tfdl_cnn_head(s->acc, P + s->off_b1d, P + s->off_w2d, P + s->off_b2d, s->labels, s->dh, G + s->off_w2d,
                          G + s->off_b2d, G + s->off_b1d, s->metrics, s->step_out, d, st));
    if (forked) {
      TFDL_CUDA_OK(cudaEventRecord(ef, st));
      TFDL_CUDA_OK(cudaStreamWaitEvent(sd, ef, 0));
    }
    // dW1[u][f] = sum_s a2[s][f] dh[s][u]: both operands MN-major; rows = features so that the transposed
    // store walks f (the contiguous index of dW1) across the lanes of a warp (off the critical path)
    TFDL_RC(tfdl_gemm_bf16_gather_ex(s->a2, s->dh, G + s->off_w1d, nullptr, nullptr, d->F, d->D1, d->B, d->F, d->D1,
                                     d->F, 1, 1, 0, 1, 64, 0, 1, off));
    TFDL_RC(tfdl_cnn_dense1_dgrad(s->dh, H + s->off_w1d, s->a2, s->idx2, s->dy2, d, st));
  }
  const bool overlap = early != nullptr && s->use_dense_block && ar_stream != nullptr && ev_ar_fork && ev_ar_join;
  trace_mark(2, st);
  if (overlap) {
    // the dense-layer gradients (91% of the arena) are final: reduce + apply them now, on their own stream,
    // while the conv backward kernels run (bucketed overlap; README.md:67 only requires completion per batch)
    TFDL_CUDA_OK(cudaEventRecord((cudaEvent_t)ev_ar_fork, st));
    TFDL_CUDA_OK(cudaStreamWaitEvent((cudaStream_t)ar_stream, (cudaEvent_t)ev_ar_fork, 0));
    TFDL_RC(tfdl_fused_allreduce_sgd(early, early_algo, ar_stream));
    trace_mark(5, (cudaStream_t)ar_stream);
    TFDL_CUDA_OK(cudaEventRecord((cudaEvent_t)ev_ar_join, (cudaStream_t)ar_stream));
  }
  if (forked) {
    // conv2 wgrad needs dy2 (main branch) but nothing downstream needs it before the all-reduce
    TFDL_CUDA_OK(cudaEventRecord(ef, st));
    TFDL_CUDA_OK(cudaStreamWaitEvent(sd, ef, 0));
  }
  if (fusedc) {
    TFDL_RC(tfdl_cnn_conv2_wgrad_tma(s->a1, s->dy2, G + s->off_w2c, G + s->off_b2c, d, off));
    TFDL_RC(tfdl_cnn_conv21_bwd_tma(s->dy2, H + s->off_w2c, s->a1, s->idx1, s->x, G + s->off_w1c, G + s->off_b1c, nullptr,
                                    d, st));
  } else if (tma) {
    TFDL_RC(tfdl_cnn_conv2_wgrad_tma(s->a1, s->dy2, G + s->off_w2c, G + s->off_b2c, d, off));
    TFDL_RC(tfdl_cnn_conv2_dgrad_tma(s->dy2, H + s->off_w2c, s->a1, s->g1, d, st));
  } else {
    TFDL_RC(tfdl_cnn_conv2_wgrad(s->a1, s->dy2, G + s->off_w2c, G + s->off_b2c, d, off));
    TFDL_RC(tfdl_cnn_conv2_dgrad(s->dy2, H + s->off_w2c, s->a1, s->g1, d, st));
  }
  if (!fusedc) TFDL_RC(tfdl_cnn_conv1_wgrad(s->x, s->g1, s->idx1, G + s->off_w1c, G + s->off_b1c, d, st));
  trace_mark(4, st);
  if (forked) {
    trace_mark(3, sd);
    TFDL_CUDA_OK(cudaEventRecord(ej, sd));
    TFDL_CUDA_OK(cudaStreamWaitEvent(st, ej, 0));
  }
  if (overlap && join_ar) TFDL_CUDA_OK(cudaStreamWaitEvent(st, (cudaEvent_t)ev_ar_join, 0));
  return 0;
}

#define TFDL_CHUNK_MAX 16
struct TfdlStageTable {
  const void* x[TFDL_CHUNK_MAX + 1];  // [j], j = 1..K: images of step j of this chunk (K = first step of the next chunk)
  const void* y[TFDL_CHUNK_MAX + 1];
  float* out[TFDL_CHUNK_MAX];         // mapped pinned host slots for (loss, accuracy) of step j
  float* prev_out;                    // ... of the step before this chunk
};

struct TfdlCnnProgram {
  TfdlCnnStep steps[2];
  TfdlFusedArParams ar[2];
  TfdlFusedArParams ar_early[2];
  bool has_early = false;
  int algo_early = 0;
  cudaStream_t ar_stream = nullptr;
  cudaEvent_t ev_ar_fork = nullptr, ev_ar_join = nullptr;
  int algo = 0;
  int use_graph = 1;
  cudaGraphExec_t exec[2] = {nullptr, nullptr};
  cudaStream_t side = nullptr, copy = nullptr, out = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_used[2] = {nullptr, nullptr}, ev_step[2] = {nullptr, nullptr},
              ev_out[2] = {nullptr, nullptr};
  int64_t steps_done = 0;
  int device = 0;
  // chunked launches (see launch_chunk)
  int chunk_max = 16;
  cudaGraphExec_t chunk_exec[TFDL_CHUNK_MAX + 1][2] = {};
  TfdlStageTable* table_dev = nullptr;
  cudaEvent_t ev_stage_fork = nullptr, ev_stage_join = nullptr;
};

static int enqueue_step(TfdlCnnProgram* p, int par, cudaStream_t st) {
  TFDL_RC(step_launch_impl(&p->steps[par], p->has_early ? &p->ar_early[par] : nullptr, p->algo_early, st, p->side,
                           p->ar_stream, p->ev_fork, p->ev_join, p->ev_ar_fork, p->ev_ar_join, false));
  TFDL_RC(tfdl_fused_allreduce_sgd(&p->ar[par], p->algo, st));
  trace_mark(6, st);
  // the dense bucket's all-reduce (own stream, own flags) may still be running next to the conv bucket's: join now
  if (p->has_early && p->steps[par].use_dense_block) TFDL_CUDA_OK(cudaStreamWaitEvent(st, p->ev_ar_join, 0));
  return 0;
}

static int ensure_graph(TfdlCnnProgram* p, int par, cudaStream_t st) {
  if (p->exec[par]) return 0;
  // cudaFuncSetAttribute calls inside the launch helpers are not stream operations, so they are legal
  // during capture; run the step once eagerly first anyway? No: that would advance training. Capture only.
  cudaGraph_t graph = nullptr;
  TFDL_CUDA_OK(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
  int rc = enqueue_step(p, par, st);
  cudaError_t ce = cudaStreamEndCapture(st, &graph);
  if (rc != 0) {
    if (graph) cudaGraphDestroy(graph);
    return rc;
  }
  TFDL_CUDA_OK(ce);
  TFDL_CUDA_OK(cudaGraphInstantiate(&p->exec[par], graph, 0));
  TFDL_CUDA_OK(cudaGraphDestroy(graph));
  return 0;
}

extern "C" int tfdl_cnn_program_create(const TfdlCnnStep steps[2], const TfdlFusedArParams ar[2], int algo,
                                       int use_graph, TfdlCnnProgram** out) {
  return tfdl_cnn_program_create2(steps, ar, algo, nullptr, 0, use_graph, out);
}

extern "C" int tfdl_cnn_program_create2(const TfdlCnnStep steps[2], const TfdlFusedArParams ar[2], int algo,
                                        const TfdlFusedArParams ar_early[2], int algo_early, int use_graph,
                                        TfdlCnnProgram** out) {
  TFDL_REQUIRE(steps && ar && out, "cnn_program_create: null argument");
  auto* p = new TfdlCnnProgram();
  if (ar_early) {
    p->has_early = true;
    p->algo_early = algo_early;
    for (int i = 0; i < 2; ++i) {
      p->ar_early[i] = ar_early[i];
      p->ar_early[i].parity = i;
    }
    TFDL_CUDA_OK(cudaStreamCreateWithFlags(&p->ar_stream, cudaStreamNonBlocking));
    TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_ar_fork, cudaEventDisableTiming));
    TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_ar_join, cudaEventDisableTiming));
  }
  for (int i = 0; i < 2; ++i) {
    p->steps[i] = steps[i];
    p->ar[i] = ar[i];
    p->ar[i].parity = i;
  }
  p->algo = algo;
  p->use_graph = use_graph;
  TFDL_CUDA_OK(cudaGetDevice(&p->device));
  TFDL_CUDA_OK(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
  TFDL_CUDA_OK(cudaStreamCreateWithFlags(&p->copy, cudaStreamNonBlocking));
  TFDL_CUDA_OK(cudaStreamCreateWithFlags(&p->out, cudaStreamNonBlocking));
  TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
  TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
  TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_stage_fork, cudaEventDisableTiming));
  TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_stage_join, cudaEventDisableTiming));
  TFDL_CUDA_OK(cudaMalloc(&p->table_dev, sizeof(TfdlStageTable)));
  TFDL_CUDA_OK(cudaMemset(p->table_dev, 0, sizeof(TfdlStageTable)));
  if (const char* e = getenv("TFDL_STEPS_PER_GRAPH")) {  // 1, 2, 4, 8 or 16 steps per graph launch
    int k = atoi(e);
    p->chunk_max = k >= 16 ? 16 : k >= 8 ? 8 : k >= 4 ? 4 : k >= 2 ? 2 : 1;
  }
  for (int i = 0; i < 2; ++i) {
    TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_in[i], cudaEventDisableTiming));
    TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_used[i], cudaEventDisableTiming));
    TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_step[i], cudaEventDisableTiming));
    TFDL_CUDA_OK(cudaEventCreateWithFlags(&p->ev_out[i], cudaEventDisableTiming));
  }
  *out = p;
  return 0;
}

extern "C" int tfdl_cnn_program_destroy(TfdlCnnProgram* p) {
  if (!p) return 0;
  for (int i = 0; i < 2; ++i) {
    if (p->exec[i]) cudaGraphExecDestroy(p->exec[i]);
    if (p->ev_in[i]) cudaEventDestroy(p->ev_in[i]);
    if (p->ev_used[i]) cudaEventDestroy(p->ev_used[i]);
    if (p->ev_step[i]) cudaEventDestroy(p->ev_step[i]);
    if (p->ev_out[i]) cudaEventDestroy(p->ev_out[i]);
  }
  for (int k = 0; k <= TFDL_CHUNK_MAX; ++k)
    for (int i = 0; i < 2; ++i)
      if (p->chunk_exec[k][i]) cudaGraphExecDestroy(p->chunk_exec[k][i]);
  if (p->table_dev) cudaFree(p->table_dev);
  if (p->ev_stage_fork) cudaEventDestroy(p->ev_stage_fork);
  if (p->ev_stage_join) cudaEventDestroy(p->ev_stage_join);
  if (p->ev_ar_fork) cudaEventDestroy(p->ev_ar_fork);
  if (p->ev_ar_join) cudaEventDestroy(p->ev_ar_join);
  if (p->ar_stream) cudaStreamDestroy(p->ar_stream);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->side) cudaStreamDestroy(p->side);
  if (p->copy) cudaStreamDestroy(p->copy);
  if (p->out) cudaStreamDestroy(p->out);
  delete p;
  return 0;
}

static int launch_one(TfdlCnnProgram* p, cudaStream_t st) {
  const int par = (int)(p->steps_done & 1);
  if (p->use_graph) {
    TFDL_RC(ensure_graph(p, par, st));
    TFDL_CUDA_OK(cudaGraphLaunch(p->exec[par], st));
  } else {
    TFDL_RC(enqueue_step(p, par, st));
  }
  p->steps_done++;
  return 0;
}

extern "C" int tfdl_cnn_program_step(TfdlCnnProgram* p, void* stream) { return launch_one(p, (cudaStream_t)stream); }

// ---- chunks of steps as ONE graph ------------------------------------------------------------------------------
// A graph boundary is not free once kernels write peer memory: the stream cannot start the next launch before every
// NVLink write of the last all-reduce has been acknowledged (measured: +3.7 us per boundary at 2 GPUs, +3 us at one).
// So the executor launches up to TFDL_CHUNK_MAX steps per graph. Inputs cannot be memcpy nodes then (their source
// changes every launch), so each step's batch is brought in by a small copy kernel that takes its source pointer from a
// device-side table the host rewrites before every launch (stream-ordered, 300 bytes). The copy for step j+1 is a
// parallel branch of step j - software-pipelined across graphs: the last branch of a graph fetches the first batch of
// the next one - and the same kernel hands the (loss, accuracy) pair of step j-1 to the host through mapped pinned
// memory. Sources may be pinned host memory (read over PCIe by the kernel) or device memory.
__global__ void __launch_bounds__(256) stage_inputs_kernel(const TfdlStageTable* __restrict__ t, int slot, char* dst,
                                                           unsigned x_bytes, unsigned y_bytes,
                                                           const float* step_out_prev, int out_slot) {
  if (blockIdx.x == 0 && threadIdx.x == 0 && step_out_prev) {
    float* o = out_slot < 0 ? t->prev_out : t->out[out_slot];
    if (o) {
      o[0] = step_out_prev[0];
      o[1] = step_out_prev[1];
    }
  }
  const char* sx = static_cast<const char*>(t->x[slot]);
  const char* sy = static_cast<const char*>(t->y[slot]);
  if (!sx) return;
  const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  if (((reinterpret_cast<uintptr_t>(sx) | reinterpret_cast<uintptr_t>(sy) | x_bytes | y_bytes) & 15) == 0) {
    const uint4* a = reinterpret_cast<const uint4*>(sx);
    uint4* d = reinterpret_cast<uint4*>(dst);
    const unsigned nx = x_bytes >> 4, ny = y_bytes >> 4;
    for (unsigned i = tid; i < nx; i += nth * 8) {  // eight 16-byte loads in flight per thread (PCIe latency) ...
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (i + u * nth < nx) v[u] = __ldcs(a + i + u * nth);
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (i + u * nth < nx) d[i + u * nth] = v[u];
    }
    const uint4* b = reinterpret_cast<const uint4*>(sy);
    for (unsigned i = tid; i < ny; i += nth) d[nx + i] = __ldcs(b + i);
  } else {
    const unsigned* a = reinterpret_cast<const unsigned*>(sx);
    unsigned* d = reinterpret_cast<unsigned*>(dst);
    for (unsigned i = tid; i < (x_bytes >> 2); i += nth) d[i] = a[i];
    const unsigned* b = reinterpret_cast<const unsigned*>(sy);
    for (unsigned i = tid; i < (y_bytes >> 2); i += nth) d[(x_bytes >> 2) + i] = b[i];
  }
}

static int enqueue_chunk(TfdlCnnProgram* p, int K, int par0, cudaStream_t st) {
  const TfdlCnnDims& d = p->steps[0].dims;
  const unsigned x_bytes = (unsigned)((size_t)d.B * d.H0 * d.W0 * sizeof(float)), y_bytes = (unsigned)(d.B * sizeof(int));
  for (int j = 0; j < K; ++j) {
    const int par = (par0 + j) & 1;
    // branch: fetch the batch of step j+1 (other parity's input record), export the result of step j-1
    TFDL_CUDA_OK(cudaEventRecord(p->ev_stage_fork, st));
    TFDL_CUDA_OK(cudaStreamWaitEvent(p->copy, p->ev_stage_fork, 0));
    // ... from only 8 CTAs: the branch runs next to the step's compute kernels (the dense block needs its 25 CTAs
    // co-resident) and must not take SMs from them; 8 x 256 x 8 x 16 B = 256 KB in flight covers a 200 KB batch.
    stage_inputs_kernel<<<8, 256, 0, p->copy>>>(p->table_dev, j + 1,
                                                 reinterpret_cast<char*>(const_cast<float*>(p->steps[par ^ 1].x)), x_bytes,
                                                 y_bytes, p->steps[par ^ 1].step_out, j - 1);
    TFDL_CUDA_OK(cudaGetLastError());
    TFDL_CUDA_OK(cudaEventRecord(p->ev_stage_join, p->copy));
    TFDL_RC(enqueue_step(p, par, st));
    TFDL_CUDA_OK(cudaStreamWaitEvent(st, p->ev_stage_join, 0));
  }
  return 0;
}

static int ensure_chunk_graph(TfdlCnnProgram* p, int K, int par0, cudaStream_t st) {
  cudaGraphExec_t& ex = p->chunk_exec[K][par0];
  if (ex) return 0;
  cudaGraph_t graph = nullptr;
  TFDL_CUDA_OK(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
  int rc = enqueue_chunk(p, K, par0, st);
  cudaError_t ce = cudaStreamEndCapture(st, &graph);
  if (rc != 0) {
    if (graph) cudaGraphDestroy(graph);
    return rc;
  }
  TFDL_CUDA_OK(ce);
  TFDL_CUDA_OK(cudaGraphInstantiate(&ex, graph, 0));
  TFDL_CUDA_OK(cudaGraphDestroy(graph));
  TFDL_CUDA_OK(cudaGraphUpload(ex, st));
  return 0;
}

// Capture, instantiate and upload every graph the executor can launch (1/2/4/8 steps x both parities) so that
// no capture happens inside a training run: a 20-step fit must cost 20 steps. Nothing is executed.
extern "C" int tfdl_cnn_program_prepare(TfdlCnnProgram* p, void* stream) {
  TFDL_REQUIRE(p, "cnn_program_prepare: null program");
  if (!p->use_graph) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  for (int K = 1; K <= p->chunk_max; K <<= 1)
    for (int par = 0; par < 2; ++par) TFDL_RC(ensure_chunk_graph(p, K, par, st));
  return 0;
}

static int launch_chunk(TfdlCnnProgram* p, int K, cudaStream_t st) {
  const int par0 = (int)(p->steps_done & 1);
  if (p->use_graph) {
    TFDL_RC(ensure_chunk_graph(p, K, par0, st));
    cudaGraphExec_t& ex = p->chunk_exec[K][par0];
    TFDL_CUDA_OK(cudaGraphLaunch(ex, st));
  } else {
    TFDL_RC(enqueue_chunk(p, K, par0, st));
  }
  p->steps_done += K;
  return 0;
}

// Shared loop of the fed entry points. `src_of(i, &x_src, &y_src)` names step i's images and labels (pinned host or
// device memory). Everything is ordered on `st`; the caller synchronises it.
template <class SrcFn>
static int run_host_fed(TfdlCnnProgram* p, int64_t n_steps, float* host_out, cudaStream_t st, SrcFn src_of) {
  const TfdlCnnDims& d = p->steps[0].dims;
  const size_t x_bytes = (size_t)d.B * d.H0 * d.W0 * sizeof(float), y_bytes = (size_t)d.B * sizeof(int);
  TFDL_REQUIRE(reinterpret_cast<const char*>(p->steps[0].labels) == reinterpret_cast<const char*>(p->steps[0].x) + x_bytes &&
               reinterpret_cast<const char*>(p->steps[1].labels) == reinterpret_cast<const char*>(p->steps[1].x) + x_bytes,
               "cnn_program_run: each step's labels must directly follow its images on the device (one input record)");
  int64_t i = 0;
  bool prefetched = false;  // the input of step i is already in its record (fetched by the previous chunk)
  while (i < n_steps) {
    const int64_t left = n_steps - i;
    int K = 1;
    for (int k = p->chunk_max; k >= 2; k >>= 1)
      if (left >= k) { K = k; break; }
    const int par0 = (int)(p->steps_done & 1);
    const char *xs = nullptr, *ys = nullptr;
    if (!prefetched) {
      src_of(i, &xs, &ys);
      char* dst = reinterpret_cast<char*>(const_cast<float*>(p->steps[par0].x));
      if (ys == xs + x_bytes) {
        TFDL_CUDA_OK(cudaMemcpyAsync(dst, xs, x_bytes + y_bytes, cudaMemcpyDefault, st));
      } else {
        TFDL_CUDA_OK(cudaMemcpyAsync(dst, xs, x_bytes, cudaMemcpyDefault, st));
        TFDL_CUDA_OK(cudaMemcpyAsync(dst + x_bytes, ys, y_bytes, cudaMemcpyDefault, st));
      }
    }
    TfdlStageTable tab{};
    for (int j = 1; j <= K; ++j) {
      if (i + j < n_steps) {
        src_of(i + j, &xs, &ys);
        tab.x[j] = xs;
        tab.y[j] = ys;
      }  // else: nothing to fetch (a null source makes the copy kernel a no-op)
    }
    for (int j = 0; j < K; ++j) tab.out[j] = host_out ? host_out + 2 * (i + j) : nullptr;
    tab.prev_out = (host_out && i > 0) ? host_out + 2 * (i - 1) : nullptr;
    // pageable source: the runtime stages these few hundred bytes before the call returns, `tab` can go out of scope
    TFDL_CUDA_OK(cudaMemcpyAsync(p->table_dev, &tab, sizeof(tab), cudaMemcpyHostToDevice, st));
    TFDL_RC(launch_chunk(p, K, st));
    i += K;
    prefetched = i < n_steps;
  }
  if (host_out && n_steps > 0) {  // the last step's pair (every earlier one was exported by the step after it)
    const int par_last = (int)((p->steps_done - 1) & 1);
    TFDL_CUDA_OK(cudaMemcpyAsync(host_out + 2 * (n_steps - 1), p->steps[par_last].step_out, 2 * sizeof(float),
                                 cudaMemcpyDeviceToHost, st));
  }
  return 0;
}

extern "C" int tfdl_cnn_program_run(TfdlCnnProgram* p, const void* host_batches, int64_t n_records, int64_t n_steps,
                                    float* host_out, void* stream) {
  TFDL_REQUIRE(p && host_batches && n_records > 0 && n_steps >= 0, "cnn_program_run: bad arguments");
  const TfdlCnnDims& d = p->steps[0].dims;
  const size_t x_bytes = (size_t)d.B * d.H0 * d.W0 * sizeof(float), rec = x_bytes + (size_t)d.B * sizeof(int);
  const char* base = static_cast<const char*>(host_batches);
  return run_host_fed(p, n_steps, host_out, (cudaStream_t)stream, [=](int64_t i, const char** xs, const char** ys) {
    *xs = base + (size_t)(i % n_records) * rec;
    *ys = *xs + x_bytes;
  });
}

// Zero-staging variant: step i reads its images from xs[i] and its int32 labels from ys[i], both in page-locked
// host memory (cudaHostRegister'ed dataset arrays, see tfdl_host_register) - no copy into a record ring.
extern "C" int tfdl_cnn_program_run_ptrs(TfdlCnnProgram* p, const void* const* xs, const void* const* ys,
                                         int64_t n_steps, float* host_out, void* stream) {
  TFDL_REQUIRE(p && xs && ys && n_steps >= 0, "cnn_program_run_ptrs: bad arguments");
  return run_host_fed(p, n_steps, host_out, (cudaStream_t)stream, [=](int64_t i, const char** x, const char** y) {
    *x = static_cast<const char*>(xs[i]);
    *y = static_cast<const char*>(ys[i]);
  });
}

// Page-lock a host range that the caller owns (a dataset array) so the executor can DMA from it directly.
// Returns 0 also when the range is already page-locked.
extern "C" int tfdl_host_register(void* ptr, uint64_t bytes) {
  TFDL_REQUIRE(ptr && bytes > 0, "host_register: empty range");
  cudaError_t e = cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterDefault);
  if (e == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();
    return 0;
  }
  TFDL_CUDA_OK(e);
  return 0;
}
extern "C" int tfdl_host_unregister(void* ptr) {
  cudaError_t e = cudaHostUnregister(ptr);
  if (e == cudaErrorHostMemoryNotRegistered) {
    cudaGetLastError();
    return 0;
  }
  TFDL_CUDA_OK(e);
  return 0;
}

// Same loop with the batches already resident on the device (a pool larger than L2 keeps the inputs coming from
// HBM): per step one D2D copy of that step's record into this parity's input buffers, issued on the copy stream so
// that it overlaps the previous step exactly like the host-fed path, and one graph launch. This is the
// kernel-only measurement path of bench.py.
extern "C" int tfdl_cnn_program_run_device(TfdlCnnProgram* p, const void* dev_batches, int64_t n_records,
                                           int64_t n_steps, void* stream) {
  TFDL_REQUIRE(p && dev_batches && n_records > 0 && n_steps >= 0, "cnn_program_run_device: bad arguments");
  const TfdlCnnDims& d = p->steps[0].dims;
  const size_t x_bytes = (size_t)d.B * d.H0 * d.W0 * sizeof(float), rec = x_bytes + (size_t)d.B * sizeof(int);
  const char* base = static_cast<const char*>(dev_batches);
  return run_host_fed(p, n_steps, nullptr, (cudaStream_t)stream, [=](int64_t i, const char** xs, const char** ys) {
    *xs = base + (size_t)(i % n_records) * rec;
    *ys = *xs + x_bytes;
  });
}

extern "C" int64_t tfdl_cnn_program_steps_done(const TfdlCnnProgram* p) { return p ? p->steps_done : 0; }
extern "C" int64_t tfdl_cnn_program_launches_per_step(const TfdlCnnProgram* p) {
  if (!p) return 0;
  const TfdlCnnStep& s = p->steps[0];
  // input staging, conv1_fwd, conv2_fwd, [dense block | dense1_fwd, head, dW1, dense1_dgrad], conv2_wgrad, conv2_dgrad,
  // conv1_wgrad, AR(s)
  const bool fusedc = s.use_tma_conv && s.use_fused_conv;
  return 1 + (fusedc ? 1 : 2) + (s.use_dense_block ? 1 : 4) + (fusedc ? 2 : 3) + 1 + (p->has_early ? 1 : 0);
}
