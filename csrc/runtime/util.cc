// Error string, crc32c (software slicing-by-8 with an SSE4.2 fast path), data-pipeline helpers.
//
// Data helpers stand in for tf.data's C++ kernels behind the reference's
// `.shuffle(BUFFER_SIZE)` / `.batch(GLOBAL_BATCH_SIZE)` (/root/reference/tf_dist_example.py:31,33):
// the shuffle buffer runs on element *indices* and batches are assembled by one gather straight
// into a (pinned) staging buffer, optionally fusing the uint8 -> float32 /255 scale of
// tf_dist_example.py:22-25.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <condition_variable>
#include <deque>
#include <memory>
#include <vector>
#include <functional>
#include <mutex>
#include <thread>
#include "rt_common.h"
#include "../../include/tfdl_runtime.h"

#if defined(__x86_64__)
#include <nmmintrin.h>
#include <cpuid.h>
#endif

namespace tfdl_rt {

static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

// ---------------------------------------------------------------------------------------------
// crc32c (Castagnoli, reflected polynomial 0x82F63B78)
// ---------------------------------------------------------------------------------------------
static uint32_t g_tab[8][256];
static bool g_tab_init = [] {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c >> 1) ^ (0x82F63B78u & (0u - (c & 1)));
    g_tab[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t) g_tab[t][i] = (g_tab[t - 1][i] >> 8) ^ g_tab[0][g_tab[t - 1][i] & 0xFF];
  return true;
}();

#if defined(__x86_64__)
static bool has_sse42() {
  unsigned a, b, c, d;
  if (!__get_cpuid(1, &a, &b, &c, &d)) return false;
  return (c & bit_SSE4_2) != 0;
}
__attribute__((target("sse4.2"))) static uint32_t crc_hw(uint32_t crc, const uint8_t* p, size_t n) {
  uint64_t c = crc;
  while (n >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8;
    n -= 8;
  }
  while (n--) c = _mm_crc32_u8((uint32_t)c, *p++);
  return (uint32_t)c;
}
#endif

uint32_t crc32c_extend(uint32_t crc, const void* data, size_t n) {
  (void)g_tab_init;
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t c = crc ^ 0xFFFFFFFFu;
#if defined(__x86_64__)
  static const bool hw = has_sse42();
  if (hw) return crc_hw(c, p, n) ^ 0xFFFFFFFFu;
#endif
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = g_tab[7][lo & 0xFF] ^ g_tab[6][(lo >> 8) & 0xFF] ^ g_tab[5][(lo >> 16) & 0xFF] ^ g_tab[4][lo >> 24] ^
        g_tab[3][hi & 0xFF] ^ g_tab[2][(hi >> 8) & 0xFF] ^ g_tab[1][(hi >> 16) & 0xFF] ^ g_tab[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = g_tab[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

}  // namespace tfdl_rt

using namespace tfdl_rt;

extern "C" const char* tfdl_rt_last_error(void) { return last_error(); }
extern "C" int tfdl_rt_abi_version(void) { return 1; }
extern "C" uint32_t tfdl_rt_crc32c(uint32_t crc, const void* data, uint64_t n) { return crc32c_extend(crc, data, n); }
extern "C" uint32_t tfdl_rt_crc32c_mask(uint32_t crc) { return crc32c_mask(crc); }

// tf.data shuffle-buffer semantics on an index stream: hold `buffer` indices, emit a uniformly
// random slot, refill it from the stream, drain the buffer in random order at the end.
extern "C" int tfdl_rt_shuffle_indices(const int64_t* src, int64_t n, int64_t buffer, uint64_t seed, int64_t* out) {
  if (n < 0 || buffer < 1) {
    set_error("shuffle_indices: bad arguments n=%lld buffer=%lld", (long long)n, (long long)buffer);
    return -1;
  }
  Rng rng(seed);
  if (buffer >= n) {
    std::copy(src, src + n, out);
    for (int64_t i = n - 1; i > 0; --i) std::swap(out[i], out[rng.below(i + 1)]);
    return 0;
  }
  std::vector<int64_t> buf(src, src + buffer);
  int64_t o = 0;
  for (int64_t t = buffer; t < n; ++t) {
    uint64_t j = rng.below(buffer);
    out[o++] = buf[j];
    buf[j] = src[t];
  }
  for (int64_t i = buffer - 1; i > 0; --i) std::swap(buf[i], buf[rng.below(i + 1)]);
  std::copy(buf.begin(), buf.end(), out + o);
  return 0;
}

// A small persistent worker pool: the input pipeline calls into here once per chunk of steps (every ~0.8 ms of
// device time for the reference CNN), and creating + joining threads per call costs more than the copy they do.
//
// Jobs are split into many small parts that any thread claims with one atomic add. The posting thread never waits
// for a worker to wake up: in run() (and in the wait of an asynchronous job) it claims parts itself, so a job costs
// at most its single-threaded time and workers that wake late simply find nothing left. After a job the workers
// poll for the next one for a short while before they sleep: fit() posts its gathers back to back, and a
// condition-variable wake-up (30-60 us) is as long as the gather of four MNIST batches.
namespace {
inline void cpu_relax() {
#if defined(__x86_64__)
  __builtin_ia32_pause();
#else
  std::this_thread::yield();
#endif
}
inline int64_t now_us() {
  return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Job {
  std::function<void(int)> fn;  // owned: a worker may look at a job after its poster has returned
  int parts = 0;
  std::atomic<int> next{0}, pending{0};
  void help() {
    for (;;) {
      const int part = next.fetch_add(1, std::memory_order_relaxed);
      if (part >= parts) return;
      fn(part);
      pending.fetch_sub(1, std::memory_order_acq_rel);
    }
  }
  bool claimed() const { return next.load(std::memory_order_relaxed) >= parts; }
  bool done() const { return pending.load(std::memory_order_acquire) == 0; }
  void wait_done() {  // parts are short: spin, then start yielding
    const int64_t t0 = now_us();
    while (!done()) {
      if (now_us() - t0 < 200) cpu_relax();
      else std::this_thread::yield();
    }
  }
};

class Pool {
 public:
  explicit Pool(int n) {
    if (const char* e = getenv("TFDL_IO_SPIN_US")) spin_us_ = std::max(0, atoi(e));
    for (int i = 0; i < n; ++i) workers_.emplace_back([this] { loop(); });
  }
  ~Pool() {
    {
      std::lock_guard<std::mutex> g(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  int size() const { return (int)workers_.size(); }
  // hands the job to the workers and returns at once; the poster joins in with job->help()
  void post(const std::shared_ptr<Job>& job) {
    job->pending.store(job->parts, std::memory_order_relaxed);
    {
      std::lock_guard<std::mutex> g(mu_);
      jobs_.push_back(job);
      posted_.fetch_add(1, std::memory_order_release);
    }
    cv_.notify_all();
  }
  // runs fn(part) for part in [0, parts) on the workers + the calling thread; returns when all are done
  void run(int parts, std::function<void(int)> fn) {
    if (parts <= 1) {
      if (parts == 1) fn(0);
      return;
    }
    auto job = std::make_shared<Job>();
    job->fn = std::move(fn);
    job->parts = parts;
    post(job);
    job->help();
    job->wait_done();
  }

 private:
  std::shared_ptr<Job> take() {
    std::lock_guard<std::mutex> g(mu_);
    while (!jobs_.empty() && jobs_.front()->claimed()) jobs_.pop_front();
    return jobs_.empty() ? nullptr : jobs_.front();
  }
  void loop() {
    for (;;) {
      const uint64_t seen = posted_.load(std::memory_order_acquire);
      if (auto job = take()) {
        job->help();
        continue;
      }
      bool again = false;
      for (const int64_t t0 = now_us(); spin_us_ > 0 && now_us() - t0 < spin_us_;) {
        if (posted_.load(std::memory_order_acquire) != seen) {
          again = true;
          break;
        }
        cpu_relax();
      }
      if (again) continue;
      std::unique_lock<std::mutex> g(mu_);
      cv_.wait(g, [&] { return stop_ || posted_.load(std::memory_order_acquire) != seen; });
      if (stop_) return;
    }
  }
  std::vector<std::thread> workers_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::shared_ptr<Job>> jobs_;
  std::atomic<uint64_t> posted_{0};
  int spin_us_ = 150;
  bool stop_ = false;
};
Pool& pool() {
  // Workers per process: from the host's cores shared between the worker processes of this node (one per GPU under
  // torchrun / TF_CONFIG launches); TFDL_IO_THREADS overrides. Leaked on purpose: no
  // static-destruction order games at exit.
  static Pool* p = [] {
    int n = (int)std::thread::hardware_concurrency();
    if (const char* lw = getenv("LOCAL_WORLD_SIZE")) n /= std::max(1, atoi(lw));
    // half of this process's share: the workers poll between jobs, and the training thread, the heartbeat thread and
    // the driver's own threads must never have to wait for a core behind them
    n = std::max(1, std::min(7, n / 2));
    if (const char* e = getenv("TFDL_IO_THREADS")) n = std::max(1, std::min(32, atoi(e)));
    return new Pool(n);
  }();
  return *p;
}
}  // namespace

template <class F>
static void parallel_ranges(int64_t n, int parts, F fn) {
  if (parts <= 1 || n <= 1) {
    fn(0, n);
    return;
  }
  parts = (int)std::min<int64_t>(parts, n);
  const int64_t per = (n + parts - 1) / parts;
  pool().run(parts, [=](int t) {
    const int64_t lo = t * per, hi = std::min(n, lo + per);
    if (lo < hi) fn(lo, hi);
  });
}

template <class F>
static void parallel_rows(int64_t n, int threads, F fn) {
  if (threads <= 1 || n < 256) {
    fn(0, n);
    return;
  }
  parallel_ranges(n, threads, fn);
}

// dst[i, :] = src[idx[i], :]  (row_bytes each)
extern "C" int tfdl_rt_gather_rows(const void* src, int64_t num_rows, int64_t row_bytes, const int64_t* idx,
                                   int64_t n, void* dst, int threads) {
  for (int64_t i = 0; i < n; ++i)
    if (idx[i] < 0 || idx[i] >= num_rows) {
      set_error("gather_rows: index %lld out of range [0,%lld)", (long long)idx[i], (long long)num_rows);
      return -1;
    }
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  parallel_rows(n, threads, [=](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) memcpy(d + i * row_bytes, s + idx[i] * row_bytes, row_bytes);
  });
  return 0;
}

// dst[i, :] = float(src_u8[idx[i], :]) * scale   (the scale() map of the reference fused into batching)
extern "C" int tfdl_rt_gather_rows_u8_to_f32(const uint8_t* src, int64_t num_rows, int64_t row_elems,
                                             const int64_t* idx, int64_t n, float scale, float* dst, int threads) {
  for (int64_t i = 0; i < n; ++i)
    if (idx[i] < 0 || idx[i] >= num_rows) {
      set_error("gather_rows_u8_to_f32: index %lld out of range", (long long)idx[i]);
      return -1;
    }
  parallel_rows(n, threads, [=](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) {
      const uint8_t* s = src + idx[i] * row_elems;
      float* d = dst + i * row_elems;
      for (int64_t j = 0; j < row_elems; ++j) d[j] = (float)s[j] * scale;
    }
  });
  return 0;
}

// Packs k training batches into the executor's pinned record ring: record i = [x_i bytes | labels_i as int32].
// xs[i] / ys[i] point at contiguous host arrays (float32 images; int32 or int64 labels). One call per chunk of
// steps replaces 2k numpy copies on the Python side of fit() (the rebatch + staging half of SURVEY N6).
extern "C" int tfdl_rt_pack_records(const void* const* xs, const void* const* ys, int64_t k, int64_t x_bytes,
                                    int64_t n_labels, int label_bytes, void* ring, int64_t rec_bytes, int threads) {
  if (!xs || !ys || !ring || k < 0 || x_bytes < 0 || n_labels < 0 || (label_bytes != 4 && label_bytes != 8) ||
      rec_bytes < x_bytes + n_labels * 4) {
    set_error("pack_records: bad arguments (k=%lld x_bytes=%lld labels=%lld x %d rec=%lld)", (long long)k,
              (long long)x_bytes, (long long)n_labels, label_bytes, (long long)rec_bytes);
    return -1;
  }
  auto work = [=](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) {
      char* rec = static_cast<char*>(ring) + i * rec_bytes;
      std::memcpy(rec, xs[i], (size_t)x_bytes);
      int32_t* yl = reinterpret_cast<int32_t*>(rec + x_bytes);
      if (label_bytes == 4) {
        std::memcpy(yl, ys[i], (size_t)n_labels * 4);
      } else {
        const int64_t* y8 = static_cast<const int64_t*>(ys[i]);
        for (int64_t j = 0; j < n_labels; ++j) yl[j] = (int32_t)y8[j];
      }
    }
  };
  // records are large (hundreds of KB): split across threads as soon as there is ~1 MB to move
  const int64_t nt = std::min<int64_t>(std::min<int64_t>(threads, k), (k * x_bytes) >> 19);
  parallel_ranges(k, (int)nt, work);
  return 0;
}

// Builds k records straight from the (cached) dataset arrays: record r = [x[idx[r*B + b]] for b < B | labels as int32].
// This is `.shuffle(...).batch(...)` of tf_dist_example.py:31-33 done as ONE gather into the executor's pinned ring -
// no per-batch arrays, no second staging copy (shuffle / batch / rebatch kernels).
//
// The work is cut into parts of GATHER_ROWS rows of one record, claimed in record order, so the first records are
// complete first. The asynchronous form (begin / ready / wait / end) is tf.data's prefetch for this path: fit() starts
// the gather of a whole chunk, launches the first steps as soon as their records are in the ring and keeps launching
// while the pool fills the rest behind the running steps.
namespace {
constexpr int64_t GATHER_ROWS = 16;
struct GatherJob {
  const char* xs = nullptr;
  const void* ys = nullptr;
  int label_bytes = 4;
  int64_t k = 0, B = 0, x_row_bytes = 0, rec_bytes = 0, ppr = 1;  // ppr: parts per record
  char* ring = nullptr;
  std::vector<int64_t> idx;                       // owned copy: the caller's array may go away
  std::unique_ptr<std::atomic<int>[]> left;       // parts of record r still to be copied
  std::atomic<int64_t> prefix{0};                 // records [0, prefix) are known to be complete
  std::shared_ptr<Job> job;
  void copy_part(int part) {
    const int64_t r = part / ppr, b0 = (part % ppr) * GATHER_ROWS, b1 = std::min(B, b0 + GATHER_ROWS);
    char* rec = ring + r * rec_bytes;
    int32_t* yl = reinterpret_cast<int32_t*>(rec + B * x_row_bytes);
    const int64_t* id = idx.data() + r * B;
    for (int64_t b = b0; b < b1; ++b) {
      const int64_t j = id[b];
      // rows are scattered over a dataset far larger than the caches: start the DRAM fetch of a row a few rows ahead
      // (the hardware prefetcher continues it) so the copy below is not one latency stall per row
      if (b + 4 < B) {  // also across the end of this part: whoever copies the next part finds its first rows on the way
        const char* nxt = xs + id[b + 4] * x_row_bytes;
        __builtin_prefetch(nxt);
        __builtin_prefetch(nxt + 64);
        __builtin_prefetch(nxt + 128);
      }
      std::memcpy(rec + b * x_row_bytes, xs + j * x_row_bytes, (size_t)x_row_bytes);
      yl[b] = label_bytes == 8 ? (int32_t) static_cast<const int64_t*>(ys)[j]
              : label_bytes == 4 ? static_cast<const int32_t*>(ys)[j]
                                 : (int32_t) static_cast<const uint8_t*>(ys)[j];
    }
    left[r].fetch_sub(1, std::memory_order_release);
  }
  int64_t ready() {
    int64_t p = prefix.load(std::memory_order_relaxed);
    while (p < k && left[p].load(std::memory_order_acquire) == 0) ++p;
    prefix.store(p, std::memory_order_relaxed);
    return p;
  }
};

GatherJob* make_gather(const void* x_src, int64_t num_rows, int64_t x_row_bytes, const void* y_src, int label_bytes,
                       const int64_t* idx, int64_t k, int64_t B, void* ring, int64_t rec_bytes) {
  if (!x_src || !y_src || !idx || !ring || k < 0 || B < 1 || (label_bytes != 1 && label_bytes != 4 && label_bytes != 8) ||
      rec_bytes < B * x_row_bytes + B * 4 || k * ((B + GATHER_ROWS - 1) / GATHER_ROWS) > (int64_t)1 << 30) {
    set_error("gather_records: bad arguments (k=%lld B=%lld row=%lld label_bytes=%d rec=%lld)", (long long)k, (long long)B,
              (long long)x_row_bytes, label_bytes, (long long)rec_bytes);
    return nullptr;
  }
  for (int64_t i = 0; i < k * B; ++i)
    if (idx[i] < 0 || idx[i] >= num_rows) {
      set_error("gather_records: index %lld out of range [0,%lld)", (long long)idx[i], (long long)num_rows);
      return nullptr;
    }
  auto* g = new GatherJob;
  g->xs = static_cast<const char*>(x_src);
  g->ys = y_src;
  g->label_bytes = label_bytes;
  g->k = k;
  g->B = B;
  g->x_row_bytes = x_row_bytes;
  g->rec_bytes = rec_bytes;
  g->ring = static_cast<char*>(ring);
  g->ppr = (B + GATHER_ROWS - 1) / GATHER_ROWS;
  g->idx.assign(idx, idx + k * B);
  g->left.reset(new std::atomic<int>[std::max<int64_t>(k, 1)]);
  for (int64_t r = 0; r < k; ++r) g->left[r].store((int)g->ppr, std::memory_order_relaxed);
  g->job = std::make_shared<Job>();
  g->job->parts = (int)(k * g->ppr);
  g->job->fn = [g](int part) { g->copy_part(part); };
  return g;
}
}  // namespace

extern "C" int tfdl_rt_gather_records(const void* x_src, int64_t num_rows, int64_t x_row_bytes, const void* y_src,
                                      int label_bytes, const int64_t* idx, int64_t k, int64_t B, void* ring,
                                      int64_t rec_bytes, int threads) {
  GatherJob* g = make_gather(x_src, num_rows, x_row_bytes, y_src, label_bytes, idx, k, B, ring, rec_bytes);
  if (!g) return -1;
  if (threads > 1 && k * B * x_row_bytes >= (1 << 18)) pool().post(g->job);  // small jobs: not worth waking anyone
  else g->job->pending.store(g->job->parts, std::memory_order_relaxed);
  g->job->help();
  g->job->wait_done();
  delete g;
  return 0;
}

// Asynchronous form. begin() validates, hands the job to the pool and returns; the caller must keep x_src, y_src and
// ring alive until end(). Returns NULL on bad arguments (tfdl_rt_last_error()).
extern "C" void* tfdl_rt_gather_records_begin(const void* x_src, int64_t num_rows, int64_t x_row_bytes, const void* y_src,
                                              int label_bytes, const int64_t* idx, int64_t k, int64_t B, void* ring,
                                              int64_t rec_bytes) {
  GatherJob* g = make_gather(x_src, num_rows, x_row_bytes, y_src, label_bytes, idx, k, B, ring, rec_bytes);
  if (!g) return nullptr;
  pool().post(g->job);
  return g;
}
// How many leading records are complete right now (does not block).
extern "C" int64_t tfdl_rt_gather_records_ready(void* h) { return static_cast<GatherJob*>(h)->ready(); }
// Returns once records [0, n) are complete. The caller does not sleep: it copies parts itself while any are
// unclaimed, so the wait costs at most the single-threaded time of what is left.
extern "C" int tfdl_rt_gather_records_wait(void* h, int64_t n) {
  auto* g = static_cast<GatherJob*>(h);
  n = std::min(n, g->k);
  while (g->ready() < n) {
    const int part = g->job->next.fetch_add(1, std::memory_order_relaxed);
    if (part < g->job->parts) {
      g->job->fn(part);
      g->job->pending.fetch_sub(1, std::memory_order_acq_rel);
    } else {
      cpu_relax();  // everything is claimed: the last parts are in flight on the workers (a few microseconds each)
    }
  }
  return 0;
}
// Completes the job (helping like wait) and frees it.
extern "C" void tfdl_rt_gather_records_end(void* h) {
  auto* g = static_cast<GatherJob*>(h);
  if (!g) return;
  g->job->help();
  g->job->wait_done();
  delete g;
}

// Streaming form of the shuffle buffer: the same algorithm as tfdl_rt_shuffle_indices, producing its output on
// demand. A fit() of 20 steps draws 1,280 of MNIST's 60,000 elements; shuffling all of them up front cost more than
// the 20 steps themselves.
namespace {
struct Shuffler {
  std::vector<int64_t> src, buf;  // src empty = the identity stream 0..n-1 (a cached / in-memory dataset)
  int64_t n = 0, t = 0, drained = 0;  // next source element to feed the buffer; elements already drained at the end
  int64_t buffer = 0;
  bool tail_shuffled = false;
  Rng rng;
  Shuffler(const int64_t* s, int64_t n_, int64_t b, uint64_t seed) : n(n_), buffer(std::min(b, n_)), rng(seed) {
    if (s) src.assign(s, s + n);
    buf.resize(buffer);
    for (int64_t i = 0; i < buffer; ++i) buf[i] = s ? s[i] : i;
    t = buffer;
  }
  int64_t at(int64_t i) const { return src.empty() ? i : src[i]; }
  // multiply-shift range reduction: one 64x64->128 multiply instead of two 64-bit divisions per draw
  uint64_t below(uint64_t m) { return (uint64_t)(((unsigned __int128)rng.next() * m) >> 64); }
};
}  // namespace

extern "C" void* tfdl_rt_shuffler_create(const int64_t* src, int64_t n, int64_t buffer, uint64_t seed) {
  if (n < 0 || buffer < 1) {  // src == NULL: the source stream is 0..n-1
    set_error("shuffler_create: bad arguments n=%lld buffer=%lld", (long long)n, (long long)buffer);
    return nullptr;
  }
  return new Shuffler(src, n, buffer, seed);
}
// Writes up to `max` further indices to `out`; returns how many (0 = exhausted).
extern "C" int64_t tfdl_rt_shuffler_next(void* h, int64_t* out, int64_t max) {
  auto* s = static_cast<Shuffler*>(h);
  const int64_t n = s->n;
  int64_t o = 0;
  while (o < max && s->t < n) {
    const uint64_t j = s->below((uint64_t)s->buffer);
    out[o++] = s->buf[j];
    s->buf[j] = s->at(s->t++);
  }
  if (o < max && s->t >= n) {
    if (!s->tail_shuffled) {
      for (int64_t i = s->buffer - 1; i > 0; --i) std::swap(s->buf[i], s->buf[s->below((uint64_t)i + 1)]);
      s->tail_shuffled = true;
    }
    while (o < max && s->drained < s->buffer) out[o++] = s->buf[s->drained++];
  }
  return o;
}
extern "C" void tfdl_rt_shuffler_destroy(void* h) { delete static_cast<Shuffler*>(h); }
