// Race-detector driver for the host runtime: W ranks as W threads of ONE process, each with its own
// group (listener, ring / control / heartbeat sockets, heartbeat thread), hammering barrier / allgather /
// broadcast / ring all-reduce back to back, then blocking and background gathers on the shared worker pool. Built with -fsanitize=thread by tools/sanitize.sh (ThreadSanitizer
// sees every pair of threads because they share the process), and without it as a plain stress test.
//   usage: tsan_driver <world> <base_port> <iters>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "../../include/tfdl_runtime.h"

// The input pipeline's worker pool under the same detector: every rank thread posts blocking and background gathers to
// the ONE pool of the process (atomic part claiming, workers that spin before they sleep, posters that copy parts
// themselves while they wait), several jobs in flight at once.
static int run_gathers(int rank, int iters) {
  const int64_t rows = 2000, row_bytes = 512, B = 32, k = 9, rec = B * row_bytes + B * 4;
  std::vector<char> x((size_t)(rows * row_bytes));
  std::vector<int64_t> y((size_t)rows);
  for (int64_t r = 0; r < rows; ++r) {
    y[(size_t)r] = r % 10;
    for (int64_t j = 0; j < row_bytes; ++j) x[(size_t)(r * row_bytes + j)] = (char)((r * 31 + j + rank) & 0x7f);
  }
  std::vector<int64_t> idx((size_t)(k * B));
  std::vector<char> ring_a((size_t)(k * rec)), ring_b((size_t)(k * rec));
  for (int it = 0; it < iters; ++it) {
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int64_t)((i * 7919 + (size_t)it * 13 + (size_t)rank) % rows);
    void* job = tfdl_rt_gather_records_begin(x.data(), rows, row_bytes, y.data(), 8, idx.data(), k, B, ring_a.data(), rec);
    if (!job) return 20;
    if (tfdl_rt_gather_records(x.data(), rows, row_bytes, y.data(), 8, idx.data(), k, B, ring_b.data(), rec, 8)) return 21;
    if (tfdl_rt_gather_records_wait(job, 2)) return 22;
    if (tfdl_rt_gather_records_ready(job) < 2) return 23;
    for (int64_t b = 0; b < 2 * B; ++b)  // the records reported ready are complete
      if (ring_a[(size_t)((b / B) * rec + (b % B) * row_bytes + 5)] != x[(size_t)(idx[(size_t)b] * row_bytes + 5)]) return 24;
    tfdl_rt_gather_records_end(job);
    if (ring_a != ring_b) return 25;
  }
  return 0;
}

static int run_rank(int rank, int world, const std::vector<std::string>& addrs, int iters) {
  std::vector<const char*> c;
  for (auto& a : addrs) c.push_back(a.c_str());
  void* g = tfdl_rt_group_create(rank, world, c.data(), 20000, 50, 2000);
  if (!g) {
    std::fprintf(stderr, "rank %d: group_create failed: %s\n", rank, tfdl_rt_last_error());
    return 1;
  }
  int bad = 0;
  std::vector<float> buf(100003);
  std::vector<double> dbuf(1001);
  std::vector<int> gathered(world);
  for (int it = 0; it < iters && !bad; ++it) {
    for (size_t i = 0; i < buf.size(); ++i) buf[i] = (float)((rank + 1) * (int)(i % 7) + it);
    if (tfdl_rt_ring_allreduce_f32(g, buf.data(), (int64_t)buf.size())) bad = 2;
    const float tot = world * (world + 1) / 2.f;
    for (size_t i = 0; i < buf.size() && !bad; i += 997)
      if (std::fabs(buf[i] - (tot * (float)(i % 7) + (float)world * it)) > 1e-3f) bad = 3;
    for (size_t i = 0; i < dbuf.size(); ++i) dbuf[i] = rank + 0.5;
    if (!bad && tfdl_rt_ring_allreduce_f64(g, dbuf.data(), (int64_t)dbuf.size())) bad = 4;
    if (!bad && std::fabs(dbuf[7] - (world * (world - 1) / 2.0 + 0.5 * world)) > 1e-9) bad = 5;
    int mine = rank * 10 + it;
    if (!bad && tfdl_rt_allgather(g, &mine, sizeof(int), gathered.data())) bad = 6;
    for (int r = 0; r < world && !bad; ++r)
      if (gathered[r] != r * 10 + it) bad = 7;
    int word = rank == 0 ? 4242 + it : -1;
    if (!bad && tfdl_rt_broadcast0(g, &word, sizeof(int))) bad = 8;
    if (!bad && word != 4242 + it) bad = 9;
    if (!bad && tfdl_rt_barrier(g)) bad = 10;
  }
  if (!bad) bad = run_gathers(rank, iters * 4);
  if (bad) std::fprintf(stderr, "rank %d: check %d failed: %s\n", rank, bad, tfdl_rt_last_error());
  if (tfdl_rt_group_failed_rank(g) != -1) bad = bad ? bad : 11;
  tfdl_rt_group_destroy(g);
  return bad;
}

int main(int argc, char** argv) {
  const int world = argc > 1 ? std::atoi(argv[1]) : 4;
  const int base = argc > 2 ? std::atoi(argv[2]) : 24500;
  const int iters = argc > 3 ? std::atoi(argv[3]) : 20;
  std::vector<std::string> addrs;
  for (int r = 0; r < world; ++r) addrs.push_back("127.0.0.1:" + std::to_string(base + r));
  std::vector<int> rc(world, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < world; ++r) th.emplace_back([&, r] { rc[r] = run_rank(r, world, addrs, iters); });
  for (auto& t : th) t.join();
  int bad = 0;
  for (int r = 0; r < world; ++r) bad |= rc[r];
  std::printf("tsan_driver world=%d iters=%d -> %s\n", world, iters, bad ? "FAILED" : "ok");
  return bad ? 1 : 0;
}
