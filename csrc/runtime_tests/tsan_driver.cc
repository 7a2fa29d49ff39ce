// Race-detector driver for the host runtime: W ranks as W threads of ONE process, each with its own
// group (listener, ring / control / heartbeat sockets, heartbeat thread), hammering barrier / allgather /
// broadcast / ring all-reduce back to back, then blocking and background gathers on the shared worker pool. Built with -fsanitize=thread by tools/sanitize.sh (ThreadSanitizer
// sees every pair of threads because they share the process), and without it as a plain stress test.
//   usage: tsan_driver <world> <base_port> <iters> [scratch_dir]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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

// Checkpoint bundle + event file round trip (memory-error coverage under ASan / UBSan; one call, rank-private paths).
static int run_files(int rank, const std::string& dir) {
  const std::string prefix = dir + "/ck-rank" + std::to_string(rank);
  std::vector<float> a(3 * 5 * 7), b(1);
  for (size_t i = 0; i < a.size(); ++i) a[i] = (float)i * 0.25f - 3.f;
  b[0] = 42.f;
  std::vector<int64_t> it(1, 1234567);
  void* w = tfdl_rt_bundle_writer_create(prefix.c_str());
  if (!w) return 30;
  const int64_t da[3] = {3, 5, 7}, db[1] = {1};
  if (tfdl_rt_bundle_writer_add(w, "layer/kernel", 1 /* DT_FLOAT */, 3, da, a.data(), a.size() * 4)) return 31;
  if (tfdl_rt_bundle_writer_add(w, "layer/bias", 1, 1, db, b.data(), 4)) return 32;
  if (tfdl_rt_bundle_writer_add(w, "optimizer/iter", 9 /* DT_INT64 */, 0, nullptr, it.data(), 8)) return 33;
  if (tfdl_rt_bundle_writer_add_string(w, "_GRAPH", "abc\0def", 7)) return 34;
  if (tfdl_rt_bundle_writer_finish(w)) return 35;
  void* r = tfdl_rt_bundle_reader_open(prefix.c_str());
  if (!r) return 36;
  const int n = tfdl_rt_bundle_reader_num_entries(r);
  int seen = 0;
  for (int i = 0; i < n; ++i) {
    char key[256];
    int dt = 0, nd = 0;
    int64_t dims[8];
    uint64_t nbytes = 0, got = 0;
    if (tfdl_rt_bundle_reader_entry(r, i, key, sizeof key, &dt, &nd, dims, &nbytes)) return 37;
    std::vector<char> buf(nbytes + 1);
    if (tfdl_rt_bundle_reader_read(r, i, buf.data(), nbytes, &got)) return 38;
    if (dt == 7 /* DT_STRING: the payload without its length prefix and checksum */ ? got != 7 : got != nbytes) return 44;
    if (std::string(key) == "layer/kernel") {
      if (nd != 3 || dims[2] != 7 || nbytes != a.size() * 4 || std::memcmp(buf.data(), a.data(), nbytes)) return 39;
      ++seen;
    }
  }
  tfdl_rt_bundle_reader_close(r);
  if (seen != 1 || n < 4) return 40;
  void* ev = tfdl_rt_events_open((dir + "/events-rank" + std::to_string(rank)).c_str(), ".v2");
  if (!ev) return 41;
  for (int s = 0; s < 5; ++s)
    if (tfdl_rt_events_scalar(ev, "epoch_loss", 2.3 - 0.1 * s, s)) return 42;
  if (tfdl_rt_events_flush(ev)) return 43;
  tfdl_rt_events_close(ev);
  return 0;
}

static int run_rank(int rank, int world, const std::vector<std::string>& addrs, int iters, const std::string& dir) {
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
  if (!bad && !dir.empty()) bad = run_files(rank, dir);
  if (bad) std::fprintf(stderr, "rank %d: check %d failed: %s\n", rank, bad, tfdl_rt_last_error());
  if (tfdl_rt_group_failed_rank(g) != -1) bad = bad ? bad : 11;
  tfdl_rt_group_destroy(g);
  return bad;
}

int main(int argc, char** argv) {
  const int world = argc > 1 ? std::atoi(argv[1]) : 4;
  const int base = argc > 2 ? std::atoi(argv[2]) : 24500;
  const int iters = argc > 3 ? std::atoi(argv[3]) : 20;
  const std::string dir = argc > 4 ? argv[4] : "";  // where the checkpoint / event-file round trip may write (optional)
  std::vector<std::string> addrs;
  for (int r = 0; r < world; ++r) addrs.push_back("127.0.0.1:" + std::to_string(base + r));
  std::vector<int> rc(world, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < world; ++r) th.emplace_back([&, r] { rc[r] = run_rank(r, world, addrs, iters, dir); });
  for (auto& t : th) t.join();
  int bad = 0;
  for (int r = 0; r < world; ++r) bad |= rc[r];
  std::printf("tsan_driver world=%d iters=%d -> %s\n", world, iters, bad ? "FAILED" : "ok");
  return bad ? 1 : 0;
}
