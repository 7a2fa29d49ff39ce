// Mutation fuzzer for the TensorBundle reader (csrc/runtime/bundle.cc): a checkpoint is input from outside - TensorFlow
// writes the same format, and files get truncated or damaged - so a damaged `.index` or data file must end in an error
// return, never in a crash or an out-of-bounds read. Built with -fsanitize=address,undefined by tests/test_sanitizers.py.
//   usage: bundle_fuzz <scratch_dir> <iterations> <seed>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iterator>
#include <string>
#include <vector>

#include "../../include/tfdl_runtime.h"

static std::vector<char> slurp(const std::string& p) {
  std::ifstream f(p, std::ios::binary);
  return std::vector<char>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}
static void spit(const std::string& p, const std::vector<char>& b) {
  std::ofstream f(p, std::ios::binary | std::ios::trunc);
  f.write(b.data(), (std::streamsize)b.size());
}
static uint64_t rng_state = 1;
static uint64_t rnd() {
  rng_state ^= rng_state << 13;
  rng_state ^= rng_state >> 7;
  rng_state ^= rng_state << 17;
  return rng_state;
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  const std::string dir = argv[1];
  const int iters = std::atoi(argv[2]);
  rng_state = std::strtoull(argv[3], nullptr, 10) | 1;
  const std::string good = dir + "/good";
  {
    std::vector<float> a(64 * 9), b(64);
    for (size_t i = 0; i < a.size(); ++i) a[i] = (float)i;
    void* w = tfdl_rt_bundle_writer_create(good.c_str());
    if (!w) return 3;
    const int64_t da[4] = {3, 3, 1, 64}, db[1] = {64};
    int64_t step = 77;
    for (int l = 0; l < 6; ++l) {  // enough keys for several index blocks' worth of prefix compression
      const std::string k = "layer_with_weights-" + std::to_string(l);
      if (tfdl_rt_bundle_writer_add(w, (k + "/kernel/.ATTRIBUTES/VARIABLE_VALUE").c_str(), 1, 4, da, a.data(), a.size() * 4)) return 4;
      if (tfdl_rt_bundle_writer_add(w, (k + "/bias/.ATTRIBUTES/VARIABLE_VALUE").c_str(), 1, 1, db, b.data(), b.size() * 4)) return 4;
    }
    if (tfdl_rt_bundle_writer_add(w, "optimizer/iter/.ATTRIBUTES/VARIABLE_VALUE", 9, 0, nullptr, &step, 8)) return 4;
    if (tfdl_rt_bundle_writer_add_string(w, "_CHECKPOINTABLE_OBJECT_GRAPH", "\x0a\x03\x0a\x01x", 5)) return 4;
    if (tfdl_rt_bundle_writer_finish(w)) return 5;
  }
  const std::vector<char> index = slurp(good + ".index"), data = slurp(good + ".data-00000-of-00001");
  if (index.empty() || data.empty()) return 6;
  const std::string bad = dir + "/bad";
  int opened = 0, read_ok = 0;
  for (int it = 0; it < iters; ++it) {
    std::vector<char> idx = index, dat = data;
    std::vector<char>& victim = (rnd() % 4 == 0) ? dat : idx;
    switch (rnd() % 4) {
      case 0:  // flip a few bytes
        for (int k = 0, n = 1 + (int)(rnd() % 6); k < n; ++k) victim[rnd() % victim.size()] ^= (char)(1 + rnd() % 255);
        break;
      case 1:  // truncate
        victim.resize(rnd() % victim.size());
        break;
      case 2:  // overwrite a run with one byte (0x00 / 0xff / 0x80: varint edge cases)
      {
        const size_t at = rnd() % victim.size(), len = 1 + rnd() % 16;
        const char v = (char)(rnd() % 3 == 0 ? 0x00 : rnd() % 2 ? 0xff : 0x80);
        for (size_t k = at; k < victim.size() && k < at + len; ++k) victim[k] = v;
        break;
      }
      default:  // duplicate a slice into another place
      {
        const size_t from = rnd() % victim.size(), to = rnd() % victim.size(), len = 1 + rnd() % 32;
        for (size_t k = 0; k < len && from + k < victim.size() && to + k < victim.size(); ++k) victim[to + k] = victim[from + k];
      }
    }
    spit(bad + ".index", idx);
    spit(bad + ".data-00000-of-00001", dat);
    void* r = tfdl_rt_bundle_reader_open(bad.c_str());
    if (!r) continue;
    ++opened;
    const int n = tfdl_rt_bundle_reader_num_entries(r);
    for (int i = 0; i < n; ++i) {
      char key[512];
      int dt = 0, nd = 0;
      int64_t dims[8] = {0};
      uint64_t nbytes = 0, got = 0;
      if (tfdl_rt_bundle_reader_entry(r, i, key, sizeof key, &dt, &nd, dims, &nbytes)) continue;
      if (nbytes > (64u << 20)) continue;  // a damaged size field: the caller would not allocate that either
      std::vector<char> buf(nbytes + 1);
      if (tfdl_rt_bundle_reader_read(r, i, buf.data(), nbytes, &got) == 0) ++read_ok;
    }
    tfdl_rt_bundle_reader_close(r);
  }
  std::printf("bundle_fuzz iterations=%d opened=%d entries_read=%d -> ok\n", iters, opened, read_ok);
  return 0;
}
