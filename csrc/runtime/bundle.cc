// TensorBundle checkpoint writer / reader and TensorBoard event writer.
//
// The reference only says that the chief saves checkpoints and writes TensorBoard
// (/root/reference/README.md:51); in TensorFlow both are native code (tensor_bundle.cc,
// events_writer.cc), so they are native here as well. Formats are re-implemented from their
// specifications, not from TF sources:
//   <prefix>.index                 LevelDB-style sorted string table (prefix-compressed blocks with
//                                  restart points, per-block masked crc32c, 48-byte footer):
//                                  key ""  -> BundleHeaderProto, key <name> -> BundleEntryProto
//   <prefix>.data-00000-of-00001   raw little-endian tensor bytes, back to back
//   events.out.tfevents.*          TFRecord framing: u64 len, masked crc(len), payload, masked crc(payload)
// Protos are encoded by hand (a dozen varint fields), so there is no protobuf dependency.
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <map>
#include <memory>

#include "rt_common.h"
#include "../../include/tfdl_runtime.h"

using namespace tfdl_rt;

namespace {

constexpr uint64_t kTableMagic = 0xdb4775248b80fb57ull;
constexpr int kRestartInterval = 16;
constexpr size_t kBlockSize = 4096;

// ------------------------------------------------------------------------------------------
// proto helpers
// ------------------------------------------------------------------------------------------
void pb_varint_field(std::string* s, int field, uint64_t v) {
  put_varint32(s, (field << 3) | 0);
  put_varint64(s, v);
}
void pb_fixed32_field(std::string* s, int field, uint32_t v) {
  put_varint32(s, (field << 3) | 5);
  put_fixed32(s, v);
}
void pb_bytes_field(std::string* s, int field, const std::string& v) {
  put_varint32(s, (field << 3) | 2);
  put_varint64(s, v.size());
  s->append(v);
}

struct Entry {
  int dtype = 0;
  std::vector<int64_t> dims;
  int64_t offset = 0, size = 0;
  uint32_t crc = 0;  // masked
};

std::string encode_entry(const Entry& e) {
  std::string s;
  if (e.dtype) pb_varint_field(&s, 1, e.dtype);
  std::string shape;
  for (int64_t d : e.dims) {
    std::string dim;
    pb_varint_field(&dim, 1, (uint64_t)d);
    pb_bytes_field(&shape, 2, dim);
  }
  pb_bytes_field(&s, 2, shape);
  if (e.offset) pb_varint_field(&s, 4, e.offset);
  if (e.size) pb_varint_field(&s, 5, e.size);
  pb_fixed32_field(&s, 6, e.crc);
  return s;
}

bool decode_entry(const std::string& v, Entry* e) {
  const uint8_t* p = (const uint8_t*)v.data();
  const uint8_t* end = p + v.size();
  while (p < end) {
    uint64_t tag;
    if (!get_varint64(p, end, &tag)) return false;
    int field = tag >> 3, wt = tag & 7;
    if (wt == 0) {
      uint64_t x;
      if (!get_varint64(p, end, &x)) return false;
      if (field == 1) e->dtype = (int)x;
      if (field == 4) e->offset = (int64_t)x;
      if (field == 5) e->size = (int64_t)x;
    } else if (wt == 5) {
      if (end - p < 4) return false;
      uint32_t x;
      memcpy(&x, p, 4);
      p += 4;
      if (field == 6) e->crc = x;
    } else if (wt == 2) {
      uint64_t len;
      if (!get_varint64(p, end, &len) || (uint64_t)(end - p) < len) return false;
      if (field == 2) {  // TensorShapeProto
        const uint8_t* q = p;
        const uint8_t* qe = p + len;
        while (q < qe) {
          uint64_t t2;
          if (!get_varint64(q, qe, &t2)) return false;
          if ((t2 & 7) == 2) {
            uint64_t l2;
            if (!get_varint64(q, qe, &l2)) return false;
            if ((t2 >> 3) == 2) {  // Dim
              const uint8_t* r = q;
              const uint8_t* re = q + l2;
              int64_t size = 0;
              while (r < re) {
                uint64_t t3;
                if (!get_varint64(r, re, &t3)) return false;
                if ((t3 & 7) == 0) {
                  uint64_t x;
                  if (!get_varint64(r, re, &x)) return false;
                  if ((t3 >> 3) == 1) size = (int64_t)x;
                } else if ((t3 & 7) == 2) {
                  uint64_t l3;
                  if (!get_varint64(r, re, &l3)) return false;
                  r += l3;
                } else {
                  return false;
                }
              }
              e->dims.push_back(size);
            }
            q += l2;
          } else if ((t2 & 7) == 0) {
            uint64_t x;
            if (!get_varint64(q, qe, &x)) return false;
          } else {
            return false;
          }
        }
      }
      p += len;
    } else if (wt == 1) {
      p += 8;
    } else {
      return false;
    }
  }
  return true;
}

// ------------------------------------------------------------------------------------------
// sorted string table
// ------------------------------------------------------------------------------------------
struct BlockBuilder {
  std::string buf, last_key;
  std::vector<uint32_t> restarts{0};
  int counter = 0, interval;
  explicit BlockBuilder(int iv) : interval(iv) {}
  bool empty() const { return buf.empty(); }
  size_t estimate() const { return buf.size() + restarts.size() * 4 + 4; }
  void add(const std::string& key, const std::string& value) {
    size_t shared = 0;
    if (counter < interval) {
      size_t m = std::min(last_key.size(), key.size());
      while (shared < m && last_key[shared] == key[shared]) ++shared;
    } else {
      restarts.push_back((uint32_t)buf.size());
      counter = 0;
    }
    put_varint32(&buf, (uint32_t)shared);
    put_varint32(&buf, (uint32_t)(key.size() - shared));
    put_varint32(&buf, (uint32_t)value.size());
    buf.append(key, shared, std::string::npos);
    buf.append(value);
    last_key = key;
    ++counter;
  }
  std::string finish() {
    std::string out = buf;
    for (uint32_t r : restarts) put_fixed32(&out, r);
    put_fixed32(&out, (uint32_t)restarts.size());
    return out;
  }
  void reset() {
    buf.clear();
    last_key.clear();
    restarts.assign(1, 0);
    counter = 0;
  }
};

struct Handle {
  uint64_t offset = 0, size = 0;
  std::string encode() const {
    std::string s;
    put_varint64(&s, offset);
    put_varint64(&s, size);
    return s;
  }
};

Handle write_block(std::string* file, const std::string& contents) {
  Handle h{file->size(), contents.size()};
  file->append(contents);
  char type = 0;  // no compression
  uint32_t crc = crc32c_extend(crc32c_value(contents.data(), contents.size()), &type, 1);
  file->push_back(type);
  put_fixed32(file, crc32c_mask(crc));
  return h;
}

std::string build_table(const std::map<std::string, std::string>& kv) {
  std::string file;
  BlockBuilder data(kRestartInterval), index(1);
  std::string last;
  auto flush = [&] {
    if (data.empty()) return;
    Handle h = write_block(&file, data.finish());
    index.add(last, h.encode());
    data.reset();
  };
  for (auto& it : kv) {
    data.add(it.first, it.second);
    last = it.first;
    if (data.estimate() >= kBlockSize) flush();
  }
  flush();
  BlockBuilder meta(kRestartInterval);
  Handle mh = write_block(&file, meta.finish());
  Handle ih = write_block(&file, index.finish());
  std::string footer = mh.encode() + ih.encode();
  footer.resize(40, '\0');
  put_fixed64(&footer, kTableMagic);
  file.append(footer);
  return file;
}

bool parse_block(const std::string& file, Handle h, std::vector<std::pair<std::string, std::string>>* out) {
  if (h.offset + h.size + 5 > file.size()) return false;
  const char* base = file.data() + h.offset;
  uint32_t stored;
  memcpy(&stored, base + h.size + 1, 4);
  uint32_t crc = crc32c_value(base, h.size + 1);
  if (crc32c_unmask(stored) != crc) {
    set_error("index block checksum mismatch at offset %llu", (unsigned long long)h.offset);
    return false;
  }
  if (base[h.size] != 0) {
    set_error("compressed index blocks are not supported (type %d)", (int)base[h.size]);
    return false;
  }
  if (h.size < 4) return false;
  uint32_t nrestart;
  memcpy(&nrestart, base + h.size - 4, 4);
  size_t data_end = h.size - 4 - 4 * (size_t)nrestart;
  const uint8_t* p = (const uint8_t*)base;
  const uint8_t* end = p + data_end;
  std::string key;
  while (p < end) {
    uint64_t shared, non_shared, vlen;
    if (!get_varint64(p, end, &shared) || !get_varint64(p, end, &non_shared) || !get_varint64(p, end, &vlen))
      return false;
    if (shared > key.size() || (uint64_t)(end - p) < non_shared + vlen) return false;
    key.resize(shared);
    key.append((const char*)p, non_shared);
    p += non_shared;
    out->emplace_back(key, std::string((const char*)p, vlen));
    p += vlen;
  }
  return true;
}

bool read_file(const std::string& path, std::string* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out->resize(n);
  bool ok = n == 0 || fread(&(*out)[0], 1, n, f) == (size_t)n;
  fclose(f);
  return ok;
}

bool write_file_atomic(const std::string& path, const std::string& data) {
  std::string tmp = path + ".tmp" + std::to_string((long)getpid());
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return false;
  bool ok = data.empty() || fwrite(data.data(), 1, data.size(), f) == data.size();
  ok = (fflush(f) == 0) && ok;
  fsync(fileno(f));
  fclose(f);
  if (!ok || rename(tmp.c_str(), path.c_str()) != 0) {
    unlink(tmp.c_str());
    return false;
  }
  return true;
}

struct Writer {
  std::string prefix, data;
  std::map<std::string, std::string> index;  // key -> encoded BundleEntryProto
};

struct Reader {
  std::string prefix, data;
  std::vector<std::pair<std::string, Entry>> entries;
  bool data_loaded = false;
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// writer
// ---------------------------------------------------------------------------------------------
extern "C" void* tfdl_rt_bundle_writer_create(const char* prefix) {
  auto* w = new Writer();
  w->prefix = prefix;
  return w;
}

extern "C" int tfdl_rt_bundle_writer_add(void* h, const char* key, int dtype, int ndim, const int64_t* dims,
                                         const void* data, uint64_t nbytes) {
  auto* w = static_cast<Writer*>(h);
  if (w->index.count(key)) {
    set_error("bundle: duplicate key %s", key);
    return -1;
  }
  Entry e;
  e.dtype = dtype;
  e.dims.assign(dims, dims + ndim);
  e.offset = (int64_t)w->data.size();
  e.size = (int64_t)nbytes;
  e.crc = crc32c_mask(crc32c_value(data, nbytes));
  w->data.append(static_cast<const char*>(data), nbytes);
  w->index[key] = encode_entry(e);
  return 0;
}

// Scalar DT_STRING tensor (used for the _CHECKPOINTABLE_OBJECT_GRAPH entry): varint length,
// masked crc of the fixed-width length, then the bytes; the entry crc covers all three.
extern "C" int tfdl_rt_bundle_writer_add_string(void* h, const char* key, const void* data, uint64_t nbytes) {
  auto* w = static_cast<Writer*>(h);
  if (w->index.count(key)) {
    set_error("bundle: duplicate key %s", key);
    return -1;
  }
  std::string blob;
  put_varint64(&blob, nbytes);
  uint32_t len32 = (uint32_t)nbytes;
  uint32_t crc = crc32c_value(&len32, 4);
  uint32_t len_ck = crc32c_mask(crc);
  put_fixed32(&blob, len_ck);
  crc = crc32c_extend(crc, &len_ck, 4);
  blob.append(static_cast<const char*>(data), nbytes);
  crc = crc32c_extend(crc, data, nbytes);
  Entry e;
  e.dtype = 7;  // DT_STRING
  e.offset = (int64_t)w->data.size();
  e.size = (int64_t)blob.size();
  e.crc = crc32c_mask(crc);
  w->data.append(blob);
  w->index[key] = encode_entry(e);
  return 0;
}

extern "C" int tfdl_rt_bundle_writer_finish(void* h) {
  std::unique_ptr<Writer> w(static_cast<Writer*>(h));
  std::string header;  // BundleHeaderProto{num_shards=1, version{producer=1}}
  pb_varint_field(&header, 1, 1);
  std::string ver;
  pb_varint_field(&ver, 1, 1);
  pb_bytes_field(&header, 3, ver);
  w->index[""] = header;
  if (!write_file_atomic(w->prefix + ".data-00000-of-00001", w->data)) {
    set_error("bundle: cannot write %s.data-00000-of-00001", w->prefix.c_str());
    return -1;
  }
  if (!write_file_atomic(w->prefix + ".index", build_table(w->index))) {
    set_error("bundle: cannot write %s.index", w->prefix.c_str());
    return -1;
  }
  return 0;
}

extern "C" void tfdl_rt_bundle_writer_abort(void* h) { delete static_cast<Writer*>(h); }

// ---------------------------------------------------------------------------------------------
// reader
// ---------------------------------------------------------------------------------------------
extern "C" void* tfdl_rt_bundle_reader_open(const char* prefix) {
  std::string file;
  if (!read_file(std::string(prefix) + ".index", &file) || file.size() < 48) {
    set_error("bundle: cannot read %s.index", prefix);
    return nullptr;
  }
  uint64_t magic;
  memcpy(&magic, file.data() + file.size() - 8, 8);
  if (magic != kTableMagic) {
    set_error("bundle: %s.index is not a table file (bad magic)", prefix);
    return nullptr;
  }
  const uint8_t* p = (const uint8_t*)file.data() + file.size() - 48;
  const uint8_t* end = p + 40;
  Handle mh, ih;
  if (!get_varint64(p, end, &mh.offset) || !get_varint64(p, end, &mh.size) || !get_varint64(p, end, &ih.offset) ||
      !get_varint64(p, end, &ih.size)) {
    set_error("bundle: bad footer in %s.index", prefix);
    return nullptr;
  }
  std::vector<std::pair<std::string, std::string>> idx, kv;
  if (!parse_block(file, ih, &idx)) return nullptr;
  for (auto& e : idx) {
    Handle bh;
    const uint8_t* q = (const uint8_t*)e.second.data();
    const uint8_t* qe = q + e.second.size();
    if (!get_varint64(q, qe, &bh.offset) || !get_varint64(q, qe, &bh.size) || !parse_block(file, bh, &kv)) {
      if (!*last_error()) set_error("bundle: corrupt data block in %s.index", prefix);
      return nullptr;
    }
  }
  auto* r = new Reader();
  r->prefix = prefix;
  for (auto& e : kv) {
    if (e.first.empty()) continue;  // header
    Entry en;
    if (!decode_entry(e.second, &en)) {
      set_error("bundle: cannot decode entry %s", e.first.c_str());
      delete r;
      return nullptr;
    }
    r->entries.emplace_back(e.first, en);
  }
  return r;
}

extern "C" void tfdl_rt_bundle_reader_close(void* h) { delete static_cast<Reader*>(h); }
extern "C" int tfdl_rt_bundle_reader_num_entries(void* h) { return (int)static_cast<Reader*>(h)->entries.size(); }

extern "C" int tfdl_rt_bundle_reader_entry(void* h, int i, char* key, int key_cap, int* dtype, int* ndim,
                                           int64_t* dims8, uint64_t* nbytes) {
  auto* r = static_cast<Reader*>(h);
  if (i < 0 || i >= (int)r->entries.size()) {
    set_error("bundle: entry %d out of range", i);
    return -1;
  }
  auto& e = r->entries[i];
  snprintf(key, key_cap, "%s", e.first.c_str());
  *dtype = e.second.dtype;
  *ndim = (int)e.second.dims.size();
  for (int d = 0; d < *ndim && d < 8; ++d) dims8[d] = e.second.dims[d];
  *nbytes = (uint64_t)e.second.size;
  return 0;
}

// Copies the raw bytes of entry i into dst (for DT_STRING scalars: the string payload only) and
// verifies the stored crc32c.
extern "C" int tfdl_rt_bundle_reader_read(void* h, int i, void* dst, uint64_t cap, uint64_t* written) {
  auto* r = static_cast<Reader*>(h);
  if (i < 0 || i >= (int)r->entries.size()) {
    set_error("bundle: entry %d out of range", i);
    return -1;
  }
  if (!r->data_loaded) {
    if (!read_file(r->prefix + ".data-00000-of-00001", &r->data)) {
      set_error("bundle: cannot read %s.data-00000-of-00001", r->prefix.c_str());
      return -1;
    }
    r->data_loaded = true;
  }
  auto& e = r->entries[i].second;
  if ((uint64_t)(e.offset + e.size) > r->data.size()) {
    set_error("bundle: entry %s points past the end of the data file", r->entries[i].first.c_str());
    return -1;
  }
  const char* src = r->data.data() + e.offset;
  if (e.dtype == 7) {
    const uint8_t* p = (const uint8_t*)src;
    const uint8_t* end = p + e.size;
    uint64_t len;
    if (!get_varint64(p, end, &len) || (uint64_t)(end - p) < 4 + len) {
      set_error("bundle: malformed string tensor %s", r->entries[i].first.c_str());
      return -1;
    }
    uint32_t len32 = (uint32_t)len, ck;
    memcpy(&ck, p, 4);
    uint32_t crc = crc32c_value(&len32, 4);
    if (crc32c_mask(crc) != ck) {
      set_error("bundle: string length checksum mismatch in %s", r->entries[i].first.c_str());
      return -1;
    }
    crc = crc32c_extend(crc, &ck, 4);
    crc = crc32c_extend(crc, p + 4, len);
    if (crc32c_mask(crc) != e.crc) {
      set_error("bundle: checksum mismatch in %s", r->entries[i].first.c_str());
      return -1;
    }
    if (len > cap) {
      set_error("bundle: buffer too small for %s", r->entries[i].first.c_str());
      return -1;
    }
    memcpy(dst, p + 4, len);
    *written = len;
    return 0;
  }
  if (crc32c_mask(crc32c_value(src, e.size)) != e.crc) {
    set_error("bundle: checksum mismatch in %s (file is corrupt)", r->entries[i].first.c_str());
    return -1;
  }
  if ((uint64_t)e.size > cap) {
    set_error("bundle: buffer too small for %s", r->entries[i].first.c_str());
    return -1;
  }
  memcpy(dst, src, e.size);
  *written = (uint64_t)e.size;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// TensorBoard event files
// ---------------------------------------------------------------------------------------------
namespace {
struct EventWriter {
  FILE* f = nullptr;
  std::string path;
};
void write_record(FILE* f, const std::string& payload) {
  uint64_t len = payload.size();
  uint32_t c1 = crc32c_mask(crc32c_value(&len, 8));
  uint32_t c2 = crc32c_mask(crc32c_value(payload.data(), payload.size()));
  fwrite(&len, 8, 1, f);
  fwrite(&c1, 4, 1, f);
  fwrite(payload.data(), 1, payload.size(), f);
  fwrite(&c2, 4, 1, f);
}
double wall_now() {
  return std::chrono::duration<double>(std::chrono::system_clock::now().time_since_epoch()).count();
}
void pb_double_field(std::string* s, int field, double v) {
  put_varint32(s, (field << 3) | 1);
  uint64_t u;
  memcpy(&u, &v, 8);
  put_fixed64(s, u);
}
}  // namespace

extern "C" void* tfdl_rt_events_open(const char* logdir, const char* suffix) {
  mkdir(logdir, 0777);
  char host[256] = "localhost";
  gethostname(host, sizeof(host) - 1);
  char name[1024];
  snprintf(name, sizeof(name), "%s/events.out.tfevents.%010lld.%s.%d%s", logdir, (long long)wall_now(), host,
           (int)getpid(), suffix ? suffix : "");
  FILE* f = fopen(name, "wb");
  if (!f) {
    set_error("events: cannot create %s", name);
    return nullptr;
  }
  auto* w = new EventWriter{f, name};
  std::string ev;
  pb_double_field(&ev, 1, wall_now());
  pb_bytes_field(&ev, 3, "brain.Event:2");
  write_record(f, ev);
  fflush(f);
  return w;
}

extern "C" int tfdl_rt_events_scalar(void* h, const char* tag, double value, int64_t step) {
  auto* w = static_cast<EventWriter*>(h);
  std::string val;
  pb_bytes_field(&val, 1, tag);
  put_varint32(&val, (2 << 3) | 5);
  float fv = (float)value;
  uint32_t u;
  memcpy(&u, &fv, 4);
  put_fixed32(&val, u);
  std::string summary;
  pb_bytes_field(&summary, 1, val);
  std::string ev;
  pb_double_field(&ev, 1, wall_now());
  pb_varint_field(&ev, 2, (uint64_t)step);
  pb_bytes_field(&ev, 5, summary);
  write_record(w->f, ev);
  return 0;
}

extern "C" int tfdl_rt_events_flush(void* h) { return fflush(static_cast<EventWriter*>(h)->f); }
extern "C" const char* tfdl_rt_events_path(void* h) { return static_cast<EventWriter*>(h)->path.c_str(); }
extern "C" void tfdl_rt_events_close(void* h) {
  auto* w = static_cast<EventWriter*>(h);
  if (w->f) fclose(w->f);
  delete w;
}
