// libsvm -> TFRecord converter, native (reference: tools/libsvm_to_tfrecord.py, TOOL:5-37, which walks the file line by
// line in Python and builds one tf.train.Example per line through the protobuf API; 7.5 K lines/s for the Python
// restatement in utils/libsvm.py, i.e. 100 minutes for Criteo's 45 M lines).
//
// Same output, byte for byte, as utils/libsvm.py + utils/tfrecord.py::make_ctr_example: features emitted in sorted key
// order (ids, label, values), packed repeated encodings, float32 values rounded from the parsed double, rows padded with
// (pad_id, 0.0) or truncated to field_size when one is given, blank lines skipped.
//
// The input is memory-mapped and cut into blocks at line boundaries; inside a block every thread converts its own run of
// lines into a private buffer of framed records (u64 length | masked crc | payload | masked crc), and the buffers are
// appended to the output in order, so the record order equals the line order.
#include <fcntl.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <charconv>
#include <string>
#include <thread>
#include <vector>

extern "C" uint32_t dfmio_masked_crc32c(const uint8_t* p, size_t n);

namespace {

thread_local char g_cerr[512] = "";

void put_varint(std::string& out, uint64_t v) {
  while (v >= 0x80) {
    out.push_back((char)(v | 0x80));
    v >>= 7;
  }
  out.push_back((char)v);
}
void put_ld(std::string& out, int field, const std::string& payload) {
  put_varint(out, ((uint64_t)field << 3) | 2);
  put_varint(out, payload.size());
  out += payload;
}

// one map entry of Features.feature: key (field 1) + Feature (field 2) holding float_list (2) or int64_list (3)
void put_feature(std::string& entries, const char* key, int list_field, const std::string& packed, std::string& t1,
                 std::string& t2) {
  t1.clear();
  put_ld(t1, 1, packed);            // FloatList / Int64List { value = 1, packed }
  t2.clear();
  put_ld(t2, list_field, t1);       // Feature { float_list = 2 | int64_list = 3 }
  t1.clear();
  put_ld(t1, 1, std::string(key));  // map entry: key
  put_ld(t1, 2, t2);                // map entry: value
  put_ld(entries, 1, t1);           // Features.feature (repeated map entry)
}

bool parse_double(const char*& p, const char* e, double& v) {
  if (p < e && *p == '+') ++p;
  auto r = std::from_chars(p, e, v);
  if (r.ec != std::errc()) return false;
  p = r.ptr;
  return true;
}

struct Scratch {
  std::vector<int64_t> ids;
  std::vector<float> vals;
  std::string packed, entries, example, t1, t2;
};

void emit_record(double label, Scratch& s, std::string& out);

// Converts one line [p, e) (no newline) and appends the framed record to `out`. Returns false on a malformed line.
bool convert_line(const char* p, const char* e, int field_size, int64_t pad_id, Scratch& s, std::string& out) {
  while (e > p && (e[-1] == '\r' || e[-1] == '\n')) --e;
  double label;
  if (!parse_double(p, e, label) || (p < e && *p != ' ')) return false;
  s.ids.clear();
  s.vals.clear();
  while (p < e) {
    while (p < e && *p == ' ') ++p;  // single-space separated; empty tokens are skipped like the Python split does
    if (p >= e) break;
    int64_t id;
    auto r = std::from_chars(p, e, id);
    if (r.ec != std::errc() || r.ptr >= e || *r.ptr != ':') return false;
    p = r.ptr + 1;
    double v;
    if (!parse_double(p, e, v) || (p < e && *p != ' ')) return false;
    s.ids.push_back(id);
    s.vals.push_back((float)v);
  }
  if (field_size > 0) {
    s.ids.resize((size_t)field_size, pad_id);  // truncates or pads
    s.vals.resize((size_t)field_size, 0.0f);
  }
  emit_record(label, s, out);
  return true;
}

// Appends the framed record of Example{ids = s.ids, label, values = s.vals} to `out`.
void emit_record(double label, Scratch& s, std::string& out) {
  s.entries.clear();
  s.packed.clear();
  for (int64_t id : s.ids) put_varint(s.packed, (uint64_t)id);
  put_feature(s.entries, "ids", 3, s.packed, s.t1, s.t2);
  const float lab = (float)label;
  s.packed.assign((const char*)&lab, 4);
  put_feature(s.entries, "label", 2, s.packed, s.t1, s.t2);
  s.packed.assign((const char*)s.vals.data(), s.vals.size() * 4);
  put_feature(s.entries, "values", 2, s.packed, s.t1, s.t2);
  s.example.clear();
  put_ld(s.example, 1, s.entries);  // Example { features = 1 }
  const uint64_t len = s.example.size();
  const uint32_t c1 = dfmio_masked_crc32c((const uint8_t*)&len, 8);
  const uint32_t c2 = dfmio_masked_crc32c((const uint8_t*)s.example.data(), s.example.size());
  out.append((const char*)&len, 8);
  out.append((const char*)&c1, 4);
  out += s.example;
  out.append((const char*)&c2, 4);
}

bool blank(const char* p, const char* e) {
  for (; p < e; ++p)
    if (*p != ' ' && *p != '\t' && *p != '\r' && *p != '\n' && *p != '\f' && *p != '\v') return false;
  return true;
}

}  // namespace

extern "C" {

const char* dfmio_libsvm_last_error() { return g_cerr; }

// Dense arrays -> framed TFRecord bytes of n Examples (labels [n] f32, ids [n, F] i64, vals [n, F] f32), for writers that
// already hold columns (synthetic data, parquet / numpy sources). Returns the number of bytes the n records take; they are
// written to `out` only if they fit in `cap` (call with out = NULL to size the buffer).
unsigned long long dfmio_encode_ctr_batch(const float* labels, const int64_t* ids, const float* vals, long long n, int F,
                                          uint8_t* out, unsigned long long cap) {
  if (!labels || !ids || !vals || n < 0 || F <= 0) return 0;
  Scratch s;
  std::string buf;
  buf.reserve((size_t)n * (size_t)(64 + 14 * F));
  for (long long r = 0; r < n; ++r) {
    s.ids.assign(ids + (size_t)r * F, ids + (size_t)(r + 1) * F);
    s.vals.assign(vals + (size_t)r * F, vals + (size_t)(r + 1) * F);
    emit_record((double)labels[r], s, buf);
  }
  if (out && buf.size() <= cap) memcpy(out, buf.data(), buf.size());
  return buf.size();
}

// Returns the number of rows written, or -1 (message in dfmio_libsvm_last_error()). field_size <= 0: rows keep the number
// of id:value pairs they have. threads <= 0: one per hardware thread (at most 16).
long long dfmio_libsvm_convert(const char* input, const char* output, int field_size, long long pad_id, int threads) {
  g_cerr[0] = 0;
  int fd = ::open(input, O_RDONLY);
  struct stat st;
  if (fd < 0 || fstat(fd, &st) != 0) {
    snprintf(g_cerr, sizeof(g_cerr), "cannot open %s", input);
    if (fd >= 0) ::close(fd);
    return -1;
  }
  FILE* fo = fopen(output, "wb");
  if (!fo) {
    snprintf(g_cerr, sizeof(g_cerr), "cannot create %s", output);
    ::close(fd);
    return -1;
  }
  const size_t size = (size_t)st.st_size;
  if (size == 0) {
    ::close(fd);
    fclose(fo);
    return 0;
  }
  const char* base = (const char*)mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
  ::close(fd);
  if (base == MAP_FAILED) {
    snprintf(g_cerr, sizeof(g_cerr), "cannot map %s", input);
    fclose(fo);
    return -1;
  }
  madvise((void*)base, size, MADV_SEQUENTIAL);
  unsigned hw = std::thread::hardware_concurrency();
  const int T = threads > 0 ? std::min(threads, 64) : (int)std::max(1u, std::min(hw ? hw : 4u, 16u));
  const size_t kBlock = (size_t)64 << 20;  // text bytes converted per round (bounds the memory held in output buffers)
  auto next_line_start = [&](size_t pos) {  // first byte after the newline at or after pos (or size)
    const void* nl = pos < size ? memchr(base + pos, '\n', size - pos) : nullptr;
    return nl ? (size_t)((const char*)nl - base) + 1 : size;
  };
  long long rows = 0;
  bool failed = false;
  size_t block_begin = 0;
  std::vector<std::string> outs((size_t)T);
  std::vector<long long> counts((size_t)T), bad_at((size_t)T);
  while (block_begin < size && !failed) {
    const size_t block_end = block_begin + kBlock >= size ? size : next_line_start(block_begin + kBlock - 1);
    std::vector<size_t> cut((size_t)T + 1, block_end);
    cut[0] = block_begin;
    for (int t = 1; t < T; ++t) {
      const size_t want = block_begin + (block_end - block_begin) / (size_t)T * (size_t)t;
      cut[(size_t)t] = std::min(block_end, std::max(cut[(size_t)t - 1], next_line_start(want == 0 ? 0 : want - 1)));
    }
    auto work = [&](int t) {
      Scratch s;
      std::string& out = outs[(size_t)t];
      out.clear();
      counts[(size_t)t] = 0;
      bad_at[(size_t)t] = -1;
      size_t pos = cut[(size_t)t];
      const size_t end = cut[(size_t)t + 1];
      while (pos < end) {
        const size_t nxt = std::min(end, next_line_start(pos));
        if (!blank(base + pos, base + nxt)) {
          if (!convert_line(base + pos, base + nxt, field_size, (int64_t)pad_id, s, out)) {
            bad_at[(size_t)t] = (long long)pos;
            return;
          }
          ++counts[(size_t)t];
        }
        pos = nxt;
      }
    };
    if (T == 1) {
      work(0);
    } else {
      std::vector<std::thread> th;
      for (int t = 0; t < T; ++t) th.emplace_back(work, t);
      for (auto& x : th) x.join();
    }
    for (int t = 0; t < T && !failed; ++t) {
      if (!outs[(size_t)t].empty() && fwrite(outs[(size_t)t].data(), 1, outs[(size_t)t].size(), fo) != outs[(size_t)t].size()) {
        snprintf(g_cerr, sizeof(g_cerr), "short write to %s", output);
        failed = true;
      }
      rows += counts[(size_t)t];
      if (bad_at[(size_t)t] >= 0 && !failed) {
        const size_t p = (size_t)bad_at[(size_t)t];
        const size_t e = std::min(next_line_start(p), p + 80);
        snprintf(g_cerr, sizeof(g_cerr), "malformed libsvm line %lld (byte %zu): %.*s", rows + 1, p, (int)(e - p), base + p);
        failed = true;
      }
    }
    block_begin = block_end;
  }
  munmap((void*)base, size);
  if (fclose(fo) != 0 && !failed) {
    snprintf(g_cerr, sizeof(g_cerr), "cannot finish %s", output);
    failed = true;
  }
  return failed ? -1 : rows;
}

}  // extern "C"
