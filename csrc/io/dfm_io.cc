// Host-side native IO for deepfm_b200 (no CUDA, no TensorFlow):
//   * CRC32C (SSE4.2 `crc32` instruction, software table fallback)                       [reference dep N1]
//   * TFRecord framing reader/writer over files, FIFOs and pipes (read()-only: "pipe mode") [N1, N3]
//   * fixed-schema `tf.train.Example` batch parser {label f32 x1, ids i64 xF, values f32 xF} writing
//     straight into caller-provided (pinned) staging buffers — the vectorised `tf.parse_example` of
//     HVD:76-89 / PS:79-92                                                                 [N2]
//   * prefetching loader: 1 sequential reader thread (shard -> batch, like `dataset.shard().batch()`),
//     P parser threads, ordered hand-off through a ring of caller-owned batch slots          [N4]
// Wire formats: SURVEY Appendix B. C ABI at the bottom; Python binding: data/native.py.
#include <errno.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <algorithm>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

namespace dfmio {

static thread_local char g_err[512] = "";
static void set_err(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ------------------------------------------------------------------------------------------ crc32c
static uint32_t g_tbl[256];
static std::once_flag g_tbl_once;
static void init_tbl() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    g_tbl[i] = c;
  }
}
static uint32_t crc32c_sw(const uint8_t* p, size_t n, uint32_t crc) {
  std::call_once(g_tbl_once, init_tbl);
  uint32_t c = crc ^ 0xFFFFFFFFu;
  for (size_t i = 0; i < n; ++i) c = g_tbl[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}
uint32_t crc32c(const uint8_t* p, size_t n, uint32_t crc = 0) {
#if defined(__SSE4_2__)
  uint64_t c = crc ^ 0xFFFFFFFFu;
  while (n >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8;
    n -= 8;
  }
  uint32_t c32 = (uint32_t)c;
  while (n--) c32 = _mm_crc32_u8(c32, *p++);
  return c32 ^ 0xFFFFFFFFu;
#else
  return crc32c_sw(p, n, crc);
#endif
}
static inline uint32_t mask_crc(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xA282EAD8u; }

// ------------------------------------------------------------------------------------------ record reader
class RecordReader {
 public:
  ~RecordReader() { close_fd(); }
  bool open_path(const char* path) {
    close_fd();
    fd_ = ::open(path, O_RDONLY);
    if (fd_ < 0) {
      set_err("cannot open %s: %s", path, strerror(errno));
      return false;
    }
    buf_.resize(1 << 20);
    beg_ = end_ = 0;
    return true;
  }
  void close_fd() {
    if (fd_ >= 0) ::close(fd_);
    fd_ = -1;
  }
  // 1 = record appended to `out`, 0 = clean EOF, -1 = error
  int next(std::vector<uint8_t>& out, size_t* off, size_t* len, bool check_crc) {
    uint8_t hdr[12];
    int r = read_exact(hdr, 12, true);
    if (r <= 0) return r;
    uint64_t n;
    uint32_t c;
    memcpy(&n, hdr, 8);
    memcpy(&c, hdr + 8, 4);
    if (check_crc && mask_crc(crc32c(hdr, 8)) != c) {
      set_err("corrupted record length CRC");
      return -1;
    }
    if (n > (1ull << 31)) {
      set_err("record too large (%llu bytes)", (unsigned long long)n);
      return -1;
    }
    const size_t o = out.size();
    out.resize(o + n);
    if (read_exact(out.data() + o, n, false) <= 0 && n > 0) return -1;
    uint8_t tail[4];
    if (read_exact(tail, 4, false) <= 0) return -1;
    memcpy(&c, tail, 4);
    if (check_crc && mask_crc(crc32c(out.data() + o, n)) != c) {
      set_err("corrupted record data CRC");
      return -1;
    }
    *off = o;
    *len = n;
    return 1;
  }

 private:
  int read_exact(uint8_t* dst, size_t n, bool eof_ok) {
    size_t got = 0;
    while (got < n) {
      if (beg_ == end_) {
        ssize_t k = ::read(fd_, buf_.data(), buf_.size());
        if (k < 0) {
          if (errno == EINTR) continue;
          set_err("read failed: %s", strerror(errno));
          return -1;
        }
        if (k == 0) {
          if (eof_ok && got == 0) return 0;
          set_err("truncated record");
          return -1;
        }
        beg_ = 0;
        end_ = (size_t)k;
      }
      const size_t take = std::min(n - got, end_ - beg_);
      memcpy(dst + got, buf_.data() + beg_, take);
      beg_ += take;
      got += take;
    }
    return 1;
  }
  int fd_ = -1;
  std::vector<uint8_t> buf_;
  size_t beg_ = 0, end_ = 0;
};

// ------------------------------------------------------------------------------------------ proto parsing
static inline bool rd_varint(const uint8_t*& p, const uint8_t* e, uint64_t* v) {
  uint64_t r = 0;
  int s = 0;
  while (p < e && s <= 63) {
    const uint8_t b = *p++;
    r |= (uint64_t)(b & 0x7F) << s;
    if (!(b & 0x80)) {
      *v = r;
      return true;
    }
    s += 7;
  }
  return false;
}
static inline bool skip_field(const uint8_t*& p, const uint8_t* e, uint32_t wt) {
  uint64_t n;
  switch (wt) {
    case 0: return rd_varint(p, e, &n);
    case 1: p += 8; return p <= e;
    case 2: if (!rd_varint(p, e, &n)) return false; p += n; return p <= e;
    case 5: p += 4; return p <= e;
    default: return false;
  }
}
static inline bool rd_ld(const uint8_t*& p, const uint8_t* e, const uint8_t** b, const uint8_t** be) {
  uint64_t n;
  if (!rd_varint(p, e, &n) || n > (uint64_t)(e - p)) return false;
  *b = p;
  *be = p + n;
  p += n;
  return true;
}

static int parse_floats(const uint8_t* p, const uint8_t* e, float* out, int cap) {  // FloatList body
  int cnt = 0;
  while (p < e) {
    uint64_t tag;
    if (!rd_varint(p, e, &tag)) return -1;
    if ((tag >> 3) == 1 && (tag & 7) == 2) {
      const uint8_t *b, *be;
      if (!rd_ld(p, e, &b, &be)) return -1;
      const int n = (int)((be - b) / 4);
      if (cnt + n > cap) return -2;
      memcpy(out + cnt, b, (size_t)n * 4);
      cnt += n;
    } else if ((tag >> 3) == 1 && (tag & 7) == 5) {
      if (p + 4 > e) return -1;
      if (cnt + 1 > cap) return -2;
      memcpy(out + cnt, p, 4);
      p += 4;
      ++cnt;
    } else if (!skip_field(p, e, tag & 7)) {
      return -1;
    }
  }
  return cnt;
}
static int parse_ints(const uint8_t* p, const uint8_t* e, int32_t* out, int cap) {  // Int64List body -> int32
  int cnt = 0;
  while (p < e) {
    uint64_t tag;
    if (!rd_varint(p, e, &tag)) return -1;
    if ((tag >> 3) == 1 && (tag & 7) == 2) {
      const uint8_t *b, *be;
      if (!rd_ld(p, e, &b, &be)) return -1;
      while (b < be) {
        uint64_t v;
        if (!rd_varint(b, be, &v)) return -1;
        if (cnt + 1 > cap) return -2;
        if (v > 0x7FFFFFFFull) return -3;
        out[cnt++] = (int32_t)v;
      }
    } else if ((tag >> 3) == 1 && (tag & 7) == 0) {
      uint64_t v;
      if (!rd_varint(p, e, &v)) return -1;
      if (cnt + 1 > cap) return -2;
      if (v > 0x7FFFFFFFull) return -3;
      out[cnt++] = (int32_t)v;
    } else if (!skip_field(p, e, tag & 7)) {
      return -1;
    }
  }
  return cnt;
}

// One serialized Example -> label, ids[F], vals[F]. 0 on success.
static int parse_example(const uint8_t* p, size_t len, int F, float* label, int32_t* ids, float* vals) {
  const uint8_t* e = p + len;
  bool got_l = false, got_i = false, got_v = false;
  while (p < e) {
    uint64_t tag;
    if (!rd_varint(p, e, &tag)) goto malformed;
    if (tag != ((1u << 3) | 2)) {
      if (!skip_field(p, e, tag & 7)) goto malformed;
      continue;
    }
    const uint8_t *f, *fe;
    if (!rd_ld(p, e, &f, &fe)) goto malformed;
    while (f < fe) {  // Features: repeated map entries
      uint64_t t2;
      if (!rd_varint(f, fe, &t2)) goto malformed;
      if (t2 != ((1u << 3) | 2)) {
        if (!skip_field(f, fe, t2 & 7)) goto malformed;
        continue;
      }
      const uint8_t *m, *me;
      if (!rd_ld(f, fe, &m, &me)) goto malformed;
      const uint8_t* key = nullptr;
      size_t klen = 0;
      const uint8_t *val = nullptr, *vale = nullptr;
      while (m < me) {
        uint64_t t3;
        if (!rd_varint(m, me, &t3)) goto malformed;
        if ((t3 & 7) != 2) {
          if (!skip_field(m, me, t3 & 7)) goto malformed;
          continue;
        }
        const uint8_t *b, *be;
        if (!rd_ld(m, me, &b, &be)) goto malformed;
        if ((t3 >> 3) == 1) { key = b; klen = (size_t)(be - b); }
        else if ((t3 >> 3) == 2) { val = b; vale = be; }
      }
      if (!key || !val) continue;
      int which = -1;
      if (klen == 5 && !memcmp(key, "label", 5)) which = 0;
      else if (klen == 3 && !memcmp(key, "ids", 3)) which = 1;
      else if (klen == 6 && !memcmp(key, "values", 6)) which = 2;
      if (which < 0) continue;
      // Feature: oneof { bytes_list=1, float_list=2, int64_list=3 }
      const uint8_t* q = val;
      while (q < vale) {
        uint64_t t4;
        if (!rd_varint(q, vale, &t4)) goto malformed;
        if ((t4 & 7) != 2) {
          if (!skip_field(q, vale, t4 & 7)) goto malformed;
          continue;
        }
        const uint8_t *b, *be;
        if (!rd_ld(q, vale, &b, &be)) goto malformed;
        const uint32_t fld = (uint32_t)(t4 >> 3);
        if (which == 0 && fld == 2) {
          float tmp[2];
          if (parse_floats(b, be, tmp, 1) != 1) { set_err("feature 'label' must hold exactly 1 float"); return -1; }
          *label = tmp[0];
          got_l = true;
        } else if (which == 1 && fld == 3) {
          const int n = parse_ints(b, be, ids, F);
          if (n == -3) { set_err("feature id does not fit in int32"); return -1; }
          if (n != F) { set_err("feature 'ids': expected %d values per record, got %s", F, n == -2 ? "more" : "fewer/malformed"); return -1; }
          got_i = true;
        } else if (which == 2 && fld == 2) {
          const int n = parse_floats(b, be, vals, F);
          if (n != F) { set_err("feature 'values': expected %d values per record, got %s", F, n == -2 ? "more" : "fewer/malformed"); return -1; }
          got_v = true;
        }
      }
    }
  }
  if (!got_l || !got_i || !got_v) {
    set_err("Example is missing feature %s", !got_l ? "'label'" : !got_i ? "'ids'" : "'values'");
    return -1;
  }
  return 0;
malformed:
  set_err("malformed Example protobuf");
  return -1;
}

// ------------------------------------------------------------------------------------------ loader
// A regular file mapped read-only; kept alive (shared_ptr) by every raw batch that points into it.
struct MappedFile {
  const uint8_t* base = nullptr;
  size_t size = 0;
  ~MappedFile() {
    if (base) munmap(const_cast<uint8_t*>(base), size);
  }
  static std::shared_ptr<MappedFile> open_path(const char* path) {
    struct stat pre;
    // never open() a FIFO here: opening and closing its read end would break the writer feeding it (pipe mode)
    if (::stat(path, &pre) != 0 || !S_ISREG(pre.st_mode) || pre.st_size == 0) return nullptr;
    int fd = ::open(path, O_RDONLY);
    if (fd < 0) return nullptr;
    struct stat st;
    if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode) || st.st_size == 0) {
      ::close(fd);
      return nullptr;
    }
    // MAP_POPULATE: the reader thread builds the page tables once; parser threads then never take page faults (and the
    // process-wide mmap lock) while they work.
    void* p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
    ::close(fd);
    if (p == MAP_FAILED) return nullptr;
    madvise(p, (size_t)st.st_size, MADV_WILLNEED);  // (advice values are not flags: one call per advice)
    auto m = std::make_shared<MappedFile>();
    m->base = (const uint8_t*)p;
    m->size = (size_t)st.st_size;
    return m;
  }
};

struct RawBatch {
  long seq = 0;
  int rows = 0;
  // streamed sources (FIFO / pipe): record bytes are copied here and located by off/len
  std::vector<uint8_t> bytes;
  std::vector<size_t> off, len;
  // mapped files: zero-copy pointers into the mapping; the payload CRC is verified by the PARSER threads
  std::vector<const uint8_t*> ptr;
  std::vector<std::shared_ptr<MappedFile>> keep;
  bool mapped = false;
};

class Loader {
  static constexpr std::chrono::microseconds kSlotPoll{100};

 public:
  Loader(std::vector<std::string> paths, int F, int batch, int shard_n, int shard_i, int epochs, int threads,
         bool drop_remainder, bool check_crc, uint8_t* slots, int nslots, size_t slot_stride, bool raw = false,
         uint64_t shuffle_seed = 0)
      : paths_(std::move(paths)), F_(F), B_(batch), shard_n_(shard_n < 1 ? 1 : shard_n), shard_i_(shard_i),
        epochs_(epochs), drop_(drop_remainder), crc_(check_crc), slots_(slots), nslots_(nslots),
        stride_(slot_stride), raw_mode_(raw), shuffle_seed_(shuffle_seed) {
    for (int i = 0; i < nslots_; ++i) free_slots_.push_back(i);
    const char* no_index = getenv("DFM_LOADER_INDEXED");
    if (!(no_index && no_index[0] == '0') && build_index()) {
      // indexed mode (regular files only): record boundaries are walked ONCE; afterwards batch k of the sharded record
      // sequence is an O(1) lookup, so `threads` workers build batches fully in parallel — no serial reader stage
      indexed_ = true;
      perm_half_ = 1;
      while ((1ull << (2 * perm_half_)) < index_.size()) ++perm_half_;
      const long full = (long)(index_.size() / (size_t)B_);
      per_epoch_batches_ = full + ((!drop_ && index_.size() % (size_t)B_) ? 1 : 0);
      total_batches_ = per_epoch_batches_ * epochs_;
      reader_finished_ = true;
      // raw mode only copies bytes (one worker moves ~3.5 GB/s): two workers by default, `threads` if given
      const int workers = threads >= 1 ? threads : (raw_mode_ ? 2 : 1);
      for (int i = 0; i < workers; ++i) parsers_.emplace_back([this] { indexed_worker(); });
      return;
    }
    if (raw_mode_) {
      // raw mode: slots carry the serialized records themselves (offsets[B+1] | bytes); the GPU parses them
      // (csrc/parse_gpu.cu), so there are no parser threads and the reader fills the slots directly
      reader_ = std::thread([this] { reader_main_raw(); });
    } else {
      reader_ = std::thread([this] { reader_main(); });
      for (int i = 0; i < (threads < 1 ? 1 : threads); ++i) parsers_.emplace_back([this] { parser_main(); });
    }
  }
  static size_t raw_slot_bytes(int B, int max_record_bytes) {
    auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
    return al((size_t)(B + 1) * 4) + al((size_t)B * max_record_bytes);
  }
  ~Loader() {
    {
      std::lock_guard<std::mutex> l(mu_);
      stop_ = true;
      stop_flag_.store(true);
    }
    cv_work_.notify_all();
    cv_free_.notify_all();
    cv_done_.notify_all();
    cv_raw_space_.notify_all();
    if (reader_.joinable()) reader_.join();
    for (auto& t : parsers_) t.join();
  }
  static size_t slot_bytes(int B, int F) {  // labels f32[B] | ids i32[B*F] | vals f32[B*F], 256-byte aligned parts
    auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
    return al((size_t)B * 4) + 2 * al((size_t)B * F * 4);
  }
  float* slot_labels(int s) const { return reinterpret_cast<float*>(slots_ + (size_t)s * stride_); }
  int32_t* slot_ids(int s) const {
    return reinterpret_cast<int32_t*>(slots_ + (size_t)s * stride_ + (((size_t)B_ * 4 + 255) & ~size_t(255)));
  }
  float* slot_vals(int s) const {
    return reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(slot_ids(s)) + (((size_t)B_ * F_ * 4 + 255) & ~size_t(255)));
  }

  // Blocks until the next batch (in order) is parsed. Returns rows (>0), 0 at end of data, -1 on error.
  int next(int* slot_out) {
    std::unique_lock<std::mutex> l(mu_);
    for (;;) {
      if (!err_.empty()) {
        set_err("%s", err_.c_str());
        return -1;
      }
      if (!done_.empty() && done_.front().seq == next_seq_) {
        Done d = done_.front();
        done_.pop_front();
        ++next_seq_;
        *slot_out = d.slot;
        return d.rows;
      }
      if (reader_finished_ && next_seq_ >= total_batches_) return 0;
      cv_done_.wait(l);
    }
  }
  void release(int slot) {
    {
      std::lock_guard<std::mutex> l(mu_);
      free_slots_.push_back(slot);
    }
    // No notification by default: two futex wake-ups per batch cost the CONSUMER thread (the training loop) 16-25 us
    // per step. Threads that wait for a free slot poll with a timed wait (kSlotPoll) instead; a freed slot is picked up
    // at most that much later, which the depth of the slot ring hides. DFM_LOADER_NOTIFY=1 restores the wake-ups.
    if (notify_on_release_) {
      cv_free_.notify_one();
      cv_work_.notify_one();
    }
  }
  long records_read() const { return records_read_.load(); }

 private:
  struct Done { long seq; int slot; int rows; };

  void fail(const std::string& m) {
    std::lock_guard<std::mutex> l(mu_);
    if (err_.empty()) err_ = m;
    cv_done_.notify_all();
    cv_work_.notify_all();
  }

  void reader_main() {
    long seq = 0, rec_idx = 0;
    std::unique_ptr<RawBatch> cur(new RawBatch());
    RecordReader rr;
    std::vector<uint8_t> scratch;
    for (int ep = 0; ep < epochs_ && !stopped(); ++ep) {
      for (const auto& path : paths_) {
        // regular files: walk the record headers of a read-only mapping (no copies; CRC of the payload is checked in
        // parallel by the parser threads). FIFOs / pipes fall through to the streaming reader below.
        std::shared_ptr<MappedFile> mf_cached;
        {
          auto it = maps_.find(path);  // one mapping per file, reused by every epoch
          if (it != maps_.end()) mf_cached = it->second;
          else if ((mf_cached = MappedFile::open_path(path.c_str()))) maps_[path] = mf_cached;
        }
        if (std::shared_ptr<MappedFile> mf = mf_cached) {
          size_t pos = 0, pf_pos = 0;
          while (pos < mf->size) {
            if ((rec_idx & 255) == 0 && stopped()) return;
            // The walk is a chain of dependent loads (the next header's address comes from this header's length), i.e.
            // one DRAM latency per record. Streaming software prefetch 64-128 KB ahead turns it into a bandwidth-bound scan.
            if (pos + (64u << 10) > pf_pos) {
              const size_t end = std::min(mf->size, pf_pos + (64u << 10));
              for (size_t q = pf_pos; q < end; q += 64) __builtin_prefetch(mf->base + q, 0, 0);
              pf_pos = end;
            }
            if (pos + 12 > mf->size) { fail("truncated record in " + path); return; }
            uint64_t n;
            uint32_t c;
            memcpy(&n, mf->base + pos, 8);
            memcpy(&c, mf->base + pos + 8, 4);
            (void)c;  // the length CRC is verified together with the payload CRC by the parser threads
            if (n > (1ull << 31) || pos + 12 + n + 4 > mf->size) { fail("truncated or corrupted record in " + path); return; }
            const uint8_t* data = mf->base + pos + 12;
            pos += 12 + n + 4;
            const bool mine = (rec_idx % shard_n_) == shard_i_;
            ++rec_idx;
            if (!mine) continue;
            if (!cur->mapped && cur->rows > 0) {  // never mix streamed and mapped records in one batch
              fail("internal: mixed batch");
              return;
            }
            if (!cur->mapped) {
              cur->mapped = true;
              cur->ptr.reserve(B_); cur->len.reserve(B_);
            }
            cur->ptr.push_back(data);
            cur->len.push_back((size_t)n);
            if (cur->keep.empty() || cur->keep.back() != mf) cur->keep.push_back(mf);
            if (++cur->rows == B_) {
              cur->seq = seq++;
              records_read_.fetch_add(B_);
              if (!push_raw(std::move(cur))) return;
              cur.reset(new RawBatch());
            }
          }
          continue;
        }
        if (cur->mapped) {  // a partial mapped batch followed by a streamed source: materialise it
          for (int r = 0; r < cur->rows; ++r) {
            cur->off.push_back(cur->bytes.size());
            cur->bytes.insert(cur->bytes.end(), cur->ptr[r], cur->ptr[r] + cur->len[r]);
          }
          cur->ptr.clear(); cur->keep.clear(); cur->mapped = false;
        }
        if (!rr.open_path(path.c_str())) { fail(g_err); return; }
        for (;;) {
          if (stopped()) return;
          if (rec_idx % shard_n_ != shard_i_) {  // `dataset.shard(n, i)`: keep every n-th record
            scratch.clear();
            size_t o, n;
            int r = rr.next(scratch, &o, &n, false);
            if (r < 0) { fail(g_err); return; }
            if (r == 0) break;
            ++rec_idx;
            continue;
          }
          size_t o, n;
          int r = rr.next(cur->bytes, &o, &n, crc_);
          if (r < 0) { fail(std::string(g_err) + " in " + path); return; }
          if (r == 0) break;
          ++rec_idx;
          records_read_.fetch_add(1);
          cur->off.push_back(o);
          cur->len.push_back(n);
          if (++cur->rows == B_) {
            cur->seq = seq++;
            if (!push_raw(std::move(cur))) return;
            cur.reset(new RawBatch());
          }
        }
      }
      // end of one pass over the data: `batch(B, drop_remainder)` comes BEFORE `repeat(num_epochs)` in the reference
      // (HVD:122-126), so the remainder of every epoch is dropped (or emitted as a short batch), never carried over
      if (cur->rows > 0) {
        if (!drop_) {
          cur->seq = seq++;
          if (!push_raw(std::move(cur))) return;
        }
        cur.reset(new RawBatch());
      }
    }
    {
      std::lock_guard<std::mutex> l(mu_);
      total_batches_ = seq;
      reader_finished_ = true;
    }
    cv_work_.notify_all();
    cv_done_.notify_all();
  }

  // ---- indexed mode ---------------------------------------------------------------------------------------
  struct RecRef {
    const uint8_t* ptr;
    uint32_t len;
  };
  // Files are scanned in parallel (one task per file: map + populate + walk the length fields); the walk keeps only the
  // record lengths (4 bytes each), from which the second, arithmetic-only pass places every record of THIS shard at its
  // final position - global record numbers, and hence shard membership, are known once the per-file counts are.
  bool build_index() {
    const int nf = (int)paths_.size();
    std::vector<std::shared_ptr<MappedFile>> maps((size_t)nf);
    std::vector<std::vector<uint32_t>> lens((size_t)nf);
    std::atomic<bool> ok{true};
    auto scan = [&](int fi) {
      std::shared_ptr<MappedFile> mf = MappedFile::open_path(paths_[(size_t)fi].c_str());
      if (!mf) { ok.store(false); return; }  // FIFO / pipe / empty file: use the streaming reader
      maps[(size_t)fi] = mf;
      std::vector<uint32_t>& out = lens[(size_t)fi];
      out.reserve(mf->size / 256 + 16);
      size_t pos = 0;
      while (pos < mf->size) {
        if (pos + 12 > mf->size) { ok.store(false); return; }  // the streaming path reports corruption with context
        uint64_t n;
        memcpy(&n, mf->base + pos, 8);
        if (n > (1ull << 31) || pos + 16 + n > mf->size) { ok.store(false); return; }
        out.push_back((uint32_t)n);
        pos += 16 + n;
      }
    };
    unsigned hw = std::thread::hardware_concurrency();
    const int T = std::max(1, std::min({nf, (int)(hw ? hw : 4), 16}));
    auto run_parallel = [&](const std::function<void(int)>& fn) {
      if (T == 1) {
        for (int fi = 0; fi < nf; ++fi) fn(fi);
        return;
      }
      std::vector<std::thread> th;
      for (int t = 0; t < T; ++t)
        th.emplace_back([&, t] { for (int fi = t; fi < nf && ok.load(); fi += T) fn(fi); });
      for (auto& x : th) x.join();
    };
    run_parallel(scan);
    if (!ok.load()) return false;
    const long n = shard_n_, sidx = shard_i_;
    auto sel_before = [&](long b) { return b > sidx ? (b - sidx + n - 1) / n : 0L; };  // #{i < b : i % n == sidx}
    std::vector<long> base((size_t)nf + 1, 0);
    for (int fi = 0; fi < nf; ++fi) base[(size_t)fi + 1] = base[(size_t)fi] + (long)lens[(size_t)fi].size();
    index_.resize((size_t)sel_before(base[(size_t)nf]));
    run_parallel([&](int fi) {
      const uint8_t* p = maps[(size_t)fi]->base;
      long idx = base[(size_t)fi];
      size_t out = (size_t)sel_before(idx), pos = 0;
      for (uint32_t len : lens[(size_t)fi]) {
        if (idx % n == sidx) index_[out++] = RecRef{p + pos + 12, len};
        ++idx;
        pos += 16 + (size_t)len;
      }
    });
    for (int fi = 0; fi < nf; ++fi) maps_[paths_[(size_t)fi]] = maps[(size_t)fi];
    return true;
  }
  // ---- per-epoch shuffle (indexed mode): position p of epoch e reads record permute(p) of this rank's shard, where
  // permute is a keyed bijection on [0, n) - a 4-round Feistel network on 2 * perm_half_ bits with cycle walking (the
  // domain is less than 4n, so under 4 walks on average). No permutation array is stored, every worker evaluates it
  // independently, the order is reproducible from (seed, epoch), and every record is still read exactly once per epoch.
  static uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
  }
  uint64_t epoch_key_of(uint64_t epoch) const { return shuffle_seed_ ? splitmix64(shuffle_seed_ ^ splitmix64(epoch + 1)) : 0; }
  const RecRef& record_at(size_t pos, uint64_t epoch_key) const {
    if (!shuffle_seed_) return index_[pos];
    const uint64_t n = index_.size(), mask = (1ull << perm_half_) - 1;
    uint64_t x = pos;
    do {
      uint64_t l = x >> perm_half_, r = x & mask;
      for (int i = 0; i < 4; ++i) {
        const uint64_t f = splitmix64(r ^ (epoch_key + (uint64_t)i * 0xD1B54A32D192ED03ull)) & mask;
        const uint64_t nl = r;
        r = l ^ f;
        l = nl;
      }
      x = (l << perm_half_) | r;
    } while (x >= n);
    return index_[x];
  }
  bool check_record_crc(const RecRef& r) const {
    uint32_t hc, dc;
    memcpy(&hc, r.ptr - 4, 4);
    memcpy(&dc, r.ptr + r.len, 4);
    return mask_crc(crc32c(r.ptr - 12, 8)) == hc && mask_crc(crc32c(r.ptr, r.len)) == dc;
  }
  void indexed_worker() {
    const size_t off_bytes = ((size_t)(B_ + 1) * 4 + 255) & ~size_t(255);
    for (;;) {
      long k;
      int slot;
      {
        // a batch number is handed out only together with a free slot, so the oldest outstanding batch always owns a
        // slot and in-order delivery cannot deadlock
        std::unique_lock<std::mutex> l(mu_);
        while (!(stop_ || !err_.empty() || next_batch_ >= total_batches_ || !free_slots_.empty())) cv_free_.wait_for(l, kSlotPoll);
        if (stop_ || !err_.empty() || next_batch_ >= total_batches_) return;
        slot = free_slots_.front();
        free_slots_.pop_front();
        k = next_batch_++;
      }
      const size_t first = (size_t)(k % per_epoch_batches_) * (size_t)B_;
      const int rows = (int)std::min<size_t>((size_t)B_, index_.size() - first);
      const uint64_t epoch_key = epoch_key_of((uint64_t)(k / per_epoch_batches_));
      bool bad = false;
      if (raw_mode_) {
        int32_t* offs = reinterpret_cast<int32_t*>(slots_ + (size_t)slot * stride_);
        uint8_t* bytes = slots_ + (size_t)slot * stride_ + off_bytes;
        const size_t cap = stride_ - off_bytes;
        size_t used = 0;
        for (int r = 0; r < rows && !bad; ++r) {
          const RecRef& rr = record_at(first + r, epoch_key);
          if (r + 6 < rows) {
            const RecRef& pf = record_at(first + r + 6, epoch_key);
            for (uint32_t o = 0; o < pf.len + 4; o += 64) __builtin_prefetch(pf.ptr - 12 + o, 0, 0);
          }
          if (crc_ && !check_record_crc(rr)) { set_err("corrupted record CRC"); bad = true; break; }
          if (used + rr.len > cap) { set_err("raw slot overflow: raise max_record_bytes"); bad = true; break; }
          offs[r] = (int32_t)used;
          memcpy(bytes + used, rr.ptr, rr.len);
          used += rr.len;
          offs[r + 1] = (int32_t)used;
        }
      } else {
        float* lab = slot_labels(slot);
        int32_t* ids = slot_ids(slot);
        float* vals = slot_vals(slot);
        for (int r = 0; r < rows && !bad; ++r) {
          const RecRef& rr = record_at(first + r, epoch_key);
          if (r + 6 < rows) {  // records stream from DRAM exactly once: fetch a few records ahead of the parser
            const RecRef& pf = record_at(first + r + 6, epoch_key);
            for (uint32_t o = 0; o < pf.len + 4; o += 64) __builtin_prefetch(pf.ptr - 12 + o, 0, 0);
          }
          if (crc_ && !check_record_crc(rr)) { set_err("corrupted record CRC"); bad = true; break; }
          if (parse_example(rr.ptr, rr.len, F_, lab + r, ids + (size_t)r * F_, vals + (size_t)r * F_)) bad = true;
        }
      }
      if (bad) {
        fail(g_err);
        cv_free_.notify_all();
        return;
      }
      records_read_.fetch_add(rows);
      {
        std::lock_guard<std::mutex> l(mu_);
        Done d{k, slot, rows};
        auto it = done_.begin();
        while (it != done_.end() && it->seq < d.seq) ++it;
        done_.insert(it, d);
      }
      cv_done_.notify_one();
    }
  }

  // ---- raw mode -------------------------------------------------------------------------------------------
  int take_free_slot() {
    std::unique_lock<std::mutex> l(mu_);
    while (!(stop_ || !free_slots_.empty())) cv_free_.wait_for(l, kSlotPoll);
    if (stop_) return -1;
    const int s = free_slots_.front();
    free_slots_.pop_front();
    return s;
  }
  void publish(long seq, int slot, int rows) {
    {
      std::lock_guard<std::mutex> l(mu_);
      done_.push_back(Done{seq, slot, rows});
    }
    cv_done_.notify_one();
  }
  void reader_main_raw() {
    const size_t off_bytes = ((size_t)(B_ + 1) * 4 + 255) & ~size_t(255);
    const size_t cap = stride_ - off_bytes;
    long seq = 0, rec_idx = 0;
    int slot = take_free_slot();
    if (slot < 0) return;
    int rows = 0;
    size_t used = 0;
    auto offs = [&](int s) { return reinterpret_cast<int32_t*>(slots_ + (size_t)s * stride_); };
    auto bytes = [&](int s) { return slots_ + (size_t)s * stride_ + off_bytes; };
    auto append = [&](const uint8_t* hdr, const uint8_t* data, size_t n, const uint8_t* tail, const std::string& path) -> bool {
      if (crc_) {
        uint32_t hc, dc;
        memcpy(&hc, hdr + 8, 4);
        memcpy(&dc, tail, 4);
        if (mask_crc(crc32c(hdr, 8)) != hc) { fail("corrupted record length CRC in " + path); return false; }
        if (mask_crc(crc32c(data, n)) != dc) { fail("corrupted record data CRC in " + path); return false; }
      }
      if (used + n > cap) { fail("raw slot overflow: raise max_record_bytes"); return false; }
      offs(slot)[rows] = (int32_t)used;
      memcpy(bytes(slot) + used, data, n);
      used += n;
      ++rows;
      offs(slot)[rows] = (int32_t)used;
      if (rows == B_) {
        records_read_.fetch_add(rows);
        publish(seq++, slot, rows);
        slot = take_free_slot();
        if (slot < 0) return false;
        rows = 0;
        used = 0;
      }
      return true;
    };
    RecordReader rr;
    std::vector<uint8_t> tmp;
    for (int ep = 0; ep < epochs_ && !stopped(); ++ep) {
      for (const auto& path : paths_) {
        std::shared_ptr<MappedFile> mf;
        {
          auto it = maps_.find(path);
          if (it != maps_.end()) mf = it->second;
          else if ((mf = MappedFile::open_path(path.c_str()))) maps_[path] = mf;
        }
        if (mf) {
          size_t pos = 0;
          while (pos < mf->size) {
            if ((rec_idx & 255) == 0 && stopped()) return;
            if (pos + 12 > mf->size) { fail("truncated record in " + path); return; }
            uint64_t n;
            memcpy(&n, mf->base + pos, 8);
            if (n > (1ull << 31) || pos + 12 + n + 4 > mf->size) { fail("truncated or corrupted record in " + path); return; }
            const uint8_t* hdr = mf->base + pos;
            pos += 12 + n + 4;
            const bool mine = (rec_idx % shard_n_) == shard_i_;
            ++rec_idx;
            if (!mine) continue;
            if (!append(hdr, hdr + 12, (size_t)n, hdr + 12 + n, path)) return;
          }
        } else {
          if (!rr.open_path(path.c_str())) { fail(g_err); return; }
          for (;;) {
            if (stopped()) return;
            tmp.clear();
            size_t o, n;
            const bool mine = (rec_idx % shard_n_) == shard_i_;
            int r = rr.next(tmp, &o, &n, crc_ && mine);
            if (r < 0) { fail(std::string(g_err) + " in " + path); return; }
            if (r == 0) break;
            ++rec_idx;
            if (!mine) continue;
            const bool keep_crc = crc_;
            crc_ = false;  // already verified by the streaming reader
            const bool ok = append(nullptr, tmp.data() + o, n, nullptr, path);
            crc_ = keep_crc;
            if (!ok) return;
          }
        }
      }
      if (rows > 0) {  // per-epoch remainder: dropped or emitted short, never carried into the next epoch
        if (!drop_) {
          records_read_.fetch_add(rows);
          publish(seq++, slot, rows);
          slot = take_free_slot();
          if (slot < 0) return;
        }
        rows = 0;
        used = 0;
      }
    }
    {
      std::lock_guard<std::mutex> l(mu_);
      total_batches_ = seq;
      reader_finished_ = true;
    }
    cv_done_.notify_all();
  }

  bool push_raw(std::unique_ptr<RawBatch> b) {
    std::unique_lock<std::mutex> l(mu_);
    cv_raw_space_.wait(l, [this] { return stop_ || (int)raw_.size() < nslots_ + 2; });
    if (stop_) return false;
    raw_.push_back(std::move(b));
    l.unlock();
    cv_work_.notify_one();
    return true;
  }

  void parser_main() {
    for (;;) {
      std::unique_ptr<RawBatch> b;
      int slot = -1;
      {
        std::unique_lock<std::mutex> l(mu_);
        // take work only when a slot is free; the OLDEST raw batch always gets the next free slot (no deadlock)
        while (!(stop_ || !err_.empty() || (!raw_.empty() && !free_slots_.empty()) || (reader_finished_ && raw_.empty())))
          cv_work_.wait_for(l, kSlotPoll);  // woken by the reader for new work; slot frees are found by polling
        if (stop_ || !err_.empty() || (raw_.empty() && reader_finished_)) return;
        b = std::move(raw_.front());
        raw_.pop_front();
        slot = free_slots_.front();
        free_slots_.pop_front();
      }
      cv_raw_space_.notify_one();
      float* lab = slot_labels(slot);
      int32_t* ids = slot_ids(slot);
      float* vals = slot_vals(slot);
      bool bad = false;
      for (int r = 0; r < b->rows && !bad; ++r) {
        const uint8_t* rec = b->mapped ? b->ptr[r] : b->bytes.data() + b->off[r];
        if (b->mapped && crc_) {  // header (length) CRC sits 4 bytes before the payload, over the 8 length bytes
          uint32_t hc;
          memcpy(&hc, rec - 4, 4);
          if (mask_crc(crc32c(rec - 12, 8)) != hc) {
            set_err("corrupted record length CRC");
            bad = true;
            continue;
          }
        }
        uint32_t want = 0;
        if (b->mapped && crc_) memcpy(&want, rec + b->len[r], 4);  // trailing payload CRC
        if (b->mapped && crc_ && mask_crc(crc32c(rec, b->len[r])) != want) {
          set_err("corrupted record data CRC");
          bad = true;
        } else if (parse_example(rec, b->len[r], F_, lab + r, ids + (size_t)r * F_, vals + (size_t)r * F_)) {
          bad = true;
        }
      }
      if (bad) {
        fail(g_err);
        return;
      }
      {
        std::lock_guard<std::mutex> l(mu_);
        Done d{b->seq, slot, b->rows};
        auto it = done_.begin();
        while (it != done_.end() && it->seq < d.seq) ++it;
        done_.insert(it, d);
      }
      cv_done_.notify_one();  // only the consumer waits here
    }
  }
  bool stopped() { return stop_flag_.load(std::memory_order_relaxed); }  // lock-free: polled once per record

  std::vector<std::string> paths_;
  int F_, B_, shard_n_, shard_i_, epochs_;
  bool drop_, crc_;
  uint8_t* slots_;
  int nslots_;
  size_t stride_;
  bool raw_mode_ = false;
  uint64_t shuffle_seed_ = 0;  // 0 = file order
  int perm_half_ = 1;
  bool indexed_ = false;
  std::vector<RecRef> index_;
  long per_epoch_batches_ = 0, next_batch_ = 0;
  std::mutex mu_;
  std::condition_variable cv_work_, cv_free_, cv_done_, cv_raw_space_;
  const bool notify_on_release_ = [] { const char* e = getenv("DFM_LOADER_NOTIFY"); return e && e[0] == '1'; }();
  std::deque<std::unique_ptr<RawBatch>> raw_;
  std::deque<int> free_slots_;
  std::deque<Done> done_;
  long next_seq_ = 0, total_batches_ = 0;
  bool reader_finished_ = false, stop_ = false;
  std::string err_;
  std::atomic<long> records_read_{0};
  // own cache line: the reader polls it, while mu_/queues next to it are written by every other thread (false sharing
  // made each poll a cross-core miss: 240 ns per record instead of 90)
  alignas(128) std::atomic<bool> stop_flag_{false};
  char pad_[120];
  std::map<std::string, std::shared_ptr<MappedFile>> maps_;  // reader thread only
  std::thread reader_;
  std::vector<std::thread> parsers_;
};

}  // namespace dfmio

using namespace dfmio;

extern "C" {

const char* dfmio_last_error() { return g_err; }
uint32_t dfmio_crc32c(const uint8_t* p, size_t n) { return crc32c(p, n); }
uint32_t dfmio_masked_crc32c(const uint8_t* p, size_t n) { return mask_crc(crc32c(p, n)); }

// ---- writer: frames `n` payload bytes as one TFRecord and appends to fd-backed FILE*
void* dfmio_writer_open(const char* path) {
  FILE* f = fopen(path, "wb");
  if (!f) set_err("cannot open %s for writing: %s", path, strerror(errno));
  return f;
}
int dfmio_writer_write(void* w, const uint8_t* data, uint64_t n) {
  FILE* f = (FILE*)w;
  uint8_t hdr[12];
  memcpy(hdr, &n, 8);
  const uint32_t c1 = mask_crc(crc32c(hdr, 8));
  memcpy(hdr + 8, &c1, 4);
  const uint32_t c2 = mask_crc(crc32c(data, n));
  if (fwrite(hdr, 1, 12, f) != 12 || fwrite(data, 1, n, f) != n || fwrite(&c2, 1, 4, f) != 4) {
    set_err("write failed: %s", strerror(errno));
    return -1;
  }
  return 0;
}
int dfmio_writer_close(void* w) { return fclose((FILE*)w); }

// ---- sequential reader (used for tests and the Python fallback path)
void* dfmio_reader_open(const char* path) {
  RecordReader* r = new RecordReader();
  if (!r->open_path(path)) {
    delete r;
    return nullptr;
  }
  return r;
}
// Returns record length (>=0) with bytes copied into buf (cap bytes), -1 = EOF, -2 = error, -3 = buffer too small.
long dfmio_reader_next(void* rd, uint8_t* buf, size_t cap, int check_crc) {
  static thread_local std::vector<uint8_t> tmp;
  tmp.clear();
  size_t o, n;
  int r = ((RecordReader*)rd)->next(tmp, &o, &n, check_crc != 0);
  if (r == 0) return -1;
  if (r < 0) return -2;
  if (n > cap) return -3;
  memcpy(buf, tmp.data(), n);
  return (long)n;
}
void dfmio_reader_close(void* rd) { delete (RecordReader*)rd; }

// ---- record count of a TFRecord file by walking the length headers of a read-only mapping (no payload reads);
// used to agree on a common step count across ranks before an epoch (uneven shards would otherwise leave one rank
// waiting in a collective: the Horovod failure mode of PDF p22-23). Returns -1 on error.
long dfmio_count_records(const char* path) {
  std::shared_ptr<MappedFile> mf = MappedFile::open_path(path);
  if (!mf) {
    set_err("cannot map %s", path);
    return -1;
  }
  long n = 0;
  size_t pos = 0;
  while (pos < mf->size) {
    if (pos + 12 > mf->size) { set_err("truncated record in %s", path); return -1; }
    uint64_t len;
    memcpy(&len, mf->base + pos, 8);
    if (len > (1ull << 31) || pos + 16 + len > mf->size) { set_err("truncated or corrupted record in %s", path); return -1; }
    pos += 16 + len;
    ++n;
  }
  return n;
}

// ---- batch parse: `nrec` serialized Examples (concatenated in `bytes`, located by off/len) -> dense arrays
int dfmio_parse_batch(const uint8_t* bytes, const uint64_t* off, const uint64_t* len, int nrec, int F, float* labels,
                      int32_t* ids, float* vals) {
  for (int r = 0; r < nrec; ++r)
    if (parse_example(bytes + off[r], len[r], F, labels + r, ids + (size_t)r * F, vals + (size_t)r * F)) return -1;
  return 0;
}

// ---- prefetching loader
uint64_t dfmio_loader_slot_bytes(int batch, int F) { return Loader::slot_bytes(batch, F); }
uint64_t dfmio_loader_raw_slot_bytes(int batch, int max_record_bytes) { return Loader::raw_slot_bytes(batch, max_record_bytes); }
// Raw mode: slots hold `int32 offsets[batch+1]` followed (256-byte aligned) by the serialized records back to back.
void* dfmio_loader_create_raw(const char** paths, int npaths, int batch, int shard_n, int shard_i, int epochs,
                              int drop_remainder, int check_crc, uint8_t* slots, int nslots, uint64_t slot_stride) {
  if (npaths <= 0 || batch <= 0 || nslots < 2 || !slots || slot_stride < (((uint64_t)(batch + 1) * 4 + 255) & ~255ull) + 256 ||
      shard_i < 0 || shard_i >= (shard_n < 1 ? 1 : shard_n)) {
    set_err("loader_create_raw: bad arguments");
    return nullptr;
  }
  std::vector<std::string> p(paths, paths + npaths);
  return new Loader(std::move(p), 0, batch, shard_n, shard_i, epochs, 0, drop_remainder != 0, check_crc != 0, slots,
                    nslots, slot_stride, true);
}
void* dfmio_loader_create(const char** paths, int npaths, int F, int batch, int shard_n, int shard_i, int epochs,
                          int threads, int drop_remainder, int check_crc, uint8_t* slots, int nslots,
                          uint64_t slot_stride) {
  if (npaths <= 0 || F <= 0 || batch <= 0 || nslots < 2 || !slots || slot_stride < Loader::slot_bytes(batch, F) ||
      shard_i < 0 || shard_i >= (shard_n < 1 ? 1 : shard_n)) {
    set_err("loader_create: bad arguments");
    return nullptr;
  }
  std::vector<std::string> p(paths, paths + npaths);
  return new Loader(std::move(p), F, batch, shard_n, shard_i, epochs, threads, drop_remainder != 0, check_crc != 0,
                    slots, nslots, slot_stride);
}
// Same as dfmio_loader_create / dfmio_loader_create_raw (raw != 0: F and threads are ignored as there), plus a per-epoch
// record shuffle when shuffle_seed != 0. The shuffle needs random access, i.e. regular files (indexed mode): FIFO / pipe
// sources keep their stream order.
void* dfmio_loader_create_ex(const char** paths, int npaths, int F, int batch, int shard_n, int shard_i, int epochs,
                             int threads, int drop_remainder, int check_crc, uint8_t* slots, int nslots,
                             uint64_t slot_stride, int raw, uint64_t shuffle_seed) {
  const uint64_t need = raw ? (((uint64_t)(batch + 1) * 4 + 255) & ~255ull) + 256 : Loader::slot_bytes(batch, F);
  if (npaths <= 0 || (!raw && F <= 0) || batch <= 0 || nslots < 2 || !slots || slot_stride < need || shard_i < 0 ||
      shard_i >= (shard_n < 1 ? 1 : shard_n)) {
    set_err("loader_create_ex: bad arguments");
    return nullptr;
  }
  std::vector<std::string> p(paths, paths + npaths);
  return new Loader(std::move(p), raw ? 0 : F, batch, shard_n, shard_i, epochs, raw ? 0 : threads, drop_remainder != 0,
                    check_crc != 0, slots, nslots, slot_stride, raw != 0, shuffle_seed);
}
int dfmio_loader_next(void* ld, int* slot) { return ((Loader*)ld)->next(slot); }
void dfmio_loader_release(void* ld, int slot) { ((Loader*)ld)->release(slot); }
long dfmio_loader_records(void* ld) { return ((Loader*)ld)->records_read(); }
void dfmio_loader_destroy(void* ld) { delete (Loader*)ld; }

}  // extern "C"
