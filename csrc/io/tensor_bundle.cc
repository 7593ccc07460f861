// TensorFlow "tensor bundle" (checkpoint V2) reader/writer without TensorFlow              [reference dep N6]
//
// The reference checkpoints through tf.estimator (`model_dir`, PS:435 / HVD:366): `model.ckpt-N.index` (an
// immutable LevelDB-format table of BundleEntryProto keyed by variable name) + `model.ckpt-N.data-00000-of-00001`
// (raw little-endian tensor bytes). Format notes: SURVEY Appendix A.
//
//   table  = data blocks | metaindex block | index block | footer(48 B)
//   block  = entries(prefix-compressed: varint shared, non_shared, value_len, key delta, value)
//            | restart offsets (u32 each) | num_restarts (u32)        followed on disk by a 5-byte trailer:
//            compression type (0 = none) | masked crc32c(block bytes + type byte)
//   footer = metaindex handle | index handle (varint64 offset, size) | zero padding to 40 B | magic 0xdb4775248b80fb57
//   key "" -> BundleHeaderProto{num_shards=1, endianness=LITTLE, version{producer=1}}
//   key v  -> BundleEntryProto{dtype=1, shape=2, shard_id=3, offset=4, size=5, crc32c=6(fixed32, masked)}
//
// Writer: single shard, no compression, restart interval 16, 4 KiB data blocks.
// Reader: uncompressed tables only (what TF's BundleWriter emits); snappy blocks fail loudly.
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

namespace dfmio {
uint32_t crc32c(const uint8_t* p, size_t n, uint32_t crc = 0);
}

namespace {

thread_local char g_berr[512] = "";
void berr(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_berr, sizeof(g_berr), fmt, ap);
  va_end(ap);
}
inline uint32_t mask_crc(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xA282EAD8u; }
constexpr uint64_t kMagic = 0xdb4775248b80fb57ull;

void put_varint(std::string& s, uint64_t v) {
  while (v >= 0x80) {
    s.push_back((char)(v | 0x80));
    v >>= 7;
  }
  s.push_back((char)v);
}
bool get_varint(const uint8_t*& p, const uint8_t* e, uint64_t* v) {
  uint64_t r = 0;
  for (int s = 0; p < e && s <= 63; s += 7) {
    uint8_t b = *p++;
    r |= (uint64_t)(b & 0x7F) << s;
    if (!(b & 0x80)) {
      *v = r;
      return true;
    }
  }
  return false;
}
void put_u32(std::string& s, uint32_t v) { s.append((const char*)&v, 4); }

struct Entry {
  std::string name;
  int dtype = 1;  // DT_FLOAT=1, DT_INT32=3, DT_INT64=9
  std::vector<int64_t> dims;
  uint64_t offset = 0, size = 0;
  uint32_t crc = 0;  // masked
  int shard = 0;     // which <prefix>.data-SSSSS-of-NNNNN file holds the bytes
  // Partitioned variable (tf.fixed_size_partitioner and friends): the full-shape entry carries no bytes, only the list
  // of slices; each slice's bytes live under their own key (slice_key below). Per dimension (start, length), length -1 =
  // the whole dimension.
  std::vector<std::vector<std::pair<int64_t, int64_t>>> slices;
};

// ---- OrderedCode (the key encoding of slice entries: tensorflow/core/lib/strings/ordered_code.h semantics) -----
void oc_write_num_increasing(std::string& s, uint64_t v) {
  uint8_t buf[9];
  int len = 0;
  while (v) {
    buf[8 - len] = (uint8_t)(v & 0xff);
    v >>= 8;
    ++len;
  }
  buf[8 - len] = (uint8_t)len;
  s.append((const char*)buf + 8 - len, len + 1);
}
void oc_write_string(std::string& s, const std::string& v) {
  for (unsigned char c : v) {
    if (c == 0x00) s.append("\x00\xff", 2);
    else if (c == 0xff) s.append("\xff\x00", 2);
    else s.push_back((char)c);
  }
  s.append("\x00\x01", 2);
}
int oc_signed_length(uint64_t x) {  // bytes needed for magnitude x: 7 n - 1 payload bits in n bytes
  int bits = 0;
  while (x) {
    ++bits;
    x >>= 1;
  }
  for (int n = 1; n <= 10; ++n)
    if (bits <= 7 * n - 1) return n;
  return 10;
}
void oc_write_signed_increasing(std::string& s, int64_t val) {
  const uint64_t x = val < 0 ? ~(uint64_t)val : (uint64_t)val;
  const int len = oc_signed_length(x);
  uint8_t buf[10];
  const uint8_t sign = val < 0 ? 0xff : 0x00;
  buf[0] = buf[1] = sign;
  for (int i = 0; i < 8; ++i) buf[2 + i] = (uint8_t)((uint64_t)val >> (8 * (7 - i)));
  uint8_t* b = buf + 10 - len;
  // header: `len` leading one bits (spilling into the second byte for len > 8), xor-ed so negatives sort first
  for (int i = 0; i < len && i < 16; ++i) b[i / 8] ^= (uint8_t)(0x80 >> (i % 8));
  s.append((const char*)b, len);
}
std::string slice_key(const std::string& name, const std::vector<std::pair<int64_t, int64_t>>& sl) {
  std::string k;
  oc_write_num_increasing(k, 0);
  oc_write_string(k, name);
  oc_write_num_increasing(k, sl.size());
  for (auto& d : sl) {
    oc_write_signed_increasing(k, d.second < 0 ? -1 : d.first);
    oc_write_signed_increasing(k, d.second < 0 ? -1 : d.second);
  }
  return k;
}
int dtype_size(int dt) { return dt == 9 || dt == 2 ? 8 : dt == 1 || dt == 3 ? 4 : dt == 14 || dt == 19 ? 2 : dt == 6 || dt == 4 || dt == 10 ? 1 : 0; }

std::string encode_entry(const Entry& e) {
  std::string shape;
  for (int64_t d : e.dims) {
    std::string dim;
    dim.push_back((char)((1 << 3) | 0));
    put_varint(dim, (uint64_t)d);
    shape.push_back((char)((2 << 3) | 2));
    put_varint(shape, dim.size());
    shape += dim;
  }
  std::string s;
  s.push_back((char)((1 << 3) | 0));
  put_varint(s, (uint64_t)e.dtype);
  s.push_back((char)((2 << 3) | 2));
  put_varint(s, shape.size());
  s += shape;
  if (e.shard) {  // shard_id (3); 0 is omitted (proto3 default)
    s.push_back((char)((3 << 3) | 0));
    put_varint(s, (uint64_t)e.shard);
  }
  if (e.offset) {
    s.push_back((char)((4 << 3) | 0));
    put_varint(s, e.offset);
  }
  if (e.slices.empty()) {
    s.push_back((char)((5 << 3) | 0));
    put_varint(s, e.size);
    s.push_back((char)((6 << 3) | 5));
    put_u32(s, e.crc);
  }
  for (auto& sl : e.slices) {  // slices (7): TensorSliceProto { repeated Extent extent = 1 { start = 1; length = 2 } }
    std::string tsp;
    for (auto& d : sl) {
      std::string ext;
      if (d.second >= 0) {
        if (d.first) {
          ext.push_back((char)((1 << 3) | 0));
          put_varint(ext, (uint64_t)d.first);
        }
        ext.push_back((char)((2 << 3) | 0));
        put_varint(ext, (uint64_t)d.second);
      }
      tsp.push_back((char)((1 << 3) | 2));
      put_varint(tsp, ext.size());
      tsp += ext;
    }
    s.push_back((char)((7 << 3) | 2));
    put_varint(s, tsp.size());
    s += tsp;
  }
  return s;
}

std::string encode_header(int num_shards) {
  std::string s;
  s.push_back((char)((1 << 3) | 0));
  put_varint(s, (uint64_t)num_shards);
  // endianness (2) = LITTLE = 0 omitted
  std::string ver;
  ver.push_back((char)((1 << 3) | 0));
  put_varint(ver, 1);  // producer
  s.push_back((char)((3 << 3) | 2));
  put_varint(s, ver.size());
  s += ver;
  return s;
}

bool decode_entry(const uint8_t* p, const uint8_t* e, Entry* out) {
  while (p < e) {
    uint64_t tag;
    if (!get_varint(p, e, &tag)) return false;
    const uint32_t f = (uint32_t)(tag >> 3), wt = (uint32_t)(tag & 7);
    if (wt == 0) {
      uint64_t v;
      if (!get_varint(p, e, &v)) return false;
      if (f == 1) out->dtype = (int)v;
      else if (f == 3) out->shard = (int)v;
      else if (f == 4) out->offset = v;
      else if (f == 5) out->size = v;
    } else if (wt == 5) {
      if (p + 4 > e) return false;
      if (f == 6) memcpy(&out->crc, p, 4);
      p += 4;
    } else if (wt == 2) {
      uint64_t n;
      if (!get_varint(p, e, &n) || n > (uint64_t)(e - p)) return false;
      if (f == 2) {  // TensorShapeProto
        const uint8_t *q = p, *qe = p + n;
        while (q < qe) {
          uint64_t t2;
          if (!get_varint(q, qe, &t2)) return false;
          if ((t2 >> 3) == 2 && (t2 & 7) == 2) {
            uint64_t m;
            if (!get_varint(q, qe, &m) || m > (uint64_t)(qe - q)) return false;
            const uint8_t *d = q, *de = q + m;
            int64_t size = 0;
            while (d < de) {
              uint64_t t3, v;
              if (!get_varint(d, de, &t3)) return false;
              if ((t3 & 7) == 0) {
                if (!get_varint(d, de, &v)) return false;
                if ((t3 >> 3) == 1) size = (int64_t)v;
              } else if ((t3 & 7) == 2) {
                if (!get_varint(d, de, &v)) return false;
                d += v;
              } else {
                return false;
              }
            }
            out->dims.push_back(size);
            q += m;
          } else if ((t2 & 7) == 0) {
            uint64_t v;
            if (!get_varint(q, qe, &v)) return false;
          } else {
            return false;
          }
        }
      }
      if (f == 7) {  // one TensorSliceProto
        std::vector<std::pair<int64_t, int64_t>> sl;
        const uint8_t *q = p, *qe = p + n;
        while (q < qe) {
          uint64_t t2, m;
          if (!get_varint(q, qe, &t2)) return false;
          if ((t2 & 7) == 2) {
            if (!get_varint(q, qe, &m) || m > (uint64_t)(qe - q)) return false;
            if ((t2 >> 3) == 1) {  // Extent
              int64_t start = 0, length = -1;
              const uint8_t *d = q, *de = q + m;
              while (d < de) {
                uint64_t t3, v;
                if (!get_varint(d, de, &t3) || (t3 & 7) != 0 || !get_varint(d, de, &v)) return false;
                if ((t3 >> 3) == 1) start = (int64_t)v;
                else if ((t3 >> 3) == 2) length = (int64_t)v;
              }
              sl.emplace_back(start, length);
            }
            q += m;
          } else if ((t2 & 7) == 0) {
            if (!get_varint(q, qe, &m)) return false;
          } else {
            return false;
          }
        }
        out->slices.push_back(std::move(sl));
      }
      p += n;
    } else if (wt == 1) {
      if (p + 8 > e) return false;
      p += 8;
    } else {
      return false;
    }
  }
  return true;
}

// ---- table building ------------------------------------------------------------------------------------
struct BlockBuilder {
  std::string buf;
  std::vector<uint32_t> restarts{0};
  std::string last_key;
  int counter = 0;
  void add(const std::string& key, const std::string& val) {
    size_t shared = 0;
    if (counter < 16) {
      const size_t m = std::min(last_key.size(), key.size());
      while (shared < m && last_key[shared] == key[shared]) ++shared;
    } else {
      restarts.push_back((uint32_t)buf.size());
      counter = 0;
    }
    put_varint(buf, shared);
    put_varint(buf, key.size() - shared);
    put_varint(buf, val.size());
    buf.append(key, shared, std::string::npos);
    buf += val;
    last_key = key;
    ++counter;
  }
  std::string finish() {
    std::string out = buf;
    for (uint32_t r : restarts) put_u32(out, r);
    put_u32(out, (uint32_t)restarts.size());
    return out;
  }
  bool empty() const { return buf.empty(); }
};

struct Handle {
  uint64_t off, size;
};
Handle write_block(FILE* f, const std::string& block, uint64_t* pos) {
  Handle h{*pos, block.size()};
  std::string with_type = block;
  with_type.push_back((char)0);  // kNoCompression
  const uint32_t c = mask_crc(dfmio::crc32c((const uint8_t*)with_type.data(), with_type.size()));
  fwrite(with_type.data(), 1, with_type.size(), f);
  fwrite(&c, 1, 4, f);
  *pos += with_type.size() + 4;
  return h;
}

std::string shard_path(const std::string& prefix, int shard, int num_shards) {
  char suffix[64];
  snprintf(suffix, sizeof(suffix), ".data-%05d-of-%05d", shard, num_shards);
  return prefix + suffix;
}

struct Writer {
  std::string prefix;
  std::vector<FILE*> data;         // one file per shard; variables are dealt round-robin, as TF's
  std::vector<uint64_t> data_pos;  // replica_device_setter deals them to parameter-server tasks
  std::vector<Entry> entries;
};

struct Reader {
  std::string prefix;
  int num_shards = 1;
  std::vector<Entry> entries;                // named tensors (whole or partitioned)
  std::map<std::string, Entry> slice_data;   // encoded slice key -> where that slice's bytes are
};

bool read_file(const std::string& path, std::vector<uint8_t>* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) {
    berr("cannot open %s", path.c_str());
    return false;
  }
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out->resize((size_t)n);
  const size_t got = fread(out->data(), 1, (size_t)n, f);
  fclose(f);
  if (got != (size_t)n) {
    berr("short read on %s", path.c_str());
    return false;
  }
  return true;
}

bool g_check_block_crc = true;  // dfmio_bundle_check_block_crc(0): the fuzzer reaches the decoders behind the CRC

bool check_block(const std::vector<uint8_t>& file, Handle h, const uint8_t** b, const uint8_t** e) {
  if (h.off + h.size + 5 > file.size()) {
    berr("block handle out of range");
    return false;
  }
  const uint8_t* p = file.data() + h.off;
  if (p[h.size] != 0) {
    berr("compressed table block (type %d) is not supported: expected an uncompressed tensor-bundle index", (int)p[h.size]);
    return false;
  }
  uint32_t c;
  memcpy(&c, p + h.size + 1, 4);
  if (g_check_block_crc && mask_crc(dfmio::crc32c(p, h.size + 1)) != c) {
    berr("table block CRC mismatch");
    return false;
  }
  *b = p;
  *e = p + h.size;
  return true;
}

template <class Fn>
bool iterate_block(const uint8_t* b, const uint8_t* e, Fn fn) {
  if (e - b < 4) return false;
  uint32_t nrest;
  memcpy(&nrest, e - 4, 4);
  const uint8_t* lim = e - 4 - (size_t)nrest * 4;
  if (lim < b) return false;
  std::string key;
  const uint8_t* p = b;
  while (p < lim) {
    uint64_t shared, non_shared, vlen;
    if (!get_varint(p, lim, &shared) || !get_varint(p, lim, &non_shared) || !get_varint(p, lim, &vlen)) return false;
    if (shared > key.size() || non_shared + vlen > (uint64_t)(lim - p)) return false;
    key.resize(shared);
    key.append((const char*)p, non_shared);
    p += non_shared;
    if (!fn(key, p, p + vlen)) return false;
    p += vlen;
  }
  return true;
}

}  // namespace

extern "C" {

const char* dfmio_bundle_last_error() { return g_berr; }

// The index key under which one slice of a partitioned variable is stored (TF: checkpoint::EncodeTensorNameSlice, OrderedCode).
// starts[d] / lens[d] per dimension, lens[d] < 0 = the whole dimension. Writes the key bytes into out (cap bytes) and returns
// its length, or -1 if it does not fit. Exposed so the encoding can be tested against an independent derivation.
int dfmio_bundle_slice_key(const char* name, int ndim, const int64_t* starts, const int64_t* lens, uint8_t* out, int cap) {
  if (!name || ndim < 0 || (ndim > 0 && (!starts || !lens)) || !out) return -1;
  std::vector<std::pair<int64_t, int64_t>> sl;
  for (int d = 0; d < ndim; ++d) sl.emplace_back(starts[d], lens[d]);
  const std::string k = slice_key(name, sl);
  if ((int)k.size() > cap) return -1;
  memcpy(out, k.data(), k.size());
  return (int)k.size();
}

// Table-block CRC verification of the index reader, on by default. Switching it off is for fuzzing only: a mutated index
// then gets past the checksum and exercises the key / BundleEntryProto / slice decoders (csrc/io/fuzz_main.cc).
void dfmio_bundle_check_block_crc(int on) { g_check_block_crc = on != 0; }

// num_shards data files (an Estimator's Saver is sharded: one file per parameter-server task in the reference's PS mode,
// one file otherwise); variable k goes to shard k % num_shards.
void* dfmio_bundle_writer_create_sharded(const char* prefix, int num_shards) {
  if (num_shards < 1 || num_shards > 99999) {
    berr("bundle writer: num_shards out of range");
    return nullptr;
  }
  Writer* w = new Writer();
  w->prefix = prefix;
  for (int sidx = 0; sidx < num_shards; ++sidx) {
    const std::string dp = shard_path(w->prefix, sidx, num_shards);
    FILE* f = fopen(dp.c_str(), "wb");
    if (!f) {
      berr("cannot open %s for writing", dp.c_str());
      for (FILE* g : w->data) fclose(g);
      delete w;
      return nullptr;
    }
    w->data.push_back(f);
    w->data_pos.push_back(0);
  }
  return w;
}

void* dfmio_bundle_writer_create(const char* prefix) { return dfmio_bundle_writer_create_sharded(prefix, 1); }

// dtype: TF DataType enum (1 float32, 3 int32, 9 int64). Bytes are host memory, little-endian.
int dfmio_bundle_writer_add(void* wh, const char* name, int dtype, int ndim, const int64_t* dims, const void* bytes,
                            uint64_t nbytes) {
  Writer* w = (Writer*)wh;
  Entry e;
  e.name = name;
  e.dtype = dtype;
  e.dims.assign(dims, dims + ndim);
  e.shard = (int)(w->entries.size() % w->data.size());
  e.offset = w->data_pos[(size_t)e.shard];
  e.size = nbytes;
  e.crc = mask_crc(dfmio::crc32c((const uint8_t*)bytes, nbytes));
  if (nbytes && fwrite(bytes, 1, nbytes, w->data[(size_t)e.shard]) != nbytes) {
    berr("write failed for %s", name);
    return -1;
  }
  w->data_pos[(size_t)e.shard] += nbytes;
  w->entries.push_back(std::move(e));
  return 0;
}

// Writes `name` as a variable partitioned along axis 0 into `num_parts` near-equal pieces (tf.fixed_size_partitioner):
// one full-shape index entry listing the slices + one data entry per slice under its encoded slice key. The pieces are
// spread over the writer's data shards like whole tensors are.
int dfmio_bundle_writer_add_partitioned(void* wh, const char* name, int dtype, int ndim, const int64_t* dims,
                                        const void* bytes, uint64_t nbytes, int num_parts) {
  Writer* w = (Writer*)wh;
  const size_t esz = (size_t)dtype_size(dtype);
  if (ndim < 1 || num_parts < 1 || esz == 0 || dims[0] < num_parts) {
    berr("cannot partition %s into %d parts", name, num_parts);
    return -1;
  }
  uint64_t inner = 1;
  for (int d = 1; d < ndim; ++d) inner *= (uint64_t)dims[d];
  if ((uint64_t)dims[0] * inner * esz != nbytes) {
    berr("size mismatch for %s", name);
    return -1;
  }
  Entry full;
  full.name = name;
  full.dtype = dtype;
  full.dims.assign(dims, dims + ndim);
  const int64_t base = dims[0] / num_parts, extra = dims[0] % num_parts;
  int64_t start = 0;
  for (int p = 0; p < num_parts; ++p) {  // the first `extra` parts get one more row, as TF's partitioner does
    const int64_t len = base + (p < extra ? 1 : 0);
    std::vector<std::pair<int64_t, int64_t>> sl;
    sl.emplace_back(start, len);
    for (int d = 1; d < ndim; ++d) sl.emplace_back(0, -1);
    Entry e;
    e.name = slice_key(name, sl);
    e.dtype = dtype;
    e.dims.push_back(len);
    for (int d = 1; d < ndim; ++d) e.dims.push_back(dims[d]);
    e.shard = (int)(w->entries.size() % w->data.size());
    e.offset = w->data_pos[(size_t)e.shard];
    e.size = (uint64_t)len * inner * esz;
    const uint8_t* src = (const uint8_t*)bytes + (uint64_t)start * inner * esz;
    e.crc = mask_crc(dfmio::crc32c(src, e.size));
    if (e.size && fwrite(src, 1, e.size, w->data[(size_t)e.shard]) != e.size) {
      berr("write failed for %s", name);
      return -1;
    }
    w->data_pos[(size_t)e.shard] += e.size;
    w->entries.push_back(std::move(e));
    full.slices.push_back(std::move(sl));
    start += len;
  }
  w->entries.push_back(std::move(full));
  return 0;
}

int dfmio_bundle_writer_finish(void* wh) {
  Writer* w = (Writer*)wh;
  for (FILE* f : w->data) fclose(f);
  std::map<std::string, std::string> kv;
  kv[""] = encode_header((int)w->data.size());
  for (const Entry& e : w->entries) kv[e.name] = encode_entry(e);
  const std::string ip = w->prefix + ".index";
  FILE* f = fopen(ip.c_str(), "wb");
  if (!f) {
    berr("cannot open %s for writing", ip.c_str());
    delete w;
    return -1;
  }
  uint64_t pos = 0;
  BlockBuilder data, index;
  std::string last_key_in_block;
  for (auto it = kv.begin(); it != kv.end(); ++it) {
    data.add(it->first, it->second);
    last_key_in_block = it->first;
    auto nx = std::next(it);
    if (data.buf.size() >= 4096 || nx == kv.end()) {
      Handle h = write_block(f, data.finish(), &pos);
      std::string hv;
      put_varint(hv, h.off);
      put_varint(hv, h.size);
      index.add(last_key_in_block, hv);  // separator = last key of the block
      data = BlockBuilder();
    }
  }
  BlockBuilder meta;
  Handle mh = write_block(f, meta.finish(), &pos);
  Handle ih = write_block(f, index.finish(), &pos);
  std::string footer;
  put_varint(footer, mh.off);
  put_varint(footer, mh.size);
  put_varint(footer, ih.off);
  put_varint(footer, ih.size);
  footer.resize(40, '\0');
  footer.append((const char*)&kMagic, 8);
  fwrite(footer.data(), 1, footer.size(), f);
  fclose(f);
  delete w;
  return 0;
}

void* dfmio_bundle_reader_open(const char* prefix) {
  std::vector<uint8_t> file;
  if (!read_file(std::string(prefix) + ".index", &file)) return nullptr;
  if (file.size() < 48) {
    berr("index file too small");
    return nullptr;
  }
  uint64_t magic;
  memcpy(&magic, file.data() + file.size() - 8, 8);
  if (magic != kMagic) {
    berr("bad table magic in %s.index", prefix);
    return nullptr;
  }
  const uint8_t* p = file.data() + file.size() - 48;
  const uint8_t* e = p + 40;
  Handle mh, ih;
  if (!get_varint(p, e, &mh.off) || !get_varint(p, e, &mh.size) || !get_varint(p, e, &ih.off) ||
      !get_varint(p, e, &ih.size)) {
    berr("bad footer");
    return nullptr;
  }
  const uint8_t *ib, *ie;
  if (!check_block(file, ih, &ib, &ie)) return nullptr;
  Reader* r = new Reader();
  r->prefix = prefix;
  bool ok = iterate_block(ib, ie, [&](const std::string&, const uint8_t* v, const uint8_t* ve) {
    Handle dh;
    if (!get_varint(v, ve, &dh.off) || !get_varint(v, ve, &dh.size)) return false;
    const uint8_t *db, *de;
    if (!check_block(file, dh, &db, &de)) return false;
    return iterate_block(db, de, [&](const std::string& key, const uint8_t* ev, const uint8_t* eve) {
      if (key.empty()) {  // BundleHeaderProto: num_shards = field 1
        const uint8_t* hp = ev;
        uint64_t tag, v;
        while (hp < eve && get_varint(hp, eve, &tag)) {
          if ((tag & 7) != 0) break;  // later fields (endianness, version) are not needed
          if (!get_varint(hp, eve, &v)) break;
          if ((tag >> 3) == 1 && v >= 1) r->num_shards = (int)v;
        }
        return true;
      }
      Entry en;
      en.name = key;
      if (!decode_entry(ev, eve, &en)) {
        berr("malformed BundleEntryProto for %s", key.c_str());
        return false;
      }
      if (key[0] == '\0')  // OrderedCode keys start with the number 0: the bytes of one slice of a partitioned variable
        r->slice_data[key] = std::move(en);
      else
        r->entries.push_back(std::move(en));
      return true;
    });
  });
  if (!ok) {
    if (!g_berr[0]) berr("malformed table in %s.index", prefix);
    delete r;
    return nullptr;
  }
  return r;
}

int dfmio_bundle_reader_count(void* rh) { return (int)((Reader*)rh)->entries.size(); }

// Fills name (cap bytes), dtype, ndim, dims[8], nbytes for entry i.
int dfmio_bundle_reader_entry(void* rh, int i, char* name, int cap, int* dtype, int* ndim, int64_t* dims,
                              uint64_t* nbytes) {
  Reader* r = (Reader*)rh;
  if (i < 0 || i >= (int)r->entries.size()) return -1;
  const Entry& e = r->entries[i];
  snprintf(name, cap, "%s", e.name.c_str());
  *dtype = e.dtype;
  *ndim = (int)e.dims.size();
  for (size_t k = 0; k < e.dims.size() && k < 8; ++k) dims[k] = e.dims[k];
  *nbytes = e.size;
  if (!e.slices.empty()) {  // partitioned: the bytes are assembled from the slices on read
    uint64_t n = (uint64_t)dtype_size(e.dtype);
    for (int64_t d : e.dims) n *= (uint64_t)(d < 0 ? 0 : d);
    *nbytes = n;
  }
  return 0;
}

namespace {
int read_span(Reader* r, const Entry& e, void* dst, int verify_crc) {
  if (e.shard < 0 || e.shard >= r->num_shards) {
    berr("%s lives in shard %d of %d", e.name.c_str(), e.shard, r->num_shards);
    return -1;
  }
  const std::string dp = shard_path(r->prefix, e.shard, r->num_shards);
  FILE* f = fopen(dp.c_str(), "rb");
  if (!f) {
    berr("cannot open %s", dp.c_str());
    return -1;
  }
  fseek(f, (long)e.offset, SEEK_SET);
  const size_t got = fread(dst, 1, e.size, f);
  fclose(f);
  if (got != e.size) {
    berr("short read for %s", e.name.c_str());
    return -1;
  }
  if (verify_crc && mask_crc(dfmio::crc32c((const uint8_t*)dst, e.size)) != e.crc) {
    berr("tensor CRC mismatch for %s", e.name.c_str());
    return -1;
  }
  return 0;
}

// Copies one slice (contiguous row-major `src` of extent ext[]) into the full row-major tensor of shape dims[].
void scatter_slice(uint8_t* dst, const uint8_t* src, const std::vector<int64_t>& dims, const std::vector<int64_t>& start,
                   const std::vector<int64_t>& ext, size_t esz) {
  const int nd = (int)dims.size();
  if (nd == 0) {
    memcpy(dst, src, esz);
    return;
  }
  std::vector<int64_t> stride(nd, 1);
  for (int d = nd - 2; d >= 0; --d) stride[d] = stride[d + 1] * dims[d + 1];
  const int64_t row = ext[nd - 1];
  int64_t outer = 1;
  for (int d = 0; d < nd - 1; ++d) outer *= ext[d];
  std::vector<int64_t> idx(nd, 0);
  for (int64_t o = 0; o < outer; ++o) {
    int64_t off = start[nd - 1];
    for (int d = 0; d < nd - 1; ++d) off += (start[d] + idx[d]) * stride[d];
    memcpy(dst + (size_t)off * esz, src + (size_t)o * row * esz, (size_t)row * esz);
    for (int d = nd - 2; d >= 0; --d) {
      if (++idx[d] < ext[d]) break;
      idx[d] = 0;
    }
  }
}
}  // namespace

int dfmio_bundle_reader_read(void* rh, int i, void* dst, uint64_t cap, int verify_crc) {
  Reader* r = (Reader*)rh;
  if (i < 0 || i >= (int)r->entries.size()) return -1;
  const Entry& e = r->entries[i];
  if (e.slices.empty()) {
    if (e.size > cap) {
      berr("destination too small for %s", e.name.c_str());
      return -1;
    }
    return read_span(r, e, dst, verify_crc);
  }
  // partitioned variable: every listed slice must be present, inside the full shape, and together cover it exactly once
  const size_t esz = (size_t)dtype_size(e.dtype);
  const int nd = (int)e.dims.size();
  uint64_t total = 1;
  for (int64_t d : e.dims) {
    if (d < 0 || (d > 0 && total > (1ull << 40) / (uint64_t)d)) {
      berr("unreasonable shape for %s", e.name.c_str());
      return -1;
    }
    total *= (uint64_t)d;
  }
  if (esz == 0 || total * esz > cap) {
    berr("destination too small (or unknown dtype) for partitioned %s", e.name.c_str());
    return -1;
  }
  uint64_t covered = 0;
  std::vector<uint8_t> tmp;
  for (auto& sl : e.slices) {
    if ((int)sl.size() != nd) {
      berr("slice rank mismatch for %s", e.name.c_str());
      return -1;
    }
    std::vector<int64_t> start(nd), ext(nd);
    uint64_t n = 1;
    for (int d = 0; d < nd; ++d) {
      start[d] = sl[d].second < 0 ? 0 : sl[d].first;
      ext[d] = sl[d].second < 0 ? e.dims[d] : sl[d].second;
      if (start[d] < 0 || ext[d] < 0 || start[d] > e.dims[d] || ext[d] > e.dims[d] - start[d]) {
        berr("slice outside the tensor for %s", e.name.c_str());
        return -1;
      }
      n *= (uint64_t)ext[d];
    }
    auto it = r->slice_data.find(slice_key(e.name, sl));
    if (it == r->slice_data.end()) {
      berr("missing slice data for %s", e.name.c_str());
      return -1;
    }
    if (it->second.size != n * esz) {
      berr("slice size mismatch for %s", e.name.c_str());
      return -1;
    }
    tmp.resize((size_t)(n * esz));
    Entry se = it->second;
    se.name = e.name;
    if (read_span(r, se, tmp.data(), verify_crc) != 0) return -1;
    scatter_slice((uint8_t*)dst, tmp.data(), e.dims, start, ext, esz);
    covered += n;
  }
  if (covered != total) {
    berr("slices of %s cover %llu of %llu elements", e.name.c_str(), (unsigned long long)covered,
         (unsigned long long)total);
    return -1;
  }
  return 0;
}

void dfmio_bundle_reader_close(void* rh) { delete (Reader*)rh; }

}  // extern "C"
