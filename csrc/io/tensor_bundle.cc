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
};

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
  s.push_back((char)((5 << 3) | 0));
  put_varint(s, e.size);
  s.push_back((char)((6 << 3) | 5));
  put_u32(s, e.crc);
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
      p += n;
    } else if (wt == 1) {
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
  std::vector<Entry> entries;
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
  if (mask_crc(dfmio::crc32c(p, h.size + 1)) != c) {
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
  return 0;
}

int dfmio_bundle_reader_read(void* rh, int i, void* dst, uint64_t cap, int verify_crc) {
  Reader* r = (Reader*)rh;
  if (i < 0 || i >= (int)r->entries.size()) return -1;
  const Entry& e = r->entries[i];
  if (e.size > cap) {
    berr("destination too small for %s", e.name.c_str());
    return -1;
  }
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

void dfmio_bundle_reader_close(void* rh) { delete (Reader*)rh; }

}  // extern "C"
