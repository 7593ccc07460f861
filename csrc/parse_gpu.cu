// example_parse_gpu (replaces the CPU ParseExample op, N2, when the host parser is the
// bottleneck): a batch of serialized `tf.train.Example` protos is copied to the GPU as raw bytes and decoded there,
// one thread per record (records staged in shared memory per CTA), straight into the packed input slot (labels | ids | vals) the training step reads.
// Fixed schema (TOOL:25-32, HVD:78-83): "label" float x1, "ids" int64 xF, "values" float xF; packed and unpacked
// repeated encodings, any map order, unknown features skipped. Errors (missing feature, wrong count, id > int32,
// malformed varint) set *err to 1 + record index of the first failure and leave the record zero-filled.
#include "common.cuh"

namespace dfm {

struct Cur {
  const uint8_t* p;
  const uint8_t* e;
};

__device__ __forceinline__ bool rd_varint(Cur& c, uint64_t& v) {
  uint64_t r = 0;
  for (int s = 0; c.p < c.e && s <= 63; s += 7) {
    const uint8_t b = *c.p++;
    r |= (uint64_t)(b & 0x7F) << s;
    if (!(b & 0x80)) { v = r; return true; }
  }
  return false;
}
__device__ __forceinline__ bool skip_field(Cur& c, uint32_t wt) {
  uint64_t n;
  switch (wt) {
    case 0: return rd_varint(c, n);
    case 1: c.p += 8; return c.p <= c.e;
    case 2: if (!rd_varint(c, n)) return false; if (n > (uint64_t)(c.e - c.p)) return false; c.p += n; return true;
    case 5: c.p += 4; return c.p <= c.e;
    default: return false;
  }
}
__device__ __forceinline__ bool rd_ld(Cur& c, Cur& sub) {
  uint64_t n;
  if (!rd_varint(c, n) || n > (uint64_t)(c.e - c.p)) return false;
  sub.p = c.p; sub.e = c.p + n; c.p += n;
  return true;
}
__device__ __forceinline__ float ld_f32_unaligned(const uint8_t* p) {
  uint32_t u = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
  return __uint_as_float(u);
}
__device__ int parse_floats(Cur c, float* out, int cap) {
  int cnt = 0;
  while (c.p < c.e) {
    uint64_t tag;
    if (!rd_varint(c, tag)) return -1;
    if ((tag >> 3) == 1 && (tag & 7) == 2) {
      Cur s;
      if (!rd_ld(c, s)) return -1;
      const int n = (int)((s.e - s.p) >> 2);
      if (cnt + n > cap) return -2;
      for (int i = 0; i < n; ++i) out[cnt + i] = ld_f32_unaligned(s.p + 4 * i);
      cnt += n;
    } else if ((tag >> 3) == 1 && (tag & 7) == 5) {
      if (c.p + 4 > c.e) return -1;
      if (cnt + 1 > cap) return -2;
      out[cnt++] = ld_f32_unaligned(c.p);
      c.p += 4;
    } else if (!skip_field(c, (uint32_t)(tag & 7))) {
      return -1;
    }
  }
  return cnt;
}
__device__ int parse_ints(Cur c, int* out, int cap) {
  int cnt = 0;
  while (c.p < c.e) {
    uint64_t tag;
    if (!rd_varint(c, tag)) return -1;
    if ((tag >> 3) == 1 && (tag & 7) == 2) {
      Cur s;
      if (!rd_ld(c, s)) return -1;
      while (s.p < s.e) {
        uint64_t v;
        if (!rd_varint(s, v)) return -1;
        if (cnt + 1 > cap) return -2;
        if (v > 0x7FFFFFFFull) return -3;
        out[cnt++] = (int)v;
      }
    } else if ((tag >> 3) == 1 && (tag & 7) == 0) {
      uint64_t v;
      if (!rd_varint(c, v)) return -1;
      if (cnt + 1 > cap) return -2;
      if (v > 0x7FFFFFFFull) return -3;
      out[cnt++] = (int)v;
    } else if (!skip_field(c, (uint32_t)(tag & 7))) {
      return -1;
    }
  }
  return cnt;
}

__device__ bool key_is(const uint8_t* k, size_t n, const char* lit, size_t m) {
  if (n != m) return false;
  for (size_t i = 0; i < m; ++i)
    if (k[i] != (uint8_t)lit[i]) return false;
  return true;
}

__device__ bool parse_example_dev(const uint8_t* p, int len, int F, float* label, int* ids, float* vals) {
  Cur c{p, p + len};
  bool got_l = false, got_i = false, got_v = false;
  while (c.p < c.e) {
    uint64_t tag;
    if (!rd_varint(c, tag)) return false;
    if (tag != ((1u << 3) | 2)) {
      if (!skip_field(c, (uint32_t)(tag & 7))) return false;
      continue;
    }
    Cur feats;
    if (!rd_ld(c, feats)) return false;
    while (feats.p < feats.e) {
      uint64_t t2;
      if (!rd_varint(feats, t2)) return false;
      if (t2 != ((1u << 3) | 2)) {
        if (!skip_field(feats, (uint32_t)(t2 & 7))) return false;
        continue;
      }
      Cur ent;
      if (!rd_ld(feats, ent)) return false;
      const uint8_t* key = nullptr;
      size_t klen = 0;
      Cur val{nullptr, nullptr};
      while (ent.p < ent.e) {
        uint64_t t3;
        if (!rd_varint(ent, t3)) return false;
        if ((t3 & 7) != 2) {
          if (!skip_field(ent, (uint32_t)(t3 & 7))) return false;
          continue;
        }
        Cur b;
        if (!rd_ld(ent, b)) return false;
        if ((t3 >> 3) == 1) { key = b.p; klen = (size_t)(b.e - b.p); }
        else if ((t3 >> 3) == 2) val = b;
      }
      if (!key || !val.p) continue;
      int which = -1;
      if (key_is(key, klen, "label", 5)) which = 0;
      else if (key_is(key, klen, "ids", 3)) which = 1;
      else if (key_is(key, klen, "values", 6)) which = 2;
      if (which < 0) continue;
      while (val.p < val.e) {
        uint64_t t4;
        if (!rd_varint(val, t4)) return false;
        if ((t4 & 7) != 2) {
          if (!skip_field(val, (uint32_t)(t4 & 7))) return false;
          continue;
        }
        Cur b;
        if (!rd_ld(val, b)) return false;
        const uint32_t fld = (uint32_t)(t4 >> 3);
        if (which == 0 && fld == 2) {
          float tmp[1];
          if (parse_floats(b, tmp, 1) != 1) return false;
          *label = tmp[0];
          got_l = true;
        } else if (which == 1 && fld == 3) {
          if (parse_ints(b, ids, F) != F) return false;
          got_i = true;
        } else if (which == 2 && fld == 2) {
          if (parse_floats(b, vals, F) != F) return false;
          got_v = true;
        }
      }
    }
  }
  return got_l && got_i && got_v;
}

constexpr int kParseThreads = 64;         // records per CTA
constexpr int kParseStage = 40 * 1024;    // staged bytes per CTA (64 records x ~400 B at F = 39; longer spans parse from global)

// Parsing is a chain of dependent byte loads (every varint byte decides where the next one is), ~0.5 us each from
// global memory. The records of one CTA are contiguous in the blob, so the CTA first copies their span into shared
// memory with coalesced 16-byte loads (bytewise head/tail: nothing outside [begin, end) is ever read) and the threads
// then walk their records there.
__global__ void __launch_bounds__(kParseThreads)
example_parse_kernel(const uint8_t* __restrict__ blob, const int* __restrict__ offsets, int n, int F,
                     float* __restrict__ labels, int* __restrict__ ids, float* __restrict__ vals, int* __restrict__ err) {
  __shared__ __align__(16) uint8_t stage[kParseStage];
  const int r0 = blockIdx.x * kParseThreads;
  const int r1 = min(n, r0 + kParseThreads);
  const int o_begin = offsets[r0], o_end = offsets[r1];
  const uint8_t* src = blob + o_begin;
  const int span = o_end - o_begin;
  const int mis = (int)(reinterpret_cast<uintptr_t>(src) & 15);  // stage[mis + k] holds src[k]: 16-byte chunks stay aligned
  const bool staged = span >= 0 && span + mis <= kParseStage;    // CTA-uniform
  if (staged) {
    const int head = min(span, (16 - mis) & 15);           // bytes before the first 16-byte boundary
    const int nvec = (span - head) >> 4;                   // whole 16-byte chunks
    const int tail0 = head + (nvec << 4);
    for (int k = threadIdx.x; k < head; k += kParseThreads) stage[mis + k] = src[k];
    const uint4* vsrc = reinterpret_cast<const uint4*>(src + head);
    uint4* vdst = reinterpret_cast<uint4*>(stage + mis + head);
    for (int k = threadIdx.x; k < nvec; k += kParseThreads) vdst[k] = __ldg(vsrc + k);
    for (int k = tail0 + threadIdx.x; k < span; k += kParseThreads) stage[mis + k] = src[k];
  }
  __syncthreads();
  const int r = r0 + threadIdx.x;
  if (r >= n) return;
  const int o = offsets[r], len = offsets[r + 1] - o;
  const uint8_t* rec = staged ? stage + mis + (o - o_begin) : blob + o;
  float* v = vals + (size_t)r * F;
  int* d = ids + (size_t)r * F;
  // offsets are produced by the host loader; a non-monotonic table must not steer the staged pointer outside the span
  const bool inside = len >= 0 && o >= o_begin && o + len <= o_end;
  if (!inside || !parse_example_dev(rec, len, F, labels + r, d, v)) {
    atomicCAS(err, 0, r + 1);
    labels[r] = 0.f;
    for (int f = 0; f < F; ++f) { d[f] = 0; v[f] = 0.f; }
  }
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// blob: serialized Examples back to back (device); offsets[n+1]: byte offsets (device); outputs: labels [n] f32,
// ids [n,F] int32, vals [n,F] f32 (device). err: device int, 0 on entry; 1 + index of the first bad record afterwards.
int dfm_example_parse(const void* blob, const int* offsets, int n, int F, float* labels, int* ids, float* vals, int* err,
                      void* stream) {
  DFM_REQUIRE(n > 0 && F > 0, "example_parse: empty batch");
  example_parse_kernel<<<ceil_div(n, kParseThreads), kParseThreads, 0, (cudaStream_t)stream>>>((const uint8_t*)blob, offsets, n, F, labels,
                                                                            ids, vals, err);
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
