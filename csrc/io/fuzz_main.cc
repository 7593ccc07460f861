// Mutation fuzzer for the host IO library, meant to be built with AddressSanitizer + UndefinedBehaviorSanitizer
// (scripts/fuzz_io.sh). The GPU pool this project is developed on does not offer compute-sanitizer, but the host-side
// parsers are the code that eats untrusted bytes (training files, checkpoints), and they can be checked here: every
// entry point is fed valid inputs with random corruptions and must return an error code or a result - never crash, read
// out of bounds, or loop forever.
//
//   fuzz_io <iterations> [seed] [tmpdir]
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <random>
#include <string>
#include <vector>

#include "deepfm_io.h"

namespace {

std::mt19937_64 rng;

size_t rnd(size_t n) { return n ? (size_t)(rng() % n) : 0; }

void write_file(const std::string& path, const std::vector<uint8_t>& b) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(2); }
  if (!b.empty()) fwrite(b.data(), 1, b.size(), f);
  fclose(f);
}
std::vector<uint8_t> read_file(const std::string& path) {
  std::vector<uint8_t> b;
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return b;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  b.resize((size_t)n);
  if (n && fread(b.data(), 1, (size_t)n, f) != (size_t)n) b.clear();
  fclose(f);
  return b;
}

// flip bytes, overwrite runs with 0xFF / 0x00 / 0x80 (varint continuation), truncate or extend
std::vector<uint8_t> mutate(std::vector<uint8_t> b) {
  const int kinds = 1 + (int)rnd(3);
  for (int k = 0; k < kinds && !b.empty(); ++k) {
    switch (rnd(6)) {
      case 0: b[rnd(b.size())] ^= (uint8_t)(1u << rnd(8)); break;
      case 1: b[rnd(b.size())] = (uint8_t)rng(); break;
      case 2: { size_t p = rnd(b.size()), n = 1 + rnd(12); for (size_t i = p; i < b.size() && i < p + n; ++i) b[i] = 0xFF; break; }
      case 3: { size_t p = rnd(b.size()), n = 1 + rnd(12); for (size_t i = p; i < b.size() && i < p + n; ++i) b[i] = 0x80; break; }
      case 4: b.resize(rnd(b.size() + 1)); break;
      default: { size_t n = 1 + rnd(16); for (size_t i = 0; i < n; ++i) b.push_back((uint8_t)rng()); break; }
    }
  }
  return b;
}

}  // namespace

int main(int argc, char** argv) {
  const long iters = argc > 1 ? atol(argv[1]) : 200;
  rng.seed(argc > 2 ? strtoull(argv[2], nullptr, 10) : 1);
  const std::string dir = argc > 3 ? argv[3] : "/tmp";
  const std::string rec_path = dir + "/fuzz.tfrecords", txt_path = dir + "/fuzz.libsvm", out_path = dir + "/fuzz.out",
                    bprefix = dir + "/fuzz.ckpt";
  const int F = 5, n = 24;
  // ---- a valid batch of Examples, framed
  std::vector<float> labels(n), vals((size_t)n * F);
  std::vector<int64_t> ids((size_t)n * F);
  for (int i = 0; i < n; ++i) {
    labels[(size_t)i] = (float)(i % 2);
    for (int f = 0; f < F; ++f) {
      ids[(size_t)i * F + f] = (int64_t)(rng() % 1000000);
      vals[(size_t)i * F + f] = (float)(rng() % 1000) / 1000.f;
    }
  }
  const uint64_t need = dfmio_encode_ctr_batch(labels.data(), ids.data(), vals.data(), n, F, nullptr, 0);
  std::vector<uint8_t> framed((size_t)need);
  dfmio_encode_ctr_batch(labels.data(), ids.data(), vals.data(), n, F, framed.data(), need);
  // payload-only copy for the batch parser
  std::vector<uint8_t> payload;
  std::vector<uint64_t> off, len;
  for (size_t p = 0; p + 16 <= framed.size();) {
    uint64_t l;
    memcpy(&l, framed.data() + p, 8);
    off.push_back(payload.size());
    len.push_back(l);
    payload.insert(payload.end(), framed.begin() + (long)p + 12, framed.begin() + (long)(p + 12 + l));
    p += 16 + l;
  }
  // ---- a valid (sharded) tensor bundle
  {
    void* w = dfmio_bundle_writer_create_sharded(bprefix.c_str(), 2);
    const int64_t d1[1] = {7}, d2[2] = {3, 4};
    float a[12] = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12};
    int64_t step = 5;
    dfmio_bundle_writer_add(w, "fm_w", 1, 1, d1, a, 28);
    dfmio_bundle_writer_add(w, "Deep-part/mlp0/weights", 1, 2, d2, a, 48);
    dfmio_bundle_writer_add(w, "global_step", 9, 0, d1, &step, 8);
    // a partitioned variable (full-shape entry + slices under OrderedCode keys): its decoder sees the mutated index too
    dfmio_bundle_writer_add_partitioned(w, "fm_v", 1, 2, d2, a, 48, 2);
    dfmio_bundle_writer_finish(w);
  }
  const std::vector<uint8_t> index_ok = read_file(bprefix + ".index");
  long parse_ok = 0, parse_err = 0, file_ok = 0, file_err = 0, bundle_ok = 0, bundle_err = 0, svm_ok = 0, svm_err = 0;
  std::vector<float> ol((size_t)n), ov((size_t)n * F);
  std::vector<int32_t> oi((size_t)n * F);
  std::vector<uint8_t> slots;
  for (long it = 0; it < iters; ++it) {
    // (1) batch parser on corrupted payloads (lengths kept: the parser must stay inside [off, off + len))
    {
      std::vector<uint8_t> p = payload;
      for (int k = 0; k < 1 + (int)rnd(4); ++k) p[rnd(p.size())] = (uint8_t)rng();
      (dfmio_parse_batch(p.data(), off.data(), len.data(), n, F, ol.data(), oi.data(), ov.data()) == 0 ? parse_ok : parse_err)++;
    }
    // (2) record file: count, reader, loader (indexed mode with CRC checks)
    {
      write_file(rec_path, it % 4 == 0 ? framed : mutate(framed));
      dfmio_count_records(rec_path.c_str());
      if (void* r = dfmio_reader_open(rec_path.c_str())) {
        uint8_t buf[4096];
        long got, guard = 0;
        while ((got = dfmio_reader_next(r, buf, sizeof(buf), 1)) >= 0 && ++guard < 100000) {}
        dfmio_reader_close(r);
      }
      const char* paths[1] = {rec_path.c_str()};
      const uint64_t stride = dfmio_loader_slot_bytes(4, F);
      slots.assign((size_t)stride * 3, 0);
      void* ld = dfmio_loader_create_ex(paths, 1, F, 4, 1, 0, 1, 2, 0, 1, slots.data(), 3, stride, 0, it % 2 ? 9 : 0);
      if (ld) {
        int slot, rows, guard = 0;
        while ((rows = dfmio_loader_next(ld, &slot)) > 0 && ++guard < 100000) dfmio_loader_release(ld, slot);
        (rows == 0 ? file_ok : file_err)++;
        dfmio_loader_destroy(ld);
      } else {
        file_err++;
      }
    }
    // (2b) every 8th round, on the VALID file: many workers, several epochs, shuffle, parsed and raw mode, slots released
    // late - the configuration a training job uses, for the thread sanitizer's benefit - and the threaded host optimizer
    if (it % 8 == 0) {
      write_file(rec_path, framed);
      const char* paths[1] = {rec_path.c_str()};
      for (int raw = 0; raw < 2; ++raw) {
        const uint64_t stride = raw ? dfmio_loader_raw_slot_bytes(4, 512) : dfmio_loader_slot_bytes(4, F);
        slots.assign((size_t)stride * 4, 0);
        void* ld = dfmio_loader_create_ex(paths, 1, F, 4, 1, 0, 5, 4, 1, 1, slots.data(), 4, stride, raw, 3);
        if (!ld) { fprintf(stderr, "stress loader: %s\n", dfmio_last_error()); return 1; }
        int slot, rows, held[3], nheld = 0;
        long total = 0;
        while ((rows = dfmio_loader_next(ld, &slot)) > 0) {
          total += rows;
          held[nheld++] = slot;
          if (nheld == 3) { while (nheld) dfmio_loader_release(ld, held[--nheld]); }
        }
        while (nheld) dfmio_loader_release(ld, held[--nheld]);
        dfmio_loader_destroy(ld);
        if (rows != 0 || total != 5 * (n / 4) * 4) { fprintf(stderr, "stress loader: %ld rows, rc %d\n", total, rows); return 1; }
      }
      std::vector<float> tab(4096 * 8, 0.5f), m(4096 * 8, 0.f), v(4096 * 8, 0.f), g(1024 * 8, 0.01f);
      std::vector<long long> rows_idx(1024);
      for (int i = 0; i < 1024; ++i) rows_idx[(size_t)i] = 4 * i + 1;
      if (dfmio_sparse_opt_apply(tab.data(), m.data(), v.data(), 4096, 8, rows_idx.data(), 1024, g.data(), 0, 1e-3f, 0.9f,
                                 0.999f, 1e-8f, 0.95f, 1e-4f, 1 + it, 4) != 0) return 1;
    }
    // (3) checkpoint index with corruptions
    {
      write_file(bprefix + ".index", it % 4 == 0 ? index_ok : mutate(index_ok));
      dfmio_bundle_check_block_crc(it % 2);  // every other round without the block checksum: the decoders see the damage
      if (void* r = dfmio_bundle_reader_open(bprefix.c_str())) {
        const int cnt = dfmio_bundle_reader_count(r);
        for (int i = 0; i < cnt; ++i) {
          char name[256];
          int dtype, ndim;
          int64_t dims[8];
          uint64_t nbytes;
          if (dfmio_bundle_reader_entry(r, i, name, sizeof(name), &dtype, &ndim, dims, &nbytes) != 0) continue;
          if (nbytes <= (1u << 20)) {
            std::vector<uint8_t> dst((size_t)nbytes + 1);
            dfmio_bundle_reader_read(r, i, dst.data(), nbytes, 1);
          }
        }
        dfmio_bundle_reader_close(r);
        dfmio_bundle_check_block_crc(1);
        bundle_ok++;
      } else {
        dfmio_bundle_check_block_crc(1);
        bundle_err++;
      }
    }
    // (4) libsvm text with garbage
    {
      std::string text;
      for (int l = 0; l < 6; ++l) {
        text += std::to_string(l % 2);
        for (int f = 0; f < F; ++f) text += " " + std::to_string(rng() % 5000) + ":" + std::to_string((double)(rng() % 100) / 7.0);
        text += "\n";
      }
      std::vector<uint8_t> t(text.begin(), text.end());
      if (it % 3) t = mutate(t);
      write_file(txt_path, t);
      (dfmio_libsvm_convert(txt_path.c_str(), out_path.c_str(), it % 2 ? F : 0, 0, 1 + (int)rnd(3)) >= 0 ? svm_ok : svm_err)++;
    }
  }
  write_file(bprefix + ".index", index_ok);
  printf("fuzz_io: %ld iterations | parse ok/err %ld/%ld | record files ok/err %ld/%ld | bundles ok/err %ld/%ld | libsvm ok/err %ld/%ld\n",
         iters, parse_ok, parse_err, file_ok, file_err, bundle_ok, bundle_err, svm_ok, svm_err);
  unlink(rec_path.c_str()); unlink(txt_path.c_str()); unlink(out_path.c_str());
  unlink((bprefix + ".index").c_str()); unlink((bprefix + ".data-00000-of-00002").c_str()); unlink((bprefix + ".data-00001-of-00002").c_str());
  return 0;
}
