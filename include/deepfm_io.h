/* C ABI of the host IO library (data/_lib/libdeepfm_io.so, sources in csrc/io/): TFRecord framing with hardware CRC32C,
 * fixed-schema tf.train.Example batch parser, prefetching loader (streaming / indexed / raw / FIFO), TensorFlow
 * tensor-bundle (checkpoint) writer and reader. Host code only, no CUDA dependency. The reference gets all of this from
 * TensorFlow's C++ runtime; there is no counterpart source in the reference tree.
 *
 * Conventions: functions returning pointers return NULL on failure, functions returning int return 0 on success (or a
 * count where stated) and a negative value on failure; dfmio_last_error() / dfmio_bundle_last_error() hold the message.
 */
#ifndef DEEPFM_IO_H_
#define DEEPFM_IO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dfmio_last_error(void);

/* ---- CRC32C (Castagnoli), TFRecord masking: ((crc >> 15) | (crc << 17)) + 0xa282ead8 */
uint32_t dfmio_crc32c(const uint8_t* p, size_t n);
uint32_t dfmio_masked_crc32c(const uint8_t* p, size_t n);

/* ---- TFRecord writer / reader: u64 length | u32 masked crc(length) | payload | u32 masked crc(payload) */
void* dfmio_writer_open(const char* path);
int dfmio_writer_write(void* writer, const uint8_t* data, uint64_t n);
int dfmio_writer_close(void* writer);
void* dfmio_reader_open(const char* path); /* regular file (memory-mapped) or FIFO (buffered reads, never probed) */
/* next record into buf: returns its length, -1 at end of stream, -2 on corruption / IO error, -3 if cap is too small */
long dfmio_reader_next(void* reader, uint8_t* buf, size_t cap, int check_crc);
void dfmio_reader_close(void* reader);
long dfmio_count_records(const char* path); /* walks the length fields only */

/* ---- fixed-schema Example parser: label float x1, ids int64 xF (stored as int32), values float xF */
int dfmio_parse_batch(const uint8_t* bytes, const uint64_t* off, const uint64_t* len, int nrec, int F, float* labels,
                      int32_t* ids, float* vals);

/* ---- prefetching loader. The caller owns `slots` (nslots x slot_stride bytes, ideally page-locked); a parsed slot is
 * laid out labels[batch] f32 | ids[batch*F] i32 | vals[batch*F] f32 (256-byte aligned sections) so that ONE copy moves a
 * batch to the GPU; a raw slot is offsets[batch+1] i32 | serialized records (for the GPU-side parser).
 * Records are sharded as record_index % shard_n == shard_i; `epochs` <= 0 repeats forever. */
uint64_t dfmio_loader_slot_bytes(int batch, int F);
uint64_t dfmio_loader_raw_slot_bytes(int batch, int max_record_bytes);
void* dfmio_loader_create(const char** paths, int npaths, int F, int batch, int shard_n, int shard_i, int epochs,
                          int threads, int drop_remainder, int check_crc, uint8_t* slots, int nslots,
                          uint64_t slot_stride);
void* dfmio_loader_create_raw(const char** paths, int npaths, int batch, int shard_n, int shard_i, int epochs,
                              int drop_remainder, int check_crc, uint8_t* slots, int nslots, uint64_t slot_stride);
/* the two above in one call (raw != 0 = raw mode), plus a reproducible per-epoch record shuffle when shuffle_seed != 0:
 * position p of epoch e reads record perm(seed, e)(p) of this rank's shard, perm = keyed Feistel bijection (no permutation
 * array). Needs regular files; FIFO sources keep their stream order. */
void* dfmio_loader_create_ex(const char** paths, int npaths, int F, int batch, int shard_n, int shard_i, int epochs,
                             int threads, int drop_remainder, int check_crc, uint8_t* slots, int nslots,
                             uint64_t slot_stride, int raw, uint64_t shuffle_seed);
int dfmio_loader_next(void* loader, int* slot); /* rows in the slot (> 0), 0 at end of data, < 0 on error; blocks */
void dfmio_loader_release(void* loader, int slot);
long dfmio_loader_records(void* loader);
void dfmio_loader_destroy(void* loader);

/* ---- libsvm text ("label id:val id:val ...") -> TFRecord file of Examples {ids, label, values}, multi-threaded, record
 * order = line order, rows padded with (pad_id, 0) / truncated to field_size when field_size > 0. Returns the number of
 * rows, -1 on error (dfmio_libsvm_last_error()). threads <= 0: one per hardware thread. */
const char* dfmio_libsvm_last_error(void);
long long dfmio_libsvm_convert(const char* input, const char* output, int field_size, long long pad_id, int threads);
/* columns -> framed TFRecord bytes of n Examples; returns the byte count, writes to out only if it fits in cap */
unsigned long long dfmio_encode_ctr_batch(const float* labels, const int64_t* ids, const float* vals, long long n, int F,
                                          uint8_t* out, unsigned long long cap);

/* ---- host-side optimizer application for the asynchronous parameter-server store (csrc/io/host_optim.cc): TF-1 rules
 * (opt: 0 Adam, 1 Adagrad, 2 Momentum, 3 FTRL, 4 SGD), `g + l2 * w` formed in the same pass, `step` = 1-based update index,
 * work split over `threads` threads, no locking (Hogwild). s0 / s1 = slot tables (NULL where the rule has fewer slots). */
int dfmio_sparse_opt_apply(float* table, float* s0, float* s1, long long N, int K, const long long* rows, long long nrows,
                           const float* grads, int opt, float lr, float beta1, float beta2, float eps, float momentum,
                           float l2, long long step, int threads);
int dfmio_dense_opt_apply(float* w, float* s0, float* s1, const float* g, long long n, int opt, float lr, float beta1,
                          float beta2, float eps, float momentum, float l2, long long step, int threads);

/* One pushed step of a parameter-server worker in ONE call: touched rows of fm_v / fm_w (rows int32 [nrows], unique; gv
 * [nrows, K], gw [nrows]; touched-row L2) followed by nvars dense variables (pointer arrays w / s0 / s1 / g, numel; no L2). */
int dfmio_ps_apply_push(float* tab_v, float* sv0, float* sv1, float* tab_w, float* sw0, float* sw1, long long N, int K,
                        const int* rows, long long nrows, const float* gv, const float* gw, int nvars, float* const* w,
                        float* const* s0, float* const* s1, const float* const* g, const long long* numel, int opt,
                        float lr, float beta1, float beta2, float eps, float momentum, float l2, long long step,
                        int threads);

/* ---- TensorFlow tensor bundle (checkpoint V2): <prefix>.index (SSTable) + <prefix>.data-00000-of-00001.
 * dtype is TensorFlow's DataType enum (1 float32, 3 int32, 9 int64); dims holds up to 8 entries. */
const char* dfmio_bundle_last_error(void);
/* Index key of one slice of a partitioned variable (TF's OrderedCode encoding of name + per-dimension (start, length);
 * length < 0 = whole dimension); returns the key length or -1. */
int dfmio_bundle_slice_key(const char* name, int ndim, const int64_t* starts, const int64_t* lens, uint8_t* out, int cap);
void dfmio_bundle_check_block_crc(int on); /* 1 (default) = verify the index's table-block CRCs; 0 = fuzzing only */
void* dfmio_bundle_writer_create(const char* prefix);
/* num_shards data files (<prefix>.data-SSSSS-of-NNNNN), variable k in shard k % num_shards: the layout a sharded TF Saver
 * produces with one parameter-server task per shard; the reader handles any number of shards */
void* dfmio_bundle_writer_create_sharded(const char* prefix, int num_shards);
int dfmio_bundle_writer_add(void* writer, const char* name, int dtype, int ndim, const int64_t* dims, const void* bytes,
                            uint64_t nbytes);
/* A variable partitioned along axis 0 into num_parts near-equal pieces (tf.fixed_size_partitioner): a full-shape entry
 * that lists the slices + one data entry per slice under TF's OrderedCode slice key. The reader assembles partitioned
 * variables (written here or by a TF-1 parameter-server job with a partitioner) back into the full tensor. */
int dfmio_bundle_writer_add_partitioned(void* writer, const char* name, int dtype, int ndim, const int64_t* dims,
                                        const void* bytes, uint64_t nbytes, int num_parts);
int dfmio_bundle_writer_finish(void* writer); /* sorts the keys, writes both files, frees the writer */
void* dfmio_bundle_reader_open(const char* prefix);
int dfmio_bundle_reader_count(void* reader);
int dfmio_bundle_reader_entry(void* reader, int i, char* name, int cap, int* dtype, int* ndim, int64_t* dims,
                              uint64_t* nbytes);
int dfmio_bundle_reader_read(void* reader, int i, void* dst, uint64_t cap, int verify_crc);
void dfmio_bundle_reader_close(void* reader);

#ifdef __cplusplus
}
#endif
#endif /* DEEPFM_IO_H_ */
