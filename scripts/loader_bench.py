"""Host input-pipeline throughput (no GPU needed): TFRecord -> CRC32C -> batch -> Example parse -> pinned-slot layout,
native loader vs the pure-Python path, plus the pre-decoded cache."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.data.synthetic import write_synthetic_tfrecords
from deepfm_b200.data.pipeline import InputPipeline, write_cache, CachedDataset

F, N, B = 39, 117581, 1024
d = tempfile.mkdtemp()
t0 = time.time()
paths = write_synthetic_tfrecords(d, "tr", 4, 65536, F, N, seed=0)
size = sum(os.path.getsize(p) for p in paths)
print("wrote %d records, %.1f MB (%.0f B/record) in %.1f s" % (4 * 65536, size / 1e6, size / (4 * 65536), time.time() - t0))
EPOCHS = 16  # 4.2 M records per measurement: long enough that the ~20 ms start-up (map + index + threads) is not the result
for threads in (1, 2, 4, 6, 8):
    for crc in (True, False):
        t0 = time.time(); rows = 0; first = None
        for b in InputPipeline(paths, F, B, threads=threads, pin_memory=False, check_crc=crc, nslots=max(8, 3 * threads),
                               num_epochs=EPOCHS):
            if first is None:
                first = time.time() - t0
            rows += b.rows; b.release()
        dt = time.time() - t0
        print("native loader threads=%d crc=%s: steady %.2f M records/s (%.0f MB/s), first batch after %.0f ms"
              % (threads, crc, (rows - B) / (dt - first) / 1e6, EPOCHS * size / (dt - first) / 1e6, first * 1e3))
for crc in (True, False):
    t0 = time.time(); rows = 0
    for b in InputPipeline(paths, F, B, pin_memory=False, check_crc=crc, gpu_parse=True, num_epochs=EPOCHS):
        rows += b.rows; b.release()
    dt = time.time() - t0
    print("raw loader (records copied for GPU parsing) crc=%s: %.2f M records/s (%.0f MB/s)" % (crc, rows / dt / 1e6, EPOCHS * size / dt / 1e6))
t0 = time.time(); rows = 0
for b in InputPipeline(paths[:1], F, B, use_native=False):
    rows += b.rows
    if rows >= 8192: break
print("python fallback: %.4f M records/s" % (rows / (time.time() - t0) / 1e6))
cache = os.path.join(d, "train.dfmcache")
n = write_cache(cache, InputPipeline(paths, F, B, threads=4, pin_memory=False), F)
ds = CachedDataset(cache)
t0 = time.time(); rows = 0
for b in ds.batches(B):
    rows += b.rows
print("binary cache: %.2f M records/s" % (rows / (time.time() - t0) / 1e6))
