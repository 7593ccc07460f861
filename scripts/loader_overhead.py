"""Consumer-side cost of one batch through the Python iterator (no GPU needed): tiny records (F = 1, batch 64) make the
parse work negligible, so the result is the per-batch overhead the training loop's thread pays - slot hand-off,
tensor views, release - which has to stay well under a 0.1 ms training step."""
import ctypes, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepfm_b200.data import native
from deepfm_b200.data.native import NativeLoader
from deepfm_b200.data.pipeline import InputPipeline
from deepfm_b200.data.synthetic import write_synthetic_tfrecords

F, N, B = 1, 100, 64
d = tempfile.mkdtemp()
paths = write_synthetic_tfrecords(d, "tr", 1, 65536, F, N, seed=0)
t0 = time.time(); n = 0
for b in InputPipeline(paths, F, B, threads=4, pin_memory=False, check_crc=False, nslots=16, num_epochs=20):
    n += 1; b.release()
print("InputPipeline: %.1f us per batch (%d batches)" % ((time.time() - t0) / n * 1e6, n))

lib = native.load()
ld = NativeLoader(paths, F, B, num_epochs=20, threads=4, check_crc=False, nslots=16)
slot = ctypes.c_int(-1)
t_next = t_views = t_rel = 0.0; n = 0
pc = time.perf_counter
while True:
    a = pc(); rows = lib.dfmio_loader_next(ld._h, ctypes.byref(slot)); b = pc()
    if rows <= 0:
        break
    ld.slot_views(slot.value, rows); c = pc()
    lib.dfmio_loader_release(ld._h, slot.value); e = pc()
    t_next += b - a; t_views += c - b; t_rel += e - c; n += 1
print("breakdown: next %.1f us, uncached slot views %.1f us, release %.1f us" % (t_next / n * 1e6, t_views / n * 1e6, t_rel / n * 1e6))
