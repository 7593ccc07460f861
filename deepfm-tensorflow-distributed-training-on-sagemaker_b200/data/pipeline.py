"""Input pipeline: the framework's `input_fn` (reference: PS:76-133, HVD:74-133).

    TFRecord files | FIFO channel  ->  shard(n, i)  ->  batch(B, drop_remainder=True)  ->  vectorised Example parse
    ->  repeat(num_epochs)  ->  prefetch

Same stage order as the reference (batch BEFORE the parse = the "vectorized map", PDF p43). The native loader
(csrc/io/dfm_io.cc) does it with one sequential reader thread + parser threads writing straight into pinned
"packed slots" (labels | ids | vals) that go to the GPU with ONE async copy; a pure-Python path with identical
semantics is the fallback and the oracle.
"""
from __future__ import annotations

import glob
import os
import random
from dataclasses import dataclass
from typing import Callable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..utils import tfrecord as T


def discover_files(data_dir: str, prefix: str, shuffle: bool = False, seed: Optional[int] = None) -> List[str]:
    """`glob("%s/**/tr*.tfrecords", recursive=True)` (+ `random.shuffle` for training files), PS:374-377."""
    if not data_dir:
        return []
    files = sorted(glob.glob(os.path.join(data_dir, "**", "%s*.tfrecords" % prefix), recursive=True))
    if shuffle:
        random.Random(seed).shuffle(files)
    return files


def channel_fifo(base_dir: str, channel: str, epoch: int) -> str:
    """SageMaker pipe mode exposes one FIFO per channel per epoch: <base>/<channel>_<epoch> (PDF p20)."""
    return os.path.join(base_dir, "%s_%d" % (channel, epoch))


@dataclass
class Batch:
    rows: int
    labels: torch.Tensor  # [rows] f32
    ids: torch.Tensor     # [rows, F] int32
    vals: torch.Tensor    # [rows, F] f32
    packed: Optional[torch.Tensor] = None  # whole pinned slot (labels|ids|vals) when it came from the native loader
    _release: Optional[Callable[[], None]] = None
    raw: Optional[torch.Tensor] = None     # whole pinned RAW slot (offsets|record bytes) when gpu_parse is on

    def release(self):
        if self._release is not None:
            self._release()
            self._release = None

    def __iter__(self):  # lets `for ids, vals, labels in batches` work (Trainer.evaluate/predict)
        return iter((self.ids, self.vals, self.labels))

    def __getitem__(self, i):
        return (self.ids, self.vals, self.labels)[i]


class InputPipeline:
    """Iterable of `Batch`. `sources` = file paths, or in pipe mode one FIFO path per epoch."""

    def __init__(self, sources: Sequence[str], field_size: int, batch_size: int, num_epochs: int = 1,
                 shard: Tuple[int, int] = (1, 0), drop_remainder: bool = True, threads: int = 4, pin_memory: bool = True,
                 use_native: bool = True, check_crc: bool = True, nslots: int = 8, gpu_parse: bool = False,
                 max_record_bytes: int = 1024, shuffle_seed: int = 0):
        """gpu_parse=True: the host only walks/CRC-checks/copies records into pinned raw slots; the Example protos are
        decoded on the GPU (`Trainer.train_raw`). Batches then carry `raw` instead of parsed tensors."""
        self.sources = list(sources)
        self.F, self.B = field_size, batch_size
        self.num_epochs = max(1, num_epochs)
        self.shard = shard
        self.drop = drop_remainder
        self.threads, self.pin, self.crc, self.nslots = threads, pin_memory, check_crc, nslots
        self.gpu_parse, self.max_record_bytes = gpu_parse, max_record_bytes
        # != 0: every epoch visits this rank's records in a fresh reproducible pseudo-random order (native loader over
        # regular files; FIFO sources and the pure-Python fallback keep the stream order)
        self.shuffle_seed = int(shuffle_seed)
        self.native = False
        if use_native:
            try:
                from . import native
                self.native = native.available()
            except Exception:
                self.native = False

    def __iter__(self) -> Iterator[Batch]:
        if not self.sources:
            return iter(())
        if self.gpu_parse and self.native:
            return self._iter_raw()
        return self._iter_native() if self.native else self._iter_python()

    def _iter_raw(self) -> Iterator[Batch]:
        from .native import NativeRawLoader
        ld = NativeRawLoader(self.sources, self.B, self.shard[0], self.shard[1], self.num_epochs, self.drop, self.crc,
                             self.nslots, self.pin, self.max_record_bytes, self.shuffle_seed)
        self.raw_off_bytes = ld.off_bytes
        try:
            while True:
                b = ld.next()
                if b is None:
                    return
                slot, rows, raw = b
                yield Batch(rows, None, None, None, None, (lambda s=slot: ld.release(s)), raw)
        finally:
            ld.close()

    def _iter_native(self) -> Iterator[Batch]:
        from .native import NativeLoader
        ld = NativeLoader(self.sources, self.F, self.B, self.shard[0], self.shard[1], self.num_epochs, self.threads,
                          self.drop, self.crc, self.nslots, self.pin, self.shuffle_seed)
        try:
            while True:
                b = ld.next()
                if b is None:
                    return
                slot, rows, lab, ids, vals = b
                packed = ld.full_views(slot)[3] if rows == self.B else None
                yield Batch(rows, lab, ids, vals, packed, (lambda s=slot: ld.release(s)))
        finally:
            ld.close()

    def _iter_python(self) -> Iterator[Batch]:
        n, i = self.shard
        idx = 0
        for _ in range(self.num_epochs):
            buf: List[bytes] = []
            for path in self.sources:
                for rec in T.read_records(path, self.crc):
                    keep = (idx % n) == i
                    idx += 1
                    if not keep:
                        continue
                    buf.append(rec)
                    if len(buf) == self.B:
                        yield self._parse(buf)
                        buf = []
            if buf and not self.drop:  # batch() precedes repeat(): the remainder never crosses an epoch boundary
                yield self._parse(buf)

    def _parse(self, recs: List[bytes]) -> Batch:
        lab, ids, vals = T.parse_ctr_batch(recs, self.F)
        return Batch(len(recs), torch.from_numpy(lab), torch.from_numpy(ids.astype(np.int32)), torch.from_numpy(vals))


def count_records(paths: Sequence[str]) -> int:
    n = 0
    try:
        from .native import available, count_records_native
        if available():
            return sum(count_records_native(p) for p in paths)
    except Exception:
        pass
    for p in paths:
        for _ in T.read_records(p, check_crc=False):
            n += 1
    return n


def local_steps(paths: Sequence[str], batch_size: int, shard: Tuple[int, int], num_epochs: int = 1) -> int:
    """Number of full batches this worker will get from `paths` under `dataset.shard(n, i).batch(B, drop_remainder)`."""
    total = count_records(paths)
    n, i = shard
    mine = (total - i + n - 1) // n if total > i else 0
    return (mine // batch_size) * num_epochs


# ---------------------------------------------------------------------------------------------- binary cache
CACHE_MAGIC = b"DFMCACH1"


def write_cache(path: str, batches, field_size: int) -> int:
    """Pre-decoded dataset cache (the deck's `dataset.cache()` advice, PDF p43): header + labels f32[n] + ids i32[n,F]
    + vals f32[n,F]; memory-mappable, so an epoch costs one memcpy per batch instead of a protobuf parse."""
    labs, ids, vals = [], [], []
    for b in batches:
        labs.append(b.labels.clone()); ids.append(b.ids.clone()); vals.append(b.vals.clone())
        b.release()
    lab = torch.cat(labs).numpy(); i = torch.cat(ids).numpy(); v = torch.cat(vals).numpy()
    with open(path, "wb") as f:
        f.write(CACHE_MAGIC)
        f.write(np.asarray([lab.shape[0], field_size], dtype=np.int64).tobytes())
        f.write(lab.astype(np.float32).tobytes()); f.write(i.astype(np.int32).tobytes()); f.write(v.astype(np.float32).tobytes())
    return int(lab.shape[0])


class CachedDataset:
    def __init__(self, path: str):
        raw = np.memmap(path, dtype=np.uint8, mode="r")
        if bytes(raw[:8]) != CACHE_MAGIC:
            raise IOError("not a deepfm_b200 cache file: %s" % path)
        n, F = np.frombuffer(raw[8:24], dtype=np.int64)
        self.n, self.F = int(n), int(F)
        o = 24
        self.labels = np.frombuffer(raw, dtype=np.float32, count=self.n, offset=o); o += 4 * self.n
        self.ids = np.frombuffer(raw, dtype=np.int32, count=self.n * self.F, offset=o).reshape(self.n, self.F); o += 4 * self.n * self.F
        self.vals = np.frombuffer(raw, dtype=np.float32, count=self.n * self.F, offset=o).reshape(self.n, self.F)

    def batches(self, batch_size: int, num_epochs: int = 1, shard: Tuple[int, int] = (1, 0), drop_remainder: bool = True,
                packed: bool = False, nslots: int = 8):
        """packed=True: full batches are gathered straight into a ring of `nslots` pinned buffers in the loader's packed
        slot layout (labels | ids | vals), so `Trainer.train_packed` moves each with ONE async H2D copy — an epoch from the
        cache costs three strided memcpys per batch and no protobuf parse (the deck's `dataset.cache()`, PDF p43). A
        buffer is reused after `nslots` further batches: keep fewer than that in flight."""
        n, i = shard
        sel = np.arange(i, self.n, n)
        ring, k = None, 0
        if packed:
            from ..models.fused import make_packed_slot
            pin = torch.cuda.is_available()
            ring = [make_packed_slot(batch_size, self.F, "cpu", pin=pin) for _ in range(nslots)]
        contiguous = n == 1
        for _ in range(num_epochs):
            for s in range(0, len(sel), batch_size):
                idx = sel[s:s + batch_size]
                if len(idx) < batch_size and drop_remainder:
                    break
                key = slice(int(idx[0]), int(idx[0]) + len(idx)) if contiguous else idx
                if ring is not None and len(idx) == batch_size:
                    sl = ring[k % nslots]
                    k += 1
                    sl["labels"].numpy()[:] = self.labels[key]
                    sl["ids"].numpy()[:] = self.ids[key]
                    sl["vals"].numpy()[:] = self.vals[key]
                    yield Batch(batch_size, sl["labels"], sl["ids"], sl["vals"], sl["packed"])
                    continue
                yield Batch(len(idx), torch.from_numpy(np.array(self.labels[key])), torch.from_numpy(np.array(self.ids[key])),
                            torch.from_numpy(np.array(self.vals[key])))
