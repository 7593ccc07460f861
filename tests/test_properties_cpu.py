"""Property-based tests (hypothesis) for the host-side formats and pure functions."""
import io

import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from deepfm_b200.parallel.shard import hvd_shard, normalise
from deepfm_b200.utils import tfrecord as T
from deepfm_b200.utils.auc import StreamingAUC, exact_auc

ids_st = st.lists(st.integers(min_value=0, max_value=2 ** 31 - 1), min_size=1, max_size=12)


@settings(max_examples=60, deadline=None)
@given(payloads=st.lists(st.binary(min_size=0, max_size=300), min_size=0, max_size=8))
def test_tfrecord_framing_roundtrip(payloads):
    buf = io.BytesIO()
    w = T.TFRecordWriter(buf)
    for p in payloads:
        w.write(p)
    buf.seek(0)
    assert list(T.read_records(buf)) == payloads


@settings(max_examples=60, deadline=None)
@given(ids=ids_st, label=st.floats(min_value=0, max_value=1, width=32), data=st.data())
def test_example_codec_python_vs_native(ids, label, data):
    from deepfm_b200.data import native
    vals = data.draw(st.lists(st.floats(min_value=-10, max_value=10, width=32), min_size=len(ids), max_size=len(ids)))
    ex = T.make_ctr_example(label, ids, vals)
    l1, i1, v1 = T.parse_ctr_batch([ex, ex], len(ids))
    l2, i2, v2 = native.parse_batch_native([ex, ex], len(ids))
    assert np.array_equal(l1, l2) and np.array_equal(i1.astype(np.int32), i2) and np.array_equal(v1, v2)
    assert i1[0].tolist() == ids and np.allclose(v1[0], np.asarray(vals, dtype=np.float32))


@settings(max_examples=100, deadline=None)
@given(hosts=st.integers(1, 4), wph=st.integers(1, 8), pipe=st.booleans(), s3=st.booleans(), multi=st.booleans(),
       data=st.data())
def test_shard_table_covers_every_record_exactly_once_per_view(hosts, wph, pipe, s3, multi, data):
    """Within the set of workers that read the SAME stream, the shards are disjoint and complete."""
    world = hosts * wph
    rank = data.draw(st.integers(0, world - 1))
    spec = normalise(hvd_shard(pipe, s3, multi, world, rank, rank % wph, wph, hosts))
    n, i = spec
    assert 0 <= i < n
    if not pipe and not s3:
        # every worker sees every file: the world partitions the records
        seen = sorted(normalise(hvd_shard(pipe, s3, multi, world, r, r % wph, wph, hosts))[1] for r in range(world))
        assert seen == list(range(world))
    if not pipe and s3:
        # each host got its own files: the host's local workers partition them
        seen = sorted(normalise(hvd_shard(pipe, s3, multi, world, h0 + l, l, wph, hosts))[1]
                      for h0 in [0] for l in range(wph))
        assert seen == list(range(wph))


@settings(max_examples=25, deadline=None)
@given(seed=st.integers(0, 10 ** 6), n=st.integers(200, 3000))
def test_bucketed_auc_tracks_exact_auc(seed, n):
    g = torch.Generator().manual_seed(seed)
    p = torch.rand(n, generator=g)
    y = (torch.rand(n, generator=g) < p).float()
    if y.sum() in (0, n):
        return
    a = StreamingAUC()
    a.update(y, p)
    assert abs(a.result() - exact_auc(y.numpy(), p.numpy())) < 1e-2


@settings(max_examples=300, deadline=None)
@given(name=st.binary(min_size=1, max_size=24).filter(lambda b: True),
       extents=st.lists(st.one_of(st.just((-1, -1)),
                                  st.tuples(st.integers(0, 2 ** 40), st.integers(0, 2 ** 40))), min_size=0, max_size=4))
def test_slice_key_encoding_matches_an_independent_ordered_code_derivation(name, extents):
    """Partitioned variables are found in a TF checkpoint index under OrderedCode keys (name, rank, (start, length) per
    dimension). The C++ encoder (csrc/io/tensor_bundle.cc) against the Python derivation of the same rules in
    P/utils/tf_bundle.py: every magnitude class of the signed varint (1..6 bytes), the (-1, -1) "whole axis" marker and
    names with 0x00 / 0xff bytes (escaped)."""
    import ctypes
    from deepfm_b200.data import native
    from deepfm_b200.utils.tf_bundle import ordered_code_slice_key
    lib = native.load()
    n = len(extents)
    starts = (ctypes.c_int64 * max(n, 1))(*[e[0] for e in extents])
    lens = (ctypes.c_int64 * max(n, 1))(*[e[1] for e in extents])
    out = (ctypes.c_uint8 * 512)()
    # names reach C as NUL-terminated strings: a name with an embedded 0 byte cannot be passed, 0xff can
    cname = name.replace(b"\x00", b"\x01")
    lib.dfmio_bundle_slice_key.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_int]
    k = lib.dfmio_bundle_slice_key(cname, n, starts, lens, out, 512)
    assert k > 0
    expect = ordered_code_slice_key(cname, extents)
    assert bytes(out[:k]) == expect
