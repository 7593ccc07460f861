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
