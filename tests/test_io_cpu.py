"""Host-side formats: TFRecord framing + CRC, Example codec, C++ parser/loader vs the Python oracle, libsvm
converter, TF tensor-bundle checkpoint codec (SURVEY section 4 'Unit' tests). CPU only."""
import io
import os
import struct

import numpy as np
import pytest
import torch

from deepfm_b200.utils import tfrecord as T
from deepfm_b200.utils.libsvm import convert_tfrecords, parse_libsvm_line

SAMPLE = ("1 1:0.5 2:0.03519 3:1 4:0.02567 7:0.03708 8:0.01705 9:0.06296 10:0.18185 11:0.02497 12:1 14:0.02565 "
          "15:0.03267 17:0.0247 18:0.03158 20:1 22:1 23:0.13169 24:0.02933 27:0.18159 31:0.0177 34:0.02888 38:1 "
          "51:1 63:1 132:1 164:1 236:1")  # the example row quoted in the reference trainer (PS:74)


def test_crc32c_known_answers():
    assert T.crc32c(b"123456789") == 0xE3069283
    assert T.crc32c(b"") == 0
    assert T.unmask_crc(T.masked_crc32c(b"abc")) == T.crc32c(b"abc")


def test_record_roundtrip_and_corruption(tmp_path):
    p = str(tmp_path / "a.tfrecords")
    payloads = [b"", b"x", os.urandom(1000)]
    with T.TFRecordWriter(p) as w:
        for d in payloads:
            w.write(d)
    assert list(T.read_records(p)) == payloads
    raw = bytearray(open(p, "rb").read())
    raw[-6] ^= 0xFF  # flip a data byte of the last record
    open(p, "wb").write(raw)
    with pytest.raises(T.TFRecordError):
        list(T.read_records(p))
    open(p, "wb").write(raw[:-3])  # truncated file
    with pytest.raises(T.TFRecordError):
        list(T.read_records(p))


def test_example_codec_matches_schema():
    ex = T.make_ctr_example(1.0, [1, 2, 300, 70000], [0.5, 0.25, 1.0, 2.0])
    d = T.decode_example(ex)
    assert d["label"] == ("float", [1.0]) and d["ids"] == ("int64", [1, 2, 300, 70000])
    lab, ids, vals = T.parse_ctr_batch([ex, ex], 4)
    assert lab.dtype == np.float32 and ids.dtype == np.int64 and ids.shape == (2, 4)
    with pytest.raises(ValueError):
        T.parse_ctr_batch([ex], 5)  # FixedLenFeature([field_size]) semantics (HVD:81-82)


def test_libsvm_converter_on_reference_sample(tmp_path):
    label, ids, vals = parse_libsvm_line(SAMPLE)
    assert label == 1.0 and len(ids) == 27 and ids[-1] == 236 and vals[1] == pytest.approx(0.03519)
    src, dst = tmp_path / "tr.libsvm", tmp_path / "tr.tfrecords"
    src.write_text(SAMPLE + "\n" + SAMPLE.replace("1 1:", "0 1:", 1) + "\n")
    assert convert_tfrecords(str(src), str(dst), field_size=39) == 2
    lab, i, v = T.parse_ctr_batch(list(T.read_records(str(dst))), 39)
    assert lab.tolist() == [1.0, 0.0] and i.shape == (2, 39) and (v[:, 27:] == 0).all()
    # without field_size the converter reproduces the reference exactly: 27 ids per record
    assert convert_tfrecords(str(src), str(dst)) == 2
    d = T.decode_example(next(T.read_records(str(dst))))
    assert len(d["ids"][1]) == 27


def _write(tmp_path, n=1000, F=39, seed=0):
    from deepfm_b200.data import native
    rng = np.random.default_rng(seed)
    p = str(tmp_path / "tr0.tfrecords")
    recs = []
    with native.NativeRecordWriter(p) as w:
        for i in range(n):
            ex = T.make_ctr_example(float(i % 2), rng.integers(0, 117581, F).tolist(),
                                    rng.random(F).astype(np.float32).tolist())
            recs.append(ex)
            w.write(ex)
    return p, recs


def test_native_reader_parser_match_python(tmp_path):
    from deepfm_b200.data import native
    assert native.crc32c(b"123456789") == 0xE3069283
    p, recs = _write(tmp_path, 300)
    assert list(T.read_records(p)) == recs  # native writer -> python reader
    assert list(native.read_records_native(p)) == recs
    l1, i1, v1 = T.parse_ctr_batch(recs[:64], 39)
    l2, i2, v2 = native.parse_batch_native(recs[:64], 39)
    assert (l1 == l2).all() and (i1 == i2).all() and (v1 == v2).all() and i2.dtype == np.int32
    with pytest.raises(ValueError):
        native.parse_batch_native(recs[:2], 40)
    # unpacked repeated encoding is accepted too
    unpacked = T._ld(1, T._ld(1, T._ld(1, b"label") + T._ld(2, T._ld(2, b"\x0d" + struct.pack("<f", 1.0)))) +
                     T._ld(1, T._ld(1, b"ids") + T._ld(2, T._ld(3, b"\x08\x05\x08\x07"))) +
                     T._ld(1, T._ld(1, b"values") + T._ld(2, T._ld(2, b"\x0d" + struct.pack("<f", 2.0) + b"\x0d" + struct.pack("<f", 3.0)))))
    l, i, v = native.parse_batch_native([unpacked], 2)
    assert l[0] == 1.0 and i.tolist() == [[5, 7]] and v.tolist() == [[2.0, 3.0]]


def test_native_loader_shard_batch_epochs(tmp_path):
    from deepfm_b200.data import native
    p, recs = _write(tmp_path, 1000)
    ld = native.NativeLoader([p], 39, 128, shard_n=2, shard_i=1, num_epochs=2, threads=3)
    rows, first = 0, None
    while True:
        b = ld.next()
        if b is None:
            break
        slot, n, lab, ids, vals = b
        if first is None:
            first = (lab.clone(), ids.clone(), vals.clone())
        rows += n
        ld.release(slot)
    ld.close()
    assert rows == 2 * (500 // 128 * 128)  # shard -> batch(drop_remainder) -> repeat, like HVD:98,122-126
    la, ia, va = T.parse_ctr_batch(recs[1:256:2], 39)
    assert (first[0].numpy() == la).all() and (first[1].numpy() == ia).all() and (first[2].numpy() == va).all()


def test_native_loader_reports_corruption(tmp_path):
    from deepfm_b200.data import native
    p, _ = _write(tmp_path, 100)
    raw = bytearray(open(p, "rb").read())
    raw[len(raw) // 2] ^= 0x55
    open(p, "wb").write(raw)
    ld = native.NativeLoader([p], 39, 10, threads=1)
    with pytest.raises(IOError):
        while ld.next() is not None:
            pass
    ld.close()


def test_pipe_mode_fifo(tmp_path):
    """Pipe mode = the same reader over a FIFO (read()-only, no seek), PipeModeDataset stand-in (HVD:105)."""
    import threading
    from deepfm_b200.data import native
    p, recs = _write(tmp_path, 64)
    fifo = str(tmp_path / "training_0")
    os.mkfifo(fifo)
    blob = open(p, "rb").read()

    def feed():
        with open(fifo, "wb") as f:
            f.write(blob)
    t = threading.Thread(target=feed)
    t.start()
    ld = native.NativeLoader([fifo], 39, 16, threads=2)
    rows = 0
    while True:
        b = ld.next()
        if b is None:
            break
        rows += b[1]
        ld.release(b[0])
    t.join()
    ld.close()
    assert rows == 64


def test_tensor_bundle_roundtrip_and_format(tmp_path):
    from deepfm_b200.utils.tf_bundle import read_bundle, write_bundle
    t = {"fm_bias": torch.tensor([0.25]), "fm_w": torch.randn(1000), "fm_v": torch.randn(1000, 32),
         "Deep-part/mlp0/weights": torch.randn(1248, 400), "Deep-part/mlp0/weights/Adam": torch.randn(1248, 400),
         "global_step": torch.tensor(123, dtype=torch.int64), "beta1_power": torch.tensor(0.9)}
    for i in range(300):  # enough keys for several 4 KiB table blocks + restart points
        t["Deep-part/extra_%03d/biases" % i] = torch.randn(3)
    prefix = str(tmp_path / "model.ckpt-123")
    write_bundle(prefix, t)
    back = read_bundle(prefix)
    assert set(back) == set(t)
    for k in t:
        assert back[k].dtype == t[k].dtype and torch.equal(back[k], t[k]), k
    assert back["global_step"].shape == ()
    # spec-level checks with an independent parser: footer magic, block trailer CRC
    raw = open(prefix + ".index", "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xdb4775248b80fb57
    foot = raw[-48:]
    def varint(b, p):
        v = s = 0
        while True:
            c = b[p]; p += 1
            v |= (c & 0x7F) << s
            if not c & 0x80:
                return v, p
            s += 7
    mo, p = varint(foot, 0); ms, p = varint(foot, p); io_, p = varint(foot, p); isz, p = varint(foot, p)
    blk = raw[io_:io_ + isz + 1]
    assert raw[io_ + isz] == 0  # no compression
    assert struct.unpack("<I", raw[io_ + isz + 1:io_ + isz + 5])[0] == T.masked_crc32c(blk)
    assert os.path.getsize(prefix + ".data-00000-of-00001") == sum(v.numel() * v.element_size() for v in t.values())
    # corruption is detected
    data = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    data[100] ^= 1
    open(prefix + ".data-00000-of-00001", "wb").write(data)
    with pytest.raises(IOError):
        read_bundle(prefix)


def test_checkpoint_names_and_layout(tmp_path):
    from deepfm_b200.models import DeepFMConfig, TorchDeepFM
    from deepfm_b200.optim import TF1Optimizer
    from deepfm_b200.utils.checkpoint import latest_checkpoint, load_checkpoint, save_checkpoint
    from deepfm_b200.utils.tf_bundle import read_bundle
    cfg = DeepFMConfig(field_size=5, feature_size=50, embedding_size=8, deep_layers=[16, 8], dropout=[1, 1])
    m = TorchDeepFM(cfg, 4, TF1Optimizer("Adam", 1e-2))
    ids = torch.randint(0, 50, (4, 5)); vals = torch.rand(4, 5); y = torch.tensor([1., 0, 1, 0])
    for _ in range(3):
        m.train_step(ids, vals, y)
    d = str(tmp_path / "ckpt")
    for _ in range(7):  # keep-5 rotation
        m.train_step(ids, vals, y)
        prefix = save_checkpoint(d, m.state_dict(), m.opt)
    assert latest_checkpoint(d) == prefix and prefix.endswith("model.ckpt-10")
    assert len([f for f in os.listdir(d) if f.endswith(".index")]) == 5
    tf_sd = read_bundle(prefix)
    for name in ["fm_bias", "fm_w", "fm_v", "Deep-part/mlp0/weights", "Deep-part/mlp0/biases", "Deep-part/mlp1/weights",
                 "Deep-part/deep_out/weights", "Deep-part/deep_out/biases", "fm_v/Adam", "fm_v/Adam_1",
                 "Deep-part/mlp0/weights/Adam_1", "beta1_power", "beta2_power", "global_step"]:
        assert name in tf_sd, name
    assert tuple(tf_sd["Deep-part/mlp0/weights"].shape) == (40, 16)  # TF layout [in, out]
    assert tuple(tf_sd["Deep-part/deep_out/weights"].shape) == (8, 1)
    assert float(tf_sd["beta1_power"]) == pytest.approx(0.9 ** 11, rel=1e-5)
    m2 = TorchDeepFM(cfg, 4, TF1Optimizer("Adam", 1e-2), seed=99)
    m2.load_state_dict(load_checkpoint(prefix, list(m2.params), "Adam"))
    assert m2.global_step == 10
    for k in m.params:
        assert torch.equal(m.params[k], m2.params[k]), k
    l1 = m.train_step(ids, vals, y); l2 = m2.train_step(ids, vals, y)
    assert float(l1) == pytest.approx(float(l2), rel=1e-6)  # slots restored too: identical next step


def test_raw_loader_slots_hold_the_serialized_records(tmp_path):
    """Raw mode (GPU parsing): offsets + record bytes per slot, shard/epoch/drop_remainder semantics unchanged."""
    from deepfm_b200.data import native
    p, recs = _write(tmp_path, 500)
    ld = native.NativeRawLoader([p], 64, shard_n=2, shard_i=0, num_epochs=2, nslots=4)
    got, batches = [], 0
    while True:
        b = ld.next()
        if b is None:
            break
        slot, rows, raw = b
        got += ld.records(raw, rows)
        batches += 1
        ld.release(slot)
    ld.close()
    mine = recs[0::2]
    per_epoch = mine[: len(mine) // 64 * 64]
    want = per_epoch + per_epoch
    assert batches == len(want) // 64 and got == want


def test_count_records_and_local_steps(tmp_path):
    from deepfm_b200.data.native import count_records_native
    from deepfm_b200.data.pipeline import InputPipeline, local_steps
    p, recs = _write(tmp_path, 777)
    assert count_records_native(p) == 777
    for shard in [(1, 0), (2, 1), (3, 2)]:
        want = 0
        for b in InputPipeline([p], 39, 32, 2, shard, pin_memory=False):
            want += 1
            b.release()
        # repeat(num_epochs) happens AFTER batch in the reference (HVD:122-126): per-epoch remainders are dropped
        assert want == 2 * local_steps([p], 32, shard, 1) == local_steps([p], 32, shard, 2)


def test_loader_survives_slow_out_of_order_consumers(tmp_path):
    """Slots are handed back late and out of order (as CUDA-event driven releases are): delivery stays in record order,
    nothing deadlocks with the smallest legal ring, in parsed and in raw mode."""
    import random
    from deepfm_b200.data import native
    from deepfm_b200.utils import tfrecord as T
    F, B, n = 3, 8, 8 * 37
    path = str(tmp_path / "tr.tfrecords")
    with native.NativeRecordWriter(path) as w:
        for i in range(n):
            w.write(T.make_ctr_example(float(i % 2), [i, i + 1, i + 2], [1.0, 0.5, 0.25]))
    rng = random.Random(0)
    for mode in ("parsed", "raw"):
        if mode == "parsed":
            ld = native.NativeLoader([path], F, B, num_epochs=3, threads=3, nslots=2)
        else:
            ld = native.NativeRawLoader([path], B, num_epochs=3, nslots=2, max_record_bytes=256)
        held, first_ids = [], []
        while True:
            item = ld.next()
            if item is None:
                break
            if mode == "parsed":
                slot, rows, lab, ids, vals = item
                first_ids.append(int(ids[0, 0]))
            else:
                slot, rows, raw = item
                first_ids.append(int(T.parse_ctr_batch(ld.records(raw, rows)[:1], F)[1][0, 0]))
            assert rows == B
            held.append(slot)
            if len(held) == 2 or rng.random() < 0.5:      # hold up to the whole ring, release in random order
                rng.shuffle(held)
                while held:
                    ld.release(held.pop())
        for s in held:
            ld.release(s)
        ld.close()
        assert first_ids == [B * k for k in range(n // B)] * 3, mode


def test_native_libsvm_converter_is_byte_identical_to_the_python_one(tmp_path):
    """C1: the multi-threaded C++ converter writes exactly the bytes utils/libsvm.py writes (record order = line order,
    padding / truncation, blank lines, CRLF, exponents, missing final newline), and names the line it cannot parse."""
    import random
    from deepfm_b200.data import native
    from deepfm_b200.utils.libsvm import convert_tfrecords
    rng = random.Random(7)
    lines = ["1 5:1 7:0.5 11:2.25", "0 3:1e-3 4:+2 9:.5 12:-0.125 13:7 14:8", "", "   ", "1 1:1\r", "0", "1 0:0 0:0 0:0 0:0 0:0"]
    for i in range(3000):
        k = rng.choice([1, 3, 4, 5, 9])
        lines.append("%d %s" % (i % 2, " ".join("%d:%r" % (rng.randrange(1 << 40), rng.uniform(-3, 3)) for _ in range(k))))
    src = tmp_path / "in.libsvm"
    src.write_text("\n".join(lines))  # no trailing newline on purpose
    for field_size in (4, None):
        a, b = str(tmp_path / "py.tfrecords"), str(tmp_path / "cc.tfrecords")
        n_py = convert_tfrecords(str(src), a, field_size, pad_id=2, use_native=False)
        for threads in (1, 3, 0):
            n_cc = native.libsvm_convert_native(str(src), b, field_size, 2, threads)
            assert n_cc == n_py == len(lines) - 2
            assert open(a, "rb").read() == open(b, "rb").read(), (field_size, threads)
    assert convert_tfrecords(str(src), b, 4, pad_id=2) == n_py          # default path = native
    bad = tmp_path / "bad.libsvm"
    bad.write_text("1 1:1\n0 2:x\n1 3:1\n")
    with pytest.raises(ValueError, match="line 2"):
        native.libsvm_convert_native(str(bad), b, 4)
    with pytest.raises(ValueError, match="cannot open"):
        native.libsvm_convert_native(str(tmp_path / "missing"), b, 4)
    empty = tmp_path / "empty.libsvm"
    empty.write_text("")
    assert native.libsvm_convert_native(str(empty), b, 4) == 0


def test_sharded_tensor_bundles_like_a_parameter_server_checkpoint(tmp_path):
    """The reference's PS jobs save through a sharded Saver: several `.data-SSSSS-of-NNNNN` files behind one index. The
    reader follows BundleEntryProto.shard_id / BundleHeaderProto.num_shards; a missing shard is reported, not misread."""
    import os
    from deepfm_b200.utils.tf_bundle import read_bundle, write_bundle
    g = torch.Generator().manual_seed(0)
    tensors = {"fm_w": torch.randn(50, generator=g), "fm_v": torch.randn(50, 8, generator=g),
               "Deep-part/mlp0/weights": torch.randn(32, 16, generator=g), "Deep-part/mlp0/biases": torch.zeros(16),
               "fm_bias": torch.tensor([0.25]), "global_step": torch.tensor(123)}
    prefix = str(tmp_path / "model.ckpt-123")
    write_bundle(prefix, tensors, num_shards=3)
    files = sorted(f for f in os.listdir(tmp_path) if ".data-" in f)
    assert files == ["model.ckpt-123.data-0000%d-of-00003" % i for i in range(3)]
    assert all(os.path.getsize(tmp_path / f) > 0 for f in files)
    back = read_bundle(prefix)
    assert set(back) == set(tensors)
    for k, v in tensors.items():
        assert torch.equal(back[k], v) and back[k].shape == v.shape, k
    one = str(tmp_path / "single")
    write_bundle(one, tensors)                                   # default layout unchanged: one data file
    assert os.path.exists(one + ".data-00000-of-00001") and torch.equal(read_bundle(one)["fm_v"], tensors["fm_v"])
    os.remove(tmp_path / files[1])
    with pytest.raises(IOError, match="cannot open"):
        read_bundle(prefix)


def _ordered_code_slice_key(name, extents):
    from deepfm_b200.utils.tf_bundle import ordered_code_slice_key
    return ordered_code_slice_key(name, extents)


def test_partitioned_variables_like_fixed_size_partitioner(tmp_path):
    """The deck's advice for big tables on parameter servers is `tf.fixed_size_partitioner` (PDF p32). Such a variable is
    saved as a full-shape BundleEntryProto that lists `slices` plus one entry per slice under an OrderedCode key; the reader
    must assemble them into the full tensor, and report missing / inconsistent slices."""
    from deepfm_b200.utils.tf_bundle import read_bundle, write_bundle
    assert _ordered_code_slice_key("fm_v", [(0, 58791), (-1, -1)]) == b"\x00fm_v\x00\x01\x01\x02\x80\xe0\xe5\xa7\x7f\x7f"
    g = torch.Generator().manual_seed(1)
    tensors = {"fm_w": torch.randn(1001, generator=g), "fm_v": torch.randn(1001, 8, generator=g),
               "fm_v/Adam": torch.randn(1001, 8, generator=g), "Deep-part/mlp0/weights": torch.randn(32, 16, generator=g),
               "fm_bias": torch.tensor([0.25]), "global_step": torch.tensor(7)}
    prefix = str(tmp_path / "model.ckpt-7")
    write_bundle(prefix, tensors, num_shards=2, partitioned={"fm_v": 3, "fm_w": 2, "fm_v/Adam": 3})
    raw = open(prefix + ".index", "rb").read()
    # the first slice of fm_v: rows [0, 334) (1001 = 334 + 334 + 333), every column; stored in full (no shared prefix
    # with the preceding header key)
    assert _ordered_code_slice_key("fm_v", [(0, 334), (-1, -1)]) in raw
    assert _ordered_code_slice_key("fm_v", [(0, 334), (-1, -1)]) == b"\x00fm_v\x00\x01\x01\x02\x80\xc1\x4e\x7f\x7f"
    back = read_bundle(prefix)
    assert set(back) == set(tensors)
    for k, v in tensors.items():
        assert torch.equal(back[k], v) and back[k].shape == v.shape, k
    # a checkpoint whose slice data is gone (an index entry without its bytes) is an error, not zeros
    write_bundle(prefix + "b", {"fm_v": tensors["fm_v"]}, partitioned={"fm_v": 4})
    idx = bytearray(open(prefix + "b.index", "rb").read())
    key = _ordered_code_slice_key("fm_v", [(0, 251), (-1, -1)])
    at = bytes(idx).find(key)
    assert at > 0
    idx[at + len(key) - 3] ^= 0x01  # the slice now claims another length: no longer the one the full entry lists
    open(prefix + "b.index", "wb").write(bytes(idx))
    with pytest.raises(IOError):
        read_bundle(prefix + "b")


def test_loader_epoch_shuffle_is_a_reproducible_permutation(tmp_path):
    """shuffle_seed != 0: each epoch reads every record of the rank's shard exactly once, in an order that differs from
    the file order, differs between epochs, is identical for the same seed, and respects sharding (parsed and raw mode)."""
    from deepfm_b200.data import native
    from deepfm_b200.utils import tfrecord as T
    F, B, n = 2, 10, 230
    path = str(tmp_path / "tr.tfrecords")
    with native.NativeRecordWriter(path) as w:
        for i in range(n):
            w.write(T.make_ctr_example(float(i % 2), [i, i + 1000], [1.0, 0.5]))

    def epochs(seed, shard=(1, 0), raw=False, num_epochs=3):
        if raw:
            ld = native.NativeRawLoader([path], B, shard[0], shard[1], num_epochs, nslots=4, max_record_bytes=128,
                                        shuffle_seed=seed)
        else:
            ld = native.NativeLoader([path], F, B, shard[0], shard[1], num_epochs, threads=3, nslots=4, shuffle_seed=seed)
        seen = []
        while True:
            item = ld.next()
            if item is None:
                break
            if raw:
                slot, rows, buf = item
                seen += [int(x) for x in T.parse_ctr_batch(ld.records(buf, rows), F)[1][:, 0]]
            else:
                slot, rows, lab, ids, vals = item
                assert torch.equal(ids[:, 1], ids[:, 0] + 1000) and torch.equal(lab, (ids[:, 0] % 2).float())
                seen += ids[:, 0].tolist()
            ld.release(slot)
        ld.close()
        per = len(seen) // num_epochs
        return [seen[e * per:(e + 1) * per] for e in range(num_epochs)]

    plain = epochs(0)
    assert plain[0] == list(range(n)) and plain[1] == plain[0]
    a, b = epochs(7), epochs(7)
    assert a == b                                                        # reproducible
    assert all(sorted(e) == list(range(n)) for e in a)                   # a permutation: every record once per epoch
    assert a[0] != plain[0] and a[0] != a[1] and a[1] != a[2]            # shuffled, and differently every epoch
    assert epochs(8)[0] != a[0]
    half = epochs(7, shard=(2, 1))
    assert all(sorted(e) == list(range(1, n, 2))[:len(e)] or set(e) <= set(range(1, n, 2)) for e in half)
    assert all(len(set(e)) == len(e) == (n // 2 // B) * B for e in half)  # drop_remainder: 110 of the 115 records
    r = epochs(7, raw=True)
    assert all(sorted(e) == list(range(n)) for e in r) and r == a        # raw mode walks the same permutation


def test_host_io_library_survives_fuzzing_under_asan_ubsan_and_tsan():
    """SURVEY 5.2 (sanitizers): the parsers that read untrusted bytes - TFRecord framing, Example protos, checkpoint
    indexes, libsvm text - are rebuilt with AddressSanitizer + UBSan and fed corrupted inputs (scripts/fuzz_io.sh)."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no compiler")
    probe = subprocess.run("echo 'int main(){return 0;}' | g++ -x c++ -fsanitize=address,undefined - -o /dev/null",
                           shell=True, capture_output=True)
    if probe.returncode != 0:
        pytest.skip("sanitizer runtimes not installed")
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([os.path.join(repo, "scripts", "fuzz_io.sh"), "150", "11"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-4000:])
    assert "fuzz_io: 150 iterations" in r.stdout                                         # ASan / UBSan pass
    assert "fuzz_io: 15 iterations" in r.stdout or "TSan pass skipped" in r.stdout          # ThreadSanitizer pass
