"""TFRecord / tf.train.Example reader + writer (sat_b200.tfrecord) without TensorFlow: known-answer CRC-32C vectors,
a hand-derived serialized Example, record round trips in the reference's schemas, corruption detection."""
import os
import struct

import numpy as np
import pytest

from sat_b200 import tfrecord as tfr


def test_crc32c_known_answers():
    # RFC 3720 B.4 / the CRC-32C check value
    assert tfr.crc32c(b"123456789") == 0xE3069283
    assert tfr.crc32c(b"") == 0
    assert tfr.crc32c(bytes(32)) == 0x8A9136AA
    assert tfr.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert tfr.crc32c(bytes(range(32))) == 0x46DD794E
    assert tfr.crc32c(bytes(range(31, -1, -1))) == 0x113FDB5C
    # TFRecord mask: rotate right by 15, add 0xa282ead8
    c = 0xE3069283
    assert tfr.masked_crc32c(b"123456789") == (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def test_example_wire_format_known_answer():
    # Example{features{feature{key:"a" value{int64_list{value:[1]}}}}} serialized by hand from feature.proto / example.proto
    golden = bytes([0x0A, 0x0C, 0x0A, 0x0A, 0x0A, 0x01, 0x61, 0x12, 0x05, 0x1A, 0x03, 0x0A, 0x01, 0x01])
    assert tfr.encode_example({"a": [1]}) == golden
    assert tfr.decode_example(golden) == {"a": [1]}
    # unpacked (one varint per element) int64 lists, as older writers emit them, decode too
    unpacked = bytes([0x0A, 0x0C, 0x0A, 0x0A, 0x0A, 0x01, 0x61, 0x12, 0x05, 0x1A, 0x03, 0x08, 0x07, 0x08]) + bytes([0x09])
    unpacked = bytes([0x0A, 0x0D, 0x0A, 0x0B, 0x0A, 0x01, 0x61, 0x12, 0x06, 0x1A, 0x04, 0x08, 0x07, 0x08, 0x09])
    assert tfr.decode_example(unpacked) == {"a": [7, 9]}


def test_example_round_trip_all_kinds():
    ex = {"i": [0, 1, -1, 2 ** 40, -(2 ** 62)], "f": [0.5, -1.25, 3.0], "b": [b"", b"abc", bytes(range(256))], "e": []}
    got = tfr.decode_example(tfr.encode_example(ex))
    assert got["i"] == ex["i"] and got["b"] == ex["b"] and got["e"] == []
    np.testing.assert_array_equal(np.float32(got["f"]), np.float32(ex["f"]))
    with pytest.raises(TypeError):
        tfr.encode_example({"key": b"not-a-list"})


def test_record_framing_and_corruption(tmp_path):
    path = os.path.join(tmp_path, "x.tfrecord")
    payloads = [b"", b"a", os.urandom(1000)]
    tfr.write_records(path, payloads)
    assert list(tfr.read_records(path)) == payloads
    raw = bytearray(open(path, "rb").read())
    # layout of the first (empty) record: 8-byte length, 4-byte crc, 0 data bytes, 4-byte crc
    assert struct.unpack("<Q", raw[:8])[0] == 0 and len(raw) == 3 * 16 + 1 + 1000
    raw[-6] ^= 0x40
    open(path, "wb").write(raw)
    with pytest.raises(ValueError):
        list(tfr.read_records(path))
    open(path, "wb").write(raw[:-3])
    with pytest.raises(ValueError):
        list(tfr.read_records(path))


def test_reference_schemas_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    src = rng.integers(1, 71, size=37).astype(np.int64)
    mel = rng.normal(size=(123, 80)).astype(np.float32)
    ps, pt, pp = (os.path.join(tmp_path, n) for n in ("s.tfrecord", "t.tfrecord", "p.tfrecord"))
    tfr.write_source_data(7, "LJ001-0007", src, "printing, in the only sense", ps)
    tfr.write_target_data(7, "LJ001-0007", mel, pt)
    (s,), (t,) = tfr.read_source_data(ps), tfr.read_target_data(pt)
    assert (s.id, s.key, s.source_length, s.text) == (7, "LJ001-0007", 37, "printing, in the only sense")
    np.testing.assert_array_equal(s.source, src)
    assert (t.id, t.key, t.mel_width, t.target_length) == (7, "LJ001-0007", 80, 123)
    np.testing.assert_array_equal(t.mel, mel)
    tfr.write_prediction_result(7, "LJ001-0007", mel[:50], mel, s.text, src, pp)
    (p,) = tfr.read_prediction_result(pp)
    assert p.mel_length == 50 and p.ground_truth_mel_length == 123 and p.source_length == 37
    np.testing.assert_array_equal(p.mel, mel[:50])
    np.testing.assert_array_equal(p.ground_truth_mel, mel)
    # the field set of the source record is exactly what datasets/ljspeech/dataset.py:52-60 parses
    assert set(tfr.decode_example(next(tfr.read_records(ps)))) == {"id", "key", "source", "source_length", "text"}


def test_load_utterances_matches_sources_and_targets(tmp_path):
    from sat_b200.predict_mel import load_utterances
    rng = np.random.default_rng(1)
    files_s, files_t = [], []
    for i, key in enumerate(["a", "b", "c"]):
        fs, ft = os.path.join(tmp_path, f"{key}.source.tfrecord"), os.path.join(tmp_path, f"{key}.target.tfrecord")
        tfr.write_source_data(i, key, rng.integers(1, 71, size=10 + i), f"text {key}", fs)
        files_s.append(fs)
        if key != "b":
            tfr.write_target_data(i, key, rng.normal(size=(20 + i, 80)).astype(np.float32), ft)
            files_t.append(ft)
    utts = load_utterances(files_s, files_t)
    assert [u.key for u in utts] == ["a", "b", "c"] and [len(u.source) for u in utts] == [10, 11, 12]
    assert utts[0].mel.shape == (20, 80) and utts[1].mel is None and utts[2].mel.shape == (22, 80)
