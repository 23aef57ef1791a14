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
    al = rng.random(size=(37, 25)).astype(np.float32)
    tfr.write_prediction_result(7, "LJ001-0007", [al], mel[:50], mel, s.text, src, None, pp)
    (p,) = tfr.read_prediction_result(pp)
    assert p.mel_length == 50 and p.ground_truth_mel_length == 123 and p.source_length == 37
    np.testing.assert_array_equal(p.mel, mel[:50])
    np.testing.assert_array_equal(p.ground_truth_mel, mel)
    assert len(p.alignment) == 1
    np.testing.assert_array_equal(p.alignment[0], al)
    # every feature utils/tfrecord.py:183-195 (parse_prediction_result) declares as a required FixedLenFeature is
    # present, each with exactly one value (one alignment: the single-source mel models predict_mel.py serves)
    e = tfr.decode_example(next(tfr.read_records(pp)))
    required = {"id", "key", "mel", "mel_length", "mel_width", "ground_truth_mel", "ground_truth_mel_length",
                "alignment", "text", "source", "source_length"}
    assert required <= set(e) and all(len(e[k]) == 1 for k in required)
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


def test_code_record_known_answer_and_round_trips(tmp_path):
    """Code target / prediction records (utils/tfrecord.py:62-113,144-157).  The serialized Example of a 1 x 2 code
    record is derived by hand from example.proto / feature.proto: map entries in key order, BytesList = field 1,
    Int64List = field 3 (packed)."""
    codes = np.array([[0.0, 1.0]], dtype="<f4")

    def entry(name, feat):
        body = bytes([0x0A, len(name)]) + name + bytes([0x12, len(feat)]) + feat
        return bytes([0x0A, len(body)]) + body

    def blist(b):
        inner = bytes([0x0A, len(b)]) + b
        return bytes([0x0A, len(inner)]) + inner

    def ilist(v):
        return bytes([0x1A, 0x03, 0x0A, 0x01, v])

    feats = (entry(b"codes", blist(codes.tobytes())) + entry(b"codes_length", ilist(1)) + entry(b"codes_width", ilist(2)) +
             entry(b"id", ilist(5)) + entry(b"key", blist(b"k")) + entry(b"lang", blist(b"en")))
    golden = bytes([0x0A, len(feats)]) + feats
    path = os.path.join(tmp_path, "c.tfrecord")
    tfr.write_code_data(5, "k", "en", codes, path)
    assert next(tfr.read_records(path)) == golden
    (c,) = tfr.read_code_data(path)
    assert (c.id, c.key, c.lang, c.codes_length, c.codes_width) == (5, "k", "en", 1, 2)
    np.testing.assert_array_equal(c.codes, codes)
    # prediction record of the code models: exactly the ten features of utils/tfrecord.py:146-155
    rng = np.random.default_rng(3)
    pred = np.eye(6, dtype=np.float32)[rng.integers(0, 6, size=9)]
    gt = np.eye(6, dtype=np.float32)[rng.integers(0, 6, size=11)]
    src = rng.integers(1, 71, size=13).astype(np.int64)
    pp = os.path.join(tmp_path, "p.tfrecord")
    tfr.write_code_prediction_result(3, "utt", [np.zeros((13, 9), np.float32)], pred, gt, "txt", src, None, pp)
    assert set(tfr.decode_example(next(tfr.read_records(pp)))) == {
        "id", "key", "codes", "codes_length", "codes_width", "ground_truth_codes", "ground_truth_codes_length", "text",
        "source", "source_length"}
    (r,) = tfr.read_code_prediction_result(pp)
    assert (r.id, r.key, r.codes_length, r.codes_width, r.ground_truth_codes_length, r.text) == (3, "utt", 9, 6, 11, "txt")
    np.testing.assert_array_equal(r.codes, pred)
    np.testing.assert_array_equal(r.ground_truth_codes, gt)
    np.testing.assert_array_equal(r.source, src)


def test_vctk_source_record_speaker_and_phone(tmp_path):
    """VCTK / codes source records (datasets/codes/dataset.py:66-96): speaker_id, age, gender, phone sequence; with
    hparams.source == 'phone' the phone ids are the model input (:147-149)."""
    import sat_b200
    from sat_b200.predict_mel import load_utterances, selected_files
    rng = np.random.default_rng(4)
    hp = sat_b200.config("vctk_self_attention")
    root = str(tmp_path)
    keys = ["p225_001", "p226_002"]
    for i, k in enumerate(keys):
        tfr.write_source_data(i, k, rng.integers(1, 71, size=12 + i), "hello", os.path.join(root, f"{k}.{hp.source_file_extension}"),
                              speaker_id=225 + i, age=23, gender=1, phone=rng.integers(1, 60, size=7 + i), phone_txt="HH AH")
    open(os.path.join(root, "test.csv"), "w").write("\n".join(keys) + "\n")
    src, tgt = selected_files(hp, root, root, root, "test.csv")
    assert [os.path.basename(f) for f in src] == [f"{k}.source.tfrecord" for k in keys] and len(tgt) == 2
    (s0,) = tfr.read_source_data(src[0])
    assert (s0.speaker_id, s0.age, s0.gender, s0.phone_length, s0.phone_txt) == (225, 23, 1, 7, "HH AH")
    utts = load_utterances(src, tgt, hp)   # hp.source == 'phone' by default: the phone ids feed the model
    assert [len(u.source) for u in utts] == [7, 8] and [u.speaker_id for u in utts] == [225, 226]
    hp.override(source="char")
    assert [len(u.source) for u in load_utterances(src, tgt, hp)] == [12, 13]
