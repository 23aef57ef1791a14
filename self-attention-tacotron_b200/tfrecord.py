"""TFRecord / tf.train.Example files either side of the synthesis path, without TensorFlow.

What the reference reads and writes around ``predict_mel`` (all through ``tf.python_io.TFRecordWriter`` /
``tf.data.TFRecordDataset``):

* source records   -- preprocess/ljspeech.py:35-44, parsed by datasets/ljspeech/dataset.py:52-72:
  ``id`` int64, ``key`` bytes, ``source`` bytes (raw int64 ids), ``source_length`` int64, ``text`` bytes;
* target records   -- preprocess/ljspeech.py:23-32: ``id``, ``key``, ``mel`` bytes (raw float32 [T, mel_width]),
  ``target_length``, ``mel_width``;
* prediction records -- utils/tfrecord.py (``write_prediction_result`` family): id, key, mel / codes bytes with length
  and width, ground truth, text, source.

File framing (TFRecord): ``uint64 length | uint32 masked_crc32c(length) | data | uint32 masked_crc32c(data)``, little
endian, CRC-32C (Castagnoli) masked as ``((crc >> 15 | crc << 17) + 0xa282ead8) mod 2^32``.  ``data`` is a serialized
``tf.train.Example``: ``Example{1: Features{1: map<string, Feature>}}``, ``Feature`` = oneof ``BytesList`` (1) /
``FloatList`` (2, packed floats) / ``Int64List`` (3, packed varints).

This is host-side I/O: it never touches the GPU path and has no CPU stand-in for any kernel.
"""
import struct
from collections import namedtuple

import numpy as np

# ---- CRC-32C -------------------------------------------------------------------------------------------------------

_POLY = 0x82F63B78


def _make_table():
    t = []
    for n in range(256):
        c = n
        for _ in range(8):
            c = (c >> 1) ^ _POLY if c & 1 else c >> 1
        t.append(c)
    return t


_TABLE = _make_table()


def crc32c(data: bytes, crc: int = 0) -> int:
    c = crc ^ 0xFFFFFFFF
    for b in data:
        c = _TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


# ---- record framing ---------------------------------------------------------------------------------------------------

def write_records(filename, payloads):
    """Write an iterable of serialized records to a TFRecord file."""
    with open(filename, "wb") as f:
        for p in payloads:
            head = struct.pack("<Q", len(p))
            f.write(head)
            f.write(struct.pack("<I", masked_crc32c(head)))
            f.write(p)
            f.write(struct.pack("<I", masked_crc32c(p)))


def read_records(filename, check_crc=True):
    """Yield the serialized records of a TFRecord file; raises ValueError on a corrupt or truncated file."""
    with open(filename, "rb") as f:
        while True:
            head = f.read(8)
            if not head:
                return
            if len(head) != 8:
                raise ValueError(f"{filename}: truncated record header")
            crc = f.read(4)
            if len(crc) != 4 or (check_crc and struct.unpack("<I", crc)[0] != masked_crc32c(head)):
                raise ValueError(f"{filename}: corrupt record length")
            (n,) = struct.unpack("<Q", head)
            data = f.read(n)
            crc = f.read(4)
            if len(data) != n or len(crc) != 4:
                raise ValueError(f"{filename}: truncated record")
            if check_crc and struct.unpack("<I", crc)[0] != masked_crc32c(data):
                raise ValueError(f"{filename}: corrupt record data")
            yield data


# ---- minimal protobuf wire format -------------------------------------------------------------------------------------

def _varint(n: int) -> bytes:
    n &= 0xFFFFFFFFFFFFFFFF  # int64 values are written as 64-bit two's complement
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos):
    shift = result = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 70:
            raise ValueError("varint too long")


def _ld(field: int, payload: bytes) -> bytes:  # length-delimited field
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def _fields(buf):
    """Yield (field number, wire type, value) of one message; value is int for varint / fixed, bytes for wire type 2."""
    pos = 0
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 1:
            v, pos = buf[pos:pos + 8], pos + 8
        elif wt == 2:
            n, pos = _read_varint(buf, pos)
            v, pos = buf[pos:pos + n], pos + n
            if len(v) != n:
                raise ValueError("truncated field")
        elif wt == 5:
            v, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield field, wt, v


def encode_example(features: dict) -> bytes:
    """features: name -> list/tuple of bytes | of ints | of floats (numpy arrays of int / float dtype work too)."""
    entries = b""
    for name in sorted(features):
        v = features[name]
        if isinstance(v, (bytes, str)):
            raise TypeError(f"feature {name!r} must be a list (reference: assert isinstance(value, Iterable))")
        v = list(v)
        if v and isinstance(v[0], bytes):
            feat = _ld(1, b"".join(_ld(1, x) for x in v))
        elif v and isinstance(v[0], (float, np.floating)):
            feat = _ld(2, _ld(1, struct.pack(f"<{len(v)}f", *v)))
        else:
            feat = _ld(3, _ld(1, b"".join(_varint(int(x)) for x in v)) if v else b"")
        entries += _ld(1, _ld(1, name.encode("utf-8")) + _ld(2, feat))
    return _ld(1, entries)


def decode_example(buf: bytes) -> dict:
    """-> name -> list of bytes / ints / floats."""
    out = {}
    for f1, _, features in _fields(buf):
        if f1 != 1:
            continue
        for f2, _, entry in _fields(features):
            if f2 != 1:
                continue
            name, feat = None, b""
            for f3, _, v in _fields(entry):
                if f3 == 1:
                    name = bytes(v).decode("utf-8")
                elif f3 == 2:
                    feat = v
            values = []
            for kind, _, lst in _fields(feat):
                for f4, wt, v in _fields(lst):
                    if f4 != 1:
                        continue
                    if kind == 1:
                        values.append(bytes(v))
                    elif kind == 2:
                        if wt == 2:
                            values.extend(struct.unpack(f"<{len(v) // 4}f", v))
                        else:
                            values.append(struct.unpack("<f", v)[0])
                    elif kind == 3:
                        if wt == 2:
                            pos = 0
                            while pos < len(v):
                                x, pos = _read_varint(v, pos)
                                values.append(x - (1 << 64) if x >> 63 else x)
                        else:
                            values.append(v - (1 << 64) if v >> 63 else v)
            out[name] = values
    return out


# ---- the reference's record schemas -----------------------------------------------------------------------------------

SourceData = namedtuple("SourceData", ["id", "key", "source", "source_length", "text"])
TargetData = namedtuple("TargetData", ["id", "key", "mel", "mel_width", "target_length"])
PredictionResult = namedtuple("PredictionResult", ["id", "key", "mel", "mel_length", "mel_width", "ground_truth_mel",
                                                   "ground_truth_mel_length", "text", "source", "source_length"])


def write_source_data(id_, key, source, text, filename):
    """preprocess/ljspeech.py:35-44"""
    source = np.asarray(source, dtype=np.int64)
    write_records(filename, [encode_example({
        "id": [int(id_)], "key": [key.encode("utf-8")], "source": [source.tobytes()],
        "source_length": [len(source)], "text": [text.encode("utf-8")]})])


def read_source_data(filename):
    """datasets/ljspeech/dataset.py:52-72 -> list of SourceData (source as int64 ndarray)."""
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        src = np.frombuffer(e["source"][0], dtype="<i8").copy()
        if len(src) != e["source_length"][0]:
            raise ValueError(f"{filename}: source_length {e['source_length'][0]} != {len(src)} ids")
        out.append(SourceData(e["id"][0], e["key"][0].decode("utf-8"), src, e["source_length"][0],
                              e["text"][0].decode("utf-8")))
    return out


def write_target_data(id_, key, mel, filename):
    """preprocess/ljspeech.py:23-32"""
    mel = np.ascontiguousarray(mel, dtype="<f4")
    write_records(filename, [encode_example({
        "id": [int(id_)], "key": [key.encode("utf-8")], "mel": [mel.tobytes()],
        "target_length": [mel.shape[0]], "mel_width": [mel.shape[1]]})])


def read_target_data(filename):
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        w, n = e["mel_width"][0], e["target_length"][0]
        mel = np.frombuffer(e["mel"][0], dtype="<f4").reshape(n, w).copy()
        out.append(TargetData(e["id"][0], e["key"][0].decode("utf-8"), mel, w, n))
    return out


def write_prediction_result(id_, key, mel, ground_truth_mel, text, source, filename):
    """The mel twin of utils/tfrecord.py write_prediction_result (same field naming scheme, mel instead of codes)."""
    mel = np.ascontiguousarray(mel, dtype="<f4")
    gt = np.ascontiguousarray(ground_truth_mel if ground_truth_mel is not None else np.zeros((0, mel.shape[1])), dtype="<f4")
    source = np.asarray(source, dtype=np.int64)
    write_records(filename, [encode_example({
        "id": [int(id_)], "key": [key.encode("utf-8")], "mel": [mel.tobytes()], "mel_length": [mel.shape[0]],
        "mel_width": [mel.shape[1]], "ground_truth_mel": [gt.tobytes()], "ground_truth_mel_length": [gt.shape[0]],
        "text": [text.encode("utf-8")], "source": [source.tobytes()], "source_length": [len(source)]})])


def read_prediction_result(filename):
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        w = e["mel_width"][0]
        out.append(PredictionResult(
            e["id"][0], e["key"][0].decode("utf-8"),
            np.frombuffer(e["mel"][0], dtype="<f4").reshape(e["mel_length"][0], w).copy(), e["mel_length"][0], w,
            np.frombuffer(e["ground_truth_mel"][0], dtype="<f4").reshape(e["ground_truth_mel_length"][0], w).copy(),
            e["ground_truth_mel_length"][0], e["text"][0].decode("utf-8"),
            np.frombuffer(e["source"][0], dtype="<i8").copy(), e["source_length"][0]))
    return out
