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


def _crc32c_bytes(data, crc=0):
    c = crc ^ 0xFFFFFFFF
    for b in data:
        c = _TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


# Large buffers (checkpoint tensors): the buffer is cut into equal chunks whose CRCs advance in lockstep as one NumPy
# vector (one table gather per byte position), and the chunk CRCs are joined with the GF(2) "append L zero bytes"
# operator (the crc32_combine construction: crc(A || B) = shift(crc(A), len(B)) ^ crc(B)).
def _gf2_times(mat, vec):
    s, i = 0, 0
    while vec:
        if vec & 1:
            s ^= mat[i]
        vec >>= 1
        i += 1
    return s


def _gf2_compose(a, b):
    return [_gf2_times(a, col) for col in b]


_SHIFT_CACHE = {}


def _shift_operator(nbytes):
    """32 columns of the operator that maps crc(A) to crc(A || nbytes zero bytes) (finalised CRC values)."""
    if nbytes not in _SHIFT_CACHE:
        op = [0x82F63B78] + [1 << (n - 1) for n in range(1, 32)]  # one zero bit
        for _ in range(3):
            op = _gf2_compose(op, op)                             # -> one zero byte
        total, n = [1 << i for i in range(32)], nbytes
        while n:
            if n & 1:
                total = _gf2_compose(op, total)
            n >>= 1
            if n:
                op = _gf2_compose(op, op)
        _SHIFT_CACHE[nbytes] = total
    return _SHIFT_CACHE[nbytes]


_FAST_CHUNK = 4096


def crc32c(data: bytes, crc: int = 0) -> int:
    n = len(data)
    if crc != 0 or n < 64 * _FAST_CHUNK:
        return _crc32c_bytes(data, crc)
    import numpy as np
    L = _FAST_CHUNK
    chunks = n // L
    cols = np.frombuffer(data, dtype=np.uint8, count=chunks * L).reshape(chunks, L).T.copy()
    table = np.asarray(_TABLE, dtype=np.uint32)
    c = np.full(chunks, 0xFFFFFFFF, dtype=np.uint32)
    for j in range(L):
        c = table[(c ^ cols[j]) & 0xFF] ^ (c >> 8)
    parts = (c ^ np.uint32(0xFFFFFFFF)).tolist()
    op = _shift_operator(L)
    total = parts[0]
    for v in parts[1:]:
        total = _gf2_times(op, total) ^ v
    return _crc32c_bytes(memoryview(data)[chunks * L:], total)


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


def encode_example(features: dict, bytes_features=()) -> bytes:
    """features: name -> list/tuple of bytes | of ints | of floats (numpy arrays of int / float dtype work too);
    names in `bytes_features` are BytesList features even when their list is empty."""
    entries = b""
    for name in sorted(features):
        v = features[name]
        if isinstance(v, (bytes, str)):
            raise TypeError(f"feature {name!r} must be a list (reference: assert isinstance(value, Iterable))")
        v = list(v)
        if (v and isinstance(v[0], bytes)) or (not v and name in bytes_features):
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

class SourceData(namedtuple("SourceData", ["id", "key", "source", "source_length", "text", "speaker_id", "age", "gender",
                                           "phone", "phone_length", "phone_txt"])):
    """LJSpeech source records carry the first five fields (datasets/ljspeech/dataset.py:52-72); the VCTK / codes
    records add speaker_id, age, gender and the phone sequence (datasets/codes/dataset.py:66-96,
    datasets/vctk/dataset.py:64-77).  Absent features read as None."""

    def __new__(cls, id, key, source, source_length, text, speaker_id=None, age=None, gender=None, phone=None,
                phone_length=None, phone_txt=None):
        return super().__new__(cls, id, key, source, source_length, text, speaker_id, age, gender, phone, phone_length,
                               phone_txt)


TargetData = namedtuple("TargetData", ["id", "key", "mel", "mel_width", "target_length"])
# utils/tfrecord.py:20-22
CodeData = namedtuple("CodeData", ["id", "key", "lang", "codes", "codes_length", "codes_width"])
# utils/tfrecord.py:30-32 (+ the alignments the writer family stores, :176)
PredictionResult = namedtuple("PredictionResult", ["id", "key", "mel", "mel_length", "mel_width", "ground_truth_mel",
                                                   "ground_truth_mel_length", "text", "source", "source_length",
                                                   "alignment"])
CodePredictionResult = namedtuple("CodePredictionResult", ["id", "key", "codes", "codes_length", "codes_width",
                                                           "ground_truth_codes", "ground_truth_codes_length", "text",
                                                           "source", "source_length"])


def write_source_data(id_, key, source, text, filename, speaker_id=None, age=None, gender=None, phone=None,
                      phone_txt=None):
    """preprocess/ljspeech.py:35-44; with speaker_id / age / gender / phone the VCTK / codes form
    (datasets/codes/dataset.py:66-79)."""
    source = np.asarray(source, dtype=np.int64)
    feats = {"id": [int(id_)], "key": [key.encode("utf-8")], "source": [source.tobytes()],
             "source_length": [len(source)], "text": [text.encode("utf-8")]}
    if speaker_id is not None:
        feats.update(speaker_id=[int(speaker_id)], age=[int(age or 0)], gender=[int(gender or 0)])
    if phone is not None:
        phone = np.asarray(phone, dtype=np.int64)
        feats.update(phone=[phone.tobytes()], phone_length=[len(phone)], phone_txt=[(phone_txt or "").encode("utf-8")])
    write_records(filename, [encode_example(feats)])


def read_source_data(filename):
    """datasets/ljspeech/dataset.py:52-72 / datasets/codes/dataset.py:66-96 -> list of SourceData (ids as int64)."""
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        src = np.frombuffer(e["source"][0], dtype="<i8").copy()
        if len(src) != e["source_length"][0]:
            raise ValueError(f"{filename}: source_length {e['source_length'][0]} != {len(src)} ids")
        one = lambda k: e[k][0] if e.get(k) else None  # noqa: E731
        phone = np.frombuffer(e["phone"][0], dtype="<i8").copy() if e.get("phone") else None
        if phone is not None and len(phone) != one("phone_length"):
            raise ValueError(f"{filename}: phone_length {one('phone_length')} != {len(phone)} ids")
        out.append(SourceData(e["id"][0], e["key"][0].decode("utf-8"), src, e["source_length"][0],
                              e["text"][0].decode("utf-8"), one("speaker_id"), one("age"), one("gender"), phone,
                              one("phone_length"), e["phone_txt"][0].decode("utf-8") if e.get("phone_txt") else None))
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


def write_code_data(id_, key, lang, codes, filename):
    """Target record of the code models, the schema utils/tfrecord.py:62-113 parses: ``codes`` = raw float32
    [codes_length, codes_width] (one-hot rows), ``lang`` bytes."""
    codes = np.ascontiguousarray(codes, dtype="<f4")
    write_records(filename, [encode_example({
        "id": [int(id_)], "key": [key.encode("utf-8")], "lang": [lang.encode("utf-8")], "codes": [codes.tobytes()],
        "codes_length": [codes.shape[0]], "codes_width": [codes.shape[1]]})])


def read_code_data(filename):
    """utils/tfrecord.py:62-113 (parse_preprocessed_code_data + decode_preprocessed_code_data)."""
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        n, w = e["codes_length"][0], e["codes_width"][0]
        codes = np.frombuffer(e["codes"][0], dtype="<f4")
        if codes.size != n * w:
            raise ValueError(f"{filename}: codes has {codes.size} values, expected {n} x {w}")
        out.append(CodeData(e["id"][0], e["key"][0].decode("utf-8"), e["lang"][0].decode("utf-8"),
                            codes.reshape(n, w).copy(), n, w))
    return out


def write_prediction_result(id_, key, alignments, mel, ground_truth_mel, text, source, accent_type, filename):
    """predict_mel.py:72-74 calls write_prediction_result(id, key, alignments, mel, ground_truth_mel, text, source,
    accent_type, path); the record is the one utils/tfrecord.py:183-195 (parse_prediction_result) declares: mel
    / mel_length / mel_width, ground truth, ``alignment`` (bytes list, one raw float32 array per alignment like
    :176), text, source."""
    mel = np.ascontiguousarray(mel, dtype="<f4")
    gt = np.ascontiguousarray(ground_truth_mel if ground_truth_mel is not None else np.zeros((0, mel.shape[1])), dtype="<f4")
    source = np.asarray(source, dtype=np.int64)
    feats = {
        "id": [int(id_)], "key": [key.encode("utf-8")], "mel": [mel.tobytes()], "mel_length": [mel.shape[0]],
        "mel_width": [mel.shape[1]], "ground_truth_mel": [gt.tobytes()], "ground_truth_mel_length": [gt.shape[0]],
        "alignment": [np.ascontiguousarray(a, dtype="<f4").tobytes() for a in (alignments or []) if a is not None],
        "text": [text.encode("utf-8")], "source": [source.tobytes()], "source_length": [len(source)]}
    if accent_type is not None:
        feats["accent_type"] = [np.asarray(accent_type).tobytes()]
    write_records(filename, [encode_example(feats, bytes_features=("alignment",))])


def read_prediction_result(filename):
    """utils/tfrecord.py:183-215; ``alignment`` -> list of float32 arrays [source_length, -1] (memory x steps)."""
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        w = e["mel_width"][0]
        n_src = e["source_length"][0]
        aligns = []
        for a in e.get("alignment", []):
            a = np.frombuffer(a, dtype="<f4").copy()
            aligns.append(a.reshape(n_src, -1) if n_src and a.size % n_src == 0 else a)
        out.append(PredictionResult(
            e["id"][0], e["key"][0].decode("utf-8"),
            np.frombuffer(e["mel"][0], dtype="<f4").reshape(e["mel_length"][0], w).copy(), e["mel_length"][0], w,
            np.frombuffer(e["ground_truth_mel"][0], dtype="<f4").reshape(e["ground_truth_mel_length"][0], w).copy(),
            e["ground_truth_mel_length"][0], e["text"][0].decode("utf-8"),
            np.frombuffer(e["source"][0], dtype="<i8").copy(), n_src, aligns))
    return out


def write_code_prediction_result(id_, key, alignments, codes, ground_truth_codes, text, source, accent_type, filename):
    """utils/tfrecord.py:144-157 (the live writer of this fork, called by predict_code.py): the alignments and
    accent_type arguments are accepted and, as in the reference, not stored."""
    codes = np.ascontiguousarray(codes, dtype="<f4")
    gt = np.ascontiguousarray(ground_truth_codes if ground_truth_codes is not None else np.zeros((0, codes.shape[1])),
                              dtype="<f4")
    source = np.asarray(source, dtype=np.int64)
    write_records(filename, [encode_example({
        "id": [int(id_)], "key": [key.encode("utf-8")], "codes": [codes.tobytes()], "codes_length": [codes.shape[0]],
        "codes_width": [codes.shape[1]], "ground_truth_codes": [gt.tobytes()], "ground_truth_codes_length": [gt.shape[0]],
        "text": [text.encode("utf-8")], "source": [source.tobytes()], "source_length": [len(source)]})])


def read_code_prediction_result(filename):
    out = []
    for rec in read_records(filename):
        e = decode_example(rec)
        w = e["codes_width"][0]
        out.append(CodePredictionResult(
            e["id"][0], e["key"][0].decode("utf-8"),
            np.frombuffer(e["codes"][0], dtype="<f4").reshape(e["codes_length"][0], w).copy(), e["codes_length"][0], w,
            np.frombuffer(e["ground_truth_codes"][0], dtype="<f4").reshape(e["ground_truth_codes_length"][0], w).copy(),
            e["ground_truth_codes_length"][0], e["text"][0].decode("utf-8"),
            np.frombuffer(e["source"][0], dtype="<i8").copy(), e["source_length"][0]))
    return out
