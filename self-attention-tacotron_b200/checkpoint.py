"""TensorFlow checkpoint (V2 "tensor bundle") reader / writer and the checkpoint -> weight-blob importer, without
TensorFlow.

The reference restores ``--checkpoint`` (or the latest one in ``--checkpoint-dir``) through ``tf.estimator``
(/root/reference/predict_mel.py:47-54, train.py:69-77).  A checkpoint ``<prefix>`` is

* ``<prefix>.index`` -- a LevelDB-format sorted string table: key ``""`` -> ``BundleHeaderProto`` (num_shards=1,
  endianness=2, version=3), key ``<variable name>`` -> ``BundleEntryProto`` (dtype=1, shape=2, shard_id=3, offset=4,
  size=5, crc32c=6 fixed32, slices=7);
* ``<prefix>.data-SSSSS-of-NNNNN`` -- the raw little-endian tensor bytes at those offsets;
* ``checkpoint`` in the same directory -- a text proto naming ``model_checkpoint_path``.

Table format: data blocks of prefix-compressed entries (``varint shared | varint non_shared | varint value_len |
key suffix | value``) with a restart array (``uint32[] offsets, uint32 count``); every block is followed by a 1-byte
compression type (0: none -- what TensorFlow's BundleWriter uses) and a masked CRC-32C of block + type; an index block
maps a separator key to each data block's ``BlockHandle`` (varint offset, varint size); the 48-byte footer holds the
metaindex and index handles padded to 40 bytes and the magic ``0xdb4775248b80fb57``.

Parity note: there is no TensorFlow in this image and the reference ships no checkpoint, so the format code is pinned
only by its own write -> read round trip and by hand-built known-answer bytes (tests/test_checkpoint_cpu.py), and the
variable-name rules below follow the names visible in the reference tree -- names created inside the
absent ``tacotron2`` package are matched by suffix and shape, with an explicit ``name_map`` override for the rest.

Host-side I/O only: nothing here touches the GPU path or stands in for a kernel."""
import os
import re
import struct

import numpy as np

from .tfrecord import _varint, _read_varint, _fields, crc32c

_MAGIC = 0xDB4775248B80FB57
# tensorflow/core/framework/types.proto
_DTYPES = {1: np.dtype("<f4"), 2: np.dtype("<f8"), 3: np.dtype("<i4"), 4: np.dtype("u1"), 5: np.dtype("<i2"),
           6: np.dtype("i1"), 9: np.dtype("<i8"), 10: np.dtype("?"), 17: np.dtype("<u2"), 19: np.dtype("<f2"),
           22: np.dtype("<u4"), 23: np.dtype("<u8")}
_DTYPE_IDS = {v: k for k, v in _DTYPES.items()}


def _mask(c):
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


# ---- LevelDB-format table ----------------------------------------------------------------------------------------------

def _read_block(buf, offset, size, what):
    if offset + size + 5 > len(buf):
        raise ValueError(f"{what}: block handle out of range")
    body, kind = buf[offset:offset + size], buf[offset + size]
    (crc,) = struct.unpack_from("<I", buf, offset + size + 1)
    if crc != _mask(crc32c(buf[offset:offset + size + 1])):
        raise ValueError(f"{what}: block checksum mismatch")
    if kind != 0:
        raise NotImplementedError(f"{what}: compressed table block (type {kind}); TensorFlow bundles are uncompressed")
    return body


def _block_entries(block):
    """(key, value) pairs of one table block."""
    if len(block) < 4:
        raise ValueError("table block too short")
    (n_restarts,) = struct.unpack_from("<I", block, len(block) - 4)
    end = len(block) - 4 - 4 * n_restarts
    if end < 0:
        raise ValueError("bad restart array")
    pos, key = 0, b""
    while pos < end:
        shared, pos = _read_varint(block, pos)
        non_shared, pos = _read_varint(block, pos)
        vlen, pos = _read_varint(block, pos)
        if shared > len(key) or pos + non_shared + vlen > end:
            raise ValueError("corrupt table entry")
        key = key[:shared] + bytes(block[pos:pos + non_shared])
        pos += non_shared
        yield key, bytes(block[pos:pos + vlen])
        pos += vlen


def _read_table(path):
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != _MAGIC:
        raise ValueError(f"{path}: not a TensorFlow checkpoint index (bad table magic)")
    foot = buf[len(buf) - 48:]
    pos = 0
    _, pos = _read_varint(foot, pos)  # metaindex handle
    _, pos = _read_varint(foot, pos)
    ioff, pos = _read_varint(foot, pos)
    isize, pos = _read_varint(foot, pos)
    out = []
    for _, handle in _block_entries(_read_block(buf, ioff, isize, path)):
        off, p = _read_varint(handle, 0)
        size, p = _read_varint(handle, p)
        out.extend(_block_entries(_read_block(buf, off, size, path)))
    return out


def _build_block(entries, restart_interval):
    out, restarts, prev = bytearray(), [], b""
    for i, (k, v) in enumerate(entries):
        shared = 0
        if i % restart_interval == 0:
            restarts.append(len(out))
        else:
            while shared < min(len(prev), len(k)) and prev[shared] == k[shared]:
                shared += 1
        out += _varint(shared) + _varint(len(k) - shared) + _varint(len(v)) + k[shared:] + v
        prev = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        out += struct.pack("<I", r)
    out += struct.pack("<I", len(restarts))
    return bytes(out)


def _write_table(path, entries, block_bytes=4096):
    """entries: sorted list of (key bytes, value bytes)."""
    with open(path, "wb") as f:
        pos = 0

        def emit(block):
            nonlocal pos
            trailer = b"\x00"
            f.write(block + trailer + struct.pack("<I", _mask(crc32c(block + trailer))))
            handle = _varint(pos) + _varint(len(block))
            pos += len(block) + 5
            return handle

        index, cur, cur_size = [], [], 0
        for k, v in entries:
            cur.append((k, v))
            cur_size += len(k) + len(v) + 3
            if cur_size >= block_bytes:
                index.append((cur[-1][0], emit(_build_block(cur, 16))))
                cur, cur_size = [], 0
        if cur:
            index.append((cur[-1][0], emit(_build_block(cur, 16))))
        meta = emit(_build_block([], 1))
        idx = emit(_build_block(index, 1))
        foot = meta + idx
        f.write(foot + b"\x00" * (40 - len(foot)) + struct.pack("<Q", _MAGIC))


# ---- bundle protos -----------------------------------------------------------------------------------------------------

def _parse_entry(value):
    e = dict(dtype=0, shape=[], shard_id=0, offset=0, size=0, crc32c=None, sliced=False)
    for num, wt, v in _fields(value):
        if num == 1:
            e["dtype"] = v
        elif num == 2:
            for n2, _, v2 in _fields(v):
                if n2 == 2:  # TensorShapeProto.dim
                    size = 0
                    for n3, _, v3 in _fields(v2):
                        if n3 == 1:
                            size = v3 - (1 << 64) if v3 >= (1 << 63) else v3
                    e["shape"].append(size)
        elif num == 3:
            e["shard_id"] = v
        elif num == 4:
            e["offset"] = v
        elif num == 5:
            e["size"] = v
        elif num == 6:
            e["crc32c"] = v if isinstance(v, int) else struct.unpack("<I", v)[0]
        elif num == 7:
            e["sliced"] = True
    return e


def _entry_proto(dtype_id, shape, offset, size, crc):
    dims = b"".join(b"\x12" + _varint(len(d)) + d for d in (b"\x08" + _varint(int(s)) for s in shape))
    out = b"\x08" + _varint(dtype_id) + b"\x12" + _varint(len(dims)) + dims
    if offset:
        out += b"\x20" + _varint(offset)
    out += b"\x28" + _varint(size) + b"\x35" + struct.pack("<I", crc)
    return out


def checkpoint_variables(prefix):
    """-> {name: (dtype, shape)} of a checkpoint, like ``tf.train.list_variables``."""
    out = {}
    for k, v in _read_table(prefix + ".index"):
        if k:
            e = _parse_entry(v)
            out[k.decode("utf-8")] = (_DTYPES.get(e["dtype"]), tuple(e["shape"]))
    return out


def read_checkpoint(prefix, names=None, verify_crc_below=1 << 20):
    """Read the variables of checkpoint ``prefix`` -> {name: ndarray}.  ``names``: optional subset.  Tensor checksums
    are verified for tensors smaller than ``verify_crc_below`` bytes (pure-Python CRC-32C; None = verify all)."""
    table = _read_table(prefix + ".index")
    header = dict(num_shards=1, endianness=0)
    if table and table[0][0] == b"":
        for num, _, v in _fields(table[0][1]):
            if num == 1:
                header["num_shards"] = v
            elif num == 2:
                header["endianness"] = v
    if header["endianness"] != 0:
        raise NotImplementedError("big-endian checkpoint")
    shards, out = {}, {}
    for k, v in table:
        if not k:
            continue
        name = k.decode("utf-8")
        if names is not None and name not in names:
            continue
        e = _parse_entry(v)
        if e["sliced"]:
            raise NotImplementedError(f"{name}: partitioned (sliced) variable")
        if e["dtype"] not in _DTYPES:
            raise NotImplementedError(f"{name}: unsupported dtype enum {e['dtype']}")
        dt = _DTYPES[e["dtype"]]
        if e["shard_id"] not in shards:
            path = f"{prefix}.data-{e['shard_id']:05d}-of-{header['num_shards']:05d}"
            shards[e["shard_id"]] = np.memmap(path, dtype=np.uint8, mode="r") if os.path.getsize(path) else np.zeros(0, np.uint8)
        raw = shards[e["shard_id"]][e["offset"]:e["offset"] + e["size"]]
        n = int(np.prod(e["shape"])) if e["shape"] else 1
        if len(raw) != e["size"] or n * dt.itemsize != e["size"]:
            raise ValueError(f"{name}: size {e['size']} does not match shape {e['shape']} / data file")
        data = bytes(raw)
        if e["crc32c"] is not None and (verify_crc_below is None or e["size"] < verify_crc_below):
            if _mask(crc32c(data)) != e["crc32c"]:
                raise ValueError(f"{name}: tensor checksum mismatch")
        out[name] = np.frombuffer(data, dtype=dt).reshape(e["shape"]).copy()
    if names is not None:
        missing = sorted(set(names) - set(out))
        if missing:
            raise KeyError(f"not in checkpoint {prefix}: {missing}")
    return out


def write_checkpoint(prefix, tensors, update_state_file=True):
    """Write {name: ndarray} as a single-shard V2 checkpoint (``<prefix>.index`` + ``<prefix>.data-00000-of-00001``)
    and, like ``tf.train.Saver``, point the directory's ``checkpoint`` file at it."""
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    entries, offset = [(b"", b"\x08\x01\x1a\x02\x08\x01")], 0  # num_shards=1, endianness=LITTLE(0), version{producer=1}
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        for name in sorted(tensors, key=lambda s: s.encode("utf-8")):
            a = np.asarray(tensors[name])
            dt = a.dtype.newbyteorder("<") if a.dtype.byteorder == ">" else a.dtype
            if np.dtype(dt) not in _DTYPE_IDS:
                raise TypeError(f"{name}: dtype {a.dtype} has no checkpoint encoding here")
            data = a.astype(dt, copy=False).tobytes()
            f.write(data)
            entries.append((name.encode("utf-8"),
                            _entry_proto(_DTYPE_IDS[np.dtype(dt)], a.shape, offset, len(data), _mask(crc32c(data)))))
            offset += len(data)
    _write_table(prefix + ".index", entries)
    if update_state_file:
        base = os.path.basename(prefix)
        with open(os.path.join(os.path.dirname(os.path.abspath(prefix)), "checkpoint"), "w") as f:
            f.write(f'model_checkpoint_path: "{base}"\nall_model_checkpoint_paths: "{base}"\n')


def latest_checkpoint(checkpoint_dir):
    """``tf.train.latest_checkpoint``: the prefix named by ``<dir>/checkpoint`` (predict_mel.py:47-54), or None."""
    path = os.path.join(checkpoint_dir, "checkpoint")
    if not os.path.exists(path):
        return None
    with open(path, encoding="utf-8") as f:
        m = re.search(r'^model_checkpoint_path:\s*"(.*)"\s*$', f.read(), re.M)
    if not m:
        return None
    p = m.group(1)
    return p if os.path.isabs(p) else os.path.join(checkpoint_dir, p)


# ---- checkpoint variable names -> weight-blob names ------------------------------------------------------------------------
# (pattern on the TensorFlow variable name, weight-blob name template).  In-tree names come from SURVEY.md A.9:
# conv1d_K{k} / proj1 / proj2 (modules/module.py:51,60,67), attention_variable / attention_bias /
# location_features_convolution / location_features_layer / transition_factor_projection (modules/
# forward_attention.py:17-86), memory_layer / query_layer / attention_v (TF BahdanauAttention), out_projection /
# stop_token_projection under decoder/ (modules/module.py:630-637,716-723), batch_normalization gamma / beta /
# moving_mean / moving_variance, lstm_cell kernel / bias.  ``{i}`` groups are renumbered in sorted order.
_BN = {"gamma": "gamma", "beta": "beta", "moving_mean": "mean", "moving_variance": "var"}
_RULES = [
    (r"(^|/)embedding/(embedding|embeddings|kernel|table)$", "embedding"),
    (r"speaker_embedding/(embedding|embeddings|kernel|table)$", "speaker_embedding"),
    (r"conv1d_K(\d+)/.*conv1d[^/]*/kernel$", "encoder/cbhg/conv1d_K{0}/kernel"),
    (r"conv1d_K(\d+)/.*batch_normalization[^/]*/(gamma|beta|moving_mean|moving_variance)$", "encoder/cbhg/conv1d_K{0}/bn/{bn1}"),
    (r"cbhg.*/(proj[12])/.*conv1d[^/]*/kernel$", "encoder/cbhg/{0}/kernel"),
    (r"cbhg.*/(proj[12])/.*batch_normalization[^/]*/(gamma|beta|moving_mean|moving_variance)$", "encoder/cbhg/{0}/bn/{bn1}"),
    (r"out_projection/(kernel|bias)$", "decoder/out_projection/{0}"),
    (r"stop_token_projection/(kernel|bias)$", "decoder/stop_token_projection/{0}"),
]


def default_name_map(names):
    """Best-effort {checkpoint variable name: weight-blob name} for the names that the rule table recognises.
    Optimiser slots (``/Adam``, ``/Adam_1``), ``global_step`` and ``beta*_power`` are skipped."""
    out = {}
    for n in names:
        if re.search(r"/Adam(_\d+)?$|^global_step$|beta\d_power$", n):
            continue
        for pat, tmpl in _RULES:
            m = re.search(pat, n)
            if m:
                g = list(m.groups())
                g = [x for x in g if x not in ("", "/", None)] if pat.startswith("(^|/)") else g
                name = tmpl
                if "{bn1}" in name:
                    name = name.replace("{bn1}", _BN[g[1]])
                out[n] = name.format(*g)
                break
    return out


# ---- names of the variables the reference creates, reconstructed ----------------------------------------------------------
# TF 1.x naming, applied to the constructors of the reference (citations into /root/reference):
# * a tf.layers.Layer without a name opens variable_scope(None, default_name=snake_case(class)) at its first call, so
#   siblings are numbered in CALL order inside their enclosing scope: dense, dense_1, ... (MultiHeadAttention calls key,
#   value, query, output: modules/self_attention.py:103-119; MultiSpeakerPreNet calls dense0, speaker_projection, dense:
#   modules/multi_speaker_modules.py:27-30; the encoder calls pre_net, pre_net_1, zoneout_cbhg, dense,
#   self_attention_transformer*: modules/module.py:425-438);
# * explicit names are kept: conv1d_K{k} / proj1 / proj2 (modules/module.py:51,60,67), memory_layer / query_layer /
#   attention_v (TF BahdanauAttention), attention_variable / attention_bias / location_features_convolution /
#   location_features_layer / transition_factor_projection inside scope location_sensitive_attention
#   (modules/forward_attention.py:17-24,73-90), out_projection / stop_token_projection under decoder/
#   (modules/module.py:716-723);
# * tf.layers.batch_normalization -> batch_normalization/{gamma,beta,moving_mean,moving_variance}; LSTMCell ->
#   lstm_cell/{kernel,bias} (blacklist fragments of models/models.py:165-168); bidirectional_dynamic_rnn ->
#   bidirectional_rnn/{fw,bw}/...; MultiRNNCell -> multi_rnn_cell/cell_{i}/...
# Names created inside the absent tacotron2 package (Embedding, PreNet, HighwayNet H / T, ZoneoutLSTMCell, AttentionRNN,
# DecoderRNNV2, OutputAndStopTokenWrapper) are L-confidence; `resolve_names` therefore matches by FRAGMENT + SHAPE, not
# by full path, and this function is the single place that states the assumed full paths (used to test the resolver).
def tf_variable_names(hparams):
    """{weight-blob name: reconstructed TensorFlow variable name} for every weight of the model."""
    from .weights import weight_shapes
    hp = hparams
    enc = "self_attention_cbhg_encoder" if hp.encoder == "SelfAttentionCBHGEncoder" else "zoneout_encoder_v1"
    dec = {"ExtendedDecoder": "extended_decoder", "DualSourceDecoder": "dual_source_decoder",
           "DualSourceTransformerDecoder": "dual_source_transformer_decoder"}[hp.decoder]
    cbhg = f"{enc}/zoneout_cbhg"
    num = lambda base, i: base if i == 0 else f"{base}_{i}"  # noqa: E731
    bn = {"gamma": "gamma", "beta": "beta", "mean": "moving_mean", "var": "moving_variance"}
    mha = ["key", "value", "query", "output"]
    ms = hp.use_speaker_embedding and hp.speaker_embedd_to_prenet
    dual = hp.decoder != "ExtendedDecoder"
    cell = f"{dec}/decoder/" + ("dual_source_attention_rnn" if dual else "attention_rnn")
    kinds = []
    for i in range(2 if dual else 1):
        a = hp.attention if i == 0 else hp.attention2
        kinds.append("location_sensitive_attention" if a in ("forward", "location_sensitive") or a.startswith("teacher_forcing_forward")
                     else "bahdanau_attention")
    out = {}
    for k in weight_shapes(hp):
        q = k.split("/")
        if k == "embedding":
            n = "embedding/embedding"
        elif k == "speaker_embedding":
            n = "embedding_1/embedding"
        elif q[0] == "speaker_embedding_projection":
            n = f"dense/{q[1]}"
        elif q[:2] == ["encoder", "prenet"]:
            n = f"{enc}/{num('pre_net', int(q[2].split('_')[1]))}/dense/{q[3]}"
        elif q[:2] == ["encoder", "cbhg"] and (q[2].startswith("conv1d_K") or q[2] in ("proj1", "proj2")):
            n = f"{cbhg}/{q[2]}/conv1d/kernel" if q[3] == "kernel" else f"{cbhg}/{q[2]}/batch_normalization/{bn[q[4]]}"
        elif q[:3] == ["encoder", "cbhg", "adjustment"]:
            n = f"{cbhg}/dense/{q[3]}"
        elif q[:2] == ["encoder", "cbhg"] and q[2].startswith("highway_"):
            n = f"{cbhg}/{num('highway_net', int(q[2].split('_')[1]))}/{q[3]}/{q[4]}"
        elif q[:2] == ["encoder", "cbhg"] and q[2].startswith("lstm_"):
            n = f"{cbhg}/bidirectional_rnn/{q[2][5:]}/zoneout_lstm_cell/lstm_cell/{q[3]}"
        elif q[:2] == ["encoder", "sa_projection"]:
            n = f"{enc}/dense/{q[2]}"
        elif q[0] in ("encoder", "decoder") and q[1].startswith("sa_"):
            scope = (enc if q[0] == "encoder" else f"{dec}/decoder") + "/" + num("self_attention_transformer", int(q[1][3:]))
            n = (f"{scope}/self_attention/multi_head_attention/{num('dense', mha.index(q[3]))}/{q[4]}" if q[2] == "mha"
                 else f"{scope}/dense/{q[3]}")
        elif q[:3] == ["decoder", "prenet", "ms"]:
            n = f"{cell}/multi_speaker_pre_net/{num('dense', ['dense0', 'speaker_projection', 'dense'].index(q[3]))}/{q[4]}"
        elif q[:2] == ["decoder", "prenet"]:
            i = int(q[2].split("_")[1])
            n = f"{cell}/{num('pre_net', i - 1 if ms else i)}/dense/{q[3]}"
        elif q[:2] == ["decoder", "attention_lstm"]:
            n = f"{cell}/attention_wrapper/zoneout_lstm_cell/lstm_cell/{q[2]}"
        elif q[0] == "decoder" and q[1] in ("lstm_1", "lstm_2"):
            n = f"{dec}/decoder/decoder_rnn_v2/multi_rnn_cell/cell_{q[1][5:]}/zoneout_lstm_cell/lstm_cell/{q[2]}"
        elif q[0] == "decoder" and q[1].startswith("attention"):
            i = int(q[1][9:]) - 1
            att = f"{cell}/attention_wrapper/{num(kinds[i], sum(1 for j in range(i) if kinds[j] == kinds[i]))}"
            leaf = {"memory_layer": f"{dec}/{num('memory_layer', i)}/kernel", "query_layer": f"{att}/query_layer/kernel",
                    "attention_v": f"{att}/" + ("attention_variable" if kinds[i] == "location_sensitive_attention" else "attention_v"),
                    "attention_bias": f"{att}/attention_bias",
                    "location_conv": f"{att}/location_features_convolution/{q[-1]}",
                    "location_layer": f"{att}/location_features_layer/kernel",
                    "transition": f"{att}/transition_factor_projection/{q[-1]}"}
            n = leaf[q[2]]
        elif q[0] == "decoder" and q[1] in ("out_projection", "stop_token_projection"):
            if dual:
                n = f"{dec}/decoder/{q[1]}/{q[2]}"  # modules/module.py:716-723
            else:  # tacotron2 OutputAndStopTokenWrapper: dense (frames), dense_1 (stop token); models/models.py:167
                n = f"{dec}/decoder/output_and_stop_token_wrapper/{num('dense', int(q[1] == 'stop_token_projection'))}/{q[2]}"
        elif q[0] == "postnet" and q[1].startswith("conv_"):
            scope = f"post_net_v2/{num('conv1d', int(q[1][5:]))}"
            n = f"{scope}/conv1d/kernel" if q[2] == "kernel" else f"{scope}/batch_normalization/{bn[q[3]]}"
        elif q[:2] == ["postnet", "projection"]:
            n = f"post_net_v2/dense/{q[2]}"
        else:
            raise KeyError(f"no TensorFlow name rule for weight {k}")
        out[k] = n
    assert len(set(out.values())) == len(out), "reconstructed names collide"
    return out


def _natural(name):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", name)]


def resolve_names(hparams, variables):
    """{checkpoint variable name: weight-blob name} by fragment + shape, for checkpoints written by the reference.

    ``variables``: {name: (dtype, shape)} (``checkpoint_variables``).  Every weight of the model gets a fragment pattern
    (what any TF naming of that layer must contain) and its expected shape; candidates = variables that match both
    (singleton axes ignored); several weights sharing pattern and shape (prenet layers, highway layers, the four
    projections of a MultiHeadAttention, self-attention hops) take their candidates in natural name order -- the order
    TF numbers sibling scopes in.  Optimiser slots, global_step and beta powers are ignored.  Weights that stay
    ambiguous or unmatched are left out (``import_weights`` reports them)."""
    from .weights import weight_shapes
    hp = hparams
    shapes = weight_shapes(hp)
    squeeze = lambda sh: tuple(d for d in sh if d != 1)  # noqa: E731
    cand = {n: squeeze(sh) for n, (_, sh) in variables.items()
            if not re.search(r"/Adam(_\d+)?$|^global_step$|beta\d_power$|/ExponentialMovingAverage$", n)}
    enc_side = lambda n: "decoder" not in n  # noqa: E731
    dec_side = lambda n: "decoder" in n  # noqa: E731
    bnp = {"gamma": "gamma", "beta": "beta", "mean": "moving_mean", "var": "moving_variance"}

    def pattern(k):
        """-> (regex, side predicate or None) for weight-blob name k"""
        q = k.split("/")
        leaf = q[-1]
        if k in ("embedding", "speaker_embedding"):
            return r"embedding(_\d+)?/[^/]*(embedding|table|kernel)[^/]*$", None
        if q[0] == "speaker_embedding_projection":
            return rf"^(?!.*(encoder|decoder|pre_net|highway)).*dense[^/]*/{leaf}$", None
        if q[:2] == ["encoder", "prenet"]:
            return rf"pre_?net[^/]*/dense[^/]*/{leaf}$", enc_side
        if q[:2] == ["encoder", "cbhg"] and (q[2].startswith("conv1d_K") or q[2] in ("proj1", "proj2")):
            tail = "conv1d[^/]*/kernel" if q[3] == "kernel" else rf"batch_normalization[^/]*/{bnp[q[4]]}"
            return rf"(^|/){q[2]}/(.*/)?{tail}$", enc_side
        if q[:3] == ["encoder", "cbhg", "adjustment"]:
            return rf"cbhg[^/]*/dense[^/]*/{leaf}$", enc_side
        if q[:2] == ["encoder", "cbhg"] and q[2].startswith("highway_"):
            return rf"highway[^/]*/{q[3]}[^/]*/{leaf}$", enc_side
        if q[:2] == ["encoder", "cbhg"] and q[2].startswith("lstm_"):
            return rf"/{q[2][5:]}/(.*/)?{leaf}$", enc_side
        if q[:2] == ["encoder", "sa_projection"]:
            return rf"encoder[^/]*/dense[^/]*/{leaf}$", enc_side
        if q[0] in ("encoder", "decoder") and q[1].startswith("sa_"):
            side = enc_side if q[0] == "encoder" else dec_side
            if q[2] == "mha":
                return rf"multi_head_attention[^/]*/dense[^/]*/{leaf}$", side
            return rf"self_attention_transformer[^/]*/dense[^/]*/{leaf}$", side
        if q[:3] == ["decoder", "prenet", "ms"]:
            return rf"multi_speaker_pre_?net[^/]*/dense[^/]*/{leaf}$", None
        if q[:2] == ["decoder", "prenet"]:
            return rf"(^|/)pre_?net[^/]*/dense[^/]*/{leaf}$", dec_side
        if q[0] == "decoder" and q[1] in ("attention_lstm", "lstm_1", "lstm_2"):
            return rf"lstm_cell[^/]*/{leaf}$", dec_side
        if q[0] == "decoder" and q[1].startswith("attention"):
            frag = {"memory_layer": "memory_layer[^/]*/kernel", "query_layer": "query_layer[^/]*/kernel",
                    "attention_v": "(attention_v|attention_variable)", "attention_bias": "attention_bias",
                    "location_conv": rf"location_features_convolution[^/]*/{leaf}",
                    "location_layer": "location_features_layer[^/]*/kernel",
                    "transition": rf"transition_factor_projection[^/]*/{leaf}"}[q[2]]
            return rf"(^|/){frag}$", None
        if q[0] == "decoder" and q[1] == "out_projection":
            return rf"(out_projection/{leaf}|output_and_stop_token_wrapper/dense/{leaf})$", None
        if q[0] == "decoder" and q[1] == "stop_token_projection":
            return rf"(stop_token_projection/{leaf}|output_and_stop_token_wrapper/dense_1/{leaf})$", None
        if q[0] == "postnet" and q[1].startswith("conv_"):
            tail = "conv1d[^/]*/kernel" if q[2] == "kernel" else rf"batch_normalization[^/]*/{bnp[q[3]]}"
            return rf"post_?net.*/{tail}$", None
        if q[:2] == ["postnet", "projection"]:
            return rf"post_?net[^/]*/dense[^/]*/{leaf}$", None
        return None, None

    groups = {}
    for k, (shape, _) in shapes.items():
        pat, side = pattern(k)
        if pat is not None:
            groups.setdefault((pat, side, squeeze(shape)), []).append(k)
    out, taken = {}, set()
    # kernels (and everything that is not a bias) first; a bias then follows its kernel into the same TF scope, which
    # settles groups of equal-shaped biases (the three decoder LSTMs) that name order alone cannot
    ordered = sorted(groups.items(), key=lambda kv: all(k.endswith("/bias") for k in kv[1]))
    for (pat, side, shape), ks in ordered:
        if all(k.endswith("/bias") for k in ks):
            inv = {k: n for n, k in out.items()}
            sib = {k: inv[k[:-4] + "kernel"][:-6] + "bias" for k in ks
                   if k[:-4] + "kernel" in inv and inv[k[:-4] + "kernel"].endswith("kernel")}
            if len(sib) == len(ks) and all(cand.get(n) == shape and n not in taken for n in sib.values()):
                for k, n in sib.items():
                    out[n] = k
                    taken.add(n)
                continue
        rx = re.compile(pat)
        names = sorted((n for n, sh in cand.items() if sh == shape and rx.search(n) and (side is None or side(n))
                        and n not in taken), key=_natural)
        if len(names) != len(ks):
            continue  # missing or ambiguous: leave to name_map / the error report
        # blob-side order: the order weight_shapes lists them in IS layer order (prenet 0, 1; highway 0..; key, value,
        # query, output; hop 0, 1), which is the order TF numbers the sibling scopes in
        for n, k in zip(names, ks):
            out[n] = k
            taken.add(n)
    return out


def import_weights(hparams, prefix, name_map=None, strict=True):
    """Checkpoint ``prefix`` -> weight blob for ``hparams`` (float32, the layouts of include/sat.h).

    ``name_map``: {checkpoint variable name: weight-blob name} overriding / extending ``default_name_map``; variables
    whose checkpoint name already IS a weight-blob name map to themselves (checkpoints written by
    ``export_weights``).  Every mapped tensor must have the shape ``weight_shapes(hparams)`` expects (a trailing or
    middle singleton axis is squeezed / added, e.g. the location convolution kernel [K, 1, F]).  ``strict``: raise if
    any weight of the model is left unassigned, listing what is missing and which checkpoint variables were unused."""
    from .weights import weight_shapes
    shapes = weight_shapes(hparams)
    avail = checkpoint_variables(prefix)
    mapping = {n: n for n in avail if n in shapes}
    mapping.update(default_name_map(avail))
    # fragment + shape resolution of the TensorFlow-scoped names of a reference checkpoint; never overrides the above
    for n, k in resolve_names(hparams, avail).items():
        if n not in mapping and k not in mapping.values():
            mapping[n] = k
    mapping.update(name_map or {})
    bad = sorted(v for v in mapping.values() if v not in shapes)
    if bad:
        raise KeyError(f"name_map targets that are not weights of this model: {bad}")
    dup = sorted({v for v in mapping.values() if list(mapping.values()).count(v) > 1})
    if dup:
        raise ValueError(f"several checkpoint variables map to the same weight: {dup}")
    tensors = read_checkpoint(prefix, names=set(mapping))
    W = {}
    for src, dst in mapping.items():
        a, want = tensors[src], tuple(shapes[dst][0])
        if tuple(a.shape) != want:
            if int(np.prod(a.shape)) == int(np.prod(want)) and [d for d in a.shape if d != 1] == [d for d in want if d != 1]:
                a = a.reshape(want)
            else:
                raise ValueError(f"{src} -> {dst}: checkpoint shape {tuple(a.shape)} != expected {want}")
        W[dst] = np.ascontiguousarray(a, dtype=np.float32)
    missing = sorted(set(shapes) - set(W))
    if strict and missing:
        unused = sorted(set(avail) - set(mapping))
        raise KeyError(f"{len(missing)} weights not found in {prefix}: {missing[:12]}{' ...' if len(missing) > 12 else ''}; "
                       f"unused checkpoint variables: {unused[:12]}{' ...' if len(unused) > 12 else ''} "
                       f"(pass name_map={{checkpoint name: weight name}})")
    return W


def export_weights(prefix, W):
    """Weight blob -> V2 checkpoint under the weight-blob names (round-trips through ``import_weights``)."""
    write_checkpoint(prefix, {k: np.asarray(v, dtype=np.float32) for k, v in W.items()})
