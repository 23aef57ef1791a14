"""TensorFlow V2 checkpoint (tensor bundle) reader / writer and the checkpoint -> weight-blob importer
(sat_b200.checkpoint).  CPU only.  TensorFlow is not installable here and the reference ships no
checkpoint, so the format is pinned by hand-built known-answer bytes of the documented table / proto layout and by
round trips."""
import os
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sat_b200  # noqa: E402
from sat_b200 import checkpoint as ck  # noqa: E402
from sat_b200.tfrecord import crc32c  # noqa: E402


def test_bundle_entry_proto_known_answer():
    # BundleEntryProto{dtype: DT_FLOAT(1), shape{dim{size:2} dim{size:3}}, offset: 64, size: 24, crc32c: 0x01020304}
    want = bytes([0x08, 0x01, 0x12, 0x08, 0x12, 0x02, 0x08, 0x02, 0x12, 0x02, 0x08, 0x03, 0x20, 0x40, 0x28, 0x18,
                  0x35, 0x04, 0x03, 0x02, 0x01])
    assert ck._entry_proto(1, (2, 3), 64, 24, 0x01020304) == want
    e = ck._parse_entry(want)
    assert (e["dtype"], e["shape"], e["offset"], e["size"], e["crc32c"]) == (1, [2, 3], 64, 24, 0x01020304)


def test_table_block_known_answer_and_footer(tmp_path):
    # one block, restart interval 16: "ab"->"1", "abc"->"22" shares 2 key bytes
    blk = ck._build_block([(b"ab", b"1"), (b"abc", b"22")], 16)
    assert blk == b"\x00\x02\x01ab1" + b"\x02\x01\x02c22" + struct.pack("<II", 0, 1)
    assert list(ck._block_entries(blk)) == [(b"ab", b"1"), (b"abc", b"22")]
    path = str(tmp_path / "t.index")
    ck._write_table(path, [(b"ab", b"1"), (b"abc", b"22")])
    raw = open(path, "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57 and len(raw) >= 48
    # block trailer: type 0 + masked crc32c over block + type
    assert raw[len(blk)] == 0
    c = crc32c(blk + b"\x00")
    assert struct.unpack("<I", raw[len(blk) + 1:len(blk) + 5])[0] == (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF
    assert ck._read_table(path) == [(b"ab", b"1"), (b"abc", b"22")]


def test_round_trip_many_blocks_and_dtypes(tmp_path):
    rng = np.random.default_rng(0)
    tensors = {f"scope_{i // 7}/layer_{i}/kernel": rng.normal(size=(i % 5 + 1, 3)).astype(np.float32) for i in range(300)}
    tensors["global_step"] = np.asarray(12345, dtype=np.int64)
    tensors["flags"] = np.array([True, False])
    tensors["half"] = rng.normal(size=(4,)).astype(np.float16)
    tensors["empty"] = np.zeros((0, 4), dtype=np.float32)
    prefix = str(tmp_path / "model.ckpt-7")
    ck.write_checkpoint(prefix, tensors)
    assert os.path.getsize(prefix + ".index") > 2 * 4096  # several data blocks
    got = ck.read_checkpoint(prefix, verify_crc_below=None)
    assert sorted(got) == sorted(tensors)
    for k, v in tensors.items():
        assert got[k].dtype == v.dtype and got[k].shape == v.shape and np.array_equal(got[k], v), k
    assert ck.checkpoint_variables(prefix)["scope_0/layer_3/kernel"] == (np.dtype("<f4"), (4, 3))
    assert ck.latest_checkpoint(str(tmp_path)) == prefix
    assert ck.latest_checkpoint(str(tmp_path / "nowhere")) is None
    assert ck.read_checkpoint(prefix, names={"half"}).keys() == {"half"}
    with pytest.raises(KeyError):
        ck.read_checkpoint(prefix, names={"not_there"})


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "m")
    ck.write_checkpoint(prefix, {"a/kernel": np.arange(6, dtype=np.float32).reshape(2, 3)})
    data = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    data[5] ^= 0xFF
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(data))
    with pytest.raises(ValueError, match="checksum"):
        ck.read_checkpoint(prefix)
    idx = bytearray(open(prefix + ".index", "rb").read())
    idx[3] ^= 0x55
    open(prefix + ".index", "wb").write(bytes(idx))
    with pytest.raises(ValueError):
        ck.read_checkpoint(prefix)
    open(prefix + ".index", "wb").write(b"not a table")
    with pytest.raises(ValueError, match="magic"):
        ck.read_checkpoint(prefix)


@pytest.mark.parametrize("cfg", ["ljspeech_tacotron", "ljspeech_self_attention", "vctk_self_attention"])
def test_export_import_round_trip(cfg, tmp_path):
    hp = sat_b200.config(cfg)
    W = sat_b200.init_weights(hp, seed=3, stress=True)
    prefix = str(tmp_path / "model.ckpt-100")
    ck.export_weights(prefix, W)
    W2 = ck.import_weights(hp, ck.latest_checkpoint(str(tmp_path)))
    assert sorted(W2) == sorted(W)
    assert all(np.array_equal(W[k], W2[k]) for k in W)


def test_import_with_tensorflow_style_names(tmp_path):
    """Variables under TensorFlow scope names: the rule table resolves the names visible in the reference tree
    (conv1d_K*, proj1/2, batch_normalization statistics, out / stop-token projections, embedding), the rest comes from
    name_map; optimiser slots are ignored; shape mismatches and leftovers are reported."""
    hp = sat_b200.config("ljspeech_tacotron")
    W = sat_b200.init_weights(hp, seed=5, stress=True)
    tf_names, name_map = {}, {}
    bn = {"gamma": "gamma", "beta": "beta", "mean": "moving_mean", "var": "moving_variance"}
    for k, v in W.items():
        parts = k.split("/")
        if k == "embedding":
            n = "embedding/embedding"
        elif parts[:2] == ["encoder", "cbhg"] and (parts[2].startswith("conv1d_K") or parts[2] in ("proj1", "proj2")):
            scope = f"zoneout_encoder_v1/cbhg/{parts[2]}"
            n = f"{scope}/conv1d/kernel" if parts[3] == "kernel" else f"{scope}/batch_normalization/{bn[parts[4]]}"
        elif parts[1] in ("out_projection", "stop_token_projection"):
            n = f"decoder/{parts[1]}/{parts[2]}"
        else:
            n = "model/" + k.replace("/", "_x_")  # a name no rule knows
            name_map[n] = k
        tf_names[n] = v
        if n.endswith("kernel"):
            tf_names[n + "/Adam"] = np.zeros_like(v)
            tf_names[n + "/Adam_1"] = np.zeros_like(v)
    tf_names["global_step"] = np.asarray(1, dtype=np.int64)
    tf_names["beta1_power"] = np.asarray(0.9, dtype=np.float32)
    # the location convolution kernel is [K, 1, F] in the blob; a [K, F] tensor is reshaped
    lc = [n for n, k in name_map.items() if k.endswith("location_conv/kernel")][0]
    tf_names[lc] = tf_names[lc].reshape(hp.attention_kernel, hp.attention_filters)
    prefix = str(tmp_path / "model.ckpt-1")
    ck.write_checkpoint(prefix, tf_names)
    auto = ck.default_name_map(tf_names)
    assert auto["zoneout_encoder_v1/cbhg/conv1d_K7/batch_normalization/moving_variance"] == "encoder/cbhg/conv1d_K7/bn/var"
    assert auto["decoder/stop_token_projection/bias"] == "decoder/stop_token_projection/bias"
    assert not any("Adam" in n or n == "global_step" for n in auto)
    W2 = ck.import_weights(hp, prefix, name_map=name_map)
    assert sorted(W2) == sorted(W) and all(np.array_equal(W[k], W2[k]) for k in W)
    with pytest.raises(KeyError, match="weights not found"):
        ck.import_weights(hp, prefix)  # without the map the unknown names are reported
    partial = ck.import_weights(hp, prefix, strict=False)
    assert "embedding" in partial and "decoder/lstm_1/kernel" not in partial
    bad = dict(name_map)
    bad[lc] = "decoder/lstm_1/kernel"
    with pytest.raises((ValueError, KeyError)):
        ck.import_weights(hp, prefix, name_map=bad)


@pytest.mark.parametrize("cfg,extra", [("ljspeech_tacotron", {}), ("ljspeech_self_attention", dict(use_postnet_v2=True)),
                                       ("vctk_self_attention", dict(use_forward_attention_transition_agent=True)),
                                       ("codes_self_attention", {})])
def test_import_reference_scoped_checkpoint_without_name_map(cfg, extra, tmp_path):
    """A whole model written under the TensorFlow-scoped names the reference creates (checkpoint.tf_variable_names:
    layer scopes numbered in call order, lstm_cell / batch_normalization / memory_layer / ... fragments), with Adam
    slots and the step counters beside it, imports with NO name_map: every weight is found by fragment + shape, the
    blob is bit-identical."""
    hp = sat_b200.config(cfg)
    hp.override(**extra)
    W = sat_b200.init_weights(hp, seed=9, stress=True)
    names = ck.tf_variable_names(hp)
    assert sorted(names) == sorted(W)
    tensors = {}
    for k, v in W.items():
        tensors[names[k]] = v
        if k.endswith("kernel"):
            tensors[names[k] + "/Adam"] = np.zeros_like(v)
            tensors[names[k] + "/Adam_1"] = np.ones_like(v)
    tensors["global_step"] = np.asarray(7, dtype=np.int64)
    tensors["beta1_power"] = np.asarray(0.9, dtype=np.float32)
    tensors["beta2_power"] = np.asarray(0.999, dtype=np.float32)
    prefix = str(tmp_path / "model.ckpt-7")
    ck.write_checkpoint(prefix, tensors)
    W2 = ck.import_weights(hp, ck.latest_checkpoint(str(tmp_path)))
    assert sorted(W2) == sorted(W)
    for k in W:
        assert W2[k].dtype == np.float32 and np.array_equal(W[k], W2[k]), k
    # the three decoder LSTM biases share a shape: each must have followed its own kernel
    m = ck.resolve_names(hp, ck.checkpoint_variables(prefix))
    assert m[names["decoder/lstm_2/bias"]] == "decoder/lstm_2/bias"
    assert m[names["decoder/attention_lstm/bias"]] == "decoder/attention_lstm/bias"
    # a different top-level scope (e.g. a model wrapped in tf.variable_scope("model")) changes nothing
    ck.write_checkpoint(str(tmp_path / "wrapped.ckpt"), {"model/" + n: v for n, v in tensors.items()
                                                         if "Adam" not in n and "power" not in n and n != "global_step"},
                        update_state_file=False)
    W3 = ck.import_weights(hp, str(tmp_path / "wrapped.ckpt"))
    assert all(np.array_equal(W[k], W3[k]) for k in W)
