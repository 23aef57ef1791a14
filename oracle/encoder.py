"""Encoder side of the oracle (test infrastructure; see oracle/__init__.py).

Weights arrive as a flat ``{name: ndarray}`` dict using the names documented in
DESIGN.md ("weight blob").  ``hp`` is any object with the attribute names of
/root/reference/hparams.py.
"""
import numpy as np
from . import nn


def kb(W, prefix):
    return W[prefix + "/kernel"], W[prefix + "/bias"]


def bn(W, prefix):
    return (W[prefix + "/bn/gamma"], W[prefix + "/bn/beta"],
            W[prefix + "/bn/mean"], W[prefix + "/bn/var"])


def embedding(table, ids, index_offset=0):
    """tacotron2 ``Embedding`` (models/models.py:28,54; A.11): table[ids - offset]."""
    return table[np.asarray(ids) - index_offset]


def mha_weights(W, prefix):
    return {n: kb(W, f"{prefix}/{n}") for n in ("key", "value", "query", "output")}


def sat_weights(W, prefix):
    return {"mha": mha_weights(W, prefix + "/mha"), "transform": kb(W, prefix + "/transform")}


def zoneout_cbhg(x, lengths, W, hp, prefix="encoder/cbhg"):
    """modules/module.py:77-110 (ZoneoutCBHG.call)."""
    bank = [nn.conv1d_bn_act(x, W[f"{prefix}/conv1d_K{k}/kernel"], bn(W, f"{prefix}/conv1d_K{k}"), nn.relu)
            for k in range(1, hp.max_filter_width + 1)]
    conv_outputs = np.concatenate(bank, axis=-1)
    pooled = nn.maxpool1d_k2_same(conv_outputs)
    p1 = nn.conv1d_bn_act(pooled, W[f"{prefix}/proj1/kernel"], bn(W, f"{prefix}/proj1"), nn.relu)
    p2 = nn.conv1d_bn_act(p1, W[f"{prefix}/proj2/kernel"], bn(W, f"{prefix}/proj2"), None)
    hw_in = p2 + x
    if hw_in.shape[2] != hp.cbhg_out_units // 2:
        hw_in = nn.dense(hw_in, *kb(W, f"{prefix}/adjustment"))
    for i in range(hp.num_highway):
        hw_in = nn.highway(hw_in, *kb(W, f"{prefix}/highway_{i}/H"), *kb(W, f"{prefix}/highway_{i}/T"))
    zc = x.dtype.type(hp.zoneout_factor_cell)
    zh = x.dtype.type(hp.zoneout_factor_output)
    return nn.bidirectional_zoneout_lstm(hw_in, lengths, kb(W, f"{prefix}/lstm_fw"),
                                         kb(W, f"{prefix}/lstm_bw"), zc, zh)


def encoder_prenet(x, W, hp):
    layers = [kb(W, f"encoder/prenet/dense_{i}") for i in range(len(hp.encoder_prenet_out_units))]
    return nn.prenet(x, layers)


def zoneout_encoder_v1(ids, lengths, W, hp):
    """modules/module.py:336-339 (ZoneoutEncoderV1.call) after the embedding
    lookup of models/models.py:28,54.  Returns cbhg output [B,T,cbhg_out_units]."""
    x = embedding(W["embedding"], ids)
    return zoneout_cbhg(encoder_prenet(x, W, hp), lengths, W, hp)


def self_attention_cbhg_encoder(ids, lengths, W, hp):
    """modules/module.py:425-438 (SelfAttentionCBHGEncoder.call).

    Returns (lstm_output [B,T,256], self_attention_output [B,T,32],
    [alignment per head per hop], each [B,T,T]).  The padding mask is OFF
    (use_padding_mask defaults False, modules/self_attention.py:134), so padded
    positions take part in the encoder self-attention.
    """
    x = embedding(W["embedding"], ids)
    lstm_out = zoneout_cbhg(encoder_prenet(x, W, hp), lengths, W, hp)
    sa = nn.dense(lstm_out, *kb(W, "encoder/sa_projection"))
    aligns = []
    for i in range(hp.self_attention_num_hop):
        sa, p = nn.self_attention_transformer(sa, sat_weights(W, f"encoder/sa_{i}"),
                                              hp.self_attention_num_heads, causal=False, lengths=lengths)
        aligns += [p[:, h] for h in range(p.shape[1])]
    return lstm_out, sa, aligns
