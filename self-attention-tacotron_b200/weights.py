"""Weight blob: names, shapes and random initialisation.

The blob is a flat ``{name: float32 ndarray}`` dict in TensorFlow layouts
(Dense ``[in, out]``, Conv1D ``[k, in, out]``, LSTM ``[in + units, 4 * units]`` gate
order i, j, f, o) -- exactly what the C ABI takes (include/sat.h) and what a
checkpoint importer would produce.  Random init follows
SURVEY.md 8c: Glorot-uniform kernels, zero biases, highway T-bias -1, BN
gamma=1 beta=0 mean=0 var=1, seed 12345 (modules/transformer_test.py:46).
``stress=True`` additionally randomises biases and BN statistics so that parity
tests exercise every term.
"""
import numpy as np


def speaker_dim(hp):
    """width of the speaker vector the decoder sees: the embedding, or its relu projection when
    speaker_embedding_projection_out_dim > -1 (models/models.py:73-75)"""
    p = getattr(hp, "speaker_embedding_projection_out_dim", -1)
    return p if p > -1 else hp.speaker_embedding_dim


def _mem_dims(hp):
    # speaker_embedd_to_decoder: the speaker vector is tiled over time and concatenated onto both encoder outputs
    # (models/models.py:78-83), so the attention memories (and contexts) widen by its width
    use = getattr(hp, "use_speaker_embedding", False) and getattr(hp, "speaker_embedd_to_decoder", False)
    extra = speaker_dim(hp) if use else 0
    if hp.decoder == "ExtendedDecoder":
        return [hp.cbhg_out_units + extra]
    return [hp.cbhg_out_units + extra, hp.self_attention_out_units + extra]


def _att_kinds_units(hp):
    if hp.decoder == "ExtendedDecoder":
        return [(hp.attention, hp.attention_out_units)]
    return [(hp.attention, hp.attention1_out_units), (hp.attention2, hp.attention2_out_units)]


def weight_shapes(hp):
    """-> {name: (shape, kind)}; kind in glorot|zeros|ones|tbias|table."""
    s = {}

    def dense(prefix, i, o, bias=True):
        s[prefix + "/kernel"] = ((i, o), "glorot")
        if bias:
            s[prefix + "/bias"] = ((o,), "zeros")

    def conv_bn(prefix, k, i, o):
        s[prefix + "/kernel"] = ((k, i, o), "glorot")
        s[prefix + "/bn/gamma"] = ((o,), "ones")
        s[prefix + "/bn/beta"] = ((o,), "zeros")
        s[prefix + "/bn/mean"] = ((o,), "zeros")
        s[prefix + "/bn/var"] = ((o,), "ones")

    def mha(prefix, i, d):
        for n in ("key", "value", "query"):
            dense(f"{prefix}/mha/{n}", i, d)
        dense(f"{prefix}/mha/output", d, d)
        dense(f"{prefix}/transform", d, d)

    # ---- encoder (modules/module.py:30-113, 293-441) ----
    s["embedding"] = ((hp.num_symbols, hp.embedding_dim), "table")
    d = hp.embedding_dim
    for i, u in enumerate(hp.encoder_prenet_out_units):
        dense(f"encoder/prenet/dense_{i}", d, u)
        d = u
    cin, half = d, hp.cbhg_out_units // 2
    for k in range(1, hp.max_filter_width + 1):
        conv_bn(f"encoder/cbhg/conv1d_K{k}", k, cin, hp.conv_channels)
    conv_bn("encoder/cbhg/proj1", 3, hp.max_filter_width * hp.conv_channels, hp.projection1_out_channels)
    conv_bn("encoder/cbhg/proj2", 3, hp.projection1_out_channels, hp.projection2_out_channels)
    assert hp.projection2_out_channels == cin, "CBHG residual needs proj2 channels == input channels"
    if hp.projection2_out_channels != half:
        dense("encoder/cbhg/adjustment", hp.projection2_out_channels, half)
    for i in range(hp.num_highway):
        dense(f"encoder/cbhg/highway_{i}/H", half, half)
        s[f"encoder/cbhg/highway_{i}/T/kernel"] = ((half, half), "glorot")
        s[f"encoder/cbhg/highway_{i}/T/bias"] = ((half,), "tbias")
    for dr in ("fw", "bw"):
        dense(f"encoder/cbhg/lstm_{dr}", half + half, 4 * half)
    if hp.encoder == "SelfAttentionCBHGEncoder":
        dense("encoder/sa_projection", hp.cbhg_out_units, hp.self_attention_out_units)
        for i in range(hp.self_attention_num_hop):
            mha(f"encoder/sa_{i}", hp.self_attention_out_units, hp.self_attention_out_units)
    if hp.use_speaker_embedding:
        s["speaker_embedding"] = ((hp.num_speakers, hp.speaker_embedding_dim), "table")
        if hp.speaker_embedding_projection_out_dim > -1:
            dense("speaker_embedding_projection", hp.speaker_embedding_dim, hp.speaker_embedding_projection_out_dim)

    # ---- decoder (modules/module.py:562-623, 1017-1152, 1499-1562) ----
    M, r, nf = hp.num_mels, hp.outputs_per_step, hp.n_feed_frame
    pu = hp.decoder_prenet_out_units
    if hp.use_speaker_embedding and hp.speaker_embedd_to_prenet:
        dense("decoder/prenet/ms/dense0", M * nf, pu[0])
        dense("decoder/prenet/ms/speaker_projection", speaker_dim(hp), pu[0])
        dense("decoder/prenet/ms/dense", pu[0], pu[0])
        dense("decoder/prenet/dense_1", pu[0], pu[1])
    else:
        d = M * nf
        for i, u in enumerate(pu):
            dense(f"decoder/prenet/dense_{i}", d, u)
            d = u
    mem = _mem_dims(hp)
    A, Dd = hp.attention_out_units, hp.decoder_out_units
    dense("decoder/attention_lstm", pu[-1] + sum(mem) + A, 4 * A)
    for i, ((kind, U), md) in enumerate(zip(_att_kinds_units(hp), mem), start=1):
        p = f"decoder/attention{i}"
        dense(p + "/memory_layer", md, U, bias=False)
        dense(p + "/query_layer", A, U, bias=False)
        s[p + "/attention_v"] = ((U,), "glorot")
        if kind in ("forward", "location_sensitive"):
            s[p + "/attention_bias"] = ((U,), "zeros")
            s[p + "/location_conv/kernel"] = ((hp.attention_kernel, 1, hp.attention_filters), "glorot")
            s[p + "/location_conv/bias"] = ((hp.attention_filters,), "zeros")
            dense(p + "/location_layer", hp.attention_filters, U, bias=False)
            if kind == "forward" and hp.use_forward_attention_transition_agent:
                dense(p + "/transition", md + U, 1)
    dense("decoder/lstm_1", A + sum(mem) + Dd, 4 * Dd)
    dense("decoder/lstm_2", Dd + Dd, 4 * Dd)
    d_out = Dd
    if hp.decoder == "DualSourceTransformerDecoder":
        Ds = hp.decoder_self_attention_out_units
        assert Ds == Dd, "SelfAttentionTransformer residual needs self-attention units == decoder_out_units"
        for i in range(hp.decoder_self_attention_num_hop):
            mha(f"decoder/sa_{i}", Dd, Ds)
        d_out = Ds
    dense("decoder/out_projection", d_out, M * r)
    dense("decoder/stop_token_projection", d_out, 1)

    # ---- post-net v2 ----
    if hp.use_postnet_v2:
        C, k = hp.postnet_v2_out_channels, hp.postnet_v2_kernel_size
        for i in range(hp.num_postnet_v2_layers):
            conv_bn(f"postnet/conv_{i}", k, M if i == 0 else C, C)
        dense("postnet/projection", C, M)
    return s


def init_weights(hp, seed=12345, stress=False, dtype=np.float32):
    rng = np.random.default_rng(seed)
    W = {}
    for name, (shape, kind) in weight_shapes(hp).items():
        if kind in ("glorot", "table"):
            if len(shape) == 1:
                fan_in, fan_out = shape[0], 1
            elif len(shape) == 2:
                fan_in, fan_out = shape
            else:
                fan_in, fan_out = shape[0] * shape[1], shape[0] * shape[2]
            lim = np.sqrt(6.0 / (fan_in + fan_out)) if kind == "glorot" else 1.0
            w = rng.uniform(-lim, lim, size=shape)
        elif kind == "zeros":
            w = rng.normal(0, 0.1, size=shape) if stress else np.zeros(shape)
        elif kind == "ones":
            w = rng.uniform(0.5, 1.5, size=shape) if stress else np.ones(shape)
        elif kind == "tbias":
            w = np.full(shape, -1.0) + (rng.normal(0, 0.1, size=shape) if stress else 0)
        else:
            raise ValueError(kind)
        W[name] = np.ascontiguousarray(w, dtype=dtype)
    return W


def cast_weights(W, dtype):
    return {k: v.astype(dtype) for k, v in W.items()}


def decoder_param_count(hp):
    return sum(int(np.prod(sh)) for n, (sh, _) in weight_shapes(hp).items() if n.startswith("decoder/"))
