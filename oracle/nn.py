"""Primitive layers of the oracle (test infrastructure; see oracle/__init__.py).

Every function works on NumPy arrays of whatever float dtype it is handed
(float64 for goldens, float32 for the timed CPU baseline).  Layout conventions
are TensorFlow's: activations ``[B, T, C]``, Dense kernels ``[in, out]``,
Conv1D kernels ``[k, in, out]``, LSTM kernels ``[in + units, 4 * units]`` with
gate order i, j, f, o.
"""
import numpy as np

# oracle switches for semantics that live in code absent from /root/reference
# (SURVEY.md Appendix A); flip here if the real tacotron2 source disagrees.
SWITCHES = {
    "bn_eps": 1e-3,                   # tf.layers.batch_normalization default (A.2)
    "zoneout_output_is_zoned": False,  # ZoneoutLSTMCell returns the un-zoned h (A.4)
    "lstm_forget_bias": 1.0,          # tf.nn.rnn_cell.LSTMCell default (A.4)
    "postnet_last_activation": None,  # PostNetV2 last conv has no activation (A.8)
}


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def relu(x):
    return np.maximum(x, 0)


def softsign(x):
    return x / (1.0 + np.abs(x))


def softmax(x, axis=-1):
    m = np.max(x, axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0)
    e = np.exp(x - m)
    return e / np.sum(e, axis=axis, keepdims=True)


def dense(x, kernel, bias=None, act=None):
    """tf.layers.Dense: ``act(x @ kernel + bias)`` on the last axis."""
    y = x @ kernel
    if bias is not None:
        y = y + bias
    return act(y) if act is not None else y


def prenet(x, layers):
    """tacotron2 ``PreNet`` stack, inference form (dropout off): relu(Dense).

    Call sites modules/module.py:394,574-576; semantics SURVEY.md A.1.
    ``layers`` is a list of (kernel, bias).
    """
    for k, b in layers:
        x = relu(dense(x, k, b))
    return x


def multi_speaker_prenet(x, speaker_embed, w):
    """modules/multi_speaker_modules.py:27-32 (dropout off at inference).

    ``w`` has dense0, speaker_projection, dense as (kernel, bias) pairs.
    """
    d0 = relu(dense(x, *w["dense0"]))
    d0 = d0 + softsign(dense(speaker_embed, *w["speaker_projection"]))
    return relu(dense(d0, *w["dense"]))


def conv1d_same(x, kernel, bias=None):
    """tf.layers.Conv1D(padding='SAME', stride 1): cross-correlation.

    ``out[t] = sum_i w[i] . x[t + i - pad_left]``, pad_left = (k-1)//2 (even k pads
    the extra sample on the right), zero padding, no length masking (A.2).
    x ``[B, T, Cin]``, kernel ``[k, Cin, Cout]``.
    """
    k = kernel.shape[0]
    B, T, _ = x.shape
    pl = (k - 1) // 2
    pr = k - 1 - pl
    xp = np.pad(x, ((0, 0), (pl, pr), (0, 0)))
    y = np.zeros((B, T, kernel.shape[2]), dtype=x.dtype)
    for i in range(k):
        y += xp[:, i:i + T, :] @ kernel[i]
    if bias is not None:
        y = y + bias
    return y


def batch_norm_inference(x, bn):
    """tf.layers.batch_normalization(training=False); bn = (gamma, beta, mean, var)."""
    gamma, beta, mean, var = bn
    eps = x.dtype.type(SWITCHES["bn_eps"])
    return gamma * (x - mean) / np.sqrt(var + eps) + beta


def conv1d_bn_act(x, kernel, bn, act):
    """tacotron2 ``Conv1d``: Conv1D(no bias) -> BN -> activation (A.2);
    call sites modules/module.py:46-68."""
    y = batch_norm_inference(conv1d_same(x, kernel), bn)
    return act(y) if act is not None else y


def maxpool1d_k2_same(x):
    """tf.layers.MaxPooling1D(pool_size=2, strides=1, padding='SAME')
    (modules/module.py:54): out[t] = max(x[t], x[t+1]); out[T-1] = x[T-1]."""
    y = x.copy()
    y[:, :-1, :] = np.maximum(x[:, :-1, :], x[:, 1:, :])
    return y


def highway(x, wh, bh, wt, bt):
    """tacotron2 ``HighwayNet`` (A.3): H=relu(Dense), T=sigmoid(Dense), H*T + x*(1-T)."""
    h = relu(dense(x, wh, bh))
    t = sigmoid(dense(x, wt, bt))
    return h * t + x * (1.0 - t)


def lstm_cell(x, c_prev, h_prev, kernel, bias):
    """tf.nn.rnn_cell.LSTMCell (A.4): gates = [x, h] @ kernel + bias, split
    i, j, f, o; c = sigmoid(f + forget_bias) * c_prev + sigmoid(i) * tanh(j);
    h = sigmoid(o) * tanh(c)."""
    z = np.concatenate([x, h_prev], axis=-1) @ kernel + bias
    i, j, f, o = np.split(z, 4, axis=-1)
    fb = z.dtype.type(SWITCHES["lstm_forget_bias"])
    c = sigmoid(f + fb) * c_prev + sigmoid(i) * np.tanh(j)
    h = sigmoid(o) * np.tanh(c)
    return c, h


def zoneout_lstm_cell(x, state, kernel, bias, zc, zh):
    """tacotron2 ``ZoneoutLSTMCell``, inference form (A.4):
    c = c_new - zc (c_new - c_prev), h likewise; the returned OUTPUT is the
    un-zoned h_new, the carried state is the zoned (c, h)."""
    c_prev, h_prev = state
    c_new, h_new = lstm_cell(x, c_prev, h_prev, kernel, bias)
    c = c_new - zc * (c_new - c_prev)
    h = h_new - zh * (h_new - h_prev)
    out = h if SWITCHES["zoneout_output_is_zoned"] else h_new
    return out, (c, h)


def bidirectional_zoneout_lstm(x, lengths, fw, bw, zc, zh):
    """tf.nn.bidirectional_dynamic_rnn over two ZoneoutLSTMCells with
    ``sequence_length`` (modules/module.py:93-108; semantics A.12): past a row's
    length the output is zero and the state is copied through; the backward
    direction reverses each row by its own length.  fw/bw = (kernel, bias)."""
    B, T, _ = x.shape
    units = fw[1].shape[0] // 4
    lengths = np.asarray(lengths)

    def run(xs, w):
        c = np.zeros((B, units), dtype=x.dtype)
        h = np.zeros((B, units), dtype=x.dtype)
        outs = np.zeros((B, T, units), dtype=x.dtype)
        for t in range(T):
            out, (c2, h2) = zoneout_lstm_cell(xs[:, t], (c, h), w[0], w[1], zc, zh)
            live = (t < lengths)[:, None]
            outs[:, t] = np.where(live, out, 0)
            c = np.where(live, c2, c)
            h = np.where(live, h2, h)
        return outs

    def reverse_by_length(a):
        r = a.copy()
        for b in range(B):
            n = int(lengths[b])
            r[b, :n] = a[b, :n][::-1]
        return r

    out_fw = run(x, fw)
    out_bw = reverse_by_length(run(reverse_by_length(x), bw))
    return np.concatenate([out_fw, out_bw], axis=-1)


def multi_head_attention(x, w, num_heads, causal=False, lengths=None,
                         use_padding_mask=False):
    """modules/self_attention.py:108-128 (MultiHeadAttention with K=V=Q=x, i.e.
    SelfAttention :142-144) + ScaledDotProductAttentionMechanism :45-65.

    Dense K/V/Q (with bias) -> heads -> softmax(mask(QK^T / sqrt(d_head))) V ->
    merge heads -> output Dense.  Padding mask (:67-78) masks BOTH query rows
    and key columns past the length with -inf (so fully masked rows are NaN in
    the reference; no caller enables it).  Subsequent mask :80-86.
    Returns (output [B,T,D], probs [B,H,T,T]).
    """
    B, T, _ = x.shape
    k = dense(x, *w["key"])
    v = dense(x, *w["value"])
    q = dense(x, *w["query"])
    D = k.shape[-1]
    dh = D // num_heads

    def heads(a):
        return a.reshape(B, T, num_heads, dh).transpose(0, 2, 1, 3)

    k, v, q = heads(k), heads(v), heads(q)
    s = (q @ k.transpose(0, 1, 3, 2)) / np.sqrt(x.dtype.type(dh))
    if use_padding_mask:
        m = np.arange(T)[None, :] < np.asarray(lengths)[:, None]
        m2 = m[:, None, :] & m[:, :, None]
        s = np.where(m2[:, None], s, -np.inf)
    if causal:
        tri = np.tril(np.ones((T, T), dtype=bool))
        s = np.where(tri[None, None], s, -np.inf)
    p = softmax(s, axis=-1)
    ctx = (p @ v).transpose(0, 2, 1, 3).reshape(B, T, D)
    return dense(ctx, *w["output"]), p


def self_attention_transformer(x, w, num_heads, causal, lengths=None):
    """modules/module.py:363-371: SelfAttention -> Dense(tanh) -> residual add."""
    a, p = multi_head_attention(x, w["mha"], num_heads, causal=causal, lengths=lengths)
    return x + np.tanh(dense(a, *w["transform"])), p
