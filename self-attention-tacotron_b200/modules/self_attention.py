"""Mirror of /root/reference/modules/self_attention.py over sat_mha_forward."""
from .. import engine


def _no_training(is_training):
    if is_training:
        raise NotImplementedError("the training graph (dropout, gradients) is outside the synthesis hot path")


class ScaledDotProductAttentionMechanism:
    """modules/self_attention.py:13-86: ``mechanism(query, memory_sequence_length)`` -> (context, alignment) for
    projected heads [B, H, T, d_head], computed by sat_sdpa_forward (the same kernels MultiHeadAttention uses with the
    projections fused in front).  Self-attention only: keys, values and query share T."""

    def __init__(self, keys, values, num_heads, drop_rate=0.0, is_training=False, use_padding_mask=True,
                 use_subsequent_mask=False):
        _no_training(is_training)
        self._keys, self._values = keys, values
        self.num_heads = num_heads
        self.drop_rate = drop_rate
        self.use_padding_mask = use_padding_mask
        self.use_subsequent_mask = use_subsequent_mask

    @property
    def keys(self):
        return self._keys

    @property
    def values(self):
        return self._values

    def __call__(self, query, memory_sequence_length=None):
        if query.shape != self._keys.shape or query.shape[1] != self.num_heads:
            raise ValueError(f"query {tuple(query.shape)} must match keys {tuple(self._keys.shape)} = [B, num_heads, T, d_head]")
        lengths = memory_sequence_length if self.use_padding_mask else None
        if self.use_padding_mask and lengths is None:
            raise ValueError("use_padding_mask=True needs memory_sequence_length")
        return engine.sdpa_forward(query, self._keys, self._values, causal=self.use_subsequent_mask, lengths=lengths)


class MultiHeadAttention:
    """modules/self_attention.py:89-128.  ``call((key, value, query))`` requires key is value is query
    (the only way the reference ever calls it, :142-144)."""

    def __init__(self, model_dim, num_heads, drop_rate, is_training, use_padding_mask=False,
                 use_subsequent_mask=False, trainable=True, name=None, dtype=None, weights=None, prefix=None, **kwargs):
        assert model_dim % num_heads == 0
        _no_training(is_training)
        if weights is None or prefix is None:
            raise ValueError("MultiHeadAttention needs weights= and prefix= (no TF variable store here)")
        self.model_dim, self.num_heads = model_dim, num_heads
        self.head_dim = model_dim // num_heads
        self.use_padding_mask = use_padding_mask
        self.use_subsequent_mask = use_subsequent_mask
        self.weights, self.prefix = weights, prefix

    def __call__(self, inputs, memory_sequence_length=None):
        key, value, query = inputs
        if not (key is value and value is query):
            raise NotImplementedError("only self-attention (key is value is query) is on the hot path")
        lengths = memory_sequence_length if self.use_padding_mask else None
        if self.use_padding_mask and lengths is None:
            raise ValueError("use_padding_mask=True needs memory_sequence_length")
        out, probs = engine.mha_forward(self.weights, self.prefix, query, self.num_heads, self.model_dim,
                                        causal=self.use_subsequent_mask, lengths=lengths)
        return out, [probs[:, i] for i in range(self.num_heads)]

    call = __call__


class SelfAttention:
    """modules/self_attention.py:131-144."""

    def __init__(self, model_dim, num_heads, drop_rate, is_training, use_padding_mask=False,
                 use_subsequent_mask=False, trainable=True, name=None, dtype=None, weights=None, prefix=None, **kwargs):
        self.attention = MultiHeadAttention(model_dim, num_heads, drop_rate, is_training,
                                            use_padding_mask=use_padding_mask, use_subsequent_mask=use_subsequent_mask,
                                            weights=weights, prefix=prefix)

    def __call__(self, inputs, memory_sequence_length=None):
        return self.attention((inputs, inputs, inputs), memory_sequence_length=memory_sequence_length)

    call = __call__
