"""Decoder attention mechanisms of the oracle (test infrastructure).

Mirrors the ``AttentionMechanism`` contract used by the reference
(``__call__(query, state) -> (alignments, next_state)``, ``initial_state``,
``keys``, ``values``); see modules/attentions.py:25-62.
"""
import numpy as np
from . import nn


class _Bahdanau:
    """TF ``BahdanauAttention`` base: values = memory zeroed past
    the length; keys = memory_layer(values) (no bias), computed once; masked
    softmax with -inf."""

    def __init__(self, W, prefix, memory, lengths):
        self.W, self.prefix = W, prefix
        B, T, _ = memory.shape
        self.mask = np.arange(T)[None, :] < np.asarray(lengths)[:, None]
        self.values = memory * self.mask[:, :, None].astype(memory.dtype)
        self.keys = self.values @ W[prefix + "/memory_layer/kernel"]
        self.dtype = memory.dtype
        self.B, self.T = B, T

    def prob(self, score):
        return nn.softmax(np.where(self.mask, score, -np.inf), axis=-1)

    def initial_alignments(self):
        return np.zeros((self.B, self.T), dtype=self.dtype)


class Additive(_Bahdanau):
    """TF ``BahdanauAttention(normalize=False)`` (modules/attentions.py:53-57):
    e_j = sum_u v[u] tanh(keys[j,u] + (W_q query)[u])."""

    def initial_state(self):
        return self.initial_alignments()

    def __call__(self, query, state):
        pq = query @ self.W[self.prefix + "/query_layer/kernel"]
        e = np.sum(self.W[self.prefix + "/attention_v"] * np.tanh(self.keys + pq[:, None, :]), axis=2)
        a = self.prob(e)
        return a, a


def _location_energy(m, query, prev_alignments):
    """modules/forward_attention.py:88-103 + _location_sensitive_score :13-26."""
    W, p = m.W, m.prefix
    pq = query @ W[p + "/query_layer/kernel"]
    f = nn.conv1d_same(prev_alignments[:, :, None], W[p + "/location_conv/kernel"],
                       W[p + "/location_conv/bias"])
    pl = f @ W[p + "/location_layer/kernel"]
    e = np.sum(W[p + "/attention_v"] * np.tanh(m.keys + pq[:, None, :] + pl + W[p + "/attention_bias"]), axis=2)
    return e, pq


class LocationSensitive(_Bahdanau):
    """tacotron2 ``LocationSensitiveAttention`` (modules/attentions.py:35-42; A.10),
    smoothing=False always (models/attention_factories.py:16)."""

    def __init__(self, W, prefix, memory, lengths, cumulative_weights):
        super().__init__(W, prefix, memory, lengths)
        self.cumulative = cumulative_weights

    def initial_state(self):
        return self.initial_alignments()

    def __call__(self, query, state):
        e, _ = _location_energy(self, query, state)
        a = self.prob(e)
        return a, (a + state if self.cumulative else a)


class Forward(_Bahdanau):
    """modules/forward_attention.py:48-136.  State = (alignments, alpha, u)."""

    def __init__(self, W, prefix, memory, lengths, cumulative_weights, use_transition_agent):
        super().__init__(W, prefix, memory, lengths)
        self.cumulative = cumulative_weights
        self.agent = use_transition_agent

    def initial_state(self):
        a0 = self.initial_alignments()
        alpha = a0.copy()
        alpha[:, 0] = 1
        u = np.full((self.B, 1), 0.5, dtype=self.dtype)
        return a0, alpha, u

    def __call__(self, query, state):
        prev_a, prev_alpha, prev_u = state
        e, pq = _location_energy(self, query, prev_a)
        a = self.prob(e)
        shifted = np.pad(prev_alpha[:, :-1], ((0, 0), (1, 0)))
        alpha = ((1 - prev_u) * prev_alpha + prev_u * shifted + self.dtype.type(1e-7)) * a
        alpha_n = alpha / np.sum(alpha, axis=1, keepdims=True)
        if self.agent:
            ctx = np.einsum("bt,btd->bd", alpha_n, self.values)
            u = nn.sigmoid(nn.dense(np.concatenate([ctx, pq], axis=-1),
                                    self.W[self.prefix + "/transition/kernel"],
                                    self.W[self.prefix + "/transition/bias"]))
        else:
            u = prev_u
        nxt = (a + prev_a if self.cumulative else a, alpha_n, u)
        return alpha_n, nxt


class TeacherForcing(_Bahdanau):
    """modules/teacher_forcing_attention.py:29-44,63-78: replay
    teacher_alignments[:, index]; index starts at -1."""

    def __init__(self, W, prefix, memory, lengths, teacher_alignments):
        super().__init__(W, prefix, memory, lengths)
        self.teacher = teacher_alignments  # [B, S, T]

    def initial_state(self):
        return self.initial_alignments(), -1

    def __call__(self, query, state):
        _, idx = state
        idx = idx + 1
        a = self.teacher[:, idx]
        return a, (a, idx)


def attention_mechanism_factory(options):
    """modules/attentions.py:25-62.  ``options`` has the AttentionOptions fields."""

    def attention_fn(W, prefix, memory, lengths, teacher_alignments=None):
        k = options["attention"]
        if k == "forward":
            return Forward(W, prefix, memory, lengths, options["cumulative_weights"],
                           options["use_transition_agent"])
        if k == "location_sensitive":
            return LocationSensitive(W, prefix, memory, lengths, options["cumulative_weights"])
        if k in ("teacher_forcing_forward", "teacher_forcing_additive"):
            return TeacherForcing(W, prefix, memory, lengths, teacher_alignments)
        if k == "additive":
            return Additive(W, prefix, memory, lengths)
        raise ValueError(f"Unknown attention mechanism: {k}")

    return attention_fn
