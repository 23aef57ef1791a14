"""Mirror of /root/reference/modules/forward_attention.py.  The step math runs inside the persistent
decoder kernel; ``__call__`` (one stand-alone step through sat_attention_step) exists for unit parity."""
from collections import namedtuple

import torch

from .. import engine


class ForwardAttentionState(namedtuple("ForwardAttentionState", ["alignments", "alpha", "u"])):
    pass


class _BahdanauBase:
    """TF BahdanauAttention base: values = memory zeroed past the length,
    keys = memory_layer(values), computed once and lazily."""
    kind = "additive"

    def __init__(self, num_units, memory, memory_sequence_length, weights=None, prefix=None, **options):
        if weights is None or prefix is None:
            raise ValueError("attention mechanisms need weights= and prefix=")
        self.num_units = num_units
        self.memory = memory
        self.memory_sequence_length = memory_sequence_length
        self.weights, self.prefix = weights, prefix
        self.options = dict(options)
        self._kv = None

    def _prepare(self):
        if self._kv is None:
            self._kv = engine.attention_prepare(self.weights, self.prefix, self.memory, self.memory_sequence_length,
                                                self.num_units)
        return self._kv

    @property
    def keys(self):
        return self._prepare()[0]

    @property
    def values(self):
        return self._prepare()[1]

    @property
    def batch_size(self):
        return self.memory.shape[0]

    @property
    def alignments_size(self):
        return self.memory.shape[1]

    def initial_alignments(self, batch_size, dtype=torch.float32):
        return torch.zeros((batch_size, self.alignments_size), device=self.memory.device, dtype=dtype)

    def initial_state(self, batch_size, dtype=torch.float32):
        return self.initial_alignments(batch_size, dtype)

    @property
    def state_size(self):
        return self.alignments_size

    def _step(self, query, st):
        align, ctx = engine.attention_step(self.weights, self.prefix, self.kind, self.options, self.keys, self.values,
                                           self.memory_sequence_length, query, st)
        self.last_context = ctx
        return align

    def __call__(self, query, state):
        st = {"alignments": state.clone()}
        align = self._step(query, st)
        return align, align


class BahdanauAttention(_BahdanauBase):
    """tf.contrib.seq2seq.BahdanauAttention as used by modules/attentions.py:53-57."""
    kind = "additive"


class LocationSensitiveAttention(_BahdanauBase):
    """tacotron2 LocationSensitiveAttention (modules/attentions.py:35-42)."""
    kind = "location_sensitive"

    def __init__(self, num_units, memory, memory_sequence_length, attention_kernel, attention_filters,
                 smoothing=False, cumulative_weights=True, weights=None, prefix=None):
        if smoothing:
            raise NotImplementedError("smoothing is never enabled by the reference (attention_factories.py:16)")
        super().__init__(num_units, memory, memory_sequence_length, weights=weights, prefix=prefix,
                         attention_kernel=attention_kernel, attention_filters=attention_filters,
                         cumulative_weights=cumulative_weights)

    def __call__(self, query, state):
        st = {"alignments": state.clone()}
        align = self._step(query, st)
        return align, st["alignments"]


class ForwardAttention(_BahdanauBase):
    """modules/forward_attention.py:48-136."""
    kind = "forward"

    def __init__(self, num_units, memory, memory_sequence_length, attention_kernel, attention_filters,
                 use_transition_agent=False, cumulative_weights=True, name="ForwardAttention", weights=None,
                 prefix=None):
        super().__init__(num_units, memory, memory_sequence_length, weights=weights, prefix=prefix,
                         attention_kernel=attention_kernel, attention_filters=attention_filters,
                         cumulative_weights=cumulative_weights, use_transition_agent=use_transition_agent)

    def __call__(self, query, state):
        prev_alignments, prev_alpha, prev_u = state
        st = {"alignments": prev_alignments.clone(), "alpha": prev_alpha.clone(),
              "u": prev_u.reshape(-1).clone().contiguous()}
        alpha = self._step(query, st)
        return alpha, ForwardAttentionState(st["alignments"], st["alpha"], st["u"].reshape(-1, 1))

    @property
    def state_size(self):
        return ForwardAttentionState(self.alignments_size, self.alignments_size, 1)

    def initial_state(self, batch_size, dtype=torch.float32):
        a0 = self.initial_alignments(batch_size, dtype)
        alpha = torch.zeros_like(a0)
        alpha[:, 0] = 1
        u = torch.full((batch_size, 1), 0.5, device=a0.device, dtype=dtype)
        return ForwardAttentionState(a0, alpha, u)
