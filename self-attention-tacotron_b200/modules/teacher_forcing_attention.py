"""Mirror of /root/reference/modules/teacher_forcing_attention.py: replay of given alignments."""
from collections import namedtuple

import torch

from .forward_attention import _BahdanauBase


class TeacherForcingAttentionState(namedtuple("TeacherForcingAttentionState", ["alignments", "index"])):
    pass


class _TeacherForcing(_BahdanauBase):
    kind = "teacher_forcing_forward"

    def __init__(self, num_units, memory, memory_sequence_length, teacher_alignments, name=None, weights=None,
                 prefix=None):
        super().__init__(num_units, memory, memory_sequence_length, weights=weights, prefix=prefix)
        if teacher_alignments is None:
            raise ValueError("teacher-forcing attention needs teacher_alignments [B, S, T_enc]")
        self.teacher_alignments = teacher_alignments

    @property
    def values(self):
        # only `values` of the base class is used by the reference (teacher_forcing_attention.py:15-27)
        return self._prepare()[1]

    def __call__(self, query, state):
        # teacher_forcing_attention.py:29-35: index = prev_index + 1; alignments = teacher[:, index]
        _, prev_index = state
        index = prev_index + 1
        alignments = self.teacher_alignments[:, index]
        return alignments, TeacherForcingAttentionState(alignments, index)

    def initial_state(self, batch_size, dtype=torch.float32):
        return TeacherForcingAttentionState(self.initial_alignments(batch_size, dtype), -1)


class TeacherForcingForwardAttention(_TeacherForcing):
    kind = "teacher_forcing_forward"


class TeacherForcingAdditiveAttention(_TeacherForcing):
    kind = "teacher_forcing_additive"
