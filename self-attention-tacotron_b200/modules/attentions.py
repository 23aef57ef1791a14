"""Mirror of /root/reference/modules/attentions.py:15-62."""
from collections import namedtuple

from .forward_attention import ForwardAttention, LocationSensitiveAttention, BahdanauAttention
from .teacher_forcing_attention import TeacherForcingForwardAttention, TeacherForcingAdditiveAttention


class AttentionOptions(namedtuple("AttentionOptions", ["attention", "num_units", "attention_kernel",
                                                       "attention_filters", "smoothing", "cumulative_weights",
                                                       "use_transition_agent"])):
    pass


def attention_mechanism_factory(options: AttentionOptions, weights=None, prefix=None):
    """-> attention_fn(memory, memory_sequence_length, teacher_alignments=None).  ``weights``/``prefix`` may
    also be given per call (the decoders pass their own blob and ``decoder/attention{1,2}``)."""

    def attention_fn(memory, memory_sequence_length, teacher_alignments=None, weights=weights, prefix=prefix):
        kw = dict(weights=weights, prefix=prefix)
        if options.attention == "forward":
            mechanism = ForwardAttention(num_units=options.num_units, memory=memory,
                                         memory_sequence_length=memory_sequence_length,
                                         attention_kernel=options.attention_kernel,
                                         attention_filters=options.attention_filters,
                                         use_transition_agent=options.use_transition_agent,
                                         cumulative_weights=options.cumulative_weights, **kw)
        elif options.attention == "location_sensitive":
            mechanism = LocationSensitiveAttention(num_units=options.num_units, memory=memory,
                                                   memory_sequence_length=memory_sequence_length,
                                                   attention_kernel=options.attention_kernel,
                                                   attention_filters=options.attention_filters,
                                                   smoothing=options.smoothing,
                                                   cumulative_weights=options.cumulative_weights, **kw)
        elif options.attention == "teacher_forcing_forward":
            mechanism = TeacherForcingForwardAttention(num_units=options.num_units, memory=memory,
                                                       memory_sequence_length=memory_sequence_length,
                                                       teacher_alignments=teacher_alignments, **kw)
        elif options.attention == "teacher_forcing_additive":
            mechanism = TeacherForcingAdditiveAttention(num_units=options.num_units, memory=memory,
                                                        memory_sequence_length=memory_sequence_length,
                                                        teacher_alignments=teacher_alignments, **kw)
        elif options.attention == "additive":
            mechanism = BahdanauAttention(num_units=options.num_units, memory=memory,
                                          memory_sequence_length=memory_sequence_length, **kw)
        else:
            raise ValueError(f"Unknown attention mechanism: {options.attention}")
        return mechanism

    attention_fn.options = options
    return attention_fn
