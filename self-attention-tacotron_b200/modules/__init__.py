"""Host-side mirror of the reference's ``modules/`` call surface (same class names, constructor
and call keyword arguments, error behaviour).  Every class only marshals torch CUDA tensors into
libsat.so; there is no arithmetic here and no CPU fallback.

Differences forced by leaving TensorFlow: variables do not live in a graph, so every layer takes
``weights`` (the flat ``{name: float32 ndarray}`` blob of sat_b200.weights) and a ``prefix``;
``is_training=True`` raises NotImplementedError (the training graph is out of scope)."""
from .self_attention import SelfAttention, MultiHeadAttention, ScaledDotProductAttentionMechanism  # noqa: F401
from .forward_attention import ForwardAttention, ForwardAttentionState  # noqa: F401
from .attentions import AttentionOptions, attention_mechanism_factory  # noqa: F401
from .teacher_forcing_attention import TeacherForcingForwardAttention, TeacherForcingAdditiveAttention  # noqa: F401
from .multi_speaker_modules import MultiSpeakerPreNet  # noqa: F401
from .module import (PreNet, Embedding, ZoneoutEncoderV1, SelfAttentionCBHGEncoder, SelfAttentionTransformer,  # noqa: F401
                     ExtendedDecoder, DualSourceDecoder, DualSourceTransformerDecoder, PostNetV2)
from . import rnn_wrappers, helpers  # noqa: F401
