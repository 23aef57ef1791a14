from .attention_factories import (attention_factory, dual_source_attention_factory,  # noqa: F401
                                  force_alignment_attention_factory, force_alignment_dual_source_attention_factory)
from .models import encoder_factory, decoder_factory, SynthesisModel, tacotron_model_factory  # noqa: F401
