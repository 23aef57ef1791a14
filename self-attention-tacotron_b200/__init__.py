"""B200-native mel-synthesis hot path of self-attention Tacotron.

Host side = thin Python mirror of the reference's ``modules/`` call surface over
the C ABI of ``libsat.so`` (include/sat.h); all arithmetic runs in hand-written
sm_100a CUDA kernels (csrc/).  There is no CPU fallback: any op raises if the
library or a CUDA device is missing.
"""
from .hparams import HParams, config, hparams_debug_string  # noqa: F401
from .weights import init_weights, weight_shapes, decoder_param_count  # noqa: F401
from . import modules, models, sharding, predict_mel, checkpoint  # noqa: F401,E402
