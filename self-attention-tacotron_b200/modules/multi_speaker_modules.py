"""Mirror of /root/reference/modules/multi_speaker_modules.py:11-35."""
from .. import engine


class MultiSpeakerPreNet:

    def __init__(self, out_units, speaker_embed, is_training, drop_rate=0.5, trainable=True, name=None, dtype=None,
                 weights=None, prefix="decoder/prenet/ms", **kwargs):
        if is_training:
            raise NotImplementedError("training-mode dropout is outside the synthesis hot path")
        if weights is None:
            raise ValueError("MultiSpeakerPreNet needs weights=")
        self.out_units = out_units
        self.speaker_embed = speaker_embed
        self.weights, self.prefix = weights, prefix

    def __call__(self, inputs, **kwargs):
        W, p = self.weights, self.prefix
        dense0 = engine.dense_forward(inputs, W[p + "/dense0/kernel"], W[p + "/dense0/bias"], "relu")
        spk = engine.dense_forward(self.speaker_embed, W[p + "/speaker_projection/kernel"],
                                   W[p + "/speaker_projection/bias"], "softsign")
        dense0 = dense0 + spk  # tensor plumbing only; the fused decoder kernel adds this term itself
        return engine.dense_forward(dense0, W[p + "/dense/kernel"], W[p + "/dense/bias"], "relu")

    call = __call__
