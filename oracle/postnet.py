"""Post-net and whole-model wiring of the oracle (test infrastructure).

PostNetV2's definition is NOT in /root/reference (only its import,
models/models.py:4, its hparams hparams.py:146-150 and its consumer
predict_mel.py:60); the form below is SURVEY.md Appendix A.8 -- parity for this
stage is unpinned even in principle.
"""
import numpy as np
from . import nn
from .encoder import kb, bn, embedding, zoneout_encoder_v1, self_attention_cbhg_encoder
from .decoder import decoder_run


def postnet_v2(mel, W, hp):
    """A.8: layers 1..n-1 = Conv1d(k, C, tanh) (no bias, BN), layer n = Conv1d(k, C, None),
    then Dense(num_mels); mel_postnet = mel + projected."""
    x = mel
    n = hp.num_postnet_v2_layers
    last_act = {None: None, "tanh": np.tanh}[nn.SWITCHES["postnet_last_activation"]]
    for i in range(n):
        act = np.tanh if i < n - 1 else last_act
        x = nn.conv1d_bn_act(x, W[f"postnet/conv_{i}/kernel"], bn(W, f"postnet/conv_{i}"), act)
    return mel + nn.dense(x, *kb(W, "postnet/projection"))


def synthesize(ids, lengths, W, hp, *, target=None, speaker_ids=None, mode="free_run",
               stop_mode="all_rows", max_iters=None, sa_form="cached"):
    """Embedding -> encoder -> decoder (-> post-net) in the wiring order of
    models/models.py:28-148.

    mode: 'teacher_forced' (is_validation=True, teacher_forcing=True),
          'validation'     (is_validation=True, teacher_forcing=False),
          'free_run'       (stop-token inference),
          'forced_alignment' (two passes, models/models.py:86-97,118-134).
    """
    speaker_embed = None
    if getattr(hp, "use_speaker_embedding", False) and speaker_ids is not None:
        speaker_embed = embedding(W["speaker_embedding"], speaker_ids, hp.speaker_embedding_offset)
        if getattr(hp, "speaker_embedding_projection_out_dim", -1) > -1:  # models/models.py:73-75
            speaker_embed = nn.relu(nn.dense(speaker_embed, *kb(W, "speaker_embedding_projection")))
    if hp.encoder == "SelfAttentionCBHGEncoder":
        lstm_out, sa_out, enc_align = self_attention_cbhg_encoder(ids, lengths, W, hp)
        sources = (lstm_out, sa_out)
    elif hp.encoder == "ZoneoutEncoderV1":
        sources = zoneout_encoder_v1(ids, lengths, W, hp)
        enc_align = []
    else:
        raise ValueError(f"Unknown encoder: {hp.encoder}")
    if speaker_embed is not None and getattr(hp, "speaker_embedd_to_decoder", False):  # models/models.py:78-83
        tile = lambda s: np.concatenate([s, np.repeat(speaker_embed[:, None, :], s.shape[1], axis=1)], axis=-1)  # noqa: E731
        sources = tuple(tile(s) for s in sources) if isinstance(sources, tuple) else tile(sources)
    if speaker_embed is not None and not hp.speaker_embedd_to_prenet:
        speaker_embed = None
    kw = dict(speaker_embed=speaker_embed, stop_mode=stop_mode, max_iters=max_iters, sa_form=sa_form)
    if mode == "teacher_forced":
        res = decoder_run(sources, lengths, W, hp, hp.decoder, target=target, is_validation=True,
                          teacher_forcing=True, **kw)
    elif mode == "validation":
        res = decoder_run(sources, lengths, W, hp, hp.decoder, target=target, is_validation=True,
                          teacher_forcing=False, **kw)
    elif mode == "free_run":
        res = decoder_run(sources, lengths, W, hp, hp.decoder, **kw)
    elif mode == "forced_alignment":
        p1 = decoder_run(sources, lengths, W, hp, hp.decoder, target=target, is_validation=True,
                         teacher_forcing=True, **kw)
        res = decoder_run(sources, lengths, W, hp, hp.decoder, target=target, is_validation=True,
                          teacher_forcing=False, teacher_alignments=tuple(p1["alignments"]),
                          forced_alignment=True, **kw)
        res["pass1"] = p1
    else:
        raise ValueError(mode)
    res["encoder_alignments"] = enc_align
    res["sources"] = sources
    if getattr(hp, "use_postnet_v2", False):
        res["mel_postnet"] = postnet_v2(res["mel"], W, hp)
    return res
