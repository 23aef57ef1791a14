"""Plain-Python hyper-parameter set for the mel-synthesis hot path.

Same key names and default values as the reference's global
``tf.contrib.training.HParams`` object (/root/reference/hparams.py:11-225), and
the same two override entry points the reference CLIs use
(``parse_json(file)`` then ``parse("k=v,...")``; train.py:110-115,
predict_mel.py:95-100) -- without TensorFlow.  Only keys that the synthesis
path reads are listed; unknown keys in a JSON file are kept (so the example
JSONs load unchanged) but nothing on the path looks at them.
"""
import ast
import copy
import json

_DEFAULTS = dict(
    # audio / targets
    num_mels=1025, frame_shift_ms=12.5, silence_mel_level_db=-3.0,
    outputs_per_step=1, n_feed_frame=1,
    # symbols / embeddings
    num_symbols=256, embedding_dim=256,
    # model selection
    tacotron_model="DualSourceSelfAttentionTacotronModel",
    encoder="SelfAttentionCBHGEncoder", decoder="DualSourceTransformerDecoder",
    # encoder (CBHG)
    encoder_prenet_drop_rate=0.5, encoder_prenet_out_units=(256, 128),
    cbhg_out_units=256, conv_channels=128, max_filter_width=16,
    projection1_out_channels=128, projection2_out_channels=128, num_highway=4,
    # encoder self-attention
    self_attention_out_units=32, self_attention_num_heads=2, self_attention_num_hop=1,
    self_attention_drop_rate=0.05,
    self_attention_transformer_num_conv_layers=1, self_attention_transformer_kernel_size=5,
    # decoder attention
    attention="additive", forced_alignment_attention="teacher_forcing_additive",
    attention2="additive", forced_alignment_attention2="teacher_forcing_additive",
    attention1_out_units=224, attention2_out_units=32,
    attention_kernel=31, attention_filters=32, cumulative_weights=False,
    use_forward_attention_transition_agent=False,
    # decoder cells
    decoder_prenet_drop_rate=0.5, apply_dropout_on_inference=False,
    decoder_prenet_out_units=(256, 128), attention_out_units=256, decoder_out_units=256,
    decoder_version="v1", zoneout_factor_cell=0.1, zoneout_factor_output=0.1,
    use_zoneout_at_encoder=False,
    # decoder self-attention
    decoder_self_attention_out_units=256, decoder_self_attention_num_heads=2,
    decoder_self_attention_num_hop=1, decoder_self_attention_drop_rate=0.05,
    # speakers
    use_speaker_embedding=False, use_external_speaker_embedding=False,
    speaker_embedding_projection_out_dim=-1, num_speakers=1, speaker_embedding_dim=16,
    speaker_embedding_offset=0, speaker_for_synthesis=-1,
    speaker_embedd_to_prenet=True, speaker_embedd_to_decoder=False, speaker_embedd_to_postnet=False,
    # post-net v2
    use_postnet_v2=False, num_postnet_v2_layers=5, postnet_v2_kernel_size=5,
    postnet_v2_out_channels=512, postnet_v2_drop_rate=0.5,
    # inference
    batch_size=32, max_iters=450, use_forced_alignment_mode=False,
    predicted_mel_extension="mfbsp",
    # bucketing (datasets/*/dataset.py group_by_batch)
    approx_min_target_length=100, batch_bucket_width=50, batch_num_buckets=50,
)


class HParams:
    """Attribute bag with the reference's names (hparams.py:11-225)."""

    def __init__(self, **overrides):
        self.__dict__.update(copy.deepcopy(_DEFAULTS))
        self.override(**overrides)

    def override(self, **kw):
        for k, v in kw.items():
            if isinstance(v, list) and isinstance(_DEFAULTS.get(k), tuple):
                v = tuple(v)
            setattr(self, k, v)
        return self

    def parse_json(self, path_or_text):
        """HParams.parse_json: values from a JSON file (or a JSON string)."""
        text = path_or_text
        if not text.lstrip().startswith("{"):
            with open(path_or_text, encoding="utf-8") as f:
                text = f.read()
        return self.override(**json.loads(text))

    def parse(self, spec):
        """HParams.parse: ``"name=value,name2=value2"`` ad-hoc overrides; an
        unknown name raises ValueError like tf.contrib.training.HParams."""
        spec = (spec or "").strip()
        if not spec:
            return self
        for item in _split_top_level(spec):
            name, _, raw = item.partition("=")
            name = name.strip()
            if not hasattr(self, name):
                raise ValueError(f"Unknown hyperparameter type for {name}")
            cur = getattr(self, name)
            raw = raw.strip()
            if isinstance(cur, bool):
                val = raw.lower() in ("1", "true", "yes")
            elif isinstance(cur, (int, float)):
                val = type(cur)(ast.literal_eval(raw))
            elif isinstance(cur, (tuple, list)):
                val = tuple(ast.literal_eval(raw))
            else:
                val = raw
            setattr(self, name, val)
        return self

    def values(self):
        return dict(self.__dict__)

    def copy(self):
        return HParams(**self.values())


def _split_top_level(spec):
    out, depth, cur = [], 0, ""
    for ch in spec:
        if ch in "([":
            depth += 1
        elif ch in ")]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    if cur:
        out.append(cur)
    return out


def hparams_debug_string(hp):
    vals = hp.values()
    return "Hyperparameters:\n" + "\n".join(f"  {k}: {vals[k]}" for k in sorted(vals))


# ---- the configurations BASELINE.json names ------------------
_LJ_SELF_ATTENTION = dict(
    num_mels=80, outputs_per_step=2, max_iters=500, attention="forward", cumulative_weights=False,
    attention_kernel=10, attention_filters=5, use_zoneout_at_encoder=True, decoder_version="v2",
    tacotron_model="DualSourceSelfAttentionTacotronModel", encoder="SelfAttentionCBHGEncoder",
    decoder="DualSourceTransformerDecoder")


def config(name):
    """Named hyper-parameter sets:
    'ljspeech_self_attention' = examples/ljspeech/self-attention-tacotron.json (cfg 1/4),
    'ljspeech_tacotron'       = examples/ljspeech/tacotron.json (cfg 2),
    'vctk_self_attention'     = examples/vctk/self-attention-tacotron.json (cfg 3),
    'codes_self_attention'    = examples/codes/self-attention-tacotron.json (cfg 5)."""
    if name == "ljspeech_self_attention":
        return HParams(**_LJ_SELF_ATTENTION)
    if name == "ljspeech_tacotron":
        return HParams(**dict(_LJ_SELF_ATTENTION, tacotron_model="ExtendedTacotronV1Model",
                              encoder="ZoneoutEncoderV1", decoder="ExtendedDecoder"))
    if name == "vctk_self_attention":
        return HParams(**dict(_LJ_SELF_ATTENTION, use_speaker_embedding=True, num_speakers=152,
                              speaker_embedding_offset=225))
    if name == "codes_self_attention":
        return HParams(**dict(_LJ_SELF_ATTENTION, num_mels=1025, outputs_per_step=1, max_iters=450,
                              use_zoneout_at_encoder=False, num_speakers=152, speaker_embedding_offset=225))
    raise ValueError(f"Unknown config: {name}")
