"""Plain-Python hyper-parameter set for the mel-synthesis hot path.

Same key names and default values as the reference's global
``tf.contrib.training.HParams`` object (/root/reference/hparams.py:11-225), and
the same two override entry points the reference CLIs use
(``parse_json(file)`` then ``parse("k=v,...")``; train.py:110-115,
predict_mel.py:95-100) -- without TensorFlow.  ``_DEFAULTS`` holds the keys
the synthesis path reads; ``_INERT`` carries every other key of the reference
(training schedule, preprocessing, dataset pipeline, vocoder-parameter
decoders) with its default so that any reference command line or JSON parses
here exactly as it does there -- nothing on the path looks at those values.
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


# keys of hparams.py:11-225 that the synthesis path never reads (kept for config compatibility)
_INERT = dict(
    # audio analysis / feature statistics
    num_mgcs=60, num_freq=2049, sample_rate=48000, frame_length_ms=50.0, ref_level_db=20,
    average_mel_level_db=(0.0,), stddev_mel_level_db=(0.0,), min_mel_level_db=(0.0,),
    mgc_dim=60, mgc_alpha=0.77, mgc_gamma=0.0, mgc_fft_len=4096,
    num_lf0s=256, f0_max=529.0, f0_min=66.0, lf0_loss_factor=0.5,
    trim_top_db=30, trim_frame_length=1024, trim_hop_length=256, num_silent_frames=0,
    # dataset / text front end
    dataset="codes.dataset.DatasetSource", source="phone", source_file_extension="source.tfrecord",
    target_file_extension="target.tfrecord", phoneme="flite",
    flite_binary_path="/home/s1738075/taco_modified/self_attention_tacotron/flite",
    phoneset_path="/home/s1738075/taco_modified/self_attention_tacotron/extensions/phoneset/uscmu_phoneset.json",
    embedding_file="",
    # accent-type encoders, EncoderV2, CBHG post-net, self-attention encoder variants (not on the hot path)
    use_accent_type=False, accent_type_embedding_dim=32, num_accent_type=129, accent_type_offset=0x3100,
    accent_type_unknown=0x3180, accent_type_prenet_out_units=(32, 16), encoder_prenet_out_units_if_accent=(224, 112),
    encoder_v2_num_conv_layers=3, encoder_v2_kernel_size=5, encoder_v2_out_units=512, encoder_v2_drop_rate=0.5,
    self_attention_encoder_out_units=32,
    post_net_cbhg_out_units=256, post_net_conv_channels=128, post_net_max_filter_width=8,
    post_net_projection1_out_channels=256, post_net_projection2_out_channels=80, post_net_num_highway=4,
    # optimiser / schedule / losses
    initial_learning_rate=0.002, decay_learning_rate=True, learning_rate_step_factor=1,
    adam_beta1=0.9, adam_beta2=0.999, adam_eps=1e-8, use_l2_regularization=False, l2_regularization_weight=1e-7,
    code_loss_type="l1",
    # estimator run configuration
    save_summary_steps=50, save_checkpoints_steps=50, keep_checkpoint_max=20000, keep_checkpoint_every_n_hours=1,
    log_step_count_steps=1, alignment_save_steps=50, save_training_time_metrics=False,
    num_evaluation_steps=5, keep_eval_results_max_epoch=10, eval_start_delay_secs=120, eval_throttle_secs=600,
    record_profile=False, profile_steps=50, logfile="log.txt",
    warm_start=False, ckpt_to_initialize_from="", vars_to_warm_start=(".*",),
    # tf.data input pipeline
    interleave_cycle_length_cpu_factor=1.0, interleave_cycle_length_min=4, interleave_cycle_length_max=16,
    interleave_buffer_output_elements=200, interleave_prefetch_input_elements=200, prefetch_buffer_size=4,
    use_cache=False, cache_file_name="", suffle_buffer_size=64,
)


class HParams:
    """Attribute bag with the reference's names and defaults (hparams.py:11-225): all of its keys."""

    def __init__(self, **overrides):
        self.__dict__.update(copy.deepcopy(_INERT))
        self.__dict__.update(copy.deepcopy(_DEFAULTS))
        self.override(**overrides)

    def override(self, **kw):
        for k, v in kw.items():
            if isinstance(v, list) and isinstance(_DEFAULTS.get(k, _INERT.get(k)), tuple):
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
                val = ast.literal_eval(raw)
                val = tuple(val) if isinstance(val, (tuple, list)) else (val,)
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
