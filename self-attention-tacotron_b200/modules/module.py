"""Mirror of the hot-path classes of /root/reference/modules/module.py (+ the tacotron2 layers they
import): encoders, decoders, SelfAttentionTransformer, PreNet, Embedding, PostNetV2.

Constructor and call keyword arguments follow the reference; every class additionally takes
``weights`` (flat blob, see sat_b200.weights) because there is no TF variable store."""
import copy
from collections import namedtuple

import torch

from .. import engine, _lib
from . import rnn_wrappers, helpers
from ..hparams import HParams


def _no_training(is_training):
    assert is_training is not None
    if is_training:
        raise NotImplementedError("the training graph is outside the synthesis hot path (no silent fallback)")


class Embedding:
    """tacotron2 Embedding (models/models.py:28,43-50): table[ids - index_offset]."""

    def __init__(self, num_symbols, embedding_dim, index_offset=0, weights=None, name="embedding"):
        if weights is None:
            raise ValueError("Embedding needs weights=")
        self.table = weights[name]
        assert self.table.shape == (num_symbols, embedding_dim)
        self.index_offset = index_offset

    def __call__(self, ids):
        return engine.embedding_lookup(self.table, ids, self.index_offset)


class PreNet:
    """tacotron2 PreNet (call sites modules/module.py:394,574-576), inference form."""

    def __init__(self, out_units, is_training, drop_rate=0.5, apply_dropout_on_inference=False, dtype=None,
                 weights=None, prefix=None):
        _no_training(is_training)
        if apply_dropout_on_inference:
            raise NotImplementedError("apply_dropout_on_inference=True is random; not on the parity path")
        self.out_units = out_units
        self.weights, self.prefix = weights, prefix

    def __call__(self, inputs):
        W, p = self.weights, self.prefix
        return engine.dense_forward(inputs, W[p + "/kernel"], W[p + "/bias"], "relu")


class SelfAttentionTransformer:
    """modules/module.py:345-371."""

    def __init__(self, is_training, out_units=32, num_conv_layers=1, kernel_size=5, self_attention_out_units=256,
                 self_attention_num_heads=2, self_attention_drop_rate=0.05, use_subsequent_mask=False,
                 trainable=True, name=None, dtype=None, weights=None, prefix=None, **kwargs):
        _no_training(is_training)
        if out_units != self_attention_out_units:
            raise ValueError("the residual add needs out_units == self_attention_out_units")
        self.num_heads = self_attention_num_heads
        self.use_subsequent_mask = use_subsequent_mask
        self.weights, self.prefix = weights, prefix

    def __call__(self, inputs, memory_sequence_length=None):
        out, probs = engine.sat_forward(self.weights, self.prefix, inputs, self.num_heads,
                                        causal=self.use_subsequent_mask)
        return out, [probs[:, i] for i in range(self.num_heads)]

    call = __call__


def _encoder_hparams(kind, cbhg_out_units, conv_channels, max_filter_width, projection1_out_channels,
                     projection2_out_channels, num_highway, prenet_out_units, zoneout_factor_cell,
                     zoneout_factor_output, embedding_dim, num_symbols, **sa):
    return HParams(encoder=kind, cbhg_out_units=cbhg_out_units, conv_channels=conv_channels,
                   max_filter_width=max_filter_width, projection1_out_channels=projection1_out_channels,
                   projection2_out_channels=projection2_out_channels, num_highway=num_highway,
                   encoder_prenet_out_units=tuple(prenet_out_units), zoneout_factor_cell=zoneout_factor_cell,
                   zoneout_factor_output=zoneout_factor_output, embedding_dim=embedding_dim, num_symbols=num_symbols,
                   decoder_version="v2", **sa)


class ZoneoutEncoderV1:
    """modules/module.py:293-342.  ``inputs`` is either the embedded sequence [B,T,E] (reference contract) or
    int64 ids [B,T] (embedding gather fused into the first GEMM)."""

    def __init__(self, is_training, cbhg_out_units=256, conv_channels=128, max_filter_width=16,
                 projection1_out_channels=128, projection2_out_channels=128, num_highway=4,
                 prenet_out_units=(256, 128), drop_rate=0.5, use_zoneout=False, zoneout_factor_cell=0.0,
                 zoneout_factor_output=0.0, lstm_impl=None, trainable=True, name=None, dtype=None, weights=None,
                 **kwargs):
        _no_training(is_training)
        if weights is None:
            raise ValueError("ZoneoutEncoderV1 needs weights=")
        if not use_zoneout:
            zoneout_factor_cell = zoneout_factor_output = 0.0  # plain CBHG == zoneout factors 0 at inference
        table = weights["embedding"]
        self.hp = _encoder_hparams("ZoneoutEncoderV1", cbhg_out_units, conv_channels, max_filter_width,
                                   projection1_out_channels, projection2_out_channels, num_highway, prenet_out_units,
                                   zoneout_factor_cell, zoneout_factor_output, table.shape[1], table.shape[0])
        self._h = engine.EncoderHandle(self.hp, weights)

    mask_padding = False

    def set_mask_padding(self, on):
        """see engine.EncoderHandle.set_mask_padding"""
        self._h.set_mask_padding(on)
        self.mask_padding = bool(on)

    def check(self):
        self._h.check()

    def poll_error(self):
        self._h.poll_error()

    def __call__(self, inputs, input_lengths=None, **kwargs):
        if input_lengths is None:
            raise ValueError("input_lengths is required")
        return self._h.forward(inputs, input_lengths)[0]

    call = __call__


class SelfAttentionCBHGEncoder:
    """modules/module.py:374-441 -> (lstm_output, self_attention_output, [alignment per head per hop])."""

    def __init__(self, is_training, cbhg_out_units=224, conv_channels=128, max_filter_width=16,
                 projection1_out_channels=128, projection2_out_channels=128, num_highway=4,
                 self_attention_out_units=32, self_attention_num_heads=2, self_attention_num_hop=1,
                 self_attention_transformer_num_conv_layers=1, self_attention_transformer_kernel_size=5,
                 prenet_out_units=(256, 128), drop_rate=0.5, zoneout_factor_cell=0.0, zoneout_factor_output=0.0,
                 self_attention_drop_rate=0.1, lstm_impl=None, trainable=True, name=None, dtype=None, weights=None,
                 **kwargs):
        _no_training(is_training)
        if weights is None:
            raise ValueError("SelfAttentionCBHGEncoder needs weights=")
        table = weights["embedding"]
        self.hp = _encoder_hparams("SelfAttentionCBHGEncoder", cbhg_out_units, conv_channels, max_filter_width,
                                   projection1_out_channels, projection2_out_channels, num_highway, prenet_out_units,
                                   zoneout_factor_cell, zoneout_factor_output, table.shape[1], table.shape[0],
                                   self_attention_out_units=self_attention_out_units,
                                   self_attention_num_heads=self_attention_num_heads,
                                   self_attention_num_hop=self_attention_num_hop)
        self._h = engine.EncoderHandle(self.hp, weights)

    mask_padding = False

    def set_mask_padding(self, on):
        """see engine.EncoderHandle.set_mask_padding"""
        self._h.set_mask_padding(on)
        self.mask_padding = bool(on)

    def check(self):
        self._h.check()

    def poll_error(self):
        self._h.poll_error()

    def __call__(self, inputs, input_lengths=None, want_alignments=True, **kwargs):
        if input_lengths is None:
            raise ValueError("input_lengths is required")
        lstm_out, sa_out, probs = self._h.forward(inputs, input_lengths, want_probs=want_alignments)
        aligns = [probs[i] for i in range(probs.shape[0])] if probs is not None else []
        return lstm_out, sa_out, aligns

    call = __call__


class PostNetV2:
    """tacotron2 PostNetV2 (models/models.py:4; definition absent from the reference tree)."""

    def __init__(self, out_units, num_postnet_layers, kernel_size, out_channels, is_training, drop_rate=0.5,
                 weights=None, **kwargs):
        _no_training(is_training)
        hp = HParams(num_mels=out_units, num_postnet_v2_layers=num_postnet_layers, postnet_v2_kernel_size=kernel_size,
                     postnet_v2_out_channels=out_channels, decoder_version="v2", use_postnet_v2=True)
        self._h = engine.PostnetHandle(hp, weights)

    def __call__(self, inputs):
        return self._h.forward(inputs)


# ---- decoders ---------------------------------------------------------------------------------

class AttentionHistoryState(namedtuple("AttentionHistoryState", ["alignment_history", "stop_steps", "n_steps",
                                                                 "alignments"])):
    """What the hot path reads out of the reference's final decoder state (models/models.py:107-115,143-144):
    per-mechanism alignment history, here already stacked as [B, S, T_enc], and ``alignments`` = the decoder
    self-attention probabilities per hop and head, each [B, S, S] (``decoder_state.alignments`` of the reference)."""


class _DecoderBase:
    kind = None
    n_mech = 1

    def __init__(self, prenet_out_units=(256, 128), drop_rate=0.5, attention_out_units=256, decoder_version="v1",
                 decoder_out_units=256, num_mels=80, outputs_per_step=2, max_iters=200, n_feed_frame=1,
                 zoneout_factor_cell=0.0, zoneout_factor_output=0.0, lstm_impl=None, trainable=True, name=None,
                 dtype=None, weights=None, attention_rnn_out_units=None, **kwargs):
        if weights is None:
            raise ValueError(f"{type(self).__name__} needs weights=")
        # the dual-source decoders of the reference call this width attention_rnn_out_units (modules/module.py:1054,1458)
        if attention_rnn_out_units is not None:
            attention_out_units = attention_rnn_out_units
        extra = kwargs.pop("_extra", {})
        if kwargs:
            raise TypeError(f"{type(self).__name__}() got unexpected keyword arguments {sorted(kwargs)}")
        self._base = dict(decoder=self.kind, decoder_prenet_out_units=tuple(prenet_out_units),
                          attention_out_units=attention_out_units, decoder_version=decoder_version,
                          decoder_out_units=decoder_out_units, num_mels=num_mels, outputs_per_step=outputs_per_step,
                          max_iters=max_iters, n_feed_frame=n_feed_frame, zoneout_factor_cell=zoneout_factor_cell,
                          zoneout_factor_output=zoneout_factor_output)
        self._base.update(extra)
        self.weights = weights
        self.want_self_attention_alignments = False  # [B,H,S,S] per hop: only on request
        self.max_iters = max_iters
        self.outputs_per_step = outputs_per_step
        self._handles = {}
        self.stop_mode = "all_rows"  # reference rule (helpers.py:156-160); "per_row" retires rows individually
        self.async_results = False   # True: a free run returns full-length tensors and n_steps as a device tensor
        self.retire_rows = True      # "per_row": stopped rows leave the row-split phases (False: they run on; same results)

    def _handle(self, mechs, speaker_embed):
        kinds = tuple(m.kind for m in mechs)
        o = mechs[0].options
        key = (kinds, tuple(m.num_units for m in mechs), tuple(sorted(o.items())), speaker_embed is not None,
               tuple(m.memory.shape[2] for m in mechs))
        if key not in self._handles:
            real = [k for k in kinds if not k.startswith("teacher_forcing")]
            hp = HParams(**self._base)
            o = dict(o)
            for i in (1, 2):  # teacher-forcing mechanisms carry no conv options: read them off the blob
                lc = self.weights.get(f"decoder/attention{i}/location_conv/kernel")
                if lc is not None:
                    o.setdefault("attention_kernel", lc.shape[0])
                    o.setdefault("attention_filters", lc.shape[2])
            # the weight blob fixes the trained mechanism kinds; teacher forcing only switches the run mode
            kw = dict(attention=kinds[0] if real and not kinds[0].startswith("teacher_forcing") else
                      self._trained_kind(1),
                      attention_kernel=int(o.get("attention_kernel", hp.attention_kernel)),
                      attention_filters=int(o.get("attention_filters", hp.attention_filters)),
                      cumulative_weights=bool(o.get("cumulative_weights", False)),
                      use_forward_attention_transition_agent=bool(o.get("use_transition_agent", False)),
                      use_speaker_embedding=speaker_embed is not None, speaker_embedd_to_prenet=True)
            if self.n_mech == 1:
                kw.update(attention_out_units=mechs[0].num_units, cbhg_out_units=mechs[0].memory.shape[2])
            else:
                kw.update(attention2=kinds[1] if not kinds[1].startswith("teacher_forcing") else self._trained_kind(2),
                          attention1_out_units=mechs[0].num_units, attention2_out_units=mechs[1].num_units,
                          cbhg_out_units=mechs[0].memory.shape[2], self_attention_out_units=mechs[1].memory.shape[2])
            if speaker_embed is not None:
                kw.update(speaker_embedding_dim=speaker_embed.shape[1])
            hp.override(**kw)
            self._handles[key] = (engine.DecoderHandle(hp, self.weights), hp)
        return self._handles[key]

    def _cell_stack(self, mechs):
        """The RNNCell wrapper stack of the reference for this decoder (modules/rnn_wrappers.py), as a declarative
        description: rnn_wrappers.compose() maps it to the persistent-kernel form that implements it."""
        cell = "DecoderRNNV2"
        if self.kind == "DualSourceTransformerDecoder":
            cell = rnn_wrappers.TransformerWrapper(rnn_wrappers.RNNStateHistoryWrapper(cell, self.max_iters),
                                                   self._base.get("decoder_self_attention_num_hop", 1),
                                                   mechs[0].memory_sequence_length)
        return rnn_wrappers.OutputAndStopTokenTransparentWrapper(cell, self._base["num_mels"] * self.outputs_per_step,
                                                                 "out_projection", "stop_token_projection")

    def _trained_kind(self, i):
        p = f"decoder/attention{i}"
        if p + "/location_conv/kernel" in self.weights:
            return "location_sensitive"  # same weights as "forward"; the replay never evaluates them
        return "additive"

    def _run(self, mechs, speaker_embed, is_training, is_validation, teacher_forcing, target, softmax_feed):
        _no_training(is_training)
        h, hp = self._handle(mechs, speaker_embed)
        forced = all(m.kind.startswith("teacher_forcing") for m in mechs)
        if any(m.kind.startswith("teacher_forcing") for m in mechs) and not forced:
            raise ValueError("teacher-forcing attention must be selected for every source or for none")
        r = hp.outputs_per_step
        kw = dict(speaker_embed=speaker_embed, forced_alignment=forced, want_sample_ids=True,
                  want_self_attention_probs=self.want_self_attention_alignments)
        if forced:
            kw["teacher_alignments"] = [m.teacher_alignments for m in mechs]
        # the cell composition the reference would build (modules/module.py:767-775 / :599-600) selects the kernel form
        assert rnn_wrappers.compose(self._cell_stack(mechs)) in (None, self.kind)
        # the seq2seq Helper the reference would construct (modules/module.py:604-617, 728-741) carries the feed and
        # stop rule of the run
        B = mechs[0].memory.shape[0]
        if is_validation:
            if target is None:
                raise ValueError("validation mode needs target")
            cls = helpers.OneHotValidationHelper if softmax_feed else helpers.ValidationHelper
            helper = cls(target, B, hp.num_mels, r, n_feed_frame=hp.n_feed_frame, teacher_forcing=teacher_forcing)
            out = h.run([m.memory for m in mechs], mechs[0].memory_sequence_length, target=target,
                        feed_mode=helper.feed_mode, stop_mode=helper.stop_mode,
                        max_steps=min(helper.num_steps, self.max_iters), **kw)
        else:
            helper = helpers.StopTokenBasedInferenceHelper(B, hp.num_mels, r, n_feed_frame=hp.n_feed_frame)
            stop = helper.stop_mode if self.stop_mode == "all_rows" else _lib.STOP_PER_ROW
            h.set_row_retirement(self.retire_rows)
            out = h.run([m.memory for m in mechs], mechs[0].memory_sequence_length, feed_mode=helper.feed_mode,
                        stop_mode=stop, max_steps=self.max_iters, min_iters=helper.min_iters, **kw)
        if is_validation:
            n = n_ret = min(helper.num_steps, self.max_iters)  # fixed step count: no host sync needed
        elif self.async_results:
            # free run without a host sync: full-length buffers, the step count stays on the device (frames past a
            # row's stop step / past n_steps are unspecified); the caller slices after its own read-back
            n, n_ret = self.max_iters, out["n_steps"]
        else:
            n = n_ret = int(out["n_steps"].item())  # the one host sync of a run: how many steps the stop rule allowed
        mel = out["mel"][:, :n * r]
        stop_token = out["stop"][:, :n, None]
        sa = [p[:, hd, :n, :n] for p in out["self_attention_probs"] for hd in range(p.shape[1])]
        state = AttentionHistoryState([a[:, :n] for a in out["alignments"]], out["stop_steps"], n_ret, sa)
        return mel, stop_token, out["sample_ids"][:, :n], state


class ExtendedDecoder(_DecoderBase):
    """modules/module.py:530-623 -> (mel_output [B,T_dec,M], stop_token [B,S,1], final_state)."""
    kind = "ExtendedDecoder"
    n_mech = 1

    def __call__(self, source, attention_fn=None, speaker_embed=None, is_training=None, is_validation=None,
                 teacher_forcing=False, memory_sequence_length=None, target_sequence_length=None, target=None,
                 teacher_alignments=None, apply_dropout_on_inference=None):
        assert is_training is not None
        assert attention_fn is not None
        m = attention_fn(source, memory_sequence_length, teacher_alignments, weights=self.weights,
                         prefix="decoder/attention1")
        mel, stop, _, state = self._run([m], speaker_embed, is_training, is_validation, teacher_forcing, target, False)
        return mel, stop, state

    call = __call__


class DualSourceDecoder(_DecoderBase):
    """modules/module.py:1051-1152."""
    kind = "DualSourceDecoder"
    n_mech = 2
    _softmax_feed = False

    def __call__(self, sources, attention1_fn=None, attention2_fn=None, speaker_embed=None, is_training=None,
                 is_validation=None, teacher_forcing=False, memory_sequence_length=None,
                 memory2_sequence_length=None, target_sequence_length=None, target=None,
                 teacher_alignments=(None, None), apply_dropout_on_inference=None):
        assert is_training is not None
        assert attention1_fn is not None
        assert attention2_fn is not None
        source1, source2 = sources
        m1 = attention1_fn(source1, memory_sequence_length, teacher_alignments[0], weights=self.weights,
                           prefix="decoder/attention1")
        m2 = attention2_fn(source2, memory2_sequence_length if memory2_sequence_length is not None
                           else memory_sequence_length, teacher_alignments[1], weights=self.weights,
                           prefix="decoder/attention2")
        return self._finish(self._run([m1, m2], speaker_embed, is_training, is_validation, teacher_forcing, target,
                                      self._softmax_feed))

    def _finish(self, res):
        mel, stop, _, state = res
        return mel, stop, state

    call = __call__


class DualSourceTransformerDecoder(DualSourceDecoder):
    """modules/module.py:1455-1562 (+ RNNTransformer :678-784, KV-cached decoder self-attention)
    -> (output [B,T_dec,M], stop_token, predicted_samples [B,S,r] int32, final_state)."""
    kind = "DualSourceTransformerDecoder"
    _softmax_feed = True  # OneHotValidationHelper feeds softmax probabilities (helpers.py:98-106)

    def __init__(self, *args, self_attention_out_units=256, self_attention_num_heads=2, self_attention_num_hop=1,
                 self_attention_transformer_num_conv_layers=1, self_attention_transformer_kernel_size=5,
                 self_attention_drop_rate=0.05, **kwargs):
        kwargs["_extra"] = dict(decoder_self_attention_out_units=self_attention_out_units,
                                decoder_self_attention_num_heads=self_attention_num_heads,
                                decoder_self_attention_num_hop=self_attention_num_hop)
        super().__init__(*args, **kwargs)

    def _finish(self, res):
        return res


def copy_hparams(hp, **kw):
    h = copy.deepcopy(hp)
    return h.override(**kw)
