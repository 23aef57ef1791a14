"""The wiring of /root/reference/models/models.py:28-148 (embedding -> encoder -> speaker embedding ->
decoder -> forced-alignment second pass -> post-net) as a plain callable, without tf.estimator.

``SynthesisModel.predict`` is what ``estimator.predict`` yields per batch in the reference
(models/models.py:252-277 / predict_mel.py:49-54): mel, mel_postnet, alignments, ..."""
import torch

from ..modules.module import (ZoneoutEncoderV1, SelfAttentionCBHGEncoder, ExtendedDecoder, DualSourceDecoder,
                              DualSourceTransformerDecoder, PostNetV2, Embedding)
from .attention_factories import (attention_factory, dual_source_attention_factory,
                                  force_alignment_attention_factory, force_alignment_dual_source_attention_factory)


def encoder_factory(params, is_training, weights):
    """models/models.py:325-346 (+ ZoneoutEncoderV1 for the baseline-Tacotron JSONs)."""
    if params.encoder == "SelfAttentionCBHGEncoder":
        return SelfAttentionCBHGEncoder(is_training, cbhg_out_units=params.cbhg_out_units,
                                        conv_channels=params.conv_channels, max_filter_width=params.max_filter_width,
                                        projection1_out_channels=params.projection1_out_channels,
                                        projection2_out_channels=params.projection2_out_channels,
                                        num_highway=params.num_highway,
                                        self_attention_out_units=params.self_attention_out_units,
                                        self_attention_num_heads=params.self_attention_num_heads,
                                        self_attention_num_hop=params.self_attention_num_hop,
                                        prenet_out_units=params.encoder_prenet_out_units,
                                        drop_rate=params.encoder_prenet_drop_rate,
                                        zoneout_factor_cell=params.zoneout_factor_cell,
                                        zoneout_factor_output=params.zoneout_factor_output,
                                        self_attention_drop_rate=params.self_attention_drop_rate, weights=weights)
    if params.encoder == "ZoneoutEncoderV1":
        return ZoneoutEncoderV1(is_training, cbhg_out_units=params.cbhg_out_units, conv_channels=params.conv_channels,
                                max_filter_width=params.max_filter_width,
                                projection1_out_channels=params.projection1_out_channels,
                                projection2_out_channels=params.projection2_out_channels,
                                num_highway=params.num_highway, prenet_out_units=params.encoder_prenet_out_units,
                                drop_rate=params.encoder_prenet_drop_rate, use_zoneout=True,
                                zoneout_factor_cell=params.zoneout_factor_cell,
                                zoneout_factor_output=params.zoneout_factor_output, weights=weights)
    raise ValueError(f"Unknown encoder: {params.encoder}")


MAX_BATCH = 256  # rows of one persistent-kernel launch (csrc/decoder_dev.cuh MAX_B); larger batches run in chunks


def decoder_factory(params, weights):
    """models/models.py:349-368 (+ ExtendedDecoder / DualSourceDecoder)."""
    kw = dict(prenet_out_units=params.decoder_prenet_out_units, drop_rate=params.decoder_prenet_drop_rate,
              attention_out_units=params.attention_out_units, decoder_version=params.decoder_version,
              decoder_out_units=params.decoder_out_units, num_mels=params.num_mels,
              outputs_per_step=params.outputs_per_step, max_iters=params.max_iters, n_feed_frame=params.n_feed_frame,
              zoneout_factor_cell=params.zoneout_factor_cell, zoneout_factor_output=params.zoneout_factor_output,
              weights=weights)
    if params.decoder == "DualSourceTransformerDecoder":
        return DualSourceTransformerDecoder(self_attention_out_units=params.decoder_self_attention_out_units,
                                            self_attention_num_heads=params.decoder_self_attention_num_heads,
                                            self_attention_num_hop=params.decoder_self_attention_num_hop,
                                            self_attention_drop_rate=params.decoder_self_attention_drop_rate, **kw)
    if params.decoder == "DualSourceDecoder":
        return DualSourceDecoder(**kw)
    if params.decoder == "ExtendedDecoder":
        return ExtendedDecoder(**kw)
    raise ValueError(f"Unknown decoder: {params.decoder}")


class SynthesisModel:
    """PREDICT / EVAL branch of model_fn (models/models.py:23-148, 252-277)."""

    def __init__(self, params, weights):
        self.params, self.weights = params, weights
        self.encoder = encoder_factory(params, False, weights)
        self.decoder = decoder_factory(params, weights)
        self.dual = params.decoder != "ExtendedDecoder"
        self.speaker_embedding = None
        if params.use_speaker_embedding:
            self.speaker_embedding = Embedding(params.num_speakers, params.speaker_embedding_dim,
                                               index_offset=params.speaker_embedding_offset, weights=weights,
                                               name="speaker_embedding")
        self.postnet = PostNetV2(params.num_mels, params.num_postnet_v2_layers, params.postnet_v2_kernel_size,
                                 params.postnet_v2_out_channels, False, weights=weights) if params.use_postnet_v2 else None

    def _predict_in_chunks(self, source, source_length, target, speaker_id, is_validation, teacher_forcing,
                           want_enc, want_dec):
        """Batches beyond one launch's row limit: consecutive chunks of <= MAX_BATCH rows (rows are independent; the
        all_rows stop rule then applies per chunk), results concatenated with the step / memory axes zero-padded to the
        longest chunk.  n_steps = the largest chunk's; stop_steps stay per row."""
        outs = []
        for b0 in range(0, source.shape[0], MAX_BATCH):
            sl = slice(b0, b0 + MAX_BATCH)
            # the padded length is kept: padded frames are part of the reference's batched graph
            outs.append(self.predict(source[sl], source_length[sl], None if target is None else target[sl],
                                     None if speaker_id is None else speaker_id[sl], is_validation, teacher_forcing,
                                     want_enc, want_dec))
        res = {}
        for k, v in outs[0].items():
            vals = [o[k] for o in outs]
            if k == "n_steps":
                res[k] = max(int(x.item()) if torch.is_tensor(x) else int(x) for x in vals)
            elif k == "encoder_alignments":
                res[k] = [_cat_padded([o[k][i] for o in outs], [1, 2]) for i in range(len(v))]
            elif torch.is_tensor(v):
                res[k] = torch.cat(vals, 0) if v.dim() == 1 else _cat_padded(vals, list(range(1, v.dim())))
        return res

    def predict(self, source, source_length, target=None, speaker_id=None, is_validation=False,
                teacher_forcing=False, want_encoder_alignments=False, want_decoder_self_attention_alignments=False):
        """source int64 ids [B,T_enc] (device), source_length [B]; target [B,T_dec,M] for validation /
        forced alignment.  Returns a dict of device tensors."""
        p = self.params
        if source.shape[0] > MAX_BATCH:
            return self._predict_in_chunks(source, source_length, target, speaker_id, is_validation, teacher_forcing,
                                           want_encoder_alignments, want_decoder_self_attention_alignments)
        self.decoder.want_self_attention_alignments = want_decoder_self_attention_alignments
        if self.dual:
            lstm_out, sa_out, enc_align = self.encoder(source, input_lengths=source_length,
                                                       want_alignments=want_encoder_alignments)
            sources = (lstm_out, sa_out)
        else:
            sources = self.encoder(source, input_lengths=source_length)
            enc_align = []
        spk = None
        if self.speaker_embedding is not None:
            ids = speaker_id
            if p.speaker_for_synthesis > -1:
                ids = torch.full_like(speaker_id, p.speaker_for_synthesis)
            spk = self.speaker_embedding(ids)
            if p.speaker_embedding_projection_out_dim > -1:
                # models/models.py:73-75: resize the speaker embedding with a relu Dense
                from .. import engine
                spk = engine.dense_forward(spk, self.weights["speaker_embedding_projection/kernel"],
                                           self.weights["speaker_embedding_projection/bias"], "relu")
            if p.speaker_embedd_to_decoder:
                # models/models.py:78-83: tile over time, concatenate onto the encoder outputs (tensor plumbing only)
                tiled = spk[:, None, :].expand(-1, sources[0].shape[1], -1) if self.dual else \
                    spk[:, None, :].expand(-1, sources.shape[1], -1)
                sources = tuple(torch.cat((s_, tiled), dim=-1) for s_ in sources) if self.dual else \
                    torch.cat((sources, tiled), dim=-1)
        spk_prenet = spk if p.speaker_embedd_to_prenet else None
        forced = p.use_forced_alignment_mode
        kw = dict(speaker_embed=spk_prenet, is_training=False, is_validation=is_validation or forced,
                  teacher_forcing=teacher_forcing or forced, target=target,
                  apply_dropout_on_inference=p.apply_dropout_on_inference)
        if self.dual:
            a1, a2 = dual_source_attention_factory(p)
            res = self.decoder(sources, attention1_fn=a1, attention2_fn=a2, memory_sequence_length=source_length,
                               memory2_sequence_length=source_length, **kw)
        else:
            res = self.decoder(sources, attention_fn=attention_factory(p), memory_sequence_length=source_length, **kw)
        state = res[-1]
        if forced:
            # second pass: replay pass-1 alignments, feed back own outputs (models/models.py:118-134)
            kw.update(is_validation=True, teacher_forcing=False)
            ta = tuple(state.alignment_history)
            if self.dual:
                f1, f2 = force_alignment_dual_source_attention_factory(p)
                res = self.decoder(sources, attention1_fn=f1, attention2_fn=f2, memory_sequence_length=source_length,
                                   memory2_sequence_length=source_length, teacher_alignments=ta, **kw)
            else:
                res = self.decoder(sources, attention_fn=force_alignment_attention_factory(p),
                                   memory_sequence_length=source_length, teacher_alignments=ta[0], **kw)
            state = res[-1]
        mel, stop = res[0], res[1]
        out = {"mel": mel, "stop_token": stop, "stop_steps": state.stop_steps, "n_steps": state.n_steps,
               "source": source, "encoder_alignments": enc_align}
        # alignments arranged (B, T_memory, T_query) like models/models.py:104-115
        for i, a in enumerate(state.alignment_history):
            out["alignment" if i == 0 else f"alignment{i + 1}"] = a.transpose(1, 2)
        # decoder self-attention alignments -> alignment3/4, encoder self-attention -> alignment5..7, each
        # (B, T_key, T_query) like models/models.py:105,110-111,265-269
        for i, a in enumerate(state.alignments[:2]):
            out[f"alignment{3 + i}"] = a.transpose(1, 2)
        for i, a in enumerate(enc_align[:3]):
            out[f"alignment{5 + i}"] = a.transpose(1, 2)
        if len(res) == 4:
            out["predicted_samples"] = res[2]
        if self.postnet is not None:
            out["mel_postnet"] = self.postnet(mel)
        return out


def _cat_padded(ts, dims):
    """concatenate along the batch axis tensors whose axes `dims` may differ in length (zero padded to the longest)"""
    tgt = [max(t.shape[d] for t in ts) for d in dims]
    out = []
    for t in ts:
        pad = []
        for d in range(t.dim() - 1, 0, -1):  # F.pad lists the last axis first
            pad += [0, tgt[dims.index(d)] - t.shape[d] if d in dims else 0]
        out.append(torch.nn.functional.pad(t, pad) if any(pad) else t)
    return torch.cat(out, dim=0)


def tacotron_model_factory(hparams, weights):
    """models/models.py:371-378: model selection by name; returns the synthesis callable."""
    if hparams.tacotron_model in ("DualSourceSelfAttentionTacotronModel", "ExtendedTacotronV1Model",
                                  "DualSourceTacotronModel"):
        return SynthesisModel(hparams, weights)
    raise ValueError(f"Unknown Tacotron model: {hparams.tacotron_model}")
