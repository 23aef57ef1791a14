"""Autoregressive decoder of the oracle (test infrastructure).

One generic ``decoder_run`` restates the three decoders the hot path needs:

* ``ExtendedDecoder``              modules/module.py:562-623   (single source)
* ``DualSourceDecoder``            modules/module.py:1083-1152
* ``DualSourceTransformerDecoder`` modules/module.py:1499-1562 + RNNTransformer :726-784

Step order follows SURVEY.md section 3.2 (AttentionWrapper / DecoderPreNetWrapper /
ConcatOutputAndAttentionWrapper / DecoderRNNV2 / wrappers of modules/rnn_wrappers.py).
"""
import numpy as np
from . import nn
from .encoder import kb, sat_weights
from .attention import attention_mechanism_factory

MIN_ITERS = 10  # StopTokenBasedInferenceHelper(min_iters=10), modules/helpers.py:113


def attention_options(hp, which, forced=False):
    """models/attention_factories.py:11-66.  which = 0 (single source), 1 or 2."""
    if which == 0:
        kind = hp.forced_alignment_attention if forced else hp.attention
        units = hp.attention_out_units
    elif which == 1:
        kind = hp.forced_alignment_attention if forced else hp.attention
        units = hp.attention1_out_units
    else:
        kind = hp.forced_alignment_attention2 if forced else hp.attention2
        units = hp.attention2_out_units
    return {"attention": kind, "num_units": units,
            "attention_kernel": hp.attention_kernel, "attention_filters": hp.attention_filters,
            "smoothing": False, "cumulative_weights": hp.cumulative_weights,
            "use_transition_agent": hp.use_forward_attention_transition_agent}


def _decoder_prenet(x, W, hp, speaker_embed):
    """modules/module.py:1511-1517 / :568-577."""
    if speaker_embed is not None:
        w = {n: kb(W, f"decoder/prenet/ms/{n}") for n in ("dense0", "speaker_projection", "dense")}
        x = nn.multi_speaker_prenet(x, speaker_embed, w)
        return nn.prenet(x, [kb(W, "decoder/prenet/dense_1")])
    return nn.prenet(x, [kb(W, f"decoder/prenet/dense_{i}") for i in range(len(hp.decoder_prenet_out_units))])


def _mha_last_row_cached(cache, d_out, w, num_heads):
    """KV-cached evaluation of the last row of a causal SelfAttentionTransformer
    (identical to re-running modules/rnn_wrappers.py:111-124 over the whole
    history and keeping row -1, because the causal mask makes row t depend only
    on rows <= t)."""
    k = nn.dense(d_out, *w["mha"]["key"])
    v = nn.dense(d_out, *w["mha"]["value"])
    q = nn.dense(d_out, *w["mha"]["query"])
    cache["k"].append(k)
    cache["v"].append(v)
    K = np.stack(cache["k"], axis=1)  # [B,t,D]
    V = np.stack(cache["v"], axis=1)
    B, t, D = K.shape
    dh = D // num_heads
    Kh = K.reshape(B, t, num_heads, dh)
    Vh = V.reshape(B, t, num_heads, dh)
    qh = q.reshape(B, num_heads, dh)
    s = np.einsum("bhd,bthd->bht", qh, Kh) / np.sqrt(d_out.dtype.type(dh))
    p = nn.softmax(s, axis=-1)
    ctx = np.einsum("bht,bthd->bhd", p, Vh).reshape(B, D)
    a = nn.dense(ctx, *w["mha"]["output"])
    return d_out + np.tanh(nn.dense(a, *w["transform"])), p


def decoder_run(sources, lengths, W, hp, decoder="DualSourceTransformerDecoder", *,
                target=None, is_validation=False, teacher_forcing=False,
                teacher_alignments=None, speaker_embed=None, forced_alignment=False,
                stop_mode="all_rows", sa_form="cached", max_iters=None, trace=None):
    """Run the decoder.

    sources: memory [B,T,D] (ExtendedDecoder) or (memory1, memory2).
    lengths: memory_sequence_length (both sources share it, models/models.py:92-93).
    target: [B,T_dec,M] when is_validation.
    sa_form: 'cached' | 'history' (literal O(S^3) reference form,
             modules/rnn_wrappers.py:74-80,111-124) | 'batched' (training-branch
             form, modules/module.py:743-765; teacher forcing only).
    stop_mode (free-run only): 'all_rows' = modules/helpers.py:156-160;
             'per_row' = each row's first t > MIN_ITERS with sigmoid(stop) > 0.5.
    Returns dict(mel [B,S*r,M], stop [B,S,1], alignments [list of [B,S,T]],
                 sa_probs, sample_ids [B,S,r], n_steps, stop_steps [B]).
    """
    dual = decoder != "ExtendedDecoder"
    transformer = decoder == "DualSourceTransformerDecoder"
    if not dual:
        sources = (sources,)
    dt = sources[0].dtype
    B = sources[0].shape[0]
    M, r, nf = hp.num_mels, hp.outputs_per_step, hp.n_feed_frame
    max_iters = hp.max_iters if max_iters is None else max_iters
    zc, zh = dt.type(hp.zoneout_factor_cell), dt.type(hp.zoneout_factor_output)
    assert hp.decoder_version == "v2", "oracle restates DecoderRNNV2 only"

    if teacher_alignments is None:
        teacher_alignments = (None,) * len(sources)
    mechs = []
    for i, src in enumerate(sources):
        opt = attention_options(hp, (i + 1) if dual else 0, forced=forced_alignment)
        mechs.append(attention_mechanism_factory(opt)(W, f"decoder/attention{i + 1}", src, lengths,
                                                      teacher_alignments[i]))
    mstates = [m.initial_state() for m in mechs]
    A, Dd = hp.attention_out_units, hp.decoder_out_units
    z = lambda n: np.zeros((B, n), dtype=dt)  # noqa: E731
    st_a, st_1, st_2 = (z(A), z(A)), (z(Dd), z(Dd)), (z(Dd), z(Dd))
    attention = z(sum(m.values.shape[-1] for m in mechs))

    if is_validation:
        S = target.shape[1] // r
        targets = target.reshape(B, S, M * r)
        n_max = min(S, max_iters)
    else:
        n_max = max_iters
    hops = hp.decoder_self_attention_num_hop if transformer else 0
    heads = hp.decoder_self_attention_num_heads
    sa_w = [sat_weights(W, f"decoder/sa_{i}") for i in range(hops)]
    caches = [{"k": [], "v": []} for _ in range(hops)]
    history = []
    # feed rule: OneHotValidationHelper softmaxes (modules/helpers.py:98-106) only for the
    # transformer decoder in validation mode; tacotron2 Validation/StopToken helpers feed raw.
    softmax_feed = transformer and is_validation and not teacher_forcing

    x = z(M * nf)
    outs, stops, aligns, d_outs, sa_probs = [], [], [[] for _ in mechs], [], []
    stop_steps = np.full(B, -1, dtype=np.int64)
    t = 0
    while t < n_max:
        p = _decoder_prenet(x, W, hp, speaker_embed)
        cell_in = np.concatenate([p, attention], axis=-1)
        query, st_a = nn.zoneout_lstm_cell(cell_in, st_a, *kb(W, "decoder/attention_lstm"), zc, zh)
        ctxs = []
        for i, m in enumerate(mechs):
            al, mstates[i] = m(query, mstates[i])
            aligns[i].append(al)
            ctxs.append(np.einsum("bt,btd->bd", al, m.values))
        attention = np.concatenate(ctxs, axis=-1)
        d_in = np.concatenate([query, attention], axis=-1)
        o1, st_1 = nn.zoneout_lstm_cell(d_in, st_1, *kb(W, "decoder/lstm_1"), zc, zh)
        o2, st_2 = nn.zoneout_lstm_cell(o1, st_2, *kb(W, "decoder/lstm_2"), zc, zh)
        y = o2
        d_outs.append(o2)
        if trace is not None:
            trace.append({"prenet": p, "query": query, "attention": attention, "o1": o1, "o2": o2})
        if hops and sa_form == "cached":
            pr = []
            for w, c in zip(sa_w, caches):
                y, pp = _mha_last_row_cached(c, y, w, heads)
                pr.append(pp)
            sa_probs.append(pr)
        elif hops and sa_form == "history":
            history.append(o2)
            h = np.stack(history, axis=1)
            for w in sa_w:
                h, _ = nn.self_attention_transformer(h, w, heads, causal=True)
            y = h[:, -1]
        if hops and sa_form == "batched":
            assert is_validation and teacher_forcing
            out = stop = None
        else:
            out = nn.dense(y, *kb(W, "decoder/out_projection"))
            stop = nn.dense(y, *kb(W, "decoder/stop_token_projection"))
            outs.append(out)
            stops.append(stop)
        # helper.next_inputs (modules/helpers.py:54-58,98-108,156-160)
        if is_validation:
            finished = t + 1 >= S
            if teacher_forcing:
                x = targets[:, t, -M * nf:]
            elif softmax_feed:
                x = nn.softmax(out.reshape(B, r, M), axis=-1).reshape(B, r * M)[:, -M * nf:]
            else:
                x = out[:, -M * nf:]
        else:
            term = (nn.sigmoid(stop[:, 0]) > 0.5) & (t > MIN_ITERS)
            newly = term & (stop_steps < 0)
            stop_steps[newly] = t
            if stop_mode == "never":
                finished = False
            else:
                finished = bool(np.all(term)) if stop_mode == "all_rows" else bool(np.all(stop_steps >= 0))
            x = out[:, -M * nf:]
        t += 1
        if finished:
            break

    n_steps = t
    res = {"n_steps": n_steps, "stop_steps": stop_steps,
           "alignments": [np.stack(a, axis=1) for a in aligns],
           "decoder_outputs": np.stack(d_outs, axis=1)}
    if hops and sa_form == "batched":
        h = res["decoder_outputs"]
        probs = []
        for w in sa_w:
            h, pp = nn.self_attention_transformer(h, w, heads, causal=True)
            probs.append(pp)
        out_all = nn.dense(h, *kb(W, "decoder/out_projection"))
        stop_all = nn.dense(h, *kb(W, "decoder/stop_token_projection"))
        res["sa_probs_batched"] = probs
    else:
        out_all = np.stack(outs, axis=1)
        stop_all = np.stack(stops, axis=1)
        res["sa_probs_steps"] = sa_probs
    res["mel"] = out_all.reshape(B, n_steps * r, M)
    res["stop"] = stop_all
    res["sample_ids"] = np.argmax(out_all.reshape(B, n_steps, r, M), axis=-1).astype(np.int32)
    return res
