"""GPU parity of the encoder, stand-alone attention modules and post-net (through the C ABI) against the oracle.
Tolerance 1e-3 relative (max|d| / max|ref|), BASELINE.json north_star; the oracle runs in float64."""
import numpy as np
import pytest

import sat_b200
from oracle import encoder as oenc, nn as onn, postnet as opost, attention as oatt
from tests.util import rel_err, small_hp

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _t(a):
    import torch
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")


def _w64(W):
    return {k: v.astype(np.float64) for k, v in W.items()}


def _enc_inputs(hp, B, T, seed, ragged=True):
    rng = np.random.default_rng(seed)
    ids = rng.integers(1, 71, size=(B, T)).astype(np.int64)
    lengths = np.full(B, T, dtype=np.int64)
    if ragged and B > 1:
        lengths[1:] = rng.integers(max(1, T // 3), T + 1, size=B - 1)
    for b in range(B):
        ids[b, lengths[b]:] = 0  # batch padding id (datasets/ljspeech/dataset.py:268)
    return ids, lengths


@pytest.mark.parametrize("B,T", [(1, 5), (3, 37), (9, 64)])
def test_self_attention_cbhg_encoder(B, T):
    from sat_b200.engine import EncoderHandle
    hp = small_hp("ljspeech_self_attention")
    W = sat_b200.init_weights(hp, seed=B, stress=True)
    ids, lengths = _enc_inputs(hp, B, T, seed=B)
    ref_l, ref_s, ref_a = oenc.self_attention_cbhg_encoder(ids, lengths, _w64(W), hp)
    h = EncoderHandle(hp, W)
    lstm_out, sa_out, probs = h.forward(_t(ids), _t(lengths), want_probs=True)
    assert rel_err(lstm_out.cpu().numpy(), ref_l) <= TOL
    assert rel_err(sa_out.cpu().numpy(), ref_s) <= TOL
    p = probs.cpu().numpy()
    assert p.shape[0] == len(ref_a)
    for i, a in enumerate(ref_a):
        assert rel_err(p[i], a) <= TOL
    # padded tail of the BiLSTM output is exactly zero (TF dynamic_rnn)
    lo = lstm_out.cpu().numpy()
    for b in range(B):
        assert np.all(lo[b, lengths[b]:] == 0)


@pytest.mark.parametrize("tc", ["1", "0"])
def test_encoder_tensor_core_and_simt_conv_paths(tc, monkeypatch):
    """The conv bank / max-pool / proj1 run as tcgen05 implicit GEMMs over fp16 activations by default and as fp32
    SIMT GEMMs with SAT_ENC_TC=0; both must meet the 1e-3 bar against the float64 oracle at the real layer sizes."""
    from sat_b200.engine import EncoderHandle
    monkeypatch.setenv("SAT_ENC_TC", tc)
    hp = small_hp("ljspeech_tacotron")
    W = sat_b200.init_weights(hp, seed=5, stress=True)
    ids, lengths = _enc_inputs(hp, 6, 83, seed=5)
    ref = oenc.zoneout_encoder_v1(ids, lengths, _w64(W), hp)
    h = EncoderHandle(hp, W)
    assert h.uses_tensor_cores() == (tc == "1")
    out = h.forward(_t(ids), _t(lengths))[0].cpu().numpy()
    e = rel_err(out, ref)
    print("encoder conv path tensor_cores=%s rel err %.3e" % (tc, e))
    assert e <= TOL


def test_zoneout_encoder_v1_and_embedded_input():
    import torch
    from sat_b200.engine import EncoderHandle
    hp = small_hp("ljspeech_tacotron")
    W = sat_b200.init_weights(hp, seed=2, stress=True)
    ids, lengths = _enc_inputs(hp, 4, 29, seed=2)
    ref = oenc.zoneout_encoder_v1(ids, lengths, _w64(W), hp)
    h = EncoderHandle(hp, W)
    out, sa, _ = h.forward(_t(ids), _t(lengths))
    assert sa is None
    assert rel_err(out.cpu().numpy(), ref) <= TOL
    emb = _t(W["embedding"][ids])
    out2, _, _ = h.forward(emb, _t(lengths))
    # id inputs take the per-symbol prenet table (exact), embedded inputs the two dense layers: same bar for both
    assert rel_err(out2.cpu().numpy(), ref) <= TOL
    assert rel_err(out2.cpu().numpy(), out.cpu().numpy()) <= TOL


def test_encoder_two_hops_and_adjustment_free_dims():
    from sat_b200.engine import EncoderHandle
    hp = small_hp("ljspeech_self_attention", self_attention_num_hop=2, self_attention_out_units=24,
                  self_attention_num_heads=3, max_filter_width=5, num_highway=2)
    W = sat_b200.init_weights(hp, seed=3, stress=True)
    ids, lengths = _enc_inputs(hp, 2, 21, seed=3)
    ref_l, ref_s, ref_a = oenc.self_attention_cbhg_encoder(ids, lengths, _w64(W), hp)
    h = EncoderHandle(hp, W)
    lstm_out, sa_out, probs = h.forward(_t(ids), _t(lengths), want_probs=True)
    assert rel_err(lstm_out.cpu().numpy(), ref_l) <= TOL
    assert rel_err(sa_out.cpu().numpy(), ref_s) <= TOL
    assert probs.shape[0] == 6
    for i, a in enumerate(ref_a):
        assert rel_err(probs[i].cpu().numpy(), a) <= TOL


@pytest.mark.parametrize("tc", ["1", "0"])
@pytest.mark.parametrize("D,H,T,causal", [(32, 2, 17, False), (32, 2, 40, True), (256, 2, 33, True), (24, 3, 9, False),
                                           (12, 2, 7, True), (32, 2, 150, False), (128, 2, 200, True), (96, 2, 256, False),
                                           (256, 2, 400, True), (256, 2, 130, False), (64, 2, 300, True), (192, 2, 257, False)])
def test_self_attention_module(D, H, T, causal, tc, monkeypatch):
    """modules/self_attention.py:131-144 through the SelfAttention mirror.  Head sizes 16..64 with T <= 256 take the
    single-tile tcgen05 / TMEM kernel, longer sequences and heads up to 128 (the decoder's self-attention) the
    two-pass flash-style tcgen05 kernel (SAT_MHA_TC=0 forces the SIMT kernel); other head sizes run the SIMT kernel."""
    from sat_b200.modules import SelfAttention
    monkeypatch.setenv("SAT_MHA_TC", tc)
    rng = np.random.default_rng(D + T)
    W = {}
    for n in ("key", "value", "query", "output"):
        W[f"sa/{n}/kernel"] = rng.normal(0, 0.3, size=(D, D)).astype(np.float32)
        W[f"sa/{n}/bias"] = rng.normal(0, 0.1, size=(D,)).astype(np.float32)
    x = rng.normal(0, 1, size=(3, T, D)).astype(np.float32)
    w64 = {n: (W[f"sa/{n}/kernel"].astype(np.float64), W[f"sa/{n}/bias"].astype(np.float64))
           for n in ("key", "value", "query", "output")}
    ref, ref_p = onn.multi_head_attention(x.astype(np.float64), w64, H, causal=causal)
    layer = SelfAttention(D, H, 0.0, False, use_subsequent_mask=causal, weights=W, prefix="sa")
    out, aligns = layer(_t(x))
    assert rel_err(out.cpu().numpy(), ref) <= TOL
    assert len(aligns) == H
    for i in range(H):
        assert rel_err(aligns[i].cpu().numpy(), ref_p[:, i]) <= TOL


@pytest.mark.parametrize("H,T,dh,causal,masked", [(2, 23, 16, False, False), (2, 150, 16, False, True),
                                                  (2, 70, 128, True, False), (3, 9, 8, True, True),
                                                  (2, 300, 64, True, False)])
def test_scaled_dot_product_attention_mechanism(H, T, dh, causal, masked):
    """modules/self_attention.py:13-86 at its own interface: projected heads [B, H, T, d_head] in, (context, alignment)
    out -- softmax(mask(Q K^T / sqrt(d_head))) V with the padding (:67-78) and subsequent (:80-86) masks."""
    from sat_b200.modules.self_attention import ScaledDotProductAttentionMechanism
    rng = np.random.default_rng(H * T + dh)
    B = 3
    q, k, v = (rng.normal(0, 1, size=(B, H, T, dh)).astype(np.float32) for _ in range(3))
    lengths = np.array([T, max(1, T // 2), max(1, T - 3)], dtype=np.int64)
    s = (q.astype(np.float64) @ k.astype(np.float64).transpose(0, 1, 3, 2)) / np.sqrt(dh)
    ok = np.ones((B, 1, T, T), dtype=bool)
    if masked:
        m = np.arange(T)[None, :] < lengths[:, None]
        ok = ok & (m[:, None, :, None] & m[:, None, None, :])
    if causal:
        ok = ok & np.tril(np.ones((T, T), dtype=bool))[None, None]
    s = np.where(ok, s, -np.inf)
    with np.errstate(invalid="ignore"):
        p = onn.softmax(s, axis=-1)
    ref = p @ v.astype(np.float64)
    mech = ScaledDotProductAttentionMechanism(_t(k), _t(v), H, use_padding_mask=masked, use_subsequent_mask=causal)
    ctx, align = mech(_t(q), memory_sequence_length=_t(lengths) if masked else None)
    assert tuple(ctx.shape) == (B, H, T, dh) and tuple(align.shape) == (B, H, T, T)
    ctx, align = ctx.cpu().numpy(), align.cpu().numpy()
    for b in range(B):
        n = int(lengths[b]) if masked else T  # rows past the length are NaN in the reference (0 / 0): compared apart
        assert rel_err(ctx[b, :, :n], ref[b, :, :n]) <= TOL
        assert rel_err(align[b, :, :n], p[b, :, :n]) <= TOL
        if n < T:
            assert np.isnan(ctx[b, :, n:]).all() and np.isnan(ref[b, :, n:]).all()


@pytest.mark.parametrize("D,H,T", [(16, 2, 11), (32, 2, 300), (256, 2, 70)])
def test_self_attention_padding_mask(D, H, T):
    """use_padding_mask=True (:67-78): rows/columns past the length are -inf; fully masked rows are NaN in the
    reference, so only the valid block is compared.  (SIMT, and both tensor-core kernels.)"""
    from sat_b200.modules import SelfAttention
    rng = np.random.default_rng(5)
    W = {}
    for n in ("key", "value", "query", "output"):
        W[f"sa/{n}/kernel"] = rng.normal(0, 0.3, size=(D, D)).astype(np.float32)
        W[f"sa/{n}/bias"] = rng.normal(0, 0.1, size=(D,)).astype(np.float32)
    x = rng.normal(0, 1, size=(2, T, D)).astype(np.float32)
    lengths = np.array([T, T // 2 + 1], dtype=np.int64)
    w64 = {n: (W[f"sa/{n}/kernel"].astype(np.float64), W[f"sa/{n}/bias"].astype(np.float64))
           for n in ("key", "value", "query", "output")}
    with np.errstate(invalid="ignore"):
        ref, ref_p = onn.multi_head_attention(x.astype(np.float64), w64, H, lengths=lengths, use_padding_mask=True)
    layer = SelfAttention(D, H, 0.0, False, use_padding_mask=True, weights=W, prefix="sa")
    out, aligns = layer(_t(x), memory_sequence_length=_t(lengths))
    out = out.cpu().numpy()
    for b in range(2):
        L = lengths[b]
        assert rel_err(out[b, :L], ref[b, :L]) <= TOL
        assert np.all(np.isnan(out[b, L:])) and np.all(np.isnan(ref[b, L:]))
        for i in range(H):
            assert rel_err(aligns[i].cpu().numpy()[b, :L], ref_p[b, i, :L]) <= TOL


@pytest.mark.parametrize("kind,cumulative,agent", [("forward", False, False), ("forward", True, True),
                                                   ("location_sensitive", True, False), ("additive", False, False)])
def test_attention_mechanism_step(kind, cumulative, agent):
    """ForwardAttention.__call__ (modules/forward_attention.py:88-122) & co through the mechanism mirrors,
    three consecutive stand-alone steps with state threading."""
    from sat_b200.modules import AttentionOptions, attention_mechanism_factory
    rng = np.random.default_rng(11)
    B, T, Dm, U, A = 3, 19, 20, 12, 16
    hp = small_hp("ljspeech_tacotron", attention=kind, attention_out_units=A, cbhg_out_units=Dm,
                  attention_kernel=7, attention_filters=3, cumulative_weights=cumulative,
                  use_forward_attention_transition_agent=agent)
    hp.attention_out_units = U  # num_units of the mechanism for ExtendedDecoder weights
    Wall = sat_b200.init_weights(hp, seed=4, stress=True)
    W = {k: v for k, v in Wall.items() if k.startswith("decoder/attention1/")}
    # query_layer rows = attention LSTM units; rebuild with A rows
    W["decoder/attention1/query_layer/kernel"] = rng.normal(0, 0.3, size=(A, U)).astype(np.float32)
    memory = rng.normal(0, 0.5, size=(B, T, Dm)).astype(np.float32)
    lengths = np.array([T, T - 5, T - 9], dtype=np.int64)
    opt = AttentionOptions(kind, U, 7, 3, False, cumulative, agent)
    mech = attention_mechanism_factory(opt, weights=W, prefix="decoder/attention1")(_t(memory), _t(lengths))
    omech = oatt.attention_mechanism_factory({"attention": kind, "cumulative_weights": cumulative,
                                              "use_transition_agent": agent})(
        _w64(W), "decoder/attention1", memory.astype(np.float64), lengths)
    assert rel_err(mech.keys.cpu().numpy(), omech.keys) <= TOL
    assert rel_err(mech.values.cpu().numpy(), omech.values) <= TOL
    st, ost = mech.initial_state(B), omech.initial_state()
    for step in range(3):
        q = rng.normal(0, 1, size=(B, A)).astype(np.float32)
        al, st = mech(_t(q), st)
        oal, ost = omech(q.astype(np.float64), ost)
        assert rel_err(al.cpu().numpy(), oal) <= TOL, step
        octx = np.einsum("bt,btd->bd", oal, omech.values)
        assert rel_err(mech.last_context.cpu().numpy(), octx) <= TOL


def test_multi_speaker_prenet_module():
    from sat_b200.modules import MultiSpeakerPreNet
    hp = small_hp("vctk_self_attention")
    W = sat_b200.init_weights(hp, seed=6, stress=True)
    rng = np.random.default_rng(6)
    x = rng.normal(0, 1, size=(5, hp.num_mels)).astype(np.float32)
    spk = rng.normal(0, 1, size=(5, hp.speaker_embedding_dim)).astype(np.float32)
    w = {n: (W[f"decoder/prenet/ms/{n}/kernel"].astype(np.float64), W[f"decoder/prenet/ms/{n}/bias"].astype(np.float64))
         for n in ("dense0", "speaker_projection", "dense")}
    ref = onn.multi_speaker_prenet(x.astype(np.float64), spk.astype(np.float64), w)
    out = MultiSpeakerPreNet(hp.decoder_prenet_out_units[0], _t(spk), False, weights=W)(_t(x))
    assert rel_err(out.cpu().numpy(), ref) <= TOL


@pytest.mark.parametrize("tc", ["1", "0"])
@pytest.mark.parametrize("B,T", [(1, 9), (3, 50), (5, 131)])
def test_postnet_v2(B, T, tc, monkeypatch):
    """both conv forms: tcgen05 implicit GEMM over fp16 activations (default) and fp32 SIMT (SAT_POSTNET_TC=0)"""
    from sat_b200.modules import PostNetV2
    monkeypatch.setenv("SAT_POSTNET_TC", tc)
    hp = small_hp("ljspeech_self_attention", use_postnet_v2=True)
    W = sat_b200.init_weights(hp, seed=7, stress=True)
    rng = np.random.default_rng(7)
    mel = rng.normal(0, 1, size=(B, T, hp.num_mels)).astype(np.float32)
    ref = opost.postnet_v2(mel.astype(np.float64), _w64(W), hp)
    net = PostNetV2(hp.num_mels, hp.num_postnet_v2_layers, hp.postnet_v2_kernel_size, hp.postnet_v2_out_channels,
                    False, weights=W)
    out = net(_t(mel))
    assert rel_err(out.cpu().numpy(), ref) <= TOL


def test_is_training_raises_and_unknown_names():
    from sat_b200.modules import SelfAttention, AttentionOptions, attention_mechanism_factory
    with pytest.raises(NotImplementedError):
        SelfAttention(32, 2, 0.1, True, weights={}, prefix="x")
    fn = attention_mechanism_factory(AttentionOptions("bogus", 8, 3, 2, False, False, False), weights={}, prefix="p")
    with pytest.raises(ValueError, match="Unknown attention mechanism"):
        fn(_t(np.zeros((1, 2, 3), np.float32)), _t(np.array([2])))
