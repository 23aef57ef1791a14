"""GPU parity of the persistent decoder kernel (through the C ABI) against the oracle.

Tolerance: BASELINE.json north_star -- teacher-forced mel / stop / alignments within 1e-3
relative (max|d| / max|ref|); free-run: identical stop step and alignment argmax.
The oracle runs in float64 on the same float32 weights and inputs.
"""
import numpy as np
import pytest

import sat_b200
from tests.util import rel_err, small_hp, decoder_inputs, oracle_decoder

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _torch():
    import torch
    return torch


def run_cuda(hp, W, mems, lengths, **kw):
    torch = _torch()
    from sat_b200.engine import DecoderHandle
    dev = torch.device("cuda:0")
    h = DecoderHandle(hp, W)
    t = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    if "teacher_alignments" in kw and kw["teacher_alignments"] is not None:
        kw["teacher_alignments"] = [t(a) for a in kw["teacher_alignments"]]
    for k in ("target", "speaker_embed"):
        if k in kw:
            kw[k] = t(kw[k])
    out = h.run([t(m) for m in mems], t(lengths), **kw)
    torch.cuda.synchronize()
    n = int(out["n_steps"].item())
    res = {"n_steps": n, "tensor_cores": h.used_tensor_cores(), "cluster": h.cluster_size(), "stop_steps": out["stop_steps"].cpu().numpy(),
           "mel": out["mel"].cpu().numpy()[:, :n * hp.outputs_per_step],
           "stop": out["stop"].cpu().numpy()[:, :n, None],
           "alignments": [a.cpu().numpy()[:, :n] for a in out["alignments"] if a is not None]}
    if out["sample_ids"] is not None:
        res["sample_ids"] = out["sample_ids"].cpu().numpy()[:, :n]
    if out["decoder_outputs"] is not None:
        res["decoder_outputs"] = out["decoder_outputs"].cpu().numpy()[:, :n]
    return res


def check_close(got, ref, keys=("mel", "stop"), tol=TOL):
    assert got["n_steps"] == ref["n_steps"]
    for k in keys:
        assert got[k].shape == ref[k].shape, (k, got[k].shape, ref[k].shape)
        e = rel_err(got[k], ref[k])
        assert e <= tol, f"{k}: rel err {e:.3e} > {tol}"
    for i, (a, b) in enumerate(zip(got["alignments"], ref["alignments"])):
        e = rel_err(a, b)
        assert e <= tol, f"alignment{i}: rel err {e:.3e} > {tol}"


SMALL = dict(attention_out_units=32, decoder_out_units=32, decoder_self_attention_out_units=32,
             attention1_out_units=24, attention2_out_units=8, cbhg_out_units=16, self_attention_out_units=8,
             decoder_prenet_out_units=(24, 16), num_mels=12)


@pytest.mark.parametrize("B,T,S", [(1, 7, 5), (3, 23, 12), (16, 40, 9)])
def test_extended_decoder_teacher_forced_small(B, T, S):
    hp = small_hp("ljspeech_tacotron", **SMALL)
    W = sat_b200.init_weights(hp, seed=B, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, T, S, seed=B)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=S,
                   want_decoder_outputs=True)
    check_close(got, ref)
    assert rel_err(got["decoder_outputs"], ref["decoder_outputs"]) <= TOL


def test_extended_decoder_full_dims_teacher_forced():
    hp = small_hp("ljspeech_tacotron")
    W = sat_b200.init_weights(hp, seed=7, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 2, 150, 40, seed=7)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=40)
    check_close(got, ref)


@pytest.mark.parametrize("name,B,T,S,mode", [("ljspeech_self_attention", 16, 150, 24, "teacher"),
                                             ("ljspeech_self_attention", 5, 77, 16, "teacher"),
                                             ("ljspeech_tacotron", 16, 200, 30, "free"),
                                             ("ljspeech_tacotron", 33, 61, 12, "free"),
                                             ("ljspeech_self_attention", 9, 100, 10, "replay")])
def test_cluster_split_attention_rows(name, B, T, S, mode, monkeypatch):
    """Small batches: every (row, mechanism) attention item is split over a thread-block cluster (query-layer columns,
    attention units of the energies, context columns; partial energies exchanged through distributed shared memory).
    Same bar as the one-CTA-per-row form, which SAT_CLUSTER=0 selects and which must agree to fp32 rounding."""
    hp = small_hp(name)
    W = sat_b200.init_weights(hp, seed=21, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, T, S, seed=21)
    kw = dict(max_steps=S)
    if mode == "free":
        kw.update(feed_mode=1, stop_mode=0)
        ref = oracle_decoder(hp, W, mems, lengths, max_iters=S, stop_mode="never")
    elif mode == "teacher":
        kw.update(target=target, feed_mode=0, stop_mode=0)
        ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    else:  # forced-alignment replay (models/models.py:118-134): the alignments are given, only the context is split
        p1 = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
        ta = [a.astype(np.float32) for a in p1["alignments"]]
        ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=False,
                             teacher_alignments=ta, forced_alignment=True)
        kw.update(target=target, teacher_alignments=ta, forced_alignment=True, feed_mode=2, stop_mode=0)
    got = run_cuda(hp, W, mems, lengths, **kw)
    assert got["cluster"] in (2, 4) and not got["tensor_cores"]
    check_close(got, ref)
    monkeypatch.setenv("SAT_CLUSTER", "0")
    one = run_cuda(hp, W, mems, lengths, **kw)
    assert one["cluster"] == 1
    for k in ("mel", "stop"):
        assert rel_err(got[k], one[k]) <= 2e-5, k
    print("cluster", got["cluster"], name, B, T, {k: rel_err(got[k], ref[k]) for k in ("mel", "stop")})


@pytest.mark.parametrize("name,B,mode", [("ljspeech_tacotron", 5, "free"), ("ljspeech_self_attention", 3, "teacher"),
                                         ("ljspeech_tacotron", 70, "free_stop")])
def test_decoder_step_by_step(name, B, mode, monkeypatch):
    """sat_decoder_step (the RNN-cell contract: one step per call, state kept in the handle) reproduces sat_decoder_run:
    free run bit for bit in the fp32 form, the teacher-forced run (whose one-launch form batches the tail over time)
    within the bar; at B = 70 the one-launch form runs the tensor-core gate GEMMs and the stepped one the fp32 form."""
    hp = small_hp(name, max_iters=40)
    W = sat_b200.init_weights(hp, seed=27, stress=True)
    S = 14
    mems, lengths, target, _ = decoder_inputs(hp, B, 31, S, seed=27)
    if mode == "teacher":
        kw = dict(target=target, feed_mode=0, stop_mode=0, max_steps=S)
        ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    elif mode == "free":
        kw = dict(feed_mode=1, stop_mode=0, max_steps=S)
        ref = oracle_decoder(hp, W, mems, lengths, max_iters=S, stop_mode="never")
    else:
        W["decoder/stop_token_projection/kernel"] *= 8.0
        probe = oracle_decoder(hp, W, mems, lengths, stop_mode="never", max_iters=40)
        W["decoder/stop_token_projection/bias"] -= np.float32(np.median(probe["stop"][:, 11:, 0].max(axis=1)))
        kw = dict(feed_mode=1, stop_mode=2, max_steps=40)
        ref = oracle_decoder(hp, W, mems, lengths, stop_mode="per_row", max_iters=40)
    one = run_cuda(hp, W, mems, lengths, **kw)
    stepped = run_cuda(hp, W, mems, lengths, step_by_step=True, **kw)
    assert stepped["n_steps"] == one["n_steps"] == ref["n_steps"] and not stepped["tensor_cores"]
    if mode == "free":
        np.testing.assert_array_equal(stepped["mel"], one["mel"])
        np.testing.assert_array_equal(stepped["alignments"][0], one["alignments"][0])
    if mode == "free_stop":
        np.testing.assert_array_equal(stepped["stop_steps"], ref["stop_steps"])
        n, r = stepped["n_steps"], hp.outputs_per_step
        for b, ss in enumerate(ref["stop_steps"]):  # the stepped form keeps every row running; compare up to its stop
            m = n if ss < 0 else ss + 1
            assert rel_err(stepped["mel"][b, :m * r], ref["mel"][b, :m * r]) <= TOL
    else:
        check_close(stepped, ref)


@pytest.mark.parametrize("B,T,want", [(6, 400, "2"), (4, 600, None), (3, 1000, None), (40, 90, None), (7, 64, "2")])
def test_cluster_split_edge_shapes(B, T, want, monkeypatch):
    """Cluster size 2 (SAT_CLUSTER=2), long memories (T_enc = 400 splits; from T_enc ~ 600 on the row's shared-memory
    plan no longer fits behind the query-layer arrays -> one CTA per row), and a batch too large for any split
    (B x mechanisms x 2 > SM count)."""
    if want:
        monkeypatch.setenv("SAT_CLUSTER", want)
    hp = small_hp("ljspeech_self_attention")
    W = sat_b200.init_weights(hp, seed=23, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, T, 5, seed=23)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=5)
    print("cluster size", got["cluster"], "for B, T =", B, T)
    if want:
        assert got["cluster"] == int(want)
    check_close(got, ref)


@pytest.mark.parametrize("B", [130, 256])
def test_extended_decoder_two_row_groups(B):
    hp = small_hp("ljspeech_tacotron", **SMALL)
    W = sat_b200.init_weights(hp, seed=3, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, 19, 6, seed=3)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=6)
    check_close(got, ref)


@pytest.mark.parametrize("name,B,T,S", [("ljspeech_tacotron", 64, 50, 12), ("ljspeech_tacotron", 200, 37, 8),
                                        ("ljspeech_self_attention", 96, 41, 10)])
def test_tensor_core_gate_gemms_full_dims(name, B, T, S):
    """B >= 64 switches the three ZoneoutLSTM gate GEMMs to tcgen05 (fp16 operands, fp32 accumulation, TMA-fed);
    same 1e-3 bar as the fp32 path, at the real layer sizes."""
    hp = small_hp(name)
    W = sat_b200.init_weights(hp, seed=11, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, T, S, seed=11)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=S)
    assert got["tensor_cores"]
    check_close(got, ref)
    errs = {k: rel_err(got[k], ref[k]) for k in ("mel", "stop")}
    print("tensor-core path rel err", name, B, errs)


@pytest.mark.parametrize("name,B,T,S,mode", [("ljspeech_tacotron", 200, 61, 14, "free"),
                                             ("ljspeech_tacotron", 70, 33, 9, "teacher"),
                                             ("ljspeech_self_attention", 96, 41, 10, "teacher")])
@pytest.mark.parametrize("chunks,stages", [(None, None), ("1", "3"), ("3", "2")])
def test_tensor_core_activation_ring_shapes(name, B, T, S, mode, chunks, stages, monkeypatch):
    """The activation stream of the tcgen05 gate GEMMs -- fp16 operand tiles laid out in HBM in their shared-memory
    byte order, moved by 1-D bulk copies of SAT_TC_G tiles with SAT_TC_STAGES copies in flight (default: the largest
    copies that fit, 2 x 64 KB) -- meets the bar for every ring shape, with a ragged last row block and K ranges that
    are not a multiple of the copy size."""
    if chunks is not None:
        monkeypatch.setenv("SAT_TC_G", chunks)
        monkeypatch.setenv("SAT_TC_STAGES", stages)
    hp = small_hp(name)
    W = sat_b200.init_weights(hp, seed=13, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, T, S, seed=13)
    if mode == "free":
        ref = oracle_decoder(hp, W, mems, lengths, max_iters=S, stop_mode="never")
        got = run_cuda(hp, W, mems, lengths, feed_mode=1, stop_mode=0, max_steps=S)
    else:
        ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
        got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=S)
    assert got["tensor_cores"]
    check_close(got, ref)


@pytest.mark.parametrize("name,B,T", [("ljspeech_tacotron", 256, 53), ("ljspeech_tacotron", 149, 200),
                                      ("ljspeech_self_attention", 128, 77)])
def test_attention_rows_paired_and_sequential(name, B, T, monkeypatch):
    """More attention items than CTAs: a CTA runs its two rows concurrently, one per half (named barriers,
    per-batch shared-memory plan); SAT_ATT_PAIR=0 runs them back to back on the whole CTA.  Both forms meet the
    oracle bar on ragged lengths, a T that is not a multiple of 16 and an odd number of second rows."""
    hp = small_hp(name)
    W = sat_b200.init_weights(hp, seed=5, stress=True)
    S = 5
    mems, lengths, target, _ = decoder_inputs(hp, B, T, S, seed=5)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    outs = []
    for pair in ("1", "0"):
        monkeypatch.setenv("SAT_ATT_PAIR", pair)
        got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=S)
        assert got["tensor_cores"]
        check_close(got, ref)
        outs.append(got)
    assert rel_err(outs[0]["alignments"][0], outs[1]["alignments"][0]) <= 1e-4


def test_long_memory_rows_fall_back_to_sequential_rows():
    """T_enc = 600: two rows' attention arrays no longer fit the phase scratch together, so a CTA with two rows runs
    them back to back (per-batch shared-memory plan, whole CTA per row); T_enc = 1000 uses the full-size plan."""
    hp = small_hp("ljspeech_tacotron")
    W = sat_b200.init_weights(hp, seed=9, stress=True)
    for B, T in ((150, 600), (66, 1000)):
        mems, lengths, target, _ = decoder_inputs(hp, B, T, 3, seed=9)
        ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
        got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=3)
        check_close(got, ref)


def test_small_batch_uses_fp32_simt():
    hp = small_hp("ljspeech_tacotron", **SMALL)
    W = sat_b200.init_weights(hp, seed=1)
    mems, lengths, target, _ = decoder_inputs(hp, 8, 11, 4, seed=1)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=4)
    assert not got["tensor_cores"]


@pytest.mark.parametrize("stop_mode", ["all_rows", "per_row"])
def test_extended_decoder_free_run_stop(stop_mode):
    hp = small_hp("ljspeech_tacotron", **SMALL)
    W = sat_b200.init_weights(hp, seed=11, stress=True)
    W["decoder/stop_token_projection/kernel"] *= 8.0
    mems, lengths, _, _ = decoder_inputs(hp, 6, 21, 1, seed=11)
    # the stop logit does not feed back: shift its bias so the rows stop at spread-out steps
    probe = oracle_decoder(hp, W, mems, lengths, stop_mode="never", max_iters=40)
    W["decoder/stop_token_projection/bias"] -= np.float32(np.quantile(probe["stop"][:, 11:], 0.9))
    ref = oracle_decoder(hp, W, mems, lengths, stop_mode=stop_mode, max_iters=40)
    got = run_cuda(hp, W, mems, lengths, feed_mode=1, stop_mode=1 if stop_mode == "all_rows" else 2, max_steps=40)
    assert got["n_steps"] == ref["n_steps"]
    np.testing.assert_array_equal(got["stop_steps"], ref["stop_steps"])
    n, r = got["n_steps"], hp.outputs_per_step
    if stop_mode == "per_row":
        # a stopped row is retired from the row-split phases: it emits up to its own stop step, zeros afterwards
        assert len(set(ref["stop_steps"].tolist())) > 1, "rows should stop at different steps"
        for b, ss in enumerate(ref["stop_steps"]):
            m = n if ss < 0 else ss + 1
            assert rel_err(got["mel"][b, :m * r], ref["mel"][b, :m * r]) <= TOL
            assert rel_err(got["stop"][b, :m], ref["stop"][b, :m]) <= TOL
            assert rel_err(got["alignments"][0][b, :m], ref["alignments"][0][b, :m]) <= TOL
            np.testing.assert_array_equal(np.argmax(got["alignments"][0][b, :m], -1), np.argmax(ref["alignments"][0][b, :m], -1))
            assert not got["mel"][b, m * r:].any() and not got["alignments"][0][b, m:].any()
        return
    check_close(got, ref)
    np.testing.assert_array_equal(np.argmax(got["alignments"][0][:, :n], -1), np.argmax(ref["alignments"][0][:, :n], -1))


def test_row_retirement_tensor_core_path():
    """Per-row stop rule at B = 200 (tcgen05 gate GEMMs, paired attention rows): rows stop at spread-out steps, leave the
    row-split phases, and every row's frames up to its own stop step equal the oracle's; SAT_RETIRE=0 (all rows keep
    running) gives the same frames there."""
    hp = small_hp("ljspeech_tacotron")
    W = sat_b200.init_weights(hp, seed=19, stress=True)
    W["decoder/stop_token_projection/kernel"] *= 8.0
    B, T, S = 200, 37, 30
    mems, lengths, _, _ = decoder_inputs(hp, B, T, 1, seed=19)
    probe = oracle_decoder(hp, W, mems, lengths, stop_mode="never", max_iters=S)
    W["decoder/stop_token_projection/bias"] -= np.float32(np.median(probe["stop"][:, 11:, 0].max(axis=1)))
    ref = oracle_decoder(hp, W, mems, lengths, stop_mode="per_row", max_iters=S)
    got = run_cuda(hp, W, mems, lengths, feed_mode=1, stop_mode=2, max_steps=S)
    assert got["tensor_cores"] and got["n_steps"] == ref["n_steps"]
    stops = ref["stop_steps"]
    assert 20 < int((stops >= 0).sum()) < B
    # a stop logit within fp tolerance of 0 cannot be pinned: compare the rows whose decision is clear in the oracle
    logit = ref["stop"][:, :, 0]
    scale = np.abs(logit).max()
    r = hp.outputs_per_step
    n_checked = 0
    for b in range(B):
        m = got["n_steps"] if stops[b] < 0 else stops[b] + 1
        if np.abs(logit[b, 11:m]).min() < 4 * TOL * scale:
            continue
        n_checked += 1
        assert got["stop_steps"][b] == stops[b]
        assert rel_err(got["mel"][b, :m * r], ref["mel"][b, :m * r]) <= TOL
        assert not got["mel"][b, m * r:].any()
    assert n_checked > B // 2


@pytest.mark.parametrize("kind,cumulative,agent", [("location_sensitive", False, False),
                                                   ("location_sensitive", True, False),
                                                   ("additive", False, False),
                                                   ("forward", True, True),
                                                   ("forward", False, True)])
def test_attention_kinds(kind, cumulative, agent):
    hp = small_hp("ljspeech_tacotron", attention=kind, cumulative_weights=cumulative,
                  use_forward_attention_transition_agent=agent, attention_kernel=7, attention_filters=3, **SMALL)
    W = sat_b200.init_weights(hp, seed=5, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 4, 18, 10, seed=5)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=10)
    check_close(got, ref)


def test_generic_location_filters():
    hp = small_hp("ljspeech_tacotron", attention="forward", attention_kernel=31, attention_filters=32, **SMALL)
    W = sat_b200.init_weights(hp, seed=6, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 3, 20, 6, seed=6)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=6)
    check_close(got, ref)


@pytest.mark.parametrize("B,T,S", [(2, 9, 6), (5, 30, 14)])
def test_dual_source_transformer_teacher_forced(B, T, S):
    hp = small_hp("ljspeech_self_attention", **SMALL)
    W = sat_b200.init_weights(hp, seed=B, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, B, T, S, seed=B)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=S, want_sample_ids=True)
    check_close(got, ref)
    np.testing.assert_array_equal(got["sample_ids"], ref["sample_ids"])


def test_dual_source_transformer_full_dims():
    hp = small_hp("ljspeech_self_attention")
    W = sat_b200.init_weights(hp, seed=9, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 2, 60, 30, seed=9)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=30)
    check_close(got, ref)


def test_dual_source_decoder_no_transformer():
    hp = small_hp("ljspeech_self_attention", decoder="DualSourceDecoder", **SMALL)
    W = sat_b200.init_weights(hp, seed=2, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 3, 15, 8, seed=2)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=0, stop_mode=0, max_steps=8)
    check_close(got, ref)


def test_code_decoder_softmax_feedback():
    """cfg 5 form: r=1, validation without teacher forcing feeds softmax probs (helpers.py:98-106);
    argmax code ids must match exactly."""
    hp = small_hp("codes_self_attention", **dict(SMALL, num_mels=37))
    W = sat_b200.init_weights(hp, seed=4, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 4, 16, 12, seed=4)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=False)
    got = run_cuda(hp, W, mems, lengths, target=target, feed_mode=2, stop_mode=0, max_steps=12, want_sample_ids=True)
    check_close(got, ref)
    np.testing.assert_array_equal(got["sample_ids"], ref["sample_ids"])


def test_multi_speaker_prenet():
    hp = small_hp("vctk_self_attention", **SMALL)
    W = sat_b200.init_weights(hp, seed=8, stress=True)
    mems, lengths, target, spk = decoder_inputs(hp, 5, 14, 7, seed=8)
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True, speaker_embed=spk)
    got = run_cuda(hp, W, mems, lengths, target=target, speaker_embed=spk, feed_mode=0, stop_mode=0, max_steps=7)
    check_close(got, ref)


def test_forced_alignment_second_pass():
    """models/models.py:118-134: pass 2 replays pass-1 alignments, feeding its own output."""
    hp = small_hp("ljspeech_self_attention", **SMALL)
    W = sat_b200.init_weights(hp, seed=10, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 3, 13, 9, seed=10)
    p1 = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    ta = [a.astype(np.float32) for a in p1["alignments"]]
    ref = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=False,
                         teacher_alignments=ta, forced_alignment=True)
    got = run_cuda(hp, W, mems, lengths, target=target, teacher_alignments=ta, forced_alignment=True,
                   feed_mode=2, stop_mode=0, max_steps=9)
    check_close(got, ref)
