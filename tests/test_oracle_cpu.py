"""CPU tests of the oracle itself (no GPU): the one executable pin the reference offers -- the property of
modules/transformer_test.py:44-90 (batched teacher-forced pass == step-wise inference pass of RNNTransformer) --
run inside the oracle, plus internal consistency of its three decoder self-attention forms, the golden
fixtures, and the torch-CPU port used as the timed baseline."""
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import sat_b200
from oracle import decoder as odec
from oracle.postnet import synthesize
from tests.util import small_hp, decoder_inputs, oracle_decoder, rel_err

SMALL = dict(attention_out_units=32, decoder_out_units=32, decoder_self_attention_out_units=32,
             attention1_out_units=24, attention2_out_units=8, cbhg_out_units=16, self_attention_out_units=8,
             decoder_prenet_out_units=(24, 16), num_mels=12)
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@settings(max_examples=12, deadline=None)
@given(B=st.integers(1, 3), t_factor=st.integers(2, 8), r=st.sampled_from([1, 2]), half_dim=st.integers(1, 10),
       seed=st.integers(0, 1000))
def test_equality_between_training_and_inference(B, t_factor, r, half_dim, seed):
    """transformer_test.py:18-32 generators (batch 1-3, even target_dim 2-20, t = t_factor * r, r in {1,2}, one-hot
    targets); :88-90 asserts outputs, stop tokens and argmax ids equal between the two branches."""
    M = 2 * half_dim
    hp = small_hp("ljspeech_self_attention", **dict(SMALL, num_mels=M, outputs_per_step=r, n_feed_frame=r, max_iters=100))
    W = {k: v.astype(np.float64) for k, v in sat_b200.init_weights(hp, seed=12345).items()}  # tf.set_random_seed(12345)
    rng = np.random.default_rng(seed)
    S = t_factor
    mems, lengths, _, _ = decoder_inputs(hp, B, 6, S, seed=seed, ragged=False)
    target = np.eye(M)[rng.integers(0, M, size=(B, S * r))]
    kw = dict(target=target, is_validation=True, teacher_forcing=True)
    src = tuple(m.astype(np.float64) for m in mems)
    train = odec.decoder_run(src, lengths, W, hp, sa_form="batched", **kw)
    infer = odec.decoder_run(src, lengths, W, hp, sa_form="history", **kw)
    cached = odec.decoder_run(src, lengths, W, hp, sa_form="cached", **kw)
    for other in (infer, cached):
        np.testing.assert_allclose(train["mel"], other["mel"], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(train["stop"], other["stop"], rtol=1e-6, atol=1e-6)
        np.testing.assert_array_equal(train["sample_ids"], other["sample_ids"])


def test_forward_attention_alignments_are_distributions():
    hp = small_hp("ljspeech_tacotron", **SMALL)
    W = sat_b200.init_weights(hp, seed=1, stress=True)
    mems, lengths, target, _ = decoder_inputs(hp, 4, 17, 9, seed=1)
    res = oracle_decoder(hp, W, mems, lengths, target=target, is_validation=True, teacher_forcing=True)
    a = res["alignments"][0]
    np.testing.assert_allclose(a.sum(-1), 1.0, atol=1e-12)
    for b in range(4):
        assert np.all(a[b, :, lengths[b]:] == 0)


def test_stop_rule_modes():
    """helpers.py:156-160: whole batch stops when ALL rows say stop at the same step, only after time > 10."""
    hp = small_hp("ljspeech_tacotron", **SMALL)
    W = sat_b200.init_weights(hp, seed=11, stress=True)
    W["decoder/stop_token_projection/bias"] += 50.0  # every row says "stop" at every step
    mems, lengths, _, _ = decoder_inputs(hp, 3, 9, 1, seed=11)
    res = oracle_decoder(hp, W, mems, lengths, stop_mode="all_rows", max_iters=40)
    assert res["n_steps"] == 12 and np.all(res["stop_steps"] == 11)  # first t with t > 10 is t = 11


def test_golden_fixtures_match_oracle():
    """tests/golden/*.npz were produced by tests/golden/make_golden.py from this oracle (seeded); they freeze the
    oracle's numbers so that a later change of oracle or kernels cannot drift silently."""
    from tests.golden.make_golden import CASES, run_case
    for name in CASES:
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        got = run_case(name)
        for k in z.files:
            np.testing.assert_allclose(got[k], z[k], rtol=1e-5, atol=1e-6, err_msg=f"{name}/{k}")


@pytest.mark.parametrize("cfg", ["ljspeech_tacotron", "ljspeech_self_attention"])
def test_torch_cpu_port_matches_numpy_oracle(cfg):
    from oracle import torch_cpu
    hp = sat_b200.config(cfg)
    W = sat_b200.init_weights(hp, seed=3, stress=True)
    rng = np.random.default_rng(0)
    lengths = np.array([21, 15, 9])
    ids = rng.integers(1, 71, size=(3, 21))
    ids[np.arange(21)[None, :] >= lengths[:, None]] = 0
    ref = synthesize(ids, lengths, {k: v.astype(np.float64) for k, v in W.items()}, hp, mode="free_run",
                     stop_mode="never", max_iters=8)
    got = torch_cpu.synthesize_free_run(ids, lengths, W, hp, 8).numpy()
    assert rel_err(got, ref["mel"]) <= 1e-4
