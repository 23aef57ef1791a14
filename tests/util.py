"""Shared helpers of the test-suite: seeded inputs, oracle calls, error metric."""
import numpy as np

import sat_b200
from oracle.decoder import decoder_run


def rel_err(a, ref):
    """max|a - ref| / max|ref|  -- the parity metric of BASELINE.json (tolerance 1e-3)."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-30))


def small_hp(base="ljspeech_self_attention", **kw):
    hp = sat_b200.config(base)
    hp.override(**kw)
    return hp


def decoder_inputs(hp, B, T, S, seed=0, ragged=True):
    """Synthetic encoder memories / lengths / teacher frames / speaker embeddings."""
    rng = np.random.default_rng(seed)
    dims = [hp.cbhg_out_units] if hp.decoder == "ExtendedDecoder" else [hp.cbhg_out_units, hp.self_attention_out_units]
    mems = [rng.normal(0, 0.5, size=(B, T, d)).astype(np.float32) for d in dims]
    lengths = np.full(B, T, dtype=np.int64)
    if ragged and B > 1:
        lengths[1:] = rng.integers(max(1, T // 2), T + 1, size=B - 1)
    target = rng.normal(0, 1, size=(B, S * hp.outputs_per_step, hp.num_mels)).astype(np.float32)
    spk = rng.normal(0, 1, size=(B, hp.speaker_embedding_dim)).astype(np.float32)
    return mems, lengths, target, spk


def oracle_decoder(hp, W, mems, lengths, **kw):
    W64 = {k: v.astype(np.float64) for k, v in W.items()}
    src = [m.astype(np.float64) for m in mems]
    sources = src[0] if hp.decoder == "ExtendedDecoder" else tuple(src)
    for k in ("target", "speaker_embed"):
        if kw.get(k) is not None:
            kw[k] = kw[k].astype(np.float64)
    if kw.get("teacher_alignments") is not None:
        kw["teacher_alignments"] = tuple(a.astype(np.float64) for a in kw["teacher_alignments"])
    return decoder_run(sources, lengths, W64, hp, hp.decoder, **kw)
