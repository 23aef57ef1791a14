"""Generator of the golden fixtures in this directory:  python -m tests.golden.make_golden

The reference holds no golden vectors, known-answer tests or checkpoints for this path and
cannot be executed here (TensorFlow 1.x + un-vendored tacotron2), so these vectors come from the float64 NumPy
oracle (oracle/), seeded; they pin the ORACLE (and through the GPU tests, the kernels) against drift -- they are
not outputs of the reference.  Stored as float32 to stay small."""
import os

import numpy as np

import sat_b200
from oracle.postnet import synthesize

HERE = os.path.dirname(os.path.abspath(__file__))
SMALL = dict(attention_out_units=32, decoder_out_units=32, decoder_self_attention_out_units=32,
             attention1_out_units=24, attention2_out_units=8, decoder_prenet_out_units=(24, 16), num_mels=12,
             max_filter_width=4, num_highway=2)
CASES = {
    "tacotron_teacher_forced": ("ljspeech_tacotron", "teacher_forced", {}),
    "tacotron_free_run": ("ljspeech_tacotron", "free_run", {}),
    "self_attention_teacher_forced": ("ljspeech_self_attention", "teacher_forced", {}),
    "self_attention_forced_alignment": ("ljspeech_self_attention", "forced_alignment", {}),
    "vctk_teacher_forced": ("vctk_self_attention", "teacher_forced", {}),
    "codes_validation": ("codes_self_attention", "validation", dict(num_mels=37)),
}


def case_inputs(name):
    cfg, mode, extra = CASES[name]
    hp = sat_b200.config(cfg)
    hp.override(**dict(SMALL, **extra))
    hp.override(max_iters=16)
    seed = sorted(CASES).index(name) + 12345
    W = sat_b200.init_weights(hp, seed=seed, stress=True)
    rng = np.random.default_rng(seed)
    B, T, S = 3, 19, 8
    lengths = np.array([19, 14, 8], dtype=np.int64)
    ids = rng.integers(1, 71, size=(B, T)).astype(np.int64)
    ids[np.arange(T)[None, :] >= lengths[:, None]] = 0
    target = rng.normal(0, 1, size=(B, S * hp.outputs_per_step, hp.num_mels)).astype(np.float32)
    spk = np.array([225, 300, 376], dtype=np.int64) if hp.use_speaker_embedding else None
    return hp, W, ids, lengths, target, spk, mode


def run_case(name):
    hp, W, ids, lengths, target, spk, mode = case_inputs(name)
    W64 = {k: v.astype(np.float64) for k, v in W.items()}
    res = synthesize(ids, lengths, W64, hp, target=None if mode == "free_run" else target.astype(np.float64),
                     speaker_ids=spk, mode=mode)
    out = {"mel": res["mel"], "stop": res["stop"], "n_steps": np.array(res["n_steps"]),
           "sample_ids": res["sample_ids"], "memory1": res["sources"][0] if isinstance(res["sources"], tuple) else res["sources"]}
    for i, a in enumerate(res["alignments"]):
        out[f"alignment{i + 1}"] = a
    return out


def main():
    for name in CASES:
        out = run_case(name)
        np.savez_compressed(os.path.join(HERE, name + ".npz"),
                            **{k: (v.astype(np.float32) if v.dtype == np.float64 else v) for k, v in out.items()})
        print(name, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
