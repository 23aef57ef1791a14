"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/sat.h declares, host-side
config parsing, weight-blob shapes, length bucketing / sharding incl. a world_size-2 gloo gather, and the loud
failure of the product path without CUDA."""
import ctypes
import os
import re
import subprocess
import sys
import textwrap

import numpy as np
import pytest

import sat_b200
from sat_b200.sharding import make_buckets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    header = open(os.path.join(ROOT, "include", "sat.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    names = set(re.findall(r"\b(sat_[a-z_0-9]+)\s*\(", header))
    assert len(names) >= 20
    lib = ctypes.CDLL(os.path.join(ROOT, "self-attention-tacotron_b200", "libsat.so"))
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, f"declared in sat.h but not exported: {missing}"
    lib.sat_abi_version.restype = ctypes.c_int
    assert lib.sat_abi_version() == 2


def test_hparams_struct_matches_header_field_order():
    """ctypes mirror of sat_hparams must list the same fields in the same order as include/sat.h."""
    from sat_b200._lib import SatHParams, SatDecoderIO
    header = open(os.path.join(ROOT, "include", "sat.h")).read()
    body = re.search(r"typedef struct \{([^{}]*)\} sat_hparams;", header, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = [re.sub(r"\[\d+\]", "", f).strip() for decl in body.split(";") if decl.strip()
              for f in decl.strip().split(None, 1)[1].split(",")]
    assert fields == [f[0] for f in SatHParams._fields_]
    body = re.search(r"typedef struct \{([^{}]*)\} sat_decoder_io;", header, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if decl:
            decl = re.sub(r"^(const\s+)?\w+\s*\**", "", decl, count=1)
            fields += [re.sub(r"\[\d+\]|\*", "", f).strip() for f in decl.split(",")]
    assert fields == [f[0] for f in SatDecoderIO._fields_]


def test_hparams_json_and_overrides():
    hp = sat_b200.HParams()
    hp.parse_json('{"num_mels": 80, "outputs_per_step": 2, "attention": "forward", "attention_kernel": 10, '
                  '"attention_filters": 5, "decoder_version": "v2", "decoder_prenet_out_units": [256, 128], '
                  '"some_training_only_key": 3}')
    hp.parse("max_iters=123,cumulative_weights=true,decoder_prenet_out_units=(64,32)")
    assert (hp.num_mels, hp.outputs_per_step, hp.max_iters, hp.cumulative_weights) == (80, 2, 123, True)
    assert hp.decoder_prenet_out_units == (64, 32)
    with pytest.raises(ValueError):
        hp.parse("no_such_key=1")
    with pytest.raises(ValueError, match="Unknown config"):
        sat_b200.config("nope")


def test_hparams_carry_every_reference_key_with_its_default():
    """hparams.py:11-225: the full key set and defaults (golden listing made by tests/golden/make_hparams_golden.py
    from the reference file); a reference command line that touches training-only keys must parse here too."""
    import json
    ref = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_hparams_defaults.json")))
    ours = sat_b200.HParams().values()
    assert set(ours) == set(ref) and len(ref) == 150
    norm = lambda v: list(v) if isinstance(v, (list, tuple)) else v  # noqa: E731
    assert {k for k in ref if norm(ours[k]) != ref[k]} == set()
    hp = sat_b200.HParams()
    hp.parse("batch_size=1,learning_rate_step_factor=4,vars_to_warm_start=['a','b'],use_cache=true,logfile=x.txt")
    assert (hp.batch_size, hp.learning_rate_step_factor, hp.vars_to_warm_start, hp.use_cache, hp.logfile) == \
        (1, 4, ("a", "b"), True, "x.txt")


def test_decoder_parameter_counts_match_survey():
    """decoder parameters cfg1/3/4 = 2 584 216, cfg2 = 2 197 080, cfg5 = 3 048 441."""
    assert sat_b200.decoder_param_count(sat_b200.config("ljspeech_self_attention")) == 2584216
    assert sat_b200.decoder_param_count(sat_b200.config("ljspeech_tacotron")) == 2197080
    assert sat_b200.decoder_param_count(sat_b200.config("codes_self_attention")) == 3048441


def test_unknown_names_raise_like_the_reference():
    from sat_b200._lib import make_hparams
    hp = sat_b200.config("ljspeech_tacotron")
    hp.encoder = "Bogus"
    with pytest.raises(ValueError, match="Unknown encoder"):
        make_hparams(hp)
    hp = sat_b200.config("ljspeech_tacotron")
    hp.attention = "bogus"
    with pytest.raises(ValueError, match="Unknown attention mechanism"):
        make_hparams(hp)
    from sat_b200.models import tacotron_model_factory
    hp = sat_b200.config("ljspeech_tacotron")
    hp.tacotron_model = "Other"
    with pytest.raises(ValueError, match="Unknown Tacotron model"):
        tacotron_model_factory(hp, {})


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from sat_b200 import engine
    with pytest.raises(RuntimeError, match="CUDA"):
        engine.dense_forward(torch.zeros(2, 3), np.zeros((3, 4), np.float32))
    from sat_b200.predict_mel import predict, Utterance
    with pytest.raises(RuntimeError):
        predict(sat_b200.config("ljspeech_tacotron"), {}, [Utterance(0, "k", [1, 2])], device=torch.device("cpu"))


@pytest.mark.parametrize("n,bs,world", [(0, 8, 2), (1, 8, 4), (37, 8, 1), (1024, 128, 8), (513, 64, 4)])
def test_buckets_partition_every_utterance_once(n, bs, world):
    rng = np.random.default_rng(n)
    lengths = rng.integers(50, 251, size=n)
    per_rank = make_buckets(lengths, bs, world)
    assert len(per_rank) == world
    seen = sorted(i for r in per_rank for b in r for i in b)
    assert seen == list(range(n))
    for r in per_rank:
        for b in r:
            assert 1 <= len(b) <= bs
    if n >= world * 4:
        loads = [sum(int(lengths[b].max()) * len(b) for b in r) for r in per_rank]
        assert max(loads) <= 1.25 * min(loads)


def test_cost_model_buckets_never_cost_more_than_fixed_width():
    """The DP cut minimises sum(longest * (c0 + c1 * size)) over contiguous cuts of the sorted share, so it is at
    least as cheap as the fixed-width rule on the same share; with a flat step cost it merges down to the minimum
    number of batches, with a purely per-utterance cost it isolates nothing longer than needed."""
    from sat_b200.sharding import STEP_US0, STEP_US1
    rng = np.random.default_rng(7)
    lengths = rng.integers(50, 251, size=700)

    def cost(per_rank):
        return max(sum(int(lengths[b].max()) * (STEP_US0 + STEP_US1 * len(b)) for b in r) for r in per_rank)
    for world in (1, 3, 8):
        dp, fixed = make_buckets(lengths, 128, world), make_buckets(lengths, 128, world, cost_model=None)
        assert cost(dp) <= cost(fixed) + 1e-6
        assert sorted(i for r in dp for b in r for i in b) == list(range(700))
    flat = make_buckets(lengths, 128, 1, cost_model=(1.0, 0.0))[0]
    assert len(flat) == -(-700 // 128)
    lens4 = np.array([10, 10, 10, 200])
    per_utt = make_buckets(lens4, 4, 1, cost_model=(0.0, 1.0))[0]  # the long utterance must not drag the short ones
    assert sum(int(lens4[b].max()) * len(b) for b in per_utt) == 230 and [3] in per_utt
    # the native dynamic programme (libsat.so: sat_bucket_cuts) and its NumPy statement cut alike
    from sat_b200.sharding import _bucket_cuts, _bucket_cuts_numpy
    for n, bs in ((1, 4), (5, 8), (700, 128), (2048, 256)):
        L = np.sort(rng.integers(20, 300, size=n))[::-1].copy()
        np.testing.assert_array_equal(_bucket_cuts(L, bs, STEP_US0, STEP_US1), _bucket_cuts_numpy(L, bs, STEP_US0, STEP_US1))
    assert make_buckets([], 8, 2) == [[], []]


def test_two_rank_gloo_shard_and_gather(tmp_path):
    """world_size-2 gloo run of the host-side sharding + output gather (the only collective of the path)."""
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {ROOT!r})
        import numpy as np, torch.distributed as dist
        import sat_b200
        import torch
        from sat_b200.sharding import make_buckets, gather_results, gather_mels
        dist.init_process_group("gloo")
        rank, world = dist.get_rank(), dist.get_world_size()
        lengths = np.random.default_rng(0).integers(20, 90, size=41)
        mine = make_buckets(lengths, 8, world)[rank]
        local = [(int(i), int(lengths[i]) * 2) for b in mine for i in b]   # stand-in for per-utterance results
        allr = gather_results(local, world, rank)
        if rank == 0:
            assert sorted(i for i, _ in allr) == list(range(41)), allr
            assert all(v == int(lengths[i]) * 2 for i, v in allr)
            print("GATHER_OK", len(allr))
        # tensor gather of variable-length frames (all_reduce MAX + all_gather of the padded block)
        ids = [i for b in mine for i in b]
        mels = [torch.full((int(lengths[i]), 5), float(i)) for i in ids]
        got = gather_mels(ids, mels, world, rank)
        if rank == 0:
            assert sorted(got) == list(range(41))
            assert all(tuple(got[i].shape) == (int(lengths[i]), 5) and bool((got[i] == float(i)).all()) for i in got)
            print("MELS_OK", len(got))
        else:
            assert got is None
        dist.destroy_process_group()
    """))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29611", str(script)],
                         capture_output=True, text=True, timeout=240)
    assert "GATHER_OK 41" in out.stdout, out.stdout + out.stderr
    assert "MELS_OK 41" in out.stdout, out.stdout + out.stderr
