"""GPU parity against the committed golden fixtures (tests/golden/*.npz, produced by the float64 oracle; see
tests/golden/make_golden.py) -- nothing here reads /root/reference or runs the oracle."""
import os

import numpy as np
import pytest

from tests.util import rel_err
from tests.golden.make_golden import CASES, case_inputs

pytestmark = pytest.mark.gpu
TOL = 1e-3
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", sorted(CASES))
def test_golden_case(name):
    import torch
    from sat_b200.models import tacotron_model_factory
    hp, W, ids, lengths, target, spk, mode = case_inputs(name)
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    t = lambda a: None if a is None else torch.from_numpy(a).to("cuda:0")  # noqa: E731
    if mode == "forced_alignment":
        hp.override(use_forced_alignment_mode=True)
    out = tacotron_model_factory(hp, W).predict(
        t(ids), t(lengths), target=None if mode == "free_run" else t(target), speaker_id=t(spk),
        is_validation=mode in ("teacher_forced", "validation"), teacher_forcing=mode == "teacher_forced")
    assert out["n_steps"] == int(z["n_steps"])
    assert rel_err(out["mel"].cpu().numpy(), z["mel"]) <= TOL
    assert rel_err(out["stop_token"].cpu().numpy(), z["stop"]) <= TOL
    assert rel_err(out["alignment"].cpu().numpy(), z["alignment1"].transpose(0, 2, 1)) <= TOL
    if "alignment2" in z.files:
        assert rel_err(out["alignment2"].cpu().numpy(), z["alignment2"].transpose(0, 2, 1)) <= TOL
    if name == "codes_validation":  # cfg 5 form: argmax code ids must match exactly
        np.testing.assert_array_equal(out["predicted_samples"].cpu().numpy(), z["sample_ids"])
