"""CPU oracle: a NumPy restatement of the mel-synthesis hot path of
rhoposit/self-attention-tacotron.

THIS IS TEST INFRASTRUCTURE.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it.  The
product package must never route through it.

PARITY UNPINNED: the reference is TensorFlow 1.x + the un-vendored
``nii-yamagishilab/tacotron2`` @ 6af04c7f (WORKSPACE:9-13); neither can run in
this image, and the reference tree holds no golden vectors or checkpoints.  The
only executable pin the reference offers is the *property* asserted by
``modules/transformer_test.py:44-90`` (batched teacher-forced pass == step-wise
inference pass); ``tests/test_oracle_property.py`` runs that property inside
this oracle.  Fidelity of every other function is by construction from the
cited lines (and SURVEY.md Appendix A for the absent third-party code).

All citations ``file:line`` are into /root/reference.
"""
from . import nn, attention, encoder, decoder, postnet  # noqa: F401
