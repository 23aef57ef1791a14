"""sat_mha_forward_noalloc: the allocation-free variant of sat_mha_forward (no counterpart in the reference).
Called through the C ABI and checked against the existing entry point on two
calls in a row that share the fixed device scratch array; sizes beyond the maximum must be rejected.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import sat_b200
from tests.util import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _mha_weights(rng, Din, D):
    W = {}
    for n, i in (("key", Din), ("value", Din), ("query", Din), ("output", D)):
        W[f"mha/{n}/kernel"] = rng.uniform(-0.3, 0.3, size=(i, D)).astype(np.float32)
        W[f"mha/{n}/bias"] = rng.normal(0, 0.1, size=(D,)).astype(np.float32)
    return W


def _call_noalloc(lib, dW, x, H, D, causal, lengths=None):
    import torch
    from sat_b200._lib import ptr, stream_ptr
    B, T, Din = x.shape
    out = torch.full((B, T, D), float("nan"), device=x.device, dtype=torch.float32)
    probs = torch.full((B, H, T, T), float("nan"), device=x.device, dtype=torch.float32)
    rc = lib.sat_mha_forward_noalloc(ptr(dW["wqkv"]), ptr(dW["bqkv"]), ptr(dW["wo"]), ptr(dW["bo"]), ptr(x),
                                     B, T, Din, D, H, int(causal), ptr(lengths), ptr(out), ptr(probs), stream_ptr())
    return rc, out, probs


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    import torch
    from sat_b200 import engine
    from sat_b200._lib import lib
    L = lib()
    L.sat_mha_noalloc_max_floats.restype = C.c_int64
    max_floats = int(L.sat_mha_noalloc_max_floats())
    assert max_floats == 2 * 1024 * 1024
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(77)
    # two different problems back to back: the second is smaller, so it runs over the first call's leftovers
    for (B, T, Din, D, H, causal) in ((5, 150, 32, 32, 2, False), (3, 67, 256, 256, 2, True)):
        assert B * T * 4 * D <= max_floats
        W = _mha_weights(rng, Din, D)
        x = torch.from_numpy(rng.normal(0, 1, size=(B, T, Din)).astype(np.float32)).to(dev)
        ref_out, ref_probs = engine.mha_forward(W, "mha", x, H, D, causal=causal)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
        dW = {"wqkv": t(np.concatenate([W["mha/key/kernel"], W["mha/value/kernel"], W["mha/query/kernel"]], axis=1)),
              "bqkv": t(np.concatenate([W["mha/key/bias"], W["mha/value/bias"], W["mha/query/bias"]])),
              "wo": t(W["mha/output/kernel"]), "bo": t(W["mha/output/bias"])}
        for _ in range(2):
            rc, out, probs = _call_noalloc(L, dW, x, H, D, causal)
            assert rc == 0, L.sat_last_error()
            torch.cuda.synchronize()
            assert rel_err(out.cpu().numpy(), ref_out.cpu().numpy()) <= TOL
            assert rel_err(probs.cpu().numpy(), ref_probs.cpu().numpy()) <= TOL


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from sat_b200._lib import lib
    L = lib()
    dev = torch.device("cuda:0")
    B, T, Din, D, H = 64, 1024, 16, 16, 2   # 64 * 1024 * 4 * 16 = 4 Mi floats > 2 Mi
    z = torch.zeros(4, device=dev)
    x = torch.zeros((1, 1, Din), device=dev)  # never read: the size check comes first
    rc = L.sat_mha_forward_noalloc(C.c_void_p(z.data_ptr()), C.c_void_p(z.data_ptr()), C.c_void_p(z.data_ptr()),
                                   C.c_void_p(z.data_ptr()), C.c_void_p(x.data_ptr()), B, T, Din, D, H, 0,
                                   C.c_void_p(0), C.c_void_p(z.data_ptr()), C.c_void_p(0), C.c_void_p(0))
    assert rc == -1  # SAT_EINVAL
    msg = L.sat_last_error().decode()
    assert "exceed the fixed maximum" in msg and "SAT_MHA_NOALLOC_MAX_FLOATS" in msg
