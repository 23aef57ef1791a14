"""compute-sanitizer driver: small runs of every kernel through the public handles, one
group per process, so that `compute-sanitizer --tool memcheck|racecheck|synccheck python scripts/sanitize_cases.py
<group>` covers the hand-rolled grid barrier, the named-barrier half-CTA attention rows, the cluster-split rows, the
TMA / mbarrier rings and the tcgen05 paths at sizes the tools finish in minutes.  Outputs are only checked for
finiteness here (parity is what tests/ is for)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import sat_b200
from sat_b200.engine import DecoderHandle, EncoderHandle


def dec(cfg, B, T, S, feed, stop=0, env=None, **kw):
    for k, v in (env or {}).items():
        os.environ[k] = v
    hp = sat_b200.config(cfg)
    W = sat_b200.init_weights(hp, seed=1)
    rng = np.random.default_rng(0)
    dims = [hp.cbhg_out_units] if hp.decoder == "ExtendedDecoder" else [hp.cbhg_out_units, hp.self_attention_out_units]
    dev = torch.device("cuda:0")
    mems = [torch.from_numpy(rng.normal(0, .5, size=(B, T, d)).astype(np.float32)).to(dev) for d in dims]
    lengths = torch.from_numpy(rng.integers(max(1, T // 2), T + 1, size=B).astype(np.int64)).to(dev)
    target = torch.from_numpy(rng.normal(0, 1, size=(B, S * hp.outputs_per_step, hp.num_mels)).astype(np.float32)).to(dev)
    h = DecoderHandle(hp, W)
    out = h.run(mems, lengths, target=target if feed == 0 else None, feed_mode=feed, stop_mode=stop, max_steps=S, **kw)
    torch.cuda.synchronize()
    n = int(out["n_steps"].item())
    assert bool(torch.isfinite(out["mel"][:, :n * hp.outputs_per_step]).all())
    print(f"  ok decoder {cfg} B={B} T={T} S={S} feed={feed} stop={stop} tc={h.used_tensor_cores()} "
          f"cluster={h.cluster_size()} env={env}", flush=True)
    for k in (env or {}):
        os.environ.pop(k)


def enc(cfg, B, T):
    hp = sat_b200.config(cfg)
    W = sat_b200.init_weights(hp, seed=2)
    rng = np.random.default_rng(1)
    dev = torch.device("cuda:0")
    ids = torch.from_numpy(rng.integers(1, 70, size=(B, T)).astype(np.int64)).to(dev)
    lengths = torch.from_numpy(rng.integers(T // 2, T + 1, size=B).astype(np.int64)).to(dev)
    for mask in (False, True):
        h = EncoderHandle(hp, W)
        h.set_mask_padding(mask)
        o = h.forward(ids, lengths, want_probs=True)
        torch.cuda.synchronize()
        assert bool(torch.isfinite(o[0]).all())
    print(f"  ok encoder {cfg} B={B} T={T} tc={h.uses_tensor_cores()}", flush=True)


GROUPS = {
    # fp32 SIMT form: cluster-split rows (default at small B), one CTA per row (SAT_CLUSTER=0), free run with both stop
    # rules, forced-alignment replay, the time-batched teacher-forced tail (mha_flash_tc_kernel, gemm_kernel)
    "decoder_simt": lambda: (dec("ljspeech_self_attention", 3, 23, 4, 0),
                             dec("ljspeech_self_attention", 3, 23, 3, 0, env={"SAT_CLUSTER": "0"}),
                             dec("ljspeech_tacotron", 4, 30, 13, 1, stop=2),
                             dec("ljspeech_tacotron", 37, 19, 3, 1, stop=1),
                             dec("codes_self_attention", 2, 17, 3, 2)),
    # tensor-core form: tcgen05 gate GEMMs with the TMA ring, paired attention rows (B > 148), two activation ring shapes
    "decoder_tc": lambda: (dec("ljspeech_tacotron", 64, 20, 3, 1),
                           dec("ljspeech_tacotron", 150, 24, 2, 1),
                           dec("ljspeech_self_attention", 66, 18, 2, 1),
                           dec("ljspeech_tacotron", 130, 20, 2, 1, env={"SAT_TC_G": "1", "SAT_TC_STAGES": "3"})),
    # encoder: conv_tc_kernel (fused bank, proj, highway, LSTM projection), bilstm_kernel, mha_tc_kernel
    "encoder": lambda: (enc("ljspeech_self_attention", 2, 40), enc("ljspeech_tacotron", 3, 33)),
}

if __name__ == "__main__":
    for g in sys.argv[1:]:
        print("group", g, flush=True)
        GROUPS[g]()
