"""Secondary measurement (not the driver's bench contract): BASELINE.json configs[0] / SURVEY.md 8d cfg 1 -- the
dual-source self-attention Tacotron, teacher-forced, B=16, T_enc=150, 800 target frames (S=400, r=2) -- through
SynthesisModel.predict with device-resident inputs.  north_star target: >= 1e6 teacher-forced mel frames/s.
Prints one JSON line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200
from sat_b200.models import tacotron_model_factory

def main(B=16, T=150, S=400, steps=10, warmup=3):
    hp = sat_b200.config("ljspeech_self_attention")
    W = sat_b200.init_weights(hp, seed=12345)
    rng = np.random.default_rng(12345)
    dev = torch.device("cuda:0")
    ids = torch.from_numpy(rng.integers(1, 71, size=(B, T)).astype(np.int64)).to(dev)
    lengths = torch.full((B,), T, dtype=torch.int64, device=dev)
    target = torch.from_numpy(rng.normal(0, 1, size=(B, S * hp.outputs_per_step, hp.num_mels)).astype(np.float32)).to(dev)
    model = tacotron_model_factory(hp, W)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    def step():
        flush.fill_(1)
        return model.predict(ids, lengths, target=target, is_validation=True, teacher_forcing=True)
    for _ in range(warmup):
        out = step()
    torch.cuda.synchronize()
    assert int(out["n_steps"]) == S and bool(torch.isfinite(out["mel"]).all())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    frames = B * S * hp.outputs_per_step
    print(json.dumps({"metric": "mel_frames_per_sec_teacher_forced", "value": frames / (ms * 1e-3), "unit": "frames/s",
                      "ms_per_step": ms, "config": {"workload": f"cfg1: dual-source self-attention Tacotron, teacher-forced, B={B}, "
                      f"T_enc={T}, {S * hp.outputs_per_step} frames/utterance; encoder + decoder (+ L2 flush) per step"},
                      "steps": steps, "warmup": warmup, "data": "synthetic"}))

if __name__ == "__main__":
    main()
