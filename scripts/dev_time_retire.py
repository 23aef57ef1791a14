"""Developer aid: per-phase cycles of the decoder kernel under the all_rows rule (decoder_persistent_kernel<false>) and
the per_row rule (<true>, row retirement) in ONE process, interleaved, with no row ever stopping (stop bias -20): the
pure cost of the retirement instantiation."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200
from sat_b200.engine import DecoderHandle
hp = sat_b200.config("ljspeech_tacotron")
W = sat_b200.init_weights(hp, seed=12345)
W["decoder/stop_token_projection/bias"] = W["decoder/stop_token_projection/bias"] - np.float32(float(os.environ.get("STOP_BIAS", "20")))
rng = np.random.default_rng(0)
dev = torch.device("cuda:0")
B, T, S = 256, 200, 500
mem = torch.from_numpy(rng.normal(0, .5, size=(B, T, 256)).astype(np.float32)).to(dev)
lengths = torch.from_numpy(rng.integers(100, 201, size=B).astype(np.int64)).to(dev)
h = DecoderHandle(hp, W)
for sm in (1, 2):
    h.run([mem], lengths, feed_mode=1, stop_mode=sm, max_steps=S)
torch.cuda.synchronize()
for rep in range(3):
    for sm in (1, 2):
        h.profile(reset=True)
        o = h.run([mem], lengths, feed_mode=1, stop_mode=sm, max_steps=S)
        torch.cuda.synchronize()
        prof = h.profile()
        ms = h.kernel_ms()
        print(f"stop_mode {sm}: {ms:.2f} ms, rows stopped {int((o['stop_steps'] >= 0).sum().item())}, kcycles/step per phase",
              " ".join(f"{p / S / 1e3:.2f}" for p in prof[:16] if p), "sum %.1f" % (sum(prof[:16]) / S / 1e3),
              "| attention stages", " ".join(f"{p / S / 1e3:.2f}" for p in prof[16:24]),
              "| frame stages", " ".join(f"{p / S / 1e3:.2f}" for p in prof[32:38]))
