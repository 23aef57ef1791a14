import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch, sat_b200
from sat_b200.engine import DecoderHandle
hp = sat_b200.config("ljspeech_tacotron"); W = sat_b200.init_weights(hp, seed=12345)
rng = np.random.default_rng(0); dev = torch.device("cuda:0"); B, T, S = 256, 200, 500
mem = torch.from_numpy(rng.normal(0, .5, size=(B, T, 256)).astype(np.float32)).to(dev)
lengths = torch.from_numpy(rng.integers(100, 201, size=B).astype(np.int64)).to(dev)
h = DecoderHandle(hp, W)
for sm in (1, 2, 1, 2):
    h.run([mem], lengths, feed_mode=1, stop_mode=sm, max_steps=S); torch.cuda.synchronize()
    ms = []
    for _ in range(3):
        o = h.run([mem], lengths, feed_mode=1, stop_mode=sm, max_steps=S); ms.append(h.kernel_ms())
    print("stop_mode", sm, "kernel ms", np.round(ms, 3), "n_steps", int(o["n_steps"].item()), "rows stopped", int((o["stop_steps"] >= 0).sum().item()))
