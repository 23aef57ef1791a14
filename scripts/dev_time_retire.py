"""Developer aid: kernel time of the decoder under the all_rows rule (decoder_persistent_kernel<false>) and the per_row rule
(<true>, row retirement) in ONE process, interleaved, with no row ever stopping (STOP_BIAS=20, the default) or with the
random weights' own stops (STOP_BIAS=0): the cost / gain of the retirement instantiation.  (The phase counters run in an
instantiation of their own without retirement, so this script reads the CUDA-event time only.)"""
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
        o = h.run([mem], lengths, feed_mode=1, stop_mode=sm, max_steps=S)
        torch.cuda.synchronize()
        print(f"stop_mode {sm}: {h.kernel_ms():.2f} ms, rows stopped {int((o['stop_steps'] >= 0).sum().item())}")
