"""Developer aid: time the encoder (CUDA events) with the conv bank / proj1 on the tensor cores vs the fp32 SIMT
form (SAT_ENC_TC=0) and report the difference of their outputs.  Not part of the bench contract."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200
from sat_b200.engine import EncoderHandle

def run(cfg, B, T, reps=5):
    hp = sat_b200.config(cfg)
    W = sat_b200.init_weights(hp, seed=12345)
    rng = np.random.default_rng(0)
    dev = torch.device("cuda:0")
    ids = torch.from_numpy(rng.integers(1, 70, size=(B, T)).astype(np.int64)).to(dev)
    lengths = torch.from_numpy(rng.integers(T * 2 // 3, T + 1, size=B).astype(np.int64)).to(dev)
    outs = {}
    for mode in ("1", "0"):
        os.environ["SAT_ENC_TC"] = mode
        h = EncoderHandle(hp, W)
        o = h.forward(ids, lengths); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            o = h.forward(ids, lengths)
        e1.record(); torch.cuda.synchronize()
        outs[mode] = o[0].float().cpu().numpy()
        print(f"{cfg} B={B} T={T} tensor_cores={mode}: {e0.elapsed_time(e1)/reps:.3f} ms/forward")
    a, b = outs["1"], outs["0"]
    print("  max|tc - simt| / max|simt| = %.3e" % (np.abs(a - b).max() / np.abs(b).max()))

if __name__ == "__main__":
    run("ljspeech_tacotron", 256, 200)
    run("ljspeech_self_attention", 16, 150)
