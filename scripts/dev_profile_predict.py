"""Developer aid: host-side profile (cProfile) of predict_mel.predict on the bench workload (k batches of 256)."""
import cProfile, pstats, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200, bench
from sat_b200.models import tacotron_model_factory
from sat_b200.predict_mel import predict, Utterance
hp = sat_b200.config("ljspeech_tacotron")
W = sat_b200.init_weights(hp, seed=12345)
ids, lens = bench.workload(hp)
B, k = ids.shape[0], 4
utts = [Utterance(b + j * B, f"u{b + j * B}", ids[b, :lens[b]]) for j in range(k) for b in range(B)]
model = tacotron_model_factory(hp, W)
dev = torch.device("cuda:0")
predict(hp, W, utts[:2 * B], batch_size=B, device=dev, model=model)
torch.cuda.synchronize()
t0 = time.perf_counter()
pr = cProfile.Profile()
pr.enable()
res = predict(hp, W, utts, batch_size=B, device=dev, model=model)
pr.disable()
torch.cuda.synchronize()
print(f"{(time.perf_counter() - t0) * 1e3 / k:.1f} ms per batch")
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
