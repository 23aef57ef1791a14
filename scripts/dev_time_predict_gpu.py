"""Developer aid: GPU-side cadence of predict_mel.predict on the bench workload (k batches of 256): CUDA events at the
start of every batch on the compute stream, next to model.predict() alone in the same modes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200, bench
from sat_b200.models import tacotron_model_factory
from sat_b200.predict_mel import predict, Utterance, _pad_batch
hp = sat_b200.config("ljspeech_tacotron")
W = sat_b200.init_weights(hp, seed=12345)
ids, lens = bench.workload(hp)
B, k = ids.shape[0], 8
utts = [Utterance(b + j * B, f"u{b + j * B}", ids[b, :lens[b]]) for j in range(k) for b in range(B)]
model = tacotron_model_factory(hp, W)
dev = torch.device("cuda:0")
evs, host_t = [], []


def hook(i):
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    evs.append(e)
    host_t.append(time.perf_counter())


kw = dict(batch_size=B, device=dev, model=model, batch_hook=hook, pinned_results=True)
predict(hp, W, utts[:B], **kw)
predict(hp, W, utts, **kw)
for rep in range(2):
    evs.clear(); host_t.clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = predict(hp, W, utts, **kw)
    end = torch.cuda.Event(enable_timing=True); end.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    print(f"predict: {(t1 - t0) * 1e3 / k:.2f} ms per batch wall; GPU cadence",
          " ".join(f"{evs[i].elapsed_time(evs[i + 1]):.1f}" for i in range(k - 1)),
          f"last->end {evs[-1].elapsed_time(end):.1f}; first hook at {(host_t[0] - t0) * 1e3:.1f} ms; host cadence",
          " ".join(f"{(host_t[i + 1] - host_t[i]) * 1e3:.1f}" for i in range(k - 1)))

# model.predict alone on resident inputs, in predict()'s modes and in the bench's, interleaved (clocks drift)
x, l, _, _ = _pad_batch(utts[:B], hp, dev)
modes = (("bench modes (all_rows, unmasked)", False, "all_rows"), ("per_row, unmasked", False, "per_row"),
         ("per_row, masked (predict's)", True, "per_row"))
model.decoder.async_results = True
acc = {m[0]: [] for m in modes}
kacc = {m[0]: [] for m in modes}
for rep in range(4):
    for name, mask, stop in modes:
        model.encoder.set_mask_padding(mask)
        model.decoder.stop_mode = stop
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        o = model.predict(x, l)
        b.record()
        torch.cuda.synchronize()
        if rep:
            acc[name].append(a.elapsed_time(b))
            kacc[name].append([h.kernel_ms() for h, _ in model.decoder._handles.values()][-1])
for name in acc:
    print(f"model.predict {name}: {np.round(acc[name], 2)} ms, decoder kernel {np.round(kacc[name], 2)} ms; rows stopped",
          int((o["stop_steps"] >= 0).sum().item()) if torch.is_tensor(o.get("stop_steps")) else None)
