"""CLI twin of /root/reference/predict_mel.py:1-15 over the CUDA hot path.

Usage: predict_mel.py --hparam-json-file=<path> [--hparams=<k=v,...>] --output-dir=<path>
                      [--weights=<npz>] [--sources=<npz>] [--batch-size=<n>]

Inputs are an .npz of utterances (keys ``<key>/source`` int64 ids, optional ``<key>/mel``) because the
TFRecord reader is a "next" row.  Without --weights the model is random-initialised
(seed 12345); without --sources a small synthetic list is synthesised."""
import argparse

import numpy as np

import sat_b200
from sat_b200.predict_mel import predict, Utterance


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hparam-json-file")
    ap.add_argument("--hparams", default="")
    ap.add_argument("--output-dir", required=True)
    ap.add_argument("--weights")
    ap.add_argument("--sources")
    ap.add_argument("--batch-size", type=int, default=None)
    args = ap.parse_args()
    hp = sat_b200.HParams()
    if args.hparam_json_file:
        hp.parse_json(args.hparam_json_file)
    hp.parse(args.hparams)
    print(sat_b200.hparams_debug_string(hp))
    W = dict(np.load(args.weights)) if args.weights else sat_b200.init_weights(hp)
    utts = []
    if args.sources:
        z = np.load(args.sources)
        keys = sorted({k.split("/")[0] for k in z.files})
        for i, k in enumerate(keys):
            utts.append(Utterance(i, k, z[k + "/source"], mel=z[k + "/mel"] if k + "/mel" in z.files else None))
    else:
        rng = np.random.default_rng(12345)
        utts = [Utterance(i, f"synthetic_{i:04d}", rng.integers(1, 71, size=int(rng.integers(20, 60)))) for i in range(4)]
    res = predict(hp, W, utts, output_dir=args.output_dir, batch_size=args.batch_size)
    for v in res:
        print(v.key, v.predicted_mel.shape)


if __name__ == "__main__":
    main()
