"""CLI twin of /root/reference/predict_mel.py:1-15 over the CUDA hot path.

Usage (the reference's own options, predict_mel.py:4-14, so its Slurm scripts run unchanged):
    predict_mel.py --source-data-root=<dir> --target-data-root=<dir> --selected-list-dir=<dir>
                   [--selected-list-filename=<name>] --checkpoint-dir=<dir> [--checkpoint=<path>]
                   --hparam-json-file=<path> [--hparams=<k=v,...>] --output-dir=<path>
  -> synthesises <source-data-root>/<key>.<source_file_extension> for every key of <selected-list-dir>/<test.csv>
     (predict_mel.py:84-119); ground truth from <target-data-root>/<key>.<target_file_extension> when present.
Additional options of this implementation:
    [--name-map=<json>] [--weights=<npz>] [--sources=<npz>] [--source-tfrecords=<glob>] [--target-tfrecords=<glob>]
    [--batch-size=<n>] [--mode=free_run|validation|teacher_forced]

Weights: ``--checkpoint`` (a TensorFlow V2 checkpoint prefix) or the latest checkpoint of ``--checkpoint-dir`` as in
predict_mel.py:47-54, read without TensorFlow by sat_b200.checkpoint (``--name-map``: JSON {checkpoint variable:
weight name} for variables the built-in rules do not recognise); or ``--weights``, an .npz weight blob.

Inputs: the reference's preprocessed source / target TFRecord files (--source-tfrecords / --target-tfrecords,
preprocess/ljspeech.py:23-44; read without TensorFlow by sat_b200.tfrecord) or an .npz of utterances (keys
``<key>/source`` int64 ids, optional ``<key>/mel``).  Outputs per utterance, like predict_mel.py:56-74: ``<key>.mfbsp``
(raw '<f4' frames) and ``<key>.tfrecord`` (prediction record).  Without --weights the model is random-initialised
(seed 12345); without any source a small synthetic list is synthesised."""
import glob
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
    ap.add_argument("--checkpoint")
    ap.add_argument("--checkpoint-dir")
    ap.add_argument("--name-map")
    ap.add_argument("--sources")
    ap.add_argument("--source-tfrecords")
    ap.add_argument("--target-tfrecords")
    ap.add_argument("--batch-size", type=int, default=None)
    ap.add_argument("--source-data-root")
    ap.add_argument("--target-data-root")
    ap.add_argument("--selected-list-dir")
    ap.add_argument("--selected-list-filename", default="test.csv")
    ap.add_argument("--mode", default="free_run", choices=["free_run", "validation", "teacher_forced"])
    args = ap.parse_args()
    hp = sat_b200.HParams()
    if args.hparam_json_file:
        hp.parse_json(args.hparam_json_file)
    hp.parse(args.hparams)
    print(sat_b200.hparams_debug_string(hp))
    if args.checkpoint or args.checkpoint_dir:
        import json
        from sat_b200 import checkpoint
        prefix = args.checkpoint or checkpoint.latest_checkpoint(args.checkpoint_dir)
        if prefix is None:
            raise SystemExit(f"no checkpoint found in {args.checkpoint_dir}")
        name_map = json.load(open(args.name_map)) if args.name_map else None
        W = checkpoint.import_weights(hp, prefix, name_map=name_map)
    else:
        W = dict(np.load(args.weights)) if args.weights else sat_b200.init_weights(hp)
    utts = []
    if args.source_data_root:
        # predict_mel.py:104-112: key list -> <root>/<key>.<extension> files
        from sat_b200.predict_mel import load_utterances, selected_files
        if not args.selected_list_dir:
            raise SystemExit("--source-data-root needs --selected-list-dir (directory holding test.csv)")
        src, tgt = selected_files(hp, args.source_data_root, args.target_data_root, args.selected_list_dir,
                                  args.selected_list_filename)
        print(f"A selected list file to use: {args.selected_list_dir}/{args.selected_list_filename} ({len(src)} keys)")
        utts = load_utterances(src, tgt, hp)
    elif args.source_tfrecords:
        from sat_b200.predict_mel import load_utterances
        utts = load_utterances(sorted(glob.glob(args.source_tfrecords)),
                               sorted(glob.glob(args.target_tfrecords)) if args.target_tfrecords else None, hp)
    elif args.sources:
        z = np.load(args.sources)
        keys = sorted({k.split("/")[0] for k in z.files})
        for i, k in enumerate(keys):
            utts.append(Utterance(i, k, z[k + "/source"], mel=z[k + "/mel"] if k + "/mel" in z.files else None))
    else:
        rng = np.random.default_rng(12345)
        utts = [Utterance(i, f"synthetic_{i:04d}", rng.integers(1, 71, size=int(rng.integers(20, 60)))) for i in range(4)]
    res = predict(hp, W, utts, output_dir=args.output_dir, batch_size=args.batch_size, mode=args.mode)
    for v in res:
        print(v.key, v.predicted_mel.shape)


if __name__ == "__main__":
    main()
