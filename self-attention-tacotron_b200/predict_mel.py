"""predict_mel entry point: the surface of /root/reference/predict_mel.py (``PredictedMel`` fields :28-33,
``predict`` :36-74, raw little-endian float32 ``.mfbsp`` output :58-62) fed by the CUDA hot path.

Differences from the reference: batched (the reference forces batch_size=1, :43-44), utterances are
length-bucketed and sharded over the GPUs of one box (``sharding.py``), inputs arrive as arrays or through
``load_utterances`` (the reference's source / target TFRecord files, read by ``tfrecord.py`` without TensorFlow), and no
PNG plot is written (no plotting stack).  Next to every ``.mfbsp`` a prediction TFRecord is written (:70-74)."""
import os
from collections import namedtuple
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from . import tfrecord
from .models import tacotron_model_factory
from .sharding import make_buckets


class PredictedMel(
    namedtuple("PredictedMel",
               ["id", "key", "predicted_mel", "predicted_mel_postnet", "predicted_mel_width", "predicted_target_length",
                "ground_truth_mel", "alignment", "alignment2", "alignment3", "alignment4", "alignment5", "alignment6",
                "source", "text", "accent_type"])):
    pass


class Utterance(namedtuple("Utterance", ["id", "key", "source", "text", "mel", "speaker_id"])):
    """One synthesis request: ``source`` int64 symbol ids [T_enc]; ``mel`` optional ground truth [T_dec, M]."""

    def __new__(cls, id, key, source, text="", mel=None, speaker_id=None):
        return super().__new__(cls, id, key, np.asarray(source, dtype=np.int64), text, mel, speaker_id)


def _pad_batch(utts, hp, device):
    B = len(utts)
    T = max(len(u.source) for u in utts)
    ids = np.zeros((B, T), dtype=np.int64)  # batch padding id 0 (datasets/ljspeech/dataset.py:268)
    lengths = np.zeros(B, dtype=np.int64)
    for i, u in enumerate(utts):
        ids[i, :len(u.source)] = u.source
        lengths[i] = len(u.source)
    target = None
    if all(u.mel is not None for u in utts):
        r = hp.outputs_per_step
        Td = max(u.mel.shape[0] for u in utts)
        Td = (Td + r - 1) // r * r
        # r-padding with silence (datasets/ljspeech/dataset.py:134-135,275)
        tgt = np.full((B, Td, hp.num_mels), hp.silence_mel_level_db, dtype=np.float32)
        for i, u in enumerate(utts):
            tgt[i, :u.mel.shape[0]] = u.mel
        target = torch.from_numpy(tgt).pin_memory().to(device, non_blocking=True)
    spk = None
    if hp.use_speaker_embedding:
        if hp.speaker_for_synthesis > -1:
            sid = [int(hp.speaker_for_synthesis)] * B
        elif any(u.speaker_id is None for u in utts):
            missing = [u.key for u in utts if u.speaker_id is None][:3]
            raise ValueError(f"use_speaker_embedding is set but utterances {missing} carry no speaker_id "
                             "(VCTK / codes source records have one: datasets/vctk/dataset.py:70)")
        else:
            sid = [int(u.speaker_id) for u in utts]
        spk = torch.tensor(sid, dtype=torch.int64).to(device)
    return (torch.from_numpy(ids).pin_memory().to(device, non_blocking=True),
            torch.from_numpy(lengths).pin_memory().to(device, non_blocking=True), target, spk)


COPY_THREADS = max(2, min(8, (os.cpu_count() or 8) // 2))  # host threads that copy results out of the pinned buffers


class _HostSlot:
    """Pinned host buffers for the results of one batch, reused every other batch (two slots = the read-back of batch
    i overlaps the compute of batch i+1)."""

    def __init__(self):
        self.bufs = {}
        self.copied = torch.cuda.Event()

    def fetch(self, name, dev_tensor):
        """enqueue dev_tensor -> pinned host copy on the current (copy) stream; returns the host tensor"""
        n = dev_tensor.numel()
        buf = self.bufs.get(name)
        if buf is None or buf.numel() < n or buf.dtype != dev_tensor.dtype:
            buf = torch.empty(max(n, 1), dtype=dev_tensor.dtype, pin_memory=True)
            self.bufs[name] = buf
        host = buf[:n].view(dev_tensor.shape)
        host.copy_(dev_tensor, non_blocking=True)
        return host


RETIRE_SPREAD = 0.7  # predict(): rows are retired in batches whose shortest source is below this fraction of the longest


def predict(hparams, weights, utterances, output_dir=None, batch_size=None, device=None, rank=0, world_size=1,
            mode="free_run", model=None, batch_hook=None, pinned_results=False, retire_rows=None):
    """Synthesise ``utterances`` (list of Utterance) -> list[PredictedMel] for the ones this rank owns.

    mode: 'free_run' (stop-token inference, PREDICT), 'validation' / 'teacher_forced' (EVAL forms).
    With ``hparams.use_forced_alignment_mode`` the two-pass forced alignment of models/models.py:118-134 runs.

    Batches are pipelined: the results of batch i (mel, post-net mel, stop steps, alignment histories) are read back
    into pinned host memory on a copy stream while batch i+1 is padded, uploaded and computed; slicing per utterance
    and the file writes of batch i happen on the host during that compute.  ``batch_hook(i)`` (optional) runs on the
    compute stream before batch i is launched (bench.py flushes L2 there).

    ``retire_rows``: None (default) -- per batch, by the spread of the source lengths (RETIRE_SPREAD); True / False force it.
    The results are the same either way (frames up to each row's own stop step).

    ``pinned_results``: False (default) -- two pinned buffer sets are reused and every batch's results are copied out
    of them into ordinary arrays (host threads, in row blocks); right for a one-shot run that keeps all its results.
    True -- every batch gets its own pinned tensors and the returned arrays are VIEWS of them (no second host copy);
    right for a serving loop that drops results after use: torch's caching host allocator then recycles the blocks,
    whereas a caller that accumulates results pays a cudaHostAlloc (~80 ms per 184 MB) per batch."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    if device.type != "cuda":
        raise RuntimeError("predict_mel runs on CUDA devices only (no CPU fallback)")
    model = model or tacotron_model_factory(hparams, weights)
    # The reference's predict forces batch_size=1 (predict_mel.py:43-44), so an utterance never sees padding and stops on
    # its own stop token.  Batches here keep that per-utterance result: padded encoder frames are masked, and every
    # row stops by its own rule (the run ends when the last row has stopped).
    saved = (model.encoder.mask_padding, model.decoder.stop_mode, model.decoder.async_results, model.decoder.retire_rows)
    model.encoder.set_mask_padding(True)
    model.decoder.stop_mode = "per_row"
    model.decoder.async_results = not hparams.use_forced_alignment_mode  # no host sync inside model.predict()
    batch_size = batch_size or hparams.batch_size
    r = hparams.outputs_per_step
    results = []
    buckets = make_buckets([len(u.source) for u in utterances], batch_size, world_size,
                           bucket_width=hparams.batch_bucket_width)
    copy_stream = torch.cuda.Stream(device=device)
    slots = [_HostSlot(), _HostSlot()]
    pool = ThreadPoolExecutor(max_workers=COPY_THREADS)

    def launch(i, idxs):
        utts = [utterances[j] for j in idxs]
        ids, lengths, target, spk = _pad_batch(utts, hparams, device)
        need_target = mode != "free_run" or hparams.use_forced_alignment_mode
        if need_target and target is None:
            raise ValueError(f"mode {mode!r} needs ground-truth mel for every utterance")
        if batch_hook is not None:
            batch_hook(i)
        # Row retirement costs ~10 % per step until rows stop and pays once they stop far apart.  Output length follows
        # source length, and the buckets hold utterances of similar length: retire only in batches whose shortest source
        # is less than RETIRE_SPREAD of the longest (ragged tails, small corpora, fixed-width buckets).
        model.decoder.retire_rows = (retire_rows if retire_rows is not None else
                                     min(len(u.source) for u in utts) < RETIRE_SPREAD * max(len(u.source) for u in utts))
        out = model.predict(ids, lengths, target=target if need_target else None, speaker_id=spk,
                            is_validation=mode != "free_run", teacher_forcing=mode == "teacher_forced")
        done = torch.cuda.Event()
        done.record()
        slot = _HostSlot() if pinned_results else slots[i % 2]
        host = {}
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(done)
            for k in ("mel", "mel_postnet", "stop_steps", "n_steps", "alignment", "alignment2"):
                v = out.get(k)
                if not torch.is_tensor(v):
                    host[k] = v  # n_steps of a fixed-length run is a Python int
                    continue
                if k.startswith("alignment"):
                    v = v.transpose(1, 2)  # back to the contiguous [B, S, T] history the kernel wrote
                host[k] = slot.fetch(k, v)
                v.record_stream(copy_stream)
            slot.copied.record()
        return utts, [len(u.source) for u in utts], host, slot

    def finish(utts, src_len, host, slot):
        slot.copied.synchronize()
        # a symbol id outside the embedding table raises here (tacotron2 Embedding asserts it); no synchronisation:
        # the batch's forward has finished, the read-back waited for it
        model.encoder.poll_error()
        stop_steps = host["stop_steps"].numpy().copy()
        n_steps = host["n_steps"]
        if torch.is_tensor(n_steps):
            n_steps = int(n_steps.numpy().reshape(-1)[0])
        # ONE copy per result tensor out of the pinned buffers (they are reused two batches later); the utterances of the
        # batch then share these arrays as views -- 256 separate copies cost 3x as much in page faults
        # (NumPy copies release the GIL: the tensors are copied side by side on a small thread pool)
        jobs = {"mel": host["mel"].numpy()[:, :n_steps * r]}
        views = pinned_results
        if host.get("mel_postnet") is not None:
            jobs["post"] = host["mel_postnet"].numpy()[:, :n_steps * r]
        for k in ("alignment", "alignment2"):
            if host.get(k) is not None:
                jobs[k] = host[k].numpy()[:, :n_steps]  # [B, S, T]
        if views:
            copies = jobs  # the arrays keep the batch's pinned tensors alive
        else:
            copies = {k: np.empty(v.shape, dtype=v.dtype) for k, v in jobs.items()}
            parts = []
            for k, v in jobs.items():  # row blocks of every tensor, so that all pool threads fault in and fill pages
                step = max(1, -(-v.shape[0] // COPY_THREADS))
                parts += [(copies[k][a:a + step], v[a:a + step]) for a in range(0, v.shape[0], step)]
            list(pool.map(lambda dv: np.copyto(dv[0], dv[1]), parts))
        mel, post = copies["mel"], copies.get("post")
        aligns = [copies[k] for k in ("alignment", "alignment2") if k in copies]
        for i, u in enumerate(utts):
            # frames kept per utterance: up to and including its own stop step (free run), else all steps
            n = n_steps if (mode != "free_run" or stop_steps[i] < 0) else int(stop_steps[i]) + 1
            L = src_len[i]
            # (T_memory, T_query) like models/models.py:104-115: a transposed VIEW of the batch's history array
            al = [a[i, :n, :L].T for a in aligns] + [None] * (6 - len(aligns))
            m = mel[i, :n * r]
            pm = PredictedMel(u.id, u.key, m, post[i, :n * r] if post is not None else None, m.shape[1],
                              m.shape[0], u.mel, al[0], al[1], al[2], al[3], al[4], al[5], u.source, u.text, None)
            results.append(pm)
            if output_dir is not None:
                write_predicted_mel(pm, hparams, output_dir)

    try:
        pending = None
        for i, idxs in enumerate(buckets[rank]):
            cur = launch(i, idxs)  # slot i % 2 is free: batch i - 2 was finished before batch i - 1 was launched
            if pending is not None:
                finish(*pending)   # host work of batch i - 1 while batch i computes
            pending = cur
        if pending is not None:
            finish(*pending)
    finally:
        pool.shutdown(wait=True)
        model.encoder.set_mask_padding(saved[0])
        model.decoder.stop_mode, model.decoder.async_results, model.decoder.retire_rows = saved[1], saved[2], saved[3]
    return results


def write_predicted_mel(v, hparams, output_dir):
    """predict_mel.py:56-62: ``<key>.<predicted_mel_extension>`` holding raw '<f4' frames [T_dec, num_mels]."""
    key = v.key.decode("utf-8") if isinstance(v.key, bytes) else str(v.key)
    mel = v.predicted_mel_postnet if hparams.use_postnet_v2 else v.predicted_mel
    assert mel.shape[1] == hparams.num_mels
    os.makedirs(output_dir, exist_ok=True)
    path = os.path.join(output_dir, f"{key}.{hparams.predicted_mel_extension}")
    np.ascontiguousarray(mel, dtype="<f4").tofile(path)
    # predict_mel.py:70-74: the prediction record next to it (TFRecord, tf.train.Example; see tfrecord.py)
    text = v.text.decode("utf-8") if isinstance(v.text, bytes) else str(v.text or "")
    alignments = [a for a in (v.alignment, v.alignment2, v.alignment3, v.alignment4, v.alignment5, v.alignment6)
                  if a is not None]
    tfrecord.write_prediction_result(v.id, key, alignments, mel, v.ground_truth_mel, text, v.source, v.accent_type,
                                     os.path.join(output_dir, f"{key}.tfrecord"))
    return path


def load_utterances(source_files, target_files=None, hparams=None):
    """Source (and optional target) TFRecord files written by the reference's preprocessing
    (preprocess/ljspeech.py:23-44; VCTK / codes form datasets/codes/dataset.py:66-96) -> list[Utterance]; targets are
    matched to sources by key.  With ``hparams.source == 'phone'`` the phone sequence of a VCTK / codes record is
    the model input (datasets/codes/dataset.py:147-149); the record's speaker_id travels with the utterance."""
    targets = {}
    for f in target_files or []:
        if not os.path.exists(f):
            continue
        for t in tfrecord.read_target_data(f):
            targets[t.key] = t.mel
    utts = []
    for f in source_files:
        for s in tfrecord.read_source_data(f):
            ids, text = s.source, s.text
            if hparams is not None and getattr(hparams, "source", None) == "phone" and s.phone is not None:
                ids, text = s.phone, s.phone_txt
            utts.append(Utterance(s.id, s.key, ids, text=text, mel=targets.get(s.key), speaker_id=s.speaker_id))
    return utts


def load_key_list(filename, in_dir):
    """predict_mel.py:77-81: one utterance key per line."""
    with open(os.path.join(in_dir, filename), mode="r", encoding="utf-8") as f:
        return [line.rstrip("\n") for line in f if line.rstrip("\n")]


def selected_files(hparams, source_data_root, target_data_root, selected_list_dir, selected_list_filename="test.csv"):
    """predict_mel.py:106-112: <root>/<key>.<source_file_extension> / <key>.<target_file_extension> for every key of
    the selected list."""
    keys = load_key_list(selected_list_filename or "test.csv", selected_list_dir)
    src = [os.path.join(source_data_root, f"{k}.{hparams.source_file_extension}") for k in keys]
    tgt = [os.path.join(target_data_root, f"{k}.{hparams.target_file_extension}") for k in keys] if target_data_root else []
    return src, tgt
