"""Host-side length bucketing and sharding of utterances over the GPUs of one box.

Inference is embarrassingly data-parallel: each rank owns whole utterances, weights are replicated and
there is no collective on the compute path.  The reference's own bucketing (datasets/ljspeech/dataset.py:
235-286, group_by_window + padded_batch) effectively puts every utterance in bucket 0 because of
``tf.minimum(..., 0)`` at :241-244; this module does real bucketing by source length."""
import numpy as np


# measured on B200 (profiles/README.md): one decoder step costs about STEP_US0 + STEP_US1 * B microseconds for
# B <= 256 (grid-barrier / latency floor plus a small per-utterance term), and the number of steps of a batch is
# set by its longest utterance
STEP_US0, STEP_US1 = 30.0, 0.27


def _bucket_cuts_numpy(L, batch_size, c0, c1):
    """the dynamic programme of make_buckets, one NumPy row per utterance (what sat_bucket_cuts does natively)"""
    n = len(L)
    L = L.astype(np.float64)
    best = np.zeros(n + 1)
    cut = np.zeros(n, dtype=np.int32)
    for i in range(n - 1, -1, -1):
        ks = np.arange(1, min(batch_size, n - i) + 1)
        c = L[i] * (c0 + c1 * ks) + best[i + ks]
        j = int(np.argmin(c))
        best[i], cut[i] = c[j], ks[j]
    return cut


def _bucket_cuts(L, batch_size, c0, c1):
    """cut[i] = size of the batch starting at i; libsat.so's sat_bucket_cuts (the programme is O(n * batch_size) and sits
    in front of the first launch of every predict() call: 10 ms per 2048 utterances in NumPy, 0.2 ms natively)"""
    import ctypes as C
    from . import _lib
    L = np.ascontiguousarray(L, dtype=np.int64)
    cut = np.zeros(len(L), dtype=np.int32)
    _lib.check(_lib.lib().sat_bucket_cuts(L.ctypes.data_as(C.c_void_p), C.c_int64(len(L)), C.c_int32(batch_size),
                                          C.c_double(c0), C.c_double(c1), cut.ctypes.data_as(C.c_void_p)))
    return cut


def make_buckets(lengths, batch_size, world_size=1, bucket_width=50, cost_model=(STEP_US0, STEP_US1)):
    """-> per-rank list of batches; each batch is a list of utterance indices of similar length.

    Utterances are sorted by length (longest first) and dealt round-robin to the ranks, so every rank
    receives the same length distribution and therefore the same amount of work.  Each rank then cuts
    its own share, still sorted, into batches of at most ``batch_size``:

    * ``cost_model=(c0, c1)`` (default): the cuts minimise the modelled decode time
      ``sum_batches longest(batch) * (c0 + c1 * len(batch))`` by dynamic programming over the sorted
      share -- a decoder step costs nearly the same for 16 and for 128 utterances, so a short tail is
      only split off when the steps it saves outweigh the extra batch;
    * ``cost_model=None``: fixed-width buckets -- a batch is closed when it is full or, once it is at
      least half full, at a ``bucket_width`` boundary.

    Deterministic; every index appears exactly once."""
    lengths = np.asarray(lengths, dtype=np.int64)
    per_rank = [[] for _ in range(world_size)]
    if len(lengths) == 0:
        return per_rank
    order = np.argsort(-lengths, kind="stable")
    width = max(int(bucket_width), 1)
    batch_size = max(int(batch_size), 1)
    for k in range(world_size):
        share = order[k::world_size]
        n = len(share)
        if n == 0:
            continue
        if cost_model is not None:
            cut = _bucket_cuts(lengths[share], batch_size, *cost_model)
            i = 0
            while i < n:
                per_rank[k].append(share[i:i + cut[i]].tolist())
                i += int(cut[i])
            continue
        cur, cur_bucket = [], None
        for i in share:
            b = int(lengths[i]) // width
            if cur and (len(cur) >= batch_size or (b != cur_bucket and len(cur) >= max(1, batch_size // 2))):
                per_rank[k].append(cur)
                cur = []
            if not cur:
                cur_bucket = b
            cur.append(int(i))
        if cur:
            per_rank[k].append(cur)
    return per_rank


def gather_results(local, world_size, rank):
    """Gather per-rank python results on rank 0 with torch.distributed (NCCL/gloo object gather); the
    only collective of the inference path, after all compute is done."""
    if world_size == 1:
        return local
    import torch.distributed as dist
    out = [None] * world_size if rank == 0 else None
    dist.gather_object(local, out, dst=0)
    if rank != 0:
        return None
    return [x for part in out for x in part]


def gather_mels(ids, mels, world_size, rank, device=None):
    """Gather variable-length per-utterance frames [T_i, M] of every rank on rank 0 with tensor collectives
    (SURVEY.md 8e): one all_reduce(MAX) for the padded shape, then all_gather of the padded block
    [n_max, T_max, M] and of (utterance id, frame count) -- NCCL over NVLink for CUDA tensors, gloo for CPU
    tensors.  The only collective of the inference path; it runs after all compute is done.

    ids: utterance ids owned by this rank; mels: list of torch tensors [T_i, M] (same device, fp32).
    -> on rank 0 a dict id -> tensor [T_i, M] (on ``device``) over all ranks; None elsewhere."""
    import torch
    if world_size == 1:
        return {int(i): m for i, m in zip(ids, mels)}
    import torch.distributed as dist
    device = device or (mels[0].device if mels else torch.device("cpu"))
    M = int(mels[0].shape[1]) if mels else 0
    shape = torch.tensor([len(mels), max([int(m.shape[0]) for m in mels], default=0), M], dtype=torch.int64, device=device)
    dist.all_reduce(shape, op=dist.ReduceOp.MAX)
    n_max, t_max, M = (int(x) for x in shape.tolist())
    block = torch.zeros((n_max, t_max, M), dtype=torch.float32, device=device)
    meta = torch.full((n_max, 2), -1, dtype=torch.int64, device=device)
    for k, (i, m) in enumerate(zip(ids, mels)):
        block[k, :m.shape[0]] = m
        meta[k, 0], meta[k, 1] = int(i), int(m.shape[0])
    blocks = [torch.empty_like(block) for _ in range(world_size)]
    metas = [torch.empty_like(meta) for _ in range(world_size)]
    dist.all_gather(blocks, block)
    dist.all_gather(metas, meta)
    if rank != 0:
        return None
    out = {}
    for blk, mt in zip(blocks, metas):
        for k, (i, n) in enumerate(mt.tolist()):
            if i >= 0:
                out[int(i)] = blk[k, :n]
    return out
