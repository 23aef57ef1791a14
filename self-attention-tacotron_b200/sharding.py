"""Host-side length bucketing and sharding of utterances over the GPUs of one box.

Inference is embarrassingly data-parallel: each rank owns whole utterances, weights are replicated and
there is no collective on the compute path.  The reference's own bucketing (datasets/ljspeech/dataset.py:
235-286, group_by_window + padded_batch) effectively puts every utterance in bucket 0 because of
``tf.minimum(..., 0)`` at :241-244; this module does real bucketing by source length."""
import numpy as np


def make_buckets(lengths, batch_size, world_size=1, bucket_width=50):
    """-> per-rank list of batches; each batch is a list of utterance indices of similar length.

    Utterances are sorted by length (longest first) and dealt round-robin to the ranks, so every rank
    receives the same length distribution and therefore the same amount of work; each rank then cuts
    its own share, still sorted, into batches of at most ``batch_size`` whose lengths span at most
    ``bucket_width`` symbols where the share allows it (a batch is only closed early at a bucket
    boundary once it is at least half full).  Deterministic; every index appears exactly once."""
    lengths = np.asarray(lengths, dtype=np.int64)
    per_rank = [[] for _ in range(world_size)]
    if len(lengths) == 0:
        return per_rank
    order = np.argsort(-lengths, kind="stable")
    width = max(int(bucket_width), 1)
    for k in range(world_size):
        cur, cur_bucket = [], None
        for i in order[k::world_size]:
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
