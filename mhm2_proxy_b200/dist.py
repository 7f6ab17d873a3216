"""Rank plumbing for the one-process-per-GPU launch (torch.distributed is used for bootstrap only: the supermer
exchange itself is NCCL inside the CUDA library).

The reference splits its input by file offset per UPC++ rank (fastq.cpp) and shards the table by
quick_hash(minimizer) % rank_n (kmer_dht.cpp:192-196); here reads are split by index and the library does the rest.
"""


def split_range(n, rank, world):
    """Contiguous, balanced slice [begin, end) of n items for `rank` of `world` (every item in exactly one slice)."""
    return n * rank // world, n * (rank + 1) // world


def share_unique_id(rank, make_id, src=0):
    """Rank `src` creates the 128-byte NCCL unique id (make_id()), everyone receives it over the bootstrap group."""
    import torch.distributed as dist
    box = [make_id() if rank == src else None]
    dist.broadcast_object_list(box, src=src)
    uid = box[0]
    if not isinstance(uid, (bytes, bytearray)) or len(uid) != 128:
        raise RuntimeError("bad unique id received")
    return bytes(uid)
