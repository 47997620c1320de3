"""Host-side sharding plan for the multi-GPU paths (one process per GPU, torch.distributed for the plumbing).

* Batched extraction shards by frame: rank r takes frames r, r+W, r+2W, ... -- independent units, no collective.
* Brute-force map matching shards the descriptor database by contiguous index range; each rank searches its range with
  `orbm_bruteforce(..., idx_offset=lo)`, the per-query (best, idx, second) triples are all-gathered (rank order =
  ascending index range) and merged by `orbm_merge_partials`, which reproduces the sequential scan of the whole
  database (first minimum wins; second = second smallest of the multiset).
"""


def frame_shard(n_frames, rank, world):
    """Indices of the frames rank `rank` extracts."""
    return list(range(rank, n_frames, world))


def db_shard(n_rows, rank, world):
    """Contiguous [lo, hi) database range of rank `rank`; ranges are ordered by rank and cover [0, n_rows)."""
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi
