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


# ---- host-side model of the fused exchange (k_merge_exchange, csrc/orbx_match.cu): used by the CPU protocol tests ----
EPOCH_BITS = 14


def pack_exchange_word(epoch, best, second, idx):
    """64-bit slot word: epoch (14 bits) << 50 | best (9) << 41 | second (9) << 32 | index (32, -1 = none).
    numpy arrays in, uint64 array out; payload and the 'ready' tag travel in one word (one 64-bit store on the GPU)."""
    import numpy as np

    best = np.asarray(best).astype(np.uint64)
    second = np.asarray(second).astype(np.uint64)
    idx = np.asarray(idx).astype(np.int64).astype(np.uint64) & np.uint64(0xFFFFFFFF)
    ep = np.uint64(epoch & ((1 << EPOCH_BITS) - 1))
    return (ep << np.uint64(50)) | (best << np.uint64(41)) | (second << np.uint64(32)) | idx


def unpack_exchange_word(word):
    """-> (epoch, best, second, idx) as int arrays (idx -1 = none)."""
    import numpy as np

    word = np.asarray(word, np.uint64)
    epoch = (word >> np.uint64(50)).astype(np.int64)
    best = ((word >> np.uint64(41)) & np.uint64(0x1FF)).astype(np.int32)
    second = ((word >> np.uint64(32)) & np.uint64(0x1FF)).astype(np.int32)
    idx = (word & np.uint64(0xFFFFFFFF)).astype(np.uint32).astype(np.int64)
    idx = np.where(idx >= 1 << 31, idx - (1 << 32), idx).astype(np.int32)
    return epoch, best, second, idx


def exchange_slot_offset(epoch, src_rank, world, nq_max):
    """index (in 64-bit words) of the first slot rank `src_rank` writes for call `epoch`: two slot sets alternate by parity"""
    return ((epoch & 1) * world + src_rank) * nq_max
