"""Host-side sharding rules of the two paths over one 8 x B200 box (one process per GPU).

Distance: tile rows of the upper triangle are dealt in pairs (r, R-1-r) -- every pair holds R+1
tiles, so every rank gets the same triangle area -- and no collective runs until the gather
(the C side, lzg_distance_shard in csrc/lzg_api.cu, applies the same rule; this mirror is what
tests and bench.py use to predict which rows a rank owns).

NNI: every rank holds the whole tree, scores a contiguous slice of the host's candidate
enumeration and the ranks min-all-reduce one 64-bit key (score << 32 | global index), which
reproduces the reference's "strictly smaller, first in enumeration order" rule
(/root/reference/src/copy_number.cxx:353-357) across ranks; greedy mode reduces the index alone.
"""
import numpy as np

NO_MOVE = (1 << 63) - 1  # int64 max: "no improving candidate on this rank"


def tile_row_owner(n_cells, n_shards, bm=128):
    """Owner rank of every tile row."""
    R = (n_cells + bm - 1) // bm
    r = np.arange(R)
    return np.minimum(r, R - 1 - r) % n_shards


def shard_pairs(n_cells, shard, n_shards, bm=128):
    """Cell pairs (i <= j, as the reference's loop counts them) computed by one shard."""
    owner = tile_row_owner(n_cells, n_shards, bm)
    pairs = 0
    for r in np.flatnonzero(owner == shard):
        r0, r1 = r * bm, min(n_cells, (r + 1) * bm)
        rows = r1 - r0
        pairs += rows * (rows + 1) // 2 + rows * (n_cells - r1)
    return int(pairs)


def shard_rows(n_cells, shard, n_shards, bm=128):
    """Boolean mask of the matrix rows whose upper-triangle part this shard writes."""
    owner = tile_row_owner(n_cells, n_shards, bm)
    return np.repeat(owner == shard, bm)[:n_cells]


def candidate_slice(n_candidates, rank, world):
    """Contiguous slice [lo, hi) of the enumeration scored by `rank`.  Boundaries are even: the
    enumeration lists the two candidates of an edge back to back (src/copy_number.cxx:340-346)
    and the GPU pass scores such a pair as one work item."""
    pairs = (n_candidates + 1) // 2
    lo, hi = 2 * (pairs * rank // world), 2 * (pairs * (rank + 1) // world)
    return min(lo, n_candidates), min(hi, n_candidates)


def pack_key(score, global_index, greedy=False):
    """64-bit reduction key; NO_MOVE when the rank found nothing (index < 0)."""
    if global_index < 0:
        return NO_MOVE
    return int(global_index) if greedy else (int(score) << 32) | int(global_index)


def unpack_key(key, greedy=False):
    """(global index, score or None) of a reduced key; (-1, None) for NO_MOVE."""
    key = int(key)
    if key == NO_MOVE:
        return -1, None
    if greedy:
        return key, None
    return key & 0xFFFFFFFF, key >> 32


def allreduce_best(local_index, local_score, lo, greedy=False, device="cpu"):
    """Min-all-reduce of the ranks' best candidates.  local_index is relative to the rank's slice
    (-1 = none).  Returns (global index or -1, score or None).  Needs an initialised
    torch.distributed process group (NCCL over NVLink on the GPU box, gloo in CPU tests)."""
    import torch
    import torch.distributed as dist
    key = pack_key(local_score, lo + local_index if local_index >= 0 else -1, greedy)
    t = torch.tensor([key], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return unpack_key(int(t.item()), greedy)
