"""Host-side sharding rules of the two paths over one 8 x B200 box (one process per GPU).

Distance: tile rows of the upper triangle are dealt in pairs (r, R-1-r) -- every pair holds R+1
tiles, so every rank gets the same triangle area -- and no collective runs until the gather
(the C side, lzg_distance_shard in csrc/lzg_api.cu, applies the same rule; this mirror is what
tests and bench.py use to predict which rows a rank owns).

NNI, positions sharded (PositionShardedTree, the one that scales): every quantity of a
neighbourhood pass is a sum over positions (src/copy_number.cxx:227-277 merges position by
position), so every rank keeps the same topology over its own column slice of the profiles, runs
the whole pass on 1/world of the bytes, and the ranks sum-all-reduce ONE int32 vector
{root score, candidate scores} (NCCL over NVLink, ordered on the context's stream -- no host wait)
before the arg-min.  Moves are applied by every rank; nothing else is exchanged.

NNI, candidates sharded (the first cut, kept for comparison): every rank holds the whole tree, scores a contiguous slice of the host's candidate
enumeration and the ranks min-all-reduce one 64-bit key (score << 32 | global index), which
reproduces the reference's "strictly smaller, first in enumeration order" rule
(/root/reference/src/copy_number.cxx:353-357) across ranks; greedy mode reduces the index alone.
"""
import ctypes as C
import os

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


def position_slice(n_pos, rank, world, align=16):
    """Columns [lo, hi) of the breakpoint profiles held by `rank`; boundaries are multiples of
    `align` (the device rows are padded to 16-byte chunks) except the last one."""
    units = (n_pos + align - 1) // align
    lo, hi = align * (units * rank // world), align * (units * (rank + 1) // world)
    return min(lo, n_pos), min(hi, n_pos)


class _DeviceVector:
    """int32 device memory owned by the C library, seen through __cuda_array_interface__."""

    def __init__(self, address, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (address, False), "version": 2}


def device_vector(address, n, device):
    """torch view (no copy) of n int32 at a device address."""
    import torch
    return torch.as_tensor(_DeviceVector(address, n), device="cuda:%d" % device)


class PositionShardedTree:
    """This rank's part of a tree whose positions are dealt over the ranks of a torch.distributed
    group.  Same calls as api.Tree; scores are global (summed over the ranks)."""

    def __init__(self, ctx, bp, child0, child1, leaf_row, rank=None, world=None, group=None):
        import torch
        import torch.distributed as dist
        from .api import Profiles, Tree
        self.ctx, self.group = ctx, group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.lo, self.hi = position_slice(bp.shape[1], self.rank, self.world)
        if self.hi <= self.lo:
            raise ValueError("more ranks than 16-position chunks")
        self.profiles = Profiles(ctx, np.ascontiguousarray(bp[:, self.lo:self.hi]))
        self.tree = Tree(ctx, self.profiles, child0, child1, leaf_row)
        self._stream = torch.cuda.ExternalStream(ctx.stream(), device=ctx.device)
        self.root = None
        # The all-reduce inside the library (lzg_nni_best_sharded: launch, ncclAllReduce and arg-min queued
        # back to back by one C call).  The ranks agree on an NCCL unique id through the process group;
        # LZG_SHARD_NCCL=0, or a libnccl the library cannot open, leaves the exchange to torch.distributed.
        self.comm = None
        if os.environ.get("LZG_SHARD_NCCL", "1") != "0":
            ident = [torch.cuda.nccl.unique_id() if self.rank == 0 else None]
            if self.world > 1:
                dist.broadcast_object_list(ident, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            handle = C.c_void_p()
            rc = ctx.lib.lzg_comm_init(ctx.h, self.rank, self.world, ident[0], C.byref(handle))
            ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device="cuda:%d" % ctx.device)
            if self.world > 1:
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)   # every rank or none
            if int(ok.item()) == 1:
                self.comm = handle
            elif rc == 0:
                ctx.lib.lzg_comm_free(handle)
        self._vec, self._vec_key = None, None    # torch view of the exchange vector (same address while the move count is)

    def free(self):
        if self.comm is not None:
            self.ctx.lib.lzg_comm_free(self.comm)
            self.comm = None
        self.tree.free()
        self.profiles.free()

    def _sum(self, tensor):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            with torch.cuda.stream(self._stream):
                dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=self.group)

    def small_rectilinear(self):
        import torch
        part = self.tree.small_rectilinear()
        t = torch.tensor([part], dtype=torch.int32, device="cuda:%d" % self.ctx.device)
        self._stream.wait_stream(torch.cuda.current_stream(self.ctx.device))
        self._sum(t)
        self._stream.synchronize()
        self.root = int(t.item())
        return self.root

    def edges(self):
        return self.tree.edges()

    def enumerate_moves(self, eu, ev, order):
        return self.tree.enumerate_moves(eu, ev, order)

    def nni_best(self, moves, greedy=False, want_all=False):
        import time
        if self.comm is not None:
            moves = np.ascontiguousarray(moves, np.int32).reshape(-1, 4)
            if len(moves) == 0:
                return -1, self.root, (np.zeros(0, np.int32) if want_all else None)
            bi, bs = C.c_int32(0), C.c_int32(0)
            allsc = np.zeros(len(moves), np.int32) if want_all else None
            self.ctx.check(self.ctx.lib.lzg_nni_best_sharded(self.ctx.h, self.comm, self.tree.h, len(moves), moves.ctypes.data_as(C.c_void_p),
                                                             int(greedy), C.byref(bi), C.byref(bs),
                                                             allsc.ctypes.data_as(C.c_void_p) if want_all else None))
            return bi.value, bs.value, allsc
        trace = os.environ.get("LZG_SHARD_TRACE")
        t0 = time.perf_counter()
        address, n = self.tree.nni_scores_launch(moves)
        t1 = time.perf_counter()
        if n:
            if self._vec is None or self._vec_key != (address, n):
                self._vec, self._vec_key = device_vector(address, n, self.ctx.device), (address, n)
            self._sum(self._vec)
            t2 = time.perf_counter()
            out = self.tree.nni_scores_finish(greedy, want_all)
            if trace:
                t3 = time.perf_counter()
                print("[shard %d] launch %.0f us, all-reduce call %.0f us, finish %.0f us" % (self.rank, 1e6 * (t1 - t0), 1e6 * (t2 - t1), 1e6 * (t3 - t2)), flush=True)
            return out
        return -1, self.root, (np.zeros(0, np.int32) if want_all else None)

    def hill_climb(self, edge_order, greedy=False, max_moves=2 ** 31 - 1, want_moves=True):
        """The whole climb with the ranks in lock step (lzg_hill_climb_sharded): per iteration one NCCL
        sum all-reduce of {root score, candidate scores} on the library stream, nothing else exchanged,
        no host decision -- every rank derives the same winner from the same reduced vector."""
        if self.comm is None:
            raise RuntimeError("the sharded climb needs the in-library NCCL communicator (LZG_SHARD_NCCL=1)")
        out = self.tree.hill_climb(edge_order, greedy, max_moves, want_moves, comm=self.comm)
        self.root = out["true_score"]
        return out

    def apply_nni(self, u, w, v, z, new_root=None):
        """Every rank applies the move to its columns.  The new global root score is the accepted
        candidate's score (pass it as new_root to skip the all-reduce of one int)."""
        part = self.tree.apply_nni(u, w, v, z)
        if new_root is None:
            import torch
            t = torch.tensor([part], dtype=torch.int32, device="cuda:%d" % self.ctx.device)
            self._stream.wait_stream(torch.cuda.current_stream(self.ctx.device))
            self._sum(t)
            self._stream.synchronize()
            new_root = int(t.item())
        self.root = new_root
        return new_root

    def last_nni_traffic(self):
        return self.tree.last_nni_traffic()
