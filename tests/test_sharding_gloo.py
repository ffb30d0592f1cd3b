"""CPU: the multi-rank host logic with world_size 2 over gloo -- candidate slices + 64-bit min
all-reduce reproduce the single-rank choice (strict '<', first wins), and the distance shards of
all ranks tile the triangle exactly once.  The per-rank scores come from the oracle here (no GPU)."""
import os
import socket

import numpy as np
import pytest

from lazac_copy_number_b200 import sharding, synth


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n,world", [(1, 1), (127, 2), (128, 2), (1000, 8), (20000, 8), (777, 3)])
def test_shards_tile_the_triangle(n, world):
    total = sum(sharding.shard_pairs(n, s, world) for s in range(world))
    assert total == n * (n + 1) // 2
    masks = np.array([sharding.shard_rows(n, s, world) for s in range(world)])
    assert (masks.sum(0) == 1).all()
    if n >= 128 * 2 * world and (n + 127) // 128 % (2 * world) == 0:
        per = [sharding.shard_pairs(n, s, world) for s in range(world)]
        assert max(per) - min(per) <= 128 * 128  # paired tile rows: equal areas up to the ragged last row


def test_key_order_is_the_reference_tie_break():
    # smaller score wins; equal scores: smaller enumeration index wins; nothing found loses
    assert sharding.pack_key(10, 500) < sharding.pack_key(11, 3)
    assert sharding.pack_key(10, 3) < sharding.pack_key(10, 500)
    assert sharding.pack_key(10, 3) < sharding.pack_key(0, -1) == sharding.NO_MOVE
    assert sharding.unpack_key(sharding.pack_key(154, 77)) == (77, 154)
    assert sharding.unpack_key(sharding.NO_MOVE) == (-1, None)
    assert sharding.pack_key(99, 5, greedy=True) < sharding.pack_key(1, 6, greedy=True)


def _worker(rank, world, port, greedy, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import Port, PortTree
    n_leaves = 24
    prof = synth.make_profiles(n_leaves, 120, 2, seed=4)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(9))
    n = 2 * n_leaves - 1
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    pt = PortTree(Port(), c0, c1, leaf_bp)
    base = pt.small_rectilinear()
    steps = []
    for _ in range(4):
        eu, ev = pt.edges()
        order = np.random.default_rng(1).permutation(len(eu)).astype(np.int32)
        full = pt.greedy_nni(eu, ev, order)               # every candidate's score, enumeration order
        scores, moves = full["scores"], full["moves"]
        lo, hi = sharding.candidate_slice(len(scores), rank, world)
        # what lzg_nni_best returns for the slice: first strictly-best below the current root score
        better = np.flatnonzero(scores[lo:hi] < base)
        if len(better) == 0:
            li, ls = -1, base
        elif greedy:
            li, ls = int(better[0]), int(scores[lo + better[0]])
        else:
            li = int(better[np.argmin(scores[lo:hi][better])])
            ls = int(scores[lo + li])
        gi, gs = sharding.allreduce_best(li, ls, lo, greedy)
        want = pt.greedy_nni(eu, ev, order, greedy=greedy, want_all=False)
        assert gi == want["index"], (gi, want)
        if gi < 0:
            break
        if not greedy:
            assert gs == want["score"]
        base = pt.apply_nni(*[int(x) for x in moves[gi]])  # every rank applies the same move
        steps.append(gi)
    if rank == 0:
        out.put(steps)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("greedy", [False, True])
def test_two_ranks_choose_the_reference_move(greedy):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, greedy, out)) for r in range(2)]
    for p in procs:
        p.start()
    steps = out.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert len(steps) >= 1


@pytest.mark.parametrize("n_pos,world", [(1, 1), (16, 2), (17, 2), (20000, 8), (10000, 3), (33, 8)])
def test_position_slices_tile_the_columns(n_pos, world):
    cuts = [sharding.position_slice(n_pos, r, world) for r in range(world)]
    assert cuts[0][0] == 0 and cuts[-1][1] == n_pos
    for r in range(world - 1):
        assert cuts[r][1] == cuts[r + 1][0] and cuts[r][1] % 16 == 0 or cuts[r][1] == n_pos


def _position_worker(rank, world, port, out):
    """Position sharding with the oracle in place of the GPU pass: every rank scores the whole
    neighbourhood over its column slice, the ranks sum-all-reduce {root, scores}, and the arg-min
    of the sum is the single-rank choice on the whole profiles."""
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import Port, PortTree
    n_leaves = 20
    prof = synth.make_profiles(n_leaves, 90, 2, seed=11)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(5))
    n = 2 * n_leaves - 1
    lo, hi = sharding.position_slice(bp.shape[1], rank, world)
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    whole = PortTree(Port(), c0, c1, leaf_bp)
    mine = PortTree(Port(), c0, c1, np.ascontiguousarray(leaf_bp[:, lo:hi]))
    whole.small_rectilinear()
    mine.small_rectilinear()
    steps = []
    for it in range(4):
        eu, ev = whole.edges()
        order = np.random.default_rng(it).permutation(len(eu)).astype(np.int32)
        part = mine.greedy_nni(eu, ev, order)
        vec = torch.tensor(np.concatenate([[mine.score[0]], part["scores"]]).astype(np.int32))
        dist.all_reduce(vec, op=dist.ReduceOp.SUM)
        vec = vec.numpy()
        want = whole.greedy_nni(eu, ev, order)
        assert vec[0] == whole.score[0]
        assert np.array_equal(vec[1:], want["scores"])
        better = np.flatnonzero(vec[1:] < vec[0])
        gi = int(better[np.argmin(vec[1:][better])]) if len(better) else -1
        assert gi == want["index"]
        if gi < 0:
            break
        m = [int(x) for x in want["moves"][gi]]
        whole.apply_nni(*m)
        mine.apply_nni(*m)
        steps.append(gi)
    if rank == 0:
        out.put(steps)
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_position_sharded_pass():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_position_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    steps = out.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert len(steps) >= 1
