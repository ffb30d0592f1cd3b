"""GPU parity: small parsimony, NNI neighbourhood scoring, move application and labeling through
the C ABI vs the oracle (bit-exact intervals, scores, chosen moves, labels)."""
import numpy as np
import pytest

from lazac_copy_number_b200 import synth
from lazac_copy_number_b200._capi import LazacGpuError
from lazac_copy_number_b200.api import Profiles, Tree
from oracle import PortTree

pytestmark = pytest.mark.gpu


def make_case(n_leaves, bins, alleles, seed, shuffle_tree=True):
    prof = synth.make_profiles(n_leaves, bins, alleles, seed=seed, want_tree=True)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    rng = np.random.default_rng(seed + 100)
    c0, c1 = (synth.random_tree(n_leaves, rng) if shuffle_tree else (prof["child0"], prof["child1"]))
    n = 2 * n_leaves - 1
    leaf_row = np.full(n, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    return bp, c0, c1, leaf_row, leaf_bp


def assert_same_state(gt, pt):
    S, E, sc = gt.intervals()
    assert np.array_equal(sc, pt.score)
    assert np.array_equal(S, pt.start)
    assert np.array_equal(E, pt.end)


@pytest.mark.parametrize("n_leaves,bins,alleles", [(2, 5, 1), (3, 40, 1), (10, 255, 1), (40, 300, 2), (150, 1000, 2)])
def test_full_score_and_intervals(ctx, port, n_leaves, bins, alleles):
    bp, c0, c1, leaf_row, leaf_bp = make_case(n_leaves, bins, alleles, seed=n_leaves)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    assert_same_state(gt, pt)
    gt.free()
    prof.free()


@pytest.mark.parametrize("n_leaves,bins,alleles,greedy", [(4, 30, 1, False), (10, 255, 1, False), (60, 500, 2, False),
                                                          (60, 500, 2, True), (200, 1200, 1, False)])
def test_neighbourhood_scores_and_choice(ctx, port, n_leaves, bins, alleles, greedy):
    bp, c0, c1, leaf_row, leaf_bp = make_case(n_leaves, bins, alleles, seed=3 * n_leaves)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    gt.small_rectilinear()
    pt.small_rectilinear()
    eu, ev = pt.edges()
    geu, gev = gt.edges()
    assert np.array_equal(eu, geu) and np.array_equal(ev, gev)
    order = np.random.default_rng(1).permutation(len(eu)).astype(np.int32)
    full = pt.greedy_nni(eu, ev, order, greedy=False)           # every candidate's score
    want = pt.greedy_nni(eu, ev, order, greedy=greedy, want_all=False)
    moves = gt.enumerate_moves(eu, ev, order)
    assert np.array_equal(moves, full["moves"])
    idx, score, allsc = gt.nni_best(moves, greedy=greedy, want_all=True)
    assert np.array_equal(allsc, full["scores"])
    assert idx == want["index"]
    if want["found"]:
        assert score == want["score"] and tuple(moves[idx]) == want["move"]
    merges, nbytes = gt.last_nni_work()
    assert merges >= 2 * len(moves) and nbytes == (merges + len(moves)) * 2 * gt.L
    cand_bytes, up_bytes = gt.last_nni_traffic()
    # the two candidates of an edge are one work item: 10 L bytes per pair, 6 L under the root
    root_moves = int((moves[:, 0] == 0).sum())
    assert len(moves) % 2 == 0 and root_moves % 2 == 0
    assert cand_bytes == (10 * len(moves) // 2 - 4 * root_moves // 2) * gt.L and up_bytes % (2 * gt.L) == 0
    gt.free()
    prof.free()


def test_hill_climb_trajectory(ctx, port):
    """Repeated greedy_nni + apply: identical move sequence, scores and intervals at every step."""
    bp, c0, c1, leaf_row, leaf_bp = make_case(50, 600, 2, seed=9)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    eu, ev = pt.edges()
    order = np.random.default_rng(4).permutation(len(eu)).astype(np.int32)
    steps = 0
    for _ in range(60):
        eu, ev = pt.edges()  # re-indexing every step is fine here: both sides use the same lists
        want = pt.greedy_nni(eu, ev, order, want_all=False)
        got = gt.greedy_nni(eu, ev, order)
        assert got == want["move"]
        if got is None:
            break
        assert gt.apply_nni(*got) == pt.apply_nni(*got)
        assert_same_state(gt, pt)
        steps += 1
    assert steps >= 3
    gt.free()
    prof.free()


def test_clone_set_topology_and_labeling(ctx, port):
    bp, c0, c1, leaf_row, leaf_bp = make_case(35, 400, 2, seed=21)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    gt.small_rectilinear()
    g2 = gt.clone()
    rng = np.random.default_rng(8)
    d0, d1 = synth.random_tree(35, rng)
    g2.set_topology(d0, d1)
    p2 = PortTree(port, d0, d1, leaf_bp)
    assert g2.small_rectilinear() == p2.small_rectilinear()
    assert_same_state(g2, p2)
    p1 = PortTree(port, c0, c1, leaf_bp)
    p1.small_rectilinear()
    assert_same_state(gt, p1)  # the clone's re-scoring did not touch the original
    for g, p in ((gt, p1), (g2, p2)):
        got, want = g.ancestral_labeling(), p.ancestral_labeling()
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
        assert int(want[3][1:].sum()) == int(p.score[0])  # score = sum of branch lengths
    gt.free()
    g2.free()
    prof.free()


def test_errors(ctx):
    bp = synth.random_bp(3, 20, seed=1)
    prof = Profiles(ctx, bp)
    c0 = np.array([1, 3, -1, -1, -1], np.int32)
    c1 = np.array([2, 4, -1, -1, -1], np.int32)
    rows = np.array([-1, -1, 0, 1, 2], np.int32)
    with pytest.raises(LazacGpuError) as e:  # unary node: the reference's logic_error
        Tree(ctx, prof, np.array([1, 3, -1, -1, -1]), np.array([2, -1, -1, -1, -1]), rows)
    assert e.value.status == 2
    t = Tree(ctx, prof, c0, c1, rows)
    with pytest.raises(LazacGpuError) as e:  # not scored yet
        t.nni_best([(0, 2, 1, 3)])
    assert e.value.status == 6
    t.small_rectilinear()
    with pytest.raises(LazacGpuError) as e:  # w is not a child of u
        t.nni_best([(0, 3, 1, 4)])
    assert e.value.status == 7
    assert t.nni_best(np.zeros((0, 4), np.int32))[0] == -1
    t.free()
    prof.free()


def test_arbitrary_move_lists(ctx, port):
    """lzg_nni_best takes any list of moves: pairs broken apart by a shuffle, duplicates, odd
    lengths, a single move -- scores and the first-wins choice still follow list order."""
    bp, c0, c1, leaf_row, leaf_bp = make_case(50, 700, 2, seed=77)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    gt.small_rectilinear()
    pt.small_rectilinear()
    eu, ev = pt.edges()
    order = np.arange(len(eu), dtype=np.int32)
    full = pt.greedy_nni(eu, ev, order)
    moves, scores = full["moves"], full["scores"]
    rng = np.random.default_rng(5)
    root = pt.score[0]
    for trial in range(6):
        if trial == 0:
            sel = rng.permutation(len(moves))                      # pairs broken apart
        elif trial == 1:
            sel = np.concatenate([np.arange(len(moves)), rng.integers(0, len(moves), 17)])  # duplicates
        elif trial == 2:
            sel = np.arange(1, len(moves))                         # odd length, shifted pairing
        elif trial == 3:
            sel = np.array([int(np.argmin(scores))])               # one move
        elif trial == 4:
            sel = np.repeat(np.arange(0, len(moves), 2), 2)        # (z, z) twins: same edge, same z
        else:
            sel = np.arange(len(moves))[::-1]                      # reversed pairs
        idx, score, allsc = gt.nni_best(moves[sel], want_all=True)
        want_sc = scores[sel]
        assert np.array_equal(allsc, want_sc), trial
        better = np.flatnonzero(want_sc < root)
        if len(better) == 0:
            assert idx == -1 and score == root
        else:
            best = better[np.argmin(want_sc[better])]              # argmin returns the first minimum
            assert idx == best and score == want_sc[best], trial
        gidx, _, _ = gt.nni_best(moves[sel], greedy=True)
        assert gidx == (better[0] if len(better) else -1)
    gt.free()
    prof.free()


def test_up_intervals_invariant_and_deep_trees(ctx, port):
    """Caterpillar (maximally deep) and random trees: every candidate's 3-merge score equals the
    reference's path re-scoring, before and after accepted moves (UP intervals rebuilt)."""
    n_leaves, bins = 64, 300
    prof = synth.make_profiles(n_leaves, bins, 2, seed=31)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    n = 2 * n_leaves - 1
    # caterpillar: internal i has children (i+1, leaf), last internal has two leaves
    c0 = np.full(n, -1, np.int32)
    c1 = np.full(n, -1, np.int32)
    for i in range(n_leaves - 1):
        c0[i] = i + 1 if i < n_leaves - 2 else n_leaves - 1 + i
        c1[i] = n_leaves - 1 + i + (1 if i == n_leaves - 2 else 0)
    leaf_row = np.full(n, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    gp = Profiles(ctx, bp)
    gt = Tree(ctx, gp, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    for step in range(6):
        eu, ev = pt.edges()
        order = np.random.default_rng(step).permutation(len(eu)).astype(np.int32)
        full = pt.greedy_nni(eu, ev, order)
        moves = gt.enumerate_moves(eu, ev, order)
        idx, score, allsc = gt.nni_best(moves, want_all=True)
        assert np.array_equal(allsc, full["scores"])
        assert idx == full["index"]
        if idx < 0:
            break
        assert gt.apply_nni(*moves[idx]) == pt.apply_nni(*full["move"])
    gt.free()
    gp.free()


@pytest.mark.parametrize("n_leaves,bins,alleles,parts,greedy", [(40, 300, 2, 2, False), (90, 700, 1, 3, False), (90, 700, 1, 2, True),
                                                                (12, 16, 1, 1, False)])
def test_position_sharded_pass(ctx, port, n_leaves, bins, alleles, parts, greedy):
    """lzg_nni_scores_launch / _finish: column slices of the profiles as separate trees (what the
    ranks of a multi-GPU run hold), their {root, scores} vectors summed on the context's stream (the
    NCCL all-reduce of sharding.PositionShardedTree), arg-min on the sum -- the oracle's choice on the
    whole profiles, along a hill climb."""
    import torch
    from lazac_copy_number_b200 import sharding
    bp, c0, c1, leaf_row, leaf_bp = make_case(n_leaves, bins, alleles, seed=7 * n_leaves)
    L = bp.shape[1]
    cuts = [sharding.position_slice(L, r, parts) for r in range(parts)]
    assert cuts[0][0] == 0 and cuts[-1][1] == L and all(cuts[r][1] == cuts[r + 1][0] for r in range(parts - 1))
    profs = [Profiles(ctx, np.ascontiguousarray(bp[:, lo:hi])) for lo, hi in cuts]
    trees = [Tree(ctx, p, c0, c1, leaf_row) for p in profs]
    pt = PortTree(port, c0, c1, leaf_bp)
    root = pt.small_rectilinear()
    assert sum(t.small_rectilinear() for t in trees) == root
    stream = torch.cuda.ExternalStream(ctx.stream(), device=ctx.device)
    for it in range(6):
        eu, ev = pt.edges()
        order = np.random.default_rng(it).permutation(len(eu)).astype(np.int32)
        full = pt.greedy_nni(eu, ev, order, greedy=False)
        want = pt.greedy_nni(eu, ev, order, greedy=greedy, want_all=False)
        moves = full["moves"]
        vecs = []
        for t in trees:
            address, n = t.nni_scores_launch(moves)
            assert n == len(moves) + 1
            vecs.append(sharding.device_vector(address, n, ctx.device))
        with torch.cuda.stream(stream):
            total = torch.stack(vecs).sum(0, dtype=torch.int32)
            for v in vecs:
                v.copy_(total)
        for t in trees:
            bi, bs, allsc = t.nni_scores_finish(greedy, want_all=True)
            assert np.array_equal(allsc, full["scores"])
            assert bi == want["index"]
            if bi >= 0:
                assert bs == full["scores"][bi]
            else:
                assert bs == root
            assert t.last_candidate_score() == full["scores"][-1]
        if want["index"] < 0:
            break
        m = [int(x) for x in moves[want["index"]]]
        root = pt.apply_nni(*m)
        assert sum(t.apply_nni(*m) for t in trees) == root
    with pytest.raises(LazacGpuError):
        trees[0].nni_scores_finish()          # nothing pending
    for t in trees:
        t.free()
    for p in profs:
        p.free()


def test_sharded_entry_point_with_one_rank(ctx, port):
    """lzg_comm_init(world = 1) + lzg_nni_best_sharded: the three-step call without an exchange is
    lzg_nni_best (the N > 1 exchange itself is exercised by bench.py --gpus N, same trajectory and scores)."""
    import ctypes as C
    bp, c0, c1, leaf_row, leaf_bp = make_case(70, 400, 2, seed=77)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    comm = C.c_void_p()
    ctx.check(ctx.lib.lzg_comm_init(ctx.h, 0, 1, None, C.byref(comm)))
    eu, ev = pt.edges()
    order = np.random.default_rng(5).permutation(len(eu)).astype(np.int32)
    full = pt.greedy_nni(eu, ev, order)
    moves = np.ascontiguousarray(full["moves"], np.int32)
    bi, bs = C.c_int32(0), C.c_int32(0)
    allsc = np.zeros(len(moves), np.int32)
    ctx.check(ctx.lib.lzg_nni_best_sharded(ctx.h, comm, gt.h, len(moves), moves.ctypes.data_as(C.c_void_p), 0, C.byref(bi), C.byref(bs),
                                           allsc.ctypes.data_as(C.c_void_p)))
    assert np.array_equal(allsc, full["scores"]) and bi.value == full["index"]
    with pytest.raises(LazacGpuError):
        ctx.check(ctx.lib.lzg_comm_init(ctx.h, 3, 2, None, C.byref(C.c_void_p())))   # rank out of range / no id
    ctx.check(ctx.lib.lzg_comm_free(comm))
    gt.free()
    prof.free()
