"""GPU parity: hill_climb (src/copy_number.cxx:375-417) run entirely on the device (lzg_hill_climb: one
CUDA graph whose WHILE node repeats accept -> UP pass -> all candidates -> arg-min) against the
oracle driven step by step with the reference's bookkeeping: one fixed order of edge indices, an
edge index stays with its child end while moves are accepted (:401-406), strict '<' first-wins
choice, greedy = first improving candidate, and t[0].score left at the last candidate's score by
the final, fruitless pass.  Compared: every accepted move and its score, the number of passes, the
returned score, and the final intervals and cumulative scores of every node."""
import numpy as np
import pytest

from lazac_copy_number_b200 import nj, synth
from lazac_copy_number_b200.api import Profiles, Tree
from oracle import PortTree

pytestmark = pytest.mark.gpu


def oracle_climb(pt, order, greedy=False, max_moves=1 << 30):
    eu, ev = pt.edges()
    eu, ev = eu.copy(), ev.copy()
    current = int(pt.score[0])
    moves, passes, reason, returned = [], 0, 0, current
    while True:
        r = pt.greedy_nni(eu, ev, order, greedy)
        passes += 1
        if not r["found"]:
            reason = 1
            returned = int(r["scores"][-1]) if r["n"] else current
            break
        u, w, v, z = r["move"]
        new = pt.apply_nni(u, w, v, z)
        i1 = int(np.flatnonzero((eu == u) & (ev == w))[0])     # (u, w) -> (v, w), (v, z) -> (u, z): same indices
        i2 = int(np.flatnonzero((eu == v) & (ev == z))[0])
        eu[i1], eu[i2] = v, u
        moves.append((u, w, v, z, new))
        if current <= new:
            reason, returned = 2, new
            break
        current = returned = new
        if len(moves) >= max_moves:
            reason = 3
            break
    pt.small_rectilinear()   # the probed paths are left un-scored by the candidate loop (:365-366)
    return {"moves": np.array(moves, np.int32).reshape(-1, 5), "n_passes": passes, "reason": reason, "returned_score": returned,
            "true_score": int(pt.score[0])}


def check_climb(ctx, port, bp, c0, c1, leaf_row, order_seed=0, greedy=False, max_moves=1 << 30, rounds=1):
    n = len(c0)
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[leaf_row >= 0] = bp[leaf_row[leaf_row >= 0]]
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    total = 0
    for r in range(rounds):
        order = np.random.default_rng(order_seed + r).permutation(n - 1).astype(np.int32)
        want = oracle_climb(pt, order, greedy, max_moves)
        got = gt.hill_climb(order, greedy, max_moves)
        assert np.array_equal(got["moves"], want["moves"]), (r, got["moves"][:4], want["moves"][:4])
        for k in ("n_passes", "reason", "returned_score", "true_score"):
            assert got[k] == want[k], (r, k, got[k], want[k])
        assert got["n_moves"] == len(want["moves"])
        S, E, sc = gt.intervals()
        assert np.array_equal(sc, pt.score) and np.array_equal(S, pt.start) and np.array_equal(E, pt.end)
        t0, t1 = gt.topology()
        assert np.array_equal(t0, pt.child0) and np.array_equal(t1, pt.child1)
        st = gt.climb_stamps()
        assert len(st) <= got["n_passes"] and (len(st) < 2 or np.all(np.diff(st.astype(np.int64)) >= 0))
        total += got["n_moves"]
    # the stepwise calls still work on the tree the climb left behind
    eu, ev = gt.edges()
    order = np.arange(len(eu), dtype=np.int32)
    moves = gt.enumerate_moves(eu, ev, order)
    if len(moves):
        idx, score, allsc = gt.nni_best(moves, want_all=True)
        full = pt.greedy_nni(*pt.edges(), order)
        assert np.array_equal(allsc, full["scores"]) and idx == full["index"]
    gt.free()
    prof.free()
    return total


def random_case(n_leaves, bins, alleles, seed):
    prof = synth.make_profiles(n_leaves, bins, alleles, seed=seed)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(seed + 100))
    leaf_row = np.full(2 * n_leaves - 1, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    return bp, c0, c1, leaf_row


@pytest.mark.parametrize("n_leaves,bins,alleles,greedy", [(2, 20, 1, False), (3, 40, 1, False), (4, 30, 1, False), (10, 255, 1, False),
                                                          (50, 600, 2, False), (50, 600, 2, True), (200, 1200, 1, False)])
def test_climb_matches_oracle(ctx, port, n_leaves, bins, alleles, greedy):
    bp, c0, c1, leaf_row = random_case(n_leaves, bins, alleles, seed=9 + n_leaves)
    moves = check_climb(ctx, port, bp, c0, c1, leaf_row, greedy=greedy)
    assert n_leaves < 50 or moves >= 3


def test_climb_in_queued_batches(ctx, port, monkeypatch):
    """LZG_CLIMB_GRAPH=0: the same kernels queued in batches by the host (what the sharded climb uses)."""
    monkeypatch.setenv("LZG_CLIMB_GRAPH", "0")
    bp, c0, c1, leaf_row = random_case(80, 700, 2, seed=4)
    assert check_climb(ctx, port, bp, c0, c1, leaf_row) >= 3


def test_climb_with_a_short_move_log(ctx, port, monkeypatch):
    """The device logs accepted moves; a full log stops the loop, the host drains it and the climb goes on."""
    monkeypatch.setenv("LZG_CLIMB_LOGCAP", "2")
    bp, c0, c1, leaf_row = random_case(80, 700, 2, seed=5)
    assert check_climb(ctx, port, bp, c0, c1, leaf_row) >= 5


def test_climb_with_a_move_budget_and_repeated_climbs(ctx, port):
    """max_moves cuts a climb short (reason 3); the next climb re-indexes the edges of the tree as it is then."""
    bp, c0, c1, leaf_row = random_case(60, 500, 2, seed=6)
    check_climb(ctx, port, bp, c0, c1, leaf_row, max_moves=2, rounds=4)


def test_climb_on_deep_trees(ctx, port):
    """A perturbed caterpillar (one long heavy path, many light heads) and an NJ seed tree."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    from sanitize_case import caterpillar
    n_leaves = 150
    bp = synth.random_bp(n_leaves, 500, seed=3, density=0.2)
    c0, c1, leaf_row = caterpillar(n_leaves)
    c0, c1 = synth.perturb_tree(c0, c1, 60, np.random.default_rng(4))
    assert check_climb(ctx, port, bp, c0, c1, leaf_row) >= 1
    prof = synth.make_profiles(120, 800, 2, seed=12)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    gp = Profiles(ctx, bp)
    perm = np.random.default_rng(0).permutation(120).astype(np.int32)
    left, right = ctx.nj_seed_tree(gp.distance_matrix(), perm)
    gp.free()
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
    c0, c1 = synth.perturb_tree(c0, c1, 40, np.random.default_rng(5))
    assert check_climb(ctx, port, bp, c0, c1, leaf_row, rounds=2) >= 1


def test_climb_rejects_a_bad_order(ctx):
    from lazac_copy_number_b200._capi import LazacGpuError
    bp, c0, c1, leaf_row = random_case(10, 100, 1, seed=1)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    with pytest.raises(LazacGpuError):                       # not scored yet
        gt.hill_climb(np.arange(18, dtype=np.int32))
    gt.small_rectilinear()
    with pytest.raises(LazacGpuError):
        gt.hill_climb(np.zeros(18, np.int32))                # not a permutation
    with pytest.raises(LazacGpuError):
        gt.hill_climb(np.arange(5, dtype=np.int32))          # wrong length
    gt.free()
    prof.free()


def test_library_hill_climb_equals_reference_hill_climb(ctx, ref):
    """search.hill_climb (host/search.cpp behind liblazac_host.so: std::shuffle with std::ranlux48_base, then
    lzg_hill_climb) against the reference's own hill_climb with the same seed: same final topology, same
    returned t[0].score (the stale last-candidate value included), same generator state afterwards -- on
    the leaves of the paper simulation n100_l1000_s0 (golden: hill_climb(ranlux48_base(7))) and on seeded cases."""
    import os
    from lazac_copy_number_b200 import search
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sim_n100_l1000_s0.npz"))
    bp, c0, c1 = gold["bp"].astype(np.int32), gold["child0"], gold["child1"]
    n_leaves = bp.shape[0]
    leaf_row = np.full(2 * n_leaves - 1, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    assert gt.small_rectilinear() == int(gold["score_before"])
    g = search.Generator(int(gold["hc_seed"]))
    out = search.hill_climb(ctx, gt, g, greedy=False)
    assert np.array_equal(out["child0"], gold["hc_child0"]) and np.array_equal(out["child1"], gold["hc_child1"])
    assert g.next() == int(gold["gen_next_after"])
    t0, t1 = gt.topology()
    assert np.array_equal(t0, gold["hc_child0"]) and np.array_equal(t1, gold["hc_child1"])
    gt.set_topology(t0, t1)
    assert gt.small_rectilinear() == int(gold["hc_score"])
    gt.free()
    prof.free()
    for n_leaves, seed, greedy in ((30, 11, False), (30, 12, True), (90, 13, False)):
        bp, c0, c1, leaf_row = random_case(n_leaves, 400, 2, seed)
        leaf_bp = np.zeros((2 * n_leaves - 1, bp.shape[1]), np.int32)
        leaf_bp[n_leaves - 1:] = bp
        prof = Profiles(ctx, bp)
        gt = Tree(ctx, prof, c0, c1, leaf_row)
        rt = ref.tree(c0, c1, leaf_bp)
        assert gt.small_rectilinear() == rt.small_rectilinear()
        g, rg = search.Generator(seed), ref.gen(seed)
        out = search.hill_climb(ctx, gt, g, greedy=greedy)
        rt.hill_climb(rg, greedy)
        r0, r1 = rt.topology()
        assert np.array_equal(out["child0"], r0) and np.array_equal(out["child1"], r1)
        assert out["score"] == rt.node(0)[2]                      # t[0].data.score as the reference's hill_climb leaves it
        assert g.next() == rg.next()
        gt.free()
        prof.free()
