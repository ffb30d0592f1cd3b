"""CPU: liblazac_host.so (host/capi.cpp) -- the host-side search layer for library users -- loads, exports what
lazac_copy_number_b200/search.py binds, and its RNG-driven parts reproduce the reference's: the
std::ranlux48_base sequence, stochastic_nni (src/copy_number.cxx:419-462) draw for draw, the Clearcut taxon
shuffle and NJ (src/clearcut.cxx) against the reference CLI's own seed trees."""
import os

import numpy as np
import pytest

from lazac_copy_number_b200 import nj, search, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def test_library_loads_and_exports_every_bound_symbol():
    lib = search.load()
    for name, _, _ in search.SYMBOLS:
        assert hasattr(lib, name), name


def test_generator_is_ranlux48_base(ref):
    for seed in (0, 1, 7, 12345):
        g, r = search.Generator(seed), ref.gen(seed)
        assert [g.next() for _ in range(50)] == [r.next() for _ in range(50)]


@pytest.mark.parametrize("n_leaves,seed,aggression", [(3, 1, 1.0), (10, 2, 0.5), (40, 3, 0.7), (40, 4, 1.75), (200, 5, 0.25), (200, 6, 0.0)])
def test_stochastic_nni_equals_reference(ref, n_leaves, seed, aggression):
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(seed))
    n = 2 * n_leaves - 1
    rt = ref.tree(c0, c1, np.zeros((n, 3), np.int32))
    g, rg = search.Generator(seed), ref.gen(seed)
    o0, o1 = search.stochastic_nni(g, c0, c1, aggression)
    r0, r1 = rt.stochastic_nni(rg, aggression).topology()
    assert np.array_equal(o0, r0) and np.array_equal(o1, r1)
    assert g.next() == rg.next()          # the same number of draws was consumed


def newick_of(left, right, names_at_position):
    def rec(u):
        return names_at_position[u] if left[u] < 0 else "(" + rec(left[u]) + "," + rec(right[u]) + ")"
    import sys
    sys.setrecursionlimit(10000)
    return rec(len(left) - 1) + ";\n"


@pytest.mark.parametrize("case,prefix,seed", [("ptx008_cli", "PTX008", 0), ("synth_cli", "full", 3), ("synth_cli", "greedy", 5), ("sim_cli", "sim", 1)])
def test_cpu_seed_tree_equals_reference_cli(case, prefix, seed):
    """The reference CLI's PREFIX_dist_matrix.txt -> its PREFIX_nj_tree.newick, through the library's shuffle + NJ."""
    d = os.path.join(GOLD, case)
    with open(os.path.join(d, prefix + "_dist_matrix.txt")) as f:
        n = int(f.readline())
        rows = [f.readline().split() for _ in range(n)]
    names = [r[0] for r in rows]
    D = np.array([[int(x) for x in r[1:]] for r in rows], np.int32)
    perm = search.clearcut_shuffle(n, seed)
    assert sorted(perm.tolist()) == list(range(n))
    left, right = search.nj_seed_tree_cpu(D, perm)
    got = newick_of(left, right, [names[p] for p in perm])
    with open(os.path.join(d, prefix + "_nj_tree.newick")) as f:
        assert got == f.read()
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)      # the layout lzg_tree_create takes
    assert c0[0] >= 0 and (leaf_row[n - 1:] == perm).all()
