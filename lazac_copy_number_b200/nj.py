"""Host-side neighbour joining for seed trees of synthetic benchmarks (numpy, float64).

This is NOT the byte-exact Clearcut restatement (that is lzg_nj_seed_tree on the GPU and
host/seed_tree.cpp on the CPU; tree_from_seed below adapts the GPU's output); neighbor_joining
only provides an NJ-shaped seed topology (deep, unbalanced, like the reference's
/root/reference/src/clearcut.cxx:925 NJ_neighbor_joining produces) for bench.py and tests, where
the GPU path and the oracle are both driven from the same topology arrays.
"""
import numpy as np


def neighbor_joining(D):
    """Classic NJ on a symmetric matrix D [n][n].  Returns (child0, child1) over 2n-1 nodes with
    root = 0 and leaf i at node id (n-1)+i, like synth.random_tree.  The final three-way join is
    resolved into a binary root (the reference's output tree is binary at the root as well)."""
    n = D.shape[0]
    if n == 1:
        return np.array([-1], np.int32), np.array([-1], np.int32)
    n_nodes = 2 * n - 1
    child0 = np.full(n_nodes, -1, np.int32)
    child1 = np.full(n_nodes, -1, np.int32)
    d = D.astype(np.float64).copy()
    ids = np.arange(n - 1, n_nodes)          # node id of every active row
    active = np.ones(n, bool)
    R = d.sum(1)
    nxt = n - 2                              # internal ids are handed out counting down to 1
    big = np.float64(np.inf)
    for r in range(n, 2, -1):
        Q = (r - 2) * d - R[:, None] - R[None, :]
        Q[~active, :] = big
        Q[:, ~active] = big
        np.fill_diagonal(Q, big)
        k = int(np.argmin(Q))
        i, j = divmod(k, n)
        if i > j:
            i, j = j, i
        a, b = int(ids[i]), int(ids[j])
        child0[nxt], child1[nxt] = min(a, b), max(a, b)
        dij = d[i, j]
        new = 0.5 * (d[i] + d[j] - dij)
        R -= d[i] + d[j]
        d[i, :] = new
        d[:, i] = new
        d[i, i] = 0.0
        active[j] = False
        d[j, :] = 0.0
        d[:, j] = 0.0
        R += new
        R[i] = new[active].sum()
        R[j] = 0.0
        ids[i] = nxt
        nxt -= 1
    rest = [int(x) for x in ids[active]]
    assert len(rest) == 2 and nxt == 0
    child0[0], child1[0] = min(rest), max(rest)
    return child0, child1


def tree_from_seed(left, right, perm):
    """(left, right) of lzg_nj_seed_tree (Context.nj_seed_tree: leaves 0 .. n-1 = shuffled positions,
    internal nodes n .. 2n-2 in creation order, root last) -> (child0, child1, leaf_row) in the
    layout lzg_tree_create takes: root = node 0, child0 < child1, leaf at node id (n-1) + position
    with leaf_row = the taxon (profile row) at that position."""
    total = len(left)
    n = (total + 1) // 2
    new_id = np.empty(total, np.int64)
    new_id[:n] = n - 1 + np.arange(n)
    new_id[n:] = total - 1 - np.arange(n, total)         # the root 2n-2 becomes 0
    child0 = np.full(total, -1, np.int32)
    child1 = np.full(total, -1, np.int32)
    a, b = new_id[left[n:]], new_id[right[n:]]
    child0[new_id[n:]] = np.minimum(a, b)
    child1[new_id[n:]] = np.maximum(a, b)
    leaf_row = np.full(total, -1, np.int32)
    leaf_row[n - 1:] = np.asarray(perm, np.int32)
    return child0, child1, leaf_row


def depths(child0, child1):
    """Depth of every node below the root (root = 0)."""
    n = len(child0)
    dep = np.zeros(n, np.int32)
    stack = [0]
    while stack:
        u = stack.pop()
        for c in (child0[u], child1[u]):
            if c >= 0:
                dep[c] = dep[u] + 1
                stack.append(int(c))
    return dep
