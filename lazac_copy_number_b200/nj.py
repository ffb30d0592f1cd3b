"""Adapters around the NJ seed tree (lzg_nj_seed_tree on the GPU, host/seed_tree.cpp on the CPU -- both
Clearcut-exact, /root/reference/src/clearcut.cxx:925): array layouts and depths."""
import numpy as np


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
