"""oracle.ref -- ctypes view of oracle/_ref/libref_shim.so (the unmodified reference functions)
and a runner for oracle/_ref/lazac_ref (the reference CLI).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

from . import REF_BIN, REF_SO, _i32, _i32p


class Ref:
    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " (run `make -C oracle` where /root/reference exists)")
        self.lib = lib = C.CDLL(REF_SO)
        vp, ci = C.c_void_p, C.c_int
        lib.ref_bp_from_cn.argtypes = [ci] + [_i32p] * 7
        lib.ref_cn_from_bp.argtypes = [ci] + [_i32p] * 7
        lib.ref_distance_matrix.argtypes = [ci, ci, _i32p, _i32p]
        lib.ref_tree_new.argtypes = [ci, ci, _i32p, _i32p, _i32p]
        lib.ref_tree_new.restype = vp
        _i8p = np.ctypeslib.ndpointer(dtype=np.int8, flags="C_CONTIGUOUS")
        lib.ref_tree_new_i8.argtypes = [ci, ci, _i32p, _i32p, _i32p, _i8p]
        lib.ref_tree_new_i8.restype = vp
        lib.ref_distance_rows_i8.argtypes = [ci, ci, _i8p, ci, _i32p, _i32p]
        lib.ref_get_scores.argtypes = [vp, _i32p]
        lib.ref_tree_clone.argtypes = [vp]
        lib.ref_tree_clone.restype = vp
        lib.ref_tree_free.argtypes = [vp]
        lib.ref_small_rectilinear.argtypes = [vp, C.POINTER(C.c_int32)]
        lib.ref_unvisit_all.argtypes = [vp]
        lib.ref_get_node.argtypes = [vp, ci, _i32p, _i32p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        lib.ref_get_topology.argtypes = [vp, _i32p, _i32p]
        lib.ref_edges.argtypes = [vp, _i32p, _i32p]
        lib.ref_apply_nni.argtypes = [vp, ci, ci, ci, ci]
        lib.ref_candidate_scores.argtypes = [vp, ci, _i32p, _i32p, ci, _i32p, _i32p, _i32p, ci]
        lib.ref_greedy_nni.argtypes = [vp, ci, _i32p, _i32p, ci, _i32p, ci, _i32p]
        lib.ref_gen_new.argtypes = [ci]
        lib.ref_gen_new.restype = vp
        lib.ref_gen_free.argtypes = [vp]
        lib.ref_gen_next.argtypes = [vp]
        lib.ref_gen_next.restype = C.c_uint64
        lib.ref_hill_climb.argtypes = [vp, vp, ci]
        lib.ref_stochastic_nni.argtypes = [vp, vp, C.c_float]
        lib.ref_stochastic_nni.restype = vp
        lib.ref_ancestral_labeling.argtypes = [vp, _i32p, _i32p, _i32p, _i32p]

    def bp_from_cn(self, chrom, allele, start, end, cn):
        chrom, allele, start, end, cn = map(_i32, (chrom, allele, start, end, cn))
        bp, perm = np.empty_like(cn), np.empty_like(cn)
        assert self.lib.ref_bp_from_cn(len(cn), chrom, allele, start, end, cn, bp, perm) == 0
        return bp, perm

    def cn_from_bp(self, chrom, allele, start, end, bp):
        chrom, allele, start, end, bp = map(_i32, (chrom, allele, start, end, bp))
        cn, perm = np.empty_like(bp), np.empty_like(bp)
        assert self.lib.ref_cn_from_bp(len(bp), chrom, allele, start, end, bp, cn, perm) == 0
        return cn, perm

    def distance_matrix(self, bp):
        bp = _i32(bp)
        n, L = bp.shape
        D = np.zeros((n, n), np.int32)
        self.lib.ref_distance_matrix(n, L, bp, D)
        return D

    def distance_rows_i8(self, bp, rows):
        """The reference's pair arithmetic (operator- + breakpoint_magnitude) for the given rows
        against every cell; bp is int8 [n][L].  Returns int32 [len(rows)][n]."""
        bp = np.ascontiguousarray(bp, dtype=np.int8)
        rows = _i32(rows)
        n, L = bp.shape
        D = np.zeros((len(rows), n), np.int32)
        self.lib.ref_distance_rows_i8(n, L, bp, len(rows), rows, D)
        return D

    def gen(self, seed):
        return RefGen(self, seed)

    def tree_i8(self, child0, child1, leaf_row, bp):
        """Tree whose leaves take their profiles from rows of a compact int8 matrix."""
        bp = np.ascontiguousarray(bp, dtype=np.int8)
        h = self.lib.ref_tree_new_i8(len(child0), bp.shape[1], _i32(child0), _i32(child1), _i32(leaf_row), bp)
        return RefTree(self, handle=h, n=len(child0), L=bp.shape[1])

    def tree(self, child0, child1, leaf_bp):
        return RefTree(self, child0, child1, leaf_bp)


class RefGen:
    def __init__(self, ref, seed):
        self.ref = ref
        self.h = ref.lib.ref_gen_new(int(seed))

    def next(self):
        return int(self.ref.lib.ref_gen_next(self.h))

    def __del__(self):
        if getattr(self, "h", None):
            self.ref.lib.ref_gen_free(self.h)
            self.h = None


class RefTree:
    def __init__(self, ref, child0=None, child1=None, leaf_bp=None, handle=None, n=None, L=None):
        self.ref = ref
        if handle is not None:
            self.h, self.n, self.L = handle, n, L
            return
        leaf_bp = _i32(leaf_bp)
        self.n, self.L = leaf_bp.shape
        self.h = ref.lib.ref_tree_new(self.n, self.L, _i32(child0), _i32(child1), leaf_bp)

    def __del__(self):
        if getattr(self, "h", None):
            self.ref.lib.ref_tree_free(self.h)
            self.h = None

    def clone(self):
        return RefTree(self.ref, handle=self.ref.lib.ref_tree_clone(self.h), n=self.n, L=self.L)

    def small_rectilinear(self):
        s = C.c_int32(0)
        rc = self.ref.lib.ref_small_rectilinear(self.h, C.byref(s))
        if rc == 1:
            raise ValueError("every child must have exactly two children")
        return s.value

    def unvisit_all(self):
        self.ref.lib.ref_unvisit_all(self.h)

    def node(self, u):
        s, e = np.zeros(self.L, np.int32), np.zeros(self.L, np.int32)
        sc, vis = C.c_int32(0), C.c_int32(0)
        rc = self.ref.lib.ref_get_node(self.h, u, s, e, C.byref(sc), C.byref(vis))
        return (s, e, sc.value, vis.value) if rc == 0 else (None, None, sc.value, vis.value)

    def state(self):
        """(start[n][L], end[n][L], score[n]) of every node."""
        S, E = np.zeros((self.n, self.L), np.int32), np.zeros((self.n, self.L), np.int32)
        sc = np.zeros(self.n, np.int32)
        for u in range(self.n):
            s, e, k, _ = self.node(u)
            if s is not None:
                S[u], E[u] = s, e
            sc[u] = k
        return S, E, sc

    def scores(self):
        sc = np.zeros(self.n, np.int32)
        self.ref.lib.ref_get_scores(self.h, sc)
        return sc

    def topology(self):
        c0, c1 = np.zeros(self.n, np.int32), np.zeros(self.n, np.int32)
        self.ref.lib.ref_get_topology(self.h, c0, c1)
        return c0, c1

    def edges(self):
        eu, ev = np.zeros(self.n, np.int32), np.zeros(self.n, np.int32)
        k = self.ref.lib.ref_edges(self.h, eu, ev)
        return eu[:k].copy(), ev[:k].copy()

    def apply_nni(self, u, w, v, z):
        self.ref.lib.ref_apply_nni(self.h, u, w, v, z)

    def candidate_scores(self, eu, ev, order):
        eu, ev, order = _i32(eu), _i32(ev), _i32(order)
        cap = 2 * len(order) + 2
        moves, scores = np.zeros((cap, 4), np.int32), np.zeros(cap, np.int32)
        k = self.ref.lib.ref_candidate_scores(self.h, len(eu), eu, ev, len(order), order,
                                              moves, scores, cap)
        assert k >= 0
        return moves[:k], scores[:k]

    def greedy_nni(self, eu, ev, order, greedy=False):
        eu, ev, order = _i32(eu), _i32(ev), _i32(order)
        mv = np.zeros(4, np.int32)
        found = self.ref.lib.ref_greedy_nni(self.h, len(eu), eu, ev, len(order), order,
                                            int(greedy), mv)
        return tuple(int(x) for x in mv) if found else None

    def hill_climb(self, gen, greedy=False):
        self.ref.lib.ref_hill_climb(self.h, gen.h, int(greedy))

    def stochastic_nni(self, gen, aggression):
        h = self.ref.lib.ref_stochastic_nni(self.h, gen.h, float(aggression))
        return RefTree(self.ref, handle=h, n=self.n, L=self.L)

    def ancestral_labeling(self):
        old, par = np.zeros(self.n, np.int32), np.zeros(self.n, np.int32)
        lab, br = np.zeros((self.n, self.L), np.int32), np.zeros(self.n, np.int32)
        k = self.ref.lib.ref_ancestral_labeling(self.h, old, par, lab, br)
        assert k == self.n
        return old, par, lab, br


def run_ref_cli(args, cwd=None, timeout=600):
    """Run the reference CLI.  `nni` must always get an explicit -a (stale-errno bug,
    /root/reference/src/dmat.cxx:573-577 -- SURVEY.md F14)."""
    return subprocess.run([REF_BIN] + list(args), cwd=cwd, timeout=timeout, check=True,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
