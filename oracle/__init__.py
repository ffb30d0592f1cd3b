"""oracle -- TEST INFRASTRUCTURE ONLY (checker, never the thing shipped or measured).

ctypes bindings for
  * ``libzcnt_oracle.so``  -- our plain-C restatement (oracle/zcnt_oracle.c), class ``Port``;
  * ``_ref/libref_shim.so`` -- the UNMODIFIED reference functions behind a C ABI
    (oracle/ref_shim.cpp + /root/reference/src/copy_number.cxx), class ``Ref``;
  * ``_ref/lazac_ref``     -- the reference CLI built from its own sources.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  Parity is pinned: see the header of zcnt_oracle.c.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libzcnt_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_shim.so")
REF_BIN = os.path.join(HERE, "_ref", "lazac_ref")

_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(quiet=True):
    """Run oracle/Makefile (compiles the port always, the reference when it is present)."""
    subprocess.run(["make", "-C", HERE, "-j4"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def have_ref():
    return os.path.exists(REF_SO)


def have_ref_bin():
    return os.path.exists(REF_BIN)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _opt(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Port:
    """The C restatement on flat arrays (oracle/zcnt_oracle.c)."""

    def __init__(self):
        if not os.path.exists(PORT_SO):
            build()
        self.lib = lib = C.CDLL(PORT_SO)
        lib.zo_bp_from_cn.argtypes = [C.c_int, _i32p, _i32p, _i32p]
        lib.zo_cn_from_bp.argtypes = [C.c_int, _i32p, _i32p, _i32p]
        lib.zo_distance_matrix.argtypes = [C.c_int, C.c_int, _i32p, _i32p]
        lib.zo_distance_rows.argtypes = [C.c_int, C.c_int, _i32p, C.c_int, C.c_int, _i32p]
        lib.zo_distance_rows_i8.argtypes = [C.c_int, C.c_int, np.ctypeslib.ndpointer(dtype=np.int8, flags="C_CONTIGUOUS"),
                                            C.c_int, _i32p, _i32p]
        lib.zo_merge.argtypes = [C.c_int] + [_i32p] * 6
        lib.zo_merge.restype = C.c_int32
        lib.zo_small_rectilinear.argtypes = [C.c_int, C.c_int] + [_i32p] * 6
        lib.zo_greedy_nni.argtypes = ([C.c_int, C.c_int] + [_i32p] * 6 + [C.c_int, _i32p, _i32p,
                                      C.c_int, _i32p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                      _i32p])
        lib.zo_apply_nni.argtypes = [C.c_int, C.c_int] + [_i32p] * 6 + [C.c_int] * 4
        lib.zo_ancestral_labeling.argtypes = [C.c_int, C.c_int] + [_i32p] * 8

    def bp_from_cn(self, seg, cn):
        seg, cn = _i32(seg), _i32(cn)
        bp = np.empty_like(cn)
        self.lib.zo_bp_from_cn(len(seg) - 1, seg, cn, bp)
        return bp

    def cn_from_bp(self, seg, bp):
        seg, bp = _i32(seg), _i32(bp)
        cn = np.empty_like(bp)
        self.lib.zo_cn_from_bp(len(seg) - 1, seg, bp, cn)
        return cn

    def distance_matrix(self, bp):
        bp = _i32(bp)
        n, L = bp.shape
        D = np.zeros((n, n), np.int32)
        self.lib.zo_distance_matrix(n, L, bp, D)
        return D

    def distance_rows(self, bp, r0, r1):
        bp = _i32(bp)
        n, L = bp.shape
        D = np.zeros((r1 - r0, n), np.int32)
        self.lib.zo_distance_rows(n, L, bp, r0, r1, D)
        return D

    def distance_rows_i8(self, bp, rows):
        """Rows `rows` of the matrix (against every cell) from an int8 matrix: int32 [len(rows)][n]."""
        bp = np.ascontiguousarray(bp, dtype=np.int8)
        rows = _i32(rows)
        n, L = bp.shape
        D = np.zeros((len(rows), n), np.int32)
        self.lib.zo_distance_rows_i8(n, L, bp, len(rows), rows, D)
        return D

    def merge(self, us, ue, vs, ve):
        us, ue, vs, ve = map(_i32, (us, ue, vs, ve))
        so, eo = np.empty_like(us), np.empty_like(us)
        cost = self.lib.zo_merge(len(us), us, ue, vs, ve, so, eo)
        return so, eo, int(cost)


class PortTree:
    """Flat-array tree state driven by the C restatement."""

    def __init__(self, port, child0, child1, leaf_bp):
        self.port = port
        c0, c1 = _i32(child0).copy(), _i32(child1).copy()
        lo, hi = np.minimum(c0, c1), np.maximum(c0, c1)
        self.child0, self.child1 = _i32(lo), _i32(hi)
        leaf_bp = _i32(leaf_bp)
        self.n, self.L = leaf_bp.shape
        self.start = leaf_bp.copy()
        self.end = leaf_bp.copy()
        internal = self.child0 >= 0
        self.start[internal] = 0
        self.end[internal] = 0
        self.score = np.zeros(self.n, np.int32)
        self.visited = np.zeros(self.n, np.int32)

    def clone(self):
        import copy
        return copy.deepcopy(self)

    def _state(self):
        return (self.child0, self.child1, self.start, self.end, self.score, self.visited)

    def small_rectilinear(self):
        rc = self.port.lib.zo_small_rectilinear(self.n, self.L, *self._state())
        if rc == 1:
            raise ValueError("every child must have exactly two children")
        return int(self.score[0])

    def unvisit_all(self):
        self.visited[:] = 0

    def edges(self):
        """digraph::edges() order (include/digraph.hpp:62-70): by parent id, then child id."""
        eu, ev = [], []
        for u in range(self.n):
            for c in (self.child0[u], self.child1[u]):
                if c >= 0:
                    eu.append(u)
                    ev.append(int(c))
        return _i32(eu), _i32(ev)

    def greedy_nni(self, eu, ev, order, greedy=False, want_all=True):
        eu, ev, order = _i32(eu), _i32(ev), _i32(order)
        cap = 2 * len(order) + 2
        moves = np.zeros((cap, 4), np.int32) if want_all else None
        scores = np.zeros(cap, np.int32) if want_all else None
        best = np.zeros(6, np.int32)
        k = self.port.lib.zo_greedy_nni(self.n, self.L, *self._state(), len(eu), eu, ev,
                                        len(order), order, int(greedy), _opt(moves), _opt(scores),
                                        cap, best)
        found = best[5] >= 0
        res = {"n": k, "found": bool(found),
               "move": tuple(int(x) for x in best[:4]) if found else None,
               "score": int(best[4]) if found else None, "index": int(best[5])}
        if want_all:
            res["moves"], res["scores"] = moves[:k], scores[:k]
        return res

    def apply_nni(self, u, w, v, z):
        self.port.lib.zo_apply_nni(self.n, self.L, *self._state(), u, w, v, z)
        return int(self.score[0])

    def ancestral_labeling(self):
        old = np.zeros(self.n, np.int32)
        par = np.zeros(self.n, np.int32)
        lab = np.zeros((self.n, self.L), np.int32)
        br = np.zeros(self.n, np.int32)
        k = self.port.lib.zo_ancestral_labeling(self.n, self.L, self.child0, self.child1,
                                                self.start, self.end, old, par, lab, br)
        assert k == self.n
        return old, par, lab, br
