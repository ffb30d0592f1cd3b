"""The host-side search layer of `lazac nni` for library users: the same C++ code the `lazac` binary runs
(host/search.cpp, host/seed_tree.cpp) behind liblazac_host.so.

Mirrors /root/reference/src/copy_number.cxx: stochastic_nni (:419-462), hill_climb (:375-417: the
std::shuffle of the edge indices with the caller's std::ranlux48_base, then the device-side climb,
lzg_hill_climb), and the Clearcut seed tree (src/clearcut.cxx: MT19937 taxon shuffle, float32 NJ).
The generator is libstdc++'s std::ranlux48_base, so a seed gives the reference's draw sequence.
"""
import ctypes as C
import os

import numpy as np

from ._capi import HERE, i32, ptr

LIB_PATH = os.path.join(HERE, "liblazac_host.so")

_vp, _i32 = C.c_void_p, C.c_int32
SYMBOLS = [
    ("lzh_last_error", C.c_char_p, []),
    ("lzh_gen_new", _vp, [_i32]),
    ("lzh_gen_free", None, [_vp]),
    ("lzh_gen_next", C.c_uint64, [_vp]),
    ("lzh_gen_uniform_int", _i32, [_vp, _i32, _i32]),
    ("lzh_gen_uniform_real", C.c_double, [_vp, C.c_double, C.c_double]),
    ("lzh_stochastic_nni", C.c_int, [_vp, _i32, _vp, _vp, C.c_float, _vp, _vp]),
    ("lzh_hill_climb", C.c_int, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    ("lzh_clearcut_shuffle", C.c_int, [_i32, C.c_uint32, _vp]),
    ("lzh_nj_seed_tree_cpu", C.c_int, [_vp, _i32, _vp, _vp, _vp]),
]

_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(LIB_PATH + " is not built: run `make -C host`")
        lib = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError(load().lzh_last_error().decode())


class Generator:
    """std::ranlux48_base(seed) -- the generator of do_nni (src/lazac.cxx:237)."""

    def __init__(self, seed):
        self.lib = load()
        self.h = self.lib.lzh_gen_new(int(seed))

    def next(self):
        return int(self.lib.lzh_gen_next(self.h))

    def uniform_int(self, lo, hi):
        return int(self.lib.lzh_gen_uniform_int(self.h, lo, hi))

    def uniform_real(self, lo, hi):
        return float(self.lib.lzh_gen_uniform_real(self.h, lo, hi))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.lzh_gen_free(self.h)
            self.h = None


def stochastic_nni(gen, child0, child1, aggression):
    """stochastic_nni(t, gen, aggression): floor(#internal edges x aggression) random NNIs.  Returns (child0, child1)."""
    c0, c1 = i32(child0), i32(child1)
    o0, o1 = np.empty_like(c0), np.empty_like(c1)
    _check(load().lzh_stochastic_nni(gen.h, len(c0), ptr(c0), ptr(c1), float(aggression), ptr(o0), ptr(o1)))
    return o0, o1


def hill_climb(ctx, tree, gen, greedy=False, score=None):
    """hill_climb(t, gen, greedy) on a scored api.Tree.  Returns dict(score = what the reference's returned tree
    reports as t[0].data.score, child0, child1, passes, candidates, moves)."""
    s = C.c_int32(tree.small_rectilinear() if score is None else int(score))
    c0, c1 = np.zeros(tree.n, np.int32), np.zeros(tree.n, np.int32)
    stats = np.zeros(3, np.int64)
    _check(load().lzh_hill_climb(ctx.h, tree.h, gen.h, tree.n, int(greedy), C.byref(s), ptr(c0), ptr(c1), ptr(stats)))
    return {"score": s.value, "child0": c0, "child1": c1, "passes": int(stats[0]), "candidates": int(stats[1]), "moves": int(stats[2])}


def clearcut_shuffle(n, seed):
    """Taxon order of NJ_shuffle_distance_matrix (src/clearcut.cxx:1644) for init_genrand(seed): position i holds taxon perm[i]."""
    perm = np.zeros(n, np.int32)
    _check(load().lzh_clearcut_shuffle(int(n), int(seed), ptr(perm)))
    return perm


def nj_seed_tree_cpu(D, perm):
    """Clearcut-exact neighbour joining on the CPU: the same (left, right) arrays as Context.nj_seed_tree."""
    D = i32(D)
    n = D.shape[0]
    perm = i32(perm)
    left, right = np.zeros(2 * n - 1, np.int32), np.zeros(2 * n - 1, np.int32)
    _check(load().lzh_nj_seed_tree_cpu(ptr(D), n, ptr(perm), ptr(left), ptr(right)))
    return left, right
