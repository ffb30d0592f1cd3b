"""Host-side mirror of the reference's hot-path interface over the C ABI.

Names follow the reference (namespace copynumber, /root/reference/include/copy_number.hpp and
build_distance_matrix in /root/reference/src/lazac.cxx:63): build_distance_matrix,
small_rectilinear, greedy_nni, nni (apply), ancestral_labeling, convert_to_breakpoint_profile.
All arithmetic happens on the GPU through liblazac_b200.so; this module only marshals arrays.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import LazacGpuError, i32, ptr


class Context:
    """One GPU context (one per process and device)."""

    def __init__(self, device=0):
        self.lib = _capi.load()
        h = C.c_void_p()
        rc = self.lib.lzg_init(int(device), C.byref(h))
        if rc != 0:
            raise LazacGpuError(rc, self.lib.lzg_last_error(None).decode())
        self.h = h
        self.device = device

    def check(self, rc):
        if rc != 0:
            raise LazacGpuError(rc, self.lib.lzg_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.lzg_shutdown(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- measurement helpers ------------------------------------------------------------------
    def launch_count(self):
        n = C.c_int64(0)
        self.check(self.lib.lzg_launch_count(self.h, C.byref(n)))
        return n.value

    def timer_start(self):
        self.check(self.lib.lzg_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        self.check(self.lib.lzg_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def sync(self):
        self.check(self.lib.lzg_sync(self.h))

    def stream(self):
        """cudaStream_t (as an int) every call on this context is ordered on."""
        out = C.c_void_p()
        self.check(self.lib.lzg_stream(self.h, C.byref(out)))
        return out.value or 0

    def flush_l2(self):
        self.check(self.lib.lzg_flush_l2(self.h))

    def pinned(self, shape, dtype):
        """numpy array backed by pinned host memory (freed with the context's process)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = C.c_void_p()
        self.check(self.lib.lzg_host_alloc(self.h, max(n, 1), C.byref(p)))
        buf = (C.c_uint8 * max(n, 1)).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def nj_seed_tree(self, D, perm):
        """Clearcut-exact neighbour joining of the integer matrix D on the GPU (the seed tree of
        do_nni, src/lazac.cxx:155-200); perm = the taxon shuffle (position i holds taxon perm[i]).
        Returns (left, right): int32 [2N-1], -1 for the N leaves, root = 2N-2."""
        D = i32(D)
        n = D.shape[0]
        perm = i32(perm)
        left, right = np.zeros(2 * n - 1, np.int32), np.zeros(2 * n - 1, np.int32)
        self.check(self.lib.lzg_nj_seed_tree(self.h, ptr(D), n, ptr(perm), ptr(left), ptr(right)))
        return left, right

    def nj_seed_tree_device(self, d_ptr, n, perm):
        """nj_seed_tree on a matrix already in HBM (raw device pointer to n*n int32, not modified)."""
        perm = i32(perm)
        left, right = np.zeros(2 * n - 1, np.int32), np.zeros(2 * n - 1, np.int32)
        self.check(self.lib.lzg_nj_seed_tree_device(self.h, C.c_void_p(d_ptr), n, ptr(perm), ptr(left), ptr(right)))
        return left, right

    # -- reference-named entry points -------------------------------------------------------------
    def upload_profiles(self, bp):
        return Profiles(self, bp)

    def build_distance_matrix(self, bp_profiles):
        """build_distance_matrix(std::map<std::string, breakpoint_profile>), src/lazac.cxx:63-95:
        names sorted byte-lexicographically, full symmetric int matrix."""
        names = sorted(bp_profiles, key=lambda s: s.encode())
        L = len(next(iter(bp_profiles.values()))) if names else 0
        bp = np.zeros((len(names), L), np.int32)
        for k, nm in enumerate(names):
            bp[k] = bp_profiles[nm]
        prof = Profiles(self, bp)
        try:
            return names, prof.distance_matrix()
        finally:
            prof.free()


class MultiContext:
    """Several GPUs of one box driven from this process (lzg_multi_*): build_distance_matrix with the
    whole matrix gathered into one host buffer."""

    def __init__(self, devices):
        self.lib = _capi.load()
        dev = i32(list(devices))
        h = C.c_void_p()
        rc = self.lib.lzg_multi_init(ptr(dev), len(dev), C.byref(h))
        if rc != 0:
            raise LazacGpuError(rc, self.lib.lzg_last_error(None).decode())
        self.h, self.devices = h, [int(d) for d in dev]

    def close(self):
        if getattr(self, "h", None):
            self.lib.lzg_multi_shutdown(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def distance_matrix(self, bp, out=None):
        """bp: int8 or int32 [n][L] (pinned buffers make the copies asynchronous).  Returns (D, pairs)."""
        bp = np.asarray(bp)
        n, L = bp.shape
        D = np.empty((n, n), np.int32) if out is None else out
        pairs = C.c_int64(0)
        if bp.dtype == np.int8:
            bp = np.ascontiguousarray(bp)
            rc = self.lib.lzg_distance_matrix_multi_i8(self.h, ptr(bp), n, L, ptr(D), C.byref(pairs))
        else:
            bp = i32(bp)
            rc = self.lib.lzg_distance_matrix_multi(self.h, ptr(bp), n, L, ptr(D), C.byref(pairs))
        if rc != 0:
            raise LazacGpuError(rc, self.lib.lzg_multi_last_error(self.h).decode())
        return D, pairs.value


class Profiles:
    """Breakpoint vectors of N cells resident in HBM (int8 + 128 bias)."""

    def __init__(self, ctx, bp):
        self.ctx = ctx
        bp = np.asarray(bp)
        if bp.ndim != 2:
            raise ValueError("bp must be [n_cells][n_positions]")
        h = C.c_void_p()
        if bp.dtype == np.int8:
            bp = np.ascontiguousarray(bp)
            rc = ctx.lib.lzg_profiles_upload_i8(ctx.h, ptr(bp), bp.shape[0], bp.shape[1], C.byref(h))
        else:
            bp = i32(bp)
            rc = ctx.lib.lzg_profiles_upload(ctx.h, ptr(bp), bp.shape[0], bp.shape[1], C.byref(h))
        ctx.check(rc)
        self.h = h
        self.n, self.L = bp.shape

    def free(self):
        if getattr(self, "h", None):
            self.ctx.lib.lzg_profiles_free(self.h)
            self.h = None

    def distance_matrix(self, out=None):
        D = np.empty((self.n, self.n), np.int32) if out is None else out
        self.ctx.check(self.ctx.lib.lzg_distance_matrix(self.ctx.h, self.h, ptr(D)))
        return D

    def distance_shard(self, shard, n_shards, out):
        pairs = C.c_int64(0)
        self.ctx.check(self.ctx.lib.lzg_distance_shard(self.ctx.h, self.h, shard, n_shards, ptr(out),
                                                       C.byref(pairs)))
        return pairs.value

    def distance_device_async(self, d_ptr, shard=0, n_shards=1):
        """Result stays in HBM at raw device pointer d_ptr (N*N int32)."""
        pairs = C.c_int64(0)
        self.ctx.check(self.ctx.lib.lzg_distance_shard_device_async(
            self.ctx.h, self.h, shard, n_shards, C.c_void_p(d_ptr), C.byref(pairs)))
        return pairs.value

    def nj_seed_tree(self, perm):
        """lzg_nj_seed_tree_profiles: NJ on these profiles' distance matrix, which stays in HBM."""
        perm = i32(perm)
        left, right = np.zeros(2 * self.n - 1, np.int32), np.zeros(2 * self.n - 1, np.int32)
        self.ctx.check(self.ctx.lib.lzg_nj_seed_tree_profiles(self.ctx.h, self.h, ptr(perm), ptr(left), ptr(right)))
        return left, right

    def distance_path(self):
        """{'tensor_core': bool, 'K': summed value range, 'k_pad': code width, 'launches': k_tc_distance launches so far}."""
        info = (C.c_int64 * 8)()
        self.ctx.check(self.ctx.lib.lzg_profiles_distance_path(self.h, info))
        out = {"tensor_core": bool(info[0]), "K": int(info[1]), "k_pad": int(info[2]), "launches": int(info[3])}
        if info[4] or info[5]:
            out.update(prepare_us=int(info[4]), kernel_us=int(info[5]), tiles=int(info[6]))
        return out

    def distance_noalloc(self, d_ptr):
        """lzg_distance_matrix_noalloc: the whole matrix to raw device pointer d_ptr (N*N int32) without any allocation --
        the kernel's work plan goes into the library's static device array (1 MiB; larger plans raise LazacGpuError)."""
        self.ctx.check(self.ctx.lib.lzg_distance_matrix_noalloc(self.ctx.h, self.h, C.c_void_p(d_ptr)))


class Tree:
    """digraph<rectilinear_vertex_data> resident on the GPU (include/copy_number.hpp:98-105)."""

    def __init__(self, ctx, profiles=None, child0=None, child1=None, leaf_row=None, _handle=None,
                 _n=None, _L=None):
        self.ctx = ctx
        if _handle is not None:
            self.h, self.n, self.L = _handle, _n, _L
            return
        child0, child1, leaf_row = i32(child0), i32(child1), i32(leaf_row)
        self.n, self.L = len(child0), profiles.L
        h = C.c_void_p()
        ctx.check(ctx.lib.lzg_tree_create(ctx.h, profiles.h, self.n, ptr(child0), ptr(child1),
                                          ptr(leaf_row), C.byref(h)))
        self.h = h

    def free(self):
        if getattr(self, "h", None):
            self.ctx.lib.lzg_tree_free(self.h)
            self.h = None

    def clone(self):
        h = C.c_void_p()
        self.ctx.check(self.ctx.lib.lzg_tree_clone(self.ctx.h, self.h, C.byref(h)))
        return Tree(self.ctx, _handle=h, _n=self.n, _L=self.L)

    def set_topology(self, child0, child1):
        child0, child1 = i32(child0), i32(child1)
        self.ctx.check(self.ctx.lib.lzg_tree_set_topology(self.ctx.h, self.h, ptr(child0), ptr(child1)))

    def topology(self):
        c0, c1 = np.zeros(self.n, np.int32), np.zeros(self.n, np.int32)
        self.ctx.check(self.ctx.lib.lzg_tree_get_topology(self.h, ptr(c0), ptr(c1)))
        return c0, c1

    def edges(self):
        """digraph::edges() order (include/digraph.hpp:62-70)."""
        c0, c1 = self.topology()
        eu, ev = [], []
        for u in range(self.n):
            if c0[u] >= 0:
                eu += [u, u]
                ev += [int(c0[u]), int(c1[u])]
        return i32(eu), i32(ev)

    def small_rectilinear(self):
        """small_rectilinear(t, 0), src/copy_number.cxx:227-277 -> t[0].data.score."""
        s = C.c_int32(0)
        self.ctx.check(self.ctx.lib.lzg_tree_score(self.ctx.h, self.h, C.byref(s)))
        return s.value

    def enumerate_moves(self, eu, ev, order):
        """Candidate enumeration of greedy_nni, src/copy_number.cxx:331-347."""
        c0, c1 = self.topology()
        mv = []
        for idx in order:
            u, v = int(eu[idx]), int(ev[idx])
            if c0[v] < 0:
                continue
            w = int(c1[u]) if c0[u] == v else int(c0[u])
            mv.append((u, w, v, int(c0[v])))
            mv.append((u, w, v, int(c1[v])))
        return i32(mv).reshape(-1, 4)

    def nni_best(self, moves, greedy=False, want_all=False):
        moves = i32(moves).reshape(-1, 4)
        bi, bs = C.c_int32(0), C.c_int32(0)
        allsc = np.zeros(len(moves), np.int32) if want_all else None
        self.ctx.check(self.ctx.lib.lzg_nni_best(self.ctx.h, self.h, len(moves), ptr(moves), int(greedy),
                                                 C.byref(bi), C.byref(bs), ptr(allsc)))
        return bi.value, bs.value, allsc

    def nni_scores_launch(self, moves):
        """First half of nni_best for position-sharded trees (lzg_nni_scores_launch): queues the
        pass and returns (device address, length) of {root score, candidate scores}; the caller
        sum-reduces that vector over the ranks on ctx.stream(), then calls nni_scores_finish."""
        moves = i32(moves).reshape(-1, 4)
        d, n = C.c_void_p(), C.c_int32(0)
        self.ctx.check(self.ctx.lib.lzg_nni_scores_launch(self.ctx.h, self.h, len(moves), ptr(moves), C.byref(d), C.byref(n)))
        self._pending_moves = len(moves)
        return d.value or 0, n.value

    def nni_scores_finish(self, greedy=False, want_all=False):
        bi, bs = C.c_int32(0), C.c_int32(0)
        allsc = np.zeros(self._pending_moves, np.int32) if want_all else None
        self.ctx.check(self.ctx.lib.lzg_nni_scores_finish(self.ctx.h, self.h, int(greedy), C.byref(bi), C.byref(bs), ptr(allsc)))
        return bi.value, bs.value, allsc

    def greedy_nni(self, eu, ev, order, greedy=False):
        """greedy_nni(t, indexed_edges, edge_indices, greedy), src/copy_number.cxx:322-373:
        returns the chosen (u, w, v, z) or None."""
        moves = self.enumerate_moves(eu, ev, order)
        idx, _, _ = self.nni_best(moves, greedy)
        return None if idx < 0 else tuple(int(x) for x in moves[idx])

    def apply_nni(self, u, w, v, z):
        """nni + unvisit + small_rectilinear, src/copy_number.cxx:397,408-409."""
        s = C.c_int32(0)
        self.ctx.check(self.ctx.lib.lzg_tree_apply_nni(self.ctx.h, self.h, u, w, v, z, C.byref(s)))
        return s.value

    def hill_climb(self, edge_order, greedy=False, max_moves=2 ** 31 - 1, want_moves=True, comm=None):
        """hill_climb(t, gen, greedy), src/copy_number.cxx:375-417, with the shuffle done by the caller:
        edge_order = random_indices (a permutation of the indices of edges() of the CURRENT topology).
        The whole climb runs on the device (lzg_hill_climb).  Returns a dict: moves [k][5] =
        (u, w, v, z, score after), n_moves, n_passes, reason, true_score, returned_score, candidates."""
        order = i32(edge_order)
        res = _capi.ClimbResult()
        cap = min(int(max_moves), 1 << 20) if want_moves else 0
        moves = np.zeros((cap, 5), np.int32) if cap else None
        if comm is None:
            rc = self.ctx.lib.lzg_hill_climb(self.ctx.h, self.h, len(order), ptr(order), int(greedy), int(max_moves), cap, ptr(moves), C.byref(res))
        else:
            rc = self.ctx.lib.lzg_hill_climb_sharded(self.ctx.h, comm, self.h, len(order), ptr(order), int(greedy), int(max_moves), cap, ptr(moves),
                                                     C.byref(res))
        self.ctx.check(rc)
        out = {k: getattr(res, k) for k, _ in _capi.ClimbResult._fields_}
        out["moves"] = moves[:min(cap, res.n_moves)] if cap else None
        return out

    def climb_stamps(self):
        """Device time stamps (ns) at the end of every pass of the last hill_climb on this tree."""
        n = C.c_int32(0)
        self.ctx.check(self.ctx.lib.lzg_tree_last_climb_stamps(self.h, 0, None, C.byref(n)))
        st = np.zeros(n.value, np.uint64)
        if n.value:
            self.ctx.check(self.ctx.lib.lzg_tree_last_climb_stamps(self.h, n.value, ptr(st), C.byref(n)))
        return st

    def climb_pass0_ms(self):
        """Device ms of the five kernels of the last climb's first (full) pass, or None when it ran as a graph."""
        ms = np.zeros(5, np.float32)
        rc = self.ctx.lib.lzg_tree_last_climb_pass0_ms(self.h, ptr(ms))
        return None if rc != 0 else [float(x) for x in ms]

    def intervals(self):
        S = np.zeros((self.n, self.L), np.int32)
        E = np.zeros((self.n, self.L), np.int32)
        sc = np.zeros(self.n, np.int32)
        self.ctx.check(self.ctx.lib.lzg_tree_get_intervals(self.ctx.h, self.h, ptr(S), ptr(E), ptr(sc)))
        return S, E, sc

    def node_intervals(self, ids):
        """(start, end) int32 [len(ids)][L] of the listed nodes; scores() has the cumulative scores."""
        ids = i32(ids)
        S = np.zeros((len(ids), self.L), np.int32)
        E = np.zeros((len(ids), self.L), np.int32)
        self.ctx.check(self.ctx.lib.lzg_tree_get_node_intervals(self.ctx.h, self.h, len(ids), ptr(ids), ptr(S), ptr(E)))
        return S, E

    def scores(self):
        sc = np.zeros(self.n, np.int32)
        self.ctx.check(self.ctx.lib.lzg_tree_get_intervals(self.ctx.h, self.h, None, None, ptr(sc)))
        return sc

    def ancestral_labeling(self):
        """ancestral_labeling(t, 0, bins), src/copy_number.cxx:187-225, including the pre-order
        renumbering (children pushed ascending, so the larger id is numbered first).
        Returns (old_id, parent_new, labels, branch) indexed by the new ids."""
        lab = np.zeros((self.n, self.L), np.int32)
        br = np.zeros(self.n, np.int32)
        self.ctx.check(self.ctx.lib.lzg_tree_ancestral(self.ctx.h, self.h, ptr(lab), ptr(br)))
        c0, c1 = self.topology()
        old, par, stack = [], [], [(0, -1)]
        while stack:
            u, p = stack.pop()
            k = len(old)
            old.append(u)
            par.append(p)
            if c0[u] >= 0:
                stack.append((int(c0[u]), k))
                stack.append((int(c1[u]), k))
        old = i32(old)
        return old, i32(par), lab[old], br[old]

    def last_nni_work(self):
        m, b = C.c_int64(0), C.c_int64(0)
        self.ctx.check(self.ctx.lib.lzg_tree_last_nni_work(self.h, C.byref(m), C.byref(b)))
        return m.value, b.value

    def last_nni_traffic(self):
        """(candidate bytes, UP-pass bytes) of the last nni_best as executed on the GPU."""
        c, u = C.c_int64(0), C.c_int64(0)
        self.ctx.check(self.ctx.lib.lzg_tree_last_nni_traffic(self.h, C.byref(c), C.byref(u)))
        return c.value, u.value

    def last_candidate_score(self):
        """t[0].data.score as the reference's greedy_nni leaves it after a pass that finds nothing."""
        s = C.c_int32(0)
        self.ctx.check(self.ctx.lib.lzg_tree_last_candidate_score(self.h, C.byref(s)))
        return s.value
