#!/usr/bin/env python
"""bench.py -- ZCNT hot-path throughput on B200 (contract: one JSON line on stdout from rank 0).

    python bench.py --gpus 1 --steps K --warmup W            # 1 GPU
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...                     # the reference's own CPU code

Workloads (BASELINE.json `configs`); a step = one whole distance matrix:
  cfg3  20,000 cells x 10,000 bins x 2 alleles (L = 20,000)  the headline at EVERY --gpus N: the
        matrix BASELINE.json's scaling target is quoted on; it fits one GPU (0.4 GB in, 1.6 GB
        out), so N = 1 runs the same matrix and the 1/2/4/8 series is one strong-scaling series
        (tile rows of the ONE matrix dealt to the ranks, no data-path collective; every rank
        ends with its own row/column strips)
  cfg2  1,000 cells x 5,000 bins x 2 alleles (L = 10,000)   configs[1]; reported in the same
        line at --gpus 1 as "cfg2" (same calls, same keys), or alone with --workload cfg2
The NNI path is reported in the same line at --gpus 1: "nni" = cfg4 (1,000 cells x 5,000 bins, NJ
seed tree), "nni_large" = cfg5's shape (10,000 cells x 20,000 positions) on the one GPU; a step is
one hill-climb iteration (accept a move + score the whole new neighbourhood).

value   = cell pairs (n(n+1)/2, as the reference's loop counts them, src/lazac.cxx:81-91) per
          second with the packed profiles already resident in HBM; device time (CUDA events on
          the library's stream), summed over the K steps, max over ranks; L2 flushed between steps.
e2e     = the same through the host-buffer C-ABI calls a caller of build_distance_matrix would
          make (lzg_profiles_upload_i8 + lzg_distance_matrix / lzg_distance_shard), pinned host
          buffers, H2D and D2H inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    "cfg2": dict(cells=1000, bins=5000, alleles=2),
    "cfg3": dict(cells=20000, bins=10000, alleles=2),
    "tiny": dict(cells=200, bins=500, alleles=2),       # CPU-side plumbing tests only
}
NNI_WORKLOADS = {
    "cfg4": dict(cells=1000, bins=5000, alleles=1),
    "cfg5": dict(cells=10000, bins=10000, alleles=2),
    "tiny": dict(cells=60, bins=300, alleles=1),
}
# VABSDIFF4.U8.ACC issue rate measured on this pool's B200 (tools/microbench.cu,
# profiles/r01_microbench_int_pipes.jsonl): 62.94 lane-instr/clk/SM x 148 SMs x 1.965 GHz.
SAD4_PEAK_LANE_INSTR_PER_S = 1.8304e13


# DRAM traffic per launch / per pass from the committed ncu --set full captures (bytes); None elsewhere
NCU_DRAM_BYTES = {
    ("distance", "cfg2"): 14432768,                                   # k_zcnt_distance: 14.43 MB read, 0 written (14.0 MB algorithmic)
    ("distance", "cfg3"): 7726542000 + 1607675000,                    # 7.7 GB read + 1.6 GB written (2.0 GB algorithmic); 54.75 + 1.60 GB with the row-major tile list
    ("nni", "cfg5 on 1 GPU"): 513864000 + 364247296 + 1283499000 + 4908800,  # k_up_pass (r + w) + k_nni_items (r + w); 2.93 GB algorithmic
}


NCU_DIST_FILES = {"cfg2": "r01_ncu_dist_cfg2.txt", "cfg3": "r01_ncu_dist_cfg3_v3.txt"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local, world


def make_workload(name):
    from lazac_copy_number_b200 import synth
    w = WORKLOADS[name]
    prof = synth.make_profiles(w["cells"], w["bins"], w["alleles"], seed=0)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"]).astype(np.int8)  # |bp| <= 8: int8 exact
    return bp


def bp_int8_chunked(cn, seg, rows=2000):
    from lazac_copy_number_b200 import synth
    out = np.empty(cn.shape, np.int8)
    for r0 in range(0, cn.shape[0], rows):
        out[r0:r0 + rows] = synth.bp_from_cn(cn[r0:r0 + rows], seg).astype(np.int8)
    return out


def cpu_distance_sample(bp, n_sample):
    """The reference's pair loop on the first n_sample cells, single-threaded (the reference has
    no threads).  Uses oracle/_ref (the unmodified reference functions) when it was built, else
    the C port.  Returns (pairs/s, kind, seconds, pairs)."""
    from oracle import Port, have_ref
    sub = np.ascontiguousarray(bp[:n_sample].astype(np.int32))
    if have_ref():
        from oracle.ref import Ref
        fn, kind = Ref().distance_matrix, "reference"
    else:
        fn, kind = Port().distance_matrix, "port"
    t0 = time.perf_counter()
    fn(sub)
    dt = time.perf_counter() - t0
    pairs = n_sample * (n_sample + 1) // 2
    return pairs / dt, kind, dt, pairs


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host
    cores (1 thread: the reference is single-threaded), each step a bounded sample."""
    rank, local, world = dist_env()
    if rank != 0:
        return
    name = args.workload or "cfg3"
    w = WORKLOADS[name]
    from lazac_copy_number_b200 import synth
    # the sample only needs the first cells: generate a smaller tree of the same shape per cell
    n_sample = min(w["cells"], args.ref_cells or (160 if w["bins"] * w["alleles"] <= 10000 else 110))
    prof = synth.make_profiles(n_sample, w["bins"], w["alleles"], seed=0)
    bp = bp_int8_chunked(prof["cn"], prof["seg"])
    for _ in range(args.warmup):
        cpu_distance_sample(bp, max(8, n_sample // 4))
    tot_pairs, tot_t, kind = 0, 0.0, "port"
    for _ in range(args.steps):
        v, kind, dt, pairs = cpu_distance_sample(bp, n_sample)
        tot_pairs += pairs
        tot_t += dt
    value = tot_pairs / tot_t
    sample = "%d cells x %d positions of %s (%d pairs per step)" % (n_sample, bp.shape[1], name, n_sample * (n_sample + 1) // 2)
    line = {
        "impl": "reference", "metric": "zcnt_cell_pairs_per_sec", "value": value, "unit": "cell-pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int32", "data": "synthetic",
        "config": {"workload": name, "cells": w["cells"], "bins": w["bins"], "alleles": w["alleles"],
                   "positions": bp.shape[1], "sample": sample},
        "cpu_baseline": {"value": value, "unit": "cell-pairs/s", "cores": 1, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "cell-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def nni_iterations(ctx, t, moves, order, steps, warmup, apply=None, align=None):
    """Times hill-climb iterations on tree t: accept a move (lzg_tree_apply_nni), then score the
    whole new neighbourhood (lzg_nni_best).  The tree alternates between T0 and T1 = T0 + one
    move (the inverse of nni(u, w, v, z) is nni(u, z, v, w)), so every step sees stale UP
    intervals exactly as in hill_climb (src/copy_number.cxx:393-409).  Device time, CUDA events
    on the library stream, L2 flushed before every step.  `apply` replaces t.apply_nni and `align`
    (a rank barrier) runs before every timed region when t is one rank's part of a sharded tree."""
    apply = apply or t.apply_nni
    align = align or ctx.sync      # the L2 flush has drained before the clock starts: host preparation is not hidden behind it
    m = tuple(int(x) for x in moves[len(moves) // 2])
    inv = (m[0], m[3], m[2], m[1])
    t.nni_best(moves)
    apply(*m)
    eu1, ev1 = t.edges()
    moves1 = t.enumerate_moves(eu1, ev1, order)
    apply(*inv)
    lists = (moves1, moves)
    for k in range(2 * max(2, (warmup + 1) // 2)):
        apply(*(m if k % 2 == 0 else inv))
        t.nni_best(lists[k % 2])
    n_it = steps + steps % 2                    # an even number of steps: the tree ends as T0
    # (1) the pass on its own, as lzg_nni_best runs it on stale UP intervals: UP pass + items + argmin
    os.environ["LZG_EAGER_UP"] = "0"
    stale_ms = 0.0
    cand = 0
    cb = ub = 0
    for k in range(n_it):
        apply(*(m if k % 2 == 0 else inv))   # untimed
        ctx.flush_l2()
        align()
        ctx.timer_start()
        t.nni_best(lists[k % 2])
        stale_ms += ctx.timer_stop()
        cand += len(lists[k % 2])
        c, u = t.last_nni_traffic()
        cb, ub = cb + c, ub + u
    del os.environ["LZG_EAGER_UP"]
    # (2) whole iterations: the move application queues the UP pass behind itself; ONE timer over both
    # calls (a stop between them would wait for the queued UP pass and expose the host's packing)
    it_ms = wall = 0.0
    for k in range(n_it):
        ctx.flush_l2()
        ctx.sync()
        align()
        t0 = time.perf_counter()
        ctx.timer_start()
        apply(*(m if k % 2 == 0 else inv))
        t.nni_best(lists[k % 2])
        it_ms += ctx.timer_stop()
        wall += time.perf_counter() - t0
    hot = 0.0
    for _ in range(n_it):                        # UP intervals still valid: scoring only
        ctx.flush_l2()
        align()
        ctx.timer_start()
        idx, score, _ = t.nni_best(moves)
        hot += ctx.timer_stop()
    return {"iterations": n_it, "candidates": cand, "stale_pass_ms": stale_ms / n_it, "iteration_ms": it_ms / n_it, "wall_s": wall,
            "scoring_only_ms": hot / n_it, "cand_bytes": cb // n_it, "up_bytes": ub // n_it, "best_index": int(idx), "best_score": int(score),
            "moves_bytes": int(moves.nbytes)}


def nni_trajectory(ctx, t, order, max_steps, apply=None, align=None):
    """A real hill climb from the tree's current topology (hill_climb, src/copy_number.cxx:375-417:
    one fixed shuffle of the edge indices, every pass scores the whole neighbourhood, the best
    strictly improving move is accepted, until none is left or max_steps).  Every iteration
    (accept the previous winner + pass) is timed on the device; the host's re-enumeration of the
    move list in Python between iterations is not part of the library and is left out."""
    apply = apply or t.apply_nni
    align = align or ctx.sync
    eu, ev = t.edges()
    moves = t.enumerate_moves(eu, ev, order)
    idx, score, _ = t.nni_best(moves)           # first pass: untimed (UP intervals of the starting tree)
    first = score
    ms, cand, accepted = [], 0, 0
    while idx >= 0 and accepted < max_steps:
        m = tuple(int(x) for x in moves[idx])
        # the move list of the tree AFTER the move has to exist before the clock starts
        # (hill_climb patches four entries of its edge index in place, src/copy_number.cxx:401-406):
        # apply on the host's copy of the topology by applying for real, enumerating, undoing
        apply(*m)
        eu, ev = t.edges()
        nxt = t.enumerate_moves(eu, ev, order)
        apply(m[0], m[3], m[2], m[1])
        t.nni_best(moves)                       # back in the state right after the last pass
        ctx.flush_l2()
        align()
        ctx.timer_start()
        apply(*m)
        idx, score, _ = t.nni_best(nxt)
        ms.append(ctx.timer_stop())
        cand += len(nxt)
        moves = nxt
        accepted += 1
    if not ms:
        return {"iterations": 0}
    ms_sorted = sorted(ms)
    return {"iterations": accepted, "median_ms": ms_sorted[len(ms) // 2], "mean_ms": sum(ms) / len(ms), "max_ms": ms_sorted[-1],
            "value": cand / (sum(ms) * 1e-3), "unit": "candidates/s", "first_accepted_score": int(first), "last_score": int(score),
            "ended": "no improving move left" if idx < 0 else "step limit",
            "note": "real hill climb from the seed tree: every iteration = accept the previous winner (lzg_tree_apply_nni) + score the whole new "
                    "neighbourhood (lzg_nni_best), device time, L2 flushed before each"}


def nni_report(name, r, t, moves, dep, root, seed_tree, peaks, peaks_kind, n_leaves, L, note):
    n = 2 * n_leaves - 1
    C_ = len(moves)
    it_ms = r["iteration_ms"]
    ps_ms = r["stale_pass_ms"]
    nbytes = r["cand_bytes"] + r["up_bytes"]
    merges, ref_bytes = t.last_nni_work()
    return {
        "metric": "nni_candidates_scored_per_sec", "value": C_ / (ps_ms * 1e-3), "unit": "candidates/s",
        "ms_per_step": ps_ms,
        "e2e": {"value": r["candidates"] / r["wall_s"], "unit": "candidates/s", "h2d_bytes_per_step": r["moves_bytes"] + 16, "d2h_bytes_per_step": 72,
                "call": "whole hill-climb iterations, lzg_tree_apply_nni + lzg_nni_best with a host move list, wall clock"},
        "iteration": {"value": C_ / (it_ms * 1e-3), "unit": "candidates/s", "ms_per_step": it_ms,
                      "note": "accept a move (path re-scoring; the UP pass is queued behind it) + score the new neighbourhood, device time, one timer over both calls"},
        "scoring_only": {"value": C_ / (r["scoring_only_ms"] * 1e-3), "unit": "candidates/s", "ms_per_step": r["scoring_only_ms"],
                         "note": "UP intervals already valid (no move accepted since the last pass)"},
        "config": {"workload": name, "cells": n_leaves, "positions": L, "nodes": n, "candidates": C_, "seed_tree": seed_tree,
                   "step": "one neighbourhood pass as lzg_nni_best runs it after an accepted move: host packing of the move list, UP pass over "
                           "the internal nodes, scoring of every candidate (the two candidates of an edge = one work item), argmin, read-back",
                   "mean_depth_v": float(dep[moves[:, 2]].mean()), "max_depth_v": int(dep[moves[:, 2]].max()),
                   "reference_algorithm_merges_per_pass": int(merges), "reference_algorithm_bytes_per_pass": int(ref_bytes),
                   "root_score": int(root), "best_index": r["best_index"], "best_score": r["best_score"], "l2": "flushed between steps"},
        "roofline": {"bound": "hbm", "achieved": nbytes / (ps_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": nbytes / (ps_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "traffic": NCU_DRAM_BYTES.get(("nni", name)), "peak_kind": peaks_kind,
                     "traffic_source": "k_up_pass + k_nni_items, dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full on a random tree of this "
                                       "shape (profiles/r01_ncu_nni_v5_cfg5.txt)" if ("nni", name) in NCU_DRAM_BYTES else None,
                     "algorithmic_bytes_per_pass": int(nbytes), "candidate_bytes": int(r["cand_bytes"]), "up_pass_bytes": int(r["up_bytes"]),
                     "scoring_only_frac": r["cand_bytes"] / (r["scoring_only_ms"] * 1e-3) / 1e9 / peaks["hbm_gbs"], "note": note},
    }


def nj_seed(ctx, gp, n_leaves):
    """The seed tree of do_nni (src/lazac.cxx:155-200) for the profiles gp: GPU distance matrix, then
    Clearcut-exact NJ on the GPU (the MT19937 taxon shuffle replaced by a numpy one, same on every rank)."""
    from lazac_copy_number_b200 import nj
    t0 = time.perf_counter()
    D = gp.distance_matrix()
    t1 = time.perf_counter()
    perm = np.random.default_rng(0).permutation(n_leaves).astype(np.int32)
    left, right = ctx.nj_seed_tree(D, perm)
    t2 = time.perf_counter()
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
    return c0, c1, leaf_row, "lzg_nj_seed_tree (Clearcut-exact NJ on the GPU, %.0f ms) of the GPU distance matrix (%.0f ms to the host)" % (
        1e3 * (t2 - t1), 1e3 * (t1 - t0))


def nni_section(ctx, args, peaks, peaks_kind):
    """cfg4: hill-climb iterations on an NJ seed tree (1 GPU)."""
    from lazac_copy_number_b200 import nj, synth
    from lazac_copy_number_b200.api import Profiles, Tree
    name = "tiny" if args.workload == "tiny" else "cfg4"
    w = NNI_WORKLOADS[name]
    prof = synth.make_profiles(w["cells"], w["bins"], w["alleles"], seed=0)
    bp = bp_int8_chunked(prof["cn"], prof["seg"])
    n_leaves, L = bp.shape
    gp = Profiles(ctx, bp)
    D = gp.distance_matrix()
    # the seed tree of do_nni (src/lazac.cxx:155-200): Clearcut-exact NJ on the GPU, seed-0 shuffle replaced by a numpy one
    perm = np.random.default_rng(0).permutation(n_leaves).astype(np.int32)
    t0 = time.perf_counter()
    left, right = ctx.nj_seed_tree(D, perm)
    nj_s = time.perf_counter() - t0
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
    n = 2 * n_leaves - 1
    t = Tree(ctx, gp, c0, c1, leaf_row)
    root = t.small_rectilinear()
    eu, ev = t.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = t.enumerate_moves(eu, ev, order)
    dep = nj.depths(c0, c1)
    r = nni_iterations(ctx, t, moves, order, args.steps, args.warmup)
    out = nni_report(name, r, t, moves, dep, root, "lzg_nj_seed_tree (Clearcut-exact NJ on the GPU, %.1f ms) of the GPU distance matrix" % (1e3 * nj_s),
                     peaks, peaks_kind, n_leaves, L,
                     "tree state (%.1f MB incl. UP intervals) fits the 126 MB L2 at this size: the step is launch/latency bound, not HBM bound"
                     % (4 * n * L / 1e6))
    C_ = len(moves)
    if not args.no_cpu:
        # the reference's candidate loop (src/copy_number.cxx:347-366) on the same tree, 1 thread
        from oracle import Port, PortTree, have_ref
        leaf_bp = np.zeros((n, L), np.int32)
        leaf_bp[n_leaves - 1:] = bp[perm]
        if have_ref():
            from oracle.ref import Ref
            rt = Ref().tree(c0, c1, leaf_bp)
            rt.small_rectilinear()
            t0 = time.perf_counter()
            mv, sc = rt.candidate_scores(eu, ev, order)
            dt = time.perf_counter() - t0
            kind = "reference"
        else:
            pt = PortTree(Port(), c0, c1, leaf_bp)
            pt.small_rectilinear()
            t0 = time.perf_counter()
            rr = pt.greedy_nni(eu, ev, order)
            dt = time.perf_counter() - t0
            sc, kind = rr["scores"], "port"
        _, _, allsc = t.nni_best(moves, want_all=True)
        out["cpu_baseline"] = {"value": C_ / dt, "unit": "candidates/s", "cores": 1, "kind": kind, "sample": "the same full pass (%d candidates)" % C_,
                               "scores_equal_gpu": bool(np.array_equal(sc, allsc))}
    out["trajectory"] = nni_trajectory(ctx, t, order, max(20, args.steps))
    t.free()
    gp.free()
    return out


def nni_section_large(ctx, args, peaks, peaks_kind):
    """cfg5's shape on ONE GPU (tree state 0.8 GB + 0.8 GB of UP intervals: beyond L2, the regime the
    HBM roofline describes)."""
    from lazac_copy_number_b200 import nj, synth
    from lazac_copy_number_b200.api import Profiles, Tree
    w = NNI_WORKLOADS["cfg5"]
    prof = synth.make_profiles(w["cells"], w["bins"], w["alleles"], seed=0, want_tree=True)
    bp = bp_int8_chunked(prof["cn"], prof["seg"])
    n_leaves, L = bp.shape
    n = 2 * n_leaves - 1
    gp = Profiles(ctx, bp)
    c0, c1, leaf_row, seed_note = nj_seed(ctx, gp, n_leaves)
    t = Tree(ctx, gp, c0, c1, leaf_row)
    root = t.small_rectilinear()
    eu, ev = t.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = t.enumerate_moves(eu, ev, order)
    dep = nj.depths(c0, c1)
    r = nni_iterations(ctx, t, moves, order, min(args.steps, 10), args.warmup)
    out = nni_report("cfg5 on 1 GPU", r, t, moves, dep, root, seed_note,
                     peaks, peaks_kind, n_leaves, L, "tree state %.1f GB incl. UP intervals" % (4 * n * L / 1e9))
    out["trajectory"] = nni_trajectory(ctx, t, order, 20)
    t.free()
    gp.free()
    return out


def nni_section_sharded(ctx, args, rank, world, peaks, peaks_kind):
    """cfg5 (N > 1), positions sharded: every rank holds the same topology over its column slice of
    the profiles, runs the whole pass (UP pass + all candidates) on 1/N of the bytes, and the ranks
    sum-all-reduce the int32 vector {root score, candidate scores} over NCCL before the arg-min.
    Same steps as the one-GPU sections (nni_iterations): the pass after an accepted move, whole
    iterations, scoring only."""
    import torch
    import torch.distributed as dist
    from lazac_copy_number_b200 import nj, sharding, synth
    from lazac_copy_number_b200.api import Profiles
    name = "tiny" if args.workload == "tiny" else "cfg5"
    w = NNI_WORKLOADS[name]
    prof = synth.make_profiles(w["cells"], w["bins"], w["alleles"], seed=0, want_tree=True)
    bp = bp_int8_chunked(prof["cn"], prof["seg"])
    n_leaves, L = bp.shape
    n = 2 * n_leaves - 1
    gp = Profiles(ctx, bp)                      # whole rows once, for the seed tree (the same on every rank)
    c0, c1, leaf_row, seed_note = nj_seed(ctx, gp, n_leaves)
    gp.free()
    st = sharding.PositionShardedTree(ctx, bp, c0, c1, leaf_row, rank, world)
    root = st.small_rectilinear()
    eu, ev = st.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = st.enumerate_moves(eu, ev, order)
    C_ = len(moves)
    dep = nj.depths(c0, c1)

    def align():
        ctx.sync()
        torch.cuda.synchronize()
        dist.barrier()

    # the new global root score of an accepted move is the winner's score: no collective in the move application
    r = nni_iterations(ctx, st, moves, order, args.steps, args.warmup, apply=lambda *m: st.apply_nni(*m, new_root=0), align=align)
    keys = ("stale_pass_ms", "iteration_ms", "scoring_only_ms", "wall_s")
    tt = torch.tensor([r[k] for k in keys], dtype=torch.float64, device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    r.update({k: float(v) for k, v in zip(keys, tt)})
    bt = torch.tensor([float(r["cand_bytes"] + r["up_bytes"])], dtype=torch.float64, device="cuda")
    dist.all_reduce(bt, op=dist.ReduceOp.SUM)
    ps_ms, it_ms = r["stale_pass_ms"], r["iteration_ms"]
    out = {
        "metric": "nni_candidates_scored_per_sec", "value": C_ / (ps_ms * 1e-3), "unit": "candidates/s", "n_gpus": world,
        "ms_per_step": ps_ms, "scaling": "strong",
        "e2e": {"value": r["candidates"] / r["wall_s"], "unit": "candidates/s", "h2d_bytes_per_step": r["moves_bytes"] + 16, "d2h_bytes_per_step": 72,
                "call": "whole hill-climb iterations on every rank (apply + launch + all-reduce + finish), wall clock, max over ranks"},
        "iteration": {"value": C_ / (it_ms * 1e-3), "unit": "candidates/s", "ms_per_step": it_ms},
        "scoring_only": {"value": C_ / (r["scoring_only_ms"] * 1e-3), "unit": "candidates/s", "ms_per_step": r["scoring_only_ms"]},
        "config": {"workload": name, "cells": n_leaves, "positions": L, "nodes": n, "candidates": C_,
                   "seed_tree": seed_note, "positions_per_rank": st.hi - st.lo,
                   "step": "per rank, after an accepted move: UP pass + all candidates over its column slice; one NCCL sum all-reduce of %d int32 "
                           "on the library stream; argmin; read-back" % (C_ + 1),
                   "mean_depth_v": float(dep[moves[:, 2]].mean()), "max_depth_v": int(dep[moves[:, 2]].max()),
                   "root_score": int(root), "best_index": r["best_index"], "best_score": r["best_score"],
                   "timing": "CUDA events on the library stream (the all-reduce is ordered on it), ranks aligned by a barrier before every step, max over ranks",
                   "l2": "flushed between steps"},
        "roofline": {"bound": "hbm", "achieved": float(bt[0]) / world / (ps_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s per GPU",
                     "frac": float(bt[0]) / world / (ps_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "traffic": None, "peak_kind": peaks_kind,
                     "algorithmic_bytes_per_pass_all_ranks": int(bt[0])},
    }
    out["trajectory"] = nni_trajectory(ctx, st, order, 20, apply=lambda *m: st.apply_nni(*m, new_root=0), align=align)
    st.free()
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    from lazac_copy_number_b200.api import Context, Profiles

    rank, local, world = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
    if not torch.cuda.is_available():
        sys.exit("no GPU: the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL prints its version banner on stdout at NCCL_DEBUG=VERSION; stdout carries one JSON line only
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/nccl_bench_%h_%p.log")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peaks, peaks_kind = measured_peaks()
    ctx = Context(local)

    name = args.workload or "cfg3"
    line = distance_section(ctx, args, name, rank, local, world, peaks, peaks_kind, cpu=not args.no_cpu)
    if world == 1 and name == "cfg3" and not args.no_cfg2:
        # BASELINE.json configs[1] beside the headline: same kernel, same calls, a matrix of 36 tiles
        line["cfg2"] = distance_section(ctx, args, "cfg2", rank, local, world, peaks, peaks_kind, cpu=not args.no_cpu)
    if world == 1 and not args.no_nni:
        line["nni"] = nni_section(ctx, args, peaks, peaks_kind)
        if args.workload != "tiny" and not args.no_nni_large:
            line["nni_large"] = nni_section_large(ctx, args, peaks, peaks_kind)
    if world > 1 and not args.no_nni:
        nni = nni_section_sharded(ctx, args, rank, world, peaks, peaks_kind)
        if rank == 0:
            line["nni"] = nni
    ctx.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def distance_section(ctx, args, name, rank, local, world, peaks, peaks_kind, cpu):
    """One distance workload: device-resident steps (value, roofline), then the host-buffer calls
    (e2e), then the reference's pair loop on a bounded sample (rank 0, one GPU only)."""
    import torch
    import torch.distributed as dist
    from lazac_copy_number_b200.api import Profiles

    w = WORKLOADS[name]
    from lazac_copy_number_b200 import synth
    prof = synth.make_profiles(w["cells"], w["bins"], w["alleles"], seed=0)   # same seed on every rank
    bp = bp_int8_chunked(prof["cn"], prof["seg"])
    del prof
    n, L = bp.shape
    P = n * (n + 1) // 2

    gp = Profiles(ctx, bp)
    d_D = torch.empty((n, n), dtype=torch.int32, device="cuda")
    shard, n_shards = rank, world

    def barrier():
        ctx.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step_device():
        ctx.flush_l2()
        ctx.timer_start()
        pairs = gp.distance_device_async(d_D.data_ptr(), shard, n_shards)
        return ctx.timer_stop(), pairs

    for _ in range(max(3, args.warmup)):
        step_device()
    n1_same = None
    if world > 1 and rank == 0:
        # the same matrix on ONE GPU (all shards on rank 0), for the strong-scaling denominator
        ts = []
        for _ in range(3):
            ctx.flush_l2()
            ctx.timer_start()
            gp.distance_device_async(d_D.data_ptr(), 0, 1)
            ts.append(ctx.timer_stop())
        n1_same = {"ms_per_step": float(np.median(ts)), "value": P / (float(np.median(ts)) * 1e-3), "unit": "cell-pairs/s",
                   "note": "this workload on one GPU of the same box, same kernel (measured before the timed region)"}
    barrier()
    launches0 = ctx.launch_count()
    clocks = ClockSampler(local)
    clocks.start()
    t_wall0 = time.perf_counter()
    ms_list, my_pairs = [], 0
    for _ in range(args.steps):
        ms, my_pairs = step_device()
        ms_list.append(ms)
    barrier()
    wall = time.perf_counter() - t_wall0
    clk = clocks.stop()
    launches = ctx.launch_count() - launches0
    dev_ms = float(sum(ms_list))
    tt = torch.tensor([dev_ms, float(my_pairs)], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = tt.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms_max, pairs_total = float(mx[0]), int(sm[1])
    else:
        dev_ms_max, pairs_total = dev_ms, my_pairs
    assert pairs_total == P, (pairs_total, P)
    value = P * args.steps / (dev_ms_max * 1e-3)

    # ---- e2e: host buffers through the C ABI (upload + compute + download), pinned memory ----
    h_bp = ctx.pinned(bp.shape, np.int8)
    h_bp[...] = bp
    h_D = ctx.pinned((n, n), np.int32)
    def step_e2e():
        p = Profiles(ctx, h_bp)
        if world == 1:
            p.distance_matrix(h_D)
        else:
            p.distance_shard(shard, n_shards, h_D)
        p.free()
    for _ in range(3):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0])
    d2h = 4 * n * n if world == 1 else None

    line = None
    if rank == 0:
        ms_step = dev_ms_max / args.steps
        lane_instr = P * (L / 4.0)                       # one VABSDIFF4.ACC per 4 positions of a pair
        achieved = lane_instr / world / (ms_step * 1e-3)  # per GPU
        hbm_bytes = n * L + 4 * n * n
        line = {
            "metric": "zcnt_cell_pairs_per_sec", "value": value, "unit": "cell-pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8 (int8 + 128 bias), int32 accumulate",
            "data": "synthetic",
            "config": {"workload": name, "cells": n, "bins": w["bins"], "alleles": w["alleles"], "positions": L, "pairs": P,
                       "position_ops_per_step": P * L, "sharding": "tile rows dealt in pairs (r, R-1-r) to %d rank(s)" % world,
                       "l2": "flushed between timed steps (256 MiB write)", "timing": "CUDA events on the library stream, summed over steps, max over ranks"},
            "e2e": {"value": P * args.steps / e2e_s, "unit": "cell-pairs/s", "h2d_bytes_per_step": int(n * L),
                    "d2h_bytes_per_step": int(d2h) if d2h is not None else int(4 * n * n // world),
                    "ms_per_step": 1e3 * e2e_s / args.steps,
                    "call": "lzg_profiles_upload_i8 + %s + lzg_profiles_free, pinned host buffers" % ("lzg_distance_matrix" if world == 1 else "lzg_distance_shard (every rank keeps its strips)")},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "int_alu", "kernel": "k_zcnt_distance", "achieved": achieved / 1e9, "peak": SAD4_PEAK_LANE_INSTR_PER_S / 1e9,
                         "unit": "G lane-instr/s (VABSDIFF4.U8.ACC, 4 positions each) per GPU", "frac": achieved / SAD4_PEAK_LANE_INSTR_PER_S,
                         "peak_kind": "measured (tools/microbench.cu on this pool, profiles/r01_microbench_int_pipes.jsonl)",
                         "traffic": NCU_DRAM_BYTES.get(("distance", name)),
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full (profiles/%s)" % NCU_DIST_FILES.get(name)
                                           if ("distance", name) in NCU_DRAM_BYTES else None,
                         "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / world / (ms_step * 1e-3) / 1e9, "peak_gbs": peaks["hbm_gbs"],
                                 "frac": hbm_bytes / world / (ms_step * 1e-3) / 1e9 / peaks["hbm_gbs"], "peak_kind": peaks_kind}},
            "wall_s_timed_region": wall,
        }
        if n1_same:
            line["n1_same_workload"] = n1_same
    if world == 1 and cpu:
        n_sample = args.ref_cells or (500 if L <= 10000 else 360)   # about 10 s of the reference's pair loop
        v, kind, dt, pairs = cpu_distance_sample(bp, min(n, n_sample))
        line["cpu_baseline"] = {"value": v, "unit": "cell-pairs/s", "cores": 1, "kind": kind,
                                "sample": "first %d cells of the same input (%d pairs, %.1f s)" % (min(n, n_sample), pairs, dt)}
    gp.free()
    del d_D, h_bp, h_D
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--ref-cells", type=int, default=0, help="cells in the CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-cfg2", action="store_true", help="skip the cfg2 section beside the cfg3 headline at --gpus 1")
    ap.add_argument("--no-nni", action="store_true", help="skip the NNI section")
    ap.add_argument("--no-nni-large", action="store_true", help="skip the cfg5-shaped NNI section at --gpus 1")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
