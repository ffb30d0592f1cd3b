#!/usr/bin/env python
"""bench.py -- ZCNT hot-path throughput on B200 (contract: one JSON line on stdout from rank 0).

    python bench.py --gpus 1 --steps K --warmup W            # 1 GPU
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...                     # the reference's own CPU code

Workloads (BASELINE.json `configs`); a step = one whole distance matrix:
  cfg3  20,000 cells x 10,000 bins x 2 alleles (L = 20,000)  the headline at EVERY --gpus N: the
        matrix BASELINE.json's scaling target is quoted on; it fits one GPU (0.4 GB in, 1.6 GB
        out), so N = 1 runs the same matrix and the 1/2/4/8 series is one strong-scaling series
  cfg2  1,000 cells x 5,000 bins x 2 alleles (L = 10,000)   configs[1]; reported in the same
        line at --gpus 1 as "cfg2" (same calls, same keys), or alone with --workload cfg2
The NNI path is reported in the same line: "nni" = cfg4 (1,000 cells x 5,000 bins, NJ seed tree),
"nni_large" = cfg5's shape on the one GPU; at --gpus N > 1 "nni" = cfg5 with the positions sharded
over the ranks.  A step is one hill-climb iteration of lzg_hill_climb (accept the previous
winner + score the neighbourhood of the new tree), timed on the device by the kernels' own stamps.

value   = cell pairs (n(n+1)/2, as the reference's loop counts them, src/lazac.cxx:81-91) per
          second with the packed profiles already resident in HBM; device time (CUDA events on
          the library's stream), summed over the K steps, max over ranks; L2 flushed between steps.
e2e     = the same through the host-buffer C-ABI calls a caller of build_distance_matrix makes:
          pinned host profiles in, the WHOLE matrix in one pinned host buffer out.  N = 1:
          lzg_profiles_upload_i8 + lzg_distance_matrix; N > 1: rank 0 drives all N GPUs from one
          process (lzg_distance_matrix_multi: per-GPU upload of a row block, NVLink all-gather,
          mirrors written into the owners' memory, per-GPU copy-out), the other ranks idle.
Every roofline denominator and DRAM byte count is PARSED from a file under profiles/ that the JSON
line names (integer-pipe microbenchmark, ncu --set full summaries); MEASURED_PEAKS.json gives the
HBM peak.
"""
import argparse
import glob
import hashlib
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PROFILES = os.path.join(ROOT, "profiles")

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
PERTURB = 0.25      # the climbs start from the NJ seed tree after floor(0.25 x internal edges) random NNIs -- do_nni's candidate
                    # set is the seed tree perturbed with aggressions 0 .. 1.75 (src/lazac.cxx:243-248)


# ---- roofline inputs: parsed from profiles/, never pasted ---------------------------------------------
def latest(pattern):
    """Newest round's file matching profiles/<pattern> (rNN_ prefix sorts by round)."""
    hits = sorted(glob.glob(os.path.join(PROFILES, pattern)))
    return hits[-1] if hits else None


def microbench_rate(op_prefix):
    """lane-instructions / s of one instruction from the committed run of tools/microbench.cu."""
    path = latest("r*_microbench_int_pipes.jsonl")
    if path:
        with open(path) as f:
            for line in f:
                try:
                    row = json.loads(line)
                except ValueError:
                    continue
                if str(row.get("op", "")).startswith(op_prefix) and "lane_instr_per_s" in row:
                    return float(row["lane_instr_per_s"]), os.path.relpath(path, ROOT), row
    return None, None, None


_UNITS = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def ncu_value(path, metric, kernel_substr=None):
    """One metric of a tools/ncu_summary.py text file: value of the first launch whose kernel name contains
    kernel_substr (files with several launches list them side by side).  Byte units are converted to bytes."""
    if not path or not os.path.exists(path):
        return None
    kernels, col = None, 0
    with open(path) as f:
        lines = f.read().splitlines()
    for ln in lines:
        if ln.startswith("kernels:"):
            kernels = [k.strip() for k in ln[len("kernels:"):].split("|")]
    if kernel_substr and kernels:
        hits = [i for i, k in enumerate(kernels) if kernel_substr in k]
        if not hits:
            return None
        col = hits[0]
    for ln in lines:
        parts = ln.split()
        if parts and parts[0] == metric:
            rest = parts[1:]
            unit = None
            if rest and not re.match(r"^-?[0-9.,eE+]+$", rest[0]):
                unit, rest = rest[0], rest[1:]
            try:
                v = float(rest[col].replace(",", ""))
            except (IndexError, ValueError):
                return None
            return v * _UNITS.get(unit, 1.0)
    return None


def ncu_dram_bytes(path, kernel_substr=None):
    r, w = ncu_value(path, "dram__bytes_read.sum", kernel_substr), ncu_value(path, "dram__bytes_write.sum", kernel_substr)
    return None if r is None or w is None else r + w


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


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


def bp_int8_chunked(cn, seg, rows=2000):
    from lazac_copy_number_b200 import synth
    out = np.empty(cn.shape, np.int8)
    for r0 in range(0, cn.shape[0], rows):
        out[r0:r0 + rows] = synth.bp_from_cn(cn[r0:r0 + rows], seg).astype(np.int8)
    return out


def workload_bp(w):
    from lazac_copy_number_b200 import synth
    prof = synth.make_profiles(w["cells"], w["bins"], w["alleles"], seed=0)   # same seed on every rank
    return bp_int8_chunked(prof["cn"], prof["seg"])


# ---- the reference's CPU code (cpu_baseline legs and --impl reference) -------------------------------------
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


def seed_tree_arrays(bp, D=None):
    """The seed tree of the NNI sections, identical in both arms: Clearcut-exact NJ of the distance matrix with
    the taxon order of NJ_shuffle_distance_matrix for seed 0 (src/lazac.cxx:171-187).  On the CPU (liblazac_host.so,
    the code `lazac` links) when no GPU matrix is given -- the reference arm."""
    from lazac_copy_number_b200 import nj, search
    n = bp.shape[0]
    perm = search.clearcut_shuffle(n, 0)
    if D is None:
        from oracle import Port
        D = Port().distance_matrix(bp.astype(np.int32))
    left, right = search.nj_seed_tree_cpu(D, perm)
    return nj.tree_from_seed(left, right, perm), perm


def perturbed(c0, c1):
    """The starting tree of the measured climbs: floor(PERTURB x internal edges) random NNIs on the seed tree."""
    from lazac_copy_number_b200 import synth
    n_internal_edges = int(((c0 >= 0) & (np.arange(len(c0)) != 0)).sum())
    return synth.perturb_tree(c0, c1, int(PERTURB * n_internal_edges), np.random.default_rng(1))


def cpu_candidate_loop(bp, c0, c1, leaf_row, order, max_edges=None):
    """The reference's candidate loop (src/copy_number.cxx:347-366: nni, unvisit, small_rectilinear, undo per
    candidate) on one thread over the first max_edges entries of the shuffled edge order.  Returns
    (candidates/s, kind, seconds, candidates, scores, moves)."""
    from oracle import Port, PortTree, have_ref
    sub = order if max_edges is None else order[:max_edges]
    if have_ref():
        from oracle.ref import Ref
        rt = Ref().tree_i8(c0, c1, leaf_row, bp)
        rt.small_rectilinear()
        eu, ev = rt.edges()
        t0 = time.perf_counter()
        mv, sc = rt.candidate_scores(eu, ev, sub)
        dt = time.perf_counter() - t0
        return len(sc) / dt, "reference", dt, len(sc), sc, mv
    n = len(c0)
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[leaf_row >= 0] = bp[leaf_row[leaf_row >= 0]]
    pt = PortTree(Port(), c0, c1, leaf_bp)
    pt.small_rectilinear()
    eu, ev = pt.edges()
    t0 = time.perf_counter()
    rr = pt.greedy_nni(eu, ev, sub)
    dt = time.perf_counter() - t0
    return len(rr["scores"]) / dt, "port", dt, len(rr["scores"]), rr["scores"], rr["moves"]


def run_reference(args):
    """--impl reference: the reference's CPU implementation of both paths on this box's host cores
    (1 thread: the reference is single-threaded), each step a bounded sample."""
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
                   "positions": bp.shape[1], "sample": sample,
                   "note": "a per-pair rate at the workload's row length on a sub-sample of its cells: the whole cfg3 matrix is ~14 h of this loop"},
        "cpu_baseline": {"value": value, "unit": "cell-pairs/s", "cores": 1, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "cell-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_nni:
        # NNI candidates scored / s: the reference's candidate loop over whole neighbourhoods of cfg4's tree (the
        # same seed tree, perturbation and edge order as the B200 arm's "nni" section)
        nw = NNI_WORKLOADS["tiny" if name == "tiny" else "cfg4"]
        prof = synth.make_profiles(nw["cells"], nw["bins"], nw["alleles"], seed=0)
        nbp = bp_int8_chunked(prof["cn"], prof["seg"])
        (c0, c1, leaf_row), _ = seed_tree_arrays(nbp)
        c0, c1 = perturbed(c0, c1)
        order = np.random.default_rng(0).permutation(len(c0) - 1).astype(np.int32)
        tot_c, tot_t = 0, 0.0
        for _ in range(max(1, min(args.steps, 5))):
            v, kind, dt, cnt, _, _ = cpu_candidate_loop(nbp, c0, c1, leaf_row, order)
            tot_c += cnt
            tot_t += dt
        line["nni"] = {"metric": "nni_candidates_scored_per_sec", "value": tot_c / tot_t, "unit": "candidates/s", "ms_per_step": 1e3 * tot_t / max(1, min(args.steps, 5)),
                       "config": {"workload": "cfg4", "cells": nw["cells"], "positions": nbp.shape[1], "candidates": cnt,
                                  "seed_tree": "Clearcut-exact NJ (CPU, liblazac_host.so) of the C port's distance matrix, %g x internal edges random NNIs" % PERTURB,
                                  "step": "one whole neighbourhood through the reference's candidate loop (nni, unvisit, small_rectilinear, undo per candidate)"},
                       "cpu_baseline": {"value": tot_c / tot_t, "unit": "candidates/s", "cores": 1, "kind": kind, "sample": "whole neighbourhoods of cfg4's tree"}}
    emit(line)


# ---- NNI: the device-side hill climb ------------------------------------------------------------------------
def iteration_stats(stamps_list):
    """Per-iteration device times (us) from the %globaltimer stamps of one or more climbs."""
    it = np.concatenate([np.diff(s.astype(np.int64)) / 1e3 for s in stamps_list if len(s) > 1]) if stamps_list else np.zeros(0)
    if not len(it):
        return None
    return {"iterations": int(len(it)), "median_us": float(np.median(it)), "mean_us": float(it.mean()), "min_us": float(it.min()),
            "p90_us": float(np.percentile(it, 90)), "max_us": float(it.max())}


def climb_section(ctx, name, bp, c0, c1, leaf_row, seed_note, reps, max_moves, peaks, peaks_kind, cpu_edges, want_cpu):
    """One NNI workload on one GPU: real climbs (lzg_hill_climb) from the perturbed seed tree."""
    from lazac_copy_number_b200 import nj
    from lazac_copy_number_b200.api import Profiles, Tree
    n_leaves, L = bp.shape
    n = 2 * n_leaves - 1
    gp = Profiles(ctx, bp)
    p0, p1 = perturbed(c0, c1)
    base = Tree(ctx, gp, p0, p1, leaf_row)
    root = base.small_rectilinear()
    order = np.random.default_rng(0).permutation(n - 1).astype(np.int32)
    n_cand = 2 * int(((p0 >= 0) & (np.arange(n) != 0)).sum())
    dep = nj.depths(p0, p1)
    internal = np.flatnonzero((p0 >= 0) & (np.arange(n) != 0))
    # warm-up climb (graph instantiation, pools), then the measured ones; every climb starts from a fresh clone
    w = base.clone()
    w.hill_climb(order, max_moves=3, want_moves=False)
    w.free()
    stamps, walls, results = [], [], []
    for _ in range(reps):
        w = base.clone()
        ctx.flush_l2()
        ctx.sync()
        t0 = time.perf_counter()
        r = w.hill_climb(order, max_moves=max_moves, want_moves=True)
        walls.append(time.perf_counter() - t0)
        stamps.append(w.climb_stamps())
        results.append(r)
        w.free()
    r = results[-1]
    st = iteration_stats(stamps)
    # the first (full) pass kernel by kernel: one climb as queued launches with events between the kernels
    os.environ["LZG_CLIMB_GRAPH"] = "0"
    w = base.clone()
    ctx.flush_l2()
    ctx.sync()
    w.hill_climb(order, max_moves=1, want_moves=False)
    pass0 = w.climb_pass0_ms()
    w.free()
    del os.environ["LZG_CLIMB_GRAPH"]
    # one iteration per call with the L2 flushed before each (max_moves = 1: first pass + accept), CUDA events around the call
    flushed = []
    w = base.clone()
    for _ in range(6):
        ctx.flush_l2()
        ctx.sync()
        ctx.timer_start()
        rr = w.hill_climb(order, max_moves=1, want_moves=False)
        flushed.append(ctx.timer_stop())
        if rr["reason"] == 1:
            break
    w.free()
    # the older call-per-pass form for comparison (lzg_nni_best + lzg_tree_apply_nni, host enumeration between them)
    w = base.clone()
    eu, ev = w.edges()
    moves = w.enumerate_moves(eu, ev, order)
    idx, _, _ = w.nni_best(moves)
    step_ms = None
    if idx >= 0:
        m = tuple(int(x) for x in moves[idx])
        w.apply_nni(*m)
        eu1, ev1 = w.edges()
        # edge indices stay with their child ends: re-enumerate from the unchanged child list of the first topology
        moves1 = w.enumerate_moves(eu1, ev1, order)
        w.apply_nni(m[0], m[3], m[2], m[1])
        w.nni_best(moves)
        ctx.flush_l2()
        ctx.sync()
        ctx.timer_start()
        w.apply_nni(*m)
        w.nni_best(moves1)
        step_ms = ctx.timer_stop()
    w.free()
    per_unit_bytes = 10 * L / max(1, r["units_per_full_pass"] // max(1, n_cand // 2))   # 10 rows of one chunk group
    median_us = st["median_us"] if st else None
    iters = sum(len(s) - 1 for s in stamps if len(s) > 1)
    inc_units = (sum(x["item_units"] for x in results) - reps * r["units_per_full_pass"]) / max(1, iters)
    up_units = sum(x["up_units"] for x in results) / max(1, iters)
    items_full_bytes = (10 * (n_cand // 2) - 4 * int((p0[0] >= 0) * sum(1 for c in (p0[0], p1[0]) if p0[c] >= 0))) * L
    up_full_bytes = (4 * len(internal)) * L
    out = {
        "metric": "nni_candidates_scored_per_sec", "value": n_cand / (median_us * 1e-6) if median_us else None, "unit": "candidates/s",
        "ms_per_step": median_us / 1e3 if median_us else None,
        "iteration": st,
        "e2e": {"value": sum(x["candidates"] for x in results) / sum(walls), "unit": "candidates/s",
                "h2d_bytes_per_step": int(4 * (n - 1) + 44 * (n_cand // 2) + 8 * n) // max(1, r["n_passes"]), "d2h_bytes_per_step": int(64 + 20 * r["n_moves"]) // max(1, r["n_passes"]),
                "ms_per_climb": 1e3 * float(np.mean(walls)), "passes_per_climb": r["n_passes"],
                "call": "lzg_hill_climb per climb, wall clock: host layout of the work items, their upload with the caller's edge order, ONE graph launch (WHILE node), "
                        "read-back of the move log; candidates = passes x neighbourhood size"},
        "climb": {"moves": r["n_moves"], "passes": r["n_passes"], "reason": r["reason"], "start_score": int(root), "end_score": r["true_score"],
                  "returned_score": r["returned_score"], "max_moves": max_moves,
                  "moves_sha1": hashlib.sha1(np.ascontiguousarray(r["moves"]).tobytes()).hexdigest()[:16]},
        "incremental": {"item_units_per_iteration": float(inc_units), "units_per_full_pass": int(r["units_per_full_pass"]),
                        "fraction_rescored": float(inc_units / max(1, r["units_per_full_pass"])), "up_units_per_iteration": float(up_units),
                        "note": "an iteration re-scores only the (work item, chunk group) pairs that read a row the accepted move changed, and recomputes only "
                                "the UP rows whose inputs changed; every candidate's score is still produced every pass and is identical to a full pass "
                                "(tests/test_gpu_climb.py: whole trajectories against the oracle)"},
        "full_pass": None if not pass0 else {
            "ms": float(sum(pass0)), "value": n_cand / (sum(pass0) * 1e-3), "unit": "candidates/s",
            "kernels_ms": dict(zip(("k_up_df", "k_dirty_scan", "k_items_wl", "k_cand_assemble", "k_nni_argmin"), pass0)),
            "note": "the first pass of a climb: full UP pass + every work item, L2 flushed, CUDA events between the queued kernels"},
        "flushed_iteration_call_ms": [float(x) for x in flushed],
        "stepwise_iteration_ms": step_ms,
        "config": {"workload": name, "cells": n_leaves, "positions": L, "nodes": n, "candidates": n_cand, "seed_tree": seed_note,
                   "start": "%g x internal edges random NNIs on the seed tree (do_nni climbs from perturbed candidates, src/lazac.cxx:243-286)" % PERTURB,
                   "step": "one iteration of hill_climb on the device: accept the winner (k_apply2), incremental UP pass (k_up_delta), dirty scan, re-scoring of "
                           "the dirty work items (k_items_wl), score assembly for all candidates, arg-min; median over real climbs",
                   "timing": "%globaltimer stamps written by the arg-min kernel of every pass (the loop runs inside one CUDA graph: no host in it)",
                   "l2": "tree state %.0f MB incl. UP rows vs 126 MB of L2: %s; flushed_iteration_call_ms / full_pass are measured behind an L2 flush"
                         % (4 * n * L / 1e6, "far beyond L2" if 4 * n * L > 3e8 else "resident in L2 across iterations, as in the product's loop"),
                   "mean_depth_v": float(dep[internal].mean()), "max_depth_v": int(dep[internal].max())},
    }
    # rooflines: the full pass's dominant kernel (k_items_wl over every item) against the HBM peak on algorithmic bytes and, from the
    # committed ncu capture, against the integer ALU pipe; the incremental iteration is a latency chain, which neither describes
    ncu_file = latest("r*_ncu_climb_%s.txt" % ("cfg5" if n_leaves >= 5000 else "cfg4"))
    alu_pct = ncu_value(ncu_file, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "k_items_wl")
    traffic = ncu_dram_bytes(ncu_file, "k_items_wl")
    if pass0:
        t_items = pass0[2] * 1e-3
        out["roofline"] = {
            "kernel": "k_items_wl (first pass: every work item)", "bound": "int_alu" if (alu_pct or 0) > 60 else "hbm",
            "achieved": items_full_bytes / t_items / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": items_full_bytes / t_items / 1e9 / peaks["hbm_gbs"],
            "peak_kind": peaks_kind, "algorithmic_bytes": int(items_full_bytes), "traffic": traffic,
            "alu_pipe_pct_of_peak": alu_pct, "source": os.path.relpath(ncu_file, ROOT) if ncu_file else None,
            "note": "frac = algorithmic bytes (10 L per work item = per pair of candidates) / kernel time / HBM peak; alu_pipe_pct_of_peak is the integer ALU pipe "
                    "utilisation ncu measured for the same launch shape (the kernel is issue bound where it exceeds the HBM fraction)",
            "up_pass": {"kernel": "k_up_df", "achieved": up_full_bytes / (pass0[0] * 1e-3) / 1e9, "frac": up_full_bytes / (pass0[0] * 1e-3) / 1e9 / peaks["hbm_gbs"],
                        "algorithmic_bytes": int(up_full_bytes)},
            "iteration": {"bound": "latency", "algorithmic_bytes": float(inc_units * per_unit_bytes + up_units * 6 * per_unit_bytes / 10),
                          "note": "six dependent kernels on a few hundred rows: the accept walk (depth x row latency) dominates; see profiles/ launch lists"}}
    if want_cpu:
        v, kind, dt, cnt, sc, mv = cpu_candidate_loop(bp, p0, p1, leaf_row, order, cpu_edges)
        # the same candidates on the GPU (stepwise call with all scores) must give the same numbers
        w = base.clone()
        eu, ev = w.edges()
        moves = w.enumerate_moves(eu, ev, order if cpu_edges is None else order[:cpu_edges])
        _, _, allsc = w.nni_best(moves, want_all=True)
        w.free()
        out["cpu_baseline"] = {"value": v, "unit": "candidates/s", "cores": 1, "kind": kind,
                               "sample": "%d candidates of the same neighbourhood (%.1f s of the reference's candidate loop)" % (cnt, dt),
                               "scores_equal_gpu": bool(np.array_equal(sc, allsc) and np.array_equal(mv, moves))}
    # the calls around the climb: the bottom-up scoring pass (lzg_tree_score) and the ancestral labelling + download
    # (lzg_tree_ancestral), each behind an L2 flush; bytes = rows read + rows written of the device format (1 byte per value here)
    import ctypes as C
    ms_score = []
    for _ in range(3):
        base.set_topology(p0, p1)          # same topology: drops the cached scores so that the pass runs again
        ctx.flush_l2()
        ctx.sync()
        ctx.timer_start()
        base.small_rectilinear()
        ms_score.append(ctx.timer_stop())
    lab = np.zeros((n, L), np.int32)   # touched: the page faults of a fresh buffer are not the library's
    br = np.zeros(n, np.int32)
    wall_anc = []
    for _ in range(3):
        ctx.flush_l2()
        ctx.sync()
        t0 = time.perf_counter()
        ctx.check(ctx.lib.lzg_tree_ancestral(ctx.h, base.h, lab.ctypes.data_as(C.c_void_p), br.ctypes.data_as(C.c_void_p)))
        wall_anc.append(time.perf_counter() - t0)
    n_int = len(internal) + 1
    score_bytes = n_int * 6 * L            # two children's start/end rows read, the node's start/end written
    anc_dev_bytes = n * 4 * L              # per node: its interval (2 rows) and the parent's label read, its label written
    out["tail"] = {
        "lzg_tree_score": {"ms": float(min(ms_score)), "algorithmic_bytes": int(score_bytes),
                           "roofline": {"bound": "hbm", "achieved": score_bytes / (min(ms_score) * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                        "frac": score_bytes / (min(ms_score) * 1e-3) / 1e9 / peaks["hbm_gbs"],
                                        "note": "one kernel per tree level (a level waits for the one below): %d levels" % (int(dep.max()) + 1)}},
        "lzg_tree_ancestral": {"ms": 1e3 * float(min(wall_anc)), "d2h_bytes": int(n * L), "host_int32_bytes_written": int(4 * n * L),
                               "device_algorithmic_bytes": int(anc_dev_bytes),
                               "note": "wall clock of the call: labelling kernels level by level, the labels cross PCIe as bytes into pinned staging and "
                                       "host threads widen them into the caller's int32 matrix while the next block is copied"}}
    base.free()
    gp.free()
    return out


def nni_sections(ctx, args, line, peaks, peaks_kind):
    from lazac_copy_number_b200 import nj, search, synth
    from lazac_copy_number_b200.api import Profiles
    for key, name, reps, cap, cpu_edges in (("nni", "tiny" if args.workload == "tiny" else "cfg4", 4, 400, None), ("nni_large", "cfg5", 2, 80, 2500)):
        if key == "nni_large" and (args.workload == "tiny" or args.no_nni_large):
            continue
        w = NNI_WORKLOADS[name]
        bp = workload_bp(w)
        gp = Profiles(ctx, bp)
        t0 = time.perf_counter()
        D = gp.distance_matrix()
        t1 = time.perf_counter()
        perm = search.clearcut_shuffle(bp.shape[0], 0)
        left, right = ctx.nj_seed_tree(D, perm)
        t2 = time.perf_counter()
        # the same tree from the matrix the distance call left in HBM (what `lazac nni` does): no upload of D
        l2, r2 = gp.nj_seed_tree(perm)
        t3 = time.perf_counter()
        assert np.array_equal(l2, left) and np.array_equal(r2, right), "NJ from the device matrix differs"
        gp.free()
        c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
        n_taxa = bp.shape[0]
        scan_bytes = 4.0 * n_taxa ** 3 / 6.0     # the reference's n^3 / 6 float32 reads of the remaining triangle, one scan per join
        line.setdefault("nj", {})[name] = {
            "taxa": int(n_taxa), "ms_host_matrix": 1e3 * (t2 - t1), "ms_device_matrix": 1e3 * (t3 - t2),
            "roofline": {"bound": "hbm", "algorithmic_bytes": scan_bytes, "achieved": scan_bytes / (t3 - t2) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": scan_bytes / (t3 - t2) / 1e9 / peaks["hbm_gbs"], "peak_kind": peaks_kind,
                         "note": "wall clock of lzg_nj_seed_tree_profiles (gather, row sums, all joins in one cooperative launch, read-back); the joins "
                                 "are a chain: per join a grid barrier or two and n sequential float32 adds for the joined node's row sum "
                                 "(n^2 / 2 dependent adds in all), so small n sits far below the scan's bandwidth bound"}}
        note = "lzg_nj_seed_tree (Clearcut-exact NJ on the GPU, %.0f ms; taxon order of seed 0) of the GPU distance matrix (%.0f ms to the host)" % (
            1e3 * (t2 - t1), 1e3 * (t1 - t0))
        if name != "cfg5" and not args.no_cpu:   # the CPU twin the reference arm uses must give the same tree
            cl, cr = search.nj_seed_tree_cpu(D, perm)
            assert np.array_equal(cl, left) and np.array_equal(cr, right), "GPU and CPU NJ seed trees differ"
        line[key] = climb_section(ctx, name if name != "cfg5" else "cfg5 on 1 GPU", bp, c0, c1, leaf_row, note, reps, cap, peaks, peaks_kind,
                                  cpu_edges, not args.no_cpu)


def nni_section_sharded(ctx, args, rank, world, peaks, peaks_kind):
    """cfg5 (N > 1), positions sharded: every rank holds the same topology over its column slice, runs every
    iteration on 1/N of the bytes, and the ranks sum-all-reduce {root score, candidate scores} (NCCL, on the
    library stream) between the score assembly and the arg-min -- lzg_hill_climb_sharded.  Rank 0 also runs the
    same climb unsharded and the results must be identical."""
    import torch
    import torch.distributed as dist
    from lazac_copy_number_b200 import nj, search, sharding
    from lazac_copy_number_b200.api import Profiles, Tree
    name = "tiny" if args.workload == "tiny" else "cfg5"
    w = NNI_WORKLOADS[name]
    bp = workload_bp(w)
    n_leaves, L = bp.shape
    n = 2 * n_leaves - 1
    gp = Profiles(ctx, bp)                      # whole rows once, for the seed tree (the same on every rank)
    D = gp.distance_matrix()
    perm = search.clearcut_shuffle(n_leaves, 0)
    left, right = ctx.nj_seed_tree(D, perm)
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
    p0, p1 = perturbed(c0, c1)
    order = np.random.default_rng(0).permutation(n - 1).astype(np.int32)
    cap = 80
    one = None
    if rank == 0:                               # the 1-rank answer
        t1 = Tree(ctx, gp, p0, p1, leaf_row)
        t1.small_rectilinear()
        one = t1.hill_climb(order, max_moves=cap)
        t1.free()
    gp.free()
    st = sharding.PositionShardedTree(ctx, bp, p0, p1, leaf_row, rank, world)
    root = st.small_rectilinear()
    n_cand = 2 * int(((p0 >= 0) & (np.arange(n) != 0)).sum())
    c0g, c1g = st.tree.topology()
    stamps, walls, res = [], [], None
    for rep in range(3):
        st.tree.set_topology(p0, p1)
        st.small_rectilinear()
        ctx.flush_l2()
        ctx.sync()
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        res = st.hill_climb(order, max_moves=cap)
        walls.append(time.perf_counter() - t0)
        if rep > 0:
            stamps.append(st.tree.climb_stamps())
    s = iteration_stats(stamps)
    tt = torch.tensor([s["median_us"], s["mean_us"], s["max_us"], float(np.mean(walls[1:]))], dtype=torch.float64, device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    med, mean, mx, wall = (float(x) for x in tt)
    same = None
    if rank == 0:
        same = bool(one["n_moves"] == res["n_moves"] and one["true_score"] == res["true_score"] and one["returned_score"] == res["returned_score"]
                    and np.array_equal(one["moves"], res["moves"]))
        assert same, ("the sharded climb differs from the 1-rank climb", one["n_moves"], res["n_moves"], one["true_score"], res["true_score"])
    pass0 = st.tree.climb_pass0_ms()
    out = {
        "metric": "nni_candidates_scored_per_sec", "value": n_cand / (med * 1e-6), "unit": "candidates/s", "n_gpus": world,
        "ms_per_step": med / 1e3, "scaling": "strong",
        "iteration": {"median_us": med, "mean_us": mean, "max_us": mx, "iterations": s["iterations"], "note": "max over ranks of each rank's statistic"},
        "e2e": {"value": res["candidates"] / wall, "unit": "candidates/s", "ms_per_climb": 1e3 * wall, "passes_per_climb": res["n_passes"],
                "h2d_bytes_per_step": int(4 * (n - 1) + 44 * (n_cand // 2) + 8 * n) // max(1, res["n_passes"]), "d2h_bytes_per_step": int(64 + 20 * res["n_moves"]) // max(1, res["n_passes"]),
                "call": "lzg_hill_climb_sharded on every rank, wall clock, max over ranks"},
        "climb": {"moves": res["n_moves"], "passes": res["n_passes"], "reason": res["reason"], "start_score": int(root), "end_score": res["true_score"],
                  "returned_score": res["returned_score"], "max_moves": cap, "equal_to_1_rank_climb": same,
                  "moves_sha1": hashlib.sha1(np.ascontiguousarray(res["moves"]).tobytes()).hexdigest()[:16]},
        "full_pass": None if not pass0 else {"ms": float(sum(pass0)), "kernels_ms": dict(zip(("k_up_df", "k_dirty_scan", "k_items_wl", "k_cand_assemble", "k_nni_argmin"), pass0))},
        "config": {"workload": name, "cells": n_leaves, "positions": L, "nodes": n, "candidates": n_cand, "positions_per_rank": st.hi - st.lo,
                   "step": "per rank and iteration: accept, incremental UP pass, dirty scan, dirty items over its column slice, score assembly; one NCCL sum "
                           "all-reduce of %d int32 on the library stream; arg-min.  The kernels are queued in batches by every rank's host (a collective cannot "
                           "sit inside the WHILE graph); no host decision between them" % (n_cand + 1),
                   "timing": "%globaltimer stamps of every pass on each rank, median per rank, max over ranks", "l2": "tree state far beyond L2"},
    }
    st.free()
    return out


# ---- distance ---------------------------------------------------------------------------------------------------------
def distance_section(ctx, args, name, rank, local, world, peaks, peaks_kind, cpu, host_pg=None):
    """One distance workload: device-resident steps (value, roofline), then the host-buffer calls
    (e2e), then the reference's pair loop on a bounded sample (rank 0, one GPU only)."""
    import torch
    import torch.distributed as dist
    from lazac_copy_number_b200.api import MultiContext, Profiles

    w = WORKLOADS[name]
    bp = workload_bp(w)
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
    # which kernel ran (dispatch by value range), and -- outside the timed region -- the tensor-core kernel's own duration and
    # the other kernel on the same input
    path = gp.distance_path()
    tc_detail, alu_ms = None, None
    if path["tensor_core"]:
        os.environ["LZG_TC_TIMING"] = "1"
        prep, kern, tiles = [], [], 0
        for _ in range(5):
            ctx.flush_l2()
            gp.distance_device_async(d_D.data_ptr(), shard, n_shards)
            ctx.sync()
            q = gp.distance_path()
            prep.append(q.get("prepare_us", 0) / 1e3)
            kern.append(q.get("kernel_us", 0) / 1e3)
            tiles = q.get("tiles", 0)
        del os.environ["LZG_TC_TIMING"]
        tc_detail = {"prepare_ms": float(np.median(prep)), "kernel_ms": float(np.median(kern)), "tiles": int(tiles), "K": path["K"], "k_pad": path["k_pad"]}
        if rank == 0 and world == 1:
            os.environ["LZG_TC"] = "0"
            q = Profiles(ctx, bp)
            ts = []
            for _ in range(4):
                ctx.flush_l2()
                ctx.timer_start()
                q.distance_device_async(d_D.data_ptr(), 0, 1)
                ts.append(ctx.timer_stop())
            q.free()
            del os.environ["LZG_TC"]
            alu_ms = float(np.median(ts[1:]))
    del d_D

    # ---- e2e: host buffers through the C ABI, the WHOLE matrix in one pinned host buffer at the end of every step ----
    h_bp = ctx.pinned(bp.shape, np.int8)
    h_bp[...] = bp
    h_D = ctx.pinned((n, n), np.int32)
    e2e_extra = {}
    if world == 1:
        def step_e2e():
            p = Profiles(ctx, h_bp)
            p.distance_matrix(h_D)
            p.free()
        for _ in range(3):
            step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        barrier()
        e2e_s = time.perf_counter() - t0
        call = "lzg_profiles_upload_i8 + lzg_distance_matrix + lzg_profiles_free, pinned host buffers"
    else:
        # rank 0 drives all N GPUs from one process; the other ranks wait ON THE HOST (a gloo barrier: an NCCL barrier would
        # keep a spinning kernel on the very GPUs rank 0 is about to use)
        e2e_s = 0.0
        barrier()
        if rank == 0:
            gp.distance_matrix(h_D)                                    # the 1-GPU bytes
            want = hashlib.sha1(h_D.tobytes()).hexdigest()
            with MultiContext(list(range(world))) as m:
                h_D[...] = -1
                for _ in range(3):
                    m.distance_matrix(h_bp, out=h_D)
                got = hashlib.sha1(h_D.tobytes()).hexdigest()
                assert got == want, "the %d-GPU matrix differs from the 1-GPU matrix" % world
                t0 = time.perf_counter()
                for _ in range(args.steps):
                    m.distance_matrix(h_bp, out=h_D)
                e2e_s = time.perf_counter() - t0
            e2e_extra = {"matrix_equal_to_1gpu": True, "sha1": got[:16]}
        torch.cuda.set_device(local)
        dist.barrier(group=host_pg)
        call = ("lzg_distance_matrix_multi_i8 from rank 0's process over all %d GPUs (row block per GPU over its own PCIe link, NVLink all-gather, mirrors "
                "written into the owners' memory, per-GPU copy-out into the one pinned host matrix); the other ranks idle" % world)
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0])

    line = None
    if rank == 0:
        ms_step = dev_ms_max / args.steps
        lane_instr = P * (L / 4.0)                       # one VABSDIFF4.ACC per 4 positions of a pair
        achieved = lane_instr / world / (ms_step * 1e-3)  # per GPU
        hbm_bytes = n * L + 4 * n * n
        sad_peak, sad_file, sad_row = microbench_rate("VABSDIFF4")
        ncu_file = latest("r*_ncu_dist_%s%s*.txt" % ("tc_" if tc_detail else "", name)) or latest("r*_ncu_dist_%s*.txt" % name)
        traffic = ncu_dram_bytes(ncu_file)
        line = {
            "metric": "zcnt_cell_pairs_per_sec", "value": value, "unit": "cell-pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8 (int8 + 128 bias), int32 accumulate",
            "data": "synthetic",
            "config": {"workload": name, "cells": n, "bins": w["bins"], "alleles": w["alleles"], "positions": L, "pairs": P,
                       "position_ops_per_step": P * L, "sharding": "tile rows dealt in pairs (r, R-1-r) to %d GPU(s)" % world,
                       "l2": "flushed between timed steps (256 MiB write)", "timing": "CUDA events on the library stream, summed over steps, max over ranks"},
            "e2e": dict({"value": P * args.steps / e2e_s, "unit": "cell-pairs/s", "h2d_bytes_per_step": int(n * L),
                         "d2h_bytes_per_step": int(4 * n * n), "ms_per_step": 1e3 * e2e_s / args.steps, "call": call}, **e2e_extra),
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "int_alu", "kernel": "k_zcnt_distance", "achieved": achieved / 1e9, "peak": (sad_peak or 0) / 1e9,
                         "unit": "G lane-instr/s (VABSDIFF4.U8.ACC, 4 positions each) per GPU", "frac": achieved / sad_peak if sad_peak else None,
                         "peak_kind": "measured: tools/microbench.cu on this pool", "peak_source": sad_file,
                         "peak_lanes_per_clk_per_sm": sad_row.get("lane_instr_per_clk_per_sm") if sad_row else None,
                         "traffic": traffic, "traffic_source": os.path.relpath(ncu_file, ROOT) if ncu_file else None,
                         "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / world / (ms_step * 1e-3) / 1e9, "peak_gbs": peaks["hbm_gbs"],
                                 "frac": hbm_bytes / world / (ms_step * 1e-3) / 1e9 / peaks["hbm_gbs"], "peak_kind": peaks_kind}},
            "wall_s_timed_region": wall,
        }
        if tc_detail:
            # the exact tensor-core product: algorithmic operations = 2 x pairs x K (K = summed value range of the positions);
            # executed = 2 x tiles x 128 x 256 x k_pad (whole tiles along the diagonal, K padded to 128)
            # MEASURED_PEAKS.json holds a cuBLAS bf16 rate, no int8 one.  Twice that rate (int8 issues at twice the bf16 rate) is
            # EXCEEDED by this kernel, so the denominator is the nominal dense int8 peak of B200_PROFILING.md's table (2 x 2.25
            # PFLOP/s); the measured-based figure is reported beside it.
            int8_from_bf16 = 2.0 * peaks.get("bf16_tflops", 0.0)
            int8_peak = 4500.0
            k_s = tc_detail["kernel_ms"] * 1e-3
            alg_ops = 2.0 * P * tc_detail["K"] / world
            exe_ops = 2.0 * tc_detail["tiles"] * 128 * 256 * tc_detail["k_pad"]
            line["dtype"] = "u8 thermometer code x u8 -> int32 (tcgen05.mma kind::i8), int32 epilogue"
            line["roofline"] = {
                "bound": "tensor", "kernel": "k_tc_distance2 (CTA pairs)" if os.environ.get("LZG_TC_2CTA", "1") != "0" else "k_tc_distance", "achieved": alg_ops / k_s / 1e12 if k_s else None, "peak": int8_peak, "unit": "TOP/s (int8, dense) per GPU",
                "frac": alg_ops / k_s / 1e12 / int8_peak if k_s and int8_peak else None,
                "peak_kind": "nominal dense int8 (2 x 2.25 PFLOP/s bf16, B200_PROFILING.md): MEASURED_PEAKS.json has no int8 figure and 2 x its bf16 rate is exceeded",
                "peak_2x_measured_bf16": int8_from_bf16, "frac_of_2x_measured_bf16": alg_ops / k_s / 1e12 / int8_from_bf16 if k_s and int8_from_bf16 else None,
                "measured_peaks": peaks_kind,
                "kernel_ms": tc_detail["kernel_ms"], "prepare_ms": tc_detail["prepare_ms"],
                "algorithmic_ops": alg_ops, "executed_ops": exe_ops, "executed_frac": exe_ops / k_s / 1e12 / int8_peak if k_s and int8_peak else None,
                "K": tc_detail["K"], "k_pad": tc_detail["k_pad"], "K_per_position": tc_detail["K"] / float(L), "tiles": tc_detail["tiles"],
                "traffic": traffic, "traffic_source": os.path.relpath(ncu_file, ROOT) if ncu_file else None,
                "note": "ms_per_step (and value) time the whole call: value ranges + thermometer code (prepare_ms) + the product (kernel_ms, CUDA events "
                        "inside the library around the launch, measured after the timed region); dispatched by value range beside k_zcnt_distance",
                "hbm": line["roofline"]["hbm"]}
            if alu_ms:
                alu_ach = P * (L / 4.0) / (alu_ms * 1e-3)
                line["roofline"]["k_zcnt_distance_same_input"] = {
                    "ms": alu_ms, "bound": "int_alu", "achieved": alu_ach / 1e9, "peak": (sad_peak or 0) / 1e9, "frac": alu_ach / sad_peak if sad_peak else None,
                    "unit": "G lane-instr/s (VABSDIFF4.U8.ACC)", "peak_source": sad_file, "note": "LZG_TC=0: the kernel that runs for wide value ranges"}
        if n1_same:
            line["n1_same_workload"] = n1_same
    if world == 1 and cpu:
        n_sample = args.ref_cells or (500 if L <= 10000 else 360)   # about 10 s of the reference's pair loop
        v, kind, dt, pairs = cpu_distance_sample(bp, min(n, n_sample))
        line["cpu_baseline"] = {"value": v, "unit": "cell-pairs/s", "cores": 1, "kind": kind,
                                "sample": "first %d cells of the same input (%d pairs, %.1f s)" % (min(n, n_sample), pairs, dt)}
    gp.free()
    del h_bp, h_D
    return line


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries ONE JSON line: file descriptor 1 is pointed at stderr for everything else in the process
    (NCCL prints a version banner on fd 1 from C; libraries may print warnings), the line itself goes to the saved fd."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from lazac_copy_number_b200.api import Context

    rank, local, world = dist_env()
    host_pg = None
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
        host_pg = dist.new_group(backend="gloo")   # host-side waits (ranks that must leave their GPU idle)
    peaks, peaks_kind = measured_peaks()
    ctx = Context(local)

    name = args.workload or "cfg3"
    line = distance_section(ctx, args, name, rank, local, world, peaks, peaks_kind, cpu=not args.no_cpu, host_pg=host_pg if world > 1 else None)
    if world == 1 and name == "cfg3" and not args.no_cfg2:
        # BASELINE.json configs[1] beside the headline: same kernel, same calls, a matrix of 36 tiles
        line["cfg2"] = distance_section(ctx, args, "cfg2", rank, local, world, peaks, peaks_kind, cpu=not args.no_cpu)
    if world == 1 and not args.no_nni:
        nni_sections(ctx, args, line, peaks, peaks_kind)
    if world > 1 and not args.no_nni:
        nni = nni_section_sharded(ctx, args, rank, world, peaks, peaks_kind)
        if rank == 0:
            line["nni"] = nni
    ctx.close()
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--ref-cells", type=int, default=0, help="cells in the CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-cfg2", action="store_true", help="skip the cfg2 section beside the cfg3 headline at --gpus 1")
    ap.add_argument("--no-nni", action="store_true", help="skip the NNI sections")
    ap.add_argument("--no-nni-large", action="store_true", help="skip the cfg5-shaped NNI section at --gpus 1")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
