#!/usr/bin/env python
"""oracle/make_golden.py -- TEST INFRASTRUCTURE ONLY.

Generates the golden vectors under tests/golden/ from the UNMODIFIED reference:
oracle/_ref/libref_shim.so (reference functions compiled from /root/reference/src/copy_number.cxx)
and oracle/_ref/lazac_ref (the reference CLI).  Run in the build container (needs
/root/reference); the outputs are committed because the reference does not travel to the GPU box.

    python oracle/make_golden.py

Files written
  ptx008.npz            examples/PTX008_cn_profile.csv (cn_a, as the reference reads it,
                        src/lazac.cxx:41-61): bins, cn, bp, D, names + tree results
  ptx008_cli/*          byte-exact outputs of `lazac_ref distance` and `lazac_ref nni -a 2 -s 0`
  tree_<tag>.npz        seeded synthetic cases: intervals/scores after small_rectilinear,
                        every candidate's score in enumeration order, greedy_nni's choice in both
                        modes, the state after applying it, ancestral labeling
  sim_n100_l1000_s0.npz leaves of data/simulations/n100_l1000_s0: bp, D, a whole
                        hill_climb(ranlux48_base(7)) result (topology + score)
"""
import csv
import gzip
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from lazac_copy_number_b200 import synth  # noqa: E402  (generator only: no GPU code is touched)
from oracle import build, have_ref, have_ref_bin  # noqa: E402
from oracle.ref import Ref, run_ref_cli  # noqa: E402

REFROOT = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def read_cn_csv(path, opener=open):
    """node -> list of (chrom, start, end, cn_a) in file order."""
    cells = {}
    with opener(path, "rt") as f:
        for row in csv.DictReader(f):
            cells.setdefault(row["node"], []).append(
                (int(row["chrom"]), int(row["start"]), int(row["end"]), int(row["cn_a"])))
    return cells


def bp_matrix(ref, cells, names):
    """Reference breakpoint vectors of `names` (rows) in the reference's position order."""
    rows, perm0, bins0 = [], None, None
    for nm in names:
        rec = np.array(cells[nm], np.int32)
        chrom, start, end, cn = rec[:, 0], rec[:, 1], rec[:, 2], rec[:, 3]
        bp, perm = ref.bp_from_cn(chrom, np.zeros_like(chrom), start, end, cn)
        rows.append(bp)
        if perm0 is None:
            perm0, bins0 = perm, rec[perm][:, :3]
    return np.array(rows, np.int32), bins0, perm0


def tree_case(ref, n_leaves, bins, alleles, seed):
    prof = synth.make_profiles(n_leaves, bins, alleles, seed=seed)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    rng = np.random.default_rng(seed + 100)
    c0, c1 = synth.random_tree(n_leaves, rng)
    n = 2 * n_leaves - 1
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    t = ref.tree(c0, c1, leaf_bp)
    out = {"bp": bp.astype(np.int8), "child0": c0, "child1": c1, "cn": prof["cn"], "seg": prof["seg"]}
    out["root_score"] = np.int32(t.small_rectilinear())
    S, E, sc = t.state()
    out["start"], out["end"], out["score"] = S.astype(np.int8), E.astype(np.int8), sc
    # labeling first: greedy_nni below leaves stale, un-visited intervals on the paths it probed
    # (src/copy_number.cxx:365-366), which the reference always re-scores before using them
    old, par, lab, br = t.ancestral_labeling()
    out["lab_old"], out["lab_parent"], out["lab_labels"], out["lab_branch"] = old, par, lab.astype(np.int8), br
    eu, ev = t.edges()
    order = np.random.default_rng(seed + 1).permutation(len(eu)).astype(np.int32)
    out["eu"], out["ev"], out["order"] = eu, ev, order
    moves, scores = t.candidate_scores(eu, ev, order)
    out["moves"], out["cand_scores"] = moves, scores
    for tag, greedy in (("best", False), ("first", True)):
        mv = t.clone().greedy_nni(eu, ev, order, greedy)
        out["move_" + tag] = np.array(mv if mv else (-1, -1, -1, -1), np.int32)
    mv = t.greedy_nni(eu, ev, order, False)
    if mv:
        t2 = t.clone()
        t2.apply_nni(*mv)
        out["applied_root_score"] = np.int32(t2.small_rectilinear())
        S2, E2, sc2 = t2.state()
        out["applied_start"], out["applied_end"], out["applied_score"] = S2.astype(np.int8), E2.astype(np.int8), sc2
    # the distance matrix of the leaves, by the reference's pair loop
    out["D"] = ref.distance_matrix(bp)
    return out


def sim_cli_case(sim_name, out_dir, flags, keep_small=True):
    """Leaves of one paper simulation through the reference CLI: the input travels gzipped, the
    outputs as SHA-256 (plus the small files themselves)."""
    import hashlib
    sim = os.path.join(REFROOT, "data", "simulations", sim_name)
    with open(sim + "_edgelist.csv") as f:
        edges = [(r["src"], r["dst"]) for r in csv.DictReader(f)]
    srcs = {a for a, _ in edges}
    leafset = {b for _, b in edges if b not in srcs}
    shutil.rmtree(out_dir, ignore_errors=True)
    os.makedirs(out_dir)
    with tempfile.TemporaryDirectory() as tmp:
        src3 = os.path.join(tmp, "sim.csv")
        with gzip.open(sim + "_full_cn_profiles.csv.gz", "rt") as f, open(src3, "w") as o:
            o.write("node,chrom,start,end,cn_a\n")
            for r in csv.DictReader(f):
                if r["node"] in leafset:  # leaves only; names must not look like numbers
                    o.write("c%s,%s,%s,%s,%s\n" % (r["node"], r["chrom"], r["start"], r["end"], r["cn_a"]))
        run_ref_cli(["nni", src3] + flags + ["-o", os.path.join(tmp, "sim")], timeout=3600)
        with open(src3, "rb") as f, gzip.GzipFile(os.path.join(out_dir, "sim_cn_profile.csv.gz"), "wb", mtime=0) as o:
            o.write(f.read())
        hashes = {}
        for fn in sorted(os.listdir(tmp)):
            if fn == "sim.csv":
                continue
            with open(os.path.join(tmp, fn), "rb") as f:
                data = f.read()
            hashes[fn] = hashlib.sha256(data).hexdigest()
            if keep_small and len(data) < 100000:
                shutil.copy(os.path.join(tmp, fn), os.path.join(out_dir, fn))
        with open(os.path.join(out_dir, "SHA256.txt"), "w") as f:
            f.write("# %s leaves: lazac_ref nni sim.csv %s -o sim   (sim.csv = gunzip sim_cn_profile.csv.gz)\n" % (sim_name, " ".join(flags)))
            for fn, h in hashes.items():
                f.write("%s  %s\n" % (h, fn))


def main():
    build()
    if not (have_ref() and have_ref_bin()):
        sys.exit("the reference is not built (needs /root/reference)")
    if "--n600" in sys.argv:   # the 600-cell paper simulation alone (about three minutes of the reference CLI)
        sim_cli_case("n600_l1000_s0", os.path.join(OUT, "sim600_cli"), ["-a", "1", "-i", "1", "-s", "1"], keep_small=True)
        return
    ref = Ref()
    os.makedirs(OUT, exist_ok=True)

    # ---- PTX008 ----------------------------------------------------------------------------
    src = os.path.join(REFROOT, "examples", "PTX008_cn_profile.csv")
    cells = read_cn_csv(src)
    names = sorted(cells, key=lambda s: s.encode())
    bp, bins, perm = bp_matrix(ref, cells, names)
    cn = np.array([[r[3] for r in cells[nm]] for nm in names], np.int32)
    D = ref.distance_matrix(bp)
    cli = os.path.join(OUT, "ptx008_cli")
    shutil.rmtree(cli, ignore_errors=True)
    os.makedirs(cli)
    with tempfile.TemporaryDirectory() as tmp:
        run_ref_cli(["distance", src, "-o", os.path.join(tmp, "PTX008")])
        run_ref_cli(["nni", src, "-a", "2", "-s", "0", "-o", os.path.join(tmp, "PTX008")])
        for f in sorted(os.listdir(tmp)):
            shutil.copy(os.path.join(tmp, f), os.path.join(cli, f))
    # the CLI's matrix must equal the shim's (same reference functions)
    with open(os.path.join(cli, "PTX008_dist_matrix.csv")) as f:
        rows = list(csv.reader(f))
    assert rows[0] == names  # header row: names only (src/lazac.cxx:108-112)
    assert np.array_equal(np.array([[int(x) for x in r[1:]] for r in rows[1:]]), D)
    np.savez_compressed(os.path.join(OUT, "ptx008.npz"), names=np.array(names), cn_file_order=cn,
                        perm=perm, bins=bins, bp=bp, D=D)

    # ---- synthetic tree cases ----------------------------------------------------------------
    for tag, args in (("n10_b255_a1", (10, 255, 1, 1)), ("n40_b300_a2", (40, 300, 2, 2)),
                      ("n120_b800_a2", (120, 800, 2, 3)), ("n3_b30_a1", (3, 30, 1, 4))):
        np.savez_compressed(os.path.join(OUT, "tree_%s.npz" % tag), **tree_case(ref, *args))

    # ---- one paper simulation ----------------------------------------------------------------
    sim = os.path.join(REFROOT, "data", "simulations", "n100_l1000_s0")
    cells = read_cn_csv(sim + "_full_cn_profiles.csv.gz", gzip.open)
    with open(sim + "_edgelist.csv") as f:
        edges = [(r["src"], r["dst"]) for r in csv.DictReader(f)]
    srcs = {a for a, _ in edges}
    leaves = sorted({b for _, b in edges if b not in srcs}, key=lambda s: ("c" + s).encode())
    bp, bins, perm = bp_matrix(ref, cells, leaves)
    D = ref.distance_matrix(bp)
    n_leaves = len(leaves)
    rng = np.random.default_rng(0)
    c0, c1 = synth.random_tree(n_leaves, rng)
    n = 2 * n_leaves - 1
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    t = ref.tree(c0, c1, leaf_bp)
    s0 = t.small_rectilinear()
    gen = ref.gen(7)
    t.hill_climb(gen, False)
    s1 = t.small_rectilinear()
    h0, h1 = t.topology()
    np.savez_compressed(os.path.join(OUT, "sim_n100_l1000_s0.npz"), names=np.array(["c" + s for s in leaves]),
                        bp=bp.astype(np.int8), D=D, child0=c0, child1=c1, score_before=np.int32(s0),
                        hc_seed=np.int32(7), hc_child0=h0, hc_child1=h1, hc_score=np.int32(s1),
                        gen_next_after=np.uint64(gen.next()))
    # ---- CLI runs on a synthetic single-allele input (what the reference's loader reads) -------
    cli2 = os.path.join(OUT, "synth_cli")
    shutil.rmtree(cli2, ignore_errors=True)
    os.makedirs(cli2)
    prof = synth.make_profiles(30, 150, 1, seed=11)
    src2 = os.path.join(cli2, "synth_cn_profile.csv")
    synth.write_cn_csv(src2, prof)
    shutil.copy(src, os.path.join(cli, "PTX008_cn_profile.csv"))   # inputs travel with the goldens
    runs = {"full": ["-a", "1", "-i", "4", "-s", "3"], "greedy": ["-a", "1.5", "-i", "4", "-s", "5", "-g"]}
    for tag, flags in runs.items():
        run_ref_cli(["nni", src2] + flags + ["-o", os.path.join(cli2, tag)])
    # a run from a given seed tree (-t): the NJ tree of the first run, other seed
    run_ref_cli(["nni", src2, "-t", os.path.join(cli2, "full_nj_tree.newick"), "-a", "2", "-i", "3", "-s", "9",
                 "-o", os.path.join(cli2, "fromtree")])
    run_ref_cli(["distance", src2, "-o", os.path.join(cli2, "dist")])
    with open(os.path.join(cli2, "COMMANDS.txt"), "w") as f:
        f.write("lazac_ref nni synth_cn_profile.csv -a 1 -i 4 -s 3 -o full\n"
                "lazac_ref nni synth_cn_profile.csv -a 1.5 -i 4 -s 5 -g -o greedy\n"
                "lazac_ref nni synth_cn_profile.csv -t full_nj_tree.newick -a 2 -i 3 -s 9 -o fromtree\n"
                "lazac_ref distance synth_cn_profile.csv -o dist\n"
                "(ptx008_cli: lazac_ref distance / nni -a 2 -s 0 on PTX008_cn_profile.csv)\n")
    # ---- CLI run on the leaves of one paper simulation (100 cells x 1000 loci): a long trajectory ----
    cli3 = os.path.join(OUT, "sim_cli")
    shutil.rmtree(cli3, ignore_errors=True)
    os.makedirs(cli3)
    import hashlib
    leafset = set(leaves)
    with tempfile.TemporaryDirectory() as tmp:
        src3 = os.path.join(tmp, "sim.csv")
        with gzip.open(sim + "_full_cn_profiles.csv.gz", "rt") as f, open(src3, "w") as o:
            o.write("node,chrom,start,end,cn_a\n")
            for r in csv.DictReader(f):
                if r["node"] in leafset:  # leaves only; names must not look like numbers
                    o.write("c%s,%s,%s,%s,%s\n" % (r["node"], r["chrom"], r["start"], r["end"], r["cn_a"]))
        run_ref_cli(["nni", src3, "-a", "1", "-i", "2", "-s", "1", "-o", os.path.join(tmp, "sim")])
        with open(src3, "rb") as f, gzip.GzipFile(os.path.join(cli3, "sim_cn_profile.csv.gz"), "wb", mtime=0) as o:
            o.write(f.read())
        hashes = {}
        for fn in sorted(os.listdir(tmp)):
            if fn == "sim.csv":
                continue
            with open(os.path.join(tmp, fn), "rb") as f:
                data = f.read()
            hashes[fn] = hashlib.sha256(data).hexdigest()
            if len(data) < 100000:
                shutil.copy(os.path.join(tmp, fn), os.path.join(cli3, fn))
        with open(os.path.join(cli3, "SHA256.txt"), "w") as f:
            f.write("# lazac_ref nni sim.csv -a 1 -i 2 -s 1 -o sim   (sim.csv = gunzip sim_cn_profile.csv.gz)\n")
            for fn, h in hashes.items():
                f.write("%s  %s\n" % (h, fn))
    total = sum(os.path.getsize(os.path.join(dp, f)) for dp, _, fs in os.walk(OUT) for f in fs)
    print("golden vectors written to %s (%.1f KiB)" % (OUT, total / 1024))


if __name__ == "__main__":
    main()
