"""CPU: the GPU-free parts of the C++ host (host/lazac): the Clearcut-exact NJ seed tree and the
CSV -> breakpoint-vector ingest, against the reference CLI's own output files (tests/golden)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
LAZAC = os.path.join(ROOT, "host", "lazac")


@pytest.fixture(scope="module")
def lazac():
    # make is a no-op when the binary is current
    subprocess.run(["make", "-C", os.path.join(ROOT, "lazac_copy_number_b200", "csrc"), "-j4"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["make", "-C", os.path.join(ROOT, "host")], check=True, stdout=subprocess.DEVNULL)
    return LAZAC


@pytest.mark.parametrize("matrix,seed,tree", [
    ("ptx008_cli/PTX008_dist_matrix.txt", 0, "ptx008_cli/PTX008_nj_tree.newick"),
    ("synth_cli/full_dist_matrix.txt", 3, "synth_cli/full_nj_tree.newick"),
    ("synth_cli/greedy_dist_matrix.txt", 5, "synth_cli/greedy_nj_tree.newick"),
])
def test_nj_seed_tree_is_byte_identical(lazac, tmp_path, matrix, seed, tree):
    out = tmp_path / "nj.newick"
    subprocess.run([lazac, "nj", os.path.join(GOLD, matrix), "-s", str(seed), "-o", str(out)], check=True)
    with open(os.path.join(GOLD, tree), "rb") as f:
        assert out.read_bytes() == f.read()


def test_breakpoint_ingest_matches_reference(lazac, tmp_path):
    g = np.load(os.path.join(GOLD, "ptx008.npz"))
    out = tmp_path / "bp.txt"
    subprocess.run([lazac, "bp", os.path.join(GOLD, "ptx008_cli", "PTX008_cn_profile.csv"), "-o", str(out)], check=True)
    lines = out.read_text().split("\n")
    n, L = (int(x) for x in lines[0].split())
    assert (n, L) == g["bp"].shape
    for i in range(n):
        tok = lines[1 + i].split()
        assert tok[0] == str(g["names"][i])
        assert [int(x) for x in tok[1:]] == list(g["bp"][i])
    bins = [ln.split() for ln in lines[1 + n:1 + n + L]]
    assert [[int(b[0]), int(b[2]), int(b[3])] for b in bins] == g["bins"].tolist()
    assert all(b[1] == "cn_a" for b in bins)


def test_usage_errors(lazac):
    r = subprocess.run([lazac, "nni", "x.csv"], capture_output=True, text=True)
    assert r.returncode == 1 and "-o: required." in r.stderr
    r = subprocess.run([lazac, "frobnicate"], capture_output=True, text=True)
    assert r.returncode == 1
    r = subprocess.run([lazac, "bp", "/nonexistent.csv", "-o", "/dev/null"], capture_output=True, text=True)
    assert r.returncode == 2 and "Cannot open" in r.stderr


def test_indexed_stochastic_nni_equals_the_list_rebuilding_procedure(lazac):
    """host/search.cpp: stochastic_nni picks the k-th internal edge through a Fenwick tree instead of
    rebuilding the edge list before every draw (src/copy_number.cxx:419-462); `lazac selftest` runs both
    on random trees (2 .. 1000 leaves, aggressions 0 .. 1.75) and compares topologies and generator states,
    also for the draws-only variant used where do_nni drops the perturbed tree (src/lazac.cxx:284)."""
    r = subprocess.run([lazac, "selftest"], capture_output=True, text=True)
    assert r.returncode == 0 and "selftest ok" in r.stdout, r.stdout + r.stderr


def test_tree_from_seed_layout():
    """nj.tree_from_seed: lzg_nj_seed_tree's (left, right) -> lzg_tree_create's arrays."""
    import numpy as np
    from lazac_copy_number_b200 import nj
    left = np.array([-1, -1, -1, 0, 3], np.int32)    # node 3 = (leaf 0, leaf 2), root 4 = (3, leaf 1)
    right = np.array([-1, -1, -1, 2, 1], np.int32)
    c0, c1, leaf_row = nj.tree_from_seed(left, right, [2, 0, 1])
    assert c0.tolist() == [1, 2, -1, -1, -1] and c1.tolist() == [3, 4, -1, -1, -1]
    assert leaf_row.tolist() == [-1, -1, 2, 0, 1]
    assert nj.depths(c0, c1).tolist() == [0, 1, 2, 1, 2]


def test_csv_fast_path_equals_general_reader(lazac, tmp_path):
    """The in-place CSV splitter (plain files) and the RFC 4180 reader (quotes, stray CR) give the
    same breakpoint vectors; CRLF, blank lines, a missing final newline, extra columns, cells whose
    rows are interleaved and bins listed in different orders per cell are all covered."""
    rows = [("b2", "2", 10, 20, 3), ("a1", "1", 0, 10, 2), ("a1", "1", 10, 20, 4), ("b2", "1", 0, 10, 1),
            ("a1", "2", 10, 20, 2), ("b2", "1", 10, 20, 2), ("a1", "2", 0, 10, 5), ("b2", "2", 0, 10, 0),
            ("c3", "10", 0, 5, 7), ("c3", "1", 0, 10, 2), ("c3", "1", 10, 20, 2), ("c3", "2", 0, 10, 2), ("c3", "2", 10, 20, 1)]
    # c3 has one more bin: give the others that bin too (all cells must cover the same bins)
    rows += [("a1", "10", 0, 5, 2), ("b2", "10", 0, 5, 3)]

    def text(nl, quote, final_nl=True, blank=False):
        q = (lambda s: '"%s"' % s) if quote else (lambda s: s)
        lines = ["node,chrom,start,end,cn_a,cn_b,extra"]
        for k, (node, chrom, s, e, cn) in enumerate(rows):
            lines.append("%s,%s,%d,%d,%d,1,x" % (q(node), q(chrom), s, e, cn))
            if blank and k == 3:
                lines.append("")
        return nl.join(lines) + (nl if final_nl else "")

    outs = []
    for name, body in [("plain", text("\n", False)), ("crlf", text("\r\n", False)), ("quoted", text("\n", True)),
                       ("nofinal", text("\n", False, final_nl=False)), ("blank", text("\n", False, blank=True)),
                       ("quoted_crlf", text("\r\n", True))]:
        src = tmp_path / (name + ".csv")
        src.write_bytes(body.encode())
        out = tmp_path / (name + ".txt")
        subprocess.run([lazac, "bp", str(src), "-o", str(out)], check=True)
        outs.append(out.read_text())
    assert all(o == outs[0] for o in outs)
    first = outs[0].splitlines()
    assert first[0] == "3 5" and first[1].split()[0] == "a1"
    # a1: chrom "1": cn 2,4 -> [0, -2]; chrom "10": cn 2 -> single bin = first rule [0]; chrom "2": 5,2 -> [3, 0]
    assert [int(x) for x in first[1].split()[1:]] == [0, -2, 0, 3, 0]


def test_binary_profile_cache(tmp_path):
    """--cache: the second run reads CSV.lzbp instead of the CSV and yields the same breakpoint
    matrix; a changed CSV invalidates the cache."""
    import shutil
    import subprocess
    import time
    src = tmp_path / "c.csv"
    shutil.copy(os.path.join(ROOT, "tests", "golden", "synth_cli", "synth_cn_profile.csv"), src)
    lazac = os.path.join(ROOT, "host", "lazac")
    env = dict(os.environ, LAZAC_TIMING="1")

    def run(out):
        r = subprocess.run([lazac, "bp", str(src), "-o", str(tmp_path / out), "--cache"], capture_output=True, text=True, env=env, check=True)
        return r.stderr, (tmp_path / out).read_bytes()
    err1, b1 = run("b1")
    assert "read_cn_profiles" in err1 and (tmp_path / "c.csv.lzbp").exists()
    err2, b2 = run("b2")
    assert "profile cache" in err2 and "read_cn_profiles" not in err2 and b1 == b2
    plain = subprocess.run([lazac, "bp", str(src), "-o", str(tmp_path / "b0")], capture_output=True, text=True, check=True)
    assert (tmp_path / "b0").read_bytes() == b1 and plain.returncode == 0
    time.sleep(0.01)
    text = src.read_text().splitlines()
    text[1] = text[1].rsplit(",", 1)[0] + ",7"          # another copy number in the first row
    src.write_text("\n".join(text) + "\n")
    err3, b3 = run("b3")
    assert "read_cn_profiles" in err3 and b3 != b1        # stale cache ignored, rewritten
    err4, b4 = run("b4")
    assert "profile cache" in err4 and b4 == b3


def test_matrix_writers_across_thread_blocks(lazac, tmp_path):
    """write_matrix_csv / _txt format blocks of rows on several threads and write them in order: the files of a matrix that
    spans several blocks (2,500 rows of ~30 KB) must be what a plain loop writes (`lazac writebench` fills the matrix with
    a linear congruential sequence, restated here)."""
    n = 2500
    out = subprocess.run([lazac, "writebench", str(n), "-o", str(tmp_path / "wb")], capture_output=True, text=True, check=True)
    assert '"n": %d' % n in out.stdout
    x = np.empty(n * n, np.uint64)
    v = 12345
    a, c, m = 1664525, 1013904223, (1 << 32) - 1
    # the recurrence x <- a x + c (mod 2^32), vectorised by jumping: x_k = A_k x_0 + C_k
    A, Cc = np.empty(n * n, np.uint64), np.empty(n * n, np.uint64)
    ak, ck = 1, 0
    for k in range(n * n):
        ak, ck = (ak * a) & m, (ck * a + c) & m
        A[k], Cc[k] = ak, ck
    x = (A * np.uint64(v) + Cc) & np.uint64(m)
    D = (x >> np.uint64(20)).astype(np.int64).reshape(n, n)
    names = ["cell_%06d" % i for i in range(n)]
    csv = ",".join(names) + "\n" + "".join(names[i] + "," + ",".join(map(str, D[i].tolist())) + "\n" for i in range(n))
    txt = "%d\n" % n + "".join(names[i] + " " + " ".join(map(str, D[i].tolist())) + "\n" for i in range(n))
    assert (tmp_path / "wb_dist_matrix.csv").read_text() == csv
    assert (tmp_path / "wb_dist_matrix.txt").read_text() == txt
