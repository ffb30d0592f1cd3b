"""GPU: the C++ host `lazac distance` / `lazac nni` end to end against the reference CLI's output
files (tests/golden/*_cli, produced by oracle/make_golden.py): every file byte for byte --
distance matrices, NJ seed tree, final tree with branch lengths, the per-iteration score log
(= the whole RNG-driven search trajectory) and the ancestral copy-number profiles."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
LAZAC = os.path.join(ROOT, "host", "lazac")
if not os.path.exists(LAZAC):  # normally built by __graft_entry__.build() and shipped with the snapshot
    subprocess.run(["make", "-C", os.path.join(ROOT, "host")], check=False, stdout=subprocess.DEVNULL)


def run(args):
    return subprocess.run([LAZAC] + args, check=True, capture_output=True, text=True)


def same(path, golden):
    with open(path, "rb") as a, open(golden, "rb") as b:
        return a.read() == b.read()


def test_ptx008_distance(tmp_path):
    src = os.path.join(GOLD, "ptx008_cli", "PTX008_cn_profile.csv")
    run(["distance", src, "-o", str(tmp_path / "P")])
    for f in ("_dist_matrix.csv", "_dist_matrix.txt"):
        assert same(tmp_path / ("P" + f), os.path.join(GOLD, "ptx008_cli", "PTX008" + f)), f


def test_ptx008_nni(tmp_path):
    """BASELINE.json configs[0]: lazac nni examples/PTX008_cn_profile.csv -a 2 (seed 0)."""
    src = os.path.join(GOLD, "ptx008_cli", "PTX008_cn_profile.csv")
    r = run(["nni", src, "-a", "2", "-s", "0", "-o", str(tmp_path / "P")])
    assert "Candidate tree scores @ iteration 0: [177, 166, 166, 165, 163, 161, 156, 155]" in r.stdout
    for f in ("_dist_matrix.txt", "_nj_tree.newick", "_tree.newick", "_info.json", "_ancestral_cn_profile.csv"):
        assert same(tmp_path / ("P" + f), os.path.join(GOLD, "ptx008_cli", "PTX008" + f)), f


@pytest.mark.parametrize("tag,flags", [
    ("full", ["-a", "1", "-i", "4", "-s", "3"]),
    ("greedy", ["-a", "1.5", "-i", "4", "-s", "5", "-g"]),
])
def test_synthetic_nni(tmp_path, tag, flags):
    src = os.path.join(GOLD, "synth_cli", "synth_cn_profile.csv")
    run(["nni", src] + flags + ["-o", str(tmp_path / tag)])
    for f in ("_dist_matrix.txt", "_nj_tree.newick", "_tree.newick", "_info.json", "_ancestral_cn_profile.csv"):
        assert same(tmp_path / (tag + f), os.path.join(GOLD, "synth_cli", tag + f)), f


def test_synthetic_nni_from_seed_tree(tmp_path):
    src = os.path.join(GOLD, "synth_cli", "synth_cn_profile.csv")
    tree = os.path.join(GOLD, "synth_cli", "full_nj_tree.newick")
    run(["nni", src, "-t", tree, "-a", "2", "-i", "3", "-s", "9", "-o", str(tmp_path / "fromtree")])
    for f in ("_tree.newick", "_info.json", "_ancestral_cn_profile.csv"):
        assert same(tmp_path / ("fromtree" + f), os.path.join(GOLD, "synth_cli", "fromtree" + f)), f
    assert not os.path.exists(tmp_path / "fromtree_nj_tree.newick")


def test_synthetic_distance(tmp_path):
    src = os.path.join(GOLD, "synth_cli", "synth_cn_profile.csv")
    run(["distance", src, "-o", str(tmp_path / "dist")])
    for f in ("_dist_matrix.csv", "_dist_matrix.txt"):
        assert same(tmp_path / ("dist" + f), os.path.join(GOLD, "synth_cli", "dist" + f)), f


def test_devices_flag(tmp_path):
    """--devices: the distance matrix on every GPU of the box (one GPU: the same single-device path), same bytes."""
    import torch
    g = torch.cuda.device_count()
    src = os.path.join(GOLD, "synth_cli", "synth_cn_profile.csv")
    run(["distance", src, "-o", str(tmp_path / "dist"), "--devices", "0-%d" % (g - 1)])
    for f in ("_dist_matrix.csv", "_dist_matrix.txt"):
        assert same(tmp_path / ("dist" + f), os.path.join(GOLD, "synth_cli", "dist" + f)), f
    run(["nni", src, "-a", "1", "-i", "4", "-s", "3", "-o", str(tmp_path / "full"), "--devices", ",".join(str(d) for d in range(g))])
    for f in ("_dist_matrix.txt", "_nj_tree.newick", "_tree.newick", "_info.json", "_ancestral_cn_profile.csv"):
        assert same(tmp_path / ("full" + f), os.path.join(GOLD, "synth_cli", "full" + f)), f


def test_no_gpu_flag_errors_cleanly(tmp_path):
    src = os.path.join(GOLD, "synth_cli", "synth_cn_profile.csv")
    r = subprocess.run([LAZAC, "distance", src, "-o", str(tmp_path / "x"), "--device", "99"], capture_output=True, text=True)
    assert r.returncode == 2 and "cannot open CUDA device" in r.stderr


def hashed_simulation_case(tmp_path, case_dir, flags, n_files):
    import gzip
    import hashlib
    d = os.path.join(GOLD, case_dir)
    src = tmp_path / "sim.csv"
    with gzip.open(os.path.join(d, "sim_cn_profile.csv.gz"), "rb") as f:
        src.write_bytes(f.read())
    run(["nni", str(src)] + flags + ["-o", str(tmp_path / "sim")])
    want = {}
    with open(os.path.join(d, "SHA256.txt")) as f:
        for line in f:
            if not line.startswith("#"):
                h, fn = line.split()
                want[fn] = h
    assert len(want) == n_files
    for fn, h in want.items():
        assert hashlib.sha256((tmp_path / fn).read_bytes()).hexdigest() == h, fn


def test_simulation_nni_long_trajectory(tmp_path):
    """Leaves of the paper simulation n100_l1000_s0: hundreds of neighbourhood passes; every output
    file must hash to what the reference CLI wrote."""
    hashed_simulation_case(tmp_path, "sim_cli", ["-a", "1", "-i", "2", "-s", "1"], 5)


def test_simulation_n600_hash_only(tmp_path):
    """Leaves of n600_l1000_s0 (1,467 neighbourhood passes, 1.75 M candidates, 170 s
    of the reference CLI on one core): `nni -a 1 -i 1 -s 1`, every output file by SHA-256."""
    hashed_simulation_case(tmp_path, "sim600_cli", ["-a", "1", "-i", "1", "-s", "1"], 5)
