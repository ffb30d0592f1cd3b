"""Seeded synthetic copy-number profiles of the BASELINE.json shapes.

n leaves named cell_%06d; 22 chromosomes "1".."22" with B bins split as evenly as possible
(bin k of a chromosome has start = end = k); alleles cn_a (and cn_b when alleles == 2); a random
binary tree by uniform coalescence; root profile all 2; Poisson(3) events per edge, each a +-1
on a uniform interval of one (allele, chromosome), clipped to [0, 8].

Positions are emitted directly in the reference's breakpoint-vector order: segments sorted by
(chromosome string, allele string) as the std::map at /root/reference/src/copy_number.cxx:26
iterates, bins ascending inside a segment.
"""
import numpy as np

N_CHROM = 22
ALLELES = ("cn_a", "cn_b")


def segments(bins, alleles):
    """Segment table in reference order: list of (chrom_str, allele_idx, n_bins) and offsets."""
    per = [bins // N_CHROM + (1 if c < bins % N_CHROM else 0) for c in range(N_CHROM)]
    segs = []
    for c in range(N_CHROM):
        if per[c] == 0:
            continue
        for a in range(alleles):
            segs.append((str(c + 1), a, per[c]))
    segs.sort(key=lambda s: (s[0], ALLELES[s[1]]))
    off = np.zeros(len(segs) + 1, np.int32)
    off[1:] = np.cumsum([s[2] for s in segs])
    return segs, off


def random_tree(n_leaves, rng):
    """Uniform-coalescence rooted binary tree.  Returns (child0, child1) over 2n-1 nodes with
    root = 0, internal nodes 0 .. n-2 (parents have smaller ids than their children) and
    leaves n-1 .. 2n-2."""
    n_nodes = 2 * n_leaves - 1
    child0 = np.full(n_nodes, -1, np.int32)
    child1 = np.full(n_nodes, -1, np.int32)
    active = list(range(n_leaves - 1, n_nodes))
    nxt = n_leaves - 2
    while len(active) > 1:
        i, j = rng.choice(len(active), size=2, replace=False)
        a, b = active[i], active[j]
        child0[nxt], child1[nxt] = min(a, b), max(a, b)
        for k in sorted((i, j), reverse=True):
            active.pop(k)
        active.append(nxt)
        nxt -= 1
    return child0, child1


def make_profiles(n_leaves, bins, alleles=1, seed=0, events_per_edge=3.0, want_tree=False):
    """Returns dict(cn=int8[n][L], seg=offsets, segs=table, names=[...]) (+ tree arrays)."""
    rng = np.random.default_rng(seed)
    segs, off = segments(bins, alleles)
    L = int(off[-1])
    child0, child1 = random_tree(n_leaves, rng)
    n_nodes = 2 * n_leaves - 1
    cn = np.empty((n_leaves, L), np.int8)
    live = {0: np.full(L, 2, np.int8)}
    nseg = len(segs)
    # ids increase away from the root (parents are created after children, counting down),
    # so ascending id order visits parents first.
    for u in range(n_nodes):
        prof = live.pop(u)
        if child0[u] < 0:
            cn[u - (n_leaves - 1)] = prof
            continue
        for c in (child0[u], child1[u]):
            p = prof.copy()
            for _ in range(rng.poisson(events_per_edge)):
                s = int(rng.integers(nseg))
                lo, hi = int(off[s]), int(off[s + 1])
                a, b = sorted(int(x) for x in rng.integers(lo, hi, size=2))
                d = 1 if rng.random() < 0.5 else -1
                p[a:b + 1] = np.clip(p[a:b + 1].astype(np.int16) + d, 0, 8).astype(np.int8)
            live[int(c)] = p
    out = {"cn": cn, "seg": off, "segs": segs,
           "names": ["cell_%06d" % i for i in range(n_leaves)], "L": L}
    if want_tree:
        out["child0"], out["child1"] = child0, child1
    return out


def bp_from_cn(cn, seg):
    """Vectorised breakpoint transform with the reference's last-bin rule
    (/root/reference/src/copy_number.cxx:43-55): bp[0] = cn[0]-2; bp[i] = cn[i]-cn[i-1];
    bp[n-1] = 2-cn[n-1] (for n == 1 the first rule wins)."""
    cn = np.asarray(cn)
    wide = cn.astype(np.int32)
    bp = np.empty_like(wide)
    bp[:, 1:] = wide[:, 1:] - wide[:, :-1]
    seg = np.asarray(seg)
    first, last = seg[:-1], seg[1:] - 1
    multi = last > first
    bp[:, last[multi]] = 2 - wide[:, last[multi]]
    bp[:, first] = wide[:, first] - 2
    return bp


def random_bp(n, L, seed=0, lo=-8, hi=8, density=0.1):
    """Sparse random breakpoint vectors for kernel-level tests."""
    rng = np.random.default_rng(seed)
    v = rng.integers(lo, hi + 1, size=(n, L), dtype=np.int32)
    mask = rng.random((n, L)) < density
    return np.where(mask, v, 0).astype(np.int32)


def write_cn_csv(path, prof, both_alleles_columns=True):
    """CSV in the reference's input format (/root/reference/src/lazac.cxx:41-61):
    node,chrom,start,end,cn_a[,cn_b].  Only single-allele data round-trips through the
    reference CLI (it reads cn_a only)."""
    segs, off, cn = prof["segs"], prof["seg"], prof["cn"]
    by_chrom = {}
    for k, (ch, a, nb) in enumerate(segs):
        by_chrom.setdefault(ch, {})[a] = (int(off[k]), nb)
    with open(path, "w") as f:
        two = any(len(v) == 2 for v in by_chrom.values())
        f.write("node,chrom,start,end,cn_a" + (",cn_b" if two else "") + "\n")
        for i, name in enumerate(prof["names"]):
            for ch in sorted(by_chrom, key=int):
                lo_a, nb = by_chrom[ch][0]
                for k in range(nb):
                    row = "%s,%s,%d,%d,%d" % (name, ch, k, k, cn[i, lo_a + k])
                    if two:
                        row += ",%d" % cn[i, by_chrom[ch][1][0] + k]
                    f.write(row + "\n")


def perturb_tree(child0, child1, n_moves, rng):
    """n_moves random NNIs on a (child0, child1) tree with root 0: a cheap stand-in for a seed tree
    that is close to, but not at, a local optimum.  Returns new arrays."""
    c0, c1 = np.array(child0, np.int32), np.array(child1, np.int32)
    n = len(c0)
    parent = np.full(n, -1, np.int32)
    for u in range(n):
        if c0[u] >= 0:
            parent[c0[u]] = u
            parent[c1[u]] = u
    internal = np.flatnonzero((c0 >= 0) & (np.arange(n) != 0))
    if len(internal) == 0:
        return c0, c1
    for _ in range(n_moves):
        v = int(internal[rng.integers(len(internal))])
        u = int(parent[v])
        w = int(c1[u] if c0[u] == v else c0[u])
        z = int(c0[v] if rng.random() < 0.5 else c1[v])
        y = int(c1[v] if c0[v] == z else c0[v])
        c0[v], c1[v] = min(w, y), max(w, y)
        c0[u], c1[u] = min(v, z), max(v, z)
        parent[w], parent[z] = v, u
    return c0, c1
