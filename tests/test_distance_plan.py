"""CPU: the work plan of the distance kernel (lzg_distance_plan, csrc/lzg_distance.cu) covers every
16-byte chunk of every tile of the reference's i <= j loop (src/lazac.cxx:81-91) exactly once --
whatever the tile order (row-major, super-block, phase-aligned), shard count or row group -- and
keeps the CTAs' loads equal.  No GPU: the plan is host code."""
import ctypes as C

import numpy as np
import pytest

from lazac_copy_number_b200 import _capi

INT32_MAX = 2**31 - 1


def plan(n, L, sms=148, shard=0, n_shards=1, row_lo=0, row_hi=INT32_MAX):
    lib = _capi.load()
    info = (C.c_int32 * 4)()
    lib.lzg_distance_plan(n, L, sms, shard, n_shards, row_lo, row_hi, None, 0, None, 0, info)
    n_ctas, n_segs, bm, K = list(info)
    segs = np.zeros((max(n_segs, 1), 4), np.int32)
    off = np.zeros(n_ctas + 1, np.int32)
    rc = lib.lzg_distance_plan(n, L, sms, shard, n_shards, row_lo, row_hi,
                               segs.ctypes.data_as(C.c_void_p), n_segs, off.ctypes.data_as(C.c_void_p), n_ctas, info)
    assert rc == 0
    return segs[:n_segs], off, bm, K


def check_cover(segs, bm, K, n, owned_rows):
    R = (n + bm - 1) // bm
    cover = {}
    for r, c, k0, k1 in segs.tolist():
        assert 0 <= r <= c < R and 0 <= k0 < abs(k1) <= K
        cover.setdefault((r, c), []).append((k0, abs(k1), k1 < 0))
    assert set(cover) == {(r, c) for r in owned_rows for c in range(r, R)}
    for parts in cover.values():
        parts.sort()
        assert parts[0][0] == 0 and parts[-1][1] == K
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))       # no gap, no overlap
        assert all(p[2] == (len(parts) > 1) for p in parts)              # cut tiles accumulate with atomics
    return cover


def loads(segs, off):
    cost = np.where(segs[:, 0] == segs[:, 1], 3, 4) * (np.abs(segs[:, 3]) - segs[:, 2]) + 26
    return np.array([cost[off[b]:off[b + 1]].sum() for b in range(len(off) - 1)])


@pytest.fixture(params=[("0", "0"), ("12", "0"), ("12", "1"), ("5", "1")], ids=["row-major", "super-block", "aligned", "aligned-sb5"])
def order(request, monkeypatch):
    monkeypatch.setenv("LZG_DIST_SB", request.param[0])
    monkeypatch.setenv("LZG_DIST_ALIGN", request.param[1])
    return request.param


@pytest.mark.parametrize("n,L", [(20000, 20000), (5000, 700), (1000, 10000), (129, 40), (64, 16), (1, 1)])
def test_whole_matrix_is_covered_once_and_balanced(order, n, L):
    segs, off, bm, K = plan(n, L)
    R = (n + bm - 1) // bm
    check_cover(segs, bm, K, n, range(R))
    ld = loads(segs, off)
    if len(ld) == 148 and R * (R + 1) // 2 >= 148:
        assert ld.max() <= 1.02 * ld.mean() + 4 * 8 + 26                  # equal up to one minimum segment


@pytest.mark.parametrize("n_shards", [2, 8])
def test_shards_partition_the_tile_rows(order, n_shards):
    n, L = 20000, 20000
    seen = set()
    for sh in range(n_shards):
        segs, off, bm, K = plan(n, L, shard=sh, n_shards=n_shards)
        R = (n + bm - 1) // bm
        rows = [r for r in range(R) if min(r, R - 1 - r) % n_shards == sh]
        cover = check_cover(segs, bm, K, n, rows)
        assert not (seen & set(cover))
        seen |= set(cover)
    assert len(seen) == R * (R + 1) // 2


def test_row_groups_as_lzg_distance_matrix_cuts_them(order):
    n, L = 20000, 20000
    R = (n + 127) // 128
    total = 0
    for g in range(16):
        ra, rb = R * g // 16, R * (g + 1) // 16
        segs, off, bm, K = plan(n, L, row_lo=ra, row_hi=rb)
        total += len(check_cover(segs, bm, K, n, range(ra, rb)))
    assert total == R * (R + 1) // 2


def test_super_block_order_keeps_the_cuts_and_shares_strips(monkeypatch):
    """Off-diagonal tiles all cost the same, so permuting them leaves every CTA's load where the
    row-major list put it; what changes is how many distinct operand strips the CTAs hold at the
    same step (what the L2 has to serve)."""
    def strips_per_step(segs, off):
        out = []
        for step in range(0, 80, 8):
            strips = set()
            for b in range(len(off) - 1):
                whole = [s for s in segs[off[b]:off[b + 1]].tolist() if s[0] != s[1]]
                if step < len(whole):
                    strips |= {("r", whole[step][0]), ("c", whole[step][1])}
            out.append(len(strips))
        return float(np.mean(out))
    monkeypatch.setenv("LZG_DIST_ALIGN", "0")
    monkeypatch.setenv("LZG_DIST_SB", "0")
    s0, o0, _, _ = plan(20000, 20000)
    monkeypatch.setenv("LZG_DIST_SB", "12")
    s1, o1, _, _ = plan(20000, 20000)
    assert np.array_equal(loads(s0, o0), loads(s1, o1))
    assert strips_per_step(s0, o0) > 150 and strips_per_step(s1, o1) < 60
    monkeypatch.setenv("LZG_DIST_ALIGN", "1")
    s2, o2, _, K = plan(20000, 20000)
    assert strips_per_step(s2, o2) < 40
    # aligned: every CTA starts with the same number of whole off-diagonal tiles
    lead = []
    for b in range(148):
        q = s2[o2[b]:o2[b + 1]]
        not_whole = (q[:, 2] != 0) | (q[:, 3] != K) | (q[:, 0] == q[:, 1])
        lead.append(int(np.argmax(not_whole)) if not_whole.any() else len(q))
    assert max(lead) - min(lead) <= 1 and min(lead) > 70      # (a CTA whose tail begins with an uncut tile counts one more)
