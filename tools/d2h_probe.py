#!/usr/bin/env python
"""Aggregate device-to-host bandwidth of G GPUs writing disjoint slices of ONE pinned host buffer (what the tail of
lzg_distance_matrix_multi does), under three page placements of that buffer: default (first touch by the allocating thread),
interleaved over the NUMA nodes (set_mempolicy around the allocation), and one buffer per GPU.  Also host-to-device."""
import ctypes as C
import glob
import subprocess
import sys
import time

import torch

G = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
TOTAL = 1600 * 1000 * 1000
libc = C.CDLL(None, use_errno=True)
nodes = sorted(int(p.rsplit("node", 1)[1]) for p in glob.glob("/sys/devices/system/node/node[0-9]*"))
print("numa nodes:", nodes, flush=True)
try:
    print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=60).stdout[:3000], flush=True)
except Exception as e:  # noqa: BLE001
    print("topo:", e)


def mempolicy(mode, mask):
    m = C.c_ulong(mask)
    rc = libc.syscall(238, mode, C.byref(m), 64)  # SYS_set_mempolicy on x86_64
    return rc, C.get_errno()


def run(tag, bufs, d2h=True):
    per = TOTAL // G // 4
    dev = [torch.empty(per, dtype=torch.int32, device="cuda:%d" % g).fill_(g) for g in range(G)]
    streams = [torch.cuda.Stream(device=g) for g in range(G)]
    best = 1e9
    for it in range(4):
        for g in range(G):
            torch.cuda.synchronize(g)
        t0 = time.perf_counter()
        for g in range(G):
            with torch.cuda.stream(streams[g]):
                if d2h:
                    bufs[g].copy_(dev[g], non_blocking=True)
                else:
                    dev[g].copy_(bufs[g], non_blocking=True)
        for g in range(G):
            streams[g].synchronize()
        best = min(best, time.perf_counter() - t0)
    print("%-44s %s  %6.2f ms  %6.1f GB/s aggregate" % (tag, "D2H" if d2h else "H2D", best * 1e3, per * 4 * G / best / 1e9), flush=True)


def one_buffer():
    per = TOTAL // G // 4
    big = torch.empty(per * G, dtype=torch.int32).pin_memory()
    return [big[g * per:(g + 1) * per] for g in range(G)]


for g in range(G):
    torch.cuda.init()
    torch.empty(1, device="cuda:%d" % g)
b = one_buffer()
run("one pinned buffer, default placement", b)
run("one pinned buffer, default placement", b, d2h=False)
del b
if len(nodes) > 1:
    rc = mempolicy(3, sum(1 << n for n in nodes))  # MPOL_INTERLEAVE
    print("set_mempolicy(interleave):", rc, flush=True)
    b = one_buffer()
    mempolicy(0, 0)
    run("one pinned buffer, interleaved over the nodes", b)
    run("one pinned buffer, interleaved over the nodes", b, d2h=False)
    del b
    for n in nodes:
        rc = mempolicy(2, 1 << n)  # MPOL_BIND
        b = one_buffer()
        mempolicy(0, 0)
        run("one pinned buffer, bound to node %d" % n, b)
        del b
per = TOTAL // G // 4
b = [torch.empty(per, dtype=torch.int32).pin_memory() for g in range(G)]
run("one pinned buffer per GPU", b)
for k in (1, 2, 4):
    if k < G:
        G_save, G = G, k
        b = one_buffer()
        run("%d GPU(s) only, one buffer" % k, b)
        G = G_save
