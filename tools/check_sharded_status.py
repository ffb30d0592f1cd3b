#!/usr/bin/env python
"""Two (or more) ranks, positions sharded: lzg_nni_best_sharded gives the unsharded scores, and when ONE rank's launch fails
(a quadruple that is no NNI move) every rank returns an error instead of waiting in the all-reduce; the trees stay usable.
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tools/check_sharded_status.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200._capi import LazacGpuError  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles, Tree  # noqa: E402
from lazac_copy_number_b200.sharding import PositionShardedTree  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
prof = synth.make_profiles(60, 400, 2, seed=3)
bp = synth.bp_from_cn(prof["cn"], prof["seg"]).astype(np.int32)
c0, c1 = synth.random_tree(60, np.random.default_rng(5))
leaf_row = np.full(2 * 60 - 1, -1, np.int32)
leaf_row[59:] = np.arange(60)
with Context(local) as ctx:
    whole = Tree(ctx, Profiles(ctx, bp), c0, c1, leaf_row)
    whole.small_rectilinear()
    eu, ev = whole.edges()
    order = np.arange(len(eu), dtype=np.int32)
    moves = whole.enumerate_moves(eu, ev, order)
    _, _, want = whole.nni_best(moves, want_all=True)
    st = PositionShardedTree(ctx, bp, c0, c1, leaf_row)
    st.small_rectilinear()
    assert st.comm is not None, "no in-library NCCL communicator"
    bi, bs, got = st.nni_best(moves, want_all=True)
    assert np.array_equal(got, want), "sharded scores differ"
    bad = moves.copy()
    if rank == world - 1:
        bad[3] = (0, 0, 0, 0)   # not a move, on the last rank only
    failed = False
    try:
        st.nni_best(bad, want_all=True)
    except LazacGpuError as e:
        failed = True
        print("[rank %d] error as expected: status %d, %s" % (rank, e.status, str(e)[:90]), flush=True)
    assert failed, "rank %d did not see the failure" % rank
    bi2, bs2, got2 = st.nni_best(moves, want_all=True)   # still usable
    assert np.array_equal(got2, want) and (bi2, bs2) == (bi, bs)
    st.free()
    whole.free()
dist.barrier()
if rank == 0:
    print("sharded status agreement OK on %d ranks" % world, flush=True)
dist.destroy_process_group()
