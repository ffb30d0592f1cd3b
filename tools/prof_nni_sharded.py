#!/usr/bin/env python
"""The position-sharded NNI section of bench.py alone (cfg5 over the ranks of a torchrun launch,
world 1 included), with LZG_SHARD_TRACE / LZG_PASS_DEBUG=4 tracing if set.
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
           --master-port P tools/prof_nni_sharded.py [steps]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from lazac_copy_number_b200.api import Context  # noqa: E402

rank, local, world = bench.dist_env()
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
peaks, kind = bench.measured_peaks()
args = argparse.Namespace(workload=os.environ.get("WORKLOAD") or None, steps=int(sys.argv[1]) if len(sys.argv) > 1 else 10, warmup=3)
ctx = Context(local)
out = bench.nni_section_sharded(ctx, args, rank, world, peaks, kind)
ctx.close()
if rank == 0:
    print(json.dumps(out), flush=True)
dist.destroy_process_group()
