"""Run under torchrun on N GPUs: a tree-sharded fit must give the 1-process oracle forest bit for bit.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dist_fit_check.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from euromillioner_b200 import synth  # noqa: E402
from euromillioner_b200.classifier import Dataset, RandomForestClassifier  # noqa: E402
from euromillioner_b200.dist import env_rank, make_context  # noqa: E402

rank, local_rank, world = env_rank()
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
ctx = make_context(dist)
X, y = synth.flat(50000, 32, 50)
rf = RandomForestClassifier(ctx).setNumTrees(13).setMaxDepth(8).setNumClasses(50).setMaxMemoryInMB(16)
model = rf.fit(Dataset(X, y, num_partitions=32))
e = model.export()
if rank == 0:
    from oracle import binding as ob
    o = ob.fit(ob.make_params(num_trees=13, max_depth=8, num_classes=50, max_memory_mb=16), X, y,
               ob.partition_offsets(len(X), 32)).export()
    for k in e:
        assert np.array_equal(e[k], o[k]), k
    print(f"DIST_FIT_OK world={world} nodes={model.totalNumNodes}")
# ---- row-sharded: each rank passes only its rows; node histograms are all-reduced per group
P = 32
goff = np.array([(i * len(X)) // P for i in range(P + 1)], np.int64)
p_lo, p_hi = (rank * P) // world, ((rank + 1) * P) // world
r_lo, r_hi = int(goff[p_lo]), int(goff[p_hi])
Xb, yb = synth.flat(200000, 16, 50)    # third case: nodes above and below 65535 rows in one group (packed 16-bit transport)
for kw in ({}, {"n": 9000}, {"big": True}):            # second case: N <= 10000 → findSplits keeps every row
    n = len(Xb) if kw.get("big") else kw.get("n", len(X))
    Xs, ys = (Xb, yb) if kw.get("big") else (X[:n], y[:n])
    goff = np.array([(i * n) // P for i in range(P + 1)], np.int64)
    r_lo, r_hi = int(goff[p_lo]), int(goff[p_hi])
    rf2 = RandomForestClassifier(ctx).setNumTrees(7).setMaxDepth(7).setNumClasses(50).setMaxMemoryInMB(16)
    m2 = rf2.fit(Dataset.shard(Xs[r_lo:r_hi], ys[r_lo:r_hi], goff, p_lo, p_hi - p_lo))
    e2 = m2.export()
    if rank == 0:
        o2 = ob.fit(ob.make_params(num_trees=7, max_depth=7, num_classes=50, max_memory_mb=16), Xs, ys, goff).export()
        for k in e2:
            assert np.array_equal(e2[k], o2[k]), ("row-sharded", n, k)
        print(f"DIST_ROWS_OK world={world} n={n} nodes={m2.totalNumNodes}")
    e = e2
# every rank must hold the same forest
digest = torch.tensor([float(e["threshold"].sum()), float(e["stats"].sum()), float(len(e["left"]))], device="cuda",
                      dtype=torch.float64)
lo, hi = digest.clone(), digest.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN)
dist.all_reduce(hi, op=dist.ReduceOp.MAX)
assert torch.equal(lo, hi), "ranks disagree on the forest"
dist.barrier()
dist.destroy_process_group()
