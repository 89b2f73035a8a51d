#!/usr/bin/env python
"""kNN shard scan rate: nq queries against n_db device-resident rows (default 2000 x 1.25 M = one GPU's share of config 5)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lpslam_b200 import knn  # noqa: E402

n_db = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
g = torch.Generator(device="cuda").manual_seed(5)
db = torch.randint(0, 256, (n_db, 32), dtype=torch.uint8, device="cuda", generator=g)
q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device="cuda", generator=g)
d = knn.KnnDatabase(n_db)
d.add_dev(db.data_ptr(), n_db)
keys = torch.empty((nq, 2), dtype=torch.int64, device="cuda")
for _ in range(3):
    d.query_dev(q.data_ptr(), nq, keys.data_ptr())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e0.record()
for _ in range(reps):
    d.query_dev(q.data_ptr(), nq, keys.data_ptr())
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"{nq} x {n_db}: {ms:.3f} ms  {nq * n_db / ms / 1e6:.1f} G comparisons/s")
