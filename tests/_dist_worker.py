"""world_size-2 gloo worker for tests/test_dist.py: the N>1 host logic of bench.py's batch arm
(contiguous partition, max/sum over ranks, all-gather of the 40-byte summaries) without a GPU."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                   # noqa: E402
from qpdo_b200 import batch as qbatch          # noqa: E402

B = 1001                                       # ragged on purpose
dist = bench.Dist(2, backend="gloo")
lo, hi = qbatch.partition(B, dist.world, dist.rank)
summ = np.zeros(hi - lo, dtype=qbatch.SUMMARY_DTYPE)
summ["status_val"] = 1
summ["iterations"] = np.arange(lo, hi) % 23
summ["objective"] = np.arange(lo, hi) * 0.5
pad = np.zeros(((B + dist.world - 1) // dist.world) * 40, dtype=np.uint8)
pad[:summ.size * 40] = summ.view(np.uint8)
parts = []
for r, a in enumerate(dist.gather_bytes(pad)):
    l2, h2 = qbatch.partition(B, dist.world, r)
    parts.append(a[:(h2 - l2) * 40].view(qbatch.SUMMARY_DTYPE))
allsum = np.concatenate(parts)
assert allsum.size == B
assert np.array_equal(allsum["iterations"], np.arange(B) % 23)
assert np.array_equal(allsum["objective"], np.arange(B) * 0.5)
mx = dist.reduce([float(dist.rank + 1), 10.0 - dist.rank], "max")
assert mx == [2.0, 10.0], mx
sm = dist.reduce([float(hi - lo)], "sum")
assert sm == [float(B)], sm
dist.barrier()
dist.close()
if dist.rank == 0:
    print("DIST_OK")
