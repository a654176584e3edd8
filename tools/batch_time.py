"""quick throughput probe of the batch kernel (host-buffer entry): python tools/batch_time.py [B]"""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qpdo_b200 import problems, batch as qbatch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
t0 = time.time(); b = problems.mpc_batch(B); print("gen %.1fs" % (time.time() - t0))
bt, vals = qbatch.from_mpc(b)
print(bt.info())
for rep in range(3):
    t0 = time.time(); x, y, s, a = bt.solve(*vals, want_active=False); dt = time.time() - t0
    print("rep %d: %.3f s  -> %.0f QP/s; status==1: %d/%d; iters mean %.2f max %d" %
          (rep, dt, B / dt, int((s["status_val"] == 1).sum()), B, s["iterations"].mean(), s["iterations"].max()))
