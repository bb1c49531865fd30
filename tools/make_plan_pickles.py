"""Writes data/plans/trajectory.dat and odometry.dat: the reference's committed plan (its two Python-2
pickles, hw2_astar.py:198-204) re-dumped at full precision as protocol-2 pickles of plain lists, so that
tools/mc_simulation.py can run where /root/reference does not exist.  Dev container only."""
import os
import pickle

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for name in ("trajectory.dat", "odometry.dat"):
    with open(os.path.join("/root/reference", name), "rb") as f:
        rows = pickle.load(f, encoding="latin1")
    rows = [[float(v) for v in np.asarray(r).ravel()] for r in rows]
    with open(os.path.join(ROOT, "data", "plans", name), "wb") as f:
        pickle.dump(rows, f, protocol=2)
    print(name, len(rows))
