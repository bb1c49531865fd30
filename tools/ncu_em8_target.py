"""ncu target: EM E-step at K = 8 (one-warp kernel) and K = 16 (split kernel), 8 fits x 10^6 samples, 2 iterations each."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
ctx = abi.Context(0)
probs = [pcoll_b200.inputs.synthetic_em_problem(9000 + f, 10 ** 6, 16) for f in range(8)]
x = np.stack([q[0] for q in probs])
first = True
for K in (8, 16, 4, 12):
    h = np.stack([np.full(K, 1.0 / K) for q in probs]); m = np.stack([q[2][:K] for q in probs]); c = np.stack([q[3][:K] for q in probs])
    r = ctx.em_fit(x if first else None, h, m, c, 2)
    first = False
    print("K=%d kernel %.3f ms per iteration" % (K, r["kernel_ms"] / 2))
ctx.close()
