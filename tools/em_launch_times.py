#!/usr/bin/env python
"""Per-launch device times of the EM kernels at the sharded sizes (run under `ncu --metrics gpu__time_duration.sum`):
one fit per (K, n) with 3 iterations on device 0."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcoll_b200  # noqa: E402

abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
ctx = abi.Context(0)
for K in (1, 3, 16):
    for n in (1_000_000, 500_000, 125_000):
        x, h, m, c = inp.synthetic_em_problem(9000 + K, n, K)
        r = ctx.em_fit(x, h, m, c, 3)
        print(K, n, r["kernel_ms"] / 3)
