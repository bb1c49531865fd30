#!/usr/bin/env python
"""Sharded EM fit through the device group (mcsim_cuda_group_em_fit): device time per iteration with the statistics
exchange fused into the kernels (peer-memory push, M-step gathers) against one ncclAllReduce per iteration.

    python tools/em_group_quick.py [--n 1000000] [--iters 10] [--reps 5]

Prints one line per (devices, K, mode): ms per iteration (kernel_ms of the call = max over the devices, events on each
device's stream around all E / exchange / M launches), and the device-time of the LAST iteration's exchange + M-step."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcoll_b200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--lib", help="a variant of libmcsim_cuda.so (tools/build_variant.sh)")
    ap.add_argument("--groups", default="1,2,4,8")
    a = ap.parse_args()
    abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
    if a.lib:
        abi.LIB_PATH = os.path.abspath(a.lib)
    ndev = abi.lib().mcsim_cuda_device_count()
    out = []
    for G in [g for g in (int(v) for v in a.groups.split(",")) if g <= ndev]:
        grp = abi.Group(list(range(G)))
        try:
            for K in (1, 3, 16):
                x, h, m, c = inp.synthetic_em_problem(9000 + K, a.n, K)
                modes = [("single", abi.EM_EXCHANGE_AUTO)] if G == 1 else [("peer", abi.EM_EXCHANGE_PEER), ("nccl", abi.EM_EXCHANGE_NCCL)]
                ref = None
                for name, mode in modes:
                    if G > 1:
                        try:
                            grp.set_em_exchange(mode)
                        except abi.McsimCudaError as e:
                            print("G=%d %s: %s" % (G, name, e))
                            continue
                    r = grp.em_fit(x, h, m, c, a.iters)          # upload + warm-up
                    ms, tail, wall = [], [], []
                    for _ in range(a.reps):
                        t0 = time.perf_counter()
                        r = grp.em_fit(None, h, m, c, a.iters)
                        wall.append((time.perf_counter() - t0) * 1e3 / a.iters)
                        ms.append(r["kernel_ms"] / a.iters)
                        tail.append(grp.collective_stats()[0])
                    if ref is None:
                        ref = r
                    dev = float(np.max(np.abs(r["means"] - ref["means"])))
                    rec = dict(devices=G, K=K, mode=name, n=a.n, iters=a.iters, ms_per_iter=float(np.median(ms)),
                               ms_per_iter_min=float(np.min(ms)), wall_ms_per_iter=float(np.median(wall)), last_exchange_ms=float(np.median(tail)), max_mean_diff_vs_first_mode=dev,
                               iters_run=int(r["iters_run"][0]))
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
        finally:
            grp.close()
    return out


if __name__ == "__main__":
    main()
