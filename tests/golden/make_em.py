"""Generates tests/golden/em_K*.npz: EM known-answer vectors from the reference repository's vendored
Armadillo 8.400.0 gmm_full (oracle/_ref, built from /root/reference by oracle/Makefile).  Run in the
dev container only:  python tests/golden/make_em.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_ref  # noqa: E402
import pcoll_b200  # noqa: E402

N = 4096
for K in (1, 3, 8, 16):
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(5000 + K, N, K)
    X = x.T.astype(np.float64)
    out = dict(x=x, hefts0=h, means0=m, fcovs0=c)
    for it in (1, 5, 20):
        mm, cc, hh, ll = oracle_ref.em_fit(X, m, c.reshape(K, 9), h, it)
        out["means_%d" % it], out["fcovs_%d" % it], out["hefts_%d" % it], out["avg_log_p_after_%d" % it] = mm, cc.reshape(K, 3, 3), hh, ll
    np.savez_compressed(os.path.join(HERE, "em_K%d.npz" % K), **out)
    print(K, out["hefts_20"], out["avg_log_p_after_20"])
