"""Round-2 golden vectors, all from the reference's own code run here (oracle/_ref = unmodified
MCSimulator.h / GM_Model.h / vendored Armadillo, built by oracle/Makefile).  Dev container only:

    python tests/golden/make_round2.py [dist] [score] [em]

Outputs (committed; the GPU box has no /root/reference):
  tests/golden/mc_dist_pr2test2.json   run-to-run distribution of the reference's runSimulation() estimate,
                                       2,000 runs (mt19937_64 seeds 0..1999) x 1,000 particles: the 99 % CI of its
                                       mean has half-width < 0.003 (Philox-mode anchor, tests/test_mc_gpu.py)
  tests/golden/score_oracle.json       the same for 3 synthetic plans of the BASELINE config-4 generator against the
                                       256-box synthetic field: 256 runs x 1,000 particles per plan
                                       (batched-scoring anchor, tests/test_score_gpu.py)
  tests/golden/em_1e6_K{1,3,16}.npz    arma::gmm_full::learn(keep_existing) after 3 iterations on the config-5
                                       size (10^6 samples; the samples are regenerated from the seed by
                                       pcoll_b200.inputs.synthetic_em_problem, only the parameters are stored)
"""
import json
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLD = os.path.join(ROOT, "tests", "golden")

import pcoll_b200  # noqa: E402
from pcoll_b200 import inputs  # noqa: E402
import oracle_ref  # noqa: E402

SCORE_SEED, SCORE_T, SCORE_CLEAR = 4104, 120, -0.15


def _mc_job(args):
    plan, boxes, hulls, n, seed = args
    return oracle_ref.mc_run(plan, boxes, hulls, n, mt_seed=seed, record=False)["P"]


def score_problem():
    """The config-4 generator's field (256 boxes, PCG64(4096)) and 3 plans that graze it (P spans (0, 1))."""
    boxes = inputs.synthetic_obstacle_field()
    tr, od = inputs.synthetic_plans(SCORE_SEED, 3, SCORE_T, boxes, clear=SCORE_CLEAR)
    return boxes, tr, od


def main(what):
    plan = pcoll_b200.load_plan_json(os.path.join(GOLD, "plan_committed.json"))
    sc = np.load(os.path.join(GOLD, "scene_pr2test2.npz"))
    pool = mp.Pool(min(8, os.cpu_count() or 1))
    if "dist" in what:
        R, Nd = 2000, 1000
        Ps = pool.map(_mc_job, [(plan, sc["boxes"], sc["hulls"], Nd, s) for s in range(R)], chunksize=8)
        with open(os.path.join(GOLD, "mc_dist_pr2test2.json"), "w") as f:
            json.dump(dict(runs=R, particles=Nd, P=Ps, mean=float(np.mean(Ps)), std=float(np.std(Ps, ddof=1))), f)
        print("MC dist: mean %.5f std %.5f half-width(99%%) %.5f" % (np.mean(Ps), np.std(Ps, ddof=1), 2.576 * np.std(Ps, ddof=1) / np.sqrt(R)))
    if "score" in what:
        boxes, tr, od = score_problem()
        R, Nd = 256, 1000
        out = dict(seed=SCORE_SEED, T=SCORE_T, clear=SCORE_CLEAR, runs=R, particles=Nd, plans=[])
        for q in range(tr.shape[0]):
            pl = dict(plan)
            pl["trajectory"], pl["odometry"] = np.ascontiguousarray(tr[q]), np.ascontiguousarray(od[q])
            Ps = pool.map(_mc_job, [(pl, boxes, sc["hulls"], Nd, 1000 * q + s) for s in range(R)], chunksize=4)
            out["plans"].append(dict(mean=float(np.mean(Ps)), std=float(np.std(Ps, ddof=1)), P=Ps,
                                     start=[float(v) for v in tr[q, :, 0]]))
            print("score plan %d: reference mean %.4f std %.4f" % (q, np.mean(Ps), np.std(Ps, ddof=1)))
        with open(os.path.join(GOLD, "score_oracle.json"), "w") as f:
            json.dump(out, f)
    if "em" in what:
        for K in (1, 3, 16):
            x, h, m, c = inputs.synthetic_em_problem(9000 + K, 1_000_000, K)
            mm, cc, hh, ll = oracle_ref.em_fit(x.T.astype(np.float64), m, c.reshape(K, 9), h, 3)
            np.savez_compressed(os.path.join(GOLD, "em_1e6_K%d.npz" % K), seed=9000 + K, n=1_000_000, iters=3,
                                means=mm, fcovs=cc.reshape(K, 3, 3), hefts=hh, avg_log_p=ll)
            print("EM 1e6 K=%d: hefts %s avg_log_p %.6f" % (K, np.round(hh, 4), ll))
    pool.close()


if __name__ == "__main__":
    main(set(sys.argv[1:]) or {"dist", "score", "em"})
