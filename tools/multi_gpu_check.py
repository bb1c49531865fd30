"""Multi-GPU check (run under torchrun, one rank per GPU, NCCL): the three sharded drivers of
pcoll_b200.sharding against the same job on one GPU.  Prints one JSON line on rank 0.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402
from pcoll_b200 import sharding  # noqa: E402

abi = pcoll_b200.cuda_abi
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(local)
ctx.set_scene(sc["boxes"], sc["hulls"])
res = {"world": world}

# MC: job-wide counters must equal the single-GPU run of the same ids
N = 2_000_003
multi = sharding.mc_convergence_run(ctx, abi, plan, N)
p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False)
single = ctx.mc_run(p, want_hist=True)
res["mc_hits_multi"], res["mc_hits_single"] = multi["hits"], single["hits"]
assert multi["hits"] == single["hits"] and multi["executed_steps"] == single["executed_steps"]
assert np.array_equal(multi["first_hit_hist"], single["first_hit_hist"])

# scoring: plans sharded, P gathered
boxes = pcoll_b200.inputs.synthetic_obstacle_field()
tr, od = pcoll_b200.inputs.synthetic_plans(7, 10, 80, boxes, clear=-0.15)
ctx.set_scene(boxes, sc["hulls"])
P_multi = sharding.score_plans_sharded(ctx, plan, tr, od, 8192)
P_single = ctx.score_plans(plan, tr, od, 8192)["P"]
assert np.array_equal(P_multi, P_single), (P_multi, P_single)
res["score_P"] = P_multi.round(4).tolist()

# EM: samples sharded, statistics all-reduced every iteration
x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(5, 400_000, 5)
lo, hi = sharding.shard_range(x.shape[1], rank, world)
em = sharding.em_fit_sharded(ctx, np.ascontiguousarray(x[None, :, lo:hi]), x.shape[1], h[None], m[None], c[None], 6)
whole = ctx.em_fit(x, h, m, c, 6)
np.testing.assert_allclose(em["means"], whole["means"], rtol=1e-6)
np.testing.assert_allclose(em["hefts"], whole["hefts"], rtol=1e-5)
np.testing.assert_allclose(em["fcovs"], whole["fcovs"], rtol=1e-4, atol=1e-9)
gathered = [None] * world
dist.all_gather_object(gathered, em["means"].tolist())
assert all(g == gathered[0] for g in gathered), "ranks disagree on the fitted means"
res["em_means_max_rel_diff_vs_single"] = float(np.max(np.abs(em["means"] - whole["means"]) / np.maximum(np.abs(whole["means"]), 1)))
if rank == 0:
    print(json.dumps(res))
ctx.close()
dist.destroy_process_group()
