"""One process per GPU: how the three operators shard across ranks (SURVEY.md 8e / DESIGN.md 5).

  MC convergence run   rollout ids [lo, hi) per rank, Philox counters from the GLOBAL id; one all-reduce
                       (sum) of the uint64 counter block {hits, near, steps, tests, first_hit_hist[T]}
  batched scoring      plans [lo, hi) per rank (plan id in the Philox counter); all-gather of P
  EM fit               samples [lo, hi) per rank, parameters replicated; one all-reduce (sum, FP64) of the
                       K * 10 + 1 sufficient statistics per fit and iteration (the reference's own
                       map / reduce shape, gmm_full_meat.hpp:2509-2548), identical M-step everywhere

The collectives are torch.distributed calls on the tensors the context exposes (NCCL on the GPU box,
issued on the library's stream; gloo in the CPU tests, which drive these functions with a stand-in
context).  No arithmetic of the path lives here."""
import contextlib

import numpy as np


def shard_range(total, rank, world):
    """Contiguous, exhaustive, disjoint: [lo, hi) of `total` units for `rank` of `world`."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("rank %r of world %r" % (rank, world))
    return (total * rank) // world, (total * (rank + 1)) // world


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _world():
    d = _dist()
    return (d.get_rank(), d.get_world_size()) if d else (0, 1)


@contextlib.contextmanager
def _on_stream(ctx):
    """Issue torch work on the context's own CUDA stream (no-op for CPU stand-ins)."""
    stream = getattr(ctx, "torch_stream", None)
    if stream is None:
        yield
    else:
        import torch
        with torch.cuda.stream(stream):
            yield


def mc_convergence_run(ctx, abi, plan, n_total, seed=0x5AFEC0DE, early_exit=False, estimator=0):
    """BASELINE config 3: n_total rollouts of one plan over all ranks.  Returns the job-wide counters
    (identical on every rank, independent of the world size)."""
    rank, world = _world()
    lo, hi = shard_range(n_total, rank, world)
    T = plan["trajectory"].shape[1]
    p, keep = abi.make_mc_params(plan, n_rollouts=hi - lo, rollout_offset=lo, mode=abi.MODE_PHILOX,
                                 early_exit=early_exit, estimator=estimator, seed=seed)
    ctx.mc_prepare(p)
    ctx.mc_launch(hi - lo, lo)
    d = _dist()
    if d and world > 1:
        with _on_stream(ctx):
            d.all_reduce(ctx.mc_counters_tensor(T))  # int64 view of the uint64 block: sum
    out = ctx.mc_collect(T, want_hist=True)
    out["P"] = out["hits"] / n_total if n_total else 0.0
    return out


def score_plans_sharded(ctx, model, trajectories, odometry, rollouts, **kw):
    """BASELINE config 4: plans split over ranks, P all-gathered.  Returns P[n_plans] on every rank."""
    import torch
    rank, world = _world()
    n = trajectories.shape[0]
    lo, hi = shard_range(n, rank, world)
    P = np.zeros(n)
    if hi > lo:
        P[lo:hi] = ctx.score_plans(model, trajectories[lo:hi], odometry[lo:hi], rollouts, plan_offset=lo, **kw)["P"]
    d = _dist()
    if d and world > 1:
        # ranks own disjoint slices and the rest is zero: a sum is the gather (16 KB for 4,096 plans)
        dev = "cuda:%d" % ctx.device if getattr(ctx, "device", None) is not None and torch.cuda.is_available() else "cpu"
        t = torch.from_numpy(P).to(dev)
        d.all_reduce(t)
        P = t.cpu().numpy()
    return P


def em_fit_sharded(ctx, x_local, n_total, hefts, means, fcovs, iters, var_floor=1e-10):
    """BASELINE config 5: x_local [F, 3, n_local] is this rank's slice of every fit's samples.
    Returns the fitted parameters (identical on every rank)."""
    d = _dist()
    _, world = _world()
    ctx.em_prepare(x_local, hefts, means, fcovs, iters, var_floor=var_floor, n_total=n_total)
    for _ in range(iters):
        ctx.em_estep()
        if d and world > 1:
            with _on_stream(ctx):
                d.all_reduce(ctx.em_stats_tensor())
        ctx.em_mstep()
    return ctx.em_collect()
