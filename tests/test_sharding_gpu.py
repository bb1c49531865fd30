"""Multi-GPU behind the C ABI (mcsim_cuda_group_*): no torch, no torch.distributed -- what a C++ host of the plugin
gets.  Needs >= 2 GPUs (skipped on a one-GPU box; `gpurun --gpus 2` runs it)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _devices():
    import pcoll_b200
    n = pcoll_b200.cuda_abi.lib().mcsim_cuda_device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    return list(range(n))


def test_group_of_one_is_the_plain_path(ctx, plan, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    g = abi.Group([0])
    try:
        g.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
        ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
        p, keep = abi.make_mc_params(plan, n_rollouts=100001, mode=abi.MODE_PHILOX, early_exit=False)
        a, b = g.mc_run(p, want_counts=True, want_hist=True), ctx.mc_run(p, want_counts=True, want_hist=True)
        np.testing.assert_array_equal(a["counts"], b["counts"])
        assert a["hits"] == b["hits"] and g.collective_stats()[1] == 0
    finally:
        g.close()


def test_group_mc_hits_do_not_depend_on_the_device_count(ctx, plan, scenes):
    """BASELINE config 3's property: rollout ids split over the devices, Philox counters from the global id, one
    ncclAllReduce of the counter block -> hits, histogram, executed steps and per-rollout counts equal the one-device
    run's; with early exit (survivor compaction per device) and in replay mode with injected particles too."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    devs = _devices()
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    g = abi.Group(devs)
    try:
        g.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
        N = 1_000_003
        for ee in (False, True):
            p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=ee, rollout_offset=99)
            one = ctx.mc_run(p, want_counts=True, want_hist=True)
            many = g.mc_run(p, want_counts=True, want_hist=True)
            np.testing.assert_array_equal(many["counts"], one["counts"])
            np.testing.assert_array_equal(many["first_hit_hist"], one["first_hit_hist"])
            assert many["hits"] == one["hits"] and many["executed_steps"] == one["executed_steps"]
        ms, n_coll = g.collective_stats()
        assert n_coll == 2 and ms >= 0.0
        print("%d devices: all-reduce of the counter block %.3f ms" % (len(devs), ms))
        # replay mode: shared applied-command table, injected start poses sliced per device
        T = plan["trajectory"].shape[1]
        Nr = 20001
        draws = np.random.Generator(np.random.PCG64(5)).standard_normal(3 * Nr + 11 * (T - 1))
        ht = abi.host_thread(plan, draws[3 * Nr:])
        p, keep = abi.make_mc_params(plan, n_rollouts=Nr, mode=abi.MODE_REPLAY, applied=ht["applied_noisy"], z0=draws[:3 * Nr], early_exit=False)
        np.testing.assert_array_equal(g.mc_run(p, want_counts=True)["counts"], ctx.mc_run(p, want_counts=True)["counts"])
    finally:
        g.close()


def test_group_score_em_and_gmm_batch_equal_one_device(ctx, plan, scenes):
    import pcoll_b200
    abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
    devs = _devices()
    g = abi.Group(devs)
    try:
        boxes = inp.synthetic_obstacle_field()
        hulls = scenes["pr2test2"]["hulls"]
        ctx.set_scene(boxes, hulls)
        g.set_scene(boxes, hulls)
        tr, od = inp.synthetic_plans(7, 13, 60, boxes, clear=-0.15)      # 13 plans: uneven split
        a, b = g.score_plans(plan, tr, od, 8192), ctx.score_plans(plan, tr, od, 8192)
        np.testing.assert_array_equal(a["hits"], b["hits"])
        assert a["executed_steps"] == b["executed_steps"] and a["pair_tests"] == b["pair_tests"]
        # EM: samples split, statistics all-reduced every iteration
        for K, n in ((3, 100003), (16, 40000)):
            probs = [inp.synthetic_em_problem(60 + f, n, K, centre=(0.4 * f, 1.0, 0.2)) for f in range(2)]
            x = np.stack([q[0] for q in probs]); h = np.stack([q[1] for q in probs])
            m = np.stack([q[2] for q in probs]); c = np.stack([q[3] for q in probs])
            one, many = ctx.em_fit(x, h, m, c, 4), g.em_fit(x, h, m, c, 4)
            np.testing.assert_allclose(many["means"], one["means"], rtol=1e-6, atol=1e-9)
            np.testing.assert_allclose(many["hefts"], one["hefts"], rtol=1e-6)
            np.testing.assert_allclose(many["fcovs"], one["fcovs"], rtol=1e-4, atol=1e-10)
            np.testing.assert_allclose(many["avg_log_p"], one["avg_log_p"], rtol=1e-6)
            # the two ways the statistics travel (fused peer-memory push / ncclAllReduce) give the same fit, and the fused
            # one again on the resident samples (FP64 atomics inside a shard are unordered: rounding-level differences)
            assert g.em_exchange in (abi.EM_EXCHANGE_PEER, abi.EM_EXCHANGE_NCCL)
            g.set_em_exchange(abi.EM_EXCHANGE_NCCL)
            nccl = g.em_fit(x, h, m, c, 4)
            g.set_em_exchange(abi.EM_EXCHANGE_AUTO)
            again = g.em_fit(None, h, m, c, 4)
            for other in (nccl, again):
                np.testing.assert_allclose(other["means"], many["means"], rtol=1e-9, atol=1e-12)
                np.testing.assert_allclose(other["fcovs"], many["fcovs"], rtol=1e-6, atol=1e-12)
                np.testing.assert_allclose(other["hefts"], many["hefts"], rtol=1e-9)
                np.testing.assert_array_equal(other["iters_run"], many["iters_run"])
        print("EM exchange of this group:", "peer memory" if g.em_exchange == abi.EM_EXCHANGE_PEER else "NCCL")
        assert g.collective_stats()[1] == 24
        # batched GMM runs: members split over the devices
        ctx.set_scene(scenes["pr2test2"]["boxes"], hulls)
        g.set_scene(scenes["pr2test2"]["boxes"], hulls)
        a, b = g.gmm_run_batch(plan, 3, 4000, 9, stream0=50), ctx.gmm_run_batch(plan, 3, 4000, 9, stream0=50)
        np.testing.assert_allclose(a["probabilities"], b["probabilities"], rtol=1e-9, atol=1e-12)
    finally:
        g.close()
