"""GPU tests of the drop-in boundary: the MCSimulation.py command sequence against libmcsim.so, checked against the
reference estimator's golden vectors (deterministic mode) and its run-to-run distribution (shared / philox modes)."""
import json
import os

import numpy as np
import pytest

from conftest import GOLD, ROOT, golden

pytestmark = pytest.mark.gpu


def _module(plan, scene_xml_name, n, **kw):
    import pcoll_b200
    mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", scene_xml_name + ".env.xml"), device=0)
    pcoll_b200.push_parameters(mod, plan, num_particles=n, **kw)
    return mod


@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_run_simulation_deterministic_matches_reference_text(plan, name, tmp_path):
    g = golden("mc_replay_%s.npz" % name)
    N = g["counts"].shape[0]
    mod = _module(plan, name, N)
    f = tmp_path / "draws.f64"
    g["draws"].astype("<f8").tofile(f)
    mod.SendCommand("setMode deterministic")
    mod.SendCommand("setDrawsFile %s" % f)
    near = (np.abs(g["clearance"].astype(np.float64)) < 1e-4).any(axis=0)
    for ee in (0, 1):
        mod.SendCommand("setEarlyExit %d" % ee)
        text = mod.SendCommand("runSimulation")
        hits = int(mod.SendCommand("getStats").split()[0])
        # bit-exact hit count unless a near-contact rollout flips (none does on the committed draws)
        assert abs(hits - int((g["counts"] > 0).sum())) <= int(near.sum())
        if hits == int((g["counts"] > 0).sum()):
            assert text == "%g" % float(g["P"])          # `sout << collprop`: default ostream formatting
        hist = np.array(mod.SendCommand("getFirstHitHistogram").split(), dtype=np.int64)
        assert hist.sum() == hits and hist.shape[0] == plan["trajectory"].shape[1]
    mod.close()


@pytest.mark.parametrize("K", [1, 3])
def test_run_gmm_estimation_deterministic_matches_reference(plan, K, tmp_path):
    g = golden("gmm_replay_pr2test2_K%d.npz" % K)
    N = int(g["n_samples"])
    mod = _module(plan, "pr2test2", N, num_gaussians=K)
    g["draws"].astype("<f8").tofile(tmp_path / "d.f64")
    g["gmm_counts"].astype("<i4").tofile(tmp_path / "c.i32")
    mod.SendCommand("setMode deterministic")
    mod.SendCommand("setDrawsFile %s" % (tmp_path / "d.f64"))
    mod.SendCommand("setGMMCountsFile %s" % (tmp_path / "c.i32"))
    text = mod.SendCommand("runGMMEstimation")
    assert text == "%g" % float(g["P"])
    prob = np.array(mod.SendCommand("getWaypointProbabilities").split(), dtype=np.float64)
    np.testing.assert_array_equal(prob, g["probabilities"])
    post = np.array(mod.SendCommand("getGMM").split(), dtype=np.float64).reshape(K, 13)
    np.testing.assert_allclose(post[:, :4], g["gmm_post"][:, :4], rtol=1e-4, atol=1e-4)
    mod.close()


def test_shared_mode_reproduces_the_reference_run_to_run_law(plan):
    """Default mode: every runSimulation is one draw of the reference's estimator (one shared real robot per run), so
    the returned values scatter like the reference's own runs (tests/golden/mc_dist_pr2test2.json)."""
    dist = json.load(open(os.path.join(GOLD, "mc_dist_pr2test2.json")))
    mod = _module(plan, "pr2test2", dist["particles"])
    mod.SendCommand("setSeed 12345")
    ours = np.array([float(mod.SendCommand("runSimulation")) for _ in range(200)])
    ref = np.array(dist["P"])
    half = 2.576 * np.sqrt(ours.var(ddof=1) / len(ours) + ref.var(ddof=1) / len(ref))
    print("shared mode: mean %.4f std %.4f | reference mean %.4f std %.4f | 99%% half-width %.4f" %
          (ours.mean(), ours.std(ddof=1), ref.mean(), ref.std(ddof=1), half))
    assert abs(ours.mean() - ref.mean()) < half
    assert 0.5 * ref.std(ddof=1) < ours.std(ddof=1) < 2.0 * ref.std(ddof=1)
    assert len(set(ours)) > 20                       # successive runs use fresh draws
    mod.SendCommand("setSeed 12345")                 # same seed: same sequence of results
    again = [float(mod.SendCommand("runSimulation")) for _ in range(3)]
    assert again == list(ours[:3])
    # philox mode: expectation of the above
    mod.SendCommand("setMode philox")
    mod.SendCommand("setNumParticles 1000000")
    p = float(mod.SendCommand("runSimulation"))
    assert abs(p - ref.mean()) < 2.576 * ref.std(ddof=1) / np.sqrt(len(ref))
    mod.close()


def test_gmm_default_mode_runs_and_scatter(plan):
    mod = _module(plan, "pr2test2", 10000, num_gaussians=3)
    mod.SendCommand("setSeed 99")
    vals = [float(mod.SendCommand("runGMMEstimation")) for _ in range(5)]
    assert all(0.0 <= v <= 1.0 for v in vals) and len(set(vals)) > 1
    mod.close()


def test_run_without_parameters_is_an_error():
    import pcoll_b200
    mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=0)
    with pytest.raises(pcoll_b200.McsimError, match="setAlphas"):
        mod.SendCommand("runSimulation")
    mod.close()


def test_py3_driver_twin_mc_and_gmm(tmp_path):
    """tools/mc_simulation.py: the reference driver's loop and report files (MCSimulation.py:100-270)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("mc_simulation", os.path.join(ROOT, "tools", "mc_simulation.py"))
    drv = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(drv)
    for opt, prefix in (("MC", "simReport_"), ("GMM", "GMMsimReport_")):
        out = drv.main([opt, "--num-simulations", "3", "--num-particles", "2000", "--outdir", str(tmp_path), "--seed", "5"])
        assert len(out["proportions"]) == 3 and all(0.5 < p <= 1.0 for p in out["proportions"])
        text = open(out["report"]).read()
        assert os.path.basename(out["report"]).startswith(prefix)
        for key in ("Environment: ", "Num Landmarks: 8", "NumSimulations: 3", "Average Prob Collision: ", "Trajectory: ", "Odometry: "):
            assert key in text
        ck = open(out["checkpoint"]).read()
        assert ck.count("collProp: ") == 3 and "Simulation: 2" in ck


def test_run_gmm_estimation_batch_command_equals_the_loop():
    """runGMMEstimationBatch <n> == the next n runGMMEstimation calls (MCSimulation.py:226-246 in one launch chain)."""
    import pcoll_b200
    plan = pcoll_b200.load_plan_json(os.path.join(GOLD, "plan_committed.json"))
    outs = []
    for batched in (False, True):
        mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=0)
        pcoll_b200.push_parameters(mod, plan, num_particles=4000, num_gaussians=3, num_gmm_samples=4000)
        mod.SendCommand("setSeed 9")
        if batched:
            outs.append([float(v) for v in mod.SendCommand("runGMMEstimationBatch 5").split()])
        else:
            outs.append([float(mod.SendCommand("runGMMEstimation")) for _ in range(5)])
        mod.close()
    np.testing.assert_allclose(outs[0], outs[1], rtol=1e-5)   # text round trip: 6 significant digits
    assert len(set(outs[1])) > 1


def test_two_modules_with_different_scenes_share_one_device(plan):
    """Two MCModules (pr2test2 and pr2custom) on cuda:0, run alternately: the __constant__ scene / plan tables are per
    device, each module must get its own back at every launch (per-device ownership registry in mcsim_cuda.cu).
    Every alternating result must equal what the module returns when it has the device to itself."""
    import pcoll_b200
    alone = {}
    for name in ("pr2test2", "pr2custom"):
        mod = _module(plan, name, 50000, num_gaussians=2, num_gmm_samples=3000)
        mod.SendCommand("setSeed 77")
        mod.SendCommand("setMode philox")
        alone[name] = [mod.SendCommand("runSimulation"), mod.SendCommand("runSimulation"), mod.SendCommand("runGMMEstimation")]
        mod.close()
    assert alone["pr2test2"][0] != alone["pr2custom"][0]
    mods = {name: _module(plan, name, 50000, num_gaussians=2, num_gmm_samples=3000) for name in alone}
    for m in mods.values():
        m.SendCommand("setSeed 77")
        m.SendCommand("setMode philox")
    got = {name: [] for name in alone}
    for cmd in ("runSimulation", "runSimulation", "runGMMEstimation"):
        for name in ("pr2test2", "pr2custom"):
            got[name].append(mods[name].SendCommand(cmd))
    for m in mods.values():
        m.close()
    assert got == alone


def test_long_reply_is_not_resent():
    """A reply longer than the 64 KB buffer is fetched with mcsim_last_output, not by running the command again."""
    import pcoll_b200
    plan = pcoll_b200.load_plan_json(os.path.join(GOLD, "plan_committed.json"))
    short = dict(plan)
    short["trajectory"] = plan["trajectory"][:, :3]
    short["odometry"] = plan["odometry"][:, :2]
    mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=0)
    pcoll_b200.push_parameters(mod, short, num_particles=64, num_gaussians=1, num_gmm_samples=64)
    mod.SendCommand("setSeed 3")
    n = 9000
    reply = mod.SendCommand("runGMMEstimationBatch %d" % n)
    assert len(reply) > (1 << 16) and len(reply.split()) == n
    assert mod.SendCommand("describeRun").split()[-1] == str(n)      # the run counter advanced by n, not 2 n
    mod.close()


def _n_gpus():
    import pcoll_b200
    return pcoll_b200.cuda_abi.lib().mcsim_cuda_device_count()


def test_set_devices_single_device_equals_default(plan):
    import pcoll_b200
    outs = []
    for use in (False, True):
        mod = _module(plan, "pr2test2", 200000)
        if use:
            mod.SendCommand("setDevices 0")
            assert mod.SendCommand("getDevices") == "0"
        mod.SendCommand("setSeed 5")
        mod.SendCommand("setMode philox")
        outs.append((mod.SendCommand("runSimulation"), mod.SendCommand("getStats").split()[0], mod.SendCommand("getFirstHitHistogram")))
        mod.close()
    assert outs[0] == outs[1]


def test_multi_device_run_simulation_equals_one_device(plan):
    """setDevices all: the C++ plugin host drives every GPU of the box through the C ABI alone (no Python-side
    distribution): rollout ids split over the devices, counters all-reduced by the NCCL communicator the handle owns;
    hits, histogram and the returned text equal the one-device run (BASELINE config 3's invariance)."""
    import pcoll_b200
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    outs = []
    for devs in ("0", "all"):
        mod = _module(plan, "pr2test2", 3000001, num_gaussians=3, num_gmm_samples=5000)
        mod.SendCommand("setDevices " + devs)
        mod.SendCommand("setSeed 5")
        mod.SendCommand("setMode philox")
        res = []
        for ee in (1, 0):
            mod.SendCommand("setEarlyExit %d" % ee)
            res.append((mod.SendCommand("runSimulation"), mod.SendCommand("getStats").split()[:3], mod.SendCommand("getFirstHitHistogram")))
        res.append(mod.SendCommand("runGMMEstimationBatch 70"))
        mod.SendCommand("setMode shared")
        res.append(mod.SendCommand("runSimulation"))
        outs.append(res)
        mod.close()
    assert outs[0][0] == outs[1][0] and outs[0][1] == outs[1][1]
    np.testing.assert_allclose([float(v) for v in outs[0][2].split()], [float(v) for v in outs[1][2].split()], rtol=1e-5)
    assert outs[0][3] == outs[1][3]
