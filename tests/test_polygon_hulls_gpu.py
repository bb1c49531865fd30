"""Convex-polygon robot hulls (scene fidelity beyond boxes) on the GPU, against the CPU oracle
(oracle/sat_cpu.h, whose polygon routine is pinned by the hand-derived cases of tests/golden/sat_cases.json)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _octagon(r, z=(0.05, 0.35), offset=(0.0, 0.0), phase=22.5):
    v = [(offset[0] + r * np.cos(np.radians(phase + 45 * k)), offset[1] + r * np.sin(np.radians(phase + 45 * k))) for k in range(8)]
    return [8, z[0], z[1]] + [c for xy in v for c in xy]


def _robot():
    """PR2-like stand-in: an octagonal base (circumradius 0.362 m: inscribed in the fixture's 0.334 m half-width square)
    and the fixture's upper-body box."""
    base = _octagon(0.334 / np.cos(np.radians(22.5)))
    upper = np.array([[-0.05, 0.0, 1.0, 0.0, 0.28, 0.33, 0.35, 1.4]])
    return upper, np.array([base])


def test_device_polygon_clearance_matches_the_oracle(ctx, plan, scenes):
    import oracle_ref
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    sc = scenes["pr2test2"]
    hulls, polys = _robot()
    ctx.set_scene(sc["boxes"], hulls, polys)
    N = 4000
    one = dict(plan)
    one["trajectory"] = np.array([[0.3], [0.6], [1.0]])       # next to the MidWall door
    one["odometry"] = np.zeros((3, 0))
    one["cov0"] = np.diag([0.09, 0.09, 1.0])
    z0 = np.random.Generator(np.random.PCG64(3)).standard_normal(3 * N)
    p, keep = abi.make_mc_params(one, n_rollouts=N, mode=abi.MODE_REPLAY, z0=z0, early_exit=False)
    out = ctx.mc_run(p, want_counts=True, audit=True)
    rp, keep2 = oracle_ref.make_params(plan, sc["boxes"], hulls, poly_hulls=polys)
    pose = (C.c_double * 3)()
    ref = np.empty(N)
    for i in range(N):
        pose[0], pose[1], pose[2] = (float(v) for v in out["poses"][0, :, i])
        ref[i] = oracle_ref.lib().ref_clearance(C.byref(rp), pose)
    np.testing.assert_allclose(out["clearance"][0], ref, atol=3e-5)
    far = np.abs(ref) > 1e-4
    np.testing.assert_array_equal(out["counts"][far] > 0, ref[far] <= 0)
    assert 0.1 < (ref <= 0).mean() < 0.9
    ctx.set_scene(sc["boxes"], sc["hulls"])


def test_replay_with_a_polygon_robot_matches_the_reference_run(ctx, plan, scenes):
    """The reference's own runSimulation() with CheckCollision served by the polygon-capable CPU SAT, against the device
    replay kernel on the same injected draws: per-rollout counts bit-exact away from contact."""
    import oracle_ref
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    sc = scenes["pr2test2"]
    hulls, polys = _robot()
    T = plan["trajectory"].shape[1]
    N = 600
    draws = np.random.Generator(np.random.PCG64(21)).standard_normal(3 * N + 11 * (T - 1))
    ref = oracle_ref.mc_run(plan, sc["boxes"], hulls, N, draws=draws, poly_hulls=polys)
    ctx.set_scene(sc["boxes"], hulls, polys)
    ht = abi.host_thread(plan, draws[3 * N:])
    p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_REPLAY, applied=ht["applied_noisy"], z0=draws[:3 * N], early_exit=False)
    out = ctx.mc_run(p, want_counts=True)
    near = (np.abs(ref["clearance"]) < 1e-4).any(axis=0)
    np.testing.assert_array_equal(out["counts"][~near], ref["counts"][~near])
    assert near.mean() < 0.1 and 0 < ref["P"] <= 1
    # the octagon is inscribed in the fixture's square base: it can only collide where the square does
    ctx.set_scene(sc["boxes"], sc["hulls"])
    sq = ctx.mc_run(p, want_counts=True)
    assert np.all(out["counts"][~near] <= sq["counts"][~near]) and out["hits"] <= sq["hits"]
    print("polygon robot: P %.4f (reference %.4f), box fixture P %.4f" % (out["P"], ref["P"], sq["P"]))


def test_philox_and_gmm_paths_with_polygon_hulls(ctx, plan, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    sc = scenes["pr2test2"]
    hulls, polys = _robot()
    N = 100001
    ctx.set_scene(sc["boxes"], sc["hulls"])
    p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False, precise=True)
    square = ctx.mc_run(p, want_counts=True)
    ctx.set_scene(sc["boxes"], hulls, polys)
    p2, keep2 = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False)     # routed to the scalar kernel
    octo = ctx.mc_run(p2, want_counts=True)
    assert np.all(octo["counts"] <= square["counts"]) and 0 < octo["hits"] <= square["hits"]
    ee = ctx.mc_run(abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=True)[0])
    assert ee["hits"] == octo["hits"]
    g = ctx.gmm_run(plan, 3, 10000, mode=abi.MODE_PHILOX, stream_id=2)
    ctx.set_scene(sc["boxes"], sc["hulls"])
    gs = ctx.gmm_run(plan, 3, 10000, mode=abi.MODE_PHILOX, stream_id=2)
    assert 0.0 < g["P"] <= gs["P"] + 0.05
    print("Philox: octagon P %.4f <= square P %.4f; GMM %.4f vs %.4f" % (octo["P"], square["P"], g["P"], gs["P"]))
    ctx.set_scene(sc["boxes"], hulls, polys)
    with pytest.raises(abi.McsimCudaError, match="polygon"):
        ctx.score_plans(plan, plan["trajectory"][None], plan["odometry"][None], 1000)
    with pytest.raises(abi.McsimCudaError):
        ctx.set_scene(sc["boxes"], hulls, [[4, 0, 1, 0, 0, 1, 0, 0.2, 0.2, 0, 1] + [0] * 8])      # not convex
    ctx.set_scene(sc["boxes"], sc["hulls"])


def test_robot_fixture_with_a_prism_geom_through_the_plugin(plan, tmp_path):
    """A robot file with <Geom type="prism"> (a supplied hull set beyond boxes) loaded by the C++ scene extractor."""
    import os
    import pcoll_b200
    from conftest import ROOT
    fx = tmp_path / "fixtures"
    fx.mkdir()
    verts = " ".join("%.6f %.6f" % (0.3615 * np.cos(np.radians(22.5 + 45 * k)), 0.3615 * np.sin(np.radians(22.5 + 45 * k))) for k in range(8))
    (fx / "pr2-beta-static.fixture.xml").write_text(
        '<Robot name="PR2"><Body name="base"><Geom type="prism"><Vertices>%s</Vertices><Height>0 0.3</Height></Geom></Body>'
        '<Body name="upper"><Geom type="box"><Extents>0.28 0.33 0.525</Extents><Translation>-0.05 0 0.825</Translation></Geom></Body></Robot>' % verts)
    env = os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml")
    mod = pcoll_b200.MCModule(env, device=0, fixture_dirs=(str(fx), pcoll_b200.plugin.FIXTURE_DIR))
    scn = mod.scene()
    assert scn["poly_hulls"].shape == (1, 19) and scn["hulls"].shape == (1, 8)
    assert scn["poly_hulls"][0, 0] == 8 and abs(scn["poly_hulls"][0, 1] - 0.05) < 1e-12      # z offset of the <Robot> applied
    pcoll_b200.push_parameters(mod, plan, num_particles=20000)
    mod.SendCommand("setSeed 4")
    mod.SendCommand("setMode philox")
    P_oct = float(mod.SendCommand("runSimulation"))
    mod.SendCommand("setRobotPolygons 0")
    mod.SendCommand("setSeed 4")
    P_upper_only = float(mod.SendCommand("runSimulation"))
    assert 0.0 < P_upper_only <= P_oct <= 1.0
    mod.close()


def test_collada_model_of_the_box_fixture_gives_the_fixture_answers(plan, tmp_path):
    """The robot as a COLLADA archive (what the reference's env files name, pr2test2.env.xml:120): a .zae holding the box
    fixture's two prisms as meshes, reduced to footprints by csrc/collada_hulls.cpp, against the box fixture itself -- the
    reference's default `shared` mode draws one real-robot thread per run from the seed, so both modules see the same
    particles; the polygon SAT of a rectangle is the box SAT (tests/test_sat_known_answers.py), so the estimates agree up
    to rollouts at rounding distance from contact."""
    import os
    import pcoll_b200
    from conftest import ROOT
    from test_collada_hulls import _two_box_robot, _zae
    env = os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml")
    path = str(tmp_path / "pr2-like.zae")
    _zae(path, _two_box_robot())
    box = pcoll_b200.MCModule(env, device=0)
    dae = pcoll_b200.MCModule(env, device=0)
    assert dae.SendCommand("setRobotModel %s 0.05 0.35 1.4" % path) == "2"
    assert dae.scene()["hulls"].shape[0] == 0 and dae.scene()["poly_hulls"].shape == (2, 19)
    N = 20000
    for mod in (box, dae):
        pcoll_b200.push_parameters(mod, plan, num_particles=N)
        mod.SendCommand("setSeed 9")
    Pb = [float(box.SendCommand("runSimulation")) for _ in range(3)]
    Pd = [float(dae.SendCommand("runSimulation")) for _ in range(3)]
    assert 0.0 < min(Pb) and max(Pb) <= 1.0
    np.testing.assert_allclose(Pd, Pb, atol=3.0 / N)      # a handful of near-contact rollouts may flip between the two SAT forms
    # the same model as bounding-rectangle box hulls: the fixture's own hull rows, so the packed Philox kernel gives the
    # fixture's answer exactly
    assert dae.SendCommand("setRobotModel %s boxes 0.05 0.35 1.4" % path) == "2"
    for mod in (box, dae):
        mod.SendCommand("setMode philox")
        mod.SendCommand("setSeed 9")
    assert [box.SendCommand("runSimulation") for _ in range(2)] == [dae.SendCommand("runSimulation") for _ in range(2)]
    box.close(); dae.close()


def test_broad_phase_with_polygon_hulls_is_exact(ctx, plan, scenes):
    """The candidate lists of the exact broad phase now cover polygon hulls too (hull index: box hulls, then polygon hulls;
    reach = the largest vertex radius): per-rollout counts equal testing every live pair, in replay mode and in Philox mode
    (both run the scalar kernel when the robot has polygon hulls), with a fraction of the pair tests; the GMM estimate of the
    scalar sample kernel is unchanged too."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    sc = scenes["pr2test2"]
    hulls, polys = _robot()
    ctx.set_scene(sc["boxes"], hulls, polys)
    T = plan["trajectory"].shape[1]
    N = 30001
    outs = {}
    for ex in (False, True):
        p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False, exhaustive=ex)
        outs[ex] = ctx.mc_run(p, want_counts=True, want_hist=True)
    np.testing.assert_array_equal(outs[False]["counts"], outs[True]["counts"])
    np.testing.assert_array_equal(outs[False]["first_hit_hist"], outs[True]["first_hit_hist"])
    assert outs[True]["pair_tests"] == N * T * 63 and outs[False]["pair_tests"] < 0.2 * outs[True]["pair_tests"]
    draws = np.random.Generator(np.random.PCG64(33)).standard_normal(3 * N + 11 * (T - 1))
    ht = abi.host_thread(plan, draws[3 * N:])
    for ex in (False, True):
        p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_REPLAY, applied=ht["applied_noisy"], z0=draws[:3 * N], early_exit=False, exhaustive=ex)
        outs[ex] = ctx.mc_run(p, want_counts=True)
    np.testing.assert_array_equal(outs[False]["counts"], outs[True]["counts"])
    assert outs[False]["pair_tests"] < 0.2 * outs[True]["pair_tests"]
    # only polygon hulls (no box hull at all): the lists still work
    ctx.set_scene(sc["boxes"], np.zeros((0, 8)), polys)
    for ex in (False, True):
        p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=True, exhaustive=ex)
        outs[ex] = ctx.mc_run(p, want_counts=True)
    np.testing.assert_array_equal(outs[False]["counts"], outs[True]["counts"])
    assert 0 < outs[False]["hits"] < N
    ctx.set_scene(sc["boxes"], hulls, polys)
    a = ctx.gmm_run(plan, 3, 6000, mode=abi.MODE_PHILOX, stream_id=5)
    b = ctx.gmm_run(plan, 3, 6000, mode=abi.MODE_PHILOX, stream_id=5, exhaustive=True)
    np.testing.assert_array_equal(a["probabilities"], b["probabilities"])
    assert a["pair_tests"] < 0.3 * b["pair_tests"]
    ctx.set_scene(sc["boxes"], sc["hulls"])
