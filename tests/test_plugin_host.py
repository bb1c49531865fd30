"""CPU tests of the plugin host (libmcsim.so): C ABI export list, the C++ scene-XML extractor against the
committed scene tables and against the Python restatement (oracle/scene_xml.py), the command grammar of
mcsimplugin.cpp:49-231 on a parse-only module (device = -1), and the loud failure of run* without a device."""
import os
import re
import sys

import numpy as np
import pytest

from conftest import GOLD, ROOT

sys.path.insert(0, os.path.join(ROOT, "oracle"))


def test_capi_exports_every_declared_symbol():
    import pcoll_b200
    lib = pcoll_b200.plugin.lib()
    hdr = open(os.path.join(ROOT, "include", "mcsim.h")).read()
    names = set(re.findall(r"\b(mcsim_\w+)\s*\(", hdr))
    assert len(names) >= 8
    for n in sorted(names):
        assert hasattr(lib, n), "libmcsim.so does not export %s" % n


@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_cxx_scene_extractor_matches_committed_tables(name):
    import pcoll_b200
    g = np.load(os.path.join(GOLD, "scene_%s.npz" % name))
    mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", name + ".env.xml"), device=-1)
    sc = mod.scene()
    np.testing.assert_allclose(sc["boxes"], g["boxes"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(sc["hulls"], g["hulls"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(sc["robot_pose"], g["robot_pose"], atol=1e-12)


@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_cxx_scene_extractor_on_the_reference_files(name):
    """offsetfrom, KinBody-level transforms, external files, lower-case tags: only the reference's own files have them."""
    import pcoll_b200
    import scene_xml
    path = "/root/reference/%s.env.xml" % name
    if not os.path.isfile(path):
        pytest.skip("reference checkout not present")
    ours = pcoll_b200.MCModule(path, device=-1).scene()
    ref = scene_xml.extract_scene(path, [pcoll_b200.plugin.FIXTURE_DIR])
    np.testing.assert_allclose(ours["boxes"], ref["boxes"], rtol=0, atol=1e-14)
    np.testing.assert_allclose(ours["hulls"], ref["hulls"], rtol=0, atol=1e-14)


def test_scene_xml_features(tmp_path):
    """Synthetic env: offsetfrom chain, rotated KinBody, nested file include, comments, mixed tag case."""
    import pcoll_b200
    (tmp_path / "sub.kinbody.xml").write_text(
        "<KinBody name='t'><Body><Geom type='box'><Extents>0.5 0.25 0.1</Extents>"
        "<translation>1 0 0.1</translation></Geom></Body></KinBody>")
    (tmp_path / "robot.fixture.xml").write_text(
        "<Robot><Body><Geom type='box'><Extents>0.3 0.2 0.5</Extents><Translation>0.1 0 0.5</Translation></Geom></Body></Robot>")
    (tmp_path / "env.xml").write_text("""<?xml version="1.0"?>
<!-- comment <KinBody> -->
<Environment>
  <KinBody name="room">
    <Body name="A"><Translation>1 2 0</Translation>
      <Geom type="box"><Extents>1 1 0.1</Extents><Translation>0 0 0.1</Translation></Geom></Body>
    <Body name="B"><offsetfrom> A </offsetfrom><Translation>1 0 0</Translation>
      <Geom type="box"><Extents>0.2 0.1 0.1</Extents><Translation>0 1 0.1</Translation>
        <RotationAxis>0 0 1 30</RotationAxis></Geom></Body>
  </KinBody>
  <KinBody name="inc" file="data/sub.kinbody.xml"><RotationAxis>0 0 1 90</RotationAxis><Translation>5 5 0.4</Translation></KinBody>
  <Robot name="r" file="robots/robot.zae"><translation>-1 -2 0.05</translation></Robot>
</Environment>""")
    sc = pcoll_b200.MCModule(str(tmp_path / "env.xml"), device=-1, fixture_dirs=(str(tmp_path),)).scene()
    c30, s30 = np.cos(np.pi / 6), np.sin(np.pi / 6)
    expect = np.array([
        [1, 2, 1, 0, 1, 1, 0.0, 0.2],
        [2, 3, c30, s30, 0.2, 0.1, 0.0, 0.2],
        [5, 6, 0, 1, 0.5, 0.25, 0.4, 0.6],      # (1, 0) rotated by 90 deg about the body origin, then translated
    ])
    np.testing.assert_allclose(sc["boxes"], expect, atol=1e-12)
    np.testing.assert_allclose(sc["hulls"], [[0.1, 0, 1, 0, 0.3, 0.2, 0.05, 1.05]], atol=1e-12)
    np.testing.assert_allclose(sc["robot_pose"], [-1, -2, 0])


def test_scene_xml_errors(tmp_path):
    import pcoll_b200
    with pytest.raises(pcoll_b200.McsimError, match="cannot open"):
        pcoll_b200.MCModule(str(tmp_path / "missing.xml"), device=-1)
    (tmp_path / "bad.xml").write_text("<Environment><KinBody><Body><Geom type='sphere'/></Body></KinBody></Environment>")
    with pytest.raises(pcoll_b200.McsimError, match="only box geoms"):
        pcoll_b200.MCModule(str(tmp_path / "bad.xml"), device=-1)
    (tmp_path / "tilt.xml").write_text("<Environment><KinBody><Body><Geom type='box'><Extents>1 1 1</Extents>"
                                       "<RotationAxis>1 0 0 20</RotationAxis></Geom></Body></KinBody></Environment>")
    with pytest.raises(pcoll_b200.McsimError, match="not about z"):
        pcoll_b200.MCModule(str(tmp_path / "tilt.xml"), device=-1)
    (tmp_path / "open.xml").write_text("<Environment><KinBody>")
    with pytest.raises(pcoll_b200.McsimError, match="XML parse error"):
        pcoll_b200.MCModule(str(tmp_path / "open.xml"), device=-1)


def test_command_grammar_parse_only(plan, scenes):
    import pcoll_b200
    mod = pcoll_b200.MCModule(scenes["pr2test2"], device=-1)
    assert mod.SendCommand("MyCommand [1,2,3]") == "output"          # mcsimplugin.cpp:225-231
    assert mod.SendCommand("mycommand x") == "output"                # OpenRAVE matches names case-insensitively
    h = mod.SendCommand("help")
    for name in ("MyCommand", "ArmaCommand", "setAlphas", "setQ", "setNumLandmarks", "setLandmarks", "setNumParticles",
                 "setInitialCovariance", "setPathLength", "setTrajectory", "setOdometry", "runSimulation", "setNumGaussians",
                 "runGMMEstimation", "setNumGMMSamples"):                # the 15 RegisterCommand names, :13-44
        assert re.search(r"^%s : " % name, h, re.M), name
    pcoll_b200.push_parameters(mod, plan, num_particles=1000, num_gaussians=3)   # whole MCSimulation.py:156-207 sequence
    # ordering contract: setLandmarks reads getNumLandmarks(), setTrajectory / setOdometry read getPathLength()
    with pytest.raises(pcoll_b200.McsimError, match="expected 168 numbers"):
        mod.SendCommand("setTrajectory 1 2 3")
    with pytest.raises(pcoll_b200.McsimError, match="unknown command"):
        mod.SendCommand("noSuchCommand")
    with pytest.raises(pcoll_b200.McsimError, match="no CPU fallback"):
        mod.SendCommand("runSimulation")
    with pytest.raises(pcoll_b200.McsimError, match="no CPU fallback"):
        mod.SendCommand("runGMMEstimation")
    with pytest.raises(pcoll_b200.McsimError):
        mod.SendCommand("setMode sometimes")
    assert mod.SendCommand("setMode deterministic") == ""
    # method-variant switches (round 2) parse without a device too
    assert mod.SendCommand("setFilter deviation") == "" and mod.SendCommand("setFilter state") == ""
    assert mod.SendCommand("setRefit em 3") == "" and mod.SendCommand("setRefit moments") == ""
    with pytest.raises(pcoll_b200.McsimError, match="state \\| deviation"):
        mod.SendCommand("setFilter both")
    for name in ("setFilter", "setRefit", "setConstraints", "runLinearEstimation", "setDevices", "runGMMEstimationBatch"):
        assert re.search(r"^%s : " % name, h, re.M), name
    mod.close()


def test_run_ids_of_shards_and_runs_are_disjoint(plan, scenes):
    """Two shards (setRolloutOffset 0 / N, common seed) x three runs: the run number lives in the Philox KEY, the shard
    offset in the COUNTER, so the W x R blocks of (key, rollout id) are pairwise disjoint -- in round 1 run r + 1 of
    shard s reused the ids of run r of shard s + 1.  describeRun reports what the next runs will use (no GPU needed)."""
    import pcoll_b200
    N = 1000
    blocks = []
    gmm = []
    for shard in range(2):
        mod = pcoll_b200.MCModule(scenes["pr2test2"], device=-1)
        pcoll_b200.push_parameters(mod, plan, num_particles=N, num_gaussians=3)
        mod.SendCommand("setSeed 12345")
        mod.SendCommand("setRolloutOffset %d" % (shard * N))
        for run in range(3):
            f = mod.SendCommand("describeRun %d" % run).split()
            assert f[0] == "mc" and f[4] == "gmm"
            blocks.append((int(f[1]), int(f[2]), int(f[3])))
            gmm.append((int(f[5]), int(f[6])))
        mod.close()
    assert blocks[0] == (12345, 0, N)                      # run 0 of shard 0: the seed itself, ids from 0
    keys = {b[0] for b in blocks}
    assert len(keys) == 3                                  # one key per run number, the same on both shards
    seen = set()
    for key, first, count in blocks:
        for other_key, lo, hi in seen:
            assert key != other_key or first + count <= lo or hi <= first, "overlapping (key, id) blocks"
        seen.add((key, first, first + count))
    assert len(set(gmm)) == 6                              # GMM runs: (key salted by the shard, stream = run number)


def test_set_devices_needs_a_device_and_is_parsed(scenes):
    import pcoll_b200
    mod = pcoll_b200.MCModule(scenes["pr2test2"], device=-1)
    with pytest.raises(pcoll_b200.McsimError, match="ordinals or 'all'"):
        mod.SendCommand("setDevices zero")
    with pytest.raises(pcoll_b200.McsimError):     # no CUDA device in the CPU test environment: loud, no fallback
        mod.SendCommand("setDevices 0,1")
    assert mod.SendCommand("getDevices") == ""
    mod.close()


def test_arma_command_writes_arma_ascii(tmp_path, scenes, monkeypatch):
    import pcoll_b200
    monkeypatch.chdir(tmp_path)
    mod = pcoll_b200.MCModule(scenes["pr2test2"], device=-1)
    mod.SendCommand("setSeed 7")
    mod.SendCommand("ArmaCommand")
    for f in ("A_mat.txt", "B_mat.txt"):
        lines = (tmp_path / f).read_text().splitlines()
        assert lines[0] == "ARMA_MAT_TXT_FN008" and lines[1] == "4 4" and len(lines) == 6
    A = np.loadtxt(tmp_path / "A_mat.txt", skiprows=2)
    B = np.loadtxt(tmp_path / "B_mat.txt", skiprows=2)
    np.testing.assert_allclose(A[0, :3], A[1, :3] + A[3, :3], rtol=1e-13)   # row 0 = row 1 + row 3
    assert (A[:, 3] == 0).all()
    np.testing.assert_array_equal(B, np.diag(np.diag(A)))


def test_create_on_a_missing_device_is_loud(scenes):
    import pcoll_b200
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pcoll_b200.McsimError, match="no CPU fallback"):
        pcoll_b200.MCModule(scenes["pr2test2"], device=0)


def test_openrave_adapter_compiles_against_the_stand_in_headers():
    """OpenRAVE and Boost are not in this image: csrc/openrave_adapter.cpp is syntax-checked against tests/stub_openrave."""
    import subprocess
    csrc = os.path.join(ROOT, "probability-of-collision-for-safe-planning_b200", "csrc")
    r = subprocess.run(["make", "-C", csrc, "openrave-syntax"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
