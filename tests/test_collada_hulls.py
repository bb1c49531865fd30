"""Robot hull set from a COLLADA model (csrc/collada_hulls.cpp; SURVEY.md 8f-3: the reference's scenes name their robot as
`robots/pr2-beta-static.zae`, pr2test2.env.xml:120, which is not in its repository).  Synthetic models written here:
meshes with known footprints, node transforms, units, both up axes, zipped into a .zae the way OpenRAVE ships them;
the hulls the module reports must be the footprints, band by band.  Parse-only modules (device = -1): no GPU."""
import io
import os
import zipfile

import numpy as np
import pytest

from conftest import ROOT


def _box_mesh(gid, hx, hy, hz, prim="triangles"):
    """Axis-aligned box centred on the origin: 8 vertices, 12 triangles (or 6 quads as a <polylist>)."""
    v = [(sx * hx, sy * hy, sz * hz) for sz in (-1, 1) for sy in (-1, 1) for sx in (-1, 1)]
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    floats = " ".join("%.17g" % c for p in v for c in p)
    if prim == "triangles":
        idx = " ".join("%d %d %d %d %d %d" % (a, b, c, a, c, d) for a, b, c, d in quads)
        body = '<triangles count="12"><input semantic="VERTEX" source="#%s-vtx" offset="0"/><p>%s</p></triangles>' % (gid, idx)
    else:   # polylist with a second (unused) input at offset 1: the index stride is 2
        idx = " ".join("%d 0" % i for q in quads for i in q)
        body = ('<polylist count="6"><input semantic="VERTEX" source="#%s-vtx" offset="0"/><input semantic="NORMAL" source="#%s-nrm" offset="1"/>'
                '<vcount>4 4 4 4 4 4</vcount><p>%s</p></polylist>' % (gid, gid, idx))
    return ('<geometry id="%s"><mesh><source id="%s-pos"><float_array id="%s-arr" count="24">%s</float_array>'
            '<technique_common><accessor source="#%s-arr" count="8" stride="3"/></technique_common></source>'
            '<vertices id="%s-vtx"><input semantic="POSITION" source="#%s-pos"/></vertices>%s</mesh></geometry>'
            % (gid, gid, gid, floats, gid, gid, gid, body))


def _prism_mesh(gid, n, r, hz):
    """Regular n-gon prism, as triangles."""
    ang = [2 * np.pi * k / n for k in range(n)]
    v = [(r * np.cos(a), r * np.sin(a), z) for z in (-hz, hz) for a in ang]
    tris = []
    for k in range(n):
        a, b = k, (k + 1) % n
        tris += [(a, b, n + b), (a, n + b, n + a)]
    tris += [(0, k, k + 1) for k in range(1, n - 1)] + [(n, n + k + 1, n + k) for k in range(1, n - 1)]
    floats = " ".join("%.17g" % c for p in v for c in p)
    idx = " ".join("%d %d %d" % t for t in tris)
    return ('<geometry id="%s"><mesh><source id="%s-pos"><float_array count="%d">%s</float_array></source>'
            '<vertices id="%s-vtx"><input semantic="POSITION" source="#%s-pos"/></vertices>'
            '<triangles count="%d"><input semantic="VERTEX" source="#%s-vtx" offset="0"/><p>%s</p></triangles></mesh></geometry>'
            % (gid, gid, 6 * n, floats, gid, gid, len(tris), gid, idx))


def _dae(geoms, scene_nodes, lib_nodes="", meter=1.0, up="Z_UP"):
    return ('<?xml version="1.0" encoding="utf-8"?>\n<COLLADA xmlns="http://www.collada.org/2008/03/COLLADASchema" version="1.5.0">'
            '<asset><unit name="u" meter="%r"/><up_axis>%s</up_axis></asset>'
            '<library_geometries>%s</library_geometries><library_nodes>%s</library_nodes>'
            '<library_visual_scenes><visual_scene id="decoy"><node id="d0"/></visual_scene><visual_scene id="vs">%s</visual_scene></library_visual_scenes>'
            '<scene><instance_visual_scene url="#vs"/></scene></COLLADA>' % (meter, up, geoms, lib_nodes, scene_nodes))


def _zae(path, dae_text, name="model.dae", deflate=True):
    with zipfile.ZipFile(path, "w", zipfile.ZIP_DEFLATED if deflate else zipfile.ZIP_STORED) as z:
        z.writestr("manifest.xml", '<?xml version="1.0"?><dae_root>./%s</dae_root>' % name)
        z.writestr("readme.txt", "not the model")
        z.writestr(name, dae_text)


def _two_box_robot(scale=100.0):
    """The box fixture's robot (data/fixtures/pr2-beta-static.fixture.xml) as a COLLADA model in centimetres: base
    0.668 x 0.668 m over z in [0.05, 0.35], upper body 0.56 x 0.66 m centred at x = -0.05 over [0.35, 1.4].  The base hangs
    under two rotations that cancel, the body comes in through <library_nodes> / <instance_node> and a <matrix>."""
    s = scale
    geoms = _box_mesh("base", 0.334 * s, 0.334 * s, 0.15 * s) + _box_mesh("body", 0.28 * s, 0.33 * s, 0.525 * s, prim="polylist")
    lib = ('<node id="bodynode"><matrix>1 0 0 %r  0 1 0 0  0 0 1 %r  0 0 0 1</matrix><instance_geometry url="#body"/></node>' % (-0.05 * s, 0.875 * s))
    nodes = ('<node id="root"><rotate>0 0 1 90</rotate><node id="n1"><rotate>0 0 1 -90</rotate><translate>0 0 %r</translate>'
             '<instance_geometry url="#base"/></node></node><node id="n2"><instance_node url="#bodynode"/></node>' % (0.2 * s))
    return _dae(geoms, nodes, lib, meter=1.0 / s)


def _module():
    import pcoll_b200
    return pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=-1)


def _rows(mod):
    return mod.scene()["poly_hulls"]


def _n_box_hulls(mod):
    return mod.scene()["hulls"].shape[0]


def _same_polygon(row, want):
    n = int(row[0])
    got = row[3:3 + 2 * n].reshape(n, 2)
    want = np.asarray(want, dtype=np.float64)
    assert n == len(want)
    k = int(np.argmin(np.abs(got - want[0]).sum(axis=1)))
    np.testing.assert_allclose(np.roll(got, -k, axis=0), want, atol=1e-9)
    area = 0.5 * np.sum(got[:, 0] * np.roll(got[:, 1], -1) - np.roll(got[:, 0], -1) * got[:, 1])
    assert area > 0   # counter-clockwise


@pytest.mark.parametrize("deflate", [True, False])
def test_two_box_model_gives_the_fixture_rectangles(tmp_path, deflate):
    path = str(tmp_path / "pr2-like.zae")
    _zae(path, _two_box_robot(), deflate=deflate)
    mod = _module()
    assert mod.SendCommand("setRobotModel %s 0.05 0.35 1.4" % path) == "2"
    rows = _rows(mod)
    assert rows.shape == (2, 19) and _n_box_hulls(mod) == 0
    np.testing.assert_allclose(rows[:, 1:3], [[0.05, 0.35], [0.35, 1.4]], atol=1e-12)
    _same_polygon(rows[0], [(-0.334, -0.334), (0.334, -0.334), (0.334, 0.334), (-0.334, 0.334)])
    _same_polygon(rows[1], [(-0.33, -0.33), (0.23, -0.33), (0.23, 0.33), (-0.33, 0.33)])
    # default bands: four equal ones over the model's height; the lowest holds the base and the bottom of the body (whose
    # footprint lies inside the base's), the others the body alone
    assert mod.SendCommand("setRobotModel %s" % path) == "4"
    rows = _rows(mod)
    np.testing.assert_allclose(rows[:, 1], 0.05 + 0.3375 * np.arange(4), atol=1e-12)
    _same_polygon(rows[0], [(-0.334, -0.334), (0.334, -0.334), (0.334, 0.334), (-0.334, 0.334)])
    for r in rows[1:]:
        _same_polygon(r, [(-0.33, -0.33), (0.23, -0.33), (0.23, 0.33), (-0.33, 0.33)])
    mod.close()


def test_boxes_option_gives_the_fixture_box_hulls(tmp_path):
    """`boxes`: every band's footprint replaced by its minimum-area bounding rectangle -> box hulls (the packed kernels'
    path).  The model of the box fixture must give the fixture's own hull rows; a part turned by 30 degrees its turned box."""
    import pcoll_b200
    path = str(tmp_path / "pr2-like.zae")
    _zae(path, _two_box_robot())
    mod = _module()
    fixture = mod.scene()["hulls"].copy()
    assert mod.SendCommand("setRobotModel %s boxes 0.05 0.35 1.4" % path) == "2"
    scn = mod.scene()
    assert scn["poly_hulls"].shape[0] == 0
    np.testing.assert_allclose(scn["hulls"], fixture, atol=1e-12)
    geoms = _box_mesh("b", 0.4, 0.1, 0.25)
    nodes = '<node id="r"><translate>0.5 -0.2 0.25</translate><rotate>0 0 1 30</rotate><instance_geometry url="#b"/></node>'
    turned = str(tmp_path / "turned.dae")
    open(turned, "w").write(_dae(geoms, nodes))
    assert mod.SendCommand("setRobotModel %s boxes 0 0.5" % turned) == "1"
    h = mod.scene()["hulls"][0]
    c, s = np.cos(np.radians(30)), np.sin(np.radians(30))
    np.testing.assert_allclose(h, [0.5, -0.2, c, s, 0.4, 0.1, 0.0, 0.5], atol=1e-9)
    with pytest.raises(pcoll_b200.McsimError, match="neither"):
        mod.SendCommand("setRobotModel %s cubes" % turned)
    mod.close()


def test_bare_dae_y_up_and_a_band_through_a_tilted_part(tmp_path):
    """A .dae read directly; Y_UP document (model y is up: robot z = y, robot y = -z); a box rotated 45 degrees about the
    vertical; a band boundary cutting through it: both bands carry the full rotated footprint."""
    geoms = _box_mesh("b", 0.2, 0.5, 0.1)          # Y_UP: half extents x 0.2, up 0.5, depth 0.1
    nodes = '<node id="r"><translate>1 0.5 0</translate><rotate>0 1 0 45</rotate><instance_geometry url="#b"/></node>'
    path = str(tmp_path / "tilted.dae")
    open(path, "w").write(_dae(geoms, nodes, up="Y_UP"))
    mod = _module()
    assert mod.SendCommand("setRobotModel %s 0 0.3 1.0" % path) == "2"
    rows = _rows(mod)
    c, s = np.cos(np.pi / 4), np.sin(np.pi / 4)
    # model corners (x, z) = (+-0.2, +-0.1) rotated by +45 degrees about model y: x' = c x + s z, z' = -s x + c z; robot (x, y) = (x' + 1, -z')
    corners = [(1 + c * x + s * z, -(-s * x + c * z)) for x, z in ((-0.2, -0.1), (0.2, -0.1), (0.2, 0.1), (-0.2, 0.1))]
    hull = sorted(corners)
    for r in rows:
        n = int(r[0])
        got = sorted(map(tuple, r[3:3 + 2 * n].reshape(n, 2)))
        np.testing.assert_allclose(got, hull, atol=1e-9)
    np.testing.assert_allclose(rows[:, 1:3], [[0, 0.3], [0.3, 1.0]], atol=1e-12)
    mod.close()


def test_round_part_is_simplified_to_eight_vertices_conservatively(tmp_path):
    path = str(tmp_path / "round.dae")
    open(path, "w").write(_dae(_prism_mesh("cyl", 16, 0.3, 0.4), '<node id="r"><translate>0.1 -0.2 0.4</translate><instance_geometry url="#cyl"/></node>'))
    mod = _module()
    assert mod.SendCommand("setRobotModel %s 0 0.8" % path) == "1"
    r = _rows(mod)[0]
    assert int(r[0]) == 8
    poly = r[3:19].reshape(8, 2)
    ang = 2 * np.pi * np.arange(16) / 16
    orig = np.stack([0.1 + 0.3 * np.cos(ang), -0.2 + 0.3 * np.sin(ang)], axis=1)
    for k in range(8):   # every vertex of the 16-gon on or left of every edge of the octagon: a superset
        a, b = poly[k], poly[(k + 1) % 8]
        side = (b[0] - a[0]) * (orig[:, 1] - a[1]) - (b[1] - a[1]) * (orig[:, 0] - a[0])
        assert np.all(side > -1e-12)
    area = 0.5 * np.sum(poly[:, 0] * np.roll(poly[:, 1], -1) - np.roll(poly[:, 0], -1) * poly[:, 1])
    area16 = 0.5 * 16 * 0.3 ** 2 * np.sin(2 * np.pi / 16)
    assert area16 <= area < 1.09 * area16          # the circumscribed octagon of the 16-gon: + 8.2 %
    mod.close()


def test_env_xml_robot_file_resolves_to_the_collada_model(tmp_path):
    """<Robot file="robots/x.zae"> in an env XML: the model next to the env file is read (z offset of the <Robot> applied,
    zbands attribute honoured); without it the box fixture on the fixture path is used, as before."""
    import pcoll_b200
    env = open(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml")).read()
    assert 'file="robots/pr2-beta-static.zae"' in env
    os.makedirs(tmp_path / "robots")
    _zae(str(tmp_path / "robots" / "pr2-beta-static.zae"), _two_box_robot())
    envp = str(tmp_path / "scene.env.xml")
    open(envp, "w").write(env.replace('file="robots/pr2-beta-static.zae"', 'file="robots/pr2-beta-static.zae" zbands="0.05 0.35 1.4"'))
    for f in os.listdir(os.path.join(ROOT, "data", "scenes")):          # the kinbody files the env includes
        if f.endswith(".kinbody.xml"):
            open(str(tmp_path / f), "w").write(open(os.path.join(ROOT, "data", "scenes", f)).read())
    mod = pcoll_b200.MCModule(envp, device=-1)
    rows = _rows(mod)
    ref = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=-1)
    assert rows.shape[0] == 2 and _n_box_hulls(mod) == 0 and _n_box_hulls(ref) == 2 and _rows(ref).shape[0] == 0
    # like the fixture's boxes, the model's hulls stay in the robot frame and take the <Robot> element's z translation
    # (0.05 m in this env file; the fixture's boxes sit at 0 .. 0.3 and 0.3 .. 1.35 above the robot origin)
    box = ref.scene()["hulls"]
    np.testing.assert_allclose(box[:, 6:8], [[0.05, 0.35], [0.35, 1.4]], atol=1e-9)
    np.testing.assert_allclose(rows[:, 1:3], np.array([[0.05, 0.35], [0.35, 1.4]]) + 0.05, atol=1e-9)
    np.testing.assert_allclose(mod.scene()["robot_pose"], ref.scene()["robot_pose"])
    _same_polygon(rows[0], [(-0.334, -0.334), (0.334, -0.334), (0.334, 0.334), (-0.334, 0.334)])
    mod.close(); ref.close()


def test_collada_errors_are_loud(tmp_path):
    import pcoll_b200
    mod = _module()
    bad = str(tmp_path / "bad.zae")
    open(bad, "wb").write(b"PK\x03\x04 this is not an archive")
    with pytest.raises(pcoll_b200.McsimError, match="zip"):
        mod.SendCommand("setRobotModel " + bad)
    empty = str(tmp_path / "empty.dae")
    open(empty, "w").write(_dae("", '<node id="n"/>'))
    with pytest.raises(pcoll_b200.McsimError, match="no mesh geometry"):
        mod.SendCommand("setRobotModel " + empty)
    xup = str(tmp_path / "xup.dae")
    open(xup, "w").write(_dae(_box_mesh("b", 1, 1, 1), '<node id="n"><instance_geometry url="#b"/></node>', up="X_UP"))
    with pytest.raises(pcoll_b200.McsimError, match="X_UP"):
        mod.SendCommand("setRobotModel " + xup)
    ok = str(tmp_path / "ok.dae")
    open(ok, "w").write(_dae(_box_mesh("b", 1, 1, 1), '<node id="n"><instance_geometry url="#b"/></node>'))
    with pytest.raises(pcoll_b200.McsimError, match="increase"):
        mod.SendCommand("setRobotModel %s 0.5 0.2" % ok)
    with pytest.raises(pcoll_b200.McsimError, match="at most"):
        mod.SendCommand("setRobotModel %s 0 1 2 3 4 5" % ok)
    with pytest.raises(pcoll_b200.McsimError):
        mod.SendCommand("setRobotModel " + str(tmp_path / "missing.zae"))
    mod.close()
