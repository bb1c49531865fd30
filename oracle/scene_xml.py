"""TEST INFRASTRUCTURE ONLY (oracle/): Python restatement of the scene extraction.

The product extracts the collision geometry from the OpenRAVE env XML in C++
(csrc/scene_xml.cpp); this is an independent restatement with xml.etree used by the tests to check
that extractor, and by tests/golden/make_inputs.py to derive the committed inputs.

Semantics, all visible in the two scene files of the reference (pr2custom.env.xml:14-249,
pr2test2.env.xml:12-122):
  * <Extents> are half-extents; <Translation> places the geom centre; <RotationAxis>ax ay az deg
    rotates the geom about its own centre (only rotations about +-z are representable in 2.5-D);
    translation and rotation of one tag are independent of their order.
  * <offsetfrom>Name</offsetfrom> expresses the body in the frame of the named body.
  * KinBody-level <Translation>/<RotationAxis> move the whole body (the tables of pr2test2).
  * <KinBody file=...> / <Robot file=...> pull geometry from another file; files that are not in
    the repository (ikeatable.kinbody.xml, the PR2 .zae) resolve through a fixture
    search path.
  * Tag case varies (<translation> at pr2test2.env.xml:121), so tags are matched case-insensitively.

Output tables (float64):
  boxes[n, 8] = cx, cy, cos(phi), sin(phi), hx, hy, zlo, zhi      (world frame)
  hulls[m, 8] = ox, oy, cos(psi), sin(psi), hx, hy, zlo, zhi      (robot frame; z absolute)
"""
import math
import os
import xml.etree.ElementTree as ET

import numpy as np


class Xf:
    """Planar rigid transform + z offset: p_world = R(yaw) p + (x, y), z_world = z_local + z."""

    def __init__(self, x=0.0, y=0.0, z=0.0, yaw=0.0):
        self.x, self.y, self.z, self.yaw = x, y, z, yaw

    def __mul__(self, o):
        c, s = math.cos(self.yaw), math.sin(self.yaw)
        return Xf(self.x + c * o.x - s * o.y, self.y + s * o.x + c * o.y, self.z + o.z, self.yaw + o.yaw)


def _children(el, name):
    return [c for c in el if isinstance(c.tag, str) and c.tag.lower() == name]


def _floats(el):
    return [float(t) for t in el.text.split()]


def _local_xf(el):
    """Transform given by the <Translation>/<RotationAxis> children of el (either may be absent)."""
    xf = Xf()
    for t in _children(el, "translation"):
        v = _floats(t)
        xf.x, xf.y, xf.z = v[0], v[1], v[2]
    for r in _children(el, "rotationaxis"):
        ax, ay, az, deg = _floats(r)
        n = math.sqrt(ax * ax + ay * ay + az * az)
        if n == 0.0 or deg == 0.0:
            continue
        if abs(ax) > 1e-12 * n or abs(ay) > 1e-12 * n:
            raise ValueError("RotationAxis %r is not about z: not representable as a z-prism" % (r.text,))
        xf.yaw = math.radians(deg) * (1.0 if az > 0 else -1.0)
    return xf


def _resolve(path, base_dir, fixture_dirs):
    cands = [os.path.join(base_dir, path)]
    stem = os.path.basename(path)
    for d in fixture_dirs:
        cands.append(os.path.join(d, stem))
        cands.append(os.path.join(d, os.path.splitext(stem)[0] + ".fixture.xml"))
    for c in cands:
        if os.path.isfile(c) and not c.endswith(".zae"):
            return c
    raise FileNotFoundError("cannot resolve %r (tried %s)" % (path, cands))


def _body_prisms(kin, parent_xf):
    """All box geoms of a <KinBody>/<Robot> element as (Xf world, hx, hy, hz)."""
    out = []
    frames = {}
    for body in _children(kin, "body"):
        xf = _local_xf(body)
        off = _children(body, "offsetfrom")
        if off:
            xf = frames[off[0].text.strip()] * xf
        name = body.get("name")
        if name:
            frames[name] = xf
        for geom in _children(body, "geom"):
            if (geom.get("type") or "").lower() != "box":
                raise ValueError("only box geoms are supported, got %r" % geom.get("type"))
            hx, hy, hz = _floats(_children(geom, "extents")[0])
            out.append((parent_xf * xf * _local_xf(geom), hx, hy, hz))
    return out


def extract_scene(env_xml, fixture_dirs=()):
    """Returns dict(boxes, hulls, robot_pose, box_names)."""
    base_dir = os.path.dirname(os.path.abspath(env_xml))
    root = ET.parse(env_xml).getroot()
    boxes, names = [], []
    for kin in _children(root, "kinbody"):
        top = _local_xf(kin)
        src = kin
        if kin.get("file"):
            src = ET.parse(_resolve(kin.get("file"), base_dir, fixture_dirs)).getroot()
        for xf, hx, hy, hz in _body_prisms(src, top):
            boxes.append([xf.x, xf.y, math.cos(xf.yaw), math.sin(xf.yaw), hx, hy, xf.z - hz, xf.z + hz])
            names.append(kin.get("name") or "")
    hulls = []
    robot_pose = None
    robots = _children(root, "robot")
    if robots:
        rob = robots[0]
        top = _local_xf(rob)
        robot_pose = [top.x, top.y, top.yaw]
        src = rob
        if rob.get("file"):
            src = ET.parse(_resolve(rob.get("file"), base_dir, fixture_dirs)).getroot()
        # hulls stay in the robot frame; only the robot's z offset is applied
        for xf, hx, hy, hz in _body_prisms(src, Xf(0.0, 0.0, top.z, 0.0)):
            hulls.append([xf.x, xf.y, math.cos(xf.yaw), math.sin(xf.yaw), hx, hy, xf.z - hz, xf.z + hz])
    return dict(
        boxes=np.asarray(boxes, dtype=np.float64).reshape(-1, 8),
        hulls=np.asarray(hulls, dtype=np.float64).reshape(-1, 8),
        robot_pose=robot_pose,
        box_names=names,
    )


def live_pairs(boxes, hulls):
    """(hull, box) index pairs whose z-intervals overlap with positive length (oracle/sat_cpu.h)."""
    return [(h, b) for h in range(len(hulls)) for b in range(len(boxes))
            if hulls[h][6] < boxes[b][7] and boxes[b][6] < hulls[h][7]]
