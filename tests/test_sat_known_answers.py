"""Independent pins of the collision predicate (replaces CheckCollision, MCSimulator.h:269-285).

The reference's own checker (ODE + the PR2 mesh) is not in its repository, so the GPU SAT (csrc/sat_device.cuh) and the
oracle's CPU SAT (oracle/sat_cpu.h) are both restatements of one formula.  Two things here are NOT that formula:

* tests/golden/sat_cases.json -- 40 hand-derived known answers (10 of them for convex-polygon robot hulls) (face contact counts as a hit, 45-degree corner-to-edge
  gaps that only the box's own axis separates, z-disjoint prisms, nesting, zero-extent boxes, hull yaw / offset / robot
  pose, the pr2custom spike tip at y = -1.133), each with its derivation; checked against sat_cpu.h (CPU) and against the
  device audit path (GPU).
* a second algorithm written for this test only -- closed convex polygons intersect iff a vertex of one lies in the
  other or two edges cross (orientation predicates, no projections) -- compared with sat_cpu.h on 10^5 random
  hull / box / pose triples outside |gap| < 1e-9.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

from conftest import GOLD

CASES = json.load(open(os.path.join(GOLD, "sat_cases.json")))["cases"]
NEAR = 1e-4   # device FP32: |expected gap| below this is a contact case (flag not asserted, clearance is)


def _boxes(case):
    if "scene" in case:
        return np.load(os.path.join(GOLD, "scene_%s.npz" % case["scene"]))["boxes"]
    return np.asarray(case["boxes"], dtype=np.float64).reshape(-1, 8)


def _sat_cpu(boxes, hulls, pose, poly_hulls=None):
    """oracle/sat_cpu.h through oracle/_ref (ref_clearance)."""
    import oracle_ref
    b = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 8)
    h = np.ascontiguousarray(hulls, dtype=np.float64).reshape(-1, 8)
    p = oracle_ref.RefParams()
    p.boxes, p.n_boxes = b.ctypes.data_as(oracle_ref._dp), b.shape[0]
    p.hulls, p.n_hulls = h.ctypes.data_as(oracle_ref._dp), h.shape[0]
    if poly_hulls is not None and len(poly_hulls):
        po = np.ascontiguousarray(poly_hulls, dtype=np.float64).reshape(-1, 19)
        p.poly_hulls, p.n_poly_hulls = po.ctypes.data_as(oracle_ref._dp), po.shape[0]
    ps = (C.c_double * 3)(*[float(v) for v in pose])
    return oracle_ref.lib().ref_clearance(C.byref(p), ps)


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_sat_cpu_known_answers(case):
    import oracle_ref
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    g = _sat_cpu(_boxes(case), case["hulls"], case["pose"], case.get("poly_hulls"))
    if case["gap"] is None:
        assert g > 1e300 and not case["hit"]        # no live pair: +DBL_MAX
    else:
        assert abs(g - case["gap"]) < 1e-9, (case["name"], g, case["gap"], case["why"])
    if case["gap"] is None or abs(case["gap"]) > 1e-12:
        assert (g <= 0) == case["hit"], case["why"]
    else:
        assert abs(g) < 1e-12   # exact contact: hit by definition (gap <= 0); FP64 rounding may leave +-1e-16


# ---- second algorithm: vertex containment + edge crossing on the closed rectangles -----------------------------
def _corners(cx, cy, c, s, hx, hy):
    """[n, 4, 2] corners in counter-clockwise order."""
    sx = np.array([1, -1, -1, 1.0])
    sy = np.array([1, 1, -1, -1.0])
    ux = sx[None, :] * hx[:, None]
    uy = sy[None, :] * hy[:, None]
    return np.stack([cx[:, None] + c[:, None] * ux - s[:, None] * uy, cy[:, None] + s[:, None] * ux + c[:, None] * uy], axis=-1)


def _orient(a, b, p):
    return (b[..., 0] - a[..., 0]) * (p[..., 1] - a[..., 1]) - (b[..., 1] - a[..., 1]) * (p[..., 0] - a[..., 0])


def _polygons_intersect(A, B):
    """A, B: [n, 4, 2] convex CCW quadrilaterals (may be degenerate).  Closed sets."""
    n = A.shape[0]
    hit = np.zeros(n, dtype=bool)

    def inside(P, poly):   # every vertex of P against poly: on or left of all 4 edges
        ok = np.ones((n, 4), dtype=bool)
        for e in range(4):
            a, b = poly[:, e, None, :], poly[:, (e + 1) % 4, None, :]
            ok &= _orient(a, b, P) >= 0
        return ok.any(axis=1)

    hit |= inside(A, B) | inside(B, A)
    for i in range(4):
        a0, a1 = A[:, i], A[:, (i + 1) % 4]
        for j in range(4):
            b0, b1 = B[:, j], B[:, (j + 1) % 4]
            d1, d2 = _orient(a0, a1, b0), _orient(a0, a1, b1)
            d3, d4 = _orient(b0, b1, a0), _orient(b0, b1, a1)
            hit |= (d1 * d2 < 0) & (d3 * d4 < 0)   # proper crossing; touching is caught by the containment test
    return hit


def test_second_algorithm_agrees_with_sat_cpu_on_random_pairs():
    import oracle_ref
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.Generator(np.random.PCG64(77))
    n = 100_000
    # hull in the robot frame, robot pose, box in the world; scales of the scenes (extents 0.02 .. 1 m), close together
    ox, oy = rng.uniform(-0.3, 0.3, n), rng.uniform(-0.3, 0.3, n)
    psi, th, phi = rng.uniform(0, 2 * np.pi, n), rng.uniform(0, 2 * np.pi, n), rng.uniform(0, 2 * np.pi, n)
    hhx, hhy = rng.uniform(0.02, 0.6, n), rng.uniform(0.02, 0.6, n)
    bhx, bhy = rng.uniform(0.02, 1.0, n), rng.uniform(0.02, 1.0, n)
    x, y = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    bx, by = x + rng.normal(0, 0.8, n), y + rng.normal(0, 0.8, n)
    # a quarter of them axis-aligned boxes / headings (the scenes' common case; exercises the degenerate cross axes)
    q = n // 4
    phi[:q] = rng.integers(0, 4, q) * (np.pi / 2)
    th[:q // 2] = rng.integers(0, 4, q // 2) * (np.pi / 2)
    psi[:q] = 0.0
    gaps = np.empty(n)
    lib = oracle_ref.lib()
    p = oracle_ref.RefParams()
    bb, hh = np.zeros((1, 8)), np.zeros((1, 8))
    p.boxes, p.n_boxes, p.hulls, p.n_hulls = bb.ctypes.data_as(oracle_ref._dp), 1, hh.ctypes.data_as(oracle_ref._dp), 1
    pose = (C.c_double * 3)()
    for i in range(n):
        bb[0] = (bx[i], by[i], np.cos(phi[i]), np.sin(phi[i]), bhx[i], bhy[i], 0.0, 1.0)
        hh[0] = (ox[i], oy[i], np.cos(psi[i]), np.sin(psi[i]), hhx[i], hhy[i], 0.0, 1.0)
        pose[0], pose[1], pose[2] = x[i], y[i], th[i]
        gaps[i] = lib.ref_clearance(C.byref(p), pose)
    # world-frame hull rectangle
    ct, st = np.cos(th), np.sin(th)
    hcx, hcy = x + ct * ox - st * oy, y + st * ox + ct * oy
    A = _corners(hcx, hcy, np.cos(th + psi), np.sin(th + psi), hhx, hhy)
    B = _corners(bx, by, np.cos(phi), np.sin(phi), bhx, bhy)
    hit2 = _polygons_intersect(A, B)
    clear = np.abs(gaps) >= 1e-9
    assert clear.sum() > 0.999 * n
    sat_hit = gaps <= 0
    assert 0.2 < sat_hit.mean() < 0.8          # both outcomes well represented
    bad = np.nonzero((sat_hit != hit2) & clear)[0]
    assert bad.size == 0, "SAT and the polygon-intersection algorithm disagree on %d of %d pairs, first %s gap %g" % (
        bad.size, n, bad[:5], gaps[bad[0]])


def test_polygon_path_equals_box_path_for_rectangles():
    """A box hull handed over as a 4-vertex polygon: the polygon SAT (box axes + edge normals) must give the box SAT's gap."""
    import oracle_ref
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.Generator(np.random.PCG64(5))
    for _ in range(2000):
        ox, oy, psi = rng.uniform(-0.3, 0.3), rng.uniform(-0.3, 0.3), rng.uniform(0, 2 * np.pi)
        hx, hy = rng.uniform(0.05, 0.6, 2)
        box = [rng.normal(0, 0.8), rng.normal(0, 0.8), 0, 0, rng.uniform(0.02, 1.0), rng.uniform(0.02, 1.0), 0.0, 1.0]
        phi = rng.uniform(0, 2 * np.pi)
        box[2], box[3] = np.cos(phi), np.sin(phi)
        pose = [rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(0, 2 * np.pi)]
        c, s_ = np.cos(psi), np.sin(psi)
        verts = [(ox + c * ux * hx - s_ * uy * hy, oy + s_ * ux * hx + c * uy * hy) for ux, uy in ((1, 1), (-1, 1), (-1, -1), (1, -1))]
        po = [4, 0.0, 1.0] + [v for xy in verts for v in xy] + [0.0] * 8
        a = _sat_cpu([box], [[ox, oy, c, s_, hx, hy, 0.0, 1.0]], pose)
        b = _sat_cpu([box], np.zeros((0, 8)), pose, [po])
        assert abs(a - b) < 1e-12, (a, b)


@pytest.mark.gpu
def test_device_sat_known_answers(ctx, plan):
    """The same known answers through the device: replay mode, one rollout, one waypoint, z0 = 0 (x0 = the pose),
    audit launch (scalar SAT: clearance + flag) and the counting launch (pair_hit)."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    for case in CASES:
        one = dict(plan)
        one["trajectory"] = np.asarray(case["pose"], dtype=np.float64).reshape(3, 1)
        one["odometry"] = np.zeros((3, 0))
        ctx.set_scene(_boxes(case), np.asarray(case["hulls"], dtype=np.float64).reshape(-1, 8), case.get("poly_hulls"))
        p, keep = abi.make_mc_params(one, n_rollouts=1, mode=abi.MODE_REPLAY, z0=np.zeros(3), early_exit=False)
        aud = ctx.mc_run(p, want_counts=True, audit=True)
        run = ctx.mc_run(p, want_counts=True)
        g = float(aud["clearance"][0, 0])
        if case["gap"] is None:
            assert g > 1e30 and run["hits"] == 0 and run["pair_tests"] == 0, case["name"]
            continue
        assert abs(g - case["gap"]) < 2e-6 * max(1.0, abs(case["gap"]) * 10), (case["name"], g, case["gap"])
        if abs(case["gap"]) >= NEAR:
            assert (g <= 0) == case["hit"], case["name"]
            assert run["hits"] == int(case["hit"]) and int(run["counts"][0]) == int(case["hit"]), (case["name"], case["why"])
            assert aud["near_contact"] == 0
        else:
            assert aud["near_contact"] == 1     # reported separately, as north_star asks
        if case.get("poly_hulls") and abs(case["gap"]) >= NEAR:
            # polygon hulls route Philox mode to the scalar kernel: every rollout sits at the pose (tiny start covariance)
            tiny = dict(one)
            tiny["cov0"] = np.eye(3) * 1e-16
            pp, kk = abi.make_mc_params(tiny, n_rollouts=64, mode=abi.MODE_PHILOX, early_exit=False)
            assert ctx.mc_run(pp)["hits"] == (64 if case["hit"] else 0), case["name"]
            continue
        # the packed Philox kernel's SAT (orientation-grouped table, broad phase on and off) at the same pose:
        # a tiny start covariance puts every rollout at the pose
        tiny = dict(one)
        tiny["cov0"] = np.eye(3) * 1e-16
        if abs(case["gap"]) >= NEAR and not case.get("poly_hulls"):
            for ex in (False, True):
                pp, kk = abi.make_mc_params(tiny, n_rollouts=64, mode=abi.MODE_PHILOX, early_exit=False, exhaustive=ex)
                o = ctx.mc_run(pp)
                assert o["hits"] == (64 if case["hit"] else 0), (case["name"], ex, o["hits"])
