# writes tests/golden/sat_cases.json -- every expected value below is typed in from the hand derivation in the
# "why" string (plain interval arithmetic on the four face normals), not computed by any SAT code.
import json, math
U = [0, 0, 1, 0, 0.5, 0.5, 0.0, 1.0]          # unit-square hull, z 0..1
def box(cx, cy, hx=0.5, hy=0.5, yaw=0.0, z=(0.0, 1.0)):
    return [cx, cy, math.cos(math.radians(yaw)), math.sin(math.radians(yaw)), hx, hy, z[0], z[1]]
R2 = math.sqrt(2.0)
cases = []
def add(name, hulls, boxes, pose, hit, gap, why, scene=None):
    c = dict(name=name, hulls=hulls, pose=pose, hit=hit, gap=gap, why=why)
    if scene: c["scene"] = scene
    else: c["boxes"] = boxes
    cases.append(c)
add("face_gap", [U], [box(1.5, 0)], [0, 0, 0], False, 0.5, "x axis: |1.5| - 0.5 - 0.5 = 0.5")
add("face_touch", [U], [box(1.0, 0)], [0, 0, 0], True, 0.0, "faces coincide at x = 0.5: gap 0, touching counts as a hit (gap <= 0)")
add("face_overlap_1mm", [U], [box(0.999, 0)], [0, 0, 0], True, -0.001, "x axis: 0.999 - 1 = -0.001; y axis: -1")
add("face_gap_1mm", [U], [box(1.001, 0)], [0, 0, 0], False, 0.001, "x axis: 1.001 - 1")
add("corner_touch", [U], [box(1.0, 1.0)], [0, 0, 0], True, 0.0, "corners (0.5, 0.5) coincide: both axes 1 - 1 = 0")
add("corner_gap", [U], [box(1.002, 1.001)], [0, 0, 0], False, 0.002, "x: 0.002, y: 0.001, SAT gap = max")
add("diamond_vertex_gap", [U], [box(0.5 + R2 / 2 + 0.01, 0, yaw=45)], [0, 0, 0], False, 0.01,
    "45-degree square: radius on the x axis 0.5 (cos45 + sin45) = 0.70711; its left vertex is at x = 0.51")
add("diamond_vertex_overlap", [U], [box(0.5 + R2 / 2 - 0.01, 0, yaw=45)], [0, 0, 0], True, -0.01, "left vertex at x = 0.49, inside the hull")
add("corner_to_edge_gap", [U], [box(0.86, 0.86, yaw=45)], [0, 0, 0], False, 0.86 * R2 - 0.5 - R2 / 2,
    "only the box's own axis (1,1)/sqrt2 separates: centre 0.86 sqrt2 = 1.21622, radii 0.5 + 0.5 sqrt2 = 1.20711 -> 0.0091169; the hull axes overlap (0.86 - 1.2071 < 0)")
add("corner_to_edge_overlap", [U], [box(0.85, 0.85, yaw=45)], [0, 0, 0], True, 0.85 * R2 - 0.5 - R2 / 2,
    "0.85 sqrt2 = 1.20208 < 1.20711: the hull corner (0.5, 0.5) is 5.03 mm inside the box's edge")
add("z_touching_is_not_live", [U], [box(0.2, 0.1, z=(1.0, 2.0))], [0, 0, 0], False, None, "z intervals [0,1] and [1,2] share no positive length: the pair is not live, whatever x, y")
add("z_overlap_is_live", [U], [box(0.2, 0.1, z=(0.9, 2.0))], [0, 0, 0], True, -0.8, "z overlap 0.1; x axis 0.2 - 1, y axis 0.1 - 1 -> -0.8")
add("floor_below_robot", [[0, 0, 1, 0, 0.334, 0.334, 0.05, 0.35]], [box(0, 0, 4.0, 2.0, z=(-0.2, 0.0))], [1.0, 0.5, 0.3], False, None,
    "the scene floors (z -0.2..0) lie under the base hull (z 0.05..0.35): never a live pair")
add("nested_small_in_hull", [U], [box(0.1, 0.1, 0.1, 0.1)], [0, 0, 0], True, -0.5, "0.1 - 0.5 - 0.1 on both axes")
add("nested_hull_in_big_box", [U], [box(1.0, 1.0, 5.0, 5.0)], [0, 0, 0], True, -4.5, "1 - 0.5 - 5")
add("segment_inside", [U], [box(0.2, 0, 0.0, 0.5)], [0, 0, 0], True, -0.3, "zero-width box (a segment): x axis 0.2 - 0.5 - 0 = -0.3, y axis -1")
add("segment_outside", [U], [box(0.6, 0, 0.0, 0.5)], [0, 0, 0], False, 0.1, "x axis 0.6 - 0.5 - 0")
add("point_outside", [U], [box(0.501, 0.2, 0.0, 0.0)], [0, 0, 0], False, 0.001, "zero-extent box (a point) 1 mm right of the face")
add("point_inside", [U], [box(0.3, 0.2, 0.0, 0.0)], [0, 0, 0], True, -0.2, "x: 0.3 - 0.5, y: 0.2 - 0.5 -> max = -0.2")
H = [0.3, 0, 1, 0, 0.4, 0.2, 0.0, 1.0]
add("posed_offset_hull_gap", [H], [box(1, 1.85, 0.5, 0.1)], [1, 1, math.pi / 2], False, 0.05,
    "pose (1, 1, 90 deg): hull centre (1, 1.3), its long axis along +y: reaches y = 1.7; box bottom face y = 1.75")
add("posed_offset_hull_overlap", [H], [box(1, 1.79, 0.5, 0.1)], [1, 1, math.pi / 2], True, -0.01, "box bottom face y = 1.69 < 1.7")
Y = [0, 0, 0, 1, 0.4, 0.2, 0.0, 1.0]
add("hull_yaw_90_gap", [Y], [box(0.35, 0, 0.1, 0.1)], [0, 0, 0], False, 0.05, "hull yawed 90 deg in the robot frame: +-0.2 in x; 0.35 - 0.2 - 0.1 (ignoring the yaw would hit)")
add("hull_yaw_90_overlap", [Y], [box(0, 0.45, 0.1, 0.1)], [0, 0, 0], True, -0.05, "+-0.4 in y: 0.45 - 0.4 - 0.1 (ignoring the yaw would miss)")
B0 = [0, 0, 1, 0, 0.5, 0.5, 0.0, 0.3]
B1 = [-0.05, 0, 1, 0, 0.28, 0.33, 0.35, 1.4]
add("second_hull_only_gap", [B0, B1], [box(0.6, 0, 0.3, 0.1, z=(0.5, 0.6))], [0, 0, 0], False, 0.07,
    "box z 0.5..0.6 meets only the upper hull (centre x = -0.05, hx 0.28): 0.65 - 0.28 - 0.3; the lower hull (hx 0.5) would overlap but is not live")
add("second_hull_only_overlap", [B0, B1], [box(0.5, 0, 0.3, 0.1, z=(0.5, 0.6))], [0, 0, 0], True, -0.03, "0.55 - 0.58")
S = [0, 0, 1, 0, 0.3, 0.3, 0.05, 0.35]
tip = 0.5 * math.sin(math.radians(60)) + 0.02 * math.cos(math.radians(60))
add("pr2custom_spike_tip_gap", [S], None, [-3.15, -1.5, 0], False, 0.8 - 0.3 - tip,
    "BottomSpike 0 (centre (-3.4, -0.7), 0.5 x 0.02, -60 deg) and 1 (+60 deg) meet in the tip at (-3.15, -1.133); lowest corner y = -0.7 - 0.43301 - 0.01 = -1.14301; hull top y = -1.2 -> 0.056987",
    scene="pr2custom")
add("pr2custom_spike_tip_overlap", [S], None, [-3.15, -1.44, 0], True, 0.74 - 0.3 - tip, "hull top y = -1.14: 3.01 mm above the tip's lowest corner", scene="pr2custom")
add("rotated_pose_45_gap", [U], [box(0.5 + R2 / 2 + 0.01, 0)], [0, 0, math.pi / 4], False, 0.01, "the robot is the diamond: mirror image of diamond_vertex_gap")
d30 = 0.5 * (math.cos(math.radians(30)) + math.sin(math.radians(30)))
add("rotated_pose_30_gap", [U], [box(1.2, 0)], [0, 0, math.pi / 6], False, 1.2 - 0.5 - d30,
    "hull rotated 30 deg reaches x = 0.5 (cos30 + sin30) = 0.68301; box left face x = 0.7 -> 0.016987; hull axes overlap (1.2 cos30 - 1.18301 < 0)")
add("rotated_pose_30_overlap", [U], [box(1.18, 0)], [0, 0, math.pi / 6], True, 1.18 - 0.5 - d30, "box left face x = 0.68 < 0.68301")
# ---- convex-polygon robot hulls: {n, zlo, zhi, x0, y0, ...} counter-clockwise, robot frame
def poly(verts, z=(0.0, 1.0)):
    flat = [c for v in verts for c in v] + [0.0] * (16 - 2 * len(verts))
    return [len(verts), z[0], z[1]] + flat
OCT = poly([(0.5 * math.cos(math.radians(45 * k)), 0.5 * math.sin(math.radians(45 * k))) for k in range(8)])
TRI = poly([(0, 0), (1, 0), (0, 1)])
apo = 0.5 * math.cos(math.radians(22.5))
def addp(name, polys, boxes, pose, hit, gap, why, hulls=()):
    cases.append(dict(name=name, hulls=[list(h) for h in hulls], poly_hulls=polys, boxes=boxes, pose=pose, hit=hit, gap=gap, why=why))
addp("octagon_vertex_gap", [OCT], [box(1.01, 0)], [0, 0, 0], False, 0.01, "regular octagon, circumradius 0.5, a vertex at (0.5, 0); box face at x = 0.51: the box's x axis separates by 0.01")
addp("octagon_vertex_overlap", [OCT], [box(0.99, 0)], [0, 0, 0], True, -0.01, "box face at x = 0.49")
addp("octagon_edge_gap", [OCT], [box(0.77 * math.cos(math.radians(22.5)), 0.77 * math.sin(math.radians(22.5)), 0.3, 0.3, yaw=22.5)], [0, 0, 0], False, 0.77 - apo - 0.3,
     "box face parallel to the octagon's edge between 0 and 45 deg (normal at 22.5 deg, apothem 0.5 cos 22.5 = 0.46194): 0.77 - 0.46194 - 0.3 = 0.0080602")
addp("octagon_edge_overlap", [OCT], [box(0.75 * math.cos(math.radians(22.5)), 0.75 * math.sin(math.radians(22.5)), 0.3, 0.3, yaw=22.5)], [0, 0, 0], True, 0.75 - apo - 0.3, "0.75 - 0.76194 = -0.0119398")
addp("triangle_hypotenuse_gap", [TRI], [box(0.8, 0.8, 0.2, 0.2)], [0, 0, 0], False, 0.2 / R2 * 1.0,
     "right triangle (0,0) (1,0) (0,1): only its hypotenuse normal (1,1)/sqrt2 separates -- the box's axes overlap ([0.6, 1] inside [0, 1]); box corner (0.6, 0.6) projects to 0.848528, the hypotenuse to 0.707107: 0.141421")
addp("triangle_hypotenuse_overlap", [TRI], [box(0.69, 0.69, 0.2, 0.2)], [0, 0, 0], True, 0.98 / R2 - 1 / R2, "box corner (0.49, 0.49) is 0.014142 inside the hypotenuse")
addp("triangle_rotated_pose_gap", [TRI], [box(-0.8, 0.8, 0.2, 0.2)], [0, 0, math.pi / 2], False, 0.2 / R2, "the same triangle turned 90 deg by the robot pose: vertices (0,0) (0,1) (-1,0), mirror image of triangle_hypotenuse_gap")
addp("polygon_z_disjoint", [poly([(0, 0), (1, 0), (0, 1)], z=(0.0, 0.3))], [box(0.2, 0.2, 0.2, 0.2, z=(0.3, 1.0))], [0, 0, 0], False, None, "z ranges touch only: no live pair")
addp("box_and_polygon_hulls_gap", [poly([(0.4, -0.1), (0.7, -0.1), (0.7, 0.1), (0.4, 0.1)], z=(0.5, 0.9))], [box(1.0, 0, 0.25, 0.25, z=(0.5, 0.9))], [0, 0, 0], False, 0.05,
     "a box hull (half 0.3, z 0..0.3) that cannot see the obstacle (z 0.5..0.9) and a polygon arm x in [0.4, 0.7] at z 0.5..0.9: obstacle face at x = 0.75", hulls=[[0, 0, 1, 0, 0.3, 0.3, 0.0, 0.3]])
addp("box_and_polygon_hulls_arm_hits", [poly([(0.4, -0.1), (0.7, -0.1), (0.7, 0.1), (0.4, 0.1)], z=(0.5, 0.9))], [box(0.93, 0, 0.25, 0.25, z=(0.5, 0.9))], [0, 0, 0], True, -0.02,
     "obstacle face at x = 0.68: 2 cm inside the arm; the base hull still cannot see it", hulls=[[0, 0, 1, 0, 0.3, 0.3, 0.0, 0.3]])
json.dump(dict(note="hand-derived known answers for the 2.5-D SAT predicate (oracle/sat_cpu.h, csrc/sat_device.cuh); "
               "prisms are [cx, cy, cos, sin, hx, hy, zlo, zhi]; gap = expected signed SAT clearance (null: no live pair)",
               cases=cases), open("/root/repo/tests/golden/sat_cases.json", "w"), indent=1)
print(len(cases), "cases")
