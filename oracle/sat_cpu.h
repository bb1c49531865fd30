/* TEST INFRASTRUCTURE ONLY (oracle/): CPU 2.5-D separating-axis collision checker, FP64.
 *
 * Plays the role of OpenRAVE's EnvironmentBase::CheckCollision(robot) for the reference estimator
 * (called at mcsimplugin/MCSimulator.h:279 via checkCollision :257-285).  The reference's own
 * checker (ODE + the PR2 .zae mesh) is not in the repo, so this file follows the
 * scene XMLs (pr2custom.env.xml:14-238, pr2test2.env.xml:14-118) and the documented robot
 * fixture instead; the CUDA path uses the same table, so parity is "same geometry both sides".
 *
 * Geometry: every obstacle is a z-extruded oriented rectangle ("prism")
 *     {cx, cy, cos(phi), sin(phi), hx, hy, zlo, zhi}            (8 doubles)
 * and the robot is a set of prisms given in the robot frame
 *     {ox, oy, cos(psi), sin(psi), hx, hy, zlo, zhi}            (8 doubles, z absolute)
 * A (hull, obstacle) pair is live iff their z-intervals overlap with positive length; that test
 * does not depend on the pose, so it is done once when the table is built.
 *
 * gap(pair, pose) = max over the 4 face normals of (|projected centre distance| - r_hull - r_box);
 * the pair collides iff gap <= 0 (touching counts).  clearance(pose) = min over live pairs of gap.
 * |clearance| < eps marks a "near-contact" pose (BASELINE.json north_star).
 */
#ifndef ORACLE_SAT_CPU_H
#define ORACLE_SAT_CPU_H

#include <math.h>
#include <float.h>

#define SAT_PRISM_STRIDE 8

/* Robot hulls beyond boxes: a convex polygon of up to 8 vertices, counter-clockwise, in the robot
 * frame, extruded over [zlo, zhi]:  {n, zlo, zhi, x0, y0, ..., x7, y7}  (19 doubles).  Obstacles stay boxes. */
#define SAT_POLY_STRIDE 19
#define SAT_POLY_MAX_VERTICES 8

typedef struct {
    const double* boxes;  /* n_boxes x 8 */
    int n_boxes;
    const double* hulls;  /* n_hulls x 8 */
    int n_hulls;
    const double* poly_hulls;  /* n_poly_hulls x 19, may be NULL */
    int n_poly_hulls;
} sat_scene;

static inline int sat_z_overlap(const double* hull, const double* box) {
    return (hull[6] < box[7]) && (box[6] < hull[7]);
}

/* separation of one hull (world centre hcx,hcy, world axes ch,sh) from one box; <= 0 means overlap */
static inline double sat_pair_gap(double hcx, double hcy, double ch, double sh, double hhx, double hhy,
                                  const double* box) {
    const double dx = box[0] - hcx;
    const double dy = box[1] - hcy;
    const double cb = box[2], sb = box[3], bhx = box[4], bhy = box[5];
    /* relative rotation hull -> box */
    const double c = fabs(ch * cb + sh * sb);
    const double s = fabs(sh * cb - ch * sb);
    /* centre distance in hull frame and in box frame */
    const double drx = fabs(ch * dx + sh * dy);
    const double dry = fabs(ch * dy - sh * dx);
    const double dbx = fabs(cb * dx + sb * dy);
    const double dby = fabs(cb * dy - sb * dx);
    double g = drx - (hhx + bhx * c + bhy * s);
    double t = dry - (hhy + bhx * s + bhy * c);
    if (t > g) g = t;
    t = dbx - (bhx + hhx * c + hhy * s);
    if (t > g) g = t;
    t = dby - (bhy + hhx * s + hhy * c);
    if (t > g) g = t;
    return g;
}

/* separation of a convex polygon hull (robot frame, robot at (px, py) with heading cos ct / sin st) from one box: max over
 * the box's two axes and the polygon's edge normals (unit length) of the gap between the two projection intervals --
 * the same definition as sat_pair_gap, whose four axes are the special case of a rectangle. */
static inline double sat_poly_box_gap(const double* poly, double px, double py, double ct, double st, const double* box) {
    const int n = (int)poly[0];
    const double cb = box[2], sb = box[3], bhx = box[4], bhy = box[5];
    double bx[SAT_POLY_MAX_VERTICES], by[SAT_POLY_MAX_VERTICES];
    double minx = DBL_MAX, maxx = -DBL_MAX, miny = DBL_MAX, maxy = -DBL_MAX;
    for (int i = 0; i < n; ++i) {   /* vertices in the box frame */
        const double vx = poly[3 + 2 * i], vy = poly[4 + 2 * i];
        const double wx = px + (ct * vx - st * vy) - box[0], wy = py + (st * vx + ct * vy) - box[1];
        bx[i] = cb * wx + sb * wy;
        by[i] = cb * wy - sb * wx;
        if (bx[i] < minx) minx = bx[i];
        if (bx[i] > maxx) maxx = bx[i];
        if (by[i] < miny) miny = by[i];
        if (by[i] > maxy) maxy = by[i];
    }
    double g = minx - bhx, t = -bhx - maxx;
    if (t > g) g = t;
    t = miny - bhy; if (t > g) g = t;
    t = -bhy - maxy; if (t > g) g = t;
    for (int i = 0; i < n; ++i) {   /* edge i -> i + 1, outward normal of a counter-clockwise polygon */
        const int j = (i + 1) % n;
        const double ex = bx[j] - bx[i], ey = by[j] - by[i];
        const double len = sqrt(ex * ex + ey * ey);
        if (!(len > 0)) continue;
        const double nx = ey / len, ny = -ex / len;
        double hmin = DBL_MAX, hmax = -DBL_MAX;
        for (int k = 0; k < n; ++k) {
            const double d = nx * bx[k] + ny * by[k];
            if (d < hmin) hmin = d;
            if (d > hmax) hmax = d;
        }
        const double r = bhx * fabs(nx) + bhy * fabs(ny);
        t = hmin - r; if (t > g) g = t;
        t = -r - hmax; if (t > g) g = t;
    }
    return g;
}

/* signed clearance of the robot at pose (x, y, theta): min over live pairs of the pair gap.
 * Returns +DBL_MAX when there is no live pair. */
static inline double sat_clearance(const sat_scene* sc, const double pose[3]) {
    const double ct = cos(pose[2]), st = sin(pose[2]);
    double best = DBL_MAX;
    for (int h = 0; h < sc->n_hulls; ++h) {
        const double* hu = sc->hulls + h * SAT_PRISM_STRIDE;
        const double hcx = pose[0] + (ct * hu[0] - st * hu[1]);
        const double hcy = pose[1] + (st * hu[0] + ct * hu[1]);
        const double ch = ct * hu[2] - st * hu[3];
        const double sh = st * hu[2] + ct * hu[3];
        for (int b = 0; b < sc->n_boxes; ++b) {
            const double* bx = sc->boxes + b * SAT_PRISM_STRIDE;
            if (!sat_z_overlap(hu, bx)) continue;
            const double g = sat_pair_gap(hcx, hcy, ch, sh, hu[4], hu[5], bx);
            if (g < best) best = g;
        }
    }
    for (int h = 0; h < sc->n_poly_hulls; ++h) {
        const double* po = sc->poly_hulls + h * SAT_POLY_STRIDE;
        for (int b = 0; b < sc->n_boxes; ++b) {
            const double* bx = sc->boxes + b * SAT_PRISM_STRIDE;
            if (!((po[1] < bx[7]) && (bx[6] < po[2]))) continue;   /* z-intervals must overlap */
            const double g = sat_poly_box_gap(po, pose[0], pose[1], ct, st, bx);
            if (g < best) best = g;
        }
    }
    return best;
}

static inline int sat_collides(const sat_scene* sc, const double pose[3]) {
    return sat_clearance(sc, pose) <= 0.0;
}

#endif
