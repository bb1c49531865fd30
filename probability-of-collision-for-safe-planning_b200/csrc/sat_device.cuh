// sat_device.cuh -- 2.5-D separating-axis test of the robot hull set against the scene prisms.
//
// Replaces OpenRAVE's EnvironmentBase::CheckCollision(robot) as the reference calls it once per
// particle and waypoint (mcsimplugin/MCSimulator.h:257-285, :333-347, :241-253).  The obstacle
// prisms come from the env XML (pr2custom.env.xml:14-238, pr2test2.env.xml:14-118), the robot
// hulls from the documented fixture.  Geometry convention is the one of
// oracle/sat_cpu.h: a pair collides iff max over the 4 face normals of
// (|projected centre distance| - r_hull - r_box) <= 0.
//
// Table layout (FP32, __constant__): live (hull, box) pairs grouped by hull, 32 B per pair:
//     lo = (cx, cy, cos phi, sin phi)   hi = (hx, hy, -p, -q),  p = cb cx + sb cy,  q = cb cy - sb cx
// p and q fold the box-frame projection of the box centre, so the four centre projections cost
// 8 FFMA and no subtraction of centres (see DESIGN.md "SAT instruction count").
#ifndef MCSIM_SAT_DEVICE_CUH
#define MCSIM_SAT_DEVICE_CUH

#include "mcsim_cuda.h"

struct HullRec {
    float ox, oy, cp, sp;  // centre and axis in the robot frame
    float hx, hy;          // half extents
    int first, count;      // slice of c_pairs
    int g_first, n_groups; // slice of the orientation-grouped table (float4 units; mc_philox2_sm100.cuh)
};

// Robot hulls beyond boxes: a convex polygon of up to 8 vertices in the robot frame, extruded in z.
// Per edge the outward unit normal and the polygon's extent along it (max = the edge's own offset, min = the far side)
// are folded on the host, so the device rotates 8 normals instead of projecting 8 x 8 vertices.
#define MCSIM_POLY_MAX_VERTICES 8
struct PolyHullRec {
    int n;                 // vertices
    int first, count;      // slice of c_pairs: the boxes whose z-range meets the hull's
    int pad;
    float vx[MCSIM_POLY_MAX_VERTICES], vy[MCSIM_POLY_MAX_VERTICES];
    float nx[MCSIM_POLY_MAX_VERTICES], ny[MCSIM_POLY_MAX_VERTICES];      // outward edge normals (unit)
    float dmax[MCSIM_POLY_MAX_VERTICES], dmin[MCSIM_POLY_MAX_VERTICES];  // extent of the polygon along each normal
};

struct SceneConst {
    int n_hulls;
    int n_pairs;           // live (box hull, box) pairs: the packed kernels' tables cover exactly these
    float reach;           // max over hulls of |offset| + circumradius: robot bounding circle
    int n_poly;            // polygon hulls (scalar kernels only; their pairs follow the box hulls' in c_pairs)
    HullRec hulls[MCSIM_MAX_HULLS];
    PolyHullRec poly[MCSIM_MAX_POLY_HULLS];
};

__constant__ SceneConst c_scene;
__constant__ float4 c_pairs[2 * MCSIM_MAX_PAIRS];

struct HullPose {
    float ch, sh;    // world axis of the hull
    float ph, qh;    // ch*hcx + sh*hcy, ch*hcy - sh*hcx
    float hcx, hcy;  // world centre
    float hx, hy;
};

__device__ __forceinline__ HullPose place_hull(const HullRec& h, float x, float y, float ct, float st) {
    HullPose o;
    o.hcx = x + (ct * h.ox - st * h.oy);
    o.hcy = y + (st * h.ox + ct * h.oy);
    o.ch = ct * h.cp - st * h.sp;
    o.sh = st * h.cp + ct * h.sp;
    o.ph = o.ch * o.hcx + o.sh * o.hcy;
    o.qh = o.ch * o.hcy - o.sh * o.hcx;
    o.hx = h.hx;
    o.hy = h.hy;
    return o;
}

// gap of one pair: max over the four axes; <= 0 means overlap
__device__ __forceinline__ float pair_gap(const HullPose& h, const float4 lo, const float4 hi) {
    const float c = fabsf(h.ch * lo.z + h.sh * lo.w);
    const float s = fabsf(h.sh * lo.z - h.ch * lo.w);
    // centre distance along hull axes and along box axes
    const float drx = fabsf(fmaf(h.ch, lo.x, fmaf(h.sh, lo.y, -h.ph)));
    const float dry = fabsf(fmaf(h.ch, lo.y, fmaf(-h.sh, lo.x, -h.qh)));
    const float dbx = fabsf(fmaf(lo.z, h.hcx, fmaf(lo.w, h.hcy, hi.z)));
    const float dby = fabsf(fmaf(lo.z, h.hcy, fmaf(-lo.w, h.hcx, hi.w)));
    float g = drx - fmaf(hi.x, c, fmaf(hi.y, s, h.hx));
    g = fmaxf(g, dry - fmaf(hi.x, s, fmaf(hi.y, c, h.hy)));
    g = fmaxf(g, dbx - fmaf(h.hx, c, fmaf(h.hy, s, hi.x)));
    g = fmaxf(g, dby - fmaf(h.hx, s, fmaf(h.hy, c, hi.y)));
    return g;
}

// overlap predicate of one pair (no gap value: 4 compares instead of 4 subtract + 3 max)
__device__ __forceinline__ bool pair_hit(const HullPose& h, const float4 lo, const float4 hi) {
    const float c = fabsf(h.ch * lo.z + h.sh * lo.w);
    const float s = fabsf(h.sh * lo.z - h.ch * lo.w);
    const float drx = fabsf(fmaf(h.ch, lo.x, fmaf(h.sh, lo.y, -h.ph)));
    const float dry = fabsf(fmaf(h.ch, lo.y, fmaf(-h.sh, lo.x, -h.qh)));
    const float dbx = fabsf(fmaf(lo.z, h.hcx, fmaf(lo.w, h.hcy, hi.z)));
    const float dby = fabsf(fmaf(lo.z, h.hcy, fmaf(-lo.w, h.hcx, hi.w)));
    return (drx <= fmaf(hi.x, c, fmaf(hi.y, s, h.hx))) & (dry <= fmaf(hi.x, s, fmaf(hi.y, c, h.hy))) &
           (dbx <= fmaf(h.hx, c, fmaf(h.hy, s, hi.x))) & (dby <= fmaf(h.hx, s, fmaf(h.hy, c, hi.y)));
}

// gap of one (polygon hull, box) pair: max over the box's two axes and the polygon's edge normals of the separation of
// the projection intervals (oracle/sat_cpu.h sat_poly_box_gap); lo / hi: the pair record of the box
__device__ __forceinline__ float poly_pair_gap(const PolyHullRec& h, float x, float y, float ct, float st, const float4 lo, const float4 hi) {
    float minx = 3.4e38f, maxx = -3.4e38f, miny = 3.4e38f, maxy = -3.4e38f;
    for (int i = 0; i < h.n; ++i) {   // vertices in the box frame
        const float wx = x + (ct * h.vx[i] - st * h.vy[i]) - lo.x, wy = y + (st * h.vx[i] + ct * h.vy[i]) - lo.y;
        const float bx = fmaf(lo.z, wx, lo.w * wy), by = fmaf(lo.z, wy, -lo.w * wx);
        minx = fminf(minx, bx); maxx = fmaxf(maxx, bx);
        miny = fminf(miny, by); maxy = fmaxf(maxy, by);
    }
    float g = fmaxf(fmaxf(minx - hi.x, -hi.x - maxx), fmaxf(miny - hi.y, -hi.y - maxy));
    const float rx = lo.x - x, ry = lo.y - y;   // box centre seen from the robot origin
    for (int i = 0; i < h.n; ++i) {
        const float nwx = ct * h.nx[i] - st * h.ny[i], nwy = st * h.nx[i] + ct * h.ny[i];   // normal in the world
        const float cproj = fmaf(nwx, rx, nwy * ry);                                        // box centre along it, robot origin = 0
        const float nbx = fmaf(lo.z, nwx, lo.w * nwy), nby = fmaf(lo.z, nwy, -lo.w * nwx);  // normal in the box frame
        const float r = fmaf(hi.x, fabsf(nbx), hi.y * fabsf(nby));
        g = fmaxf(g, fmaxf(h.dmin[i] - (cproj + r), (cproj - r) - h.dmax[i]));
    }
    return g;
}
__device__ __forceinline__ float poly_clearance(float x, float y, float ct, float st, unsigned& n_tests) {
    float best = 3.402823466e+38f;
    for (int hI = 0; hI < c_scene.n_poly; ++hI) {
        const PolyHullRec& h = c_scene.poly[hI];
        const int end = h.first + h.count;
        for (int b = h.first; b < end; ++b) best = fminf(best, poly_pair_gap(h, x, y, ct, st, c_pairs[2 * b], c_pairs[2 * b + 1]));
        n_tests += h.count;
    }
    return best;
}

// Exhaustive test of every live pair: the reference semantics (one CheckCollision = robot vs all
// bodies).  Returns the collision flag; n_tests counts executed pair tests.
__device__ __forceinline__ bool robot_hits(float x, float y, float ct, float st, unsigned& n_tests) {
    bool hit = c_scene.n_poly ? poly_clearance(x, y, ct, st, n_tests) <= 0.f : false;
    for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const HullPose hp = place_hull(hr, x, y, ct, st);
        const int end = hr.first + hr.count;
#pragma unroll 4
        for (int b = hr.first; b < end; ++b) hit |= pair_hit(hp, c_pairs[2 * b], c_pairs[2 * b + 1]);
        n_tests += hr.count;
    }
    return hit;
}

// The same exhaustive test on the orientation-grouped table (built by mcsim_cuda_set_scene; layout and the
// algebra are described at robot_hits2 in mc_philox2_sm100.cuh): per class of boxes sharing an axis direction
// mod 90 degrees the relative rotation, the hull centre in the class frame and the hull's radii on the class
// axes are computed once, which leaves 12 FP32 operations + 4 compares per pair instead of 20 + 4.
// tab: global (read through the read-only path; every thread reads the same addresses) or shared memory.
struct HullFrame1 { float hcx, hcy, ch, sh, nph, nqh, hhx, hhy; };
struct ClassFrame1 { float c, s, gx, gy, rbx, rby; };
__device__ __forceinline__ HullFrame1 hull_frame1(const HullRec& hr, float x, float y, float ct, float st) {
    HullFrame1 h;
    h.hcx = fmaf(ct, hr.ox, fmaf(-st, hr.oy, x));
    h.hcy = fmaf(st, hr.ox, fmaf(ct, hr.oy, y));
    h.ch = fmaf(ct, hr.cp, -st * hr.sp);
    h.sh = fmaf(st, hr.cp, ct * hr.sp);
    h.nph = -fmaf(h.ch, h.hcx, h.sh * h.hcy);
    h.nqh = -fmaf(h.ch, h.hcy, -h.sh * h.hcx);
    h.hhx = hr.hx;
    h.hhy = hr.hy;
    return h;
}
__device__ __forceinline__ ClassFrame1 class_frame1(const HullFrame1& h, const float4 head) {
    ClassFrame1 k;
    k.c = fabsf(fmaf(h.ch, head.x, h.sh * head.y));
    k.s = fabsf(fmaf(h.sh, head.x, -h.ch * head.y));
    k.gx = fmaf(h.hcx, head.x, h.hcy * head.y);
    k.gy = fmaf(h.hcy, head.x, -h.hcx * head.y);
    k.rbx = fmaf(k.c, h.hhx, k.s * h.hhy);
    k.rby = fmaf(k.s, h.hhx, k.c * h.hhy);
    return k;
}
__device__ __forceinline__ bool class_pair_hit1(const HullFrame1& h, const ClassFrame1& k, const float4 A, const float2 B) {
    const float drx = fmaf(h.ch, A.x, fmaf(h.sh, A.y, h.nph));
    const float dry = fmaf(h.ch, A.y, fmaf(-h.sh, A.x, h.nqh));
    const float r0 = fmaf(k.c, A.z, fmaf(k.s, A.w, h.hhx));
    const float r1 = fmaf(k.s, A.z, fmaf(k.c, A.w, h.hhy));
    return (fabsf(drx) <= r0) & (fabsf(dry) <= r1) & (fabsf(k.gx + B.x) <= k.rbx + A.z) & (fabsf(k.gy + B.y) <= k.rby + A.w);
}

// SMEM = true: tab points into shared memory (plain loads); false: global memory through the read-only path
template <bool SMEM>
__device__ __forceinline__ float4 tab_ld4(const float4* p) { return SMEM ? *p : __ldg(p); }
template <bool SMEM>
__device__ __forceinline__ float2 tab_ld2(const float4* p) { return SMEM ? *reinterpret_cast<const float2*>(p) : __ldg(reinterpret_cast<const float2*>(p)); }

template <bool SMEM = false>
__device__ __forceinline__ bool robot_hits_grouped(const float4* __restrict__ tab, float x, float y, float ct, float st) {
    bool hit = false;
    for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const HullFrame1 h = hull_frame1(hr, x, y, ct, st);
        const float4* g = tab + hr.g_first;
        for (int gi = 0; gi < hr.n_groups; ++gi) {
            const float4 head = tab_ld4<SMEM>(g);
            const int cnt = __float_as_int(head.z);
            const ClassFrame1 k = class_frame1(h, head);
            ++g;
#pragma unroll 4
            for (int b = 0; b < cnt; ++b)
                if (class_pair_hit1(h, k, tab_ld4<SMEM>(g + 2 * b), tab_ld2<SMEM>(g + 2 * b + 1))) hit = true;
            g += 2 * cnt;
        }
    }
    return hit;
}

// ---- exact broad phase (candidate lists built on the host per plan waypoint: mcsim_cuda.cu build_cull_lists) -------
// Per waypoint t and hull h the pairs whose box footprint lies within reach_h + m of the nominal position x*_t,
// sorted by the smallest such m, and how many of them are needed for the margins m_k of the MC_CULL_LEVELS levels (mc_cull_margin_sq below).
// A hull whose robot pose is within m of x*_t cannot touch any other box (every hull point is within reach_h of
// the pose), so a warp whose poses are all within m_k of x*_t tests the first cnt[t][h][k] listed pairs; a warp
// with a pose further out than the last level (or a list longer than MC_CULL_CAP) tests every live pair.  The
// listed pairs go through the SAME arithmetic as the exhaustive loop (class frame + pair test on the grouped
// table), so flags and counts are bit-identical to the exhaustive test's; executed pair tests are counted.
// Table (32-bit words): cnt[T][n_hulls][MC_CULL_LEVELS] bytes (255 = list too long), padded to 4 bytes, then
// idx[T][n_hulls][MC_CULL_CAP] x (16-bit pair record offset, 16-bit class header offset) into the grouped table.
#ifndef MC_CULL_SUB
#define MC_CULL_SUB 2  // mantissa bits of the squared deviation that enter the level (2^SUB levels per octave of d^2)
#endif
#define MC_CULL_LEVELS (8 << MC_CULL_SUB)
#define MC_CULL_CAP 16
#define MC_CULL_OVERFLOW 255
// squared margin of level k = (1 + (k mod 2^SUB) / 2^SUB) * 2^(k / 2^SUB - 7) m^2: the float whose bit pattern is
// (MC_CULL_B0 + k) << (23 - SUB).  SUB = 2: 32 levels, m = 0.088, 0.099, 0.108, 0.117, 0.125, ... 1.0, 1.12, 1.22, 1.32 m
#define MC_CULL_B0 (120u << MC_CULL_SUB)
static inline double mc_cull_margin_sq(int k) { return ldexp(1.0 + double(k & ((1 << MC_CULL_SUB) - 1)) / double(1 << MC_CULL_SUB), (k >> MC_CULL_SUB) - 7); }

// level of a squared deviation = smallest k with d2max <= margin_sq(k), read off the exponent and leading mantissa
// bits (exact: no rounding between this test and the host's margins); beyond the last level, NaN and negative
// patterns -> exhaustive
__device__ __forceinline__ int cull_level(float d2max) {
    const unsigned b = __float_as_uint(d2max);
    const unsigned u = ((b ? b : 1u) - 1u) >> (23 - MC_CULL_SUB);  // patterns in ((B0 + k - 1) << s, (B0 + k) << s] -> B0 + k - 1
    const unsigned k = u >= MC_CULL_B0 ? u - (MC_CULL_B0 - 1u) : 0u;
    return k < unsigned(MC_CULL_LEVELS) ? int(k) : MC_CULL_LEVELS;
}

// listed pairs of waypoint t at level lev; cnt / idx point at the waypoint's slices ([n_hulls][LEVELS], [n_hulls][CAP])
template <bool SMEM = false>
__device__ __forceinline__ bool robot_hits_listed(const float4* __restrict__ tab, const unsigned char* cnt, const ushort2* idx, int lev,
                                                  float x, float y, float ct, float st, unsigned& tested) {
    bool hit = false;
    for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
        const int n = cnt[hI * MC_CULL_LEVELS + lev];
        if (n == 0) continue;
        const HullFrame1 h = hull_frame1(c_scene.hulls[hI], x, y, ct, st);
        for (int k = 0; k < n; ++k) {
            const ushort2 o = idx[hI * MC_CULL_CAP + k];
            const ClassFrame1 cf = class_frame1(h, tab_ld4<SMEM>(tab + o.y));
            if (class_pair_hit1(h, cf, tab_ld4<SMEM>(tab + o.x), tab_ld2<SMEM>(tab + o.x + 1))) hit = true;
        }
        tested += n;
    }
    // polygon hulls follow the box hulls in the lists; an entry is the pair's index in c_pairs
    for (int pI = 0; pI < c_scene.n_poly; ++pI) {
        const int hI = c_scene.n_hulls + pI;
        const int n = cnt[hI * MC_CULL_LEVELS + lev];
        for (int k = 0; k < n; ++k) {
            const int b = idx[hI * MC_CULL_CAP + k].x;
            if (poly_pair_gap(c_scene.poly[pI], x, y, ct, st, c_pairs[2 * b], c_pairs[2 * b + 1]) <= 0.f) hit = true;
        }
        tested += n;
    }
    return hit;
}

// Signed clearance: min over live pairs of the pair gap (+FLT_MAX when there is no pair).
__device__ __forceinline__ float robot_clearance(float x, float y, float ct, float st) {
    unsigned unused = 0;
    float best = c_scene.n_poly ? poly_clearance(x, y, ct, st, unused) : 3.402823466e+38f;
    for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const HullPose hp = place_hull(hr, x, y, ct, st);
        const int end = hr.first + hr.count;
        for (int b = hr.first; b < end; ++b) best = fminf(best, pair_gap(hp, c_pairs[2 * b], c_pairs[2 * b + 1]));
    }
    return best;
}

#endif
