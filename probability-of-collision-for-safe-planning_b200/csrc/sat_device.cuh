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

struct SceneConst {
    int n_hulls;
    int n_pairs;
    float reach;           // max over hulls of |offset| + circumradius: robot bounding circle
    float pad;
    HullRec hulls[MCSIM_MAX_HULLS];
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

// Exhaustive test of every live pair: the reference semantics (one CheckCollision = robot vs all
// bodies).  Returns the collision flag; n_tests counts executed pair tests.
__device__ __forceinline__ bool robot_hits(float x, float y, float ct, float st, unsigned& n_tests) {
    bool hit = false;
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
__device__ __forceinline__ bool robot_hits_grouped(const float4* __restrict__ tab, float x, float y, float ct, float st) {
    bool hit = false;
    for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const float hcx = fmaf(ct, hr.ox, fmaf(-st, hr.oy, x));
        const float hcy = fmaf(st, hr.ox, fmaf(ct, hr.oy, y));
        const float ch = fmaf(ct, hr.cp, -st * hr.sp);
        const float sh = fmaf(st, hr.cp, ct * hr.sp);
        const float nph = -fmaf(ch, hcx, sh * hcy);
        const float nqh = -fmaf(ch, hcy, -sh * hcx);
        const float hhx = hr.hx, hhy = hr.hy;
        const float4* g = tab + hr.g_first;
        for (int gi = 0; gi < hr.n_groups; ++gi) {
            const float4 head = __ldg(g);
            const int cnt = __float_as_int(head.z);
            const float c = fabsf(fmaf(ch, head.x, sh * head.y));
            const float s = fabsf(fmaf(sh, head.x, -ch * head.y));
            const float gx = fmaf(hcx, head.x, hcy * head.y);
            const float gy = fmaf(hcy, head.x, -hcx * head.y);
            const float rbx = fmaf(c, hhx, s * hhy);
            const float rby = fmaf(s, hhx, c * hhy);
            ++g;
#pragma unroll 4
            for (int b = 0; b < cnt; ++b) {
                const float4 A = __ldg(g + 2 * b);                                        // cx, cy, hx, hy
                const float2 B = __ldg(reinterpret_cast<const float2*>(g + 2 * b + 1));   // -p, -q
                const float drx = fmaf(ch, A.x, fmaf(sh, A.y, nph));
                const float dry = fmaf(ch, A.y, fmaf(-sh, A.x, nqh));
                const float r0 = fmaf(c, A.z, fmaf(s, A.w, hhx));
                const float r1 = fmaf(s, A.z, fmaf(c, A.w, hhy));
                if ((fabsf(drx) <= r0) & (fabsf(dry) <= r1) & (fabsf(gx + B.x) <= rbx + A.z) & (fabsf(gy + B.y) <= rby + A.w)) hit = true;
            }
            g += 2 * cnt;
        }
    }
    return hit;
}

// Signed clearance: min over live pairs of the pair gap (+FLT_MAX when there is no pair).
__device__ __forceinline__ float robot_clearance(float x, float y, float ct, float st) {
    float best = 3.402823466e+38f;
    for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const HullPose hp = place_hull(hr, x, y, ct, st);
        const int end = hr.first + hr.count;
        for (int b = hr.first; b < end; ++b) best = fminf(best, pair_gap(hp, c_pairs[2 * b], c_pairs[2 * b + 1]));
    }
    return best;
}

#endif
