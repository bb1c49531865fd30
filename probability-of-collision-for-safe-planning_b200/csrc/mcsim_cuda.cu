// mcsim_cuda.cu -- the C ABI of include/mcsim_cuda.h over the sm_100a kernels.  One translation
// unit: the kernels live in the .cuh files included below and share the __constant__ tables.
// There is no CPU path: every entry point fails with MCSIM_ERR_NODEVICE without a CUDA device.
#include "mcsim_cuda.h"

#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <thread>
#include <algorithm>
#include <chrono>
#include <string>
#include <utility>
#include <vector>

#include "mc_rollout_sm100.cuh"
#include "mc_philox2_sm100.cuh"
#include "gmm_sm100.cuh"
#include "em_sm100.cuh"
#include "score_sm100.cuh"
#include "peaks_sm100.cuh"

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) return fail(MCSIM_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// ---- ownership of the __constant__ tables ---------------------------------------------------------------------------
// c_scene / c_pairs (scene) and c_mc / c_plan2 (prepared plan) are one set of symbols per DEVICE, while every context
// (and every MCModule) has its own scene and plan.  Each context keeps host copies of what it wants in the symbols and a
// tag per table; the per-device registry remembers whose data the symbols hold.  A launch path takes the registry lock,
// re-uploads when the tags differ (after a device-wide synchronise: a kernel of the previous owner may still be reading
// them) and keeps the lock across its launches, so contexts sharing a device -- also from different host threads --
// cannot run on each other's tables.
struct DeviceSymbols {
    std::mutex m;
    uint64_t scene_tag = 0, plan_tag = 0;
    uint64_t owner = 0;            // context id of the last writer (its own stream orders its own re-uploads)
};
static DeviceSymbols g_symbols[64];
static std::atomic<uint64_t> g_next_tag{1};
#define MCSIM_STAGE_BYTES (64 * 1024)
struct PlanSymbols {           // what a prepare call wants in c_mc / c_plan2
    McPlanConst pc;
    std::vector<float> plan2;
    uint64_t tag = 0;
};

struct mcsim_cuda_ctx {
    int device = 0;
    uint64_t id = g_next_tag++;
    SceneConst h_scene;                 // host copies of this context's symbol contents
    std::vector<float4> h_pairs;
    uint64_t scene_tag = 0;
    PlanSymbols mc_sym, sc_sym;         // prepared MC plan / prepared scoring batch
    int n_sm = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t launches = 0;
    bool scene_set = false;
    bool scene_axis_aligned = false;   // every box class axis (1, 0) and every box hull's yaw 0: the packed kernels' AX instantiations
    int n_pairs = 0;
    // MC state
    DevBuf plan, z0, counters, counts, poses, clearance;
    unsigned long long* h_counters = nullptr;  // pinned
    mcsim_mc_params mc;                        // prepared parameters (pointers not retained)
    bool mc_prepared = false;
    int mc_T = 0;
    // FP64 scene for the replay-mode GMM SAT
    DevBuf hulls_d, pairs_d;
    int n_hulls = 0;
    int n_poly = 0;    // convex-polygon robot hulls (scalar kernels only)
    // orientation-grouped FP32 pair table of the packed Philox kernel (staged into shared memory)
    DevBuf sat2;
    int sat2_n = 0;  // float4 records
    // exact broad phase of the packed Philox kernel: host copy of the live pairs, per-plan candidate lists
    std::vector<double> h_pairgeom;  // per pair: cx, cy, cos, sin, hx, hy
    std::vector<int> h_pair_hull;    // per pair: hull index
    // polygon hulls' pairs (they follow the box hulls' in c_pairs): box geometry, polygon hull, index in c_pairs; reach per polygon
    std::vector<double> h_poly_pairgeom;
    std::vector<int> h_poly_pair_hull, h_poly_pair_index;
    double h_poly_reach[MCSIM_MAX_POLY_HULLS] = {0};
    std::vector<unsigned> h_pair_sat2;  // per pair: record offset | class header offset << 16 in the grouped table
    double h_reach[MCSIM_MAX_HULLS] = {0};
    DevBuf mc_cull, mc_nominal;
    bool mc_cull_fits_smem = false;
    DevBuf seg_state[2], seg_alive;  // early exit with survivor compaction: double-buffered state of the survivors, counts per segment
    int gmm_cull_words = 0, gmm_cull_words_cached = 0;  // words of ONE plan's lists (this run / on the device)
    // the lists depend on (scene, trajectory) only: repeated runs of one plan (the reference's 200-run loop) reuse them
    uint64_t scene_epoch = 0;
    std::vector<double> mc_cull_key, gmm_cull_key;   // scene epoch, T, trajectory x / y of the lists on the device
    std::vector<unsigned> mc_cull_host;
    bool mc_cull_uploaded = false;   // by the last mc_prepare (h2d byte accounting)
    DevBuf gmm_plans, gmm_robot_normals;      // batch: the plan(s) and the scratch normals of the device robot threads
    DevBuf gmm_cull, gmm_nominal, gmm_tests;  // GMM Philox mode: the plan's (or every batch member's) lists
    int mc_cull_words = 0;           // 0: no lists for the prepared plan (every live pair is tested)
    // GMM state
    DevBuf gmm_comps, gmm_stats, gmm_hist, gmm_counts, gmm_normals, gmm_steps, gmm_prob, gmm_pre, gmm_post, gmm_trunc, gmm_nout, gmm_deg;
    DevBuf gmm_free, gmm_free_u, gmm_free_flag, gmm_free_n, lin_A, lin_b;   // EM-refit variant: free samples of a waypoint; linear variant: half-planes
    double* h_gmm = nullptr;  // pinned staging
    char* h_stage = nullptr;  // pinned staging of the per-run plan upload (MCSIM_STAGE_BYTES)
    // scoring state
    DevBuf pairs_f, sc_plans, sc_hits, sc_counters;
    ScoreLaunch sc;
    bool sc_prepared = false;
    int sc_ee = 0, sc_est = 0;
    // EM state
    DevBuf em_x, em_comps, em_hefts, em_means, em_fcovs, em_stats, em_old, em_done, em_iters, em_trace;
    int em_moments = EM_MOMENTS_AUTO;   // reference point of the EM statistics (mcsim_cuda_em_set_moments)
    // E-step launch configuration of the prepared fit (kernel choice, occupancy query and wave fitting done once, not per iteration)
    void (*em_k0)(const EmLaunch) = nullptr;
    void (*em_k1)(const EmLaunch) = nullptr;
    size_t em_smem = 0;
    unsigned em_chunks = 0;
    int em_block = 256;
    DevBuf em_slots, em_flags;   // fused exchange of a sharded fit (mcsim_group.cuh): gather buffer [2][G][n_stats], counters [n_fits][G]
    EmLaunch em;
    bool em_prepared = false;
    bool em_x_valid = false;
    uint64_t em_x_n = 0;
    int em_x_fits = 0;
    uint64_t h2d_bytes = 0, d2h_bytes = 0;  // bytes copied by the last run call
};

// Makes the device's symbols hold this context's scene (and, when `plan` is given, that plan) and returns holding the
// device's registry lock; the caller launches its kernels and lets the lock go.
struct SymbolLock {
    std::unique_lock<std::mutex> lk;
};
static int bind_symbols(mcsim_cuda_ctx* c, const PlanSymbols* plan, SymbolLock& out) {
    DeviceSymbols& d = g_symbols[c->device & 63];
    out.lk = std::unique_lock<std::mutex>(d.m);
    const bool scene_stale = d.scene_tag != c->scene_tag, plan_stale = plan && d.plan_tag != plan->tag;
    if (!scene_stale && !plan_stale) return MCSIM_OK;
    if (d.owner == c->id) CK(cudaStreamSynchronize(c->stream));   // only this context's kernels can be reading them
    else CK(cudaDeviceSynchronize());
    d.owner = c->id;
    if (scene_stale) {
        CK(cudaMemcpyToSymbol(c_scene, &c->h_scene, sizeof(SceneConst)));
        if (!c->h_pairs.empty()) CK(cudaMemcpyToSymbol(c_pairs, c->h_pairs.data(), c->h_pairs.size() * sizeof(float4)));
        d.scene_tag = c->scene_tag;
    }
    if (plan_stale) {
        CK(cudaMemcpyToSymbol(c_mc, &plan->pc, sizeof(McPlanConst)));
        if (!plan->plan2.empty()) CK(cudaMemcpyToSymbol(c_plan2, plan->plan2.data(), plan->plan2.size() * sizeof(float)));
        d.plan_tag = plan->tag;
    }
    return MCSIM_OK;
}

extern "C" const char* mcsim_cuda_last_error(void) { return g_err; }

extern "C" int mcsim_cuda_create(int device, mcsim_cuda_ctx** out) {
    if (!out) return fail(MCSIM_ERR_ARG, "out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(MCSIM_ERR_NODEVICE, "no CUDA device (%s): this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
    if (device < 0 || device >= n) return fail(MCSIM_ERR_ARG, "device %d out of range (0..%d)", device, n - 1);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(MCSIM_ERR_NODEVICE, "device %d is sm_%d%d; this build is sm_100a only", device, prop.major, prop.minor);
    mcsim_cuda_ctx* c = new mcsim_cuda_ctx();
    c->device = device;
    c->n_sm = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&c->ev0));
    CK(cudaEventCreate(&c->ev1));
    CK(cudaMallocHost((void**)&c->h_counters, sizeof(unsigned long long) * (4 + MCSIM_MAX_WAYPOINTS)));
    CK(cudaMallocHost((void**)&c->h_gmm, sizeof(double) * 4096));
    CK(cudaMallocHost((void**)&c->h_stage, MCSIM_STAGE_BYTES));
    CK(c->counters.reserve(sizeof(unsigned long long) * (4 + MCSIM_MAX_WAYPOINTS)));
    CK(cudaMemset(c->counters.p, 0, sizeof(unsigned long long) * (4 + MCSIM_MAX_WAYPOINTS)));
    *out = c;
    return MCSIM_OK;
}

extern "C" void mcsim_cuda_destroy(mcsim_cuda_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->plan.release(); c->z0.release(); c->counters.release(); c->counts.release();
    c->poses.release(); c->clearance.release();
    c->seg_state[0].release(); c->seg_state[1].release(); c->seg_alive.release(); c->mc_nominal.release();
    c->hulls_d.release(); c->pairs_d.release(); c->sat2.release(); c->mc_cull.release(); c->gmm_cull.release(); c->gmm_nominal.release(); c->gmm_tests.release(); c->gmm_plans.release(); c->gmm_robot_normals.release();
    c->pairs_f.release(); c->sc_plans.release(); c->sc_hits.release(); c->sc_counters.release();
    c->em_x.release(); c->em_comps.release(); c->em_hefts.release(); c->em_means.release(); c->em_fcovs.release();
    c->em_stats.release(); c->em_old.release(); c->em_done.release(); c->em_iters.release(); c->em_trace.release();
    c->em_slots.release(); c->em_flags.release();
    c->gmm_comps.release(); c->gmm_stats.release(); c->gmm_hist.release(); c->gmm_counts.release();
    c->gmm_normals.release(); c->gmm_steps.release(); c->gmm_prob.release(); c->gmm_pre.release();
    c->gmm_post.release(); c->gmm_trunc.release(); c->gmm_nout.release(); c->gmm_deg.release();
    c->gmm_free.release(); c->gmm_free_u.release(); c->gmm_free_flag.release(); c->gmm_free_n.release(); c->lin_A.release(); c->lin_b.release();
    if (c->h_counters) cudaFreeHost(c->h_counters);
    if (c->h_gmm) cudaFreeHost(c->h_gmm);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" void* mcsim_cuda_stream(mcsim_cuda_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int mcsim_cuda_ctx_device(const mcsim_cuda_ctx* c) { return c ? c->device : -1; }
extern "C" uint64_t mcsim_cuda_launch_count(mcsim_cuda_ctx* c) { return c ? c->launches : 0; }

// ---- peaks ----------------------------------------------------------------------------------
// which: 0 = FFMA chain (TFLOP/s), 1 = MUFU chain (10^12 op/s), 2 = packed FFMA2 chain (TFLOP/s)
extern "C" int mcsim_cuda_measure_peak(mcsim_cuda_ctx* c, int which, double ms, double* out) {
    if (!c || !out || which < 0 || which > 2) return fail(MCSIM_ERR_ARG, "measure_peak: bad arguments");
    CK(cudaSetDevice(c->device));
    float* d = nullptr;
    CK(cudaMalloc(&d, 64));
    const int grid = c->n_sm * 8, block = 256;
    int iters = 4096;
    double rate = 0;
    for (int rep = 0; rep < 4; ++rep) {  // rep 0 warms up, later reps are sized to ~ms
        CK(cudaEventRecord(c->ev0, c->stream));
        if (which == 0) fma_peak_kernel<<<grid, block, 0, c->stream>>>(d, iters, 0.999f, 0.001f);
        else if (which == 1) mufu_peak_kernel<<<grid, block, 0, c->stream>>>(d, iters);
        else fma2_peak_kernel<<<grid, block, 0, c->stream>>>(d, iters, 0.999f, 0.001f);
        ++c->launches;
        CK(cudaEventRecord(c->ev1, c->stream));
        CK(cudaStreamSynchronize(c->stream));
        float t = 0;
        CK(cudaEventElapsedTime(&t, c->ev0, c->ev1));
        const double per = which == 0 ? 2.0 : (which == 1 ? 1.0 : 4.0);
        const double ops = double(grid) * block * double(iters) * 8.0 * per;
        const double r = ops / (double(t) * 1e-3) / 1e12;
        if (rep > 0 && r > rate) rate = r;
        const double scale = (t > 0 ? ms / t : 8.0);
        iters = int(iters * (scale > 8 ? 8 : (scale < 0.125 ? 0.125 : scale)));
        if (iters < 64) iters = 64;
    }
    CK(cudaFree(d));
    *out = rate;
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_measure_peaks(mcsim_cuda_ctx* c, double ms, double* fp32_tflops, double* mufu_tops) {
    double v = 0;
    int rc;
    if (fp32_tflops) { rc = mcsim_cuda_measure_peak(c, 0, ms, &v); if (rc) return rc; *fp32_tflops = v; }
    if (mufu_tops) { rc = mcsim_cuda_measure_peak(c, 1, ms, &v); if (rc) return rc; *mufu_tops = v; }
    return MCSIM_OK;
}

// ---- scene ----------------------------------------------------------------------------------
static inline float __int_as_float_host(int v) { float f; memcpy(&f, &v, sizeof(f)); return f; }
static inline int __float_as_int_host(float f) { int v; memcpy(&v, &f, sizeof(v)); return v; }
extern "C" int mcsim_cuda_set_scene(mcsim_cuda_ctx* c, const double* boxes, int n_boxes,
                                    const double* hulls, int n_hulls) {
    return mcsim_cuda_set_scene_ex(c, boxes, n_boxes, hulls, n_hulls, nullptr, 0);
}

extern "C" int mcsim_cuda_set_scene_ex(mcsim_cuda_ctx* c, const double* boxes, int n_boxes, const double* hulls, int n_hulls,
                                       const double* poly_hulls, int n_poly) {
    if (!c || n_boxes < 0 || n_hulls < 0 || n_hulls > MCSIM_MAX_HULLS || (n_hulls && !hulls) || (n_boxes && !boxes) ||
        n_poly < 0 || n_poly > MCSIM_MAX_POLY_HULLS || (n_poly && !poly_hulls) || n_hulls + n_poly < 1)
        return fail(MCSIM_ERR_ARG, "set_scene: bad arguments (n_boxes=%d n_hulls=%d n_poly=%d)", n_boxes, n_hulls, n_poly);
    CK(cudaSetDevice(c->device));
    SceneConst sc;
    memset(&sc, 0, sizeof(sc));
    std::vector<float4> pairs;
    std::vector<float4> sat2;  // per hull: per orientation class: header (cb, sb, count, 0), then 2 float4 per pair
    std::vector<HullD> hd(n_hulls);
    std::vector<PairD> pd;
    sc.n_hulls = n_hulls;
    c->h_pairgeom.clear();
    c->h_pair_hull.clear();
    c->h_poly_pairgeom.clear();
    c->h_poly_pair_hull.clear();
    c->h_poly_pair_index.clear();
    c->h_pair_sat2.clear();
    ++c->scene_epoch;
    c->mc_prepared = false;  // candidate lists of a prepared plan refer to the old scene
    for (int h = 0; h < n_hulls; ++h) {
        const double* hu = hulls + h * MCSIM_PRISM_STRIDE;
        HullRec& r = sc.hulls[h];
        r.ox = float(hu[0]); r.oy = float(hu[1]); r.cp = float(hu[2]); r.sp = float(hu[3]);
        r.hx = float(hu[4]); r.hy = float(hu[5]);
        r.first = int(pairs.size() / 2);
        for (int b = 0; b < n_boxes; ++b) {
            const double* bx = boxes + b * MCSIM_PRISM_STRIDE;
            if (!((hu[6] < bx[7]) && (bx[6] < hu[7]))) continue;  // z-intervals must overlap (oracle/sat_cpu.h)
            const double p = bx[2] * bx[0] + bx[3] * bx[1];
            const double q = bx[2] * bx[1] - bx[3] * bx[0];
            pairs.push_back(make_float4(float(bx[0]), float(bx[1]), float(bx[2]), float(bx[3])));
            pairs.push_back(make_float4(float(bx[4]), float(bx[5]), -float(p), -float(q)));
            pd.push_back(PairD{bx[0], bx[1], bx[2], bx[3], bx[4], bx[5]});
            for (int k = 0; k < 6; ++k) c->h_pairgeom.push_back(bx[k]);
            c->h_pair_hull.push_back(h);
        }
        r.count = int(pairs.size() / 2) - r.first;
        // ---- the same live pairs grouped by orientation class.  A box (phi, hx, hy) is the box
        // (phi - k 90 deg, extents swapped for odd k): the axis is canonicalised into [0, 90 deg), so all
        // axis-aligned boxes of a scene share one class per hull.  Per class the kernel computes the
        // relative rotation, the hull centre in the class frame and the hull's radii on the class axes
        // once; per pair it is left with 12 packed FP32 instructions instead of 20.
        struct Canon { double cb, sb, cx, cy, hx, hy; };
        std::vector<Canon> live;
        for (int b = 0; b < n_boxes; ++b) {
            const double* bx = boxes + b * MCSIM_PRISM_STRIDE;
            if (!((hu[6] < bx[7]) && (bx[6] < hu[7]))) continue;
            double cb = bx[2], sb = bx[3], hx = bx[4], hy = bx[5];
            if (fabs(cb) < 1e-12) cb = 0;
            if (fabs(sb) < 1e-12) sb = 0;
            for (int k = 0; k < 4 && !(cb > 0 && sb >= 0); ++k) {  // rotate the frame by +90 deg: u' = v = (-sb, cb)
                const double t = cb; cb = -sb; sb = t;
                std::swap(hx, hy);
            }
            live.push_back(Canon{cb, sb, bx[0], bx[1], hx, hy});
        }
        r.g_first = int(sat2.size());
        r.n_groups = 0;
        std::vector<char> done(live.size(), 0);
        c->h_pair_sat2.resize(size_t(r.first) + live.size(), 0u);  // live[j] is pair r.first + j
        for (size_t i = 0; i < live.size(); ++i) {
            if (done[i]) continue;
            const size_t header = sat2.size();
            sat2.push_back(make_float4(float(live[i].cb), float(live[i].sb), 0.f, 0.f));
            int cnt = 0;
            for (size_t j = i; j < live.size(); ++j) {
                if (done[j] || fabs(live[j].cb - live[i].cb) > 1e-9 || fabs(live[j].sb - live[i].sb) > 1e-9) continue;
                done[j] = 1;
                const Canon& q = live[j];
                const double pp = live[i].cb * q.cx + live[i].sb * q.cy, qq = live[i].cb * q.cy - live[i].sb * q.cx;
                c->h_pair_sat2[size_t(r.first) + j] = unsigned(sat2.size()) | (unsigned(header) << 16);
                sat2.push_back(make_float4(float(q.cx), float(q.cy), float(q.hx), float(q.hy)));
                sat2.push_back(make_float4(-float(pp), -float(qq), 0.f, 0.f));
                ++cnt;
            }
            sat2[header].z = __int_as_float_host(cnt);
            ++r.n_groups;
        }
        hd[h] = HullD{hu[0], hu[1], hu[2], hu[3], hu[4], hu[5], r.first, r.count};
        const float reach = float(sqrt(hu[0] * hu[0] + hu[1] * hu[1]) + sqrt(hu[4] * hu[4] + hu[5] * hu[5]));
        if (reach > sc.reach) sc.reach = reach;
        c->h_reach[h] = sqrt(hu[0] * hu[0] + hu[1] * hu[1]) + sqrt(hu[4] * hu[4] + hu[5] * hu[5]);
    }
    sc.n_pairs = int(pairs.size() / 2);
    // ---- polygon hulls: records + their live pairs, appended behind the box hulls' pairs (scalar kernels only)
    sc.n_poly = n_poly;
    for (int h = 0; h < n_poly; ++h) {
        const double* po = poly_hulls + size_t(h) * MCSIM_POLY_STRIDE;
        const int nv = int(po[0]);
        if (nv < 3 || nv > MCSIM_POLY_MAX_VERTICES) return fail(MCSIM_ERR_ARG, "set_scene: polygon hull %d has %d vertices (3..%d)", h, nv, MCSIM_POLY_MAX_VERTICES);
        PolyHullRec& r = sc.poly[h];
        r.n = nv;
        double reach = 0;
        for (int i = 0; i < nv; ++i) {
            const int j = (i + 1) % nv, k = (i + 2) % nv;
            const double ex = po[3 + 2 * j] - po[3 + 2 * i], ey = po[4 + 2 * j] - po[4 + 2 * i];
            const double fx = po[3 + 2 * k] - po[3 + 2 * j], fy = po[4 + 2 * k] - po[4 + 2 * j];
            const double len = sqrt(ex * ex + ey * ey);
            if (!(len > 0) || ex * fy - ey * fx < -1e-12 * len) return fail(MCSIM_ERR_ARG, "set_scene: polygon hull %d is not convex and counter-clockwise at vertex %d", h, j);
            const double nx = ey / len, ny = -ex / len;
            double dmin = 1e300, dmax = -1e300;
            for (int q = 0; q < nv; ++q) { const double d = nx * po[3 + 2 * q] + ny * po[4 + 2 * q]; dmin = d < dmin ? d : dmin; dmax = d > dmax ? d : dmax; }
            r.vx[i] = float(po[3 + 2 * i]); r.vy[i] = float(po[4 + 2 * i]);
            r.nx[i] = float(nx); r.ny[i] = float(ny); r.dmax[i] = float(dmax); r.dmin[i] = float(dmin);
            reach = fmax(reach, hypot(po[3 + 2 * i], po[4 + 2 * i]));
        }
        if (float(reach) > sc.reach) sc.reach = float(reach);
        r.first = int(pairs.size() / 2);
        for (int b = 0; b < n_boxes; ++b) {
            const double* bx = boxes + b * MCSIM_PRISM_STRIDE;
            if (!((po[1] < bx[7]) && (bx[6] < po[2]))) continue;
            const double p = bx[2] * bx[0] + bx[3] * bx[1], q = bx[2] * bx[1] - bx[3] * bx[0];
            c->h_poly_pair_index.push_back(int(pairs.size() / 2));
            c->h_poly_pair_hull.push_back(h);
            for (int k = 0; k < 6; ++k) c->h_poly_pairgeom.push_back(bx[k]);
            pairs.push_back(make_float4(float(bx[0]), float(bx[1]), float(bx[2]), float(bx[3])));
            pairs.push_back(make_float4(float(bx[4]), float(bx[5]), -float(p), -float(q)));
        }
        r.count = int(pairs.size() / 2) - r.first;
        c->h_poly_reach[h] = reach;
    }
    c->n_poly = n_poly;
    if (int(pairs.size() / 2) > MCSIM_MAX_PAIRS)
        return fail(MCSIM_ERR_ARG, "set_scene: %d live (hull, box) pairs exceed MCSIM_MAX_PAIRS=%d", int(pairs.size() / 2), MCSIM_MAX_PAIRS);
    if (sc.n_pairs > MCSIM_MAX_PAIRS)
        return fail(MCSIM_ERR_ARG, "set_scene: %d live (hull, box) pairs exceed MCSIM_MAX_PAIRS=%d", sc.n_pairs, MCSIM_MAX_PAIRS);
    CK(cudaStreamSynchronize(c->stream));
    c->h_scene = sc;                       // uploaded to c_scene / c_pairs by the next launch of this context (bind_symbols)
    c->h_pairs = pairs;
    c->scene_tag = g_next_tag++;
    c->sc_prepared = false;                // a prepared scoring batch points at the old pair table
    CK(c->hulls_d.reserve(sizeof(HullD) * hd.size()));
    CK(cudaMemcpy(c->hulls_d.p, hd.data(), sizeof(HullD) * hd.size(), cudaMemcpyHostToDevice));
    CK(c->pairs_d.reserve(sizeof(PairD) * (pd.size() ? pd.size() : 1)));
    if (!pd.empty()) CK(cudaMemcpy(c->pairs_d.p, pd.data(), sizeof(PairD) * pd.size(), cudaMemcpyHostToDevice));
    CK(c->pairs_f.reserve(sizeof(float4) * (pairs.size() ? pairs.size() : 1)));
    if (!pairs.empty()) CK(cudaMemcpy(c->pairs_f.p, pairs.data(), sizeof(float4) * pairs.size(), cudaMemcpyHostToDevice));
    CK(c->sat2.reserve(sizeof(float4) * (sat2.size() ? sat2.size() : 1)));
    if (!sat2.empty()) CK(cudaMemcpy(c->sat2.p, sat2.data(), sizeof(float4) * sat2.size(), cudaMemcpyHostToDevice));
    c->sat2_n = int(sat2.size());
    c->n_hulls = n_hulls;
    {   // axis-aligned scene? (class headers of the grouped table: exactly (1, 0); hull yaw: cos exactly 1, sin exactly 0)
        bool ax = n_hulls > 0;
        for (int h = 0; h < n_hulls && ax; ++h) {
            const HullRec& r = sc.hulls[h];
            if (!(r.cp == 1.f && r.sp == 0.f)) ax = false;
            size_t at = size_t(r.g_first);
            for (int gi = 0; gi < r.n_groups && ax; ++gi) {
                if (!(sat2[at].x == 1.f && sat2[at].y == 0.f)) ax = false;
                at += 1 + 2 * size_t(__float_as_int_host(sat2[at].z));
            }
        }
        c->scene_axis_aligned = ax;
    }
    c->scene_set = true;
    c->n_pairs = sc.n_pairs;
    return MCSIM_OK;
}

// ---- Monte-Carlo rollouts ---------------------------------------------------------------------
// lower Cholesky factor of a 3x3 SPD matrix (what mvnrnd uses, glue_mvnrnd_meat.hpp:97)
static bool chol3(const double* A, double* Lo /* l00 l10 l11 l20 l21 l22 */) {
    if (!(A[0] > 0)) return false;
    const double l00 = sqrt(A[0]);
    const double l10 = A[3] / l00, l20 = A[6] / l00;
    const double d1 = A[4] - l10 * l10;
    if (!(d1 > 0)) return false;
    const double l11 = sqrt(d1);
    const double l21 = (A[7] - l20 * l10) / l11;
    const double d2 = A[8] - l20 * l20 - l21 * l21;
    if (!(d2 > 0)) return false;
    Lo[0] = l00; Lo[1] = l10; Lo[2] = l11; Lo[3] = l20; Lo[4] = l21; Lo[5] = sqrt(d2);
    return true;
}

typedef void (*mc_kernel_t)(const McLaunch);

template <int MODE, bool AUDIT, bool EE, int EST>
static mc_kernel_t pick_nl(int L) {
    if (MODE == MCSIM_MODE_PHILOX && L == 8) return mc_rollout_kernel<MODE, AUDIT, EE, EST, 8>;
    return mc_rollout_kernel<MODE, AUDIT, EE, EST, 0>;
}
static mc_kernel_t pick_kernel(int mode, bool audit, bool ee, int est, int L) {
    if (mode == MCSIM_MODE_REPLAY) {
        if (audit) return pick_nl<MCSIM_MODE_REPLAY, true, false, MCSIM_EST_REFERENCE>(L);
        return ee ? pick_nl<MCSIM_MODE_REPLAY, false, true, MCSIM_EST_REFERENCE>(L)
                  : pick_nl<MCSIM_MODE_REPLAY, false, false, MCSIM_EST_REFERENCE>(L);
    }
    if (est == MCSIM_EST_CLOSEDLOOP) {
        if (audit) return pick_nl<MCSIM_MODE_PHILOX, true, false, MCSIM_EST_CLOSEDLOOP>(L);
        return ee ? pick_nl<MCSIM_MODE_PHILOX, false, true, MCSIM_EST_CLOSEDLOOP>(L)
                  : pick_nl<MCSIM_MODE_PHILOX, false, false, MCSIM_EST_CLOSEDLOOP>(L);
    }
    if (audit) return pick_nl<MCSIM_MODE_PHILOX, true, false, MCSIM_EST_REFERENCE>(L);
    return ee ? pick_nl<MCSIM_MODE_PHILOX, false, true, MCSIM_EST_REFERENCE>(L)
              : pick_nl<MCSIM_MODE_PHILOX, false, false, MCSIM_EST_REFERENCE>(L);
}

// ax: the scene is axis-aligned (every box class axis is (1, 0), every hull's yaw is 0): the instantiations that drop the
// rotations into the class / hull frames (mc_philox2_sm100.cuh, class_frame2<AX>; same bits, 24 packed instructions less
// per thread-step on the committed scenes)
static mc_kernel_t pick_packed(bool ee, int est, bool ax) {
    if (ax) {
        if (est == MCSIM_EST_CLOSEDLOOP) return ee ? mc_philox2_kernel<true, MCSIM_EST_CLOSEDLOOP, 8, true> : mc_philox2_kernel<false, MCSIM_EST_CLOSEDLOOP, 8, true>;
        return ee ? mc_philox2_kernel<true, MCSIM_EST_REFERENCE, 8, true> : mc_philox2_kernel<false, MCSIM_EST_REFERENCE, 8, true>;
    }
    if (est == MCSIM_EST_CLOSEDLOOP) return ee ? mc_philox2_kernel<true, MCSIM_EST_CLOSEDLOOP, 8> : mc_philox2_kernel<false, MCSIM_EST_CLOSEDLOOP, 8>;
    return ee ? mc_philox2_kernel<true, MCSIM_EST_REFERENCE, 8> : mc_philox2_kernel<false, MCSIM_EST_REFERENCE, 8>;
}

static int mc_check(const mcsim_cuda_ctx* c, const mcsim_mc_params* p) {
    if (!c || !p) return fail(MCSIM_ERR_ARG, "mc: NULL context or params");
    if (!c->scene_set) return fail(MCSIM_ERR_ARG, "mc: mcsim_cuda_set_scene has not been called");
    if (p->T < 1 || p->T > MCSIM_MAX_WAYPOINTS) return fail(MCSIM_ERR_ARG, "mc: T=%d out of range 1..%d", p->T, MCSIM_MAX_WAYPOINTS);
    if (!p->trajectory) return fail(MCSIM_ERR_ARG, "mc: trajectory is NULL");
    if (p->mode == MCSIM_MODE_REPLAY) {
        if (p->T > 1 && !p->applied) return fail(MCSIM_ERR_ARG, "mc: replay mode needs the applied-command table");
    } else if (p->mode == MCSIM_MODE_PHILOX) {
        if (p->T > 1 && !p->odometry) return fail(MCSIM_ERR_ARG, "mc: odometry is NULL");
        if (p->n_landmarks < 0 || p->n_landmarks > MCSIM_MAX_LANDMARKS || (p->n_landmarks && !p->landmarks))
            return fail(MCSIM_ERR_ARG, "mc: bad landmarks");
    } else {
        return fail(MCSIM_ERR_ARG, "mc: unknown mode %d", p->mode);
    }
    return MCSIM_OK;
}

static void cull_key(const mcsim_cuda_ctx* c, const double* trajectory, int n_plans, int T, std::vector<double>& key) {
    key.assign(3 + size_t(n_plans) * 2 * T, 0.0);
    key[0] = double(c->scene_epoch); key[1] = double(T); key[2] = double(n_plans);
    for (int b = 0; b < n_plans; ++b) memcpy(&key[3 + size_t(b) * 2 * T], trajectory + size_t(b) * 3 * T, sizeof(double) * 2 * T);  // x and y rows
}

// Candidate lists of the packed kernel's broad phase (layout and exactness argument: mc_philox2_sm100.cuh).
// need(pair) = distance from x*_t to the box footprint - reach of the pair's hull - 1 mm of slack for the FP32
// plan record and the FP32 deviation test of the kernel.
// With polygon hulls (scalar kernels only) the hull index runs over the box hulls and then the polygon hulls; a polygon
// pair's entry is its index in c_pairs, its reach the largest vertex radius.
static void build_cull_lists(const mcsim_cuda_ctx* c, const double* trajectory, int T, std::vector<unsigned>& words) {
    const int nb = c->n_hulls, nh = c->n_hulls + c->n_poly, np = int(c->h_pair_hull.size()), npp = int(c->h_poly_pair_hull.size());
    const size_t cnt_words = (size_t(T) * nh * MC_CULL_LEVELS + 3) / 4, idx_words = size_t(T) * nh * MC_CULL_CAP;
    words.assign(cnt_words + idx_words, 0u);
    unsigned char* cnt = reinterpret_cast<unsigned char*>(words.data());
    unsigned* idx = words.data() + cnt_words;  // ushort2 (record, header)
    std::vector<std::pair<double, int>> need;
    for (int t = 0; t < T; ++t) {
        const double px = trajectory[t], py = trajectory[T + t];
        bool overflow[MC_CULL_LEVELS] = {false};
        for (int h = 0; h < nh; ++h) {
            need.clear();
            const bool poly = h >= nb;
            for (int b = 0; b < (poly ? npp : np); ++b) {
                if ((poly ? c->h_poly_pair_hull[b] + nb : c->h_pair_hull[b]) != h) continue;
                const double* g = poly ? &c->h_poly_pairgeom[size_t(b) * 6] : &c->h_pairgeom[size_t(b) * 6];
                const double dx = px - g[0], dy = py - g[1];
                const double lx = fabs(g[2] * dx + g[3] * dy) - g[4], ly = fabs(g[2] * dy - g[3] * dx) - g[5];
                const double dist = hypot(lx > 0 ? lx : 0, ly > 0 ? ly : 0);
                need.emplace_back(dist - (poly ? c->h_poly_reach[h - nb] : c->h_reach[h]) - 1e-3, b);
            }
            std::sort(need.begin(), need.end());
            for (int k = 0; k < MC_CULL_LEVELS; ++k) {
                const double m = sqrt(mc_cull_margin_sq(k));
                size_t n = 0;
                while (n < need.size() && need[n].first <= m) ++n;
                if (n > MC_CULL_CAP) overflow[k] = true;
                cnt[(size_t(t) * nh + h) * MC_CULL_LEVELS + k] = (unsigned char)(n > MC_CULL_CAP ? MC_CULL_CAP : n);
            }
            for (size_t k = 0; k < need.size() && k < MC_CULL_CAP; ++k)
                idx[(size_t(t) * nh + h) * MC_CULL_CAP + k] = poly ? unsigned(c->h_poly_pair_index[need[k].second]) : c->h_pair_sat2[need[k].second];
        }
        for (int k = 0; k < MC_CULL_LEVELS; ++k)
            if (overflow[k]) for (int h = 0; h < nh; ++h) cnt[(size_t(t) * nh + h) * MC_CULL_LEVELS + k] = MC_CULL_OVERFLOW;
    }
}

// Uploads the plan table + constants; zeroes the counters.
extern "C" int mcsim_cuda_mc_prepare(mcsim_cuda_ctx* c, const mcsim_mc_params* p) {
    int rc = mc_check(c, p);
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    const int T = p->T, L = p->n_landmarks;
    McPlanConst pc;
    memset(&pc, 0, sizeof(pc));
    pc.T = T;
    double Lo[6];
    if (!chol3(p->cov0, Lo)) return fail(MCSIM_ERR_ARG, "mc: initial covariance is not symmetric positive definite");
    for (int k = 0; k < 3; ++k) { pc.mu0d[k] = p->trajectory[k * T]; pc.mu0[k] = float(pc.mu0d[k]); }
    for (int k = 0; k < 6; ++k) { pc.chold[k] = Lo[k]; pc.chol[k] = float(Lo[k]); }
    pc.sqrtQ = float(sqrt(p->Q));
    mcsim_math::NoiseModel<double> nmd;
    for (int k = 0; k < 4; ++k) { pc.nm.alpha[k] = float(p->alphas[k]); nmd.alpha[k] = p->alphas[k]; }
    pc.nm.Q = float(p->Q);
    pc.nm.L = L;
    for (int l = 0; l < L && p->mode == MCSIM_MODE_PHILOX; ++l) { pc.nm.lx[l] = float(p->landmarks[l]); pc.nm.ly[l] = float(p->landmarks[L + l]); }
    std::vector<float> plan;
    if (p->mode == MCSIM_MODE_REPLAY) {
        plan.resize(size_t(T > 1 ? T - 1 : 1) * MC_REPLAY_STRIDE, 0.f);
        for (int t = 0; t < T - 1; ++t)
            for (int k = 0; k < 3; ++k) memcpy(&plan[t * MC_REPLAY_STRIDE + 2 * k], &p->applied[k * (T - 1) + t], sizeof(double));
    } else {
        plan.resize(size_t(T) * MC_PHILOX_STRIDE, 0.f);
        for (int t = 0; t < T; ++t) {
            float* r = &plan[size_t(t) * MC_PHILOX_STRIDE];
            for (int k = 0; k < 3; ++k) r[k] = float(p->trajectory[k * T + t]);
            if (t < T - 1) {
                double us[3], M[3];
                for (int k = 0; k < 3; ++k) { us[k] = p->odometry[k * (T - 1) + t]; r[3 + k] = float(us[k]); }
                mcsim_math::motion_noise(nmd, us, M);  // generateM_EKF(u*_t), MCSimulator.h:698-701
                for (int k = 0; k < 3; ++k) r[6 + k] = float(M[k]);
            }
        }
    }
    // packed kernel's table (x*, u*, M, M0 + M2 per waypoint) in __constant__ memory
    std::vector<float> plan2;
    if (p->mode == MCSIM_MODE_PHILOX && T <= MC_PLAN2_MAX_T) {
        plan2.assign(size_t(T) * MC_PLAN2_STRIDE, 0.f);
        for (int t = 0; t < T; ++t) {
            const float* r = &plan[size_t(t) * MC_PHILOX_STRIDE];
            float* w = &plan2[size_t(t) * MC_PLAN2_STRIDE];
            for (int k = 0; k < 9; ++k) w[k] = r[k];
            w[9] = r[6] + r[8];
            w[10] = float(mcsim_math::angle_wrap(p->trajectory[2 * T + t]));
        }
    }
    std::vector<unsigned> cull;
    bool cull_cached = false;
    // candidate lists of the exact broad phase: the packed Philox kernels (lists in shared memory) and the replay kernel
    // (lists read from global memory); box hulls only
    const bool lists_for_packed = !plan2.empty() && L == 8 && c->n_poly == 0;
    const bool lists_for_replay = p->mode == MCSIM_MODE_REPLAY;
    const bool lists_for_poly = c->n_poly > 0;   // polygon hulls: the scalar kernel runs in both modes and walks the lists
    const int all_pairs = c->n_pairs + int(c->h_poly_pair_hull.size());
    if ((lists_for_packed || lists_for_replay || lists_for_poly) && !p->exhaustive && c->sat2_n < 65536 && all_pairs > 0 && all_pairs < 65536) {
        std::vector<double> key;
        cull_key(c, p->trajectory, 1, T, key);
        if (key == c->mc_cull_key) {
            cull = c->mc_cull_host;
            cull_cached = true;
        } else {
            build_cull_lists(c, p->trajectory, T, cull);
            c->mc_cull_key.swap(key);
            c->mc_cull_host = cull;
        }
    }
    // the packed kernels keep the lists in shared memory next to the pair table; too large a plan x scene goes without them
    c->mc_cull_fits_smem = (cull.size() * sizeof(unsigned) + sizeof(float4) * size_t(c->sat2_n)) * MC_P2_CTAS_PER_SM <= 200 * 1024;
    if (!lists_for_replay && !lists_for_poly && !c->mc_cull_fits_smem) cull.clear();
    c->mc_cull_words = int(cull.size());
    std::vector<float2> nominal(T);
    for (int t = 0; t < T; ++t) nominal[t] = make_float2(float(p->trajectory[t]), float(p->trajectory[T + t]));
    if (!cull.empty()) CK(c->mc_cull.reserve(cull.size() * sizeof(unsigned)));
    CK(c->mc_nominal.reserve(sizeof(float2) * size_t(T)));
    CK(c->plan.reserve(plan.size() * sizeof(float)));
    CK(c->counters.reserve(sizeof(unsigned long long) * (4 + MCSIM_MAX_WAYPOINTS)));
    CK(cudaStreamSynchronize(c->stream));
    // uploaded to c_mc / c_plan2 by the launch (bind_symbols) when the tag is new.  Repeated runs of one plan (the reference
    // driver's 200-run loop: only the applied commands of the shared real robot change, and those travel in c->plan) keep
    // their tag, so the launch does not synchronise and re-upload the symbols every call.
    if (c->mc_sym.tag == 0 || memcmp(&c->mc_sym.pc, &pc, sizeof(pc)) != 0 || c->mc_sym.plan2 != plan2) {
        c->mc_sym.pc = pc;
        c->mc_sym.plan2 = plan2;
        c->mc_sym.tag = g_next_tag++;
    }
    // plan and list centres through the context's pinned staging buffer when they fit: truly asynchronous copies, and no
    // synchronisation at the end of the call (the buffer is rewritten by the next prepare only, after this run was collected)
    const size_t plan_bytes = plan.size() * sizeof(float), nom_bytes = sizeof(float2) * size_t(T);
    const bool staged = c->h_stage && plan_bytes + nom_bytes <= MCSIM_STAGE_BYTES;
    if (staged) {
        memcpy(c->h_stage, plan.data(), plan_bytes);
        memcpy(c->h_stage + plan_bytes, nominal.data(), nom_bytes);
        CK(cudaMemcpyAsync(c->plan.p, c->h_stage, plan_bytes, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->mc_nominal.p, c->h_stage + plan_bytes, nom_bytes, cudaMemcpyHostToDevice, c->stream));
    } else {
        CK(cudaMemcpyAsync(c->plan.p, plan.data(), plan_bytes, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->mc_nominal.p, nominal.data(), nom_bytes, cudaMemcpyHostToDevice, c->stream));
    }
    c->mc_cull_uploaded = !cull.empty() && !cull_cached;
    if (!cull.empty() && !cull_cached) CK(cudaMemcpyAsync(c->mc_cull.p, cull.data(), cull.size() * sizeof(unsigned), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * (4 + T), c->stream));
    if (!staged || c->mc_cull_uploaded) CK(cudaStreamSynchronize(c->stream));  // stack vectors
    c->mc = *p;
    c->mc.trajectory = c->mc.odometry = c->mc.landmarks = c->mc.applied = c->mc.z0 = nullptr;
    c->mc_prepared = true;
    c->mc_T = T;
    return MCSIM_OK;
}

typedef void (*mc_seg_kernel_t)(const McSegLaunch);
// Early exit with survivor compaction (mc_philox2_segment_kernel): chunks of at most MC_SEG_CHUNK rollouts, one launch
// per MC_SEG waypoints; everything is queued on the context's stream, the survivor counts stay on the device.
#define MC_SEG_CHUNK (1ull << 24)
#define MC_SEG_MIN_ROLLOUTS (1ull << 18)   // below this the plain early-exit kernel is as fast (launch overheads)
static int mc_launch_segmented(mcsim_cuda_ctx* c, const McLaunch& a0, uint64_t n, uint64_t offset, unsigned* d_counts, int block, size_t smem, unsigned grid_cap) {
    const int T = c->mc_T;
    const int n_seg = (T + MC_SEG - 1) / MC_SEG;
    const uint64_t chunk = n < MC_SEG_CHUNK ? ((n + 1) & ~1ull) : MC_SEG_CHUNK;
    CK(c->seg_state[0].reserve(sizeof(float) * MC_STATE_WORDS * chunk));
    CK(c->seg_state[1].reserve(sizeof(float) * MC_STATE_WORDS * chunk));
    CK(c->seg_alive.reserve(sizeof(unsigned) * (size_t(n_seg) + 1)));
    mc_seg_kernel_t k = c->mc.estimator == MCSIM_EST_CLOSEDLOOP ? mc_philox2_segment_kernel<MCSIM_EST_CLOSEDLOOP, 8> : mc_philox2_segment_kernel<MCSIM_EST_REFERENCE, 8>;
    if (c->scene_axis_aligned)
        k = c->mc.estimator == MCSIM_EST_CLOSEDLOOP ? mc_philox2_segment_kernel<MCSIM_EST_CLOSEDLOOP, 8, true> : mc_philox2_segment_kernel<MCSIM_EST_REFERENCE, 8, true>;
    if (smem > 40 * 1024) CK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    if (d_counts) CK(cudaMemsetAsync(d_counts, 0, sizeof(unsigned) * n, c->stream));
    for (uint64_t done = 0; done < n; done += chunk) {
        const uint64_t m = n - done < chunk ? n - done : chunk;
        CK(cudaMemsetAsync(c->seg_alive.p, 0, sizeof(unsigned) * (size_t(n_seg) + 1), c->stream));
        for (int sg = 0; sg < n_seg; ++sg) {
            McSegLaunch g;
            g.base = a0;
            g.base.n = m;
            g.base.offset = offset + done;
            g.base.counts = d_counts ? d_counts + done : nullptr;
            g.t0 = sg * MC_SEG;
            g.t1 = (sg + 1) * MC_SEG < T ? (sg + 1) * MC_SEG : T;
            g.in = sg == 0 ? nullptr : (const float*)c->seg_state[(sg - 1) & 1].p;
            g.out = g.t1 < T ? (float*)c->seg_state[sg & 1].p : nullptr;
            g.n_in = (const unsigned*)c->seg_alive.p + sg;
            g.n_out = (unsigned*)c->seg_alive.p + sg + 1;
            g.cap = unsigned(chunk);
            const unsigned long long want = ((m + 1) / 2 + block - 1) / block;
            const unsigned grid = unsigned(want < grid_cap ? (want ? want : 1) : grid_cap);
            k<<<grid, block, smem, c->stream>>>(g);
            ++c->launches;
        }
        CK(cudaGetLastError());
    }
    return MCSIM_OK;
}

static int mc_launch_impl(mcsim_cuda_ctx* c, uint64_t n, uint64_t offset, const double* d_z0, unsigned* d_counts,
                          float* d_poses, float* d_clear) {
    const mcsim_mc_params& p = c->mc;
    SymbolLock sym;
    { const int rb = bind_symbols(c, &c->mc_sym, sym); if (rb) return rb; }
    const bool audit = d_poses || d_clear;
    const bool packed = p.mode == MCSIM_MODE_PHILOX && !audit && !p.precise && p.n_landmarks == 8 && c->mc_T <= MC_PLAN2_MAX_T && c->n_poly == 0;
    mc_kernel_t k = packed ? pick_packed(p.early_exit != 0, p.estimator, c->scene_axis_aligned)
                           : pick_kernel(p.mode, audit, p.early_exit != 0, p.estimator, p.n_landmarks);
    McLaunch a;
    a.plan = (const float*)c->plan.p;
    a.z0 = d_z0;
    a.n = n;
    a.offset = offset;
    a.k0 = unsigned(p.seed);
    a.k1 = unsigned(p.seed >> 32);
    a.counters = (unsigned long long*)c->counters.p;
    a.counts = d_counts;
    a.poses = d_poses;
    a.clearance = d_clear;
    a.near_eps = float(p.near_eps);
    a.sat2 = (const float4*)c->sat2.p;
    a.sat2_n = c->sat2_n;
    const bool cull = c->mc_cull_words > 0 && (packed ? c->mc_cull_fits_smem : ((p.mode == MCSIM_MODE_REPLAY || c->n_poly > 0) && !audit));
    a.cull = cull ? (const unsigned*)c->mc_cull.p : nullptr;
    a.cull_words = cull ? c->mc_cull_words : 0;
    a.nominal = (const float2*)c->mc_nominal.p;
    const int block = packed ? 128 : 256;
    // scalar kernels: plan (rounded up to 16 bytes) | pair table | lists, when the lists are used and all of it fits 96 KB
    const size_t plan_smem = (sizeof(float) * size_t(c->mc_T) * MC_PHILOX_STRIDE + 15) & ~size_t(15);
    const size_t lists_smem = sizeof(float4) * size_t(c->sat2_n) + sizeof(unsigned) * size_t(c->mc_cull_words);
    a.smem_lists = !packed && cull && plan_smem + lists_smem <= 96 * 1024;
    const size_t smem = packed ? sizeof(float4) * size_t(c->sat2_n) + (cull ? sizeof(unsigned) * size_t(c->mc_cull_words) : 0)
                               : plan_smem + (a.smem_lists ? lists_smem : 0);
    unsigned long long want = packed ? ((n + 1) / 2 + block - 1) / block : (n + block - 1) / block;
    // persistent: whole waves of resident CTAs
    const unsigned long long cap = (unsigned long long)c->n_sm * (packed ? MC_P2_CTAS_PER_SM : 8);
    const unsigned grid = unsigned(want < cap ? (want ? want : 1) : cap);
    if (packed && p.early_exit && n >= MC_SEG_MIN_ROLLOUTS && c->mc_T > MC_SEG && !p.no_compaction)
        return mc_launch_segmented(c, a, n, offset, d_counts, block, smem, unsigned(cap));
    if (smem > 40 * 1024) CK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    k<<<grid, block, smem, c->stream>>>(a);
    ++c->launches;
    CK(cudaGetLastError());
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_mc_launch(mcsim_cuda_ctx* c, uint64_t n, uint64_t offset) {
    if (!c || !c->mc_prepared) return fail(MCSIM_ERR_ARG, "mc_launch: mcsim_cuda_mc_prepare has not been called");
    if (c->mc.mode != MCSIM_MODE_PHILOX) return fail(MCSIM_ERR_ARG, "mc_launch: device-resident launches are Philox mode only");
    CK(cudaSetDevice(c->device));
    return mc_launch_impl(c, n, offset, nullptr, nullptr, nullptr, nullptr);
}

extern "C" void* mcsim_cuda_mc_counters(mcsim_cuda_ctx* c) { return c ? c->counters.p : nullptr; }

extern "C" int mcsim_cuda_mc_collect(mcsim_cuda_ctx* c, mcsim_mc_result* r) {
    if (!c || !r || !c->mc_prepared) return fail(MCSIM_ERR_ARG, "mc_collect: nothing prepared");
    CK(cudaSetDevice(c->device));
    const int T = c->mc_T;
    CK(cudaMemcpyAsync(c->h_counters, c->counters.p, sizeof(unsigned long long) * (4 + T), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * (4 + T), c->stream));
    CK(cudaStreamSynchronize(c->stream));
    r->hits = c->h_counters[CNT_HITS];
    r->near_contact = c->h_counters[CNT_NEAR];
    r->executed_steps = c->h_counters[CNT_STEPS];
    r->pair_tests = c->h_counters[CNT_TESTS];
    if (r->first_hit_hist) for (int t = 0; t < T; ++t) r->first_hit_hist[t] = c->h_counters[CNT_HIST + t];
    return MCSIM_OK;
}

// mc_run in two halves, so that several devices can be in flight at once (mcsim_group.cuh): begin = prepare, upload,
// launch, queue the per-rollout downloads; end = collect the counters (synchronises) and read the kernel time.
static int mc_run_begin(mcsim_cuda_ctx* c, const mcsim_mc_params* p, mcsim_mc_result* r) {
    if (!r) return fail(MCSIM_ERR_ARG, "mc_run: result is NULL");
    int rc = mcsim_cuda_mc_prepare(c, p);
    if (rc) return rc;
    c->h2d_bytes = sizeof(McPlanConst) + sizeof(float) * size_t(p->T) * MC_PHILOX_STRIDE +
                   ((p->mode == MCSIM_MODE_PHILOX && p->T <= MC_PLAN2_MAX_T) ? sizeof(float) * size_t(p->T) * MC_PLAN2_STRIDE : 0) +
                   (c->mc_cull_uploaded ? sizeof(unsigned) * size_t(c->mc_cull_words) : 0) + sizeof(float2) * size_t(p->T);
    c->d2h_bytes = sizeof(unsigned long long) * (4 + p->T);
    const uint64_t n = p->n_rollouts;
    const int T = p->T;
    const double* d_z0 = nullptr;
    if (p->mode == MCSIM_MODE_REPLAY) {
        if (p->z0 && n) {  // injected particles; z0 == NULL: the kernel draws them from Philox
            CK(c->z0.reserve(sizeof(double) * 3 * n));
            CK(cudaMemcpyAsync(c->z0.p, p->z0, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, c->stream));
            c->h2d_bytes += sizeof(double) * 3 * n;
            d_z0 = (const double*)c->z0.p;
        }
    }
    unsigned* d_counts = nullptr;
    float *d_poses = nullptr, *d_clear = nullptr;
    if (r->counts && n) { CK(c->counts.reserve(sizeof(unsigned) * n)); d_counts = (unsigned*)c->counts.p; }
    const bool audit = r->poses || r->clearance;
    if (audit && n) {
        CK(c->poses.reserve(sizeof(float) * 3 * T * n)); d_poses = (float*)c->poses.p;
        CK(c->clearance.reserve(sizeof(float) * T * n)); d_clear = (float*)c->clearance.p;
    }
    CK(cudaEventRecord(c->ev0, c->stream));
    if (n) { rc = mc_launch_impl(c, n, p->rollout_offset, d_z0, d_counts, d_poses, d_clear); if (rc) return rc; }
    CK(cudaEventRecord(c->ev1, c->stream));
    if (d_counts) { CK(cudaMemcpyAsync(r->counts, d_counts, sizeof(unsigned) * n, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(unsigned) * n; }
    if (d_poses && r->poses) CK(cudaMemcpyAsync(r->poses, d_poses, sizeof(float) * 3 * T * n, cudaMemcpyDeviceToHost, c->stream));
    if (d_clear && r->clearance) CK(cudaMemcpyAsync(r->clearance, d_clear, sizeof(float) * T * n, cudaMemcpyDeviceToHost, c->stream));
    return MCSIM_OK;
}
static int mc_run_end(mcsim_cuda_ctx* c, mcsim_mc_result* r) {
    const int rc = mcsim_cuda_mc_collect(c, r);
    if (rc) return rc;
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_mc_run(mcsim_cuda_ctx* c, const mcsim_mc_params* p, mcsim_mc_result* r) {
    const int rc = mc_run_begin(c, p, r);
    return rc ? rc : mc_run_end(c, r);
}

// the step draws of one Philox-mode rollout, host FP64 (csrc/philox.h: step draws)
extern "C" int mcsim_step_normals(uint64_t seed, uint64_t rollout_id, int t, double out[12]) {
    if (!out || t < 0) return MCSIM_ERR_ARG;
    const unsigned k0 = unsigned(seed), k1 = unsigned(seed >> 32);
    const mcsim_rng::u32x4 ca = {unsigned(rollout_id), unsigned(rollout_id >> 32), unsigned(t * 8), unsigned(mcsim_rng::PURPOSE_STEP)};
    const mcsim_rng::u32x4 cb = {unsigned(rollout_id), unsigned(rollout_id >> 32), unsigned(t * 8 + 1), unsigned(mcsim_rng::PURPOSE_STEP)};
    uint32_t rad[6], ang[6];
    mcsim_rng::step_bits(mcsim_rng::philox4x32_10(ca, k0, k1), mcsim_rng::philox4x32_10(cb, k0, k1), rad, ang);
    for (int j = 0; j < 6; ++j) {
        const double u = (double(rad[j]) + 0.5) / 16777216.0, th = 2 * MCSIM_PI * (double(ang[j]) + 0.5) / 65536.0;
        const double r = sqrt(-2.0 * log(u));
        out[2 * j] = r * cos(th);
        out[2 * j + 1] = r * sin(th);
    }
    return MCSIM_OK;
}

// the same draws as the packed kernel computes them on the device (FP32, MUFU): [n][12] floats to the host
extern "C" int mcsim_cuda_step_normals(mcsim_cuda_ctx* c, uint64_t seed, uint64_t id0, uint64_t n, int t, float* out) {
    if (!c || !out || t < 0 || n == 0 || n > (1ull << 28)) return fail(MCSIM_ERR_ARG, "step_normals: bad arguments");
    CK(cudaSetDevice(c->device));
    DevBuf buf;
    CK(buf.reserve(sizeof(float) * 12 * n));
    const unsigned long long thr = (n + 1) / 2;
    mc_step_normals_kernel<<<unsigned((thr + 255) / 256), 256, 0, c->stream>>>(id0, n, t, unsigned(seed), unsigned(seed >> 32), (float*)buf.p);
    ++c->launches;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, buf.p, sizeof(float) * 12 * n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    buf.release();
    if (e != cudaSuccess) return fail(MCSIM_ERR_CUDA, "step_normals: %s", cudaGetErrorString(e));
    return MCSIM_OK;
}

// ---- GMM mode -----------------------------------------------------------------------------------
// two standard normals from one Philox block, host side (the host thread of a Philox-mode GMM run)
extern "C" int mcsim_host_normals(uint64_t seed, uint32_t stream_id, int T, int per_step, double* out) {
    if (!out || T < 1 || per_step < 1) return MCSIM_ERR_ARG;
    for (int t = 0; t < T - 1; ++t) {
        for (int j = 0; j < per_step; j += 4) {
            mcsim_rng::u32x4 c = {unsigned(t), stream_id, unsigned(j >> 2), unsigned(mcsim_rng::PURPOSE_GMM_HOST)};
            c = mcsim_rng::philox4x32_10(c, unsigned(seed), unsigned(seed >> 32));
            const unsigned w[4] = {c.x, c.y, c.z, c.w};
            for (int h = 0; h < 2; ++h) {
                const double u1 = (double(w[2 * h]) + 0.5) * (1.0 / 4294967296.0);
                const double u2 = (double(w[2 * h + 1]) + 0.5) * (1.0 / 4294967296.0);
                const double r = sqrt(-2.0 * log(u1));
                const double n0 = r * cos(2 * MCSIM_PI * u2), n1 = r * sin(2 * MCSIM_PI * u2);
                if (j + 2 * h < per_step) out[size_t(t) * per_step + j + 2 * h] = n0;
                if (j + 2 * h + 1 < per_step) out[size_t(t) * per_step + j + 2 * h + 1] = n1;
            }
        }
    }
    return MCSIM_OK;
}
static void host_normals(uint64_t seed, unsigned stream_id, int T, int per_step, std::vector<double>& out) {
    out.resize(size_t(T > 1 ? T - 1 : 0) * per_step);
    if (!out.empty()) mcsim_host_normals(seed, stream_id, T, per_step, out.data());
}

static int gmm_reserve(mcsim_cuda_ctx* c, int T, int K, int B = 1) {
    CK(c->gmm_comps.reserve(sizeof(double) * MCSIM_MAX_GAUSSIANS * GMM_COMP * B));
    CK(c->gmm_stats.reserve(sizeof(double) * MCSIM_MAX_GAUSSIANS * GMM_STAT * B));
    CK(c->gmm_hist.reserve(sizeof(unsigned) * MCSIM_MAX_GAUSSIANS * B));
    CK(c->gmm_prob.reserve(sizeof(double) * T * B));
    CK(c->gmm_pre.reserve(sizeof(double) * size_t(T) * K * GMM_REC));
    CK(c->gmm_post.reserve(sizeof(double) * K * GMM_REC));
    CK(c->gmm_trunc.reserve(sizeof(double) * K * GMM_REC));
    CK(c->gmm_nout.reserve(sizeof(unsigned long long) * 2 * MCSIM_MAX_GAUSSIANS));
    CK(c->gmm_deg.reserve(sizeof(int) * B));
    CK(c->gmm_steps.reserve(sizeof(GmmStepTable) * (T > 1 ? T - 1 : 1) * B));
    CK(cudaMemsetAsync(c->gmm_stats.p, 0, sizeof(double) * MCSIM_MAX_GAUSSIANS * GMM_STAT * B, c->stream));
    CK(cudaMemsetAsync(c->gmm_hist.p, 0, sizeof(unsigned) * MCSIM_MAX_GAUSSIANS * B, c->stream));
    CK(cudaMemsetAsync(c->gmm_deg.p, 0, sizeof(int) * B, c->stream));
    return MCSIM_OK;
}

static void gmm_fill_launch(mcsim_cuda_ctx* c, GmmLaunch& a, int K, int T, int L, uint64_t n, uint64_t seed, unsigned stream_id) {
    memset(&a, 0, sizeof(a));
    a.K = K; a.T = T; a.L = L; a.n = n;
    a.comps = (double*)c->gmm_comps.p;
    a.stats = (double*)c->gmm_stats.p;
    a.hist = (unsigned*)c->gmm_hist.p;
    a.k0 = unsigned(seed); a.k1 = unsigned(seed >> 32);
    a.stream_id = stream_id;
    a.hulls_d = (const HullD*)c->hulls_d.p;
    a.pairs_d = (const PairD*)c->pairs_d.p;
    a.n_hulls = c->n_hulls;
    a.steps = (const GmmStepTable*)c->gmm_steps.p;
    a.probabilities = (double*)c->gmm_prob.p;
    a.degenerate = (int*)c->gmm_deg.p;
    a.sat2 = (const float4*)c->sat2.p;
}

// GMM Philox mode: candidate lists + list centres of n_plans plans (plan b = trajectories + b * 3 * T), uploaded
// to gmm_cull / gmm_nominal; the executed-pair-test counter is zeroed.  Returns the words per plan (0: no lists).
static int gmm_upload_cull(mcsim_cuda_ctx* c, GmmLaunch& a, const double* trajectories, int n_plans, int T, bool exhaustive) {
    CK(c->gmm_tests.reserve(sizeof(unsigned long long)));
    CK(cudaMemsetAsync(c->gmm_tests.p, 0, sizeof(unsigned long long), c->stream));
    a.tests = (unsigned long long*)c->gmm_tests.p;
    a.cull = nullptr;
    a.cull_stride = 0;
    c->gmm_cull_words = 0;
    const int all_pairs = c->n_pairs + int(c->h_poly_pair_hull.size());
    if (exhaustive || c->sat2_n >= 65536 || all_pairs == 0 || all_pairs >= 65536) return MCSIM_OK;
    if (size_t(n_plans) * T * (c->n_hulls + c->n_poly) * (MC_CULL_CAP + 1) > (size_t(64) << 20)) return MCSIM_OK;  // 256 MB of lists: go without
    std::vector<double> key;
    cull_key(c, trajectories, n_plans, T, key);
    if (key == c->gmm_cull_key && c->gmm_cull_words_cached > 0) {   // same scene, same plan(s): the lists are on the device
        a.cull = (const unsigned*)c->gmm_cull.p;
        a.nominal = (const float2*)c->gmm_nominal.p;
        a.cull_stride = n_plans > 1 ? c->gmm_cull_words_cached : 0;
        c->gmm_cull_words = c->gmm_cull_words_cached;
        return MCSIM_OK;
    }
    std::vector<unsigned> all, one;
    std::vector<float2> nominal(size_t(n_plans) * T);
    for (int b = 0; b < n_plans; ++b) {
        const double* tr = trajectories + size_t(b) * 3 * T;
        build_cull_lists(c, tr, T, one);
        all.insert(all.end(), one.begin(), one.end());
        for (int t = 0; t < T; ++t) nominal[size_t(b) * T + t] = make_float2(float(tr[t]), float(tr[T + t]));
    }
    CK(c->gmm_cull.reserve(all.size() * sizeof(unsigned)));
    CK(c->gmm_nominal.reserve(nominal.size() * sizeof(float2)));
    CK(cudaMemcpyAsync(c->gmm_cull.p, all.data(), all.size() * sizeof(unsigned), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->gmm_nominal.p, nominal.data(), nominal.size() * sizeof(float2), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // stack vectors
    c->h2d_bytes += all.size() * sizeof(unsigned) + nominal.size() * sizeof(float2);
    a.cull = (const unsigned*)c->gmm_cull.p;
    a.nominal = (const float2*)c->gmm_nominal.p;
    a.cull_stride = n_plans > 1 ? int(one.size()) : 0;
    c->gmm_cull_words = c->gmm_cull_words_cached = int(one.size());
    c->gmm_cull_key.swap(key);
    return MCSIM_OK;
}

// the three launches of one waypoint
static int gmm_launch_waypoint(mcsim_cuda_ctx* c, const GmmLaunch& a_in, int mode) {
    GmmLaunch a = a_in;
    SymbolLock sym;
    { const int rb = bind_symbols(c, nullptr, sym); if (rb) return rb; }
    const unsigned B = a.B > 1 ? unsigned(a.B) : 1u;
    int block, spread_flag = 0;
    unsigned gx;
    if (B == 1) {
        // one run, small sample counts (the reference's N = 10^4): 64-thread CTAs so that every SM gets work
        // (the Philox-mode sample kernel takes two samples per thread and pass unless that leaves SMs without work)
        const bool spread = mode == MCSIM_MODE_PHILOX && a.n < (unsigned long long)c->n_sm * 256;
        spread_flag = spread ? 1 : 0;
        const unsigned per_thread = mode == MCSIM_MODE_PHILOX && !spread ? 2 : 1;
        block = a.n < (unsigned long long)c->n_sm * 256 * per_thread ? 64 : 256;
        const unsigned long long want = (a.n + block * per_thread - 1) / (block * per_thread), cap = (unsigned long long)c->n_sm * 8;
        gx = unsigned(want < cap ? (want ? want : 1) : cap);
    } else {
        // a batch fills the GPU by its members: few 256-thread CTAs per member, many samples per thread (the CTA
        // prologue / epilogue and the warp reductions are per CTA and per component change, not per sample);
        // two waves of resident CTAs in total, the count per member chosen for the fewest idle lanes in the last pass
        block = 256;
        const unsigned per_thread = mode == MCSIM_MODE_PHILOX ? 2 : 1;
        const unsigned long long want = (a.n + block * per_thread - 1) / (block * per_thread);
        unsigned g0 = unsigned((2ull * c->n_sm * 8 + B - 1) / B);
        if (g0 < 1) g0 = 1;
        unsigned best = g0;
        unsigned long long best_waste = ~0ull;
        for (unsigned cand = g0; cand <= 2 * g0; ++cand) {
            const unsigned long long per = (unsigned long long)cand * block * per_thread, passes = (a.n + per - 1) / per;
            const unsigned long long waste = passes * per - a.n;
            if (waste < best_waste) { best_waste = waste; best = cand; }
        }
        gx = unsigned(best < want ? best : (want ? want : 1));
    }
    const dim3 grid(gx, B);
    if (mode == MCSIM_MODE_PHILOX && a.n > (1ull << 31)) return fail(MCSIM_ERR_ARG, "gmm: Philox mode takes at most 2^31 samples per waypoint");
    a.spread = spread_flag;
    if (mode == MCSIM_MODE_PHILOX && (a.free_out || c->n_poly)) {   // EM-refit variant (keeps the free samples) / polygon hulls: the scalar kernel
        gmm_count_kernel<<<grid, block, 0, c->stream>>>(a);
        gmm_sample_kernel<MCSIM_MODE_PHILOX><<<grid, block, 0, c->stream>>>(a);
        c->launches += 2;
    } else if (mode == MCSIM_MODE_PHILOX) {
        gmm_count_kernel<<<grid, block, 0, c->stream>>>(a);
        if (c->scene_axis_aligned) gmm_sample_philox_kernel<true><<<grid, block, 0, c->stream>>>(a);
        else gmm_sample_philox_kernel<false><<<grid, block, 0, c->stream>>>(a);
        c->launches += 2;
    } else {
        gmm_sample_kernel<MCSIM_MODE_REPLAY><<<grid, block, 0, c->stream>>>(a);
        c->launches += 1;
    }
    gmm_refit_kernel<<<B, 32, 0, c->stream>>>(a);
    c->launches += 1;
    CK(cudaGetLastError());
    return MCSIM_OK;
}

// one thread-block cluster per run for the whole plan (gmm_cluster_run_kernel): single runs and small batches in
// Philox mode whose samples per waypoint are a few passes of a cluster; larger sample counts keep the per-waypoint
// kernels (the whole GPU per waypoint), batches that fill the GPU by themselves the one-CTA-per-member kernel
#define GMM_CLUSTER_MAX_SAMPLES 98304ull   // measured switch-over (tools/gmm_quick.py sweep, B200): 65,536: 1.18 vs 1.41 ms; 131,072: 1.62 vs 1.46 ms
static bool gmm_cluster_fits(const mcsim_cuda_ctx* c, uint64_t n_samples, int n_runs, int kernel_choice) {
    if (kernel_choice == MCSIM_GMM_KERNEL_PER_WAYPOINT || c->n_poly) return false;
    const size_t smem = sizeof(float4) * size_t(c->sat2_n) + sizeof(unsigned) * size_t(c->gmm_cull_words);
    const uint64_t most = kernel_choice == MCSIM_GMM_KERNEL_CLUSTER ? (1ull << 31) : GMM_CLUSTER_MAX_SAMPLES;   // forced: any size
    return n_samples <= most && smem <= 64 * 1024 && n_runs >= 1 && n_runs < 64;
}
static int gmm_launch_cluster(mcsim_cuda_ctx* c, const GmmLaunch& a, int n_runs) {
    SymbolLock sym;
    { const int rb = bind_symbols(c, nullptr, sym); if (rb) return rb; }
    void (*ck)(const GmmLaunch, int, int) = c->scene_axis_aligned ? gmm_cluster_run_kernel<true> : gmm_cluster_run_kernel<false>;
    const size_t smem = sizeof(float4) * size_t(c->sat2_n) + sizeof(unsigned) * size_t(c->gmm_cull_words);
    // the kernel holds ~30 KB of static shared memory (the peers' slots): static + dynamic pass 48 KB early
    if (smem > 12 * 1024) CK(cudaFuncSetAttribute((const void*)ck, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    // up to 8 runs: 16 CTAs per cluster (one pass over 10^4 samples; a GPC holds one such cluster), else 8
    int cl = n_runs <= 8 ? 16 : GMM_CLUSTER;
    if (cl > 8) CK(cudaFuncSetAttribute((const void*)ck, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    for (;;) {
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(unsigned(n_runs) * unsigned(cl));
        cfg.blockDim = dim3(GMM_CLUSTER_THREADS);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = c->stream;
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = unsigned(cl);
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int fit = 0;
        const cudaError_t q = cudaOccupancyMaxActiveClusters(&fit, ck, &cfg);
        if (q == cudaSuccess && fit >= 1) break;
        (void)cudaGetLastError();
        if (cl == GMM_CLUSTER) return -1;   // no cluster shape fits (e.g. a partitioned device): the caller keeps the per-waypoint kernels
        cl = GMM_CLUSTER;
    }
    CK(cudaLaunchKernelEx(&cfg, ck, a, c->sat2_n, c->gmm_cull_words));
    ++c->launches;
    return MCSIM_OK;
}

// ---- method variant: EM-refit truncation -------------------------------------------------------------
// After the moment refit of truncateGMM (:592-629) the K-component mixture is refitted to ALL collision-free samples of
// the waypoint by `iters` EM iterations started from the moment-matched mixture (arma::gmm_full::learn with
// keep_existing: the EM operator of this library, em_sm100.cuh).  The free samples were kept by gmm_sample_kernel
// (a.free_out / a.free_flag, at their own index) and are packed here in sample order; the mixture makes one host round trip.
// `mix` in / out: K x GMM_REC (w, mu[3], cov[9]).  Returns the number of free samples through n_free_out.
static int gmm_em_refit(mcsim_cuda_ctx* c, int K, uint64_t N, int iters, double var_floor, double* mix, unsigned* n_free_out) {
    unsigned n_free = 0;
    gmm_compact_free_kernel<<<1, 1024, 0, c->stream>>>((const float*)c->gmm_free_u.p, (const unsigned char*)c->gmm_free_flag.p, unsigned(N), unsigned(N),
                                                       (float*)c->gmm_free.p, (unsigned*)c->gmm_free_n.p);
    ++c->launches;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(&n_free, c->gmm_free_n.p, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (n_free > N) n_free = unsigned(N);
    if (n_free_out) *n_free_out = n_free;
    if (n_free < 2 || iters < 1) return MCSIM_OK;
    CK(c->em_x.reserve(sizeof(float) * 3 * size_t(n_free)));
    for (int r = 0; r < 3; ++r)
        CK(cudaMemcpyAsync((float*)c->em_x.p + size_t(r) * n_free, (const float*)c->gmm_free.p + size_t(r) * N, sizeof(float) * n_free, cudaMemcpyDeviceToDevice, c->stream));
    c->em_x_valid = true; c->em_x_n = n_free; c->em_x_fits = 1;
    std::vector<double> h(K), m(size_t(K) * 3), f(size_t(K) * 9);
    for (int k = 0; k < K; ++k) {
        h[k] = mix[size_t(k) * GMM_REC];
        for (int i = 0; i < 3; ++i) m[size_t(k) * 3 + i] = mix[size_t(k) * GMM_REC + 1 + i];
        for (int i = 0; i < 9; ++i) f[size_t(k) * 9 + i] = mix[size_t(k) * GMM_REC + 4 + i];
    }
    mcsim_em_params ep;
    memset(&ep, 0, sizeof(ep));
    ep.K = K; ep.n_fits = 1; ep.n = n_free; ep.x = nullptr;
    ep.hefts = h.data(); ep.means = m.data(); ep.fcovs = f.data();
    ep.iters = iters; ep.var_floor = var_floor;
    mcsim_em_result er;
    memset(&er, 0, sizeof(er));
    er.hefts = h.data(); er.means = m.data(); er.fcovs = f.data();
    const uint64_t h2d = c->h2d_bytes, d2h = c->d2h_bytes;
    const int rc = mcsim_cuda_em_fit(c, &ep, &er);
    c->h2d_bytes += h2d; c->d2h_bytes += d2h;
    if (rc) return rc;
    for (int k = 0; k < K; ++k) {
        mix[size_t(k) * GMM_REC] = h[k];
        for (int i = 0; i < 3; ++i) mix[size_t(k) * GMM_REC + 1 + i] = m[size_t(k) * 3 + i];
        for (int i = 0; i < 9; ++i) mix[size_t(k) * GMM_REC + 4 + i] = f[size_t(k) * 9 + i];
    }
    return MCSIM_OK;
}
// GmmLaunch.do_ekf of a run: 1 = the reference's EKF on the state, 2 = the deviation-state filter (mcsim_mc_params.gmm_filter)
static int gmm_filter_code(const mcsim_mc_params* mc, int* code) {
    if (mc->gmm_filter != MCSIM_FILTER_STATE && mc->gmm_filter != MCSIM_FILTER_DEVIATION) return fail(MCSIM_ERR_ARG, "gmm: unknown gmm_filter %d", mc->gmm_filter);
    *code = mc->gmm_filter == MCSIM_FILTER_DEVIATION ? 2 : 1;
    return MCSIM_OK;
}
static int gmm_keep_free_samples(mcsim_cuda_ctx* c, GmmLaunch& a, uint64_t N) {
    if (N >= (1ull << 31)) return fail(MCSIM_ERR_ARG, "gmm: the EM-refit variant takes fewer than 2^31 samples per waypoint");
    CK(c->gmm_free.reserve(sizeof(float) * 3 * size_t(N)));       // packed in sample order (gmm_compact_free_kernel)
    CK(c->gmm_free_u.reserve(sizeof(float) * 3 * size_t(N)));     // as the sample kernel leaves them: at the sample's own index
    CK(c->gmm_free_flag.reserve(size_t(N)));
    CK(c->gmm_free_n.reserve(sizeof(unsigned)));
    a.free_out = (float*)c->gmm_free_u.p;
    a.free_flag = (unsigned char*)c->gmm_free_flag.p;
    a.free_n = (unsigned*)c->gmm_free_n.p;
    a.free_cap = unsigned(N);
    return MCSIM_OK;
}

// the shared real-robot thread of a run (host, FP64): applied commands, motion noise and observations of every step.
// hn: (T - 1) x (3 + L) standard normals in consumption order
static int gmm_host_steps(const mcsim_mc_params* mc, const std::vector<double>& hn, std::vector<GmmStepTable>& steps) {
    const int T = mc->T, L = mc->n_landmarks;
    std::vector<double> applied(size_t(3) * (T > 1 ? T - 1 : 1)), obs(size_t(T > 1 ? T - 1 : 1) * (L ? L : 1));
    if (T > 1) {
        const int rc = mcsim_host_thread(mc, hn.data(), hn.size(), nullptr, applied.data(), nullptr, nullptr, nullptr, obs.data());
        if (rc) return fail(rc, "host thread failed (%d)", rc);
    }
    steps.assign(T > 1 ? T - 1 : 1, GmmStepTable());
    mcsim_math::NoiseModel<double> nmd;
    for (int k = 0; k < 4; ++k) nmd.alpha[k] = mc->alphas[k];
    for (int t = 0; t < T - 1; ++t) {
        double us[3];
        for (int k = 0; k < 3; ++k) { us[k] = mc->odometry[k * (T - 1) + t]; steps[t].u[k] = applied[k * (T - 1) + t]; }
        for (int k = 0; k < 3; ++k) { steps[t].xs[k] = mc->trajectory[k * T + t]; steps[t].xn[k] = mc->trajectory[k * T + t + 1]; steps[t].us[k] = us[k]; }
        mcsim_math::motion_noise(nmd, us, steps[t].M);
        for (int l = 0; l < L; ++l) steps[t].z[l] = obs[size_t(t) * L + l];
    }
    return MCSIM_OK;
}

// ---- method variant: the GMM belief propagation with the analytic linear-constraint truncation
extern "C" int mcsim_cuda_linear_run(mcsim_cuda_ctx* c, const mcsim_mc_params* mc, const mcsim_linear_params* lp, mcsim_gmm_run_result* r) {
    if (!c || !mc || !lp || !r) return fail(MCSIM_ERR_ARG, "linear_run: NULL argument");
    const int T = mc->T, K = lp->K, L = mc->n_landmarks, n_con = lp->n_constraints;
    if (T < 1 || T > MCSIM_MAX_WAYPOINTS || !mc->trajectory || (T > 1 && !mc->odometry)) return fail(MCSIM_ERR_ARG, "linear_run: bad plan");
    if (K < 1 || K > MCSIM_MAX_GAUSSIANS) return fail(MCSIM_ERR_ARG, "linear_run: K=%d out of range 1..%d", K, MCSIM_MAX_GAUSSIANS);
    if (L < 0 || L > MCSIM_MAX_LANDMARKS || (L && !mc->landmarks)) return fail(MCSIM_ERR_ARG, "linear_run: bad landmarks");
    if (n_con < 0 || n_con > 4096 || (n_con && (!lp->A || !lp->b))) return fail(MCSIM_ERR_ARG, "linear_run: bad constraints");
    CK(cudaSetDevice(c->device));
    const int per_step = 3 + L;
    std::vector<double> hn;
    if (lp->normals) {
        if (lp->n_normals < size_t(T > 1 ? T - 1 : 0) * per_step) return fail(MCSIM_ERR_DRAWS, "linear_run: %zu normals given, %zu needed", lp->n_normals, size_t(T - 1) * per_step);
        hn.assign(lp->normals, lp->normals + size_t(T > 1 ? T - 1 : 0) * per_step);
    } else {
        host_normals(mc->seed, unsigned(mc->rollout_offset), T, per_step, hn);
    }
    std::vector<GmmStepTable> steps;
    int rc = gmm_host_steps(mc, hn, steps);
    if (rc) return rc;
    rc = gmm_reserve(c, T, K);
    if (rc) return rc;
    std::vector<double> comps(size_t(K) * GMM_COMP, 0.0);
    for (int k = 0; k < K; ++k) {   // K copies of N(x*_0, Sigma_0), weights 1 / K (GM_Model::initModel, GM_Model.h:57-77)
        double* cp = &comps[size_t(k) * GMM_COMP];
        cp[0] = 1.0 / K;
        for (int i = 0; i < 3; ++i) cp[1 + i] = mc->trajectory[i * T];
        for (int i = 0; i < 9; ++i) cp[4 + i] = mc->cov0[i];
    }
    CK(c->lin_A.reserve(sizeof(double) * 3 * size_t(n_con ? n_con : 1)));
    CK(c->lin_b.reserve(sizeof(double) * size_t(n_con ? n_con : 1)));
    CK(cudaMemcpyAsync(c->gmm_comps.p, comps.data(), sizeof(double) * comps.size(), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->gmm_steps.p, steps.data(), sizeof(GmmStepTable) * steps.size(), cudaMemcpyHostToDevice, c->stream));
    if (n_con) {
        CK(cudaMemcpyAsync(c->lin_A.p, lp->A, sizeof(double) * 3 * size_t(n_con), cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->lin_b.p, lp->b, sizeof(double) * size_t(n_con), cudaMemcpyHostToDevice, c->stream));
    }
    c->h2d_bytes = sizeof(double) * comps.size() + sizeof(GmmStepTable) * steps.size() + sizeof(double) * 4 * size_t(n_con);
    GmmLaunch a;
    gmm_fill_launch(c, a, K, T, L, 0, mc->seed, unsigned(mc->rollout_offset));
    a.Q = mc->Q;
    for (int l = 0; l < L; ++l) { a.lx[l] = mc->landmarks[l]; a.ly[l] = mc->landmarks[L + l]; }
    a.rec_pre = r->gmm_pre ? (double*)c->gmm_pre.p : nullptr;
    a.rec_post = (double*)c->gmm_post.p;
    if ((rc = gmm_filter_code(mc, &a.do_ekf))) return rc;
    CK(cudaEventRecord(c->ev0, c->stream));
    gmm_linear_run_kernel<<<1, 32, 0, c->stream>>>(a, (const double*)c->lin_A.p, (const double*)c->lin_b.p, n_con);
    ++c->launches;
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->ev1, c->stream));
    std::vector<double> prob(T);
    int deg = 0;
    CK(cudaMemcpyAsync(prob.data(), c->gmm_prob.p, sizeof(double) * T, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(&deg, c->gmm_deg.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (r->gmm_pre) CK(cudaMemcpyAsync(r->gmm_pre, c->gmm_pre.p, sizeof(double) * size_t(T) * K * GMM_REC, cudaMemcpyDeviceToHost, c->stream));
    if (r->gmm_post) CK(cudaMemcpyAsync(r->gmm_post, c->gmm_post.p, sizeof(double) * K * GMM_REC, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->d2h_bytes = sizeof(double) * T + sizeof(int);
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    double free_all = 1.0;
    for (int t = 0; t < T; ++t) free_all *= (1.0 - prob[t]);
    r->P = 1.0 - free_all;
    r->degenerate = deg;
    r->pair_tests = 0;
    if (r->probabilities) memcpy(r->probabilities, prob.data(), sizeof(double) * T);
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_gmm_run(mcsim_cuda_ctx* c, const mcsim_mc_params* mc, const mcsim_gmm_run_params* g,
                                  mcsim_gmm_run_result* r) {
    if (!c || !mc || !g || !r) return fail(MCSIM_ERR_ARG, "gmm_run: NULL argument");
    if (!c->scene_set) return fail(MCSIM_ERR_ARG, "gmm_run: mcsim_cuda_set_scene has not been called");
    const int T = mc->T, K = g->K, L = mc->n_landmarks;
    const uint64_t N = g->n_samples;
    if (T < 1 || T > MCSIM_MAX_WAYPOINTS || !mc->trajectory || (T > 1 && !mc->odometry)) return fail(MCSIM_ERR_ARG, "gmm_run: bad plan");
    if (K < 1 || K > MCSIM_MAX_GAUSSIANS) return fail(MCSIM_ERR_ARG, "gmm_run: K=%d out of range 1..%d", K, MCSIM_MAX_GAUSSIANS);
    if (L < 0 || L > MCSIM_MAX_LANDMARKS || (L && !mc->landmarks)) return fail(MCSIM_ERR_ARG, "gmm_run: bad landmarks");
    if (N < 1) return fail(MCSIM_ERR_ARG, "gmm_run: n_samples must be positive");
    const int mode = mc->mode;
    const int per_step = 3 + L;
    if (c->n_poly && mode == MCSIM_MODE_REPLAY) return fail(MCSIM_ERR_ARG, "gmm_run: the replay-mode (FP64) SAT holds box hulls only (the scene has %d polygon hulls)", c->n_poly);
    CK(cudaSetDevice(c->device));
    c->h2d_bytes = c->d2h_bytes = 0;

    // ---- the shared real-robot thread (host, FP64): applied commands and observations of every step
    std::vector<double> hn;
    if (mode == MCSIM_MODE_REPLAY) {
        if (!g->normals || !g->counts) return fail(MCSIM_ERR_ARG, "gmm_run: replay mode needs normals and counts");
        const size_t need = 3 * N + size_t(T - 1) * (per_step + 3 * N);
        if (g->n_normals < need) return fail(MCSIM_ERR_DRAWS, "gmm_run: %zu normals given, %zu needed", g->n_normals, need);
        hn.resize(size_t(T - 1) * per_step);
        for (int t = 0; t < T - 1; ++t)
            memcpy(&hn[size_t(t) * per_step], g->normals + 3 * N + size_t(t) * (per_step + 3 * N), sizeof(double) * per_step);
    } else if (mode == MCSIM_MODE_PHILOX) {
        host_normals(mc->seed, unsigned(mc->rollout_offset), T, per_step, hn);
    } else {
        return fail(MCSIM_ERR_ARG, "gmm_run: unknown mode %d", mode);
    }
    std::vector<GmmStepTable> steps;
    { const int rs = gmm_host_steps(mc, hn, steps); if (rs) return rs; }

    int rc = gmm_reserve(c, T, K);
    if (rc) return rc;
    // ---- initial mixture: K copies of N(x*_0, Sigma_0), weights 1/K (GM_Model::initModel, GM_Model.h:57-77)
    std::vector<double> comps(size_t(K) * GMM_COMP);
    double Lo[6];
    if (!gmm_chol3(mc->cov0, Lo)) return fail(MCSIM_ERR_ARG, "gmm_run: initial covariance is not positive definite");
    for (int k = 0; k < K; ++k) {
        double* cp = &comps[size_t(k) * GMM_COMP];
        cp[0] = 1.0 / K;
        for (int i = 0; i < 3; ++i) cp[1 + i] = mc->trajectory[i * T];
        for (int i = 0; i < 9; ++i) cp[4 + i] = mc->cov0[i];
        for (int i = 0; i < 6; ++i) cp[13 + i] = Lo[i];
    }
    CK(cudaMemcpyAsync(c->gmm_comps.p, comps.data(), sizeof(double) * comps.size(), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->gmm_steps.p, steps.data(), sizeof(GmmStepTable) * steps.size(), cudaMemcpyHostToDevice, c->stream));
    c->h2d_bytes += sizeof(double) * comps.size() + sizeof(GmmStepTable) * steps.size();
    std::vector<double> pre0(size_t(K) * GMM_REC);
    for (int k = 0; k < K; ++k) memcpy(&pre0[size_t(k) * GMM_REC], &comps[size_t(k) * GMM_COMP], sizeof(double) * GMM_REC);
    CK(cudaMemcpyAsync(c->gmm_pre.p, pre0.data(), sizeof(double) * pre0.size(), cudaMemcpyHostToDevice, c->stream));
    std::vector<double> sn;  // replay: the 3N sample normals of every waypoint, contiguous
    if (mode == MCSIM_MODE_REPLAY) {
        sn.resize(size_t(T) * 3 * N);
        for (int t = 0; t < T; ++t) {
            const size_t off = t == 0 ? 0 : 3 * N + size_t(t - 1) * (per_step + 3 * N) + per_step;
            memcpy(&sn[size_t(t) * 3 * N], g->normals + off, sizeof(double) * 3 * N);
        }
        CK(c->gmm_normals.reserve(sizeof(double) * sn.size()));
        CK(c->gmm_counts.reserve(sizeof(int) * size_t(T) * K));
        CK(cudaMemcpyAsync(c->gmm_normals.p, sn.data(), sizeof(double) * sn.size(), cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->gmm_counts.p, g->counts, sizeof(int) * size_t(T) * K, cudaMemcpyHostToDevice, c->stream));
        c->h2d_bytes += sizeof(double) * sn.size() + sizeof(int) * size_t(T) * K;
    }
    GmmLaunch a;
    gmm_fill_launch(c, a, K, T, L, N, mc->seed, unsigned(mc->rollout_offset));
    a.Q = mc->Q;
    for (int l = 0; l < L; ++l) { a.lx[l] = mc->landmarks[l]; a.ly[l] = mc->landmarks[L + l]; }
    a.rec_pre = (double*)c->gmm_pre.p;
    a.rec_post = (double*)c->gmm_post.p;
    int filter_code = 1;
    if ((rc = gmm_filter_code(mc, &filter_code))) return rc;
    a.do_ekf = filter_code;
    a.fast_filter = mode == MCSIM_MODE_PHILOX;
    a.counts_replay = (const int*)c->gmm_counts.p;
    if (mode == MCSIM_MODE_PHILOX) { rc = gmm_upload_cull(c, a, mc->trajectory, 1, T, mc->exhaustive != 0); if (rc) return rc; }
    const bool em_refit = g->refit == MCSIM_GMM_REFIT_EM;
    if (g->refit != MCSIM_GMM_REFIT_MOMENTS && !em_refit) return fail(MCSIM_ERR_ARG, "gmm_run: unknown refit %d", g->refit);
    if (em_refit) { rc = gmm_keep_free_samples(c, a, N); if (rc) return rc; }
    CK(cudaEventRecord(c->ev0, c->stream));
    rc = -1;
    if (em_refit) {
        // per waypoint: count / sample (free samples kept) / moment refit without the EKF step, EM on the free samples
        // (one host round trip for the K x 13 mixture), then EKFpredict + EKFupdate + Cholesky of the refitted mixture
        a.do_ekf = 0;
        a.trunc_out = (double*)c->gmm_trunc.p;
        std::vector<double> mix(size_t(K) * GMM_REC);
        for (int t = 0; t < T; ++t) {
            a.t = t;
            a.normals = mode == MCSIM_MODE_REPLAY ? (const double*)c->gmm_normals.p + size_t(t) * 3 * N : nullptr;
            CK(cudaMemsetAsync(c->gmm_free_n.p, 0, sizeof(unsigned), c->stream));
            rc = gmm_launch_waypoint(c, a, mode);
            if (rc) return rc;
            CK(cudaMemcpyAsync(mix.data(), c->gmm_trunc.p, sizeof(double) * mix.size(), cudaMemcpyDeviceToHost, c->stream));
            rc = gmm_em_refit(c, K, N, g->em_iters, 1e-10, mix.data(), nullptr);
            if (rc) return rc;
            std::vector<double> cps(size_t(K) * GMM_COMP, 0.0);
            for (int k = 0; k < K; ++k) memcpy(&cps[size_t(k) * GMM_COMP], &mix[size_t(k) * GMM_REC], sizeof(double) * GMM_REC);
            CK(cudaMemcpyAsync(c->gmm_comps.p, cps.data(), sizeof(double) * cps.size(), cudaMemcpyHostToDevice, c->stream));
            GmmLaunch pa = a;
            pa.do_ekf = filter_code;
            gmm_propagate_kernel<<<1, 32, 0, c->stream>>>(pa);
            ++c->launches;
            CK(cudaGetLastError());
            CK(cudaStreamSynchronize(c->stream));   // cps is a stack vector
        }
        rc = 0;
    } else if (mode == MCSIM_MODE_PHILOX && gmm_cluster_fits(c, N, 1, g->kernel)) {
        a.t = 0;
        rc = gmm_launch_cluster(c, a, 1);
        if (rc > 0) return rc;
    }
    if (rc < 0) {
        for (int t = 0; t < T; ++t) {
            a.t = t;
            a.normals = mode == MCSIM_MODE_REPLAY ? (const double*)c->gmm_normals.p + size_t(t) * 3 * N : nullptr;
            rc = gmm_launch_waypoint(c, a, mode);
            if (rc) return rc;
        }
    }
    CK(cudaEventRecord(c->ev1, c->stream));
    std::vector<double> prob(T);
    int deg = 0;
    CK(cudaMemcpyAsync(prob.data(), c->gmm_prob.p, sizeof(double) * T, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(&deg, c->gmm_deg.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += sizeof(double) * T + sizeof(int);
    unsigned long long tests = 0;
    if (a.tests) { CK(cudaMemcpyAsync(&tests, a.tests, sizeof(tests), cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(tests); }
    if (r->gmm_pre) { CK(cudaMemcpyAsync(r->gmm_pre, c->gmm_pre.p, sizeof(double) * size_t(T) * K * GMM_REC, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(double) * size_t(T) * K * GMM_REC; }
    if (r->gmm_post) { CK(cudaMemcpyAsync(r->gmm_post, c->gmm_post.p, sizeof(double) * K * GMM_REC, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(double) * K * GMM_REC; }
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    // P = 1 - prod_t (1 - p_t)  (MCSimulator.h:848-856)
    double free_all = 1.0;
    for (int t = 0; t < T; ++t) free_all *= (1.0 - prob[t]);
    r->P = 1.0 - free_all;
    r->degenerate = deg;
    r->pair_tests = tests;
    if (r->probabilities) memcpy(r->probabilities, prob.data(), sizeof(double) * T);
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_gmm_truncate(mcsim_cuda_ctx* c, const mcsim_gmm_truncate_params* p, mcsim_gmm_truncate_result* r) {
    if (!c || !p || !r) return fail(MCSIM_ERR_ARG, "gmm_truncate: NULL argument");
    if (!c->scene_set) return fail(MCSIM_ERR_ARG, "gmm_truncate: mcsim_cuda_set_scene has not been called");
    const int K = p->K;
    const uint64_t N = p->n_samples;
    if (K < 1 || K > MCSIM_MAX_GAUSSIANS || !p->weights || !p->means || !p->covs) return fail(MCSIM_ERR_ARG, "gmm_truncate: bad mixture");
    if (p->mode == MCSIM_MODE_REPLAY && (!p->counts || !p->normals)) return fail(MCSIM_ERR_ARG, "gmm_truncate: replay mode needs counts and normals");
    if (p->mode != MCSIM_MODE_REPLAY && p->mode != MCSIM_MODE_PHILOX) return fail(MCSIM_ERR_ARG, "gmm_truncate: unknown mode %d", p->mode);
    CK(cudaSetDevice(c->device));
    int rc = gmm_reserve(c, 1, K);
    if (rc) return rc;
    std::vector<double> comps(size_t(K) * GMM_COMP);
    uint64_t total = 0;
    for (int k = 0; k < K; ++k) {
        double* cp = &comps[size_t(k) * GMM_COMP];
        cp[0] = p->weights[k];
        for (int i = 0; i < 3; ++i) cp[1 + i] = p->means[k * 3 + i];
        for (int i = 0; i < 9; ++i) cp[4 + i] = p->covs[k * 9 + i];
        if (!gmm_chol3(cp + 4, cp + 13)) return fail(MCSIM_ERR_ARG, "gmm_truncate: covariance of component %d is not positive definite", k);
        if (p->mode == MCSIM_MODE_REPLAY) { if (p->counts[k] < 0) return fail(MCSIM_ERR_ARG, "gmm_truncate: negative count"); total += uint64_t(p->counts[k]); }
    }
    if (p->mode == MCSIM_MODE_REPLAY && total != N) return fail(MCSIM_ERR_DRAWS, "gmm_truncate: counts sum to %llu, n_samples is %llu", (unsigned long long)total, (unsigned long long)N);
    CK(cudaMemcpyAsync(c->gmm_comps.p, comps.data(), sizeof(double) * comps.size(), cudaMemcpyHostToDevice, c->stream));
    if (p->mode == MCSIM_MODE_REPLAY) {
        CK(c->gmm_normals.reserve(sizeof(double) * 3 * (N ? N : 1)));
        CK(c->gmm_counts.reserve(sizeof(int) * K));
        if (N) CK(cudaMemcpyAsync(c->gmm_normals.p, p->normals, sizeof(double) * 3 * N, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->gmm_counts.p, p->counts, sizeof(int) * K, cudaMemcpyHostToDevice, c->stream));
    }
    GmmLaunch a;
    gmm_fill_launch(c, a, K, 1, 0, N, p->seed, unsigned(p->stream_id));
    a.t = 0;
    a.counts_replay = (const int*)c->gmm_counts.p;
    a.normals = (const double*)c->gmm_normals.p;
    a.trunc_out = (double*)c->gmm_trunc.p;
    a.n_out = (unsigned long long*)c->gmm_nout.p;
    a.do_ekf = 0;
    const bool em_refit = p->refit == MCSIM_GMM_REFIT_EM;
    if (p->refit != MCSIM_GMM_REFIT_MOMENTS && !em_refit) return fail(MCSIM_ERR_ARG, "gmm_truncate: unknown refit %d", p->refit);
    if (em_refit) {
        rc = gmm_keep_free_samples(c, a, N);
        if (rc) return rc;
        CK(cudaMemsetAsync(c->gmm_free_n.p, 0, sizeof(unsigned), c->stream));
    }
    rc = gmm_launch_waypoint(c, a, p->mode);
    if (rc) return rc;
    std::vector<double> out(size_t(K) * GMM_REC);
    unsigned long long nout[2 * MCSIM_MAX_GAUSSIANS];
    double prob = 0;
    int deg = 0;
    CK(cudaMemcpyAsync(out.data(), c->gmm_trunc.p, sizeof(double) * out.size(), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(nout, c->gmm_nout.p, sizeof(unsigned long long) * 2 * K, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(&prob, c->gmm_prob.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(&deg, c->gmm_deg.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (em_refit) { rc = gmm_em_refit(c, K, N, p->em_iters, 1e-10, out.data(), nullptr); if (rc) return rc; }
    memset(r, 0, sizeof(*r));
    for (int k = 0; k < K; ++k) {
        r->weights[k] = out[size_t(k) * GMM_REC];
        for (int i = 0; i < 3; ++i) r->means[k * 3 + i] = out[size_t(k) * GMM_REC + 1 + i];
        for (int i = 0; i < 9; ++i) r->covs[k * 9 + i] = out[size_t(k) * GMM_REC + 4 + i];
        r->n_coll[k] = nout[k];
        r->n_free[k] = nout[K + k];
    }
    r->p_collision = prob;
    r->degenerate = deg;
    return MCSIM_OK;
}

/* bytes the last *_run call copied host->device and device->host (bench.py's e2e accounting) */
extern "C" void mcsim_cuda_io_bytes(mcsim_cuda_ctx* c, uint64_t* h2d, uint64_t* d2h) {
    if (h2d) *h2d = c ? c->h2d_bytes : 0;
    if (d2h) *d2h = c ? c->d2h_bytes : 0;
}

// ---- EM fit (BASELINE config 5) ---------------------------------------------------------------------
typedef void (*em_kernel_t)(const EmLaunch);
#ifndef EM_TMA_MAX_K
#define EM_TMA_MAX_K 4   // fits with at most this many components stream through the bulk-copy kernel (packed pairs)
#endif
template <int K>
static void em_pick(em_kernel_t& k0, em_kernel_t& k1, bool vec, size_t& smem, int& block) {
    k0 = em_estep_kernel<K, 0>;
    k1 = nullptr;
    smem = 0;
    block = 256;
#ifndef EM_PACKED_MAX_K
#define EM_PACKED_MAX_K 5   // packed pairs while the 10 K accumulator pairs fit the register file (K = 5: 230 registers; K = 6 spills).  On config 5
                            // (56 fits) the packed kernel beats the 128-register one-warp kernel for K = 4 (0.606 vs 0.549 of the FP32 peak) and
                            // K = 5 (0.577 vs 0.499); on 16 fits it is the other way round (profiles/r02_em_shifted_moments.log)
#endif
    if constexpr (K <= EM_PACKED_MAX_K) k0 = em_estep2_kernel<K, false>;  // packed: two samples per thread
    if constexpr (K <= EM_TMA_MAX_K) {  // fed by the bulk-copy engine: K = 1 is HBM-bound; K = 2..4 gain the prefetch registers back
        if (vec) { k0 = em_estep_tma_kernel<K>; smem = sizeof(float) * EM_TMA_STAGES * 3 * EM_TMA_TILE; }
    }
#ifndef EM_SPLIT_S_LOW
#define EM_SPLIT_S_LOW 2    // warps per group for 5 <= K <= 8
#endif
#ifndef EM_SPLIT_S_HIGH
#define EM_SPLIT_S_HIGH 2   // for K > 8 (measured: 16 components 8.08 ms in two launches, 6.07 ms with 4 warps per group, 5.38 ms with 2)
#endif
#ifndef EM_SPLIT_MIN_K
#define EM_SPLIT_MIN_K (EM_OW_HI + 1)    // 5 <= K <= 8: the one-warp kernel is still 5-15 % faster than a group of two
#endif
    // K >= 5: the components are split over the warps of a group (em_estep_split_kernel): one launch for any K <= 16
    if constexpr (K >= EM_SPLIT_MIN_K && K <= EM_OW_HI) k0 = em_estep_split_kernel<K, EM_SPLIT_S_LOW>;
    if constexpr (K > EM_OW_HI) {
        if constexpr (EM_SPLIT_MIN_K <= 16) k0 = em_estep_split_kernel<K, EM_SPLIT_S_HIGH>;
        else k1 = em_estep_kernel<K, EM_OW_HI>;
    }
    if constexpr (K >= EM_OW_LO && K > EM_PACKED_MAX_K && K <= EM_OW_HI && K < EM_SPLIT_MIN_K) block = EM_OW_THREADS68;   // the one-warp kernel's own CTA size there
}
static bool em_kernels(int K, em_kernel_t& k0, em_kernel_t& k1, bool vec, size_t& smem, int& block) {
    switch (K) {
        case 1: em_pick<1>(k0, k1, vec, smem, block); return true;    case 2: em_pick<2>(k0, k1, vec, smem, block); return true;
        case 3: em_pick<3>(k0, k1, vec, smem, block); return true;    case 4: em_pick<4>(k0, k1, vec, smem, block); return true;
        case 5: em_pick<5>(k0, k1, vec, smem, block); return true;    case 6: em_pick<6>(k0, k1, vec, smem, block); return true;
        case 7: em_pick<7>(k0, k1, vec, smem, block); return true;    case 8: em_pick<8>(k0, k1, vec, smem, block); return true;
        case 9: em_pick<9>(k0, k1, vec, smem, block); return true;    case 10: em_pick<10>(k0, k1, vec, smem, block); return true;
        case 11: em_pick<11>(k0, k1, vec, smem, block); return true;  case 12: em_pick<12>(k0, k1, vec, smem, block); return true;
        case 13: em_pick<13>(k0, k1, vec, smem, block); return true;  case 14: em_pick<14>(k0, k1, vec, smem, block); return true;
        case 15: em_pick<15>(k0, k1, vec, smem, block); return true;  case 16: em_pick<16>(k0, k1, vec, smem, block); return true;
    }
    return false;
}

extern "C" int mcsim_cuda_em_set_moments(mcsim_cuda_ctx* c, int form) {
    if (!c) return fail(MCSIM_ERR_ARG, "em_set_moments: NULL context");
    if (form != MCSIM_EM_MOMENTS_AUTO && form != MCSIM_EM_MOMENTS_CENTRED && form != MCSIM_EM_MOMENTS_SHIFTED)
        return fail(MCSIM_ERR_ARG, "em_set_moments: form %d is none of MCSIM_EM_MOMENTS_AUTO / _CENTRED / _SHIFTED", form);
    c->em_moments = form;
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_em_moments_in_use(mcsim_cuda_ctx* c, int fit, int* form) {
    if (!c || !form || !c->em_prepared) return fail(MCSIM_ERR_ARG, "em_moments_in_use: nothing prepared");
    if (fit < 0 || fit >= c->em.n_fits) return fail(MCSIM_ERR_ARG, "em_moments_in_use: fit %d out of range", fit);
    *form = MCSIM_EM_MOMENTS_CENTRED;
    if (c->em.K < 2) return MCSIM_OK;
    CK(cudaSetDevice(c->device));
    float flag = 0.f;   // header of the fit: spare floats of its first two records (em_sm100.cuh, em_header)
    CK(cudaMemcpyAsync(&flag, c->em.comps + (size_t(fit) * c->em.K + 1) * EM_COMP_STRIDE + 11, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (flag != 0.f) *form = MCSIM_EM_MOMENTS_SHIFTED;
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_em_prepare(mcsim_cuda_ctx* c, const mcsim_em_params* p) {
    if (!c || !p) return fail(MCSIM_ERR_ARG, "em: NULL context or params");
    if (p->K < 1 || p->K > MCSIM_MAX_GAUSSIANS) return fail(MCSIM_ERR_ARG, "em: K=%d out of range 1..%d", p->K, MCSIM_MAX_GAUSSIANS);
    if (p->n_fits < 1 || p->n_fits > 65535) return fail(MCSIM_ERR_ARG, "em: n_fits=%d out of range 1..65535", p->n_fits);
    if (p->iters < 0) return fail(MCSIM_ERR_ARG, "em: negative iteration count");
    if (p->n >= (1ull << 31) - (1ull << 24)) return fail(MCSIM_ERR_ARG, "em: more than 2^31 samples per fit and shard");
    if (!p->hefts || !p->means || !p->fcovs) return fail(MCSIM_ERR_ARG, "em: initial parameters are NULL (seed mode is keep_existing)");
    if (!p->x && !(c->em_x_valid && c->em_x_n == p->n && c->em_x_fits == p->n_fits))
        return fail(MCSIM_ERR_ARG, "em: x is NULL and no resident samples of this shape");
    CK(cudaSetDevice(c->device));
    const int K = p->K, F = p->n_fits;
    const size_t nx = size_t(F) * 3 * p->n;
    c->h2d_bytes = sizeof(double) * size_t(F) * K * 13;
    if (p->x) {
        CK(c->em_x.reserve(sizeof(float) * (nx ? nx : 1)));
        if (nx && p->x_row_stride && p->x_row_stride != p->n) {   // a column slice of wider rows (one shard of a sharded fit)
            if (p->x_row_stride < p->n) return fail(MCSIM_ERR_ARG, "em: x_row_stride smaller than n");
            CK(cudaMemcpy2DAsync(c->em_x.p, sizeof(float) * p->n, p->x, sizeof(float) * p->x_row_stride, sizeof(float) * p->n, size_t(F) * 3, cudaMemcpyHostToDevice, c->stream));
        } else if (nx) CK(cudaMemcpyAsync(c->em_x.p, p->x, sizeof(float) * nx, cudaMemcpyHostToDevice, c->stream));
        c->em_x_valid = true; c->em_x_n = p->n; c->em_x_fits = F;
        c->h2d_bytes += sizeof(float) * nx;
    }
    const int iters = p->iters > 0 ? p->iters : 1;
    CK(c->em_comps.reserve(sizeof(float) * size_t(F) * K * EM_COMP_STRIDE));
    CK(c->em_hefts.reserve(sizeof(double) * size_t(F) * K));
    CK(c->em_means.reserve(sizeof(double) * size_t(F) * K * 3));
    CK(c->em_fcovs.reserve(sizeof(double) * size_t(F) * K * 9));
    CK(c->em_stats.reserve(sizeof(double) * size_t(F) * (K * EM_NSTAT + 1)));
    CK(c->em_old.reserve(sizeof(double) * F));
    CK(c->em_done.reserve(sizeof(int) * F));
    CK(c->em_iters.reserve(sizeof(int) * F));
    CK(c->em_trace.reserve(sizeof(double) * size_t(F) * iters));
    CK(cudaMemcpyAsync(c->em_hefts.p, p->hefts, sizeof(double) * size_t(F) * K, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->em_means.p, p->means, sizeof(double) * size_t(F) * K * 3, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->em_fcovs.p, p->fcovs, sizeof(double) * size_t(F) * K * 9, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(c->em_stats.p, 0, sizeof(double) * size_t(F) * (K * EM_NSTAT + 1), c->stream));
    CK(cudaMemsetAsync(c->em_done.p, 0, sizeof(int) * F, c->stream));
    CK(cudaMemsetAsync(c->em_iters.p, 0, sizeof(int) * F, c->stream));
    {   // old_avg_log_p = -inf (:2436); trace = NaN
        std::vector<double> init(size_t(F) * iters + F);
        for (int f = 0; f < F; ++f) init[f] = -INFINITY;
        for (size_t i = 0; i < size_t(F) * iters; ++i) init[F + i] = NAN;
        CK(cudaMemcpyAsync(c->em_old.p, init.data(), sizeof(double) * F, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->em_trace.p, init.data() + F, sizeof(double) * size_t(F) * iters, cudaMemcpyHostToDevice, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    EmLaunch& a = c->em;
    a.x = (const float*)c->em_x.p;
    a.n = p->n;
    a.n_total = double(p->n_total ? p->n_total : p->n);
    a.K = K; a.n_fits = F; a.iters = p->iters; a.iter = 0;
    a.n_peers = 0; a.my_rank = 0; a.wait_value = 0;
    a.moments = c->em_moments;
    a.var_floor = p->var_floor > 0 ? p->var_floor : DBL_MIN;
    a.comps = (float*)c->em_comps.p; a.hefts = (double*)c->em_hefts.p; a.means = (double*)c->em_means.p;
    a.fcovs = (double*)c->em_fcovs.p; a.stats = (double*)c->em_stats.p; a.old_avg = (double*)c->em_old.p;
    a.done = (int*)c->em_done.p; a.iters_run = (int*)c->em_iters.p; a.trace = (double*)c->em_trace.p;
    em_mstep_kernel<<<F, 32, 0, c->stream>>>(a, 1);
    ++c->launches;
    CK(cudaGetLastError());
    c->em_prepared = true;
    c->em_chunks = 0;   // the next E-step picks kernel and grid for this shape
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_em_estep(mcsim_cuda_ctx* c) {
    if (!c || !c->em_prepared) return fail(MCSIM_ERR_ARG, "em_estep: mcsim_cuda_em_prepare has not been called");
    CK(cudaSetDevice(c->device));
    if (c->em_chunks) {   // configuration cached by the first E-step after mcsim_cuda_em_prepare
        const dim3 grid(c->em_chunks, (unsigned)c->em.n_fits);
        c->em_k0<<<grid, c->em_block, c->em_smem, c->stream>>>(c->em);
        ++c->launches;
        if (c->em_k1) { c->em_k1<<<grid, c->em_block, 0, c->stream>>>(c->em); ++c->launches; }
        CK(cudaGetLastError());
        return MCSIM_OK;
    }
    em_kernel_t k0, k1;
    // 128-bit loads when every row of every fit is 16-byte aligned
    const bool vec = (c->em.n % 4) == 0 && c->em.n >= 4;
    size_t smem = 0;
    int block = 256;
    if (!em_kernels(c->em.K, k0, k1, vec, smem, block)) return fail(MCSIM_ERR_ARG, "em_estep: bad K");
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute((const void*)k0, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    // chunks per fit: whole waves of resident CTAs with as little tail as possible (616 CTAs on 444 resident slots
    // ran at 70 % of what 392 do), at least ~8 samples per thread so the per-CTA reduction stays small
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)k0, block, smem));
    if (k1) { int o1 = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, (const void*)k1, block, 0)); if (o1 < occ) occ = o1; }
    if (occ < 1) occ = 1;
    const unsigned long long resident = (unsigned long long)c->n_sm * occ, F = (unsigned long long)c->em.n_fits;
    const unsigned long long per_cta = smem ? EM_TMA_TILE : (unsigned long long)block * 8;
    const unsigned long long most = (c->em.n + per_cta - 1) / per_cta ? (c->em.n + per_cta - 1) / per_cta : 1;
    unsigned long long chunks = 1;
    double best = -1.0;
    for (unsigned long long w = 1; w <= 4; ++w) {
        unsigned long long ch = w * resident / F;
        if (ch < 1) ch = 1;
        if (ch > most) ch = most;
        const unsigned long long ctas = ch * F, waves = (ctas + resident - 1) / resident;
        const double eff = double(ctas) / double(waves * resident);
        if (eff > best + 0.02) { best = eff; chunks = ch; }
    }
    const dim3 grid((unsigned)chunks, (unsigned)c->em.n_fits);
    c->em_k0 = k0; c->em_k1 = k1; c->em_smem = smem; c->em_chunks = (unsigned)chunks; c->em_block = block;
    k0<<<grid, block, smem, c->stream>>>(c->em);
    ++c->launches;
    if (k1) { k1<<<grid, block, 0, c->stream>>>(c->em); ++c->launches; }
    CK(cudaGetLastError());
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_em_mstep(mcsim_cuda_ctx* c) {
    if (!c || !c->em_prepared) return fail(MCSIM_ERR_ARG, "em_mstep: mcsim_cuda_em_prepare has not been called");
    CK(cudaSetDevice(c->device));
    em_mstep_kernel<<<c->em.n_fits, c->em.n_peers > 1 ? EM_MSTEP_FUSED_THREADS : 32, 0, c->stream>>>(c->em, 0);
    ++c->launches;
    ++c->em.iter;
    CK(cudaGetLastError());
    return MCSIM_OK;
}

extern "C" void* mcsim_cuda_em_stats(mcsim_cuda_ctx* c, uint64_t* n_doubles) {
    if (!c || !c->em_prepared) return nullptr;
    if (n_doubles) *n_doubles = uint64_t(c->em.n_fits) * (c->em.K * EM_NSTAT + 1);
    return c->em_stats.p;
}

extern "C" int mcsim_cuda_em_collect(mcsim_cuda_ctx* c, mcsim_em_result* r) {
    if (!c || !r || !c->em_prepared) return fail(MCSIM_ERR_ARG, "em_collect: nothing prepared");
    CK(cudaSetDevice(c->device));
    const size_t F = c->em.n_fits, K = c->em.K;
    c->d2h_bytes = 0;
    if (r->hefts) { CK(cudaMemcpyAsync(r->hefts, c->em_hefts.p, sizeof(double) * F * K, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(double) * F * K; }
    if (r->means) { CK(cudaMemcpyAsync(r->means, c->em_means.p, sizeof(double) * F * K * 3, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(double) * F * K * 3; }
    if (r->fcovs) { CK(cudaMemcpyAsync(r->fcovs, c->em_fcovs.p, sizeof(double) * F * K * 9, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += sizeof(double) * F * K * 9; }
    if (r->avg_log_p && c->em.iters > 0) CK(cudaMemcpyAsync(r->avg_log_p, c->em_trace.p, sizeof(double) * F * c->em.iters, cudaMemcpyDeviceToHost, c->stream));
    if (r->iters_run) CK(cudaMemcpyAsync(r->iters_run, c->em_iters.p, sizeof(int) * F, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_em_fit(mcsim_cuda_ctx* c, const mcsim_em_params* p, mcsim_em_result* r) {
    if (!r) return fail(MCSIM_ERR_ARG, "em_fit: result is NULL");
    int rc = mcsim_cuda_em_prepare(c, p);
    if (rc) return rc;
    CK(cudaEventRecord(c->ev0, c->stream));
    for (int it = 0; it < p->iters; ++it) {
        if ((rc = mcsim_cuda_em_estep(c))) return rc;
        if ((rc = mcsim_cuda_em_mstep(c))) return rc;
    }
    CK(cudaEventRecord(c->ev1, c->stream));
    rc = mcsim_cuda_em_collect(c, r);
    if (rc) return rc;
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    return MCSIM_OK;
}

// ---- batched plan scoring (BASELINE config 4) -------------------------------------------------------
typedef void (*score_kernel_t)(const ScoreLaunch);
static score_kernel_t pick_score(int est, bool ee) {
    if (est == MCSIM_EST_CLOSEDLOOP) return ee ? score_plans_kernel<MCSIM_EST_CLOSEDLOOP, 8, true> : score_plans_kernel<MCSIM_EST_CLOSEDLOOP, 8, false>;
    return ee ? score_plans_kernel<MCSIM_EST_REFERENCE, 8, true> : score_plans_kernel<MCSIM_EST_REFERENCE, 8, false>;
}

extern "C" int mcsim_cuda_score_prepare(mcsim_cuda_ctx* c, const mcsim_score_params* p) {
    if (!c || !p) return fail(MCSIM_ERR_ARG, "score: NULL context or params");
    if (!c->scene_set) return fail(MCSIM_ERR_ARG, "score: mcsim_cuda_set_scene has not been called");
    if (c->n_poly) return fail(MCSIM_ERR_ARG, "score: batched scoring runs the packed kernel, whose tables hold box hulls only (the scene has %d polygon hulls)", c->n_poly);
    if (p->T < 2 || p->T > MCSIM_MAX_WAYPOINTS) return fail(MCSIM_ERR_ARG, "score: T=%d out of range 2..%d", p->T, MCSIM_MAX_WAYPOINTS);
    if (p->n_plans < 1 || !p->trajectories || !p->odometry) return fail(MCSIM_ERR_ARG, "score: no plans");
    if (p->n_landmarks != 8 || !p->landmarks) return fail(MCSIM_ERR_ARG, "score: this path is built for 8 landmarks (got %d)", p->n_landmarks);
    if (p->rollouts > (1ull << 32)) return fail(MCSIM_ERR_ARG, "score: more than 2^32 rollouts per plan");
    CK(cudaSetDevice(c->device));
    const int T = p->T, L = p->n_landmarks, NP = p->n_plans;
    McPlanConst pc;
    memset(&pc, 0, sizeof(pc));
    pc.T = T;
    double Lo[6];
    if (!chol3(p->cov0, Lo)) return fail(MCSIM_ERR_ARG, "score: initial covariance is not symmetric positive definite");
    for (int k = 0; k < 6; ++k) { pc.chold[k] = Lo[k]; pc.chol[k] = float(Lo[k]); }
    pc.sqrtQ = float(sqrt(p->Q));
    mcsim_math::NoiseModel<double> nmd;
    for (int k = 0; k < 4; ++k) { pc.nm.alpha[k] = float(p->alphas[k]); nmd.alpha[k] = p->alphas[k]; }
    pc.nm.Q = float(p->Q);
    pc.nm.L = L;
    for (int l = 0; l < L; ++l) { pc.nm.lx[l] = float(p->landmarks[l]); pc.nm.ly[l] = float(p->landmarks[L + l]); }
    std::vector<float> tab(size_t(NP) * T * MC_PLAN2_STRIDE, 0.f);
    for (int q = 0; q < NP; ++q) {
        const double* tr = p->trajectories + size_t(q) * 3 * T;
        const double* od = p->odometry + size_t(q) * 3 * (T - 1);
        for (int t = 0; t < T; ++t) {
            float* w = &tab[(size_t(q) * T + t) * MC_PLAN2_STRIDE];
            for (int k = 0; k < 3; ++k) w[k] = float(tr[k * T + t]);
            w[10] = float(mcsim_math::angle_wrap(tr[2 * T + t]));
            if (t < T - 1) {
                double us[3], M[3];
                for (int k = 0; k < 3; ++k) { us[k] = od[k * (T - 1) + t]; w[3 + k] = float(us[k]); }
                mcsim_math::motion_noise(nmd, us, M);  // generateM_EKF(u*_t), MCSimulator.h:698-701
                for (int k = 0; k < 3; ++k) w[6 + k] = float(M[k]);
                w[9] = float(M[0]) + float(M[2]);
            }
        }
    }
    CK(c->sc_plans.reserve(tab.size() * sizeof(float)));
    CK(c->sc_hits.reserve(sizeof(unsigned long long) * NP));
    CK(c->sc_counters.reserve(sizeof(unsigned long long) * 2));
    CK(cudaStreamSynchronize(c->stream));
    c->sc_sym.pc = pc;                     // uploaded to c_mc by score_launch (bind_symbols); c_plan2 is not used by scoring
    c->sc_sym.plan2.clear();
    c->sc_sym.tag = g_next_tag++;
    CK(cudaMemcpyAsync(c->sc_plans.p, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(c->sc_hits.p, 0, sizeof(unsigned long long) * NP, c->stream));
    CK(cudaMemsetAsync(c->sc_counters.p, 0, sizeof(unsigned long long) * 2, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->h2d_bytes = sizeof(pc) + tab.size() * sizeof(float);
    ScoreLaunch& a = c->sc;
    a.plans = (const float*)c->sc_plans.p;
    a.pairs = (const float4*)c->pairs_f.p;
    a.T = T; a.n_plans = NP; a.n_pairs = c->n_pairs;
    // CTAs = plans x chunks of a plan's rollouts, in whole waves of resident CTAs: every CTA of a launch runs
    // the same number of rollout-steps, so n CTAs on s resident slots take ceil(n / s) CTA lifetimes and a
    // launch of 1.08 waves (128 plans x 5 chunks on 592 slots) runs at 54 % of the machine.  Take the chunk
    // count with the fullest last wave; a chunk keeps >= 512 rollouts (two passes of a CTA) when the plan has
    // them, so that the per-CTA prologue (tables + candidate lists) stays below 1 % of its work.  Rollout ids
    // do not depend on the split (plan id << 32 | rollout), so the hits do not either.
    {
        score_kernel_t k = pick_score(p->estimator, p->early_exit != 0);
        const size_t smem = score_smem_bytes(T, a.n_pairs, c->n_hulls);
        int per_sm = 4;
        if (smem <= 200 * 1024) {
            if (smem > 40 * 1024) CK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)k, 128, smem));
            if (per_sm < 1) per_sm = 1;
        }
        const unsigned long long slots = (unsigned long long)c->n_sm * per_sm;
        unsigned long long max_chunks = p->rollouts / 512;
        if (max_chunks < 1) max_chunks = (p->rollouts + 255) / 256;
        if (max_chunks < 1) max_chunks = 1;
        if (max_chunks > 4096) max_chunks = 4096;
        unsigned long long best_ch = 1;
        double best = -1.0;
        for (unsigned long long ch = 1; ch <= max_chunks; ++ch) {
            const unsigned long long n = (unsigned long long)NP * ch, waves = (n + slots - 1) / slots;
            // rollouts of a chunk are rounded up to whole passes of 256: count the padding as lost work too
            const unsigned long long per = (p->rollouts + ch - 1) / ch, passes = (per + 255) / 256;
            const double eff = double(p->rollouts) * NP / (double(waves) * slots * passes * 256.0);
            if (eff > best + 0.005) { best = eff; best_ch = ch; }
        }
        a.chunks = int(best_ch);
    }
    a.rollouts = p->rollouts;
    a.plan_offset = p->plan_offset;
    a.k0 = unsigned(p->seed); a.k1 = unsigned(p->seed >> 32);
    a.margin = float(p->margin);
    a.hits = (unsigned long long*)c->sc_hits.p;
    a.counters = (unsigned long long*)c->sc_counters.p;
    c->sc_ee = p->early_exit; c->sc_est = p->estimator;
    c->sc_prepared = true;
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_score_launch(mcsim_cuda_ctx* c) {
    if (!c || !c->sc_prepared) return fail(MCSIM_ERR_ARG, "score_launch: mcsim_cuda_score_prepare has not been called");
    CK(cudaSetDevice(c->device));
    const ScoreLaunch& a = c->sc;
    SymbolLock sym;
    { const int rb = bind_symbols(c, &c->sc_sym, sym); if (rb) return rb; }
    score_kernel_t k = pick_score(c->sc_est, c->sc_ee != 0);
    const size_t smem = score_smem_bytes(a.T, a.n_pairs, c->n_hulls);
    if (smem > 200 * 1024) return fail(MCSIM_ERR_ARG, "score: %zu bytes of shared memory needed (T=%d, %d pairs)", smem, a.T, a.n_pairs);
    if (smem > 40 * 1024) CK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    k<<<unsigned(a.n_plans) * unsigned(a.chunks), 128, smem, c->stream>>>(a);
    ++c->launches;
    CK(cudaGetLastError());
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_score_collect(mcsim_cuda_ctx* c, mcsim_score_result* r) {
    if (!c || !r || !c->sc_prepared) return fail(MCSIM_ERR_ARG, "score_collect: nothing prepared");
    CK(cudaSetDevice(c->device));
    const int NP = c->sc.n_plans;
    std::vector<unsigned long long> h(size_t(NP) + 2);
    CK(cudaMemcpyAsync(h.data(), c->sc_hits.p, sizeof(unsigned long long) * NP, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(h.data() + NP, c->sc_counters.p, sizeof(unsigned long long) * 2, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemsetAsync(c->sc_hits.p, 0, sizeof(unsigned long long) * NP, c->stream));
    CK(cudaMemsetAsync(c->sc_counters.p, 0, sizeof(unsigned long long) * 2, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->d2h_bytes = sizeof(unsigned long long) * (size_t(NP) + 2);
    for (int q = 0; q < NP; ++q) {
        if (r->hits) r->hits[q] = h[q];
        if (r->P) r->P[q] = c->sc.rollouts ? double(h[q]) / double(c->sc.rollouts) : 0.0;
    }
    r->executed_steps = h[NP];
    r->pair_tests = h[NP + 1];
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_score_plans(mcsim_cuda_ctx* c, const mcsim_score_params* p, mcsim_score_result* r) {
    if (!r) return fail(MCSIM_ERR_ARG, "score_plans: result is NULL");
    int rc = mcsim_cuda_score_prepare(c, p);
    if (rc) return rc;
    CK(cudaEventRecord(c->ev0, c->stream));
    if ((rc = mcsim_cuda_score_launch(c))) return rc;
    CK(cudaEventRecord(c->ev1, c->stream));
    if ((rc = mcsim_cuda_score_collect(c, r))) return rc;
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    return MCSIM_OK;
}

// ---- batched GMM mode: n_traj independent runGMMEstimation runs in the same launches ----------------
// Threads per run of the one-CTA-per-run kernel.  512 resident threads per SM (128 registers each) are 1, 2 or 4 runs in
// flight per SM; a run on fewer threads takes longer but more of them overlap their serial stretches (refit, barriers).
// Measured wave times on B200 (K = 3, 10^4 samples, 56 waypoints; profiles/r02_gmm_member_shapes.log): 1.5 / 2.45 / 4.63 ms
// for a full wave of 148 / 296 / 592 runs -- small batches want the big CTA (latency), large ones the small CTA
// (throughput: 98 / 121 / 128 thousand runs per second); the estimate below picks by whole waves.
static int gmm_member_threads(int B, int n_sm) {
    const int threads[3] = {512, 256, 128};
    const double wave_ms[3] = {1.5, 2.45, 4.63};
    int best = 256;
    double best_ms = 1e300;
    for (int i = 0; i < 3; ++i) {
        const long long slots = (long long)n_sm * (GMM_MEMBER_MAX_THREADS / threads[i]);
        const double ms = double((B + slots - 1) / slots) * wave_ms[i];
        if (ms < best_ms) { best_ms = ms; best = threads[i]; }
    }
    return best;
}

extern "C" int mcsim_cuda_gmm_run_batch(mcsim_cuda_ctx* c, const mcsim_mc_params* mc, const mcsim_gmm_batch_params* g,
                                        mcsim_gmm_batch_result* r) {
    if (!c || !mc || !g || !r || !r->P) return fail(MCSIM_ERR_ARG, "gmm_run_batch: NULL argument");
    if (!c->scene_set) return fail(MCSIM_ERR_ARG, "gmm_run_batch: mcsim_cuda_set_scene has not been called");
    if (mc->mode != MCSIM_MODE_PHILOX) return fail(MCSIM_ERR_ARG, "gmm_run_batch: Philox mode only");
    const int T = mc->T, K = g->K, L = mc->n_landmarks, B = g->n_traj;
    const uint64_t N = g->n_samples;
    if (T < 1 || T > MCSIM_MAX_WAYPOINTS || !mc->trajectory || (T > 1 && !mc->odometry)) return fail(MCSIM_ERR_ARG, "gmm_run_batch: bad plan");
    if (K < 1 || K > MCSIM_MAX_GAUSSIANS || N < 1 || B < 1 || B > 65535) return fail(MCSIM_ERR_ARG, "gmm_run_batch: bad K / n_samples / n_traj");
    if (L < 0 || L > MCSIM_MAX_LANDMARKS || (L && !mc->landmarks)) return fail(MCSIM_ERR_ARG, "gmm_run_batch: bad landmarks");
    if ((g->trajectories == nullptr) != (g->odometry == nullptr)) return fail(MCSIM_ERR_ARG, "gmm_run_batch: trajectories and odometry go together");
    CK(cudaSetDevice(c->device));
    const int per_step = 3 + L, TS = T > 1 ? T - 1 : 1;
    double Lo[6];
    if (!gmm_chol3(mc->cov0, Lo)) return fail(MCSIM_ERR_ARG, "gmm_run_batch: initial covariance is not positive definite");
    int rc = gmm_reserve(c, T, K, B);
    if (rc) return rc;
    // ---- the members' shared real-robot threads (FP64, O(T) each, independent) run on the device, one warp per
    // member (gmm_robot_thread_kernel): only the plan(s) cross PCIe, the step tables and initial mixtures are
    // written where the run reads them
    const size_t n_plans = g->trajectories ? size_t(B) : 1, tr_n = size_t(3) * T, od_n = size_t(3) * (T > 1 ? T - 1 : 0);
    CK(c->gmm_plans.reserve(sizeof(double) * n_plans * (tr_n + od_n + 1)));
    CK(c->gmm_robot_normals.reserve(sizeof(double) * size_t(B) * TS * per_step));
    double* d_tr = (double*)c->gmm_plans.p;
    double* d_od = d_tr + n_plans * tr_n;
    CK(cudaMemcpyAsync(d_tr, g->trajectories ? g->trajectories : mc->trajectory, sizeof(double) * n_plans * tr_n, cudaMemcpyHostToDevice, c->stream));
    if (od_n) CK(cudaMemcpyAsync(d_od, g->odometry ? g->odometry : mc->odometry, sizeof(double) * n_plans * od_n, cudaMemcpyHostToDevice, c->stream));
    c->h2d_bytes = sizeof(double) * n_plans * (tr_n + od_n);
    {
        RobotThreadLaunch rt;
        memset(&rt, 0, sizeof(rt));
        rt.B = B; rt.T = T; rt.L = L; rt.K = K;
        rt.k0 = unsigned(mc->seed); rt.k1 = unsigned(mc->seed >> 32); rt.stream0 = unsigned(g->stream0);
        for (int k = 0; k < 4; ++k) rt.alphas[k] = mc->alphas[k];
        rt.Q = mc->Q;
        for (int k = 0; k < 9; ++k) rt.cov0[k] = mc->cov0[k];
        for (int k = 0; k < 6; ++k) rt.chol0[k] = Lo[k];
        for (int l = 0; l < L; ++l) { rt.lx[l] = mc->landmarks[l]; rt.ly[l] = mc->landmarks[L + l]; }
        rt.trajectories = d_tr; rt.odometry = d_od; rt.per_member = g->trajectories ? 1 : 0;
        rt.normals = (double*)c->gmm_robot_normals.p;
        rt.steps = (GmmStepTable*)c->gmm_steps.p;
        rt.comps = (double*)c->gmm_comps.p;
        gmm_robot_thread_kernel<<<B, 32, 0, c->stream>>>(rt);
        ++c->launches;
        CK(cudaGetLastError());
    }
    GmmLaunch a;
    gmm_fill_launch(c, a, K, T, L, N, mc->seed, unsigned(g->stream0));
    a.Q = mc->Q;
    for (int l = 0; l < L; ++l) { a.lx[l] = mc->landmarks[l]; a.ly[l] = mc->landmarks[L + l]; }
    if ((rc = gmm_filter_code(mc, &a.do_ekf))) return rc;
    a.fast_filter = 1;   // batches are Philox mode
    a.B = B;
    rc = g->trajectories ? gmm_upload_cull(c, a, g->trajectories, B, T, mc->exhaustive != 0)
                         : gmm_upload_cull(c, a, mc->trajectory, 1, T, mc->exhaustive != 0);
    if (rc) return rc;
    if (N > (1ull << 31)) return fail(MCSIM_ERR_ARG, "gmm_run_batch: at most 2^31 samples per waypoint");
    // enough members to fill the GPU by themselves: one CTA per member for the whole run (gmm_member_run_kernel);
    // fewer: the per-waypoint kernels, several CTAs per member
    const size_t fused_smem = sizeof(float4) * size_t(c->sat2_n) + sizeof(unsigned) * size_t(c->gmm_cull_words);
    const bool fused = B >= 64 && fused_smem <= 64 * 1024 && g->kernel == MCSIM_GMM_KERNEL_AUTO && c->n_poly == 0;
    CK(cudaEventRecord(c->ev0, c->stream));
    if (fused) {
        SymbolLock sym;
        { const int rb = bind_symbols(c, nullptr, sym); if (rb) return rb; }
        void (*mk)(const GmmLaunch, int, int) = c->scene_axis_aligned ? gmm_member_run_kernel<true> : gmm_member_run_kernel<false>;
        if (fused_smem > 40 * 1024) CK(cudaFuncSetAttribute((const void*)mk, cudaFuncAttributeMaxDynamicSharedMemorySize, int(fused_smem)));
        mk<<<B, gmm_member_threads(B, c->n_sm), fused_smem, c->stream>>>(a, c->sat2_n, c->gmm_cull_words);
        ++c->launches;
        CK(cudaGetLastError());
    } else {
        rc = -1;
        if (gmm_cluster_fits(c, N, B, g->kernel)) {
            a.t = 0;
            rc = gmm_launch_cluster(c, a, B);
            if (rc > 0) return rc;
        }
        for (int t = 0; rc < 0 && t < T; ++t) {
            a.t = t;
            const int rw = gmm_launch_waypoint(c, a, MCSIM_MODE_PHILOX);
            if (rw) return rw;
        }
    }
    CK(cudaEventRecord(c->ev1, c->stream));
    std::vector<double> prob(size_t(B) * T);
    std::vector<int> deg(B);
    CK(cudaMemcpyAsync(prob.data(), c->gmm_prob.p, sizeof(double) * prob.size(), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(deg.data(), c->gmm_deg.p, sizeof(int) * B, cudaMemcpyDeviceToHost, c->stream));
    unsigned long long tests = 0;
    CK(cudaMemcpyAsync(&tests, a.tests, sizeof(tests), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    r->pair_tests = tests;
    c->d2h_bytes = sizeof(double) * prob.size() + sizeof(int) * B + sizeof(tests);
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    r->degenerate = 0;
    for (int b = 0; b < B; ++b) {
        double free_all = 1.0;  // P = 1 - prod_t (1 - p_t)  (MCSimulator.h:848-856)
        for (int t = 0; t < T; ++t) free_all *= (1.0 - prob[size_t(b) * T + t]);
        r->P[b] = 1.0 - free_all;
        r->degenerate += deg[b];
    }
    if (r->probabilities) memcpy(r->probabilities, prob.data(), sizeof(double) * prob.size());
    return MCSIM_OK;
}

// ---- EM fit that allocates nothing (mcsim_cuda_em_fit_static; no counterpart in the reference) ---------------------
// The scratch of the E-step / M-step kernels lives in this file-scope array instead of a context's cudaMalloc'ed
// buffers.  g_em_static_scratch is a HOST-side stand-in here: its device address is resolved once per device with
// cudaGetSymbolAddress (that device current) and only that pointer is handed to kernels.
static __device__ __align__(256) unsigned char g_em_static_scratch[MCSIM_EM_STATIC_SCRATCH_BYTES];

namespace em_static {
constexpr size_t al256(size_t v) { return (v + 255) & ~size_t(255); }
constexpr size_t OFF_STATS = 0;
constexpr size_t OFF_COMPS = al256(OFF_STATS + sizeof(double) * MCSIM_EM_STATIC_MAX_FITS * (MCSIM_MAX_GAUSSIANS * EM_NSTAT + 1));
constexpr size_t OFF_TRACE = al256(OFF_COMPS + sizeof(float) * MCSIM_EM_STATIC_MAX_FITS * MCSIM_MAX_GAUSSIANS * EM_COMP_STRIDE);
constexpr size_t OFF_OLD = al256(OFF_TRACE + sizeof(double) * MCSIM_EM_STATIC_MAX_FITS * MCSIM_EM_STATIC_MAX_ITERS);
constexpr size_t OFF_DONE = al256(OFF_OLD + sizeof(double) * MCSIM_EM_STATIC_MAX_FITS);
constexpr size_t OFF_ITERS = al256(OFF_DONE + sizeof(int) * MCSIM_EM_STATIC_MAX_FITS);
constexpr size_t END = al256(OFF_ITERS + sizeof(int) * MCSIM_EM_STATIC_MAX_FITS);
static_assert(END <= MCSIM_EM_STATIC_SCRATCH_BYTES, "mcsim_cuda_em_fit_static: scratch layout exceeds the stated maximum");

struct PerDevice { unsigned char* scratch = nullptr; int n_sm = 0; };
static PerDevice g_dev[64];
static std::mutex g_mu;
}  // namespace em_static

// The array keeps the previous call's contents: statistics, convergence state and the trace start from scratch here.
__global__ void __launch_bounds__(256) em_static_init_kernel(const EmLaunch a) {
    const int F = a.n_fits, nstat = F * (a.K * EM_NSTAT + 1), ntrace = F * a.iters;
    for (int i = threadIdx.x; i < nstat; i += blockDim.x) a.stats[i] = 0.0;
    for (int i = threadIdx.x; i < ntrace; i += blockDim.x) a.trace[i] = __longlong_as_double(0x7ff8000000000000ll);
    for (int i = threadIdx.x; i < F; i += blockDim.x) {
        a.old_avg[i] = __longlong_as_double(0xfff0000000000000ll);   // -inf (:2436)
        a.done[i] = 0;
        a.iters_run[i] = 0;
    }
}

extern "C" int mcsim_cuda_em_fit_static(const mcsim_em_static_params* p) {
    using namespace em_static;
    if (!p) return fail(MCSIM_ERR_ARG, "em_fit_static: NULL params");
    if (p->K < 1 || p->K > MCSIM_MAX_GAUSSIANS) return fail(MCSIM_ERR_ARG, "em_fit_static: K=%d out of range 1..%d", p->K, MCSIM_MAX_GAUSSIANS);
    if (p->n_fits < 1 || p->n_fits > MCSIM_EM_STATIC_MAX_FITS)
        return fail(MCSIM_ERR_ARG, "em_fit_static: n_fits=%d beyond the static scratch (1..%d fits; use mcsim_cuda_em_fit for more)", p->n_fits, MCSIM_EM_STATIC_MAX_FITS);
    if (p->iters < 0 || p->iters > MCSIM_EM_STATIC_MAX_ITERS)
        return fail(MCSIM_ERR_ARG, "em_fit_static: iters=%d beyond the static scratch (0..%d; use mcsim_cuda_em_fit for more)", p->iters, MCSIM_EM_STATIC_MAX_ITERS);
    if (p->n >= (1ull << 31) - (1ull << 24)) return fail(MCSIM_ERR_ARG, "em_fit_static: more than 2^31 samples per fit");
    if (!p->x || !p->hefts || !p->means || !p->fcovs) return fail(MCSIM_ERR_ARG, "em_fit_static: x / hefts / means / fcovs must be device pointers, got NULL");
    if (p->device < 0 || p->device >= 64) return fail(MCSIM_ERR_ARG, "em_fit_static: device %d out of range", p->device);
    CK(cudaSetDevice(p->device));
    PerDevice& pd = g_dev[p->device];
    {   // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lk(g_mu);
        if (!pd.scratch) {
            void* ptr = nullptr;
            CK(cudaGetSymbolAddress(&ptr, g_em_static_scratch));
            int n_sm = 0;
            CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, p->device));
            pd.n_sm = n_sm;
            pd.scratch = (unsigned char*)ptr;
        }
    }
    cudaStream_t st = (cudaStream_t)p->stream;
    const int K = p->K, F = p->n_fits;
    EmLaunch a;
    memset(&a, 0, sizeof(a));
    a.x = p->x; a.n = p->n; a.n_total = double(p->n);
    a.K = K; a.n_fits = F; a.iters = p->iters; a.iter = 0;
    a.var_floor = p->var_floor > 0 ? p->var_floor : DBL_MIN;
    a.hefts = p->hefts; a.means = p->means; a.fcovs = p->fcovs;
    a.stats = (double*)(pd.scratch + OFF_STATS);
    a.comps = (float*)(pd.scratch + OFF_COMPS);
    a.trace = p->avg_log_p ? p->avg_log_p : (double*)(pd.scratch + OFF_TRACE);
    a.old_avg = (double*)(pd.scratch + OFF_OLD);
    a.done = (int*)(pd.scratch + OFF_DONE);
    a.iters_run = p->iters_run ? p->iters_run : (int*)(pd.scratch + OFF_ITERS);
    em_static_init_kernel<<<1, 256, 0, st>>>(a);
    em_mstep_kernel<<<F, 32, 0, st>>>(a, 1);   // init_constants of the initial parameters
    CK(cudaGetLastError());
    if (p->iters == 0) return MCSIM_OK;
    // E-step kernel and grid: the choices of mcsim_cuda_em_estep (128-bit / bulk-copy path needs 16-byte aligned rows)
    em_kernel_t k0, k1;
    const bool vec = (p->n % 4) == 0 && p->n >= 4 && (reinterpret_cast<uintptr_t>(p->x) & 15) == 0;
    size_t smem = 0;
    int block = 256;
    if (!em_kernels(K, k0, k1, vec, smem, block)) return fail(MCSIM_ERR_ARG, "em_fit_static: bad K");
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute((const void*)k0, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)k0, block, smem));
    if (k1) { int o1 = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, (const void*)k1, block, 0)); if (o1 < occ) occ = o1; }
    if (occ < 1) occ = 1;
    const unsigned long long resident = (unsigned long long)pd.n_sm * occ, per_cta = smem ? EM_TMA_TILE : (unsigned long long)block * 8;
    unsigned long long most = (p->n + per_cta - 1) / per_cta;
    if (most < 1) most = 1;
    unsigned long long chunks = 1;
    double best = -1.0;
    for (unsigned long long w = 1; w <= 4; ++w) {
        unsigned long long ch = w * resident / (unsigned long long)F;
        if (ch < 1) ch = 1;
        if (ch > most) ch = most;
        const unsigned long long ctas = ch * F, waves = (ctas + resident - 1) / resident;
        const double eff = double(ctas) / double(waves * resident);
        if (eff > best + 0.02) { best = eff; chunks = ch; }
    }
    const dim3 grid((unsigned)chunks, (unsigned)F);
    for (int it = 0; it < p->iters; ++it) {
        k0<<<grid, block, smem, st>>>(a);
        if (k1) k1<<<grid, block, 0, st>>>(a);
        em_mstep_kernel<<<F, 32, 0, st>>>(a, 0);
        ++a.iter;
    }
    CK(cudaGetLastError());
    return MCSIM_OK;
}

#include "mcsim_group.cuh"
