// mcsim_cuda.cu -- the C ABI of include/mcsim_cuda.h over the sm_100a kernels.  One translation
// unit: the kernels live in the .cuh files included below and share the __constant__ tables.
// There is no CPU path: every entry point fails with MCSIM_ERR_NODEVICE without a CUDA device.
#include "mcsim_cuda.h"

#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "mc_rollout_sm100.cuh"
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

struct mcsim_cuda_ctx {
    int device = 0;
    int n_sm = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t launches = 0;
    bool scene_set = false;
    int n_pairs = 0;
    // MC state
    DevBuf plan, z0, counters, counts, poses, clearance;
    unsigned long long* h_counters = nullptr;  // pinned
    mcsim_mc_params mc;                        // prepared parameters (pointers not retained)
    bool mc_prepared = false;
    int mc_T = 0;
    // GMM / EM state
    DevBuf gmm_in, gmm_out, gmm_normals;
    double* h_gmm = nullptr;  // pinned staging
};

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
    c->gmm_in.release(); c->gmm_out.release(); c->gmm_normals.release();
    if (c->h_counters) cudaFreeHost(c->h_counters);
    if (c->h_gmm) cudaFreeHost(c->h_gmm);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" void* mcsim_cuda_stream(mcsim_cuda_ctx* c) { return c ? (void*)c->stream : nullptr; }
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
extern "C" int mcsim_cuda_set_scene(mcsim_cuda_ctx* c, const double* boxes, int n_boxes,
                                    const double* hulls, int n_hulls) {
    if (!c || n_boxes < 0 || n_hulls < 1 || n_hulls > MCSIM_MAX_HULLS || !hulls || (n_boxes && !boxes))
        return fail(MCSIM_ERR_ARG, "set_scene: bad arguments (n_boxes=%d n_hulls=%d)", n_boxes, n_hulls);
    CK(cudaSetDevice(c->device));
    SceneConst sc;
    memset(&sc, 0, sizeof(sc));
    std::vector<float4> pairs;
    sc.n_hulls = n_hulls;
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
            pairs.push_back(make_float4(float(bx[4]), float(bx[5]), float(p), float(q)));
        }
        r.count = int(pairs.size() / 2) - r.first;
        const float reach = float(sqrt(hu[0] * hu[0] + hu[1] * hu[1]) + sqrt(hu[4] * hu[4] + hu[5] * hu[5]));
        if (reach > sc.reach) sc.reach = reach;
    }
    sc.n_pairs = int(pairs.size() / 2);
    if (sc.n_pairs > MCSIM_MAX_PAIRS)
        return fail(MCSIM_ERR_ARG, "set_scene: %d live (hull, box) pairs exceed MCSIM_MAX_PAIRS=%d", sc.n_pairs, MCSIM_MAX_PAIRS);
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaMemcpyToSymbol(c_scene, &sc, sizeof(sc)));
    if (!pairs.empty()) CK(cudaMemcpyToSymbol(c_pairs, pairs.data(), pairs.size() * sizeof(float4)));
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
    CK(c->plan.reserve(plan.size() * sizeof(float)));
    CK(c->counters.reserve(sizeof(unsigned long long) * (4 + MCSIM_MAX_WAYPOINTS)));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaMemcpyToSymbolAsync(c_mc, &pc, sizeof(pc), 0, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->plan.p, plan.data(), plan.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * (4 + T), c->stream));
    CK(cudaStreamSynchronize(c->stream));  // plan is a stack vector
    c->mc = *p;
    c->mc.trajectory = c->mc.odometry = c->mc.landmarks = c->mc.applied = c->mc.z0 = nullptr;
    c->mc_prepared = true;
    c->mc_T = T;
    return MCSIM_OK;
}

static int mc_launch_impl(mcsim_cuda_ctx* c, uint64_t n, uint64_t offset, const double* d_z0, unsigned* d_counts,
                          float* d_poses, float* d_clear) {
    const mcsim_mc_params& p = c->mc;
    const bool audit = d_poses || d_clear;
    mc_kernel_t k = pick_kernel(p.mode, audit, p.early_exit != 0, p.estimator, p.n_landmarks);
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
    const int block = 256;
    const size_t smem = sizeof(float) * size_t(c->mc_T) * MC_PHILOX_STRIDE;
    unsigned long long want = (n + block - 1) / block;
    const unsigned long long cap = (unsigned long long)c->n_sm * 8;  // persistent: whole waves of resident CTAs
    const unsigned grid = unsigned(want < cap ? (want ? want : 1) : cap);
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

extern "C" int mcsim_cuda_mc_run(mcsim_cuda_ctx* c, const mcsim_mc_params* p, mcsim_mc_result* r) {
    if (!r) return fail(MCSIM_ERR_ARG, "mc_run: result is NULL");
    int rc = mcsim_cuda_mc_prepare(c, p);
    if (rc) return rc;
    const uint64_t n = p->n_rollouts;
    const int T = p->T;
    const double* d_z0 = nullptr;
    if (p->mode == MCSIM_MODE_REPLAY) {
        if (!p->z0 && n) return fail(MCSIM_ERR_ARG, "mc_run: replay mode needs z0");
        CK(c->z0.reserve(sizeof(double) * 3 * (n ? n : 1)));
        if (n) CK(cudaMemcpyAsync(c->z0.p, p->z0, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, c->stream));
        d_z0 = (const double*)c->z0.p;
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
    if (d_counts) CK(cudaMemcpyAsync(r->counts, d_counts, sizeof(unsigned) * n, cudaMemcpyDeviceToHost, c->stream));
    if (d_poses && r->poses) CK(cudaMemcpyAsync(r->poses, d_poses, sizeof(float) * 3 * T * n, cudaMemcpyDeviceToHost, c->stream));
    if (d_clear && r->clearance) CK(cudaMemcpyAsync(r->clearance, d_clear, sizeof(float) * T * n, cudaMemcpyDeviceToHost, c->stream));
    rc = mcsim_cuda_mc_collect(c, r);
    if (rc) return rc;
    CK(cudaEventElapsedTime(&r->kernel_ms, c->ev0, c->ev1));
    return MCSIM_OK;
}
