// TEST INFRASTRUCTURE ONLY (oracle/): the reference estimator itself, compiled from where it lies.
//
// This translation unit #includes the UNMODIFIED reference headers
//   /root/reference/mcsimplugin/MCSimulator.h   (class MCSimulator, :93-930)
//   /root/reference/mcsimplugin/GM_Model.h      (class GM_Model, :34-126)
//   /root/reference/armadillo-8.400.0/include   (mvnrnd, cov, mean, normalise, gmm_full)
// against oracle/stub/openrave/plugin.h, and exposes them through a small C ABI for ctypes.
// Nothing of the reference is copied into this repository; the build recipe is oracle/Makefile and
// the output goes to oracle/_ref/ (git-ignored).  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the resulting library.
//
// What is NOT the reference's: EnvironmentBase::CheckCollision (served by oracle/sat_cpu.h because
// ODE and the PR2 mesh are not in the repo) and the random draws (injected through
// Armadillo's ARMA_RNG_ALT hook, oracle/ref/arma_rng_inject.hpp).

#include <armadillo>
#include <openrave/plugin.h>
#include <unordered_set>
#include <cassert>
#include <string>
#include <vector>
#include <chrono>
#include <random>
#include <sstream>
#include <iostream>
#include <cstring>
#include <cstdint>

// Reach the estimator's private state (mcparticles, particlecollisions, mu, cov, gmm, and
// GM_Model::generator/weights) for golden dumps.  The two headers define no templates of their own.
#define private public
#define class struct
#include "MCSimulator.h"
#undef class
#undef private

#include "../sat_cpu.h"

extern "C" {

typedef struct {
    double alphas[4];       // squared alphas (gaussprop.py:36)
    double Q;               // range-sensor variance
    int n_landmarks;
    const double* landmarks;  // x_0..x_{L-1} y_0..y_{L-1}   (setLandmarks grammar, mcsimplugin.cpp:148-166)
    int n_particles;
    double cov0[9];         // row-major (mcsimplugin.cpp:121-134)
    int path_length;        // T
    const double* trajectory;  // x_0..x_{T-1} y_0.. th_0..    (mcsimplugin.cpp:83-97)
    const double* odometry;    // r1_0..r1_{T-2} tr_0.. r2_0.. (mcsimplugin.cpp:99-113)
    int n_gaussians;
    int n_gmm_samples;
    const double* boxes;    // n_boxes x 8
    int n_boxes;
    const double* hulls;    // n_hulls x 8
    int n_hulls;
    const double* poly_hulls;  // n_poly_hulls x 19 (sat_cpu.h), may be NULL
    int n_poly_hulls;
} ref_params;

typedef struct {
    // all optional (may be NULL)
    double* poses;         // [calls x 3]: pose handed to CheckCollision, in call order
    double* clearance;     // [calls]
    unsigned char* flags;  // [calls]
    size_t cap_calls;      // capacity of the three arrays above, in calls
    double* mu_path;       // [T x 3]  EKF mean at waypoint t (after the update of step t-1)
    double* cov_path;      // [T x 9]  EKF covariance, row-major
    double* gmm_pre;       // [T x K x 13]: w, mu[3], cov[9 row-major] of every component right
                           // before truncateGMM at waypoint t (GMM mode)
    int* gmm_counts;       // [T x K] multinomial counts drawn by GM_Model::sampleNPoints
    double* gmm_post;      // [K x 13] mixture after the last truncateGMM
} ref_record;

}  // extern "C"

namespace {

struct Ctx {
    sat_scene scene;
    ref_record rec;
    size_t calls;
    size_t group;  // calls per waypoint (N particles or N gmm samples)
    MCSimulator* sim;
    int T, K;
    bool gmm_mode;
    std::vector<size_t> hits;  // colliding calls per waypoint group
};

bool check_hook(void* vctx, const double pose[3]) {
    Ctx* c = static_cast<Ctx*>(vctx);
    const size_t i = c->calls++;
    if (c->sim && c->group && (i % c->group) == 0) {
        const size_t t = i / c->group;
        if (t < size_t(c->T)) {
            if (c->rec.mu_path) for (int k = 0; k < 3; ++k) c->rec.mu_path[t * 3 + k] = c->sim->mu(k, 0);
            if (c->rec.cov_path)
                for (int r = 0; r < 3; ++r)
                    for (int q = 0; q < 3; ++q) c->rec.cov_path[t * 9 + r * 3 + q] = c->sim->cov(r, q);
            if (c->gmm_mode && c->rec.gmm_pre) {
                for (int g = 0; g < c->K; ++g) {
                    double* o = c->rec.gmm_pre + (t * c->K + g) * 13;
                    o[0] = c->sim->gmm.weights[g];
                    for (int k = 0; k < 3; ++k) o[1 + k] = c->sim->gmm.means[g](k, 0);
                    for (int r = 0; r < 3; ++r)
                        for (int q = 0; q < 3; ++q) o[4 + r * 3 + q] = c->sim->gmm.covariances[g](r, q);
                }
            }
        }
    }
    const double cl = sat_clearance(&c->scene, pose);
    if (c->group && (i / c->group) < c->hits.size() && cl <= 0.0) ++c->hits[i / c->group];
    if (i < c->rec.cap_calls) {
        if (c->rec.poses) { c->rec.poses[i * 3] = pose[0]; c->rec.poses[i * 3 + 1] = pose[1]; c->rec.poses[i * 3 + 2] = pose[2]; }
        if (c->rec.clearance) c->rec.clearance[i] = cl;
        if (c->rec.flags) c->rec.flags[i] = (cl <= 0.0);
    }
    return cl <= 0.0;
}

// Push the parameters exactly as mcsimplugin.cpp's handlers do (:49-187).
void configure(MCSimulator& sim, const ref_params* p) {
    std::vector<double> a(p->alphas, p->alphas + 4);
    sim.setAlphas(a);
    sim.setQ(p->Q);
    sim.setNumLandmarks(p->n_landmarks);
    arma::Mat<double> lm = arma::zeros<arma::Mat<double> >(2, p->n_landmarks);
    for (int i = 0; i < p->n_landmarks; ++i) { lm(0, i) = p->landmarks[i]; lm(1, i) = p->landmarks[p->n_landmarks + i]; }
    sim.setLandmarks(lm);
    sim.setNumParticles(p->n_particles);
    arma::Mat<double> c0 = arma::zeros<arma::Mat<double> >(3, 3);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) c0(i, j) = p->cov0[i * 3 + j];
    sim.setInitialCovariance(c0);
    sim.setPathLength(p->path_length);
    const int T = p->path_length;
    arma::Mat<double> tr = arma::zeros<arma::Mat<double> >(3, T);
    for (int c = 0; c < 3; ++c) for (int i = 0; i < T; ++i) tr(c, i) = p->trajectory[c * T + i];
    sim.setTrajectory(tr);
    arma::Mat<double> od = arma::zeros<arma::Mat<double> >(3, T - 1);
    for (int c = 0; c < 3; ++c) for (int i = 0; i < T - 1; ++i) od(c, i) = p->odometry[c * (T - 1) + i];
    sim.setOdometry(od);
    sim.setNumGaussians(p->n_gaussians);
    sim.setNumGMMSamples(p->n_gmm_samples);
}

struct CoutGuard {
    std::streambuf* old;
    std::ostringstream sink;
    bool capture;
    std::ios_base::iostate old_state;
    explicit CoutGuard(bool capture_) : capture(capture_) {
        old_state = std::cout.rdstate();
        old = std::cout.rdbuf();
        if (capture) std::cout.rdbuf(sink.rdbuf());
        else std::cout.setstate(std::ios_base::failbit);  // inserters become no-ops ("stdout silenced")
    }
    ~CoutGuard() {
        std::cout.rdbuf(old);
        std::cout.clear(old_state);
    }
};

void seed_rng(const double* draws, size_t n_draws, uint64_t mt_seed) {
    oracle_rng_state& r = oracle_rng();
    r.draws = draws;
    r.n_draws = n_draws;
    r.pos = 0;
    r.inject = (draws != nullptr);
    r.underflow = false;
    r.engine.seed(mt_seed);
    r.normal.reset();
}

void parse_counts(const std::string& text, int K, int T, int* out) {
    // GM_Model::sampleNPoints prints "Counts Vector\n" then the K counts (GM_Model.h:95-96,26-31).
    size_t pos = 0;
    int t = 0;
    const std::string key = "Counts Vector\n";
    while (t < T && (pos = text.find(key, pos)) != std::string::npos) {
        pos += key.size();
        std::istringstream is(text.substr(pos, 64 * size_t(K)));
        for (int g = 0; g < K; ++g) { int v = 0; is >> v; out[t * K + g] = v; }
        ++t;
    }
}

}  // namespace

extern "C" {

// Runs the reference's own MCSimulator::runSimulation() (MCSimulator.h:361-365 -> :649-864).
// draws != NULL: injected standard normals (3N + 11(T-1) are consumed); else mt19937_64(mt_seed).
// counts_out[N]: per-particle number of colliding waypoints (particlecollisions, :108,344).
// final_particles[3N]: mcparticles after the last step, particle-major.
// Returns 0 on success, 1 if the injected draws ran out, 2 on an Armadillo exception.
int ref_mc_run(const ref_params* p, const double* draws, size_t n_draws, uint64_t mt_seed,
               ref_record* rec, double* P_out, unsigned* counts_out, double* final_particles) {
    try {
        OpenRAVE::EnvironmentBasePtr env(new OpenRAVE::EnvironmentBase());
        Ctx ctx;
        std::memset(&ctx.rec, 0, sizeof(ctx.rec));
        if (rec) ctx.rec = *rec;
        ctx.scene.boxes = p->boxes; ctx.scene.n_boxes = p->n_boxes;
        ctx.scene.hulls = p->hulls; ctx.scene.n_hulls = p->n_hulls;
        ctx.scene.poly_hulls = p->poly_hulls; ctx.scene.n_poly_hulls = p->n_poly_hulls;
        ctx.calls = 0; ctx.group = size_t(p->n_particles); ctx.T = p->path_length; ctx.K = 0; ctx.gmm_mode = false;
        ctx.sim = nullptr;
        env->check = check_hook; env->check_ctx = &ctx;
        CoutGuard guard(false);
        MCSimulator sim(env);
        ctx.sim = &sim;
        configure(sim, p);
        seed_rng(draws, n_draws, mt_seed);
        const double P = sim.runSimulation();
        if (P_out) *P_out = P;
        if (counts_out) for (int i = 0; i < p->n_particles; ++i) counts_out[i] = sim.particlecollisions(0, i);
        if (final_particles)
            for (int i = 0; i < p->n_particles; ++i)
                for (int k = 0; k < 3; ++k) final_particles[i * 3 + k] = sim.mcparticles(k, i);
        // EKF state at the last waypoint (the hook only sees mu_{t-1} when waypoint t is checked)
        if (ctx.rec.mu_path) for (int k = 0; k < 3; ++k) ctx.rec.mu_path[(p->path_length - 1) * 3 + k] = sim.mu(k, 0);
        if (ctx.rec.cov_path)
            for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q)
                ctx.rec.cov_path[(p->path_length - 1) * 9 + r * 3 + q] = sim.cov(r, q);
        return oracle_rng().underflow ? 1 : 0;
    } catch (const std::exception& e) {
        std::cerr << "ref_mc_run: " << e.what() << std::endl;
        return 2;
    }
}

// Runs the reference's own MCSimulator::runGMMEstimation() (MCSimulator.h:354-358 -> :559 -> :649-864).
// Normal draws consumed: 3N (truncateGMM at waypoint 0) + per step 3 + 8 + 3N.
// multinomial_seed reseeds GM_Model::generator (GM_Model.h:43,52-55; the reference seeds it from the clock).
// prob_out[T]: per-waypoint colliding fraction (probabilities, MCSimulator.h:660,678,817).
int ref_gmm_run(const ref_params* p, const double* draws, size_t n_draws, uint64_t mt_seed,
                unsigned multinomial_seed, ref_record* rec, double* P_out, double* prob_out) {
    try {
        OpenRAVE::EnvironmentBasePtr env(new OpenRAVE::EnvironmentBase());
        Ctx ctx;
        std::memset(&ctx.rec, 0, sizeof(ctx.rec));
        if (rec) ctx.rec = *rec;
        ctx.scene.boxes = p->boxes; ctx.scene.n_boxes = p->n_boxes;
        ctx.scene.hulls = p->hulls; ctx.scene.n_hulls = p->n_hulls;
        ctx.scene.poly_hulls = p->poly_hulls; ctx.scene.n_poly_hulls = p->n_poly_hulls;
        ctx.calls = 0; ctx.group = size_t(p->n_gmm_samples); ctx.T = p->path_length; ctx.K = p->n_gaussians; ctx.gmm_mode = true;
        ctx.sim = nullptr;
        env->check = check_hook; env->check_ctx = &ctx;
        const bool want_counts = rec && rec->gmm_counts;
        ctx.hits.assign(size_t(p->path_length), 0);
        std::string captured;
        {
            CoutGuard guard(want_counts);
            MCSimulator sim(env);
            ctx.sim = &sim;
            configure(sim, p);
            sim.gmm.generator.seed(multinomial_seed);
            seed_rng(draws, n_draws, mt_seed);
            const double P = sim.runGMMEstimation();
            if (P_out) *P_out = P;
            if (ctx.rec.gmm_post) {
                for (int g = 0; g < p->n_gaussians; ++g) {
                    double* o = ctx.rec.gmm_post + g * 13;
                    o[0] = sim.gmm.weights[g];
                    for (int k = 0; k < 3; ++k) o[1 + k] = sim.gmm.means[g](k, 0);
                    for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) o[4 + r * 3 + q] = sim.gmm.covariances[g](r, q);
                }
            }
            if (ctx.rec.mu_path) for (int k = 0; k < 3; ++k) ctx.rec.mu_path[(p->path_length - 1) * 3 + k] = sim.mu(k, 0);
            if (want_counts) captured = guard.sink.str();
        }
        if (want_counts) parse_counts(captured, p->n_gaussians, p->path_length, rec->gmm_counts);
        if (prob_out)
            for (int t = 0; t < p->path_length; ++t) prob_out[t] = double(ctx.hits[t]) / (1.0 * p->n_gmm_samples);
        return oracle_rng().underflow ? 1 : 0;
    } catch (const std::exception& e) {
        std::cerr << "ref_gmm_run: " << e.what() << std::endl;
        return 2;
    }
}

// The vendored EM the reference ships but never calls:
// arma::gmm_full::learn(X, K, maha_dist, keep_existing, km_iter = 0, em_iter, var_floor, false)
// (armadillo_bits/gmm_full_meat.hpp:764-890 -> em_iterate :2390-2484).
// X: 3 x N column-major.  means: K x 3, covs: K x 9 (row-major == column-major, symmetric), hefts: K.
// In/out.  avg_log_p_out: gmm_full::avg_log_p(X) of the fitted model.
int ref_em_fit(const double* X, size_t N, int K, double* means, double* covs, double* hefts,
               int em_iter, double var_floor, double* avg_log_p_out) {
    try {
        arma::mat data(const_cast<double*>(X), 3, N, false, true);
        arma::mat M(3, K);
        arma::cube C(3, 3, K);
        arma::rowvec H(K);
        for (int g = 0; g < K; ++g) {
            for (int k = 0; k < 3; ++k) M(k, g) = means[g * 3 + k];
            for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) C(r, q, g) = covs[g * 9 + r * 3 + q];
            H(g) = hefts[g];
        }
        arma::gmm_full model;
        model.set_params(M, C, H);
        const bool ok = model.learn(data, arma::uword(K), arma::maha_dist, arma::keep_existing, 0,
                                    arma::uword(em_iter), var_floor, false);
        for (int g = 0; g < K; ++g) {
            for (int k = 0; k < 3; ++k) means[g * 3 + k] = model.means(k, g);
            for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) covs[g * 9 + r * 3 + q] = model.fcovs(r, q, g);
            hefts[g] = model.hefts(g);
        }
        if (avg_log_p_out) *avg_log_p_out = model.avg_log_p(data);
        return ok ? 0 : 3;
    } catch (const std::exception& e) {
        std::cerr << "ref_em_fit: " << e.what() << std::endl;
        return 2;
    }
}

// Armadillo pieces the estimator leans on, exposed for known-answer checks against
// armadillo-8.400.0/tests/fn_cov.cpp:23-60.
void ref_cov(const double* A, int n_rows, int n_cols, double* out /* n_cols x n_cols, col-major */) {
    arma::mat a(const_cast<double*>(A), n_rows, n_cols, false, true);
    arma::mat c = arma::cov(a);
    std::memcpy(out, c.memptr(), sizeof(double) * c.n_elem);
}

double ref_clearance(const ref_params* p, const double pose[3]) {
    sat_scene sc;
    sc.boxes = p->boxes; sc.n_boxes = p->n_boxes; sc.hulls = p->hulls; sc.n_hulls = p->n_hulls;
    sc.poly_hulls = p->poly_hulls; sc.n_poly_hulls = p->n_poly_hulls;
    return sat_clearance(&sc, pose);
}

}  // extern "C"
