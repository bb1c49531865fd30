// mc_module_run.cpp -- the run* commands and the run-control additions of MCModule (see mc_module.h).
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "mc_module.h"
#include "collada_hulls.h"

namespace mcsim {
namespace {

void cuda_check(int rc, const char* what) {
    if (rc != MCSIM_OK) throw std::runtime_error(std::string(what) + ": " + mcsim_cuda_last_error());
}

template <typename T>
std::vector<T> read_binary(const std::string& path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) throw std::runtime_error("cannot open " + path);
    const std::streamsize bytes = f.tellg();
    if (bytes < 0 || bytes % std::streamsize(sizeof(T)) != 0) throw std::runtime_error(path + ": size is not a multiple of the element size");
    std::vector<T> v(size_t(bytes) / sizeof(T));
    f.seekg(0);
    if (bytes && !f.read(reinterpret_cast<char*>(v.data()), bytes)) throw std::runtime_error("short read on " + path);
    return v;
}

uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

}  // namespace

uint64_t MCModule::run_key(uint64_t run) const { return run ? seed_ ^ splitmix64(run) : seed_; }
uint64_t MCModule::gmm_key() const { return rollout_offset_ ? seed_ ^ splitmix64(~rollout_offset_) : seed_; }

void MCModule::fill_params(mcsim_mc_params& p) const {
    if (!ctx_) throw std::runtime_error("this module was created without a CUDA device (device < 0): run commands are unavailable, there is no CPU fallback");
    if (alphas_.size() < 4) throw std::runtime_error("setAlphas has not been sent");
    if (path_length_ < 1 || trajectory_.size() != size_t(3) * path_length_) throw std::runtime_error("setPathLength / setTrajectory have not been sent");
    if (odometry_.size() != size_t(3) * (path_length_ - 1)) throw std::runtime_error("setOdometry has not been sent (or was sent before a different setPathLength)");
    if (landmarks_.size() != size_t(2) * num_landmarks_) throw std::runtime_error("setNumLandmarks / setLandmarks have not been sent");
    if (!cov0_set_) throw std::runtime_error("setInitialCovariance has not been sent");
    if (scene_.n_hulls() + scene_.n_poly_hulls() < 1) throw std::runtime_error("no collision geometry: create the module with a scene, or send setScene / setGeometry");
    std::memset(&p, 0, sizeof(p));
    p.T = path_length_;
    p.trajectory = trajectory_.data();
    p.odometry = odometry_.data();
    for (int i = 0; i < 4; ++i) p.alphas[i] = alphas_[i];
    p.Q = Q_;
    p.n_landmarks = num_landmarks_;
    p.landmarks = landmarks_.data();
    std::memcpy(p.cov0, cov0_, sizeof(cov0_));
    p.estimator = estimator_;
    p.early_exit = early_exit_;
    p.exhaustive = exhaustive_;
    p.gmm_filter = gmm_filter_;
    p.seed = seed_;
    p.near_eps = 1e-4;
}

// mcsimplugin.cpp:75-81 -> MCSimulator::runSimulation (MCSimulator.h:361-365): returns P(collision) as text
bool MCModule::runSimulation(std::ostream& sout, std::istream&) {
    mcsim_mc_params p;
    fill_params(p);
    const int T = p.T, L = p.n_landmarks;
    const uint64_t N = uint64_t(num_particles_);
    p.n_rollouts = N;
    std::vector<double> applied_noisy(size_t(3) * (T > 1 ? T - 1 : 1));
    p.seed = run_key(run_index_);      // the run number is in the key, the shard offset in the counter (see run_key)
    p.rollout_offset = rollout_offset_;
    if (mode_ == Mode::kPhilox) {
        p.mode = MCSIM_MODE_PHILOX;
    } else {
        // one shared real-robot thread on the host (FP64), replayed by every rollout
        const size_t per_step = size_t(3 + L), need_host = per_step * size_t(T - 1);
        std::vector<double> hn;
        const double* host_normals = nullptr;
        if (mode_ == Mode::kDeterministic) {
            if (draws_.size() < 3 * N + need_host)
                throw std::runtime_error("runSimulation: deterministic mode needs " + std::to_string(3 * N + need_host) + " draws, " + std::to_string(draws_.size()) + " loaded");
            p.z0 = draws_.data();            // 3N particle normals first (initParticles), then 3 + L per step
            host_normals = draws_.data() + 3 * N;
        } else {
            hn.resize(need_host ? need_host : 1);
            if (T > 1) cuda_check(mcsim_host_normals(seed_, uint32_t(run_index_), T, int(per_step), hn.data()), "mcsim_host_normals");
            host_normals = hn.data();   // (the particles' start poses: per-rollout Philox blocks under this run's key)
        }
        if (T > 1) {
            const int rc = mcsim_host_thread(&p, host_normals, need_host, applied_noisy.data(), nullptr, nullptr, nullptr, nullptr, nullptr);
            if (rc) throw std::runtime_error("mcsim_host_thread failed (" + std::to_string(rc) + ")");
        }
        p.mode = MCSIM_MODE_REPLAY;
        p.applied = applied_noisy.data();
    }
    mcsim_mc_result r;
    std::memset(&r, 0, sizeof(r));
    first_hit_hist_.assign(size_t(T), 0);
    r.first_hit_hist = first_hit_hist_.data();
    cuda_check(mcsim_cuda_group_mc_run(group_, &p, &r), "mcsim_cuda_mc_run");   // all devices of the module; one device: the plain path
    ++run_index_;
    last_hits_ = r.hits; last_n_ = N; last_steps_ = r.executed_steps; last_tests_ = r.pair_tests; last_kernel_ms_ = r.kernel_ms;
    last_degenerate_ = 0;
    // getCollisionProportion (MCSimulator.h:324-330): nonzeros / (1.0 * N); N = 0 gives NaN there as well
    const double collprop = double(r.hits) / (1.0 * double(N));
    sout << collprop;  // default ostream formatting, 6 significant digits (mcsimplugin.cpp:79)
    return true;
}

// mcsimplugin.cpp:66-72 -> MCSimulator::runGMMEstimation (MCSimulator.h:354-358)
bool MCModule::runGMMEstimation(std::ostream& sout, std::istream&) {
    mcsim_mc_params p;
    fill_params(p);
    if (num_gaussians_ < 1) throw std::runtime_error("setNumGaussians has not been sent");
    if (num_gmm_samples_ < 1) throw std::runtime_error("setNumGMMSamples has not been sent");
    const int T = p.T, K = num_gaussians_;
    mcsim_gmm_run_params g;
    std::memset(&g, 0, sizeof(g));
    g.K = K;
    g.n_samples = uint64_t(num_gmm_samples_);
    g.refit = refit_;
    g.em_iters = em_iters_;
    if (mode_ == Mode::kDeterministic) {
        if (gmm_counts_.size() != size_t(T) * K) throw std::runtime_error("runGMMEstimation: deterministic mode needs T x K multinomial counts (setGMMCounts / setGMMCountsFile)");
        p.mode = MCSIM_MODE_REPLAY;
        g.normals = draws_.data();
        g.n_normals = draws_.size();
        g.counts = gmm_counts_.data();
    } else {
        p.mode = MCSIM_MODE_PHILOX;
        p.seed = gmm_key();
        p.rollout_offset = run_index_;  // Philox subsequence (stream id) of this run
    }
    probabilities_.assign(size_t(T), 0.0);
    gmm_post_.assign(size_t(K) * 13, 0.0);
    gmm_post_K_ = K;
    mcsim_gmm_run_result r;
    std::memset(&r, 0, sizeof(r));
    r.probabilities = probabilities_.data();
    r.gmm_post = gmm_post_.data();
    cuda_check(mcsim_cuda_gmm_run(ctx_, &p, &g, &r), "mcsim_cuda_gmm_run");
    ++run_index_;
    last_kernel_ms_ = r.kernel_ms;
    last_degenerate_ = r.degenerate;
    last_n_ = g.n_samples; last_hits_ = 0; last_steps_ = uint64_t(T) * g.n_samples; last_tests_ = 0;
    sout << r.P;
    return true;
}

// runGMMEstimationBatch <n>: the next n runs of runGMMEstimation (the driver's loop, MCSimulation.py:226-246) in the
// same launches; returns the n probabilities separated by spaces.  Run k of the batch is bit-for-bit the run that
// the (run_index + k)-th runGMMEstimation call would have been up to the summation order of the FP64 moments.
bool MCModule::runGMMEstimationBatch(std::ostream& sout, std::istream& sinput) {
    int n = 0;
    if (!(sinput >> n) || n < 1 || n > 65535) throw std::runtime_error("runGMMEstimationBatch: a run count 1..65535 expected");
    if (mode_ == Mode::kDeterministic) throw std::runtime_error("runGMMEstimationBatch: not available in deterministic mode (one injected draw stream = one run)");
    mcsim_mc_params p;
    fill_params(p);
    if (num_gaussians_ < 1) throw std::runtime_error("setNumGaussians has not been sent");
    if (num_gmm_samples_ < 1) throw std::runtime_error("setNumGMMSamples has not been sent");
    p.mode = MCSIM_MODE_PHILOX;
    p.seed = gmm_key();
    mcsim_gmm_batch_params g;
    std::memset(&g, 0, sizeof(g));
    g.K = num_gaussians_;
    g.n_samples = uint64_t(num_gmm_samples_);
    g.n_traj = n;
    g.stream0 = run_index_;
    std::vector<double> P(size_t(n), 0.0);
    mcsim_gmm_batch_result r;
    std::memset(&r, 0, sizeof(r));
    r.P = P.data();
    cuda_check(mcsim_cuda_group_gmm_run_batch(group_, &p, &g, &r), "mcsim_cuda_gmm_run_batch");   // members split over the module's devices
    run_index_ += uint64_t(n);
    last_kernel_ms_ = r.kernel_ms;
    last_degenerate_ = r.degenerate;
    last_n_ = g.n_samples * uint64_t(n); last_hits_ = 0; last_steps_ = uint64_t(p.T) * g.n_samples * uint64_t(n); last_tests_ = 0;
    for (int k = 0; k < n; ++k) sout << (k ? " " : "") << P[size_t(k)];
    return true;
}

// ---- additions ---------------------------------------------------------------------------------------------------
bool MCModule::setSeed(std::ostream&, std::istream& sinput) {
    unsigned long long s;
    if (!(sinput >> s)) throw std::runtime_error("setSeed: an unsigned integer expected");
    seed_ = s;
    run_index_ = 0;
    return true;
}

bool MCModule::setMode(std::ostream&, std::istream& sinput) {
    std::string m;
    sinput >> m;
    if (m == "shared") mode_ = Mode::kShared;
    else if (m == "philox") mode_ = Mode::kPhilox;
    else if (m == "deterministic") mode_ = Mode::kDeterministic;
    else throw std::runtime_error("setMode: shared | philox | deterministic expected, got '" + m + "'");
    return true;
}

bool MCModule::setDrawsFile(std::ostream&, std::istream& sinput) {
    std::string path;
    if (!(sinput >> path)) throw std::runtime_error("setDrawsFile: a path expected");
    draws_ = read_binary<double>(path);
    return true;
}

bool MCModule::setDraws(std::ostream&, std::istream& sinput) {
    size_t n;
    if (!(sinput >> n)) throw std::runtime_error("setDraws: a count expected");
    draws_.resize(n);
    for (size_t i = 0; i < n; ++i) if (!(sinput >> draws_[i])) throw std::runtime_error("setDraws: fewer numbers than announced");
    return true;
}

bool MCModule::setGMMCountsFile(std::ostream&, std::istream& sinput) {
    std::string path;
    if (!(sinput >> path)) throw std::runtime_error("setGMMCountsFile: a path expected");
    gmm_counts_ = read_binary<int>(path);
    return true;
}

bool MCModule::setGMMCounts(std::ostream&, std::istream& sinput) {
    size_t n;
    if (!(sinput >> n)) throw std::runtime_error("setGMMCounts: a count expected");
    gmm_counts_.resize(n);
    for (size_t i = 0; i < n; ++i) if (!(sinput >> gmm_counts_[i])) throw std::runtime_error("setGMMCounts: fewer numbers than announced");
    return true;
}

bool MCModule::setEarlyExit(std::ostream&, std::istream& sinput) {
    int v;
    if (!(sinput >> v)) throw std::runtime_error("setEarlyExit: 0 or 1 expected");
    early_exit_ = v != 0;
    return true;
}

bool MCModule::setBroadPhase(std::ostream&, std::istream& sinput) {
    int v;
    if (!(sinput >> v)) throw std::runtime_error("setBroadPhase: 0 or 1 expected");
    exhaustive_ = v == 0;
    return true;
}

bool MCModule::setEstimator(std::ostream&, std::istream& sinput) {
    std::string m;
    sinput >> m;
    if (m == "reference") estimator_ = MCSIM_EST_REFERENCE;
    else if (m == "closedloop") estimator_ = MCSIM_EST_CLOSEDLOOP;
    else throw std::runtime_error("setEstimator: reference | closedloop expected");
    return true;
}

bool MCModule::setRolloutOffset(std::ostream&, std::istream& sinput) {
    unsigned long long v;
    if (!(sinput >> v)) throw std::runtime_error("setRolloutOffset: an unsigned integer expected");
    rollout_offset_ = v;
    return true;
}

bool MCModule::setRefit(std::ostream&, std::istream& sinput) {
    std::string m;
    sinput >> m;
    if (m == "moments") { refit_ = MCSIM_GMM_REFIT_MOMENTS; em_iters_ = 0; return true; }
    if (m != "em") throw std::runtime_error("setRefit: moments | em [iterations] expected, got '" + m + "'");
    int it = 5;
    if (sinput >> it) { if (it < 1 || it > 1000) throw std::runtime_error("setRefit: 1..1000 EM iterations expected"); }
    refit_ = MCSIM_GMM_REFIT_EM;
    em_iters_ = it;
    return true;
}

// setFilter state | deviation: how runGMMEstimation / runLinearEstimation move a component's belief to the next waypoint --
// the reference's EKF on the state (MCSimulator.h:868-929), or a Kalman filter on the deviation from the nominal plan,
// linearised about the plan (the paper's future work, finalpaper/ajaay_paper.tex:893, 901; mcsim_mc_params.gmm_filter)
bool MCModule::setFilter(std::ostream&, std::istream& sinput) {
    std::string m;
    sinput >> m;
    if (m == "state") gmm_filter_ = MCSIM_FILTER_STATE;
    else if (m == "deviation") gmm_filter_ = MCSIM_FILTER_DEVIATION;
    else throw std::runtime_error("setFilter: state | deviation expected, got '" + m + "'");
    return true;
}

// setConstraints <n> then n rows (ax ay b xmin xmax ymin ymax): linConstraint(A, b, xmin, xmax, ymin, ymax) of
// linConstraint.py:4-11.  A state passes a constraint iff A . xy <= b AND it lies inside the window (:15-30), and collides
// when any constraint fails (gaussprop.py:63-73): the free region is an intersection of half-planes, kept as such.
bool MCModule::setConstraints(std::ostream&, std::istream& sinput) {
    int n = -1;
    if (!(sinput >> n) || n < 0 || n > 512) throw std::runtime_error("setConstraints: a constraint count 0..512 expected");
    auto number = [&](double& v) {
        std::string tok;
        if (!(sinput >> tok)) return false;
        if (tok == "inf" || tok == "+inf") { v = HUGE_VAL; return true; }
        if (tok == "-inf") { v = -HUGE_VAL; return true; }
        try { size_t used = 0; v = std::stod(tok, &used); return used == tok.size(); } catch (...) { return false; }
    };
    std::vector<double> A, b;
    auto half_plane = [&](double ax, double ay, double bb) { A.push_back(ax); A.push_back(ay); A.push_back(0.0); b.push_back(bb); };
    for (int k = 0; k < n; ++k) {
        double v[7];
        for (double& x : v) if (!number(x)) throw std::runtime_error("setConstraints: 7 numbers per constraint expected (ax ay b xmin xmax ymin ymax)");
        if (v[0] != 0.0 || v[1] != 0.0) half_plane(v[0], v[1], v[2]);
        if (std::isfinite(v[3])) half_plane(-1.0, 0.0, -v[3]);
        if (std::isfinite(v[4])) half_plane(1.0, 0.0, v[4]);
        if (std::isfinite(v[5])) half_plane(0.0, -1.0, -v[5]);
        if (std::isfinite(v[6])) half_plane(0.0, 1.0, v[6]);
    }
    con_A_.swap(A);
    con_b_.swap(b);
    return true;
}

// the GMM branch with the analytic truncation (mcsim_cuda_linear_run); deterministic mode takes the real-robot thread's
// draws from the draws file ((T - 1) x (3 + L) normals), the other modes from the module's Philox stream
bool MCModule::runLinearEstimation(std::ostream& sout, std::istream&) {
    mcsim_mc_params p;
    fill_params(p);
    if (num_gaussians_ < 1) throw std::runtime_error("setNumGaussians has not been sent");
    const int T = p.T, K = num_gaussians_;
    mcsim_linear_params lp;
    std::memset(&lp, 0, sizeof(lp));
    lp.K = K;
    lp.n_constraints = int(con_b_.size());
    lp.A = con_A_.data();
    lp.b = con_b_.data();
    if (mode_ == Mode::kDeterministic) { lp.normals = draws_.data(); lp.n_normals = draws_.size(); }
    p.mode = MCSIM_MODE_PHILOX;
    p.seed = gmm_key();
    p.rollout_offset = run_index_;
    probabilities_.assign(size_t(T), 0.0);
    gmm_post_.assign(size_t(K) * 13, 0.0);
    gmm_post_K_ = K;
    mcsim_gmm_run_result r;
    std::memset(&r, 0, sizeof(r));
    r.probabilities = probabilities_.data();
    r.gmm_post = gmm_post_.data();
    cuda_check(mcsim_cuda_linear_run(ctx_, &p, &lp, &r), "mcsim_cuda_linear_run");
    ++run_index_;
    last_kernel_ms_ = r.kernel_ms;
    last_degenerate_ = r.degenerate;
    last_n_ = 0; last_hits_ = 0; last_steps_ = 0; last_tests_ = 0;
    sout << r.P;
    return true;
}

// setDevices all | <ordinal> [<ordinal> ...] (commas are accepted as separators)
bool MCModule::setDevices(std::ostream&, std::istream& sinput) {
    std::string rest, tok;
    std::getline(sinput, rest);
    for (auto& ch : rest) if (ch == ',') ch = ' ';
    std::istringstream in(rest);
    std::vector<int> devs;
    while (in >> tok) {
        if (tok == "all") {
            const int n = mcsim_cuda_device_count();
            if (n < 1) throw std::runtime_error("setDevices: no CUDA device (there is no CPU path)");
            devs.clear();
            for (int d = 0; d < n; ++d) devs.push_back(d);
            break;
        }
        size_t used = 0;
        int v = -1;
        try { v = std::stoi(tok, &used); } catch (...) { used = 0; }
        if (used != tok.size() || v < 0) throw std::runtime_error("setDevices: device ordinals or 'all' expected, got '" + tok + "'");
        devs.push_back(v);
    }
    if (devs.empty()) throw std::runtime_error("setDevices: at least one device expected");
    open_devices(devs);
    return true;
}

bool MCModule::getDevices(std::ostream& sout, std::istream&) {
    const int n = mcsim_cuda_group_size(group_);
    for (int i = 0; i < n; ++i) sout << (i ? " " : "") << mcsim_cuda_ctx_device(mcsim_cuda_group_ctx(group_, i));
    return true;
}

bool MCModule::describeRun(std::ostream& sout, std::istream& sinput) {
    unsigned long long k = 0;
    sinput >> k;
    const uint64_t run = run_index_ + k;
    sout << "mc " << run_key(run) << " " << rollout_offset_ << " " << num_particles_ << " gmm " << gmm_key() << " " << run;
    return true;
}

bool MCModule::setGeometry(std::ostream&, std::istream& sinput) {
    int nb, nh;
    if (!(sinput >> nb >> nh) || nb < 0 || nh < 1) throw std::runtime_error("setGeometry: <n_boxes> <n_hulls> expected");
    Scene sc;
    sc.boxes.resize(size_t(nb) * 8);
    sc.hulls.resize(size_t(nh) * 8);
    for (auto& v : sc.boxes) if (!(sinput >> v)) throw std::runtime_error("setGeometry: too few numbers");
    for (auto& v : sc.hulls) if (!(sinput >> v)) throw std::runtime_error("setGeometry: too few numbers");
    scene_ = sc;
    upload_scene();
    return true;
}

bool MCModule::setRobotPolygons(std::ostream&, std::istream& sinput) {
    int n = -1;
    if (!(sinput >> n) || n < 0 || n > MCSIM_MAX_POLY_HULLS) throw std::runtime_error("setRobotPolygons: a hull count 0.." + std::to_string(MCSIM_MAX_POLY_HULLS) + " expected");
    std::vector<double> rows;
    for (int h = 0; h < n; ++h) {
        double zlo, zhi;
        int m;
        if (!(sinput >> zlo >> zhi >> m) || m < 3 || m > 8 || !(zlo < zhi)) throw std::runtime_error("setRobotPolygons: zlo zhi m (3..8) expected for every hull");
        double r[19] = {double(m), zlo, zhi};
        for (int i = 0; i < 2 * m; ++i) if (!(sinput >> r[3 + i])) throw std::runtime_error("setRobotPolygons: too few vertex coordinates");
        rows.insert(rows.end(), r, r + 19);
    }
    scene_.poly_hulls.swap(rows);
    upload_scene();
    return true;
}

// setRobotModel <path.zae | path.dae> [boxes | polygons] [z0 z1 ... zk]: the robot's hull set from a COLLADA model (what the reference's scenes
// name as their robot, pr2test2.env.xml:120) -- one convex footprint per z-band, replacing box and polygon hulls alike.
// Band boundaries are in the model frame (k <= MCSIM_MAX_POLY_HULLS bands; none given: that many equal bands).
bool MCModule::setRobotModel(std::ostream& sout, std::istream& sinput) {
    std::string path;
    if (!(sinput >> path)) throw std::runtime_error("setRobotModel: a .zae / .dae path expected");
    ColladaOptions opt;
    std::string tok;
    while (sinput >> tok) {
        if (tok == "boxes") { opt.as_boxes = true; continue; }       // minimum-area bounding rectangles: box hulls, packed kernels
        if (tok == "polygons") { opt.as_boxes = false; continue; }   // the footprints themselves (default)
        try { size_t used = 0; opt.z_edges.push_back(std::stod(tok, &used)); if (used != tok.size()) throw std::invalid_argument(tok); }
        catch (const std::exception&) { throw std::runtime_error("setRobotModel: '" + tok + "' is neither boxes | polygons nor a band boundary"); }
    }
    const int max_bands = opt.as_boxes ? MCSIM_MAX_HULLS : MCSIM_MAX_POLY_HULLS;
    opt.max_bands = opt.as_boxes ? 4 : MCSIM_MAX_POLY_HULLS;   // default split when no boundaries are given
    if (opt.z_edges.size() == 1) throw std::runtime_error("setRobotModel: one band boundary is not a band");
    if (opt.z_edges.size() > size_t(max_bands) + 1) throw std::runtime_error("setRobotModel: at most " + std::to_string(max_bands) + " bands");
    std::vector<double> rows = load_collada_hulls(path, opt);
    if (opt.as_boxes) { scene_.poly_hulls.clear(); scene_.hulls.swap(rows); }
    else { scene_.hulls.clear(); scene_.poly_hulls.swap(rows); }
    upload_scene();
    sout << (opt.as_boxes ? scene_.n_hulls() : scene_.n_poly_hulls());
    return true;
}

bool MCModule::setScene(std::ostream&, std::istream& sinput) {
    std::string path, d;
    if (!(sinput >> path)) throw std::runtime_error("setScene: a path expected");
    std::vector<std::string> dirs;
    while (sinput >> d) dirs.push_back(d);
    scene_ = load_scene_xml(path, dirs);
    if (scene_.n_hulls() + scene_.n_poly_hulls() < 1) throw std::runtime_error("setScene: " + path + " has no robot geometry");
    upload_scene();
    return true;
}

bool MCModule::getWaypointProbabilities(std::ostream& sout, std::istream&) {
    sout.precision(17);
    for (size_t i = 0; i < probabilities_.size(); ++i) sout << (i ? " " : "") << probabilities_[i];
    return true;
}

bool MCModule::getFirstHitHistogram(std::ostream& sout, std::istream&) {
    for (size_t i = 0; i < first_hit_hist_.size(); ++i) sout << (i ? " " : "") << first_hit_hist_[i];
    return true;
}

bool MCModule::getGMM(std::ostream& sout, std::istream&) {
    sout.precision(17);
    for (int g = 0; g < gmm_post_K_; ++g) {
        for (int k = 0; k < 13; ++k) sout << (k ? " " : "") << gmm_post_[size_t(g) * 13 + k];
        sout << "\n";
    }
    return true;
}

bool MCModule::getStats(std::ostream& sout, std::istream&) {
    sout << last_hits_ << " " << last_n_ << " " << last_steps_ << " " << last_tests_ << " " << last_kernel_ms_ << " " << last_degenerate_;
    return true;
}

}  // namespace mcsim
