// mc_module.cpp -- command handlers of the plugin (see mc_module.h).  Each handler cites the reference handler
// whose grammar and returned text it keeps (mcsimplugin/mcsimplugin.cpp).  The reference's handlers that fall off
// the end of a bool function (:83-172, undefined behaviour) return true here, which is what MCSimulation.py needs.
#include "mc_module.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <random>
#include <sstream>
#include <stdexcept>

// NVTX ranges, one per command (header-only NVTX 3: no-ops unless a profiler is attached)
#if defined(__has_include)
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
#define MCSIM_NVTX 1
#endif
#endif

namespace mcsim {
namespace {

struct NvtxRange {
#ifdef MCSIM_NVTX
    explicit NvtxRange(const std::string& name) { nvtxRangePushA(("mcsim:" + name).c_str()); }
    ~NvtxRange() { nvtxRangePop(); }
#else
    explicit NvtxRange(const std::string&) {}
#endif
};

std::string lower(std::string s) {
    for (auto& ch : s) ch = char(std::tolower(static_cast<unsigned char>(ch)));
    return s;
}

void cuda_check(int rc, const char* what) {
    if (rc != MCSIM_OK) throw std::runtime_error(std::string(what) + ": " + mcsim_cuda_last_error());
}

// `count` numbers from the stream; the reference's `sinput >> val` leaves val unchanged on failure, so a short
// command silently repeats the last value there.  Here a short command is an error.
void read_doubles(std::istream& in, size_t count, std::vector<double>& out, const char* what) {
    out.resize(count);
    for (size_t i = 0; i < count; ++i)
        if (!(in >> out[i])) throw std::runtime_error(std::string(what) + ": expected " + std::to_string(count) + " numbers, got " + std::to_string(i));
}

}  // namespace

const MCModule::Command MCModule::kCommands[] = {
    // names and help strings of mcsimplugin.cpp:13-44
    {"MyCommand", &MCModule::MyCommand, "This is an example command"},
    {"ArmaCommand", &MCModule::ArmaCommand, "This is testing armadillo"},
    {"setAlphas", &MCModule::setAlphas, "This is to initialize alphas from python"},
    {"setQ", &MCModule::setQ, "This is to initialize variance of sensor noise"},
    {"setNumLandmarks", &MCModule::setNumLandmarks, "This is to initialize number of landmark locations"},
    {"setLandmarks", &MCModule::setLandmarks, "This is to initialize landmark locations"},
    {"setNumParticles", &MCModule::setNumParticles, "This is to initialize number of particles for MC simulation"},
    {"setInitialCovariance", &MCModule::setInitialCovariance, "This is to initialize first state covariance uncertainty"},
    {"setPathLength", &MCModule::setPathLength, "This is to initialize the length of the path"},
    {"setTrajectory", &MCModule::setTrajectory, "This is to initialize the trajectory"},
    {"setOdometry", &MCModule::setOdometry, "This is to initialize the odometry"},
    {"runSimulation", &MCModule::runSimulation, "This is to run a MC simulation"},
    {"setNumGaussians", &MCModule::setNumGaussians, "This is to set number of Gaussians in mixture"},
    {"runGMMEstimation", &MCModule::runGMMEstimation, "Use sampling-based GMM algorithm to estimate probability of collision"},
    {"setNumGMMSamples", &MCModule::setNumGMMSamples, "This is to set number of samples for GMM collision estimation"},
    {"help", &MCModule::help, "display this message."},
    // additions
    {"runGMMEstimationBatch", &MCModule::runGMMEstimationBatch, "runGMMEstimationBatch <n>: the next n runGMMEstimation runs in the same launches; n probabilities"},
    {"setSeed", &MCModule::setSeed, "setSeed <u64>: Philox key; restarts the run counter"},
    {"setMode", &MCModule::setMode, "setMode shared|philox|deterministic: how the draws of a run are produced"},
    {"setDrawsFile", &MCModule::setDrawsFile, "setDrawsFile <path>: float64 little-endian standard normals in the reference's consumption order"},
    {"setDraws", &MCModule::setDraws, "setDraws <n> <v_0> ... <v_{n-1}>: the same, inline"},
    {"setGMMCountsFile", &MCModule::setGMMCountsFile, "setGMMCountsFile <path>: int32 little-endian multinomial counts [T x K] (deterministic GMM mode)"},
    {"setGMMCounts", &MCModule::setGMMCounts, "setGMMCounts <n> <c_0> ...: the same, inline"},
    {"setEarlyExit", &MCModule::setEarlyExit, "setEarlyExit 0|1: stop a rollout at its first hit (the returned P does not change)"},
    {"setBroadPhase", &MCModule::setBroadPhase, "setBroadPhase 0|1: 1 (default) tests only the obstacles within reach of each waypoint, 0 every obstacle at every step (the returned P does not change)"},
    {"setEstimator", &MCModule::setEstimator, "setEstimator reference|closedloop: collision tested on the particle (reference) or on the real state"},
    {"setRolloutOffset", &MCModule::setRolloutOffset, "setRolloutOffset <u64>: global id of this module's first rollout (shards of one job over several modules / processes)"},
    {"setFilter", &MCModule::setFilter, "setFilter state | deviation: runGMMEstimation / runLinearEstimation propagate each component with the reference's EKF on the state (default) or with a Kalman filter on the deviation from the nominal plan, linearised about the plan"},
    {"setRefit", &MCModule::setRefit, "setRefit moments | em [iterations]: runGMMEstimation refits each component by moment matching (the reference, default) or refits the whole mixture to the collision-free samples by EM started from the moment-matched one (default 5 iterations)"},
    {"setConstraints", &MCModule::setConstraints, "setConstraints <n> then n x (ax ay b xmin xmax ymin ymax): the free space as linear constraints ax x + ay y <= b inside a window (linConstraint.py:4-11; 'inf' / '-inf' for an open side)"},
    {"runLinearEstimation", &MCModule::runLinearEstimation, "P(collision) by propagating the mixture with the analytic (Patil) truncation against the constraints of setConstraints instead of sampling"},
    {"setDevices", &MCModule::setDevices, "setDevices all | <ordinal> [<ordinal> ...]: the CUDA devices this module runs on; runSimulation splits its rollouts over them (one NCCL all-reduce of the counters), runGMMEstimationBatch its runs"},
    {"getDevices", &MCModule::getDevices, "the CUDA ordinals in use"},
    {"describeRun", &MCModule::describeRun, "describeRun [k]: Philox key, first rollout id and rollout count of the k-th runSimulation from now (default 0 = the next one), then key and stream id of the k-th runGMMEstimation"},
    {"setGeometry", &MCModule::setGeometry, "setGeometry <n_boxes> <n_hulls> <8 numbers per prism ...>: replace the collision geometry"},
    {"setRobotModel", &MCModule::setRobotModel, "setRobotModel <file.zae | file.dae> [boxes | polygons] [z0 z1 ... zk]: robot hulls from a COLLADA model -- per z-band of the model one convex footprint (polygons, at most 8 vertices; default) or its minimum-area bounding rectangle (boxes: looser, but the packed kernels run); returns the number of hulls"},
    {"setRobotPolygons", &MCModule::setRobotPolygons, "setRobotPolygons <n> then per hull: zlo zhi m x0 y0 ... x(m-1) y(m-1): convex-polygon robot hulls (robot frame, 3..8 vertices) next to / instead of the box hulls"},
    {"setScene", &MCModule::setScene, "setScene <env.xml> [fixture_dir ...]: extract the collision geometry from an OpenRAVE env XML"},
    {"getWaypointProbabilities", &MCModule::getWaypointProbabilities, "per-waypoint colliding fraction of the last runGMMEstimation"},
    {"getFirstHitHistogram", &MCModule::getFirstHitHistogram, "rollouts whose first hit is at waypoint t, last runSimulation"},
    {"getGMM", &MCModule::getGMM, "mixture after the last truncation: K rows of w mu[3] cov[9]"},
    {"getStats", &MCModule::getStats, "hits rollouts executed_steps pair_tests kernel_ms degenerate of the last run"},
};

std::vector<std::pair<std::string, std::string>> MCModule::command_list() {
    std::vector<std::pair<std::string, std::string>> out;
    for (const auto& c : kCommands) out.emplace_back(c.name, c.help);
    return out;
}

MCModule::MCModule(const Scene& scene, int device) : scene_(scene) {
    std::random_device rd;  // the reference seeds from the clock / random device (MCSimulator.h:141 set_seed_random)
    seed_ = (uint64_t(rd()) << 32) ^ uint64_t(rd());
    if (device < 0) return;  // parse-only module: setters and queries work, run* / geometry uploads fail loudly
    open_devices(std::vector<int>{device});
}

MCModule::~MCModule() { mcsim_cuda_group_destroy(group_); }

// (re)creates the device group and gives every device the scene
void MCModule::open_devices(const std::vector<int>& devices) {
    mcsim_cuda_group* g = nullptr;
    cuda_check(mcsim_cuda_group_create(devices.data(), int(devices.size()), &g), "mcsim_cuda_group_create");
    mcsim_cuda_group* old = group_;
    group_ = g;
    ctx_ = mcsim_cuda_group_ctx(g, 0);
    try {
        if (scene_.n_hulls() + scene_.n_poly_hulls() > 0) upload_scene();
    } catch (...) {
        mcsim_cuda_group_destroy(g);
        group_ = old;
        ctx_ = old ? mcsim_cuda_group_ctx(old, 0) : nullptr;
        throw;
    }
    mcsim_cuda_group_destroy(old);
}

void MCModule::upload_scene() {
    if (!group_) return;  // parse-only module
    cuda_check(mcsim_cuda_group_set_scene_ex(group_, scene_.boxes.data(), scene_.n_boxes(), scene_.hulls.data(), scene_.n_hulls(),
                                             scene_.poly_hulls.data(), scene_.n_poly_hulls()), "mcsim_cuda_set_scene");
}

void MCModule::io_bytes(uint64_t* h2d, uint64_t* d2h) const { mcsim_cuda_io_bytes(ctx_, h2d, d2h); }

bool MCModule::SendCommand(std::ostream& sout, std::istream& sinput) {
    std::string name;
    if (!(sinput >> name)) return false;
    const std::string key = lower(name);
    for (const auto& c : kCommands)
        if (lower(c.name) == key) {
            NvtxRange range(c.name);
            return (this->*c.fn)(sout, sinput);
        }
    throw std::runtime_error("unknown command '" + name + "' (send 'help' for the list)");
}

bool MCModule::SendCommand(const std::string& line, std::string& out) {
    std::istringstream in(line);
    std::ostringstream os;
    err_.clear();
    try {
        const bool ok = SendCommand(os, in);
        out = os.str();
        if (!ok) err_ = "command returned false";
        return ok;
    } catch (const std::exception& e) {
        err_ = e.what();
        out.clear();
        return false;
    }
}

// ---- the reference's commands ---------------------------------------------------------------------------------
bool MCModule::MyCommand(std::ostream& sout, std::istream& sinput) {  // mcsimplugin.cpp:225-231
    std::string input;
    sinput >> input;
    sout << "output";
    return true;
}

// mcsimplugin.cpp:189-223: an Armadillo installation smoke test -- a random 4x4 matrix A, A(0,:) = A(1,:) + A(3,:),
// last column zeroed, B = diag(A); both saved as arma_ascii ("ARMA_MAT_TXT_FN008") to A_mat.txt / B_mat.txt.
bool MCModule::ArmaCommand(std::ostream&, std::istream&) {
    std::mt19937_64 gen(seed_ ^ 0xA53Aull);
    std::uniform_real_distribution<double> uni(0.0, 1.0);
    double A[4][4], B[4][4] = {};
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) A[r][c] = uni(gen);
    for (int c = 0; c < 4; ++c) A[0][c] = A[1][c] + A[3][c];
    for (int r = 0; r < 4; ++r) A[r][3] = 0.0;
    for (int i = 0; i < 4; ++i) B[i][i] = A[i][i];
    auto save = [](const char* path, const double M[4][4]) {
        FILE* f = std::fopen(path, "w");
        if (!f) throw std::runtime_error(std::string("cannot write ") + path);
        std::fprintf(f, "ARMA_MAT_TXT_FN008\n4 4\n");
        for (int r = 0; r < 4; ++r) {
            for (int c = 0; c < 4; ++c) std::fprintf(f, " %22.14e", M[r][c]);
            std::fprintf(f, "\n");
        }
        std::fclose(f);
    };
    save("A_mat.txt", A);
    save("B_mat.txt", B);
    return true;
}

bool MCModule::setAlphas(std::ostream&, std::istream& sinput) {  // :174-187, reads every remaining token
    std::vector<double> a;
    std::string tok;
    while (sinput >> tok) a.push_back(std::stod(tok));
    if (a.size() < 4) throw std::runtime_error("setAlphas: 4 (squared) alphas expected, got " + std::to_string(a.size()));
    alphas_ = a;
    return true;
}

bool MCModule::setQ(std::ostream&, std::istream& sinput) {  // :168-172
    if (!(sinput >> Q_)) throw std::runtime_error("setQ: a number expected");
    return true;
}

bool MCModule::setNumLandmarks(std::ostream&, std::istream& sinput) {  // :142-146
    int n;
    if (!(sinput >> n) || n < 0 || n > MCSIM_MAX_LANDMARKS) throw std::runtime_error("setNumLandmarks: 0.." + std::to_string(MCSIM_MAX_LANDMARKS) + " expected");
    num_landmarks_ = n;
    return true;
}

bool MCModule::setLandmarks(std::ostream&, std::istream& sinput) {  // :148-166: all x, then all y
    read_doubles(sinput, size_t(2) * num_landmarks_, landmarks_, "setLandmarks");
    return true;
}

bool MCModule::setNumParticles(std::ostream&, std::istream& sinput) {  // :136-140
    long long n;
    if (!(sinput >> n) || n < 0) throw std::runtime_error("setNumParticles: a non-negative integer expected");
    num_particles_ = n;
    return true;
}

bool MCModule::setInitialCovariance(std::ostream&, std::istream& sinput) {  // :121-134, row-major
    std::vector<double> v;
    read_doubles(sinput, 9, v, "setInitialCovariance");
    std::copy(v.begin(), v.end(), cov0_);
    cov0_set_ = true;
    return true;
}

bool MCModule::setPathLength(std::ostream&, std::istream& sinput) {  // :115-119
    int n;
    if (!(sinput >> n) || n < 1 || n > MCSIM_MAX_WAYPOINTS) throw std::runtime_error("setPathLength: 1.." + std::to_string(MCSIM_MAX_WAYPOINTS) + " expected");
    path_length_ = n;
    return true;
}

bool MCModule::setTrajectory(std::ostream&, std::istream& sinput) {  // :83-97, component-major; also mu0 (MCSimulator.h:164)
    if (path_length_ < 1) throw std::runtime_error("setTrajectory: setPathLength first");
    read_doubles(sinput, size_t(3) * path_length_, trajectory_, "setTrajectory");
    return true;
}

bool MCModule::setOdometry(std::ostream&, std::istream& sinput) {  // :99-113
    if (path_length_ < 1) throw std::runtime_error("setOdometry: setPathLength first");
    read_doubles(sinput, size_t(3) * (path_length_ - 1), odometry_, "setOdometry");
    return true;
}

bool MCModule::setNumGaussians(std::ostream&, std::istream& sinput) {  // :49-54
    int n;
    if (!(sinput >> n) || n < 1 || n > MCSIM_MAX_GAUSSIANS) throw std::runtime_error("setNumGaussians: 1.." + std::to_string(MCSIM_MAX_GAUSSIANS) + " expected");
    num_gaussians_ = n;
    return true;
}

bool MCModule::setNumGMMSamples(std::ostream&, std::istream& sinput) {  // :56-61
    long long n;
    if (!(sinput >> n) || n < 1) throw std::runtime_error("setNumGMMSamples: a positive integer expected");
    num_gmm_samples_ = n;
    return true;
}

bool MCModule::help(std::ostream& sout, std::istream&) {
    for (const auto& c : kCommands) sout << c.name << " : " << c.help << "\n";
    return true;
}

}  // namespace mcsim
