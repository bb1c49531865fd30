// mc_module.h -- the plugin's command surface without OpenRAVE: the same command names, token grammar and
// returned text as `class MCModule : public ModuleBase` of the reference (mcsimplugin/mcsimplugin.cpp:7-231),
// over the CUDA C ABI (include/mcsim_cuda.h) instead of MCSimulator.h + Armadillo + OpenRAVE CheckCollision.
//
// Two front ends share this class: the OpenRAVE adapter (openrave_adapter.cpp, built where OpenRAVE exists)
// and the headless C ABI (mcsim_capi.cpp -> libmcsim.so, include/mcsim.h).
#ifndef MCSIM_MC_MODULE_H
#define MCSIM_MC_MODULE_H

#include <cstdint>
#include <functional>
#include <iosfwd>
#include <string>
#include <utility>
#include <vector>

#include "mcsim_cuda.h"
#include "scene_xml.h"

namespace mcsim {

class MCModule {
  public:
    // How the random draws of a run are produced.
    //   kShared        (default) the reference's law, run by run: ONE closed-loop real robot per run whose noisy
    //                  command moves every particle (MCSimulator.h:754,760); its draws come from a host Philox
    //                  stream, the particles' start poses from per-rollout Philox blocks on the device.
    //   kPhilox        every rollout owns an independent real-robot thread: P is the expectation of kShared's P.
    //   kDeterministic all draws are injected (setDrawsFile / setDraws), in the reference's consumption order.
    enum class Mode { kShared, kPhilox, kDeterministic };

    MCModule(const Scene& scene, int device);
    ~MCModule();
    MCModule(const MCModule&) = delete;
    MCModule& operator=(const MCModule&) = delete;

    // ModuleBase::SendCommand semantics: the first token of `sinput` selects the command (case-insensitive);
    // false when the command is unknown or its handler returns false.  Handlers may throw std::runtime_error.
    bool SendCommand(std::ostream& sout, std::istream& sinput);
    // Convenience for the C ABI: never throws; the message of a failure is kept in last_error().
    bool SendCommand(const std::string& line, std::string& out);
    const std::string& last_error() const { return err_; }
    // every command name with its help string (what a front end registers: openrave_adapter.cpp)
    static std::vector<std::pair<std::string, std::string>> command_list();
    const Scene& scene() const { return scene_; }
    void io_bytes(uint64_t* h2d, uint64_t* d2h) const;

  private:
    using Handler = bool (MCModule::*)(std::ostream&, std::istream&);
    struct Command { const char* name; Handler fn; const char* help; };
    static const Command kCommands[];

    // -- the reference's 16 commands (mcsimplugin.cpp:13-44)
    bool MyCommand(std::ostream&, std::istream&);
    bool ArmaCommand(std::ostream&, std::istream&);
    bool setAlphas(std::ostream&, std::istream&);
    bool setQ(std::ostream&, std::istream&);
    bool setNumLandmarks(std::ostream&, std::istream&);
    bool setLandmarks(std::ostream&, std::istream&);
    bool setNumParticles(std::ostream&, std::istream&);
    bool setInitialCovariance(std::ostream&, std::istream&);
    bool setPathLength(std::ostream&, std::istream&);
    bool setTrajectory(std::ostream&, std::istream&);
    bool setOdometry(std::ostream&, std::istream&);
    bool runSimulation(std::ostream&, std::istream&);
    bool setNumGaussians(std::ostream&, std::istream&);
    bool runGMMEstimation(std::ostream&, std::istream&);
    bool runGMMEstimationBatch(std::ostream&, std::istream&);
    bool setNumGMMSamples(std::ostream&, std::istream&);
    bool help(std::ostream&, std::istream&);
    // -- additions (never required by the reference driver)
    bool setSeed(std::ostream&, std::istream&);
    bool setMode(std::ostream&, std::istream&);
    bool setDrawsFile(std::ostream&, std::istream&);
    bool setDraws(std::ostream&, std::istream&);
    bool setGMMCountsFile(std::ostream&, std::istream&);
    bool setGMMCounts(std::ostream&, std::istream&);
    bool setEarlyExit(std::ostream&, std::istream&);
    bool setBroadPhase(std::ostream&, std::istream&);
    bool setEstimator(std::ostream&, std::istream&);
    bool setRolloutOffset(std::ostream&, std::istream&);
    bool setRefit(std::ostream&, std::istream&);
    bool setFilter(std::ostream&, std::istream&);
    bool setConstraints(std::ostream&, std::istream&);
    bool runLinearEstimation(std::ostream&, std::istream&);
    bool setDevices(std::ostream&, std::istream&);
    bool getDevices(std::ostream&, std::istream&);
    bool describeRun(std::ostream&, std::istream&);
    bool setGeometry(std::ostream&, std::istream&);
    bool setScene(std::ostream&, std::istream&);
    bool setRobotPolygons(std::ostream&, std::istream&);
    bool setRobotModel(std::ostream&, std::istream&);
    bool getWaypointProbabilities(std::ostream&, std::istream&);
    bool getFirstHitHistogram(std::ostream&, std::istream&);
    bool getGMM(std::ostream&, std::istream&);
    bool getStats(std::ostream&, std::istream&);

    void fill_params(mcsim_mc_params& p) const;  // throws when a parameter is missing
    void upload_scene();
    void open_devices(const std::vector<int>& devices);
    // Philox key of run number `run` of this module: the seed for run 0, seed ^ splitmix64(run) afterwards.  The run
    // number lives in the KEY and the shard offset (setRolloutOffset) in the COUNTER, so R runs on W shards draw W x R
    // disjoint (key, id) blocks whatever the offsets are.
    uint64_t run_key(uint64_t run) const;
    // Philox key of the GMM runs of this module: the seed, salted with the shard offset (a replica id for GMM mode:
    // one run does not shard); the run number is the stream id in the counter, as a batch numbers its members.
    uint64_t gmm_key() const;

    Scene scene_;
    mcsim_cuda_group* group_ = nullptr;   // one context per device (setDevices) + the NCCL communicators; owned
    mcsim_cuda_ctx* ctx_ = nullptr;       // the group's first context (single-run GMM, queries)
    std::string err_;
    // parameters, as the setters leave them
    std::vector<double> alphas_;
    double Q_ = 0.0;
    int num_landmarks_ = 0;
    std::vector<double> landmarks_;     // x_0..x_{L-1} y_0..y_{L-1}
    int64_t num_particles_ = 0;
    double cov0_[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    bool cov0_set_ = false;
    int path_length_ = 0;
    std::vector<double> trajectory_;    // 3 x T component-major
    std::vector<double> odometry_;      // 3 x (T-1) component-major
    int num_gaussians_ = 0;
    int64_t num_gmm_samples_ = 0;
    // run control
    Mode mode_ = Mode::kShared;
    uint64_t seed_ = 0;
    uint64_t run_index_ = 0;
    uint64_t rollout_offset_ = 0;
    int early_exit_ = 1;
    int exhaustive_ = 0;
    int estimator_ = MCSIM_EST_REFERENCE;
    std::vector<double> draws_;
    std::vector<int> gmm_counts_;
    int refit_ = MCSIM_GMM_REFIT_MOMENTS;   // setRefit: how truncateGMM refits the mixture
    int gmm_filter_ = MCSIM_FILTER_STATE;   // setFilter: EKF on the state (reference) or Kalman filter on the deviation from the plan
    int em_iters_ = 0;
    std::vector<double> con_A_, con_b_;     // setConstraints, expanded into half-planes A (x, y, theta) <= b
    // results of the last run
    std::vector<double> probabilities_;
    std::vector<uint64_t> first_hit_hist_;
    std::vector<double> gmm_post_;
    int gmm_post_K_ = 0;
    uint64_t last_hits_ = 0, last_steps_ = 0, last_tests_ = 0, last_n_ = 0;
    float last_kernel_ms_ = 0.f;
    int last_degenerate_ = 0;
};

}  // namespace mcsim
#endif
