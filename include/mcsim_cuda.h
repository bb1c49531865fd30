/* mcsim_cuda.h -- thin C ABI between the C++ plugin (libmcsim.so) and the sm_100a kernels
 * (libmcsim_cuda.so).  Plain pointers and sizes; no exceptions, no torch types; every call
 * returns an int status (0 = ok) and leaves a message for mcsim_cuda_last_error().
 *
 * Each entry point names the reference code it replaces (paths relative to the reference
 * repository ajaayc/Probability-of-Collision-for-Safe-Planning):
 *
 *   mcsim_cuda_mc_run        mcsimplugin/MCSimulator.h:287-347 (initParticles, moveParticles,
 *                            checkParticleCollisions, getCollisionProportion) inside the loop
 *                            of EKF_GaussProp :649-864, collision test :257-285
 *   mcsim_cuda_gmm_truncate  mcsimplugin/MCSimulator.h:570-642 (truncateGMM) +
 *                            mcsimplugin/GM_Model.h:83-116 (sampleNPoints)
 *   mcsim_cuda_gmm_run       the GMM branch of EKF_GaussProp :672-678, :766-772, :804-817, :843-856
 *   mcsim_cuda_em_*          armadillo-8.400.0/include/armadillo_bits/gmm_full_meat.hpp:2390-2800
 *                            (em_iterate / em_generate_acc / em_update_params / em_fix_params)
 *   mcsim_cuda_score_plans   the MC loop above, batched over many plans (BASELINE.json config 4)
 *
 * All host arrays are caller-owned.  Unless stated otherwise, pointers are HOST pointers and the
 * call copies in / out on the context's stream and synchronises before returning.
 */
#ifndef MCSIM_CUDA_H
#define MCSIM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCSIM_PRISM_STRIDE 8      /* cx, cy, cos(phi), sin(phi), hx, hy, zlo, zhi */
#define MCSIM_MAX_HULLS 8
#define MCSIM_MAX_POLY_HULLS 4     /* convex-polygon robot hulls (mcsim_cuda_set_scene_ex) */
#define MCSIM_POLY_STRIDE 19       /* n, zlo, zhi, x0, y0, ..., x7, y7: counter-clockwise, robot frame, at most 8 vertices */
#define MCSIM_MAX_PAIRS 1024      /* live (hull, box) pairs held in __constant__ memory */
#define MCSIM_MAX_LANDMARKS 16
#define MCSIM_MAX_WAYPOINTS 1024
#define MCSIM_MAX_GAUSSIANS 16

typedef struct mcsim_cuda_ctx mcsim_cuda_ctx;

/* status codes */
enum {
    MCSIM_OK = 0,
    MCSIM_ERR_CUDA = 1,       /* a CUDA runtime call failed (message has the CUDA error string) */
    MCSIM_ERR_ARG = 2,        /* bad argument / limits exceeded */
    MCSIM_ERR_NODEVICE = 3,   /* no sm_100 device visible: there is NO CPU fallback */
    MCSIM_ERR_DRAWS = 4,      /* injected draws buffer too short */
    MCSIM_ERR_DEGENERATE = 5, /* GMM component with < 2 collision-free samples (reference: UB) */
    MCSIM_ERR_NCCL = 6        /* a multi-device group needs libnccl.so.2 and it failed (message has the NCCL error) */
};

/* ---- context ---------------------------------------------------------------------------- */
/* device: CUDA ordinal.  Creates one non-blocking stream; fails with MCSIM_ERR_NODEVICE when no
 * CUDA device is present (the product has no CPU path). */
int mcsim_cuda_create(int device, mcsim_cuda_ctx** out);
void mcsim_cuda_destroy(mcsim_cuda_ctx* ctx);
const char* mcsim_cuda_last_error(void);
/* cudaStream_t of the context as an opaque pointer (so a caller can record CUDA events on it). */
void* mcsim_cuda_stream(mcsim_cuda_ctx* ctx);
/* CUDA ordinal the context was created on (-1 for NULL). */
int mcsim_cuda_ctx_device(const mcsim_cuda_ctx* ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches). */
uint64_t mcsim_cuda_launch_count(mcsim_cuda_ctx* ctx);

/* Measures the FP32 FMA-pipe peak (TFLOP/s, FFMA = 2 flops) and the SFU peak (10^12 MUFU op/s) of
 * the device with register-resident dependent chains, timed with CUDA events on the context's
 * stream for about `ms` milliseconds each.  These are the roofline denominators of the MC path
 * (MEASURED_PEAKS.json has HBM and tensor peaks only). */
int mcsim_cuda_measure_peaks(mcsim_cuda_ctx* ctx, double ms, double* fp32_tflops, double* mufu_tops);
/* One pipe at a time: which = 0 scalar FFMA chain (TFLOP/s), 1 MUFU chain (10^12 op/s),
 * 2 packed FP32 pairs, fma.rn.f32x2 / SASS FFMA2 (TFLOP/s). */
int mcsim_cuda_measure_peak(mcsim_cuda_ctx* ctx, int which, double ms, double* out);

/* ---- scene ------------------------------------------------------------------------------ */
/* Replaces OpenRAVE's collision world for this path (MCSimulator.h:269-285).  boxes: n_boxes x 8
 * doubles in the world frame, hulls: n_hulls x 8 doubles in the robot frame (z absolute).  The
 * live (hull, box) pairs -- z-intervals overlapping with positive length -- are flattened into
 * an FP32 table in __constant__ memory, grouped by hull. */
int mcsim_cuda_set_scene(mcsim_cuda_ctx* ctx, const double* boxes, int n_boxes,
                         const double* hulls, int n_hulls);
/* The same with robot hulls beyond boxes (the PR2 model the reference loads at pr2test2.env.xml:120 is
 * not a set of boxes): poly_hulls = n_poly x MCSIM_POLY_STRIDE doubles, each a convex polygon in the robot frame
 * extruded over [zlo, zhi].  n_hulls may then be 0.  Polygon hulls are tested by the scalar kernels (2.5-D SAT on the
 * box's axes and the polygon's edge normals): MC replay / Philox (the precise kernel) and the per-waypoint GMM kernels
 * in Philox mode; the packed kernels' tables hold box hulls only, so batched scoring, batched GMM runs and the GMM replay
 * mode refuse a scene with polygon hulls (MCSIM_ERR_ARG). */
int mcsim_cuda_set_scene_ex(mcsim_cuda_ctx* ctx, const double* boxes, int n_boxes, const double* hulls, int n_hulls,
                            const double* poly_hulls, int n_poly);

/* ---- Monte-Carlo rollouts --------------------------------------------------------------- */
typedef struct {
    /* plan */
    int T;                       /* waypoints (setPathLength) */
    const double* trajectory;    /* 3 x T component-major: x_0..x_{T-1} y_0.. th_0.. (setTrajectory) */
    const double* odometry;      /* 3 x (T-1) component-major: rot1.., trans.., rot2.. (setOdometry) */
    /* noise model */
    double alphas[4];            /* squared alphas (setAlphas) */
    double Q;                    /* range variance (setQ) */
    int n_landmarks;
    const double* landmarks;     /* x_0..x_{L-1} y_0..y_{L-1} (setLandmarks) */
    double cov0[9];              /* Sigma_0 row-major (setInitialCovariance) */
    /* rollouts */
    uint64_t n_rollouts;         /* rollouts THIS call simulates */
    uint64_t rollout_offset;     /* global id of the first one (multi-GPU shards / resume) */
    /* mode */
    int mode;                    /* MCSIM_MODE_REPLAY or MCSIM_MODE_PHILOX */
    int estimator;               /* MCSIM_EST_REFERENCE (collision tested on the particle, which is
                                    moved by the real robot's noisy command: MCSimulator.h:754-761)
                                    or MCSIM_EST_CLOSEDLOOP (tested on the real state) */
    int early_exit;              /* 1: a rollout stops at its first hit (flags only);
                                    0: full per-rollout counts as the reference keeps them (:344) */
    uint64_t seed;               /* Philox key */
    /* MCSIM_MODE_REPLAY inputs */
    const double* applied;       /* 3 x (T-1) component-major noisy commands u~_t of the shared
                                    real-robot thread (host-computed, see mcsim_host_thread) */
    const double* z0;            /* 3 x n_rollouts standard normals, rollout-major (the order
                                    mvnrnd fills them: glue_mvnrnd_meat.hpp:89-147); NULL: every
                                    rollout draws its start pose from its own Philox block
                                    (always so in PHILOX mode) */
    double near_eps;             /* |clearance| < near_eps marks a near-contact pose (audit only) */
    int precise;                 /* PHILOX mode only.  0: packed two-rollouts-per-thread kernel with MUFU
                                    approximations (mc_philox2_sm100.cuh); 1: scalar kernel with IEEE
                                    division / sqrt / libm trigonometry (same Philox counters and draw
                                    order; audit launches always use it) */
    int exhaustive;              /* packed kernel only.  0: exact broad phase -- per waypoint the pairs within
                                    reach of the nominal pose are listed on the host, a warp whose rollouts
                                    are all near x*_t tests only those (same flags and counts, fewer pair
                                    tests; result.pair_tests counts the executed ones); 1: every live
                                    (hull, box) pair is tested at every step, as CheckCollision does */
    int no_compaction;           /* packed kernel with early_exit only.  0: runs of >= 2^18 rollouts are cut into segments
                                    of 4 waypoints, the survivors of a segment are compacted in device memory and the
                                    next launch runs full warps of them (same flags, hits and histogram; executed
                                    rollout-steps drop to the free ones); 1: one launch, a warp leaves when all its
                                    rollouts have hit */
    int gmm_filter;              /* GMM / linear runs only: how a component's belief moves from waypoint t to t + 1.
                                    MCSIM_FILTER_STATE (0): the reference's EKF on the state, linearised at the
                                    component mean (MCSimulator.h:868-929).  MCSIM_FILTER_DEVIATION: method variant
                                    (the paper's future work, finalpaper/ajaay_paper.tex:893, 901): a Kalman
                                    filter on the deviation from the nominal plan, motion and sensor models linearised once
                                    about (x*_t, u*_t) as in Patil et al. -- csrc/estimator_math.h lkf_*.  No reference
                                    implementation exists: PARITY UNPINNED, checked against a NumPy restatement */
} mcsim_mc_params;

enum { MCSIM_MODE_REPLAY = 0, MCSIM_MODE_PHILOX = 1 };
enum { MCSIM_EST_REFERENCE = 0, MCSIM_EST_CLOSEDLOOP = 1 };
enum { MCSIM_FILTER_STATE = 0, MCSIM_FILTER_DEVIATION = 1 };

typedef struct {
    uint64_t hits;               /* rollouts with >= 1 colliding waypoint */
    uint64_t near_contact;       /* poses with |clearance| < near_eps (audit launches only) */
    uint64_t executed_steps;     /* rollout-steps actually executed (early exit skips the rest) */
    uint64_t pair_tests;         /* SAT pair tests actually executed */
    /* optional outputs, may be NULL */
    uint64_t* first_hit_hist;    /* [T] rollouts whose first hit is at waypoint t */
    uint32_t* counts;            /* [n_rollouts] colliding waypoints per rollout (early_exit = 0)
                                    or 0/1 flags (early_exit = 1) */
    float* poses;                /* [T x 3 x n_rollouts] pose tested at every waypoint (audit) */
    float* clearance;            /* [T x n_rollouts] signed clearance (audit) */
    float kernel_ms;             /* device time of the rollout kernel (CUDA events) */
} mcsim_mc_result;

/* Runs n_rollouts rollouts over the plan.  Audit outputs (poses / clearance / near_contact) are
 * produced iff result->poses or result->clearance is non-NULL; that launch has no early exit. */
int mcsim_cuda_mc_run(mcsim_cuda_ctx* ctx, const mcsim_mc_params* p, mcsim_mc_result* r);

/* Device-resident variant used for throughput measurement: uploads the plan once ... */
int mcsim_cuda_mc_prepare(mcsim_cuda_ctx* ctx, const mcsim_mc_params* p);
/* ... then launches the rollout kernel only (no host copies inside); counters accumulate on the
 * device until mcsim_cuda_mc_collect.  Asynchronous on the context's stream. */
int mcsim_cuda_mc_launch(mcsim_cuda_ctx* ctx, uint64_t n_rollouts, uint64_t rollout_offset);
int mcsim_cuda_mc_collect(mcsim_cuda_ctx* ctx, mcsim_mc_result* r);
/* DEVICE pointer of the counter block the launches accumulate into: uint64[4 + T] =
 * {hits, near_contact, executed_steps, pair_tests, first_hit_hist[T]}.  A multi-GPU caller
 * all-reduces (sum) this buffer on the context's stream between mc_launch and mc_collect
 * (bench.py does it with torch.distributed/NCCL); allocated at mcsim_cuda_create, never moves. */
void* mcsim_cuda_mc_counters(mcsim_cuda_ctx* ctx);

/* ---- host thread of the reference estimator ---------------------------------------------- */
/* The single closed-loop "real robot" of EKF_GaussProp (MCSimulator.h:685-800): controller gain
 * (:532-553), applied control (:719-726), EKFpredict (:868-881), sampleOdometry (:391-410),
 * 8 sampleObservation draws (:383-387, :786-789), EKFupdate (:883-929); FP64, on the HOST (it is
 * O(T) scalar work, shared by all rollouts in replay mode).  normals: 11 x (T-1) standard
 * normals in consumption order (3 odometry + L observations per step).
 * Outputs (each may be NULL): applied_noisy 3 x (T-1) component-major = u~_t handed to
 * moveParticles (:760); applied 3 x (T-1) = commanded u_t; real_path 3 x T; mu_path 3 x T,
 * cov_path T x 9 row-major; obs (T-1) x L. */
int mcsim_host_thread(const mcsim_mc_params* p, const double* normals, size_t n_normals,
                      double* applied_noisy, double* applied, double* real_path,
                      double* mu_path, double* cov_path, double* obs);

/* Standard normals of a Philox-driven host thread: per_step x (T-1) doubles in consumption order,
 * counter = (step, stream_id, block, PURPOSE_GMM_HOST), key = seed; Box-Muller in FP64.  Used when
 * the shared real-robot thread is not replayed from a draws file (plugin modes "shared" and the
 * Philox GMM mode). */
int mcsim_host_normals(uint64_t seed, uint32_t stream_id, int T, int per_step, double* out);

/* The 12 standard normals one Philox-mode rollout draws for the step t -> t+1 (3 odometry, then the
 * observations), recomputed on the host in FP64 from the same bits the kernels use: two Philox4x32-10
 * blocks, counter = (rollout id lo, hi, 8 t + {0, 1}, PURPOSE_STEP), key = seed, six Box-Muller pairs at
 * 24 radius + 16 angle bits (csrc/philox.h, "step draws").  For audits of the Philox mode; no GPU. */
int mcsim_step_normals(uint64_t seed, uint64_t rollout_id, int t, double out[12]);

/* The same 12 draws for rollouts id0 .. id0 + n - 1 as the packed rollout kernel computes them ON THE DEVICE (FP32,
 * MUFU lg2 / sqrt / sin / cos; the device function mc_philox2_kernel itself calls): out = [n][12] floats on the host.
 * For distribution tests of the Philox-mode noise (Kolmogorov-Smirnov, tail mass) and audits against mcsim_step_normals. */
int mcsim_cuda_step_normals(mcsim_cuda_ctx* ctx, uint64_t seed, uint64_t id0, uint64_t n, int t, float* out);

/* ---- GMM truncation (moment matching) ---------------------------------------------------- */
typedef struct {
    int K;                       /* components (setNumGaussians) */
    const double* weights;       /* [K] */
    const double* means;         /* [K x 3] */
    const double* covs;          /* [K x 9] row-major */
    uint64_t n_samples;          /* numGMMSamples (setNumGMMSamples) */
    int mode;                    /* MCSIM_MODE_REPLAY: counts + normals injected; MCSIM_MODE_PHILOX */
    const int* counts;           /* [K] multinomial counts (replay; GM_Model.h:89-93) */
    const double* normals;       /* 3 x n_samples standard normals in consumption order: component
                                    0's samples first, sample-major (GM_Model.h:101-112) */
    uint64_t seed;               /* Philox key */
    uint64_t stream_id;          /* Philox subsequence (waypoint index / trajectory id) */
    int refit;                   /* MCSIM_GMM_REFIT_MOMENTS: the reference's per-component moment matching (truncateGMM);
                                    MCSIM_GMM_REFIT_EM: then em_iters EM iterations of the whole mixture on ALL
                                    collision-free samples, started from the moment-matched mixture (the method variant
                                    finalpaper/ajaay_paper.tex:901-903 lists as future work; oracle:
                                    arma::gmm_full::learn(keep_existing) on the reference's free samples) */
    int em_iters;
} mcsim_gmm_truncate_params;

enum { MCSIM_GMM_REFIT_MOMENTS = 0, MCSIM_GMM_REFIT_EM = 1 };

typedef struct {
    double weights[MCSIM_MAX_GAUSSIANS];
    double means[MCSIM_MAX_GAUSSIANS * 3];
    double covs[MCSIM_MAX_GAUSSIANS * 9];
    uint64_t n_coll[MCSIM_MAX_GAUSSIANS];
    uint64_t n_free[MCSIM_MAX_GAUSSIANS];
    double p_collision;          /* sum n_coll / n_samples (MCSimulator.h:633-641) */
    int degenerate;              /* components with < 2 free samples: previous moments kept, weight
                                    from n_free (reference: undefined behaviour) */
} mcsim_gmm_truncate_result;

/* One truncateGMM (MCSimulator.h:570-642): sample, collision-test, refit every component by the
 * mean / unbiased covariance of its collision-free samples, reweight by free-sample share. */
int mcsim_cuda_gmm_truncate(mcsim_cuda_ctx* ctx, const mcsim_gmm_truncate_params* p,
                            mcsim_gmm_truncate_result* r);

typedef struct {
    int K;
    uint64_t n_samples;
    /* replay mode only: the reference's draw stream for one runGMMEstimation, in consumption
     * order: 3N (waypoint 0), then per step 3 (odometry) + L (observations) + 3N (truncation). */
    const double* normals;
    size_t n_normals;
    const int* counts;           /* [T x K] multinomial counts per waypoint (replay) */
    int kernel;                  /* MCSIM_GMM_KERNEL_*: which launch shape runs a Philox-mode plan (results agree to the
                                    order of the FP64 sums; for tests and for measuring the switch-over) */
    int refit;                   /* MCSIM_GMM_REFIT_*: see mcsim_gmm_truncate_params; EM runs the per-waypoint kernels */
    int em_iters;
} mcsim_gmm_run_params;

enum {
    MCSIM_GMM_KERNEL_AUTO = 0,          /* one thread-block cluster for the whole plan up to 98,304 samples per waypoint
                                           (batches of >= 64 runs: one CTA per run), else three kernels per waypoint */
    MCSIM_GMM_KERNEL_PER_WAYPOINT = 1,  /* count / sample / refit launches per waypoint, the whole GPU per waypoint */
    MCSIM_GMM_KERNEL_CLUSTER = 2        /* the cluster kernel at any sample count */
};

typedef struct {
    double P;                    /* 1 - prod_t (1 - p_t)  (MCSimulator.h:848-856) */
    double* probabilities;       /* [T] optional */
    double* gmm_pre;             /* [T x K x 13] optional: w, mu[3], cov[9] before each truncation */
    double* gmm_post;            /* [K x 13] optional: mixture after the last truncation */
    int degenerate;
    float kernel_ms;             /* summed device time of the truncation kernels */
    uint64_t pair_tests;         /* Philox mode: SAT pair tests executed (broad phase as in mcsim_mc_params.exhaustive) */
} mcsim_gmm_run_result;

/* Whole GMM branch of EKF_GaussProp for one plan; plan / noise parameters come from mc (its
 * n_rollouts, z0 and applied fields are ignored): mc->mode selects replay / Philox, mc->seed is
 * the Philox key and mc->rollout_offset the Philox subsequence (a trajectory id). */
int mcsim_cuda_gmm_run(mcsim_cuda_ctx* ctx, const mcsim_mc_params* mc, const mcsim_gmm_run_params* g,
                       mcsim_gmm_run_result* r);

/* Batched form (Philox mode): n_traj independent runGMMEstimation runs in the same launches -- the
 * driver's 200-run loop (MCSimulation.py:226-246) or many candidate plans at once.  One run is a chain
 * of T dependent sample -> refit steps that leaves most of the GPU idle at the reference's N = 10^4;
 * the batch fills it.  Member b is exactly mcsim_cuda_gmm_run with rollout_offset = stream0 + b (and
 * plan b when trajectories / odometry are given). */
typedef struct {
    int K;
    uint64_t n_samples;
    int n_traj;
    const double* trajectories;  /* [n_traj][3][T] or NULL: every member runs mc->trajectory */
    const double* odometry;      /* [n_traj][3][T-1] or NULL */
    uint64_t stream0;
    int kernel;                  /* MCSIM_GMM_KERNEL_* (AUTO: batches of >= 64 runs take the one-CTA-per-run kernel) */
} mcsim_gmm_batch_params;

typedef struct {
    double* P;                   /* [n_traj] */
    double* probabilities;       /* [n_traj][T] optional */
    int degenerate;              /* summed over the batch */
    float kernel_ms;
    uint64_t pair_tests;         /* SAT pair tests executed by the whole batch */
} mcsim_gmm_batch_result;

int mcsim_cuda_gmm_run_batch(mcsim_cuda_ctx* ctx, const mcsim_mc_params* mc, const mcsim_gmm_batch_params* g,
                             mcsim_gmm_batch_result* r);

/* ---- EM fit of full-covariance 3-D mixtures (BASELINE.json config 5) ------------------------ */
/* Replaces arma::gmm_full::learn(X, K, dist, keep_existing, km_iter = 0, em_iter, var_floor)
 * (armadillo-8.400.0/include/armadillo_bits/gmm_full_meat.hpp:764-890): em_iterate :2390-2484,
 * init_constants :980-1046, em_generate_acc :2646-2734 (the data-parallel E-step + sufficient
 * statistics), em_update_params :2542-2639 (reduce + M-step), em_fix_params :2741-2800.
 * n_fits independent fits (one per waypoint buffer) run in the same launches. */
typedef struct {
    int K;                       /* components per fit, 1..MCSIM_MAX_GAUSSIANS */
    int n_fits;                  /* independent fits in this call */
    uint64_t n;                  /* samples per fit held by THIS context (its shard) */
    uint64_t n_total;            /* samples per fit over all shards (hefts = sum r / n_total,
                                    avg_log_p = sum lse / n_total); 0 means n */
    const float* x;              /* HOST [n_fits][3][n] FP32, component-major per fit (x.., y.., theta..);
                                    NULL: reuse the samples already resident from the previous call */
    const double* hefts;         /* [n_fits][K]    initial parameters (seed mode keep_existing) */
    const double* means;         /* [n_fits][K][3] */
    const double* fcovs;         /* [n_fits][K][9] symmetric */
    int iters;                   /* em_iter */
    double var_floor;            /* <= 0: DBL_MIN, as learn() does (:868) */
    uint64_t x_row_stride;       /* floats between the starts of consecutive rows of x; 0 means n (rows are contiguous).
                                    Larger: x points at a column slice of wider rows (a shard of a sharded fit) */
} mcsim_em_params;

typedef struct {
    double* hefts;               /* HOST out, same shapes as the inputs */
    double* means;
    double* fcovs;
    double* avg_log_p;           /* [n_fits][iters] optional: average log-likelihood under the parameters
                                    BEFORE each iteration's update (what em_iterate tests, :2447);
                                    NaN for iterations skipped after convergence */
    int* iters_run;              /* [n_fits] optional */
    float kernel_ms;             /* device time of all E/M launches of the call */
} mcsim_em_result;

/* Upload + iterate + download, single device. */
int mcsim_cuda_em_fit(mcsim_cuda_ctx* ctx, const mcsim_em_params* p, mcsim_em_result* r);
/* Split form for sharded samples (one context per GPU): prepare uploads samples and parameters and
 * computes the E-step constants; every iteration is estep (accumulates this shard's statistics),
 * an all-reduce (sum) by the caller of the FP64 block em_stats() on the context's stream, mstep
 * (identical on every shard); collect downloads the parameters.  All asynchronous on the stream
 * except prepare's and collect's host copies. */
int mcsim_cuda_em_prepare(mcsim_cuda_ctx* ctx, const mcsim_em_params* p);
int mcsim_cuda_em_estep(mcsim_cuda_ctx* ctx);
int mcsim_cuda_em_mstep(mcsim_cuda_ctx* ctx);
int mcsim_cuda_em_collect(mcsim_cuda_ctx* ctx, mcsim_em_result* r);
/* DEVICE pointer and length (doubles) of the statistics block: [n_fits][K * 10 + 1] =
 * per component {sum r, sum r d (3), sum r d d^T (xx xy xz yy yz zz)} with d = x - a reference point
 * (below), then sum of log-likelihoods.  Additive over shards: every shard uses the same reference
 * points because every shard holds the same parameters. */
void* mcsim_cuda_em_stats(mcsim_cuda_ctx* ctx, uint64_t* n_doubles);
/* Reference point of the statistics.  The reference accumulates raw sums in FP64 (em_generate_acc,
 * gmm_full_meat.hpp:2646-2734); the FP32 kernels here accumulate about a point near the data instead:
 *   MCSIM_EM_MOMENTS_CENTRED  the component's own current mean (most exact; 16 FP32 instructions per sample and
 *                             component for the accumulation),
 *   MCSIM_EM_MOMENTS_SHIFTED  one point per fit, the heft-weighted mean of the current means (x - c and its products
 *                             once per sample, 11 instructions per component; the M-step's covariance then cancels
 *                             (|mean - c|^2 + sigma^2) / sigma^2 digits-worth of the FP32 partial sums),
 *   MCSIM_EM_MOMENTS_AUTO     (default) SHIFTED for the fits and iterations whose components lie close enough together for
 *                             that loss to stay below ~1e-5 relative (decided on the device per fit and iteration,
 *                             DESIGN.md section 3), CENTRED otherwise -- and for K <= 5, whose packed-pair kernels
 *                             gain nothing from it.  K = 1 is always CENTRED.
 * Applies to the fits prepared after the call. */
enum { MCSIM_EM_MOMENTS_AUTO = 0, MCSIM_EM_MOMENTS_CENTRED = 1, MCSIM_EM_MOMENTS_SHIFTED = 2 };
int mcsim_cuda_em_set_moments(mcsim_cuda_ctx* ctx, int form);
/* Which of the two the NEXT E-step of fit `fit` will use (the decision the last M-step, or prepare, took):
 * *form = MCSIM_EM_MOMENTS_CENTRED or _SHIFTED.  Synchronises the context's stream. */
int mcsim_cuda_em_moments_in_use(mcsim_cuda_ctx* ctx, int fit, int* form);

/* ---- EM fit that allocates nothing ----------------------------------------------------------- */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same E-step / M-step kernels as mcsim_cuda_em_fit, without a context: samples, parameters and
 * results are the CALLER's device buffers, and the kernels' scratch (E-step constants, the FP64
 * sufficient statistics, convergence state; MCSIM_EM_STATIC_SCRATCH_BYTES) is a file-scope __device__
 * array of the library -- every GPU that loads the library holds it until exit.  The call makes no
 * cudaMalloc / cudaFree / cudaMallocHost and no host heap allocation; it enqueues 1 + 1 + 2 * iters (K <= 16)
 * launches on `stream` and returns without synchronising.  The first call on a device resolves the
 * array's device address (cudaGetSymbolAddress), so make that call outside a capture.
 * Limits (MCSIM_ERR_ARG beyond them): n_fits <= MCSIM_EM_STATIC_MAX_FITS, iters <= MCSIM_EM_STATIC_MAX_ITERS,
 * K <= MCSIM_MAX_GAUSSIANS, n < 2^31 - 2^24.
 * Every call on a device shares the one array: it is re-initialised by the call's first launch, and calls
 * that OVERLAP on the same GPU (two streams, two threads, two processes' worth of this library in one
 * context) race on it -- serialise them. */
#define MCSIM_EM_STATIC_MAX_FITS 8
#define MCSIM_EM_STATIC_MAX_ITERS 64
#define MCSIM_EM_STATIC_SCRATCH_BYTES 32768
typedef struct {
    int device;                  /* CUDA device ordinal; made current by the call */
    int K;                       /* components per fit */
    int n_fits;                  /* independent fits, 1..MCSIM_EM_STATIC_MAX_FITS */
    int iters;                   /* em_iter, 0..MCSIM_EM_STATIC_MAX_ITERS */
    uint64_t n;                  /* samples per fit */
    double var_floor;            /* <= 0: DBL_MIN */
    const float* x;              /* DEVICE in  [n_fits][3][n] FP32, component-major per fit */
    double* hefts;               /* DEVICE in/out [n_fits][K]: initial parameters in, fitted parameters out */
    double* means;               /* DEVICE in/out [n_fits][K][3] */
    double* fcovs;               /* DEVICE in/out [n_fits][K][9] */
    double* avg_log_p;           /* DEVICE out [n_fits][iters], optional (NaN for iterations skipped after convergence) */
    int* iters_run;              /* DEVICE out [n_fits], optional */
    void* stream;                /* cudaStream_t; NULL = the default stream */
} mcsim_em_static_params;
int mcsim_cuda_em_fit_static(const mcsim_em_static_params* p);

/* ---- batched plan scoring (BASELINE.json config 4) ------------------------------------------- */
/* The Philox-mode rollout law of mcsim_cuda_mc_run (EKF_GaussProp, MCSimulator.h:649-864, one
 * real-robot / EKF thread and one particle per rollout) for n_plans plans of T waypoints against the
 * current scene, `rollouts` rollouts per plan; rollout ids are (plan_offset + plan) << 32 | rollout, so
 * the result of a plan does not depend on how plans are split over calls or GPUs. */
typedef struct {
    int T;                       /* waypoints per plan, 2..MCSIM_MAX_WAYPOINTS */
    int n_plans;
    const double* trajectories;  /* [n_plans][3][T]   component-major per plan (setTrajectory layout) */
    const double* odometry;      /* [n_plans][3][T-1] component-major per plan (setOdometry layout) */
    double alphas[4];
    double Q;
    int n_landmarks;             /* must be 8 on this path */
    const double* landmarks;
    double cov0[9];
    uint64_t rollouts;           /* per plan, <= 2^32 */
    uint64_t plan_offset;        /* global id of plan 0 (shards of a batch) */
    int estimator;               /* MCSIM_EST_* */
    int early_exit;
    uint64_t seed;
    double margin;               /* broad phase: largest deviation from the nominal waypoint served by the
                                    candidate lists (8 levels, margin * 2^((k - 7) / 2) metres; a warp tests
                                    the boxes within reach at the level of its furthest rollout -- exact, see
                                    score_sm100.cuh; 3.0 is the measured optimum of BASELINE config 4);
                                    < 0: every live pair at every step */
} mcsim_score_params;

typedef struct {
    uint64_t* hits;              /* [n_plans] optional */
    double* P;                   /* [n_plans] optional: hits / rollouts */
    uint64_t executed_steps;     /* rollout-steps executed */
    uint64_t pair_tests;         /* SAT pair tests executed (nominal = executed_steps x live pairs) */
    float kernel_ms;
} mcsim_score_result;

int mcsim_cuda_score_plans(mcsim_cuda_ctx* ctx, const mcsim_score_params* p, mcsim_score_result* r);
/* Device-resident form: prepare uploads the plan tables, launch scores them (asynchronous, may be
 * repeated), collect downloads hits / counters and zeroes them. */
int mcsim_cuda_score_prepare(mcsim_cuda_ctx* ctx, const mcsim_score_params* p);
int mcsim_cuda_score_launch(mcsim_cuda_ctx* ctx);
int mcsim_cuda_score_collect(mcsim_cuda_ctx* ctx, mcsim_score_result* r);

/* ---- method variant: belief propagation with the analytic linear-constraint truncation ------- */
/* The GMM branch of EKF_GaussProp (MCSimulator.h:672-678, :766-772, :804-817, :843-856) with truncateGMM replaced by the
 * closed form of mcsim_truncate_linear applied per component ON THE DEVICE (one launch per run): the free region is the
 * intersection of the half-planes A[k] . (x, y, theta) <= b[k] -- the constraint list the reference only sketches
 * (gaussprop.py:58-73, linConstraint.py:15-30).  No samples, no collision geometry.  Parity unpinned. */
typedef struct {
    int K;                       /* mixture components (all start as N(x*_0, Sigma_0), as initModel does) */
    int n_constraints;
    const double* A;             /* [n_constraints][3] */
    const double* b;             /* [n_constraints] */
    const double* normals;       /* optional: (T - 1) x (3 + L) standard normals of the shared real-robot thread in
                                    consumption order; NULL: drawn from Philox (mc->seed, stream mc->rollout_offset) */
    size_t n_normals;
} mcsim_linear_params;
int mcsim_cuda_linear_run(mcsim_cuda_ctx* ctx, const mcsim_mc_params* mc, const mcsim_linear_params* lp, mcsim_gmm_run_result* r);

/* ---- analytic linear-constraint truncation (host, FP64) ------------------------ */
/* Truncates N(mean, cov) (3-D, cov row-major) against n half-spaces A[k] . x <= b[k]: per constraint the
 * 1-D marginal along A[k] is truncated and the moment changes are propagated through the conditional;
 * the changes of all constraints are computed from the untruncated belief and summed.  p_violation[k]
 * (optional) = mass of the belief violating constraint k.  The reference holds only a predicate stub
 * for this (linConstraint.py:15-30, gaussprop.py:58-73): parity unpinned. */
int mcsim_truncate_linear(const double mean[3], const double cov[9], int n, const double* A, const double* b,
                          double mean_out[3], double cov_out[9], double* p_violation);

/* ---- several devices behind one handle ---------------------------------------------------------------------------- */
/* What lets the C++ plugin run `setNumParticles 100000000` + `runSimulation` (mcsimplugin.cpp:75-81) on every GPU of a
 * box: one context per listed device and -- for more than one device -- one NCCL communicator per device
 * (ncclCommInitAll, single process), all owned by the group.  libnccl.so.2 is resolved at run time (dlopen; a copy
 * already in the process is shared), so single-device use needs no NCCL at all.
 *   group_mc_run          rollout ids split into contiguous ranges; one ncclAllReduce(sum, uint64) of the counter block per
 *                         run on the devices' streams; the result is the job's: identical for 1, 2, 4, 8 devices
 *   group_em_fit          samples split (x is ONE host array [n_fits][3][n]); one ncclAllReduce(sum, float64) of the
 *                         statistics per iteration, identical M-step on every device
 *   group_score_plans     plans split; no exchange (results land in disjoint slices of the caller's arrays)
 *   group_gmm_run_batch   batch members split (independent replicas); no exchange
 * A group of one device is the plain single-device path. */
typedef struct mcsim_cuda_group mcsim_cuda_group;
int mcsim_cuda_device_count(void);
int mcsim_cuda_group_create(const int* devices, int n_devices, mcsim_cuda_group** out);
void mcsim_cuda_group_destroy(mcsim_cuda_group* g);
int mcsim_cuda_group_size(const mcsim_cuda_group* g);
mcsim_cuda_ctx* mcsim_cuda_group_ctx(mcsim_cuda_group* g, int i);
int mcsim_cuda_group_set_scene(mcsim_cuda_group* g, const double* boxes, int n_boxes, const double* hulls, int n_hulls);
int mcsim_cuda_group_set_scene_ex(mcsim_cuda_group* g, const double* boxes, int n_boxes, const double* hulls, int n_hulls,
                                  const double* poly_hulls, int n_poly);
int mcsim_cuda_group_mc_run(mcsim_cuda_group* g, const mcsim_mc_params* p, mcsim_mc_result* r);
int mcsim_cuda_group_em_fit(mcsim_cuda_group* g, const mcsim_em_params* p, mcsim_em_result* r);
/* How group_em_fit exchanges the statistics every iteration.  PEER: fused into the kernels -- behind its E-step every
 * device pushes its block into slot [rank] of every device's gather buffer with remote stores over NVLink peer memory
 * and raises a counter there; the M-step kernel waits for the G counters and adds the slots in rank order itself
 * (no collective launch between E- and M-step; totals bit-identical on every device).  NCCL: one ncclAllReduce per
 * iteration.  AUTO (default): PEER when every device of the group can map every other one's memory and the group has
 * at most 8 devices, else NCCL.  mcsim_cuda_group_em_exchange returns the mode group_em_fit will use. */
#define MCSIM_EM_EXCHANGE_AUTO 0
#define MCSIM_EM_EXCHANGE_NCCL 1
#define MCSIM_EM_EXCHANGE_PEER 2
int mcsim_cuda_group_set_em_exchange(mcsim_cuda_group* g, int mode);
int mcsim_cuda_group_em_exchange(const mcsim_cuda_group* g);
int mcsim_cuda_group_score_plans(mcsim_cuda_group* g, const mcsim_score_params* p, mcsim_score_result* r);
int mcsim_cuda_group_gmm_run_batch(mcsim_cuda_group* g, const mcsim_mc_params* mc, const mcsim_gmm_batch_params* gp,
                                   mcsim_gmm_batch_result* r);
/* device time of the collective(s) of the last group call that had one (ms, max over devices; for EM the last
 * iteration's) and the number of collectives issued so far */
void mcsim_cuda_group_stats(const mcsim_cuda_group* g, float* collective_ms, uint64_t* collectives);

/* bytes the last *_run call copied host->device and device->host (bench.py's e2e accounting) */
void mcsim_cuda_io_bytes(mcsim_cuda_ctx* ctx, uint64_t* h2d, uint64_t* d2h);

#ifdef __cplusplus
}
#endif
#endif /* MCSIM_CUDA_H */
