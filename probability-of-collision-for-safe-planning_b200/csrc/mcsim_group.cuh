// mcsim_group.cuh -- several devices behind one handle (included at the end of mcsim_cuda.cu: it drives the contexts
// through the same internal halves the single-device entry points use).
//
// What the reference's plugin would need to run `setNumParticles 100000000` + `runSimulation`
// (mcsimplugin/mcsimplugin.cpp:75-81) on every GPU of a box while staying a C++ plugin: one context (stream, tables,
// counters) per device, work split by rollout id / plan / sample / batch member, and one NCCL communicator per device
// owned by the group (ncclCommInitAll: single process, one host thread).  Collectives, where the path has an exchange:
//   MC rollouts   one ncclAllReduce(sum, uint64) of the counter block {hits, near, steps, tests, hist[T]} per run,
//                 queued on each device's stream right behind its rollout kernel
//   EM fit        one ncclAllReduce(sum, float64) of the statistics block [n_fits][K 10 + 1] per iteration
//                 (gmm_full's own map-reduce shape, gmm_full_meat.hpp:2509-2548)
//   scoring, batched GMM runs   none: plans / batch members are independent, results land in disjoint host slices
#ifndef MCSIM_GROUP_CUH
#define MCSIM_GROUP_CUH

#include "nccl_dl.h"

struct mcsim_cuda_group {
    std::vector<mcsim_cuda_ctx*> ctx;
    std::vector<ncclComm_t> comm;      // empty for a single device
    std::vector<cudaEvent_t> ev0, ev1; // around the collectives of the last call, per device
    float last_collective_ms = 0.f;
    uint64_t collectives = 0;
    bool peer_access = false;          // every device of the group can map every other one's memory (NVLink / NVSwitch)
    int em_exchange = MCSIM_EM_EXCHANGE_AUTO;
    unsigned long long em_epoch = 0;   // pushes completed by every device so far (the flags count them)
};

#ifndef GROUP_EM_THREADS_MIN_DEVICES
#define GROUP_EM_THREADS_MIN_DEVICES 4
#endif
#define NCK(call)                                                                              \
    do {                                                                                       \
        ncclResult_t r_ = (call);                                                              \
        if (r_ != ncclSuccess) return fail(MCSIM_ERR_NCCL, "%s: %s", #call, nccl_api().GetErrorString(r_)); \
    } while (0)

extern "C" int mcsim_cuda_device_count(void) {
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

extern "C" void mcsim_cuda_group_destroy(mcsim_cuda_group* g) {
    if (!g) return;
    for (size_t d = 0; d < g->comm.size(); ++d) if (g->comm[d]) nccl_api().CommDestroy(g->comm[d]);
    for (size_t d = 0; d < g->ctx.size(); ++d) {
        if (g->ctx[d]) cudaSetDevice(g->ctx[d]->device);
        if (d < g->ev0.size()) { cudaEventDestroy(g->ev0[d]); cudaEventDestroy(g->ev1[d]); }
        mcsim_cuda_destroy(g->ctx[d]);
    }
    delete g;
}

extern "C" int mcsim_cuda_group_create(const int* devices, int n_devices, mcsim_cuda_group** out) {
    if (!out || !devices || n_devices < 1 || n_devices > 64) return fail(MCSIM_ERR_ARG, "group_create: bad arguments");
    *out = nullptr;
    for (int a = 0; a < n_devices; ++a)
        for (int b = a + 1; b < n_devices; ++b)
            if (devices[a] == devices[b]) return fail(MCSIM_ERR_ARG, "group_create: device %d listed twice", devices[a]);
    mcsim_cuda_group* g = new mcsim_cuda_group();
    for (int d = 0; d < n_devices; ++d) {
        mcsim_cuda_ctx* c = nullptr;
        const int rc = mcsim_cuda_create(devices[d], &c);
        if (rc) { mcsim_cuda_group_destroy(g); return rc; }
        g->ctx.push_back(c);
        cudaEvent_t a, b;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { mcsim_cuda_group_destroy(g); return fail(MCSIM_ERR_CUDA, "group_create: cudaEventCreate failed"); }
        g->ev0.push_back(a); g->ev1.push_back(b);
    }
    if (n_devices > 1) {
        const NcclApi& api = nccl_api();
        if (!api.ok()) { mcsim_cuda_group_destroy(g); return fail(MCSIM_ERR_NCCL, "group_create: %s", api.error.c_str()); }
        g->comm.assign(n_devices, nullptr);
        const ncclResult_t r = api.CommInitAll(g->comm.data(), n_devices, devices);
        if (r != ncclSuccess) { g->comm.clear(); mcsim_cuda_group_destroy(g); return fail(MCSIM_ERR_NCCL, "ncclCommInitAll: %s", api.GetErrorString(r)); }
    }
    if (n_devices > 1 && n_devices <= EM_MAX_PEERS) {
        // peer mappings for the fused EM exchange; without them (PCIe-only boxes, MIG) the group keeps to NCCL
        bool all = true;
        for (int a = 0; a < n_devices && all; ++a)
            for (int b = 0; b < n_devices && all; ++b) {
                int can = 0;
                if (a != b && (cudaDeviceCanAccessPeer(&can, devices[a], devices[b]) != cudaSuccess || !can)) all = false;
            }
        for (int a = 0; a < n_devices && all; ++a) {
            cudaSetDevice(devices[a]);
            for (int b = 0; b < n_devices && all; ++b) {
                if (a == b) continue;
                const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) { cudaGetLastError(); all = false; }
            }
        }
        g->peer_access = all;
    }
    *out = g;
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_group_set_em_exchange(mcsim_cuda_group* g, int mode) {
    if (!g || (mode != MCSIM_EM_EXCHANGE_AUTO && mode != MCSIM_EM_EXCHANGE_NCCL && mode != MCSIM_EM_EXCHANGE_PEER))
        return fail(MCSIM_ERR_ARG, "group_set_em_exchange: bad group or mode");
    if (mode == MCSIM_EM_EXCHANGE_PEER && !g->peer_access && g->ctx.size() > 1)
        return fail(MCSIM_ERR_ARG, "group_set_em_exchange: the devices of this group cannot map each other's memory (or more than %d devices)", EM_MAX_PEERS);
    g->em_exchange = mode;
    return MCSIM_OK;
}
extern "C" int mcsim_cuda_group_em_exchange(const mcsim_cuda_group* g) {
    if (!g || g->ctx.size() < 2) return MCSIM_EM_EXCHANGE_NCCL;
    if (g->em_exchange == MCSIM_EM_EXCHANGE_AUTO) return g->peer_access ? MCSIM_EM_EXCHANGE_PEER : MCSIM_EM_EXCHANGE_NCCL;
    return g->em_exchange;
}

extern "C" int mcsim_cuda_group_size(const mcsim_cuda_group* g) { return g ? int(g->ctx.size()) : 0; }
extern "C" mcsim_cuda_ctx* mcsim_cuda_group_ctx(mcsim_cuda_group* g, int i) { return (g && i >= 0 && i < int(g->ctx.size())) ? g->ctx[i] : nullptr; }
extern "C" void mcsim_cuda_group_stats(const mcsim_cuda_group* g, float* collective_ms, uint64_t* collectives) {
    if (collective_ms) *collective_ms = g ? g->last_collective_ms : 0.f;
    if (collectives) *collectives = g ? g->collectives : 0;
}

extern "C" int mcsim_cuda_group_set_scene_ex(mcsim_cuda_group* g, const double* boxes, int n_boxes, const double* hulls, int n_hulls,
                                             const double* poly_hulls, int n_poly) {
    if (!g) return fail(MCSIM_ERR_ARG, "group_set_scene: NULL group");
    for (mcsim_cuda_ctx* c : g->ctx) { const int rc = mcsim_cuda_set_scene_ex(c, boxes, n_boxes, hulls, n_hulls, poly_hulls, n_poly); if (rc) return rc; }
    return MCSIM_OK;
}
extern "C" int mcsim_cuda_group_set_scene(mcsim_cuda_group* g, const double* boxes, int n_boxes, const double* hulls, int n_hulls) {
    return mcsim_cuda_group_set_scene_ex(g, boxes, n_boxes, hulls, n_hulls, nullptr, 0);
}

// sum-all-reduce of one buffer per device, in place, on each device's stream (single thread: NCCL group call)
static int group_all_reduce(mcsim_cuda_group* g, void* const* bufs, size_t count, ncclDataType_t type, bool timed) {
    const NcclApi& api = nccl_api();
    const int G = int(g->ctx.size());
    if (timed) for (int d = 0; d < G; ++d) { CK(cudaSetDevice(g->ctx[d]->device)); CK(cudaEventRecord(g->ev0[d], g->ctx[d]->stream)); }
    NCK(api.GroupStart());
    for (int d = 0; d < G; ++d) {
        CK(cudaSetDevice(g->ctx[d]->device));
        NCK(api.AllReduce(bufs[d], bufs[d], count, type, ncclSum, g->comm[d], g->ctx[d]->stream));
    }
    NCK(api.GroupEnd());
    if (timed) for (int d = 0; d < G; ++d) { CK(cudaSetDevice(g->ctx[d]->device)); CK(cudaEventRecord(g->ev1[d], g->ctx[d]->stream)); }
    ++g->collectives;
    return MCSIM_OK;
}
static int group_collective_ms(mcsim_cuda_group* g) {   // after the streams were synchronised
    g->last_collective_ms = 0.f;
    for (size_t d = 0; d < g->ctx.size(); ++d) {
        float t = 0.f;
        CK(cudaSetDevice(g->ctx[d]->device));
        CK(cudaEventElapsedTime(&t, g->ev0[d], g->ev1[d]));
        if (t > g->last_collective_ms) g->last_collective_ms = t;
    }
    return MCSIM_OK;
}

// ---- MC rollouts: contiguous rollout-id ranges per device; Philox counters come from the global id, so the totals do
// not depend on the number of devices (BASELINE config 3)
extern "C" int mcsim_cuda_group_mc_run(mcsim_cuda_group* g, const mcsim_mc_params* p, mcsim_mc_result* r) {
    if (!g || !p || !r) return fail(MCSIM_ERR_ARG, "group_mc_run: NULL argument");
    const int G = int(g->ctx.size());
    if (G == 1) return mcsim_cuda_mc_run(g->ctx[0], p, r);
    if (r->poses || r->clearance) return fail(MCSIM_ERR_ARG, "group_mc_run: audit outputs (poses / clearance) are single-device");
    const uint64_t n = p->n_rollouts;
    std::vector<mcsim_mc_params> ps(G, *p);
    std::vector<mcsim_mc_result> rs(G);
    std::vector<void*> bufs(G);
    for (int d = 0; d < G; ++d) {
        const uint64_t lo = uint64_t((unsigned __int128)n * d / G), hi = uint64_t((unsigned __int128)n * (d + 1) / G);
        ps[d].n_rollouts = hi - lo;
        ps[d].rollout_offset = p->rollout_offset + lo;
        if (p->z0) ps[d].z0 = p->z0 + 3 * lo;
        memset(&rs[d], 0, sizeof(mcsim_mc_result));
        if (r->counts) rs[d].counts = r->counts + lo;
        if (d == 0) rs[d].first_hit_hist = r->first_hit_hist;
        CK(cudaSetDevice(g->ctx[d]->device));
        const int rc = mc_run_begin(g->ctx[d], &ps[d], &rs[d]);
        if (rc) return rc;
        bufs[d] = g->ctx[d]->counters.p;
    }
    int rc = group_all_reduce(g, bufs.data(), size_t(4 + p->T), ncclUint64, true);
    if (rc) return rc;
    float ms = 0.f;
    for (int d = 0; d < G; ++d) {
        CK(cudaSetDevice(g->ctx[d]->device));
        rc = mc_run_end(g->ctx[d], &rs[d]);   // every device now reads the job-wide totals
        if (rc) return rc;
        if (rs[d].kernel_ms > ms) ms = rs[d].kernel_ms;
    }
    r->hits = rs[0].hits; r->near_contact = rs[0].near_contact; r->executed_steps = rs[0].executed_steps; r->pair_tests = rs[0].pair_tests;
    r->kernel_ms = ms;
    return group_collective_ms(g);
}

// ---- batched plan scoring: plans split over the devices; plan ids are global, no exchange
extern "C" int mcsim_cuda_group_score_plans(mcsim_cuda_group* g, const mcsim_score_params* p, mcsim_score_result* r) {
    if (!g || !p || !r) return fail(MCSIM_ERR_ARG, "group_score_plans: NULL argument");
    const int G = int(g->ctx.size());
    if (G == 1) return mcsim_cuda_score_plans(g->ctx[0], p, r);
    const int NP = p->n_plans, T = p->T;
    if (NP < 1 || T < 2 || !p->trajectories || !p->odometry) return fail(MCSIM_ERR_ARG, "group_score_plans: no plans");
    std::vector<int> lo(G + 1);
    for (int d = 0; d <= G; ++d) lo[d] = int((long long)NP * d / G);
    for (int d = 0; d < G; ++d) {
        if (lo[d + 1] == lo[d]) continue;
        mcsim_score_params pd = *p;
        pd.n_plans = lo[d + 1] - lo[d];
        pd.trajectories = p->trajectories + size_t(lo[d]) * 3 * T;
        pd.odometry = p->odometry + size_t(lo[d]) * 3 * (T - 1);
        pd.plan_offset = p->plan_offset + uint64_t(lo[d]);
        mcsim_cuda_ctx* c = g->ctx[d];
        CK(cudaSetDevice(c->device));
        int rc = mcsim_cuda_score_prepare(c, &pd);
        if (rc) return rc;
        CK(cudaEventRecord(c->ev0, c->stream));
        if ((rc = mcsim_cuda_score_launch(c))) return rc;
        CK(cudaEventRecord(c->ev1, c->stream));
    }
    r->executed_steps = r->pair_tests = 0;
    r->kernel_ms = 0.f;
    for (int d = 0; d < G; ++d) {
        if (lo[d + 1] == lo[d]) continue;
        mcsim_cuda_ctx* c = g->ctx[d];
        CK(cudaSetDevice(c->device));
        mcsim_score_result rd;
        memset(&rd, 0, sizeof(rd));
        rd.hits = r->hits ? r->hits + lo[d] : nullptr;
        rd.P = r->P ? r->P + lo[d] : nullptr;
        const int rc = mcsim_cuda_score_collect(c, &rd);
        if (rc) return rc;
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, c->ev0, c->ev1));
        if (t > r->kernel_ms) r->kernel_ms = t;
        r->executed_steps += rd.executed_steps;
        r->pair_tests += rd.pair_tests;
    }
    return MCSIM_OK;
}

// One device's share of a fused sharded fit: E-step, then the M-step kernel that pushes this device's block to every
// device, waits for theirs and adds them up itself (em_sm100.cuh).  Two launches per iteration, as on one device.
static int group_em_fused_iterations(mcsim_cuda_group* g, int d, int G, int it0, int it1, int iters, uint64_t n_stats, unsigned long long epoch0) {
    mcsim_cuda_ctx* c = g->ctx[d];
    CK(cudaSetDevice(c->device));
    c->em.n_peers = G;
    c->em.my_rank = d;
    if (it0 == 0) CK(cudaEventRecord(c->ev0, c->stream));   // kernel_ms: this device's launches, from the first E-step on
    for (int it = it0; it < it1; ++it) {
        const unsigned long long epoch = epoch0 + (unsigned long long)it;
        const size_t par = size_t(epoch & 1) * size_t(G) * n_stats;
        int rc = mcsim_cuda_em_estep(c);
        if (rc) return rc;
        for (int q = 0; q < G; ++q) {
            c->em.peer_slots[q] = (double*)g->ctx[q]->em_slots.p + par;
            c->em.peer_flags[q] = (unsigned long long*)g->ctx[q]->em_flags.p;
        }
        c->em.my_slots = c->em.peer_slots[d];
        c->em.my_flags = c->em.peer_flags[d];
        c->em.wait_value = epoch + 1;
        if (it == iters - 1) CK(cudaEventRecord(g->ev0[d], c->stream));
        if ((rc = mcsim_cuda_em_mstep(c))) return rc;
        if (it == iters - 1) CK(cudaEventRecord(g->ev1[d], c->stream));
    }
    return MCSIM_OK;
}

// ---- EM fit: samples split over the devices (multiples of 4 per shard, so the 128-bit / bulk-copy kernels stay usable),
// statistics all-reduced every iteration, identical M-step everywhere
static mcsim_em_params group_em_shard(const mcsim_em_params* p, int d, int G) {
    const uint64_t n = p->n;
    const uint64_t lo = uint64_t((unsigned __int128)n * d / G) & ~3ull, hi = d + 1 == G ? n : (uint64_t((unsigned __int128)n * (d + 1) / G) & ~3ull);
    mcsim_em_params pd = *p;
    pd.n = hi - lo;
    pd.n_total = p->n_total ? p->n_total : n;
    if (p->x) { pd.x = p->x + lo; pd.x_row_stride = p->x_row_stride ? p->x_row_stride : n; }
    return pd;
}

// after the iterations of a fused fit: record the end, bring the parameters back (device 0's; they are identical everywhere)
// and look at the done flags -- a peer's block that never arrived ends the fit there with done = EM_DONE_EXCHANGE_TIMEOUT
#define GROUP_EM_DONE_CHECKED 1024   // flags looked at (ints in the context's pinned staging buffer); a stuck peer stops every fit
static int group_em_fused_finish(mcsim_cuda_group* g, int d, const mcsim_em_params* p, mcsim_em_result* r) {
    mcsim_cuda_ctx* c = g->ctx[d];
    CK(cudaSetDevice(c->device));
    CK(cudaEventRecord(c->ev1, c->stream));
    const int nd = p->n_fits < GROUP_EM_DONE_CHECKED ? p->n_fits : GROUP_EM_DONE_CHECKED;
    int* dn = reinterpret_cast<int*>(c->h_gmm);
    CK(cudaMemcpyAsync(dn, c->em_done.p, sizeof(int) * nd, cudaMemcpyDeviceToHost, c->stream));
    mcsim_em_result none;
    memset(&none, 0, sizeof(none));
    const int rc = mcsim_cuda_em_collect(c, d == 0 ? r : &none);   // synchronises the stream
    if (rc) return rc;
    for (int f = 0; f < nd; ++f)
        if (dn[f] == EM_DONE_EXCHANGE_TIMEOUT)
            return fail(MCSIM_ERR_CUDA, "group_em_fit: device %d waited for a peer's statistics for more than %lld cycles (fit %d)", c->device, (long long)EM_PEER_TIMEOUT_CYCLES, f);
    return MCSIM_OK;
}

extern "C" int mcsim_cuda_group_em_fit(mcsim_cuda_group* g, const mcsim_em_params* p, mcsim_em_result* r) {
    if (!g || !p || !r) return fail(MCSIM_ERR_ARG, "group_em_fit: NULL argument");
    const int G = int(g->ctx.size());
    if (G == 1) return mcsim_cuda_em_fit(g->ctx[0], p, r);
    const bool fused = mcsim_cuda_group_em_exchange(g) == MCSIM_EM_EXCHANGE_PEER && p->iters > 0;
    const int iters = p->iters;
    r->kernel_ms = 0.f;
    if (fused) {
        if (p->K < 1 || p->K > MCSIM_MAX_GAUSSIANS || p->n_fits < 1) return fail(MCSIM_ERR_ARG, "group_em_fit: K=%d / n_fits=%d out of range", p->K, p->n_fits);
        const uint64_t n_stats = uint64_t(p->n_fits) * (p->K * EM_NSTAT + 1);
        // gather buffers [2 parities][G slots][n_stats] and arrival counters [n_fits][G] on every device, before any device
        // starts (the counters hold the number of the last exchange that arrived: g->em_epoch + 1 and up)
        for (int d = 0; d < G; ++d) {
            mcsim_cuda_ctx* c = g->ctx[d];
            CK(cudaSetDevice(c->device));
            CK(c->em_slots.reserve(sizeof(double) * 2 * size_t(G) * n_stats));
            const size_t fl = sizeof(unsigned long long) * size_t(p->n_fits) * G;
            if (c->em_flags.cap < fl) {   // (re)allocated: counters restart below every value a call will wait for
                CK(c->em_flags.reserve(fl));
                CK(cudaMemsetAsync(c->em_flags.p, 0, c->em_flags.cap, c->stream));
                CK(cudaStreamSynchronize(c->stream));
            }
        }
        // Nothing on the host links the devices any more (the M-step kernels wait for each other on the GPU).  Small groups
        // are queued by this thread, device by device and iteration by iteration (2 launches each); from
        // GROUP_EM_THREADS_MIN_DEVICES devices on, that queueing rate is the bottleneck (8 devices: 86 us per iteration
        // against 47 us), so every device gets a host thread that uploads, queues all iterations and downloads (~60 us to
        // start and join the threads, once per call)
        const unsigned long long epoch0 = g->em_epoch;
        std::vector<int> rcs(G, 0);
        if (G < GROUP_EM_THREADS_MIN_DEVICES || iters < 2) {
            for (int d = 0; d < G; ++d) {
                const mcsim_em_params pd = group_em_shard(p, d, G);
                CK(cudaSetDevice(g->ctx[d]->device));
                const int rc = mcsim_cuda_em_prepare(g->ctx[d], &pd);
                if (rc) return rc;
            }
            for (int it = 0; it < iters; ++it)
                for (int d = 0; d < G; ++d) { const int rc = group_em_fused_iterations(g, d, G, it, it + 1, iters, n_stats, epoch0); if (rc) return rc; }
            for (int d = 0; d < G; ++d) { const int rc = group_em_fused_finish(g, d, p, r); if (rc) return rc; }
        } else {
            std::vector<std::string> errs(G);
            std::vector<std::thread> th;
            for (int d = 0; d < G; ++d)
                th.emplace_back([g, d, G, iters, n_stats, epoch0, p, r, &rcs, &errs]() {
                    const mcsim_em_params pd = group_em_shard(p, d, G);
                    int rc = cudaSetDevice(g->ctx[d]->device) == cudaSuccess ? MCSIM_OK : fail(MCSIM_ERR_CUDA, "cudaSetDevice failed");
                    if (!rc) rc = mcsim_cuda_em_prepare(g->ctx[d], &pd);
                    if (!rc) rc = group_em_fused_iterations(g, d, G, 0, iters, iters, n_stats, epoch0);
                    if (!rc) rc = group_em_fused_finish(g, d, p, r);
                    rcs[d] = rc;
                    if (rc) errs[d] = mcsim_cuda_last_error();
                });
            for (auto& t : th) t.join();
            for (int d = 0; d < G; ++d) if (rcs[d]) return fail(rcs[d], "device %d: %s", g->ctx[d]->device, errs[d].c_str());
        }
        g->em_epoch += (unsigned long long)iters;
        g->collectives += (uint64_t)iters;   // exchanges, whichever way they travel
        for (int d = 0; d < G; ++d) {
            float t = 0.f;
            CK(cudaSetDevice(g->ctx[d]->device));
            CK(cudaEventElapsedTime(&t, g->ctx[d]->ev0, g->ctx[d]->ev1));
            if (t > r->kernel_ms) r->kernel_ms = t;
        }
        return group_collective_ms(g);
    }
    std::vector<void*> bufs(G);
    uint64_t n_stats = 0;
    for (int d = 0; d < G; ++d) {
        const mcsim_em_params pd = group_em_shard(p, d, G);
        mcsim_cuda_ctx* c = g->ctx[d];
        CK(cudaSetDevice(c->device));
        const int rc = mcsim_cuda_em_prepare(c, &pd);
        if (rc) return rc;
        bufs[d] = mcsim_cuda_em_stats(c, &n_stats);
        CK(cudaEventRecord(c->ev0, c->stream));
    }
    for (int it = 0; it < iters; ++it) {
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(g->ctx[d]->device)); const int rc = mcsim_cuda_em_estep(g->ctx[d]); if (rc) return rc; }
        const int rc = group_all_reduce(g, bufs.data(), size_t(n_stats), ncclFloat64, it == iters - 1);
        if (rc) return rc;
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(g->ctx[d]->device)); const int rm = mcsim_cuda_em_mstep(g->ctx[d]); if (rm) return rm; }
    }
    for (int d = 0; d < G; ++d) {
        mcsim_cuda_ctx* c = g->ctx[d];
        CK(cudaSetDevice(c->device));
        CK(cudaEventRecord(c->ev1, c->stream));
        mcsim_em_result none;
        memset(&none, 0, sizeof(none));
        const int rc = mcsim_cuda_em_collect(c, d == 0 ? r : &none);   // identical parameters on every device
        if (rc) return rc;
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, c->ev0, c->ev1));
        if (t > r->kernel_ms) r->kernel_ms = t;
    }
    return iters > 0 ? group_collective_ms(g) : MCSIM_OK;
}

// ---- batched GMM runs: independent members (replicas); one host thread per device, each call is synchronous
extern "C" int mcsim_cuda_group_gmm_run_batch(mcsim_cuda_group* g, const mcsim_mc_params* mc, const mcsim_gmm_batch_params* gp, mcsim_gmm_batch_result* r) {
    if (!g || !mc || !gp || !r || !r->P) return fail(MCSIM_ERR_ARG, "group_gmm_run_batch: NULL argument");
    const int G = int(g->ctx.size());
    if (G == 1) return mcsim_cuda_gmm_run_batch(g->ctx[0], mc, gp, r);
    const int B = gp->n_traj, T = mc->T;
    if (B < 1) return fail(MCSIM_ERR_ARG, "group_gmm_run_batch: no members");
    std::vector<int> rcs(G, 0);
    std::vector<std::string> errs(G);
    std::vector<mcsim_gmm_batch_result> rs(G);
    std::vector<std::thread> th;
    for (int d = 0; d < G; ++d) {
        const int lo = int((long long)B * d / G), hi = int((long long)B * (d + 1) / G);
        memset(&rs[d], 0, sizeof(mcsim_gmm_batch_result));
        if (hi == lo) continue;
        th.emplace_back([=, &rcs, &errs, &rs]() {
            mcsim_gmm_batch_params pd = *gp;
            pd.n_traj = hi - lo;
            pd.stream0 = gp->stream0 + uint64_t(lo);
            if (gp->trajectories) { pd.trajectories = gp->trajectories + size_t(lo) * 3 * T; pd.odometry = gp->odometry + size_t(lo) * 3 * (T > 1 ? T - 1 : 0); }
            rs[d].P = r->P + lo;
            rs[d].probabilities = r->probabilities ? r->probabilities + size_t(lo) * T : nullptr;
            rcs[d] = mcsim_cuda_gmm_run_batch(g->ctx[d], mc, &pd, &rs[d]);
            if (rcs[d]) errs[d] = mcsim_cuda_last_error();
        });
    }
    for (auto& t : th) t.join();
    r->degenerate = 0; r->pair_tests = 0; r->kernel_ms = 0.f;
    for (int d = 0; d < G; ++d) {
        if (rcs[d]) return fail(rcs[d], "device %d: %s", g->ctx[d]->device, errs[d].c_str());
        r->degenerate += rs[d].degenerate;
        r->pair_tests += rs[d].pair_tests;
        if (rs[d].kernel_ms > r->kernel_ms) r->kernel_ms = rs[d].kernel_ms;
    }
    return MCSIM_OK;
}

#endif
