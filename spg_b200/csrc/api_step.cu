// C-ABI, part 2: state transfer, the VBD / XPBD step drivers (launch sequences, CUDA-graph capture),
// timings and the per-element evaluators.
#include <cstdlib>
#include <cstring>

#include "energies.cuh"
#include "energy_prepare.h"
#include "group.h"

using namespace spgx;

namespace
{
void requireFinalized(spgx_group *grp, bool vbd)
{
    if (!grp || !grp->finalized) throw StateError("group is not finalized (call spgx_group_finalize first)");
    if (vbd && grp->solverKind != SPGX_VBD) throw StateError("group was finalized for XPBD, not VBD");
    if (!vbd && grp->solverKind == SPGX_VBD) throw StateError("group was finalized for VBD, not XPBD");
    if (!vbd && grp->solverKind == SPGX_NEWTON) throw StateError("group was finalized for the implicit-Euler family, not XPBD");
}

unsigned kindMaskOf(const spgx_group *grp)
{
    unsigned m = 0;
    for (const auto &e : grp->energies) m |= 1u << e.kind;
    return m;
}

// the launch sequence of ONE VBD::step() (vbd.cpp:46-183)
void enqueueVbdFrame(spgx_group *grp, double dtStep, int nsub, int iters, cudaStream_t st, bool timePhases)
{
    const double dt = dtStep / nsub;
    const double invdt = 1. / dt;
    const double invdtSquared = invdt * invdt;
    VbdArgs args{};
    args.x = grp->d_x.ptr;
    args.xInertial = grp->d_xInertial.ptr;
    args.mass = grp->d_mass.ptr;
    args.nEnergies = static_cast<int>(grp->energies.size());
    for (int e = 0; e < args.nEnergies; ++e) {
        const auto &en = grp->energies[e];
        args.e[e] = {en.kind, en.n, en.d_stencils.ptr, en.d_records.ptr, en.d_rowptr.ptr, en.d_entries.ptr,
                     en.streamed ? 1 : 0};
    }
    args.nStreams = 0;
    for (int t = 0; t < kNumStreamTypes; ++t) {
        args.s[t] = {nullptr, nullptr};
        if (t < static_cast<int>(grp->streams.size()) && grp->streams[t].d_idx.ptr) {
            args.s[t] = {grp->streams[t].d_idx.ptr, grp->streams[t].d_params.ptr};
            args.nStreams++;
        }
    }
    args.sliceDesc = grp->d_sliceDesc.ptr;
    args.nbrIds = grp->d_nbrIds.ptr;
    args.maxNbr = grp->maxNbr;
    args.nbrFixed = grp->vbdNbrFixed;
    args.nSlices = grp->nSlices;
    args.lookahead = std::min(grp->vbdLookahead, std::max(grp->nSlices - 1, 0));
    args.pdl = grp->vbdPdl;
    args.carveout = grp->vbdCarveout;
    args.l2PersistBytes = 0;
    args.l2HitRatio = 1.0f;
    if (grp->vbdL2Persist > 0 && args.nStreams > 0) {
        // persisting L2 window over the positions (the start of the state arena).  Widening it to the inertial positions and
        // masses behind them measured slower (2.77 vs 2.71 ms per frame at config 3, profiles/r2i_sweep.txt: the set-aside
        // is taken from the incidence streams' share of L2)
        const size_t bytes = 3 * static_cast<size_t>(grp->nLocal) * sizeof(double);
        const size_t setAside = static_cast<size_t>(grp->vbdL2Persist);
        args.l2PersistBytes = std::min(bytes, static_cast<size_t>(grp->ctx->prop.accessPolicyMaxWindowSize));
        args.l2HitRatio = bytes <= setAside ? 1.0f : static_cast<float>(static_cast<double>(setAside) / static_cast<double>(bytes));
    }
    const unsigned mask = kindMaskOf(grp);
    const int nColors = static_cast<int>(grp->colorOffsets.size()) - 1;
    // halo exchange of a partitioned group: fused peer-memory stores + event counters (halo.cuh), or ncclSend / ncclRecv
    const bool p2p = grp->partitioned && grp->p2p.mode == SPGX_HALO_P2P && grp->p2p.connected && args.nStreams > 0;
    // SPGX_DEBUG_SKIP_HALO=1 (timing experiments only, results are wrong): run the partitioned sweep with no exchange
    static const bool skipHalo = getenv("SPGX_DEBUG_SKIP_HALO") && atoi(getenv("SPGX_DEBUG_SKIP_HALO")) != 0;
    const bool viaNccl = grp->partitioned && !p2p && !skipHalo;
    args.halo = (p2p && !skipHalo) ? grp->p2p.dev : VbdHaloDev{};
    const bool signalling = p2p && !skipHalo && args.halo.nPeers > 0;
    for (int s = 0; s < nsub; ++s) {
        // predict / finalize run over owned AND halo vertices: a halo copy goes through the same arithmetic as on its
        // owner, so no exchange is needed outside the colour passes
        launchPredict(grp->d_x.ptr, grp->d_v.ptr, grp->d_xOld.ptr, grp->d_xInertial.ptr, grp->nLocal, dt, grp->graphKey.g,
                      st);
        grp->launchCounter++;
        if (signalling && !(args.halo.debug & 8)) {
            // event 0: this rank's halo copies have been predicted; neighbours may store into them from now on
            launchHaloSignal(args.halo, 0, st);
            grp->launchCounter++;
        }
        if (timePhases) grp->markPhase(st);
        int event = 0;
        for (int it = 0; it < iters; ++it)
            for (int c = 0; c < nColors; ++c) {
                const int b = static_cast<int>(grp->colorOffsets[c]), e = static_cast<int>(grp->colorOffsets[c + 1]);
                ++event;
                args.halo.event = event;
                args.halo.pushWarps = signalling ? grp->p2p.pushWarps[c] : 0;
                if (e > b) {
                    launchVbdPass(args, mask, grp->streamMask, grp->vbdLanes, grp->vbdOcc, b, e, grp->sliceBase[c], invdtSquared, st);
                    grp->launchCounter++;
                } else if (signalling) {
                    launchHaloSignal(args.halo, event, st);  // no vertex of this colour here: the event is still owed
                    grp->launchCounter++;
                }
                // every rank takes part in the exchange of colour c, whether or not it owns vertices of that colour
                if (viaNccl) grp->launchCounter += enqueueHaloExchange(grp, c, st);
            }
        if (signalling && !(args.halo.debug & 8)) {
            // the neighbours' stores of the last pass must have landed before the halo copies are finalized; the epoch
            // advances by the events of this substep
            launchHaloWaitAdvance(args.halo, event, grp->p2p.d_sync.ptr + kMaxPeers, st);
            grp->launchCounter++;
        }
        if (timePhases) grp->markPhase(st);
        launchFinalize(grp->d_x.ptr, grp->d_xOld.ptr, grp->d_v.ptr, grp->nLocal, invdt, st);
        grp->launchCounter++;
        if (timePhases) grp->markPhase(st);
    }
}

// the launch sequence of ONE XPBD::step() (xpbd.cpp:44-123)
void enqueueXpbdFrame(spgx_group *grp, double dtStep, int nsub, int iters, cudaStream_t st, bool timePhases)
{
    const double dt = dtStep / nsub;
    const double invdt = 1 / dt;
    const bool serial = grp->solverKind == SPGX_XPBD_SERIAL;
    const bool useLambda = iters > 1;
    // domain-decomposed group (xpbd_partition.h): one one-way exchange per (energy, colour) pass -- peer-memory stores +
    // event counters (the pass kernel waits for the neighbours' previous event, a push kernel stores what the pass moved and
    // raises this one), or pack + ncclSend / ncclRecv + unpack
    const bool p2p = grp->partitioned && grp->p2p.mode == SPGX_HALO_P2P && grp->p2p.connected;
    const bool viaNccl = grp->partitioned && !p2p && !grp->peers.empty();
    VbdHaloDev halo = p2p ? grp->p2p.dev : VbdHaloDev{};
    const bool signalling = p2p && halo.nPeers > 0;
    // The velocity update of a substep and the prediction of the next are one launch (plain: on the programmatic-dependent-
    // launch chain of the passes it measured SLOWER, 1.63 vs 1.375 ms per frame at config 2, and so did predict / finalize as
    // separate launches on that chain); the phase timer keeps them apart.  SPGX_XPBD_FUSE_STATE=0: two launches.
    const bool fuseStateEnv = !(getenv("SPGX_XPBD_FUSE_STATE") && atoi(getenv("SPGX_XPBD_FUSE_STATE")) == 0);  // read when the graph is captured
    const bool fuseState = fuseStateEnv && !timePhases;
    for (int s = 0; s < nsub; ++s) {
        if (s == 0 || !fuseState) {
            launchPredict(grp->d_x.ptr, grp->d_v.ptr, grp->d_xOld.ptr, nullptr, grp->nLocal, dt, grp->graphKey.g, st);
            grp->launchCounter++;
        }
        if (timePhases) grp->markPhase(st);
        if (signalling) {
            // event 0: this rank's halo copies have been predicted; neighbours may store into them from now on
            launchHaloSignal(halo, 0, st);
            grp->launchCounter++;
        }
        if (useLambda)
            for (auto &en : grp->energies)
                if (en.d_lambda.ptr) SPGX_CUDA(cudaMemsetAsync(en.d_lambda.ptr, 0, en.d_lambda.count * sizeof(double), st));
        int event = 0;
        for (int it = 0; it < iters; ++it) {
            int pass = 0;  // energy-major, colour-minor over the GLOBAL colours (xpbd_partition.h)
            for (auto &en : grp->energies) {
                const int nColors = static_cast<int>(en.colorOffsets.size()) - 1;
                for (int c = 0; c < nColors; ++c, ++pass) {
                    const int b = static_cast<int>(en.colorOffsets[c]), e = static_cast<int>(en.colorOffsets[c + 1]);
                    ++event;
                    halo.event = event;
                    if (e > b) {
                        launchXpbdPass(en.kind, grp->d_x.ptr, en.d_stencilW.ptr, en.d_stencils.ptr, en.d_records.ptr,
                                       useLambda ? en.d_lambda.ptr : nullptr, b, e, dt, serial, st, signalling ? &halo : nullptr);
                        grp->launchCounter++;
                    }
                    if (signalling) {
                        // what the pass moved goes to the neighbours that cannot compute it; the event is owed even when
                        // this rank held no stencil of the pass
                        launchXpbdHaloPush(halo, grp->d_x.ptr, static_cast<int>(grp->xh.pushOffsets[pass]),
                                           static_cast<int>(grp->xh.pushOffsets[pass + 1]), event, st);
                        grp->launchCounter++;
                    } else if (viaNccl) {
                        grp->launchCounter += enqueueXpbdHaloExchange(grp, pass, st);
                    }
                }
            }
        }
        if (signalling) {
            // the neighbours' stores of the last pass must have landed before the halo copies are finalized
            launchHaloWaitAdvance(halo, event, grp->p2p.d_sync.ptr + kMaxPeers, st);
            grp->launchCounter++;
        }
        if (timePhases) grp->markPhase(st);
        if (fuseState && s + 1 < nsub)
            launchFinalizePredict(grp->d_x.ptr, grp->d_v.ptr, grp->d_xOld.ptr, grp->nLocal, dt, invdt, grp->graphKey.g, st);
        else
            launchFinalize(grp->d_x.ptr, grp->d_xOld.ptr, grp->d_v.ptr, grp->nLocal, invdt, st);
        grp->launchCounter++;
        if (timePhases) grp->markPhase(st);
    }
}

void runStep(spgx_group *grp, int solver, double dt, int nsub, int iters, const double *gravity, int nFrames, int sync)
{
    if (nsub < 1 || iters < 0 || nFrames < 0) throw std::runtime_error("invalid step parameters");
    SPGX_CUDA(cudaSetDevice(grp->ctx->device));
    cudaStream_t st = grp->ctx->stream;
    throwIfHaloTimedOut(grp);  // of an earlier step enqueued with sync = 0
    spgx_group::GraphKey key;
    key.solver = solver;
    key.nsub = nsub;
    key.iters = iters;
    key.dt = dt;
    static const double defaultG[3] = {0, -9.8, 0};  // baseSolver.h:39
    const double *g = gravity ? gravity : defaultG;
    key.g[0] = g[0];
    key.g[1] = g[1];
    key.g[2] = g[2];
    const bool keyChanged = !(key == grp->graphKey);
    grp->graphKey = key;
    grp->launchCounter = 0;
    const bool timePhases = !grp->graphMode && sync;
    grp->phaseCursor = 0;
    auto enqueue = [&] {
        if (timePhases) grp->markPhase(st);
        if (solver == SPGX_VBD)
            enqueueVbdFrame(grp, dt, nsub, iters, st, timePhases);
        else
            enqueueXpbdFrame(grp, dt, nsub, iters, st, timePhases);
    };
    if (grp->partitioned && !grp->commWarm && !(grp->p2p.mode == SPGX_HALO_P2P && grp->p2p.connected)) {
        // NCCL sets up its peer connections on first use (allocations, not capturable): exchange every colour once,
        // eagerly.  Halo copies already equal their owners' values, so this changes nothing.
        if (solver == SPGX_VBD) {
            const int nColors = static_cast<int>(grp->colorOffsets.size()) - 1;
            for (int c = 0; c < nColors; ++c) enqueueHaloExchange(grp, c, st);
        } else {
            for (int pass = 0; pass < grp->xh.nPasses; ++pass) enqueueXpbdHaloExchange(grp, pass, st);
        }
        SPGX_CUDA(cudaStreamSynchronize(st));
        grp->commWarm = true;
    }
    SPGX_CUDA(cudaEventRecord(grp->evStart, st));
    if (grp->graphMode) {
        if (!grp->graphExec || keyChanged) {
            if (grp->graphExec) {
                cudaGraphExecDestroy(grp->graphExec);
                grp->graphExec = nullptr;
            }
            cudaGraph_t graph = nullptr;
            SPGX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            try {
                enqueue();
            } catch (...) {
                cudaStreamEndCapture(st, &graph);
                if (graph) cudaGraphDestroy(graph);
                throw;
            }
            SPGX_CUDA(cudaStreamEndCapture(st, &graph));
            cudaError_t err = cudaGraphInstantiate(&grp->graphExec, graph, 0);
            cudaGraphDestroy(graph);
            if (err != cudaSuccess) throw CudaError(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(err));
            grp->graphLaunches = grp->launchCounter;
        }
        for (int f = 0; f < nFrames; ++f) SPGX_CUDA(cudaGraphLaunch(grp->graphExec, st));
        grp->launchCounter = grp->graphLaunches * nFrames;
    } else {
        for (int f = 0; f < nFrames; ++f) enqueue();
    }
    SPGX_CUDA(cudaEventRecord(grp->evStop, st));
    SPGX_CUDA(cudaGetLastError());
    grp->timings.kernel_launches = grp->launchCounter;
    if (sync) {
        SPGX_CUDA(cudaStreamSynchronize(st));
        throwIfHaloTimedOut(grp);
        float ms = 0;
        SPGX_CUDA(cudaEventElapsedTime(&ms, grp->evStart, grp->evStop));
        grp->timings.total_ms = ms;
        grp->timings.inertial_ms = grp->timings.solve_ms = grp->timings.update_ms = 0;
        if (timePhases) {
            // events per frame: [frame start] then per substep (after predict, after sweeps, after finalize)
            size_t q = 0;
            for (int f = 0; f < nFrames; ++f) {
                cudaEvent_t prev = grp->phaseEvents[q++];
                for (int s = 0; s < nsub; ++s)
                    for (int ph = 0; ph < 3; ++ph) {
                        cudaEvent_t cur = grp->phaseEvents[q++];
                        float pms = 0;
                        SPGX_CUDA(cudaEventElapsedTime(&pms, prev, cur));
                        (ph == 0 ? grp->timings.inertial_ms : ph == 1 ? grp->timings.solve_ms : grp->timings.update_ms) += pms;
                        prev = cur;
                    }
            }
        }
    }
}
}  // namespace

extern "C" {

int spgx_download(spgx_group *grp, double *x, double *v)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        const size_t n3 = 3 * static_cast<size_t>(grp->nVerts);
        if (!grp->d_x.ptr) {  // nothing on the device yet: the pending host state is the state
            if (grp->xPending.size() != n3) throw StateError("group has no state");
            if (x) std::copy(grp->xPending.begin(), grp->xPending.end(), x);
            if (v) std::copy(grp->vPending.begin(), grp->vPending.end(), v);
            return;
        }
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        const bool identity = grp->identityOrder;
        for (int k = 0; k < 2; ++k) {
            double *dst = k == 0 ? x : v;
            const double *src = k == 0 ? grp->d_x.ptr : grp->d_v.ptr;
            if (!dst) continue;
            if (identity) {
                SPGX_CUDA(cudaMemcpyAsync(dst, src, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
            } else {
                // stage[vertex] = src[slot] for the owned slots, then the owned index range goes to the host
                launchPermute3(src, grp->d_stage.ptr, grp->d_slot2vert.ptr, grp->nOwned, /*gather=*/false, st);
                const size_t lo = 3 * static_cast<size_t>(grp->ownedLo), cnt = 3 * static_cast<size_t>(grp->ownedHi - grp->ownedLo);
                if (grp->ownedContiguous) {
                    if (cnt) SPGX_CUDA(cudaMemcpyAsync(dst + lo, grp->d_stage.ptr + lo, cnt * sizeof(double), cudaMemcpyDeviceToHost, st));
                } else {
                    // scattered ownership: through a host copy of the range, keeping the entries of other ranks intact
                    std::vector<double> tmp(cnt);
                    if (cnt) SPGX_CUDA(cudaMemcpyAsync(tmp.data(), grp->d_stage.ptr + lo, cnt * sizeof(double), cudaMemcpyDeviceToHost, st));
                    SPGX_CUDA(cudaStreamSynchronize(st));
                    for (int s = 0; s < grp->nOwned; ++s) {
                        const size_t o = 3 * static_cast<size_t>(grp->slot2vert[s]);
                        dst[o] = tmp[o - lo], dst[o + 1] = tmp[o - lo + 1], dst[o + 2] = tmp[o - lo + 2];
                    }
                }
            }
            SPGX_CUDA(cudaStreamSynchronize(st));
        }
        throwIfHaloTimedOut(grp);  // steps enqueued with sync = 0 have completed by now: stale halo data must not pass silently
    });
}

int spgx_upload(spgx_group *grp, const double *x, const double *v)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        const size_t n3 = 3 * static_cast<size_t>(grp->nVerts);
        if (!grp->finalized) {
            if (x) grp->xPending.assign(x, x + n3);
            if (v) grp->vPending.assign(v, v + n3);
            return;
        }
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        const bool identity = grp->identityOrder;
        for (int k = 0; k < 2; ++k) {
            const double *src = k == 0 ? x : v;
            double *dst = k == 0 ? grp->d_x.ptr : grp->d_v.ptr;
            if (!src) continue;
            if (identity) {
                SPGX_CUDA(cudaMemcpyAsync(dst, src, n3 * sizeof(double), cudaMemcpyHostToDevice, st));
            } else {
                // only the index range the local vertices (owned + halo) span crosses the bus
                const size_t lo = 3 * static_cast<size_t>(grp->localLo), cnt = 3 * static_cast<size_t>(grp->localHi - grp->localLo);
                if (cnt) SPGX_CUDA(cudaMemcpyAsync(grp->d_stage.ptr + lo, src + lo, cnt * sizeof(double), cudaMemcpyHostToDevice, st));
                // dst[slot] = stage[slot2vert[slot]]
                launchPermute3(grp->d_stage.ptr, dst, grp->d_slot2vert.ptr, grp->nLocal, /*gather=*/true, st);
            }
            SPGX_CUDA(cudaStreamSynchronize(st));
        }
    });
}

int spgx_reset(spgx_group *grp)
{
    return guarded([&] {
        std::vector<double> zeros(3 * static_cast<size_t>(grp->nVerts), 0.0);
        grp->newton.bdf2Init = false;  // BDF2::reset (bdf2.cpp:78-83)
        grp->newton.currentDt = -1.0;  // ImplicitEulerNewtonRobust::reset (implicitEulerNewtonRobust.cpp:132-137)
        grp->newton.consecutiveSuccessfulSteps = 0;
        if (spgx_upload(grp, grp->xInitial.data(), zeros.data()) != SPGX_OK) throw CudaError(spgx_last_error());
    });
}

int spgx_scale_masses(spgx_group *grp, double factor)
{
    return guarded([&] {
        for (auto &m : grp->mass)
            if (m != 0) m *= factor;
        if (grp->finalized) {
            SPGX_CUDA(cudaSetDevice(grp->ctx->device));
            std::vector<double> ms(grp->nLocal), ws(grp->nLocal);
            for (int s = 0; s < grp->nLocal; ++s) {
                const double m = grp->mass[grp->slot2vert[s]];
                ms[s] = m;
                ws[s] = (m == 0) ? 0 : 1 / m;
            }
            grp->d_mass.upload(ms, grp->ctx->stream);
            grp->d_invMass.upload(ws, grp->ctx->stream);
            for (auto &e : grp->energies)  // the per-stencil copies the XPBD passes read
                if (e.d_stencilW.ptr)
                    launchGatherScalar(grp->d_invMass.ptr, e.d_stencils.ptr, static_cast<long long>(e.d_stencilW.count), e.d_stencilW.ptr,
                                       grp->ctx->stream);
            SPGX_CUDA(cudaStreamSynchronize(grp->ctx->stream));
            if (grp->graphExec) {  // buffers were re-allocated
                cudaGraphExecDestroy(grp->graphExec);
                grp->graphExec = nullptr;
            }
        }
    });
}

int spgx_vbd_step(spgx_group *grp, double dt, int nsubsteps, int iterations, const double *gravity, int nFrames, int sync)
{
    return guarded([&] {
        requireFinalized(grp, true);
        runStep(grp, SPGX_VBD, dt, nsubsteps, iterations, gravity, nFrames, sync);
    });
}

int spgx_xpbd_step(spgx_group *grp, double dt, int nsubsteps, int iterations, const double *gravity, int nFrames,
                   int sync)
{
    return guarded([&] {
        requireFinalized(grp, false);
        if (iterations < 1) throw std::runtime_error("XPBD needs at least one iteration");
        for (const auto &e : grp->energies)  // staticLoadEnergy.cpp:8-13: the reference throws from projectPosition
            if (e.kind == K_STATIC_LOAD) throw std::runtime_error("Constraint-based load energy not implemented");
        runStep(grp, grp->solverKind, dt, nsubsteps, iterations, gravity, nFrames, sync);
    });
}

int spgx_check_finite(spgx_group *grp, int64_t *out3)
{
    return guarded([&] {
        if (!grp || !grp->finalized || !out3) throw StateError("group is not finalized");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        DevBuf<unsigned long long> d;
        d.alloc(3);
        const unsigned long long init[3] = {0ull, 0ull, ~0ull};
        SPGX_CUDA(cudaMemcpyAsync(d.ptr, init, sizeof(init), cudaMemcpyHostToDevice, st));
        launchCheckFinite(grp->d_x.ptr, grp->d_v.ptr, grp->d_slot2vert.ptr, grp->nOwned, d.ptr, st);
        unsigned long long h[3];
        SPGX_CUDA(cudaMemcpyAsync(h, d.ptr, sizeof(h), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        out3[0] = static_cast<int64_t>(h[0]);
        out3[1] = static_cast<int64_t>(h[1]);
        out3[2] = h[2] == ~0ull ? -1 : static_cast<int64_t>(h[2] - 1);
    });
}

int spgx_last_timings(spgx_group *grp, spgx_timings *out)
{
    return guarded([&] {
        if (!grp || !out) throw StateError("null argument");
        *out = grp->timings;
    });
}

int spgx_set_graph_mode(spgx_group *grp, int enabled)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        grp->graphMode = enabled != 0;
    });
}

int spgx_eval_vertex_force_hessian(spgx_group *grp, int energyId, int stencil, int local, int handDerived,
                                   double *force3, double *hess9)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("group is not finalized");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        auto &e = grp->energies.at(energyId);
        if (stencil < 0 || stencil >= e.n || local < 0 || local >= e.S) throw std::runtime_error("stencil out of range");
        const int pos = e.positionOf.empty() ? stencil : e.positionOf[stencil];
        if (pos < 0) throw std::runtime_error("stencil is not held by this rank");
        cudaStream_t st = grp->ctx->stream;
        launchEvalVertexForceHessian(e.kind, grp->d_x.ptr, e.d_stencils.ptr + static_cast<size_t>(pos) * e.S,
                                     e.d_records.ptr + static_cast<size_t>(pos) * e.NP, local, grp->d_eval.ptr, handDerived,
                                     st);
        double out[12];
        SPGX_CUDA(cudaMemcpyAsync(out, grp->d_eval.ptr, sizeof(out), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        std::copy(out, out + 3, force3);
        std::copy(out + 3, out + 12, hess9);
    });
}

int spgx_eval_constraints(spgx_group *grp, int energyId, int stencil, double *C, double *grad)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("group is not finalized");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        auto &e = grp->energies.at(energyId);
        if (stencil < 0 || stencil >= e.n) throw std::runtime_error("stencil out of range");
        const int pos = e.positionOf.empty() ? stencil : e.positionOf[stencil];
        if (pos < 0) throw std::runtime_error("stencil is not held by this rank");
        cudaStream_t st = grp->ctx->stream;
        launchEvalConstraints(e.kind, grp->d_x.ptr, e.d_stencils.ptr + static_cast<size_t>(pos) * e.S,
                              e.d_records.ptr + static_cast<size_t>(pos) * e.NP, grp->d_eval.ptr, grp->d_eval.ptr + 4, st);
        double out[64];
        SPGX_CUDA(cudaMemcpyAsync(out, grp->d_eval.ptr, sizeof(out), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        std::copy(out, out + e.NC, C);
        std::copy(out + 4, out + 4 + 3 * e.S * e.NC, grad);
    });
}

}  // extern "C"

namespace
{
void ensureXfer(spgx_group *grp)
{
    spgx_ctx *ctx = grp->ctx;
    if (!ctx->h2dStream) SPGX_CUDA(cudaStreamCreateWithFlags(&ctx->h2dStream, cudaStreamNonBlocking));
    if (!ctx->d2hStream) SPGX_CUDA(cudaStreamCreateWithFlags(&ctx->d2hStream, cudaStreamNonBlocking));
    auto &x = grp->xfer;
    const size_t n3 = 3 * static_cast<size_t>(grp->nVerts);
    if (x.d_inX.count != n3) {
        x.d_inX.alloc(n3), x.d_inV.alloc(n3), x.d_outX.alloc(n3), x.d_outV.alloc(n3);
        x.inUsed = x.outUsed = false;
    }
    for (cudaEvent_t *e : {&x.uploaded, &x.inFree, &x.outReady, &x.downloaded})
        if (!*e) SPGX_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
}
}  // namespace

extern "C" {

// Asynchronous forms of spgx_upload / spgx_download for a frame loop whose host I/O should not serialise with the solve:
//   H2D runs on its own stream into a staging pair (vertex order) while earlier steps still run; the scatter into slot
//   order is enqueued on the compute stream behind them.  The download gathers into a second staging pair on the compute
//   stream (behind the steps enqueued so far) and copies out on a third stream while later steps run.
// Host buffers must be page-locked (spgx_host_alloc / spgx_host_register) and stay untouched until spgx_ctx_sync (or
// spgx_transfer_wait) returns.
int spgx_upload_async(spgx_group *grp, const double *x, const double *v)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("group is not finalized");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        ensureXfer(grp);
        auto &xf = grp->xfer;
        cudaStream_t st = grp->ctx->stream, cs = grp->ctx->h2dStream;
        const bool identity = grp->identityOrder;
        const size_t lo = identity ? 0 : 3 * static_cast<size_t>(grp->localLo);
        const size_t cnt = identity ? 3 * static_cast<size_t>(grp->nVerts) : 3 * static_cast<size_t>(grp->localHi - grp->localLo);
        if (xf.inUsed) SPGX_CUDA(cudaStreamWaitEvent(cs, xf.inFree, 0));  // the previous scatter has read the staging pair
        if (x && cnt) SPGX_CUDA(cudaMemcpyAsync(xf.d_inX.ptr + lo, x + lo, cnt * sizeof(double), cudaMemcpyHostToDevice, cs));
        if (v && cnt) SPGX_CUDA(cudaMemcpyAsync(xf.d_inV.ptr + lo, v + lo, cnt * sizeof(double), cudaMemcpyHostToDevice, cs));
        SPGX_CUDA(cudaEventRecord(xf.uploaded, cs));
        SPGX_CUDA(cudaStreamWaitEvent(st, xf.uploaded, 0));
        for (int k = 0; k < 2; ++k) {
            const double *src = k == 0 ? (x ? xf.d_inX.ptr : nullptr) : (v ? xf.d_inV.ptr : nullptr);
            double *dst = k == 0 ? grp->d_x.ptr : grp->d_v.ptr;
            if (!src) continue;
            if (identity)
                SPGX_CUDA(cudaMemcpyAsync(dst, src, cnt * sizeof(double), cudaMemcpyDeviceToDevice, st));
            else
                launchPermute3(src, dst, grp->d_slot2vert.ptr, grp->nLocal, /*gather=*/true, st);
        }
        SPGX_CUDA(cudaEventRecord(xf.inFree, st));
        xf.inUsed = true;
    });
}

int spgx_download_async(spgx_group *grp, double *x, double *v)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("group is not finalized");
        if (!grp->ownedContiguous) throw StateError("asynchronous download needs a contiguous owned index range");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        ensureXfer(grp);
        auto &xf = grp->xfer;
        cudaStream_t st = grp->ctx->stream, cs = grp->ctx->d2hStream;
        const bool identity = grp->identityOrder;
        const size_t lo = identity ? 0 : 3 * static_cast<size_t>(grp->ownedLo);
        const size_t cnt = identity ? 3 * static_cast<size_t>(grp->nVerts) : 3 * static_cast<size_t>(grp->ownedHi - grp->ownedLo);
        if (xf.outUsed) SPGX_CUDA(cudaStreamWaitEvent(st, xf.downloaded, 0));  // the previous copy-out has left the staging pair
        for (int k = 0; k < 2; ++k) {
            double *dst = k == 0 ? (x ? xf.d_outX.ptr : nullptr) : (v ? xf.d_outV.ptr : nullptr);
            const double *src = k == 0 ? grp->d_x.ptr : grp->d_v.ptr;
            if (!dst) continue;
            if (identity)
                SPGX_CUDA(cudaMemcpyAsync(dst, src, cnt * sizeof(double), cudaMemcpyDeviceToDevice, st));
            else
                launchPermute3(src, dst, grp->d_slot2vert.ptr, grp->nOwned, /*gather=*/false, st);
        }
        SPGX_CUDA(cudaEventRecord(xf.outReady, st));
        SPGX_CUDA(cudaStreamWaitEvent(cs, xf.outReady, 0));
        if (x && cnt) SPGX_CUDA(cudaMemcpyAsync(x + lo, xf.d_outX.ptr + lo, cnt * sizeof(double), cudaMemcpyDeviceToHost, cs));
        if (v && cnt) SPGX_CUDA(cudaMemcpyAsync(v + lo, xf.d_outV.ptr + lo, cnt * sizeof(double), cudaMemcpyDeviceToHost, cs));
        SPGX_CUDA(cudaEventRecord(xf.downloaded, cs));
        xf.outUsed = true;
    });
}

// block until the asynchronous transfers enqueued so far have completed (the host buffers are the caller's again)
int spgx_transfer_wait(spgx_group *grp)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        if (grp->ctx->h2dStream) SPGX_CUDA(cudaStreamSynchronize(grp->ctx->h2dStream));
        if (grp->ctx->d2hStream) SPGX_CUDA(cudaStreamSynchronize(grp->ctx->d2hStream));
        throwIfHaloTimedOut(grp);
    });
}

}  // extern "C"
