// C-ABI, part 3: the implicit-Euler family (ImplicitEulerBaraffWitkin, ImplicitEulerNewtonDx, ImplicitEulerNewtonDv, BDF2,
// StaticNewton, QuasiStaticNewton) and symplectic Euler on a group finalized with SPGX_NEWTON.
//
// Reference: spg/source/sim/solver/implicitEulerBase.cpp (system force / stiffness / mass, linear solve),
// implicitEulerBaraffWitkin.cpp:31-54, implicitEulerNewtonDx.cpp:31-59, implicitEulerNewtonDv.cpp:30-62, bdf2.cpp:31-70,
// staticNewton.cpp:29-45, quasiStaticNewton.cpp:31-49, simplecticEuler.cpp:12-62.  The reference rebuilds a triplet list, two Eigen sparse matrices and a fresh CG object
// every substep on the host; here the sparse pattern is fixed at finalize time and a substep is: element Hessians ->
// block gather -> device-resident CG (newton.cu), with the host only polling the convergence flag.
#include <algorithm>
#include <cstring>

#include "energies.cuh"
#include "group.h"
#include "host_build.h"

using namespace spgx;

namespace spgx
{
void newtonBuild(spgx_group *grp, cudaStream_t st)
{
    const int V = grp->nVerts;
    auto &nw = grp->newton;
    const size_t nE = grp->energies.size();
    // vertex -> (stencil, local) incidence per energy, ascending flat index (the order vbd.cpp:231-268 keeps too)
    std::vector<Incidence> inc;
    for (auto &e : grp->energies) {
        if (static_cast<uint64_t>(e.n) >= (1ull << 30)) throw std::runtime_error("too many stencils in one energy");
        inc.push_back(buildIncidence(V, {e.stencils.data(), static_cast<int64_t>(e.stencils.size()), e.S}));
    }
    for (size_t ei = 0; ei < nE; ++ei) {
        auto &e = grp->energies[ei];
        std::vector<uint32_t> rowptr(static_cast<size_t>(V) + 1, 0), entries(inc[ei].flat.size());
        for (int v = 0; v <= V; ++v) rowptr[v] = static_cast<uint32_t>(inc[ei].rowptr[v]);
        for (size_t q = 0; q < entries.size(); ++q) {
            const int64_t flat = inc[ei].flat[q];
            entries[q] = (static_cast<uint32_t>(flat / e.S) << 2) | static_cast<uint32_t>(flat % e.S);
        }
        e.d_rowptr.upload(rowptr, st);
        e.d_entries.upload(entries, st);
        SPGX_CUDA(cudaStreamSynchronize(st));
    }
    // sparse pattern: block row v = v itself and every vertex sharing a stencil with it, ascending
    nw.bsrRowptr.assign(static_cast<size_t>(V) + 1, 0);
    nw.bsrCol.clear();
    std::vector<int32_t> bsrRow, cols;
    std::vector<uint32_t> diagPos(static_cast<size_t>(V), 0);
    for (int v = 0; v < V; ++v) {
        cols.clear();
        cols.push_back(v);
        for (size_t ei = 0; ei < nE; ++ei) {
            const auto &e = grp->energies[ei];
            for (int64_t q = inc[ei].rowptr[v]; q < inc[ei].rowptr[v + 1]; ++q) {
                const int64_t base = (inc[ei].flat[q] / e.S) * e.S;
                for (int l = 0; l < e.S; ++l) cols.push_back(e.stencils[base + l]);
            }
        }
        std::sort(cols.begin(), cols.end());
        cols.erase(std::unique(cols.begin(), cols.end()), cols.end());
        for (int32_t c : cols) {
            if (c == v) diagPos[v] = static_cast<uint32_t>(nw.bsrCol.size());
            nw.bsrCol.push_back(c);
            bsrRow.push_back(v);
        }
        if (nw.bsrCol.size() >= (1ull << 31)) throw std::runtime_error("sparse pattern too large");
        nw.bsrRowptr[v + 1] = static_cast<uint32_t>(nw.bsrCol.size());
    }
    nw.nnzb = static_cast<int64_t>(nw.bsrCol.size());
    nw.d_bsrRowptr.upload(nw.bsrRowptr, st);
    nw.d_bsrCol.upload(nw.bsrCol, st);
    nw.d_bsrRow.upload(bsrRow, st);
    nw.d_diagPos.upload(diagPos, st);
    SPGX_CUDA(cudaStreamSynchronize(st));
    nw.d_bsrH.alloc(9 * static_cast<size_t>(nw.nnzb));
    // element Hessians: one arena, energy ei owning the blocks [baseBlk[ei], baseBlk[ei + 1]) (a block = 9 doubles)
    std::vector<uint64_t> baseBlk(nE + 1, 0);
    nw.elemHessianBytes = 0;
    for (size_t ei = 0; ei < nE; ++ei) {
        const auto &e = grp->energies[ei];
        baseBlk[ei + 1] = baseBlk[ei] + static_cast<uint64_t>(e.n) * e.S * e.S;
        baseBlk[ei + 1] += baseBlk[ei + 1] & 1ull;  // every energy starts on 16 bytes (a block is 72): hingeHessianKernel stores double2
        nw.elemHessianBytes += 8.0 * e.n * 9 * e.S * e.S;
    }
    if (baseBlk[nE] >= (1ull << 32)) throw std::runtime_error("too many element Hessian blocks for 32-bit gather lists");
    nw.d_elemArena.alloc(9 * static_cast<size_t>(baseBlk[nE]));
    nw.d_elemH.assign(nE, nullptr);
    for (size_t ei = 0; ei < nE; ++ei) nw.d_elemH[ei] = nw.d_elemArena.ptr + 9 * static_cast<size_t>(baseBlk[ei]);
    // gather lists: block (v, c) of the sparse Hessian sums, energies in insertion order and the row vertex's incidences
    // ascending (the order getSystemStiffnessMatrix's triplets are summed in is not defined; this one is fixed), the
    // element blocks (la, lb) of the stencils of v whose vertex lb is c
    {
        std::vector<uint32_t> asmPtr(static_cast<size_t>(nw.nnzb) + 1, 0), asmSrc;
        asmSrc.reserve(static_cast<size_t>(baseBlk[nE]));
        std::vector<std::vector<uint32_t>> perCol;  // of the current row, indexed by position in the row
        for (int v = 0; v < V; ++v) {
            const uint32_t rb = nw.bsrRowptr[v], re = nw.bsrRowptr[v + 1];
            perCol.assign(re - rb, {});
            for (size_t ei = 0; ei < nE; ++ei) {
                const auto &e = grp->energies[ei];
                for (int64_t q = inc[ei].rowptr[v]; q < inc[ei].rowptr[v + 1]; ++q) {
                    const int64_t flat = inc[ei].flat[q];
                    const int64_t stl = flat / e.S;
                    const int la = static_cast<int>(flat % e.S);
                    for (int lb = 0; lb < e.S; ++lb) {
                        const int32_t c = e.stencils[stl * e.S + lb];
                        const auto it = std::lower_bound(nw.bsrCol.begin() + rb, nw.bsrCol.begin() + re, c);
                        perCol[static_cast<size_t>(it - (nw.bsrCol.begin() + rb))].push_back(
                            static_cast<uint32_t>(baseBlk[ei] + static_cast<uint64_t>(stl) * e.S * e.S + la * e.S + lb));
                    }
                }
            }
            for (uint32_t t = rb; t < re; ++t) {
                asmSrc.insert(asmSrc.end(), perCol[t - rb].begin(), perCol[t - rb].end());
                asmPtr[t + 1] = static_cast<uint32_t>(asmSrc.size());
            }
        }
        nw.d_asmPtr.upload(asmPtr, st);
        nw.d_asmSrc.upload(asmSrc, st);
        SPGX_CUDA(cudaStreamSynchronize(st));
    }
    const size_t n3 = 3 * static_cast<size_t>(V);
    for (auto *b : {&nw.d_f, &nw.d_rhs, &nw.d_sol, &nw.d_r, &nw.d_p, &nw.d_z, &nw.d_tmp, &nw.d_invdiag, &nw.d_x0, &nw.d_v0, &nw.d_xm1, &nw.d_vm1,
                    &nw.d_xi})
        b->alloc(n3);
    nw.d_partials.alloc(3 * kCgBlocks);
    nw.d_state.alloc(2);
    if (!nw.hostState) SPGX_CUDA(cudaMallocHost(&nw.hostState, sizeof(CgState)));
    if (!nw.hostScalar) SPGX_CUDA(cudaMallocHost(&nw.hostScalar, sizeof(double)));
    nw.d_scalar.alloc(1);
    nw.currentDt = -1.0;
    nw.consecutiveSuccessfulSteps = 0;
    for (auto &e : nw.ev)
        if (!e) SPGX_CUDA(cudaEventCreate(&e));
    nw.assembled = false;
    nw.bdf2Init = false;
    nw.cgIterations.clear();
}
}  // namespace spgx

namespace
{
void requireNewton(spgx_group *grp)
{
    if (!grp || !grp->finalized) throw StateError("group is not finalized (call spgx_group_finalize first)");
    if (grp->solverKind != SPGX_NEWTON) throw StateError("group was not finalized with SPGX_NEWTON");
    if (cudaSetDevice(grp->ctx->device) != cudaSuccess) throw CudaError("cudaSetDevice failed");
}

NewtonArgs newtonArgs(spgx_group *grp)
{
    NewtonArgs a{};
    auto &nw = grp->newton;
    a.x = grp->d_x.ptr;
    a.mass = grp->d_mass.ptr;
    a.nVerts = grp->nVerts;
    a.nEnergies = static_cast<int>(grp->energies.size());
    if (a.nEnergies > kMaxEnergies) throw std::runtime_error("too many energies in one group");
    for (int e = 0; e < a.nEnergies; ++e) {
        const auto &en = grp->energies[e];
        a.e[e] = {en.kind, en.n, en.d_stencils.ptr, en.d_records.ptr, en.d_rowptr.ptr, en.d_entries.ptr, nw.d_elemH[e]};
    }
    a.bsrRowptr = nw.d_bsrRowptr.ptr;
    a.bsrCol = nw.d_bsrCol.ptr;
    a.bsrRow = nw.d_bsrRow.ptr;
    a.diagPos = nw.d_diagPos.ptr;
    a.bsrH = nw.d_bsrH.ptr;
    a.nnzb = nw.nnzb;
    a.asmPtr = nw.d_asmPtr.ptr;
    a.asmSrc = nw.d_asmSrc.ptr;
    a.elemArena = nw.d_elemArena.ptr;
    return a;
}

CgArgs cgArgs(spgx_group *grp, double cM, double cH)
{
    auto &nw = grp->newton;
    CgArgs c{};
    c.bsrRowptr = nw.d_bsrRowptr.ptr;
    c.bsrCol = nw.d_bsrCol.ptr;
    c.bsrH = nw.d_bsrH.ptr;
    c.nnzb = nw.nnzb;
    c.diagPos = nw.d_diagPos.ptr;
    c.mass = grp->d_mass.ptr;
    c.cM = cM;
    c.cH = cH;
    c.nVerts = grp->nVerts;
    c.x = nw.d_sol.ptr;
    c.r = nw.d_r.ptr;
    c.p = nw.d_p.ptr;
    c.z = nw.d_z.ptr;
    c.tmp = nw.d_tmp.ptr;
    c.invdiag = nw.d_invdiag.ptr;
    c.partials = nw.d_partials.ptr;
    c.state = nw.d_state.ptr;
    return c;
}

// element Hessians + block gather of the current positions
void assembleHessian(spgx_group *grp, cudaStream_t st)
{
    const NewtonArgs a = newtonArgs(grp);
    launchElementHessians(a, st);
    launchBsrAssemble(a, st);
    grp->launchCounter += a.nEnergies + 1;
    grp->newton.assembled = true;
}

// solve (cM M + cH H) sol = b on the device; returns the iteration count (cg.iterations())
int conjugateGradient(spgx_group *grp, double cM, double cH, const double *b, double tol, int maxIters, cudaStream_t st)
{
    auto &nw = grp->newton;
    const CgArgs c = cgArgs(grp, cM, cH);
    launchCgInit(c, b, tol, st);
    grp->launchCounter += 3;
    constexpr int kChunk = 8;  // iterations enqueued between two looks at the convergence flag
    int it = 0;
    for (;;) {
        SPGX_CUDA(cudaMemcpyAsync(nw.hostState, nw.d_state.ptr + (it & 1), sizeof(CgState), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        if (nw.hostState->done || it >= maxIters) break;
        const int n = std::min(kChunk, maxIters - it);
        for (int k = 0; k < n; ++k) launchCgIteration(c, it + k, st);
        grp->launchCounter += 3 * n;
        it += n;
    }
    SPGX_CUDA(cudaGetLastError());
    return nw.hostState->iters;
}

void copyD2D(double *dst, const double *src, size_t n, cudaStream_t st)
{
    if (n) SPGX_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
}

// out = c1 f - c2 M (a - b), returns |out| (one stream synchronisation: the line search branches on it)
double residualNorm(spgx_group *grp, double *out, double c1, double c2, const double *a, const double *b, cudaStream_t st)
{
    auto &nw = grp->newton;
    launchResidualNorm(out, c1, nw.d_f.ptr, c2, grp->d_mass.ptr, a, b, grp->nVerts, nw.d_partials.ptr, nw.d_scalar.ptr, st);
    grp->launchCounter += 2;
    SPGX_CUDA(cudaMemcpyAsync(nw.hostScalar, nw.d_scalar.ptr, sizeof(double), cudaMemcpyDeviceToHost, st));
    SPGX_CUDA(cudaStreamSynchronize(st));
    return *nw.hostScalar;
}

// One substep of ImplicitEulerNewtonRobust::step (implicitEulerNewtonRobust.cpp:31-127) or QuasiStaticNewtonRobust::step
// (quasiStaticNewtonRobust.cpp:33-125): the reference's control flow on the host, every vector on the device.
void robustSubstep(spgx_group *grp, bool quasi, double maxDt, const double *g, double cgTol, int cgMaxIters, cudaStream_t st)
{
    auto &nw = grp->newton;
    const long long n3 = 3LL * grp->nVerts;
    double *x = grp->d_x.ptr, *v = grp->d_v.ptr;
    const double newtonThreshold = quasi ? 1e-2 : 1e-5;  // quasiStaticNewtonRobust.h:20 / implicitEulerNewtonRobust.h:19
    constexpr int kMaxNewton = 5, kMaxLineSearch = 5, kRequiredGoodSteps = 5;
    constexpr double kSplit = 0.75;
    bool successful = true;
    if (nw.currentDt == -1. || nw.currentDt > maxDt) nw.currentDt = maxDt;
    const double dt = nw.currentDt, invdt = 1 / dt, invdtSquared = invdt * invdt;
    copyD2D(nw.d_x0.ptr, x, n3, st);
    if (!quasi) {
        copyD2D(nw.d_v0.ptr, v, n3, st);
        launchAxpby(x, 1.0, x, dt, v, n3, st);  // initial guess: the inertial position
        grp->launchCounter++;
    }
    // the accepted iterate: velocities vi (Newton) / positions xi (quasi-static)
    copyD2D(nw.d_xi.ptr, quasi ? nw.d_x0.ptr : nw.d_v0.ptr, n3, st);
    double stepResidual = 0;
    int newtonIterations = 0;
    do {
        const NewtonArgs a = newtonArgs(grp);
        launchSystemForce(a, g, nw.d_f.ptr, st);
        assembleHessian(grp, st);
        grp->launchCounter++;
        double initialResidual;
        if (quasi) {
            launchAxpby(v, invdt, x, -invdt, nw.d_x0.ptr, n3, st);  // computeSystemIntegratedVelocities
            grp->launchCounter++;
            initialResidual = residualNorm(grp, nw.d_rhs.ptr, 1.0, invdtSquared, x, nw.d_x0.ptr, st);
            nw.cgIterations.push_back(conjugateGradient(grp, invdtSquared, 1.0, nw.d_rhs.ptr, cgTol, cgMaxIters, st));
        } else {
            initialResidual = residualNorm(grp, nw.d_rhs.ptr, dt, 1.0, nw.d_xi.ptr, nw.d_v0.ptr, st);
            nw.cgIterations.push_back(conjugateGradient(grp, 1.0, dt * dt, nw.d_rhs.ptr, cgTol, cgMaxIters, st));
        }
        stepResidual = initialResidual;
        int lineSearch = 0;
        double alpha = 1.;
        while (stepResidual >= initialResidual && lineSearch < kMaxLineSearch) {
            if (quasi) {
                launchAxpby(x, 1.0, nw.d_xi.ptr, alpha, nw.d_sol.ptr, n3, st);   // x = xi + alpha dx
                launchAxpby(v, invdt, x, -invdt, nw.d_x0.ptr, n3, st);          // v = (x - x0) / dt
            } else {
                launchAxpby(v, 1.0, nw.d_xi.ptr, alpha, nw.d_sol.ptr, n3, st);   // v = vi + alpha dv
                launchAxpby(x, 1.0, nw.d_x0.ptr, dt, v, n3, st);                // x = x0 + v dt
            }
            launchSystemForce(newtonArgs(grp), g, nw.d_f.ptr, st);
            grp->launchCounter += 3;
            stepResidual = quasi ? residualNorm(grp, nw.d_tmp.ptr, 1.0, invdtSquared, x, nw.d_x0.ptr, st)
                                 : residualNorm(grp, nw.d_tmp.ptr, dt, 1.0, v, nw.d_v0.ptr, st);
            if (stepResidual < initialResidual) copyD2D(nw.d_xi.ptr, quasi ? x : v, n3, st);
            alpha *= 0.5;
            lineSearch++;
        }
        if (stepResidual >= initialResidual && stepResidual > newtonThreshold) successful = false;
        newtonIterations++;
    } while (stepResidual > newtonThreshold && successful && newtonIterations < kMaxNewton);
    if (stepResidual > newtonThreshold) successful = false;
    if (!successful) {
        copyD2D(x, nw.d_x0.ptr, n3, st);
        if (!quasi) copyD2D(v, nw.d_v0.ptr, n3, st);
        nw.currentDt *= kSplit;
        nw.consecutiveSuccessfulSteps = 0;
    } else {
        nw.consecutiveSuccessfulSteps++;
        if (nw.consecutiveSuccessfulSteps > kRequiredGoodSteps && nw.currentDt < maxDt) nw.currentDt /= kSplit;
    }
}
}  // namespace

extern "C" {

int spgx_implicit_step(spgx_group *grp, int implicitKind, double dt, int nsubsteps, const double *gravity,
                       double cgTolerance, int cgMaxIterations, int nFrames, int /*sync*/)
{
    return guarded([&] {
        requireNewton(grp);
        if (implicitKind < SPGX_IE_BARAFF_WITKIN || implicitKind > SPGX_IE_QUASI_STATIC_NEWTON_ROBUST) throw std::runtime_error("unknown implicit solver kind");
        if (nsubsteps < 1 || nFrames < 0 || cgMaxIterations < 0) throw std::runtime_error("invalid step parameters");
        cudaStream_t st = grp->ctx->stream;
        auto &nw = grp->newton;
        const double g[3] = {gravity ? gravity[0] : 0.0, gravity ? gravity[1] : -9.8, gravity ? gravity[2] : 0.0};
        const double h = dt / nsubsteps, invh = 1. / h;
        const long long n3 = 3LL * grp->nVerts;
        double *x = grp->d_x.ptr, *v = grp->d_v.ptr;
        nw.cgIterations.clear();
        grp->launchCounter = 0;
        double assembleMs = 0, solveMs = 0;
        // the CG loop synchronises with the stream anyway, so per-phase times cost nothing extra: inertial_ms = system
        // force + Hessian assembly ("f computation" + "H computation / assembly" of implicitEulerBase.cpp:139, :215-219),
        // solve_ms = the linear solve (:233-236)
        auto phase = [&](int k) { SPGX_CUDA(cudaEventRecord(nw.ev[k], st)); };
        auto collect = [&] {
            float a = 0, b = 0;
            SPGX_CUDA(cudaEventSynchronize(nw.ev[2]));
            SPGX_CUDA(cudaEventElapsedTime(&a, nw.ev[0], nw.ev[1]));
            SPGX_CUDA(cudaEventElapsedTime(&b, nw.ev[1], nw.ev[2]));
            assembleMs += a;
            solveMs += b;
        };
        SPGX_CUDA(cudaEventRecord(grp->evStart, st));
        for (int f = 0; f < nFrames; ++f)
            for (int s = 0; s < nsubsteps; ++s) {
                phase(0);
                if (implicitKind == SPGX_IE_NEWTON_ROBUST || implicitKind == SPGX_IE_QUASI_STATIC_NEWTON_ROBUST) {
                    robustSubstep(grp, implicitKind == SPGX_IE_QUASI_STATIC_NEWTON_ROBUST, h, g, cgTolerance, cgMaxIterations, st);
                    phase(1);
                    phase(2);
                    collect();
                } else if (implicitKind == SPGX_IE_BARAFF_WITKIN) {
                    // f, K at the current state; (M - h^2 K) dv = h (f + h K v0); v = v0 + dv; x += v h
                    const NewtonArgs a = newtonArgs(grp);
                    launchSystemForce(a, g, nw.d_f.ptr, st);
                    assembleHessian(grp, st);
                    launchBsrSpmv(cgArgs(grp, 0.0, 1.0), v, nw.d_tmp.ptr, st);  // H v0 = -K v0
                    launchAxpby(nw.d_rhs.ptr, h, nw.d_f.ptr, -h * h, nw.d_tmp.ptr, n3, st);
                    grp->launchCounter += 3;
                    phase(1);
                    nw.cgIterations.push_back(conjugateGradient(grp, 1.0, h * h, nw.d_rhs.ptr, cgTolerance, cgMaxIterations, st));
                    phase(2);
                    collect();
                    launchAxpby(v, 1.0, v, 1.0, nw.d_sol.ptr, n3, st);
                    launchAxpby(x, 1.0, x, h, v, n3, st);
                    grp->launchCounter += 2;
                } else if (implicitKind == SPGX_IE_NEWTON_DX) {
                    // x0 = x; x += v h; f, K there; (M / h^2 - K) dx = f; x += dx; v = (x - x0) / h
                    copyD2D(nw.d_x0.ptr, x, n3, st);
                    launchAxpby(x, 1.0, x, h, v, n3, st);
                    const NewtonArgs a = newtonArgs(grp);
                    launchSystemForce(a, g, nw.d_f.ptr, st);
                    assembleHessian(grp, st);
                    grp->launchCounter += 2;
                    phase(1);
                    nw.cgIterations.push_back(conjugateGradient(grp, invh * invh, 1.0, nw.d_f.ptr, cgTolerance, cgMaxIterations, st));
                    phase(2);
                    collect();
                    launchApplyDx(x, v, nw.d_sol.ptr, nw.d_x0.ptr, invh, n3, st);
                    grp->launchCounter += 1;
                } else if (implicitKind == SPGX_IE_BDF2) {
                    // (M - 4/9 h^2 K) dv = 1/3 M (v0 - v-1) + 2/3 h (f + 1/3 K (x0 - x-1 + 2 h v0));
                    // v = v0 + dv; x = 4/3 x0 - 1/3 x-1 + 2/3 h v   (K = -H)
                    if (!nw.bdf2Init) {
                        copyD2D(nw.d_xm1.ptr, x, n3, st);
                        copyD2D(nw.d_vm1.ptr, v, n3, st);
                        nw.bdf2Init = true;
                    }
                    const NewtonArgs a = newtonArgs(grp);
                    launchSystemForce(a, g, nw.d_f.ptr, st);
                    assembleHessian(grp, st);
                    launchAxpby(nw.d_x0.ptr, 1.0, x, -1.0, nw.d_xm1.ptr, n3, st);             // w = x0 - x-1
                    launchAxpby(nw.d_x0.ptr, 1.0, nw.d_x0.ptr, 2.0 * h, v, n3, st);           //     + 2 h v0
                    launchBsrSpmv(cgArgs(grp, 0.0, 1.0), nw.d_x0.ptr, nw.d_tmp.ptr, st);       // H w = -K w
                    launchAxpby(nw.d_rhs.ptr, 2.0 / 3.0 * h, nw.d_f.ptr, -2.0 / 9.0 * h, nw.d_tmp.ptr, n3, st);
                    launchAddMassDiff(nw.d_rhs.ptr, 1.0 / 3.0, grp->d_mass.ptr, v, nw.d_vm1.ptr, grp->nVerts, st);
                    grp->launchCounter += 6;
                    phase(1);
                    nw.cgIterations.push_back(conjugateGradient(grp, 1.0, 4.0 / 9.0 * h * h, nw.d_rhs.ptr, cgTolerance, cgMaxIterations, st));
                    phase(2);
                    collect();
                    copyD2D(nw.d_vm1.ptr, v, n3, st);                                          // v-1 <- v0
                    launchAxpby(v, 1.0, v, 1.0, nw.d_sol.ptr, n3, st);                         // v = v0 + dv
                    copyD2D(nw.d_x0.ptr, x, n3, st);                                           // x0
                    launchAxpby(x, 4.0 / 3.0, nw.d_x0.ptr, -1.0 / 3.0, nw.d_xm1.ptr, n3, st);
                    launchAxpby(x, 1.0, x, 2.0 / 3.0 * h, v, n3, st);
                    copyD2D(nw.d_xm1.ptr, nw.d_x0.ptr, n3, st);                                // x-1 <- x0
                    grp->launchCounter += 3;
                } else if (implicitKind == SPGX_IE_STATIC_NEWTON || implicitKind == SPGX_IE_QUASI_STATIC_NEWTON) {
                    // static: K dx = -f, i.e. (0 M - 1 H) dx = -f; quasi-static: (M - h^2 K) dx = h^2 f; x += dx
                    const bool quasi = implicitKind == SPGX_IE_QUASI_STATIC_NEWTON;
                    const NewtonArgs a = newtonArgs(grp);
                    launchSystemForce(a, g, nw.d_f.ptr, st);
                    assembleHessian(grp, st);
                    launchAxpby(nw.d_rhs.ptr, quasi ? h * h : -1.0, nw.d_f.ptr, 0.0, nw.d_f.ptr, n3, st);
                    grp->launchCounter += 2;
                    phase(1);
                    nw.cgIterations.push_back(conjugateGradient(grp, quasi ? 1.0 : 0.0, quasi ? h * h : -1.0, nw.d_rhs.ptr, cgTolerance,
                                                                cgMaxIterations, st));
                    phase(2);
                    collect();
                    launchAxpby(x, 1.0, x, 1.0, nw.d_sol.ptr, n3, st);  // updateObjectsPositionsFromDx (implicitEulerBase.cpp:90-100)
                    grp->launchCounter += 1;
                } else {
                    // x0, v0; x += v h; f, K there; (M - h^2 K) dv = h f; x = x0; v = v0 + dv; x += v h
                    copyD2D(nw.d_x0.ptr, x, n3, st);
                    launchAxpby(x, 1.0, x, h, v, n3, st);
                    const NewtonArgs a = newtonArgs(grp);
                    launchSystemForce(a, g, nw.d_f.ptr, st);
                    assembleHessian(grp, st);
                    launchAxpby(nw.d_rhs.ptr, h, nw.d_f.ptr, 0.0, nw.d_f.ptr, n3, st);
                    grp->launchCounter += 3;
                    phase(1);
                    nw.cgIterations.push_back(conjugateGradient(grp, 1.0, h * h, nw.d_rhs.ptr, cgTolerance, cgMaxIterations, st));
                    phase(2);
                    collect();
                    launchAxpby(v, 1.0, v, 1.0, nw.d_sol.ptr, n3, st);
                    launchAxpby(x, 1.0, nw.d_x0.ptr, h, v, n3, st);
                    grp->launchCounter += 2;
                }
            }
        SPGX_CUDA(cudaEventRecord(grp->evStop, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        float ms = 0;
        SPGX_CUDA(cudaEventElapsedTime(&ms, grp->evStart, grp->evStop));
        grp->timings = spgx_timings{};
        grp->timings.inertial_ms = assembleMs;
        grp->timings.solve_ms = solveMs;
        grp->timings.update_ms = ms - assembleMs - solveMs;
        grp->timings.total_ms = ms;
        grp->timings.kernel_launches = grp->launchCounter;
    });
}

int spgx_symplectic_step(spgx_group *grp, double dt, int nsubsteps, const double *gravity, int nFrames, int sync)
{
    return guarded([&] {
        requireNewton(grp);
        if (nsubsteps < 1 || nFrames < 0) throw std::runtime_error("invalid step parameters");
        cudaStream_t st = grp->ctx->stream;
        auto &nw = grp->newton;
        const double g[3] = {gravity ? gravity[0] : 0.0, gravity ? gravity[1] : -9.8, gravity ? gravity[2] : 0.0};
        const double zero[3] = {0.0, 0.0, 0.0};
        const double h = dt / nsubsteps;
        grp->launchCounter = 0;
        SPGX_CUDA(cudaEventRecord(grp->evStart, st));
        const NewtonArgs a = newtonArgs(grp);
        for (int f = 0; f < nFrames; ++f)
            for (int s = 0; s < nsubsteps; ++s) {
                launchSystemForce(a, zero, nw.d_f.ptr, st);  // internal forces only: gravity acts on v directly
                launchSymplecticUpdate(grp->d_x.ptr, grp->d_v.ptr, nw.d_f.ptr, grp->d_invMass.ptr, grp->nVerts, h, g, st);
                grp->launchCounter += 2;
            }
        SPGX_CUDA(cudaEventRecord(grp->evStop, st));
        if (sync) {
            SPGX_CUDA(cudaStreamSynchronize(st));
            SPGX_CUDA(cudaGetLastError());
            float ms = 0;
            SPGX_CUDA(cudaEventElapsedTime(&ms, grp->evStart, grp->evStop));
            grp->timings = spgx_timings{};
            grp->timings.total_ms = ms;
            grp->timings.kernel_launches = grp->launchCounter;
        }
    });
}

int spgx_last_cg_iterations(spgx_group *grp, int32_t *out, int capacity, int *n)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        const auto &it = grp->newton.cgIterations;
        if (n) *n = static_cast<int>(it.size());
        for (int i = 0; i < static_cast<int>(it.size()) && i < capacity; ++i) out[i] = it[i];
    });
}

int spgx_system_force(spgx_group *grp, const double *gravity, double *f)
{
    return guarded([&] {
        requireNewton(grp);
        cudaStream_t st = grp->ctx->stream;
        const double g[3] = {gravity ? gravity[0] : 0.0, gravity ? gravity[1] : -9.8, gravity ? gravity[2] : 0.0};
        launchSystemForce(newtonArgs(grp), g, grp->newton.d_f.ptr, st);
        SPGX_CUDA(cudaMemcpyAsync(f, grp->newton.d_f.ptr, 3 * static_cast<size_t>(grp->nVerts) * sizeof(double),
                                  cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
    });
}

int spgx_system_stiffness(spgx_group *grp, int64_t *nBlocks, int32_t *rowptr, int32_t *cols, double *vals)
{
    return guarded([&] {
        requireNewton(grp);
        auto &nw = grp->newton;
        if (nBlocks) *nBlocks = nw.nnzb;
        if (!rowptr && !cols && !vals) return;
        cudaStream_t st = grp->ctx->stream;
        assembleHessian(grp, st);
        if (rowptr)
            for (int v = 0; v <= grp->nVerts; ++v) rowptr[v] = static_cast<int32_t>(nw.bsrRowptr[v]);
        if (cols) std::copy(nw.bsrCol.begin(), nw.bsrCol.end(), cols);
        if (vals) {
            // device storage is component-major; the caller gets block-major 3 x 3 blocks of K = -Hessian (energy.h:223)
            std::vector<double> soa(9 * static_cast<size_t>(nw.nnzb));
            SPGX_CUDA(cudaMemcpyAsync(soa.data(), nw.d_bsrH.ptr, soa.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaStreamSynchronize(st));
            const size_t nb = static_cast<size_t>(nw.nnzb);
            for (size_t t = 0; t < nb; ++t)
                for (int k = 0; k < 9; ++k) vals[9 * t + k] = -soa[k * nb + t];
        }
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
    });
}

int spgx_element_hessian(spgx_group *grp, int energyId, int stencil, double *hess)
{
    return guarded([&] {
        requireNewton(grp);
        if (energyId < 0 || energyId >= static_cast<int>(grp->energies.size())) throw std::runtime_error("energy id out of range");
        const auto &e = grp->energies[energyId];
        if (stencil < 0 || stencil >= e.n) throw std::runtime_error("stencil index out of range");
        cudaStream_t st = grp->ctx->stream;
        launchElementHessians(newtonArgs(grp), st);
        const int S = e.S, K = 3 * S;
        std::vector<double> blocks(static_cast<size_t>(9) * S * S);
        SPGX_CUDA(cudaMemcpyAsync(blocks.data(), grp->newton.d_elemH[energyId] + static_cast<size_t>(stencil) * blocks.size(),
                                  blocks.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        for (int la = 0; la < S; ++la)
            for (int lb = 0; lb < S; ++lb)
                for (int i = 0; i < 3; ++i)
                    for (int j = 0; j < 3; ++j) hess[(3 * la + i) * K + 3 * lb + j] = blocks[(la * S + lb) * 9 + 3 * i + j];
    });
}

int spgx_linear_solve(spgx_group *grp, double cM, double cH, const double *b, double *x, double cgTolerance,
                      int cgMaxIterations, int *iterations)
{
    return guarded([&] {
        requireNewton(grp);
        auto &nw = grp->newton;
        if (!nw.assembled) throw StateError("no Hessian assembled yet (call spgx_system_stiffness or spgx_implicit_step first)");
        cudaStream_t st = grp->ctx->stream;
        const size_t bytes = 3 * static_cast<size_t>(grp->nVerts) * sizeof(double);
        SPGX_CUDA(cudaMemcpyAsync(nw.d_rhs.ptr, b, bytes, cudaMemcpyHostToDevice, st));
        const int it = conjugateGradient(grp, cM, cH, nw.d_rhs.ptr, cgTolerance, cgMaxIterations, st);
        nw.cgIterations.assign(1, it);
        if (iterations) *iterations = it;
        SPGX_CUDA(cudaMemcpyAsync(x, nw.d_sol.ptr, bytes, cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
    });
}

}  // extern "C"
