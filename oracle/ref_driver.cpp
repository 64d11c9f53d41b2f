// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// C-ABI driver around the UNMODIFIED reference library (alexrodag/spg @ 75122ce), compiled by
// oracle/Makefile from the sources where they lie under /root/reference into oracle/_ref/libspgref.so.
// It is the strong oracle for the VBD / XPBD hot path: tests/ call it through ctypes to obtain the
// reference's own positions, velocities, colourings, incidence tables and per-element derivatives.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
//
// Everything in this file is glue written for this repository; no reference source is reproduced.
// The scene builders below re-state apps/solverCompare.cpp:118-189 (tet beam), :260-330 (createTetrahedralBeam),
// :332-365 (buildTriangleSquareMesh) and :367-498 (createCloth) headless, keeping their `float` arithmetic,
// and use the reference's real spg::TriangleMesh (spg/include/spg/geom/triangleMesh.h:36-50) so that the
// std::map edge order is the reference's own.
#include <spg/types.h>
#include <spg/constants.h>
#include <spg/geom/triangleMesh.h>
#include <spg/geom/tetrahedralMesh.h>
#include <spg/sim/simObject/particleGroup.h>
#include <spg/sim/simObject/rigidBodyGroup.h>
#include <spg/sim/energy/springEnergy.h>
#include <spg/sim/energy/springContinuumEnergy.h>
#include <spg/sim/energy/discreteBendingEnergy.h>
#include <spg/sim/energy/quadraticBendingEnergy.h>
#include <spg/sim/energy/membraneStvkEnergy.h>
#include <spg/sim/energy/membraneChoiEnergy.h>
#include <spg/sim/energy/membraneBaraffWitkinEnergy.h>
#include <spg/sim/energy/stvkEnergy.h>
#include <spg/sim/energy/stableNeoHookeanEnergy.h>
#include <spg/sim/energy/springAnchorEnergy.h>
#include <spg/sim/energy/baraffWitkinBendingEnergy.h>
#include <spg/sim/energy/springSquaredContinuumEnergy.h>
#include <spg/sim/energy/stableSquaredNeoHookeanEnergy.h>
#include <spg/sim/energy/staticLoadEnergy.h>
#include <spg/sim/energy/rigidBodySpringEnergy.h>
#include <spg/sim/energy/rigidBodySpringSquaredEnergy.h>
#include <spg/sim/energy/rigidBodySpringAnchorEnergy.h>
#include <spg/sim/energy/rigidBodyOrientationAlignmentEnergy.h>
#include <spg/sim/solver/vbd.h>
#include <spg/sim/solver/xpbd.h>
#include <spg/sim/solver/implicitEulerBaraffWitkin.h>
#include <spg/sim/solver/implicitEulerNewtonDx.h>
#include <spg/sim/solver/implicitEulerNewtonDv.h>
#include <spg/sim/solver/simplecticEuler.h>
#include <spg/sim/solver/bdf2.h>
#include <spg/sim/solver/staticNewton.h>
#include <spg/sim/solver/quasiStaticNewton.h>
#include <spg/sim/solver/implicitEulerNewtonRobust.h>
#include <spg/sim/solver/quasiStaticNewtonRobust.h>
#include <spg/utils/graphColoring.h>

#include <omp.h>

#include <algorithm>
#include <functional>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

namespace
{
// Energy kind numbering shared by oracle/, the product C ABI (include/spgx.h) and the tests.
enum Kind {
    K_SPRING = 0,
    K_SPRING_CONTINUUM = 1,
    K_DISCRETE_BENDING = 2,
    K_QUADRATIC_BENDING = 3,
    K_MEMBRANE_STVK = 4,
    K_MEMBRANE_CHOI = 5,
    K_MEMBRANE_BW = 6,
    K_TET_STVK = 7,
    K_TET_SNH = 8,
    K_ANCHOR = 9,
    K_BENDING_BW = 10,
    K_SPRING_SQUARED_CONTINUUM = 11,
    K_TET_SSNH = 12,
    K_STATIC_LOAD = 13,
    K_COUNT = 14
};
const int kStencilSize[K_COUNT] = {2, 2, 4, 4, 3, 3, 3, 4, 4, 1, 4, 2, 4, 1};
const int kNumModelParams[K_COUNT] = {2, 2, 2, 2, 2, 3, 3, 2, 2, 4, 2, 2, 2, 4};

thread_local std::string g_error;
}  // namespace
// a test may bring an energy of its own, written against the reference's authoring base class (tests/native/extra_energy_ref.h,
// Makefile target ref_extra): it supplies makeExtraEnergy / extraKindStencilSize / extraKindNumModelParams
#ifdef SPG_EXTRA_ENERGY_IMPL
#include SPG_EXTRA_ENERGY_IMPL
#endif
namespace
{

struct RefGroup {
    spg::ParticleGroup pg;
    std::vector<std::shared_ptr<spg::ParticleGroup::EnergyT>> energies;
    std::vector<int> kinds;
};

struct OracleVBD : spg::solver::VBD {
    void setIterations(int n) { m_iterations = n; }
    std::vector<std::vector<std::vector<int>>> &vertexGroups() { return m_simObjectsVertexGroups; }
    auto &elementsPerVertex() { return m_simObjectsElementsPerVertex; }
    void ensurePrecomputed()
    {
        if (m_precomputationUpdateRequired) {
            computeParallelVertexGroups();
            computeStencilInfoPerVertex();
            m_precomputationUpdateRequired = false;
        }
    }
    // Inject colour groups computed elsewhere (used only for meshes where the reference's O(N^2)
    // colouring is impractical, after equality has been proved on smaller meshes); the incidence
    // table is still the reference's own computeStencilInfoPerVertex.
    void injectVertexGroups(const std::vector<std::vector<int>> &groups)
    {
        m_simObjectsVertexGroups.clear();
        m_simObjectsVertexGroups.push_back(groups);
        computeStencilInfoPerVertex();
        m_precomputationUpdateRequired = false;
    }
};

struct OracleXPBD : spg::solver::XPBD {
    using spg::solver::XPBD::XPBD;
    void ensurePrecomputed()
    {
        if (m_useParallelGaussSeidel && m_parallelGroupsDirty) {
            computeParallelStencilGroups();
            m_parallelGroupsDirty = false;
        }
    }
    const std::vector<std::vector<int>> *groupsOf(void *energy)
    {
        auto it = m_stencilGroupsPerEnergy.find(energy);
        return it == m_stencilGroupsPerEnergy.end() ? nullptr : &it->second;
    }
    void injectGroups(void *energy, const std::vector<std::vector<int>> &groups)
    {
        m_stencilGroupsPerEnergy[energy] = groups;
        m_parallelGroupsDirty = false;
    }
};

// The implicit-Euler family (implicitEulerBase.h:13-25): the protected system builders and the linear solve made
// callable, nothing overridden.
template <class B>
struct OracleIE : B {
    void sysForce(spg::VectorX &f) const { this->getSystemForce(f); }
    void sysStiffness(spg::SparseMatrix &K) const { this->getSystemStiffnessMatrix(K); }
    void sysMass(spg::SparseMatrix &M) const { this->getSystemMassMatrix(M); }
    void linSolve(const spg::SparseMatrix &A, const spg::VectorX &b, spg::VectorX &x) const
    {
        this->solveLinearSystem(A, b, x);
    }
};

struct RefSolver {
    std::unique_ptr<OracleVBD> vbd;
    std::unique_ptr<OracleXPBD> xpbd;
    std::unique_ptr<spg::solver::BaseSolver> other;  // implicit-Euler family, symplectic Euler
    std::function<void(spg::VectorX &)> sysForce;
    std::function<void(spg::SparseMatrix &)> sysStiffness, sysMass;
    std::function<void(const spg::SparseMatrix &, const spg::VectorX &, spg::VectorX &)> linSolve;
    std::vector<int> cgIterations;  // of every linear solve of the last ref_solver_step call
    spg::solver::BaseSolver *base()
    {
        if (other) return other.get();
        return vbd ? static_cast<spg::solver::BaseSolver *>(vbd.get()) : xpbd.get();
    }
};

template <class B>
void makeImplicit(RefSolver &s)
{
    auto p = std::make_unique<OracleIE<B>>();
    auto *raw = p.get();
    s.sysForce = [raw](spg::VectorX &f) { raw->sysForce(f); };
    s.sysStiffness = [raw](spg::SparseMatrix &K) { raw->sysStiffness(K); };
    s.sysMass = [raw](spg::SparseMatrix &M) { raw->sysMass(M); };
    s.linSolve = [raw](const spg::SparseMatrix &A, const spg::VectorX &b, spg::VectorX &x) { raw->linSolve(A, b, x); };
    s.other = std::move(p);
}

// the reference reports the CG iteration count only through its Verbosity::Performance log line
// (implicitEulerBase.cpp:234-236): collect the numbers that follow "CG iterations: "
void parseCgIterations(const std::string &log, std::vector<int> &out)
{
    const std::string key = "CG iterations: ";
    for (size_t p = log.find(key); p != std::string::npos; p = log.find(key, p + 1))
        out.push_back(std::atoi(log.c_str() + p + key.size()));
}

template <class F>
int guarded(F &&f)
{
    try {
        f();
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return 1;
    } catch (...) {
        g_error = "unknown exception";
        return 2;
    }
}

std::shared_ptr<spg::ParticleGroup::EnergyT> makeEnergy(int kind, int n, const int32_t *st, const double *p)
{
    using namespace spg;
    switch (kind) {
    case K_SPRING: {
        auto e = std::make_shared<SpringEnergy>();
        for (int i = 0; i < n; ++i) e->addStencil({st[2 * i], st[2 * i + 1]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_SPRING_CONTINUUM: {
        auto e = std::make_shared<SpringContinuumEnergy>();
        for (int i = 0; i < n; ++i) e->addStencil({st[2 * i], st[2 * i + 1]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_DISCRETE_BENDING: {
        auto e = std::make_shared<DiscreteBendingEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[4 * i], st[4 * i + 1], st[4 * i + 2], st[4 * i + 3]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_QUADRATIC_BENDING: {
        auto e = std::make_shared<QuadraticBendingEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[4 * i], st[4 * i + 1], st[4 * i + 2], st[4 * i + 3]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_MEMBRANE_STVK: {
        auto e = std::make_shared<MembraneStvkEnergy>();
        for (int i = 0; i < n; ++i) e->addStencil({st[3 * i], st[3 * i + 1], st[3 * i + 2]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_MEMBRANE_CHOI: {
        auto e = std::make_shared<MembraneChoiEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[3 * i], st[3 * i + 1], st[3 * i + 2]}, p[3 * i], p[3 * i + 1], p[3 * i + 2]);
        return e;
    }
    case K_MEMBRANE_BW: {
        auto e = std::make_shared<MembraneBaraffWitkinEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[3 * i], st[3 * i + 1], st[3 * i + 2]}, p[3 * i], p[3 * i + 1], p[3 * i + 2]);
        return e;
    }
    case K_TET_STVK: {
        auto e = std::make_shared<StvkEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[4 * i], st[4 * i + 1], st[4 * i + 2], st[4 * i + 3]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_TET_SNH: {
        auto e = std::make_shared<StableNeoHookeanEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[4 * i], st[4 * i + 1], st[4 * i + 2], st[4 * i + 3]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_ANCHOR: {
        auto e = std::make_shared<SpringAnchorEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[i]}, Vector3(p[4 * i], p[4 * i + 1], p[4 * i + 2]), p[4 * i + 3]);
        return e;
    }
    case K_BENDING_BW: {
        auto e = std::make_shared<BaraffWitkinBendingEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[4 * i], st[4 * i + 1], st[4 * i + 2], st[4 * i + 3]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_SPRING_SQUARED_CONTINUUM: {
        auto e = std::make_shared<SpringSquaredContinuumEnergy>();
        for (int i = 0; i < n; ++i) e->addStencil({st[2 * i], st[2 * i + 1]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_TET_SSNH: {
        auto e = std::make_shared<StableSquaredNeoHookeanEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[4 * i], st[4 * i + 1], st[4 * i + 2], st[4 * i + 3]}, p[2 * i], p[2 * i + 1]);
        return e;
    }
    case K_STATIC_LOAD: {
        auto e = std::make_shared<StaticLoadEnergy>();
        for (int i = 0; i < n; ++i)
            e->addStencil({st[i]}, Vector3(p[4 * i], p[4 * i + 1], p[4 * i + 2]), p[4 * i + 3]);
        return e;
    }
    default:
#ifdef SPG_EXTRA_ENERGY_IMPL
        if (auto e = makeExtraEnergy(kind, n, st, p)) return e;
#endif
        throw std::runtime_error("unknown energy kind");
    }
}
}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI (ctypes-friendly). All functions returning int return 0 on success; ref_last_error() gives
// the std::exception message otherwise.
// ---------------------------------------------------------------------------------------------
extern "C" {

const char *ref_last_error()
{
    return g_error.c_str();
}

int ref_omp_max_threads()
{
    return omp_get_max_threads();
}

void ref_omp_set_threads(int n)
{
    omp_set_num_threads(n);
}

int ref_kind_stencil_size(int kind)
{
#ifdef SPG_EXTRA_ENERGY_IMPL
    if (kind >= K_COUNT) return extraKindStencilSize(kind);
#endif
    return (kind >= 0 && kind < K_COUNT) ? kStencilSize[kind] : -1;
}

int ref_kind_num_model_params(int kind)
{
#ifdef SPG_EXTRA_ENERGY_IMPL
    if (kind >= K_COUNT) return extraKindNumModelParams(kind);
#endif
    return (kind >= 0 && kind < K_COUNT) ? kNumModelParams[kind] : -1;
}

// ---- groups -------------------------------------------------------------------------------
void *ref_group_create(int nVerts, const double *x, const double *x0, const double *mass)
{
    auto *g = new RefGroup;
    for (int i = 0; i < nVerts; ++i) {
        g->pg.addParticle(spg::Vector3(x[3 * i], x[3 * i + 1], x[3 * i + 2]),
                          spg::Vector3(x0[3 * i], x0[3 * i + 1], x0[3 * i + 2]),
                          mass[i]);
    }
    return g;
}

void ref_group_destroy(void *h)
{
    delete static_cast<RefGroup *>(h);
}

// addStencil for every stencil with its model parameters, addEnergy, preparePrecomputations
// (the order createCloth / createTetrahedralBeam use: apps/solverCompare.cpp:460-461, :493-494, :326-328).
int ref_group_add_energy(void *h, int kind, int nStencils, const int32_t *stencils, const double *params)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        auto e = makeEnergy(kind, nStencils, stencils, params);
        g->pg.addEnergy(e);
        e->preparePrecomputations(g->pg);
        g->energies.push_back(e);
        g->kinds.push_back(kind);
    });
}

// SpringAnchorEnergy::updateAnchor (springAnchorEnergy.h:13); energies are shared_ptr, so solvers that copied the
// group see the edit (particleGroup.h:35)
int ref_update_anchor(void *h, int e, int stencil, const double *anchor3)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        auto a = std::dynamic_pointer_cast<spg::SpringAnchorEnergy>(g->energies.at(e));
        if (!a) throw std::runtime_error("energy is not a SpringAnchorEnergy");
        a->updateAnchor(stencil, spg::Vector3(anchor3[0], anchor3[1], anchor3[2]));
    });
}

int ref_group_size(void *h)
{
    return static_cast<RefGroup *>(h)->pg.size();
}

void ref_group_get_state(void *h, double *x, double *v)
{
    auto &pg = static_cast<RefGroup *>(h)->pg;
    for (int i = 0; i < pg.size(); ++i) {
        for (int d = 0; d < 3; ++d) {
            if (x) x[3 * i + d] = pg.positions()[i][d];
            if (v) v[3 * i + d] = pg.velocities()[i][d];
        }
    }
}

void ref_group_set_state(void *h, const double *x, const double *v)
{
    auto &pg = static_cast<RefGroup *>(h)->pg;
    for (int i = 0; i < pg.size(); ++i) {
        for (int d = 0; d < 3; ++d) {
            if (x) pg.positions()[i][d] = x[3 * i + d];
            if (v) pg.velocities()[i][d] = v[3 * i + d];
        }
    }
}

void ref_group_get_masses(void *h, double *m, double *w)
{
    auto &pg = static_cast<RefGroup *>(h)->pg;
    for (int i = 0; i < pg.size(); ++i) {
        if (m) m[i] = pg.masses()[i];
        if (w) w[i] = pg.invMasses()[i];
    }
}

// ---- per-element queries (energy.h:240-296) --------------------------------------------------
int ref_energy_value(void *h, int e, int stencil, double *out)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        *out = g->energies.at(e)->energy(stencil, g->pg);
    });
}

// grad: k doubles, hess: k*k doubles row-major (k = 3 * stencilSize)
int ref_energy_grad_hess(void *h, int e, int stencil, double *grad, double *hess)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        auto &en = g->energies.at(e);
        if (grad) {
            const spg::VectorX gr = en->energyGradientGeneric(stencil, g->pg);
            for (int i = 0; i < gr.size(); ++i) grad[i] = gr[i];
        }
        if (hess) {
            const spg::MatrixX H = en->energyHessianGeneric(stencil, g->pg);
            for (int i = 0; i < H.rows(); ++i)
                for (int j = 0; j < H.cols(); ++j) hess[i * H.cols() + j] = H(i, j);
        }
    });
}

// force += -grad block, hessian += diagonal block (what VBD calls per incidence, vbd.cpp:135-136)
int ref_vertex_force_hessian(void *h, int e, int stencil, int local, double *force3, double *hess9)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        auto &en = g->energies.at(e);
        spg::Vector3 f = spg::Vector3::Zero();
        spg::Matrix3 H = spg::Matrix3::Zero();
        en->accumulateVertexForce(stencil, local, g->pg, f);
        en->accumulateVertexHessian(stencil, local, g->pg, H);
        for (int i = 0; i < 3; ++i) {
            force3[i] = f[i];
            for (int j = 0; j < 3; ++j) hess9[3 * i + j] = H(i, j);
        }
    });
}

// Constraint values C (nC) and gradient (k x nC, row-major) through the public constraintsAndGradient.
extern "C++" template <class E>
bool tryConstraints(spg::ParticleGroup::EnergyT *en, spg::ParticleGroup &pg, int stencil, double *C, double *grad)
{
    auto *p = dynamic_cast<E *>(en);
    if (!p) return false;
    auto [c, gr] = p->constraintsAndGradient(stencil, pg);
    for (int i = 0; i < c.size(); ++i) C[i] = c[i];
    for (int i = 0; i < gr.rows(); ++i)
        for (int j = 0; j < gr.cols(); ++j) grad[i * gr.cols() + j] = gr(i, j);
    return true;
}

int ref_constraints_and_gradient(void *h, int e, int stencil, double *C, double *grad)
{
    return guarded([&] {
        using namespace spg;
        auto *g = static_cast<RefGroup *>(h);
        auto *en = g->energies.at(e).get();
        bool ok = tryConstraints<SpringEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<SpringContinuumEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<DiscreteBendingEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<QuadraticBendingEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<MembraneStvkEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<MembraneChoiEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<MembraneBaraffWitkinEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<StvkEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<StableNeoHookeanEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<SpringAnchorEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<BaraffWitkinBendingEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<SpringSquaredContinuumEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<StableSquaredNeoHookeanEnergy>(en, g->pg, stencil, C, grad) ||
                  tryConstraints<StaticLoadEnergy>(en, g->pg, stencil, C, grad);
#ifdef SPG_EXTRA_ENERGY_IMPL
        ok = ok || tryConstraints<SPG_EXTRA_ENERGY_TYPE>(en, g->pg, stencil, C, grad);
#endif
        if (!ok) throw std::runtime_error("unsupported energy type");
    });
}

// One XPBD projection of one stencil on the group's own state (energy.h:117-153, springEnergy.cpp:48-68).
int ref_project_position(void *h, int e, int stencil, double dt)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        g->energies.at(e)->projectPosition(stencil, g->pg, dt);
    });
}

// Effective stiffness / compliance matrices of stencil i (nC x nC row-major each).
extern "C++" template <class E>
bool tryStiffness(spg::ParticleGroup::EnergyT *en, int i, double *K, double *A)
{
    auto *p = dynamic_cast<E *>(en);
    if (!p) return false;
    const auto &k = p->effectiveStiffness()[i];
    const auto &a = p->effectiveCompliance()[i];
    for (int r = 0; r < k.rows(); ++r)
        for (int c = 0; c < k.cols(); ++c) {
            K[r * k.cols() + c] = k(r, c);
            A[r * k.cols() + c] = a(r, c);
        }
    return true;
}

int ref_effective_stiffness(void *h, int e, int stencil, double *K, double *A)
{
    return guarded([&] {
        using namespace spg;
        auto *g = static_cast<RefGroup *>(h);
        auto *en = g->energies.at(e).get();
        bool ok = tryStiffness<SpringEnergy>(en, stencil, K, A) || tryStiffness<SpringContinuumEnergy>(en, stencil, K, A) ||
                  tryStiffness<DiscreteBendingEnergy>(en, stencil, K, A) ||
                  tryStiffness<QuadraticBendingEnergy>(en, stencil, K, A) ||
                  tryStiffness<MembraneStvkEnergy>(en, stencil, K, A) ||
                  tryStiffness<MembraneChoiEnergy>(en, stencil, K, A) ||
                  tryStiffness<MembraneBaraffWitkinEnergy>(en, stencil, K, A) ||
                  tryStiffness<StvkEnergy>(en, stencil, K, A) || tryStiffness<StableNeoHookeanEnergy>(en, stencil, K, A) ||
                  tryStiffness<SpringAnchorEnergy>(en, stencil, K, A) || tryStiffness<BaraffWitkinBendingEnergy>(en, stencil, K, A) ||
                  tryStiffness<SpringSquaredContinuumEnergy>(en, stencil, K, A) ||
                  tryStiffness<StableSquaredNeoHookeanEnergy>(en, stencil, K, A) || tryStiffness<StaticLoadEnergy>(en, stencil, K, A);
#ifdef SPG_EXTRA_ENERGY_IMPL
        ok = ok || tryStiffness<SPG_EXTRA_ENERGY_TYPE>(en, stencil, K, A);
#endif
        if (!ok) throw std::runtime_error("unsupported energy type");
    });
}

// Kind-specific precomputed rest data: discrete bending -> restTheta (1); quadratic bending -> 3x12
// laplacian row-major (36); membranes -> 2x2 Dm^-1 row-major (4); tets -> 3x3 Dm^-1 row-major (9)
// (+ SNH alpha as 10th value). Returns the number of doubles written through *n.
int ref_rest_data(void *h, int e, int stencil, double *out, int *n)
{
    return guarded([&] {
        using namespace spg;
        auto *g = static_cast<RefGroup *>(h);
        auto *en = g->energies.at(e).get();
        *n = 0;
        auto put2 = [&](const Matrix2 &m) {
            for (int r = 0; r < 2; ++r)
                for (int c = 0; c < 2; ++c) out[(*n)++] = m(r, c);
        };
        auto put3 = [&](const Matrix3 &m) {
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) out[(*n)++] = m(r, c);
        };
        if (auto *p = dynamic_cast<DiscreteBendingEnergy *>(en)) {
            out[(*n)++] = p->restThetas()[stencil];
        } else if (auto *p = dynamic_cast<QuadraticBendingEnergy *>(en)) {
            const auto &L = p->laplacians()[stencil];
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 12; ++c) out[(*n)++] = L(r, c);
        } else if (auto *p = dynamic_cast<MembraneStvkEnergy *>(en)) {
            put2(p->inverseReferenceMats()[stencil]);
        } else if (auto *p = dynamic_cast<MembraneChoiEnergy *>(en)) {
            put2(p->inverseReferenceMats()[stencil]);
        } else if (auto *p = dynamic_cast<MembraneBaraffWitkinEnergy *>(en)) {
            put2(p->inverseReferenceMats()[stencil]);
        } else if (auto *p = dynamic_cast<StvkEnergy *>(en)) {
            put3(p->inverseReferenceMats()[stencil]);
        } else if (auto *p = dynamic_cast<StableNeoHookeanEnergy *>(en)) {
            put3(p->inverseReferenceMats()[stencil]);
            out[(*n)++] = p->hydrostaticCorrectionTerms()[stencil];
        } else if (auto *p = dynamic_cast<BaraffWitkinBendingEnergy *>(en)) {
            out[(*n)++] = p->restThetas()[stencil];
        } else if (auto *p = dynamic_cast<StableSquaredNeoHookeanEnergy *>(en)) {
            put3(p->inverseReferenceMats()[stencil]);
        } else if (auto *p = dynamic_cast<StaticLoadEnergy *>(en)) {
            for (int c = 0; c < 3; ++c) out[(*n)++] = p->loadDirections()[stencil][c];
        }
    });
}

// ---- colouring (graphColoring.cpp:51-84) ---------------------------------------------------
int ref_color_vertices(void *h, int32_t *colors)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        std::vector<spg::coloring::FlatStencils> set;
        for (const auto &e : g->pg.energies()) set.push_back({e->flatStencils(), e->stencilSize()});
        const auto c = spg::coloring::colorVertices(g->pg.size(), set);
        std::copy(c.begin(), c.end(), colors);
    });
}

int ref_color_stencils(void *h, int e, int32_t *colors)
{
    return guarded([&] {
        auto *g = static_cast<RefGroup *>(h);
        auto &en = g->energies.at(e);
        const auto c = spg::coloring::colorStencils({en->flatStencils(), en->stencilSize()});
        std::copy(c.begin(), c.end(), colors);
    });
}

// Raw colouring entry points on caller-provided flat stencils (no group needed).
int ref_color_vertices_flat(int nVerts, int nSets, const int32_t *const *entries, const int32_t *nEntries,
                            const int32_t *stencilSizes, int32_t *colors)
{
    return guarded([&] {
        std::vector<spg::coloring::FlatStencils> set;
        for (int s = 0; s < nSets; ++s)
            set.push_back({std::vector<int>(entries[s], entries[s] + nEntries[s]), stencilSizes[s]});
        const auto c = spg::coloring::colorVertices(nVerts, set);
        std::copy(c.begin(), c.end(), colors);
    });
}

int ref_color_stencils_flat(const int32_t *entries, int nEntries, int stencilSize, int32_t *colors)
{
    return guarded([&] {
        const auto c = spg::coloring::colorStencils({std::vector<int>(entries, entries + nEntries), stencilSize});
        std::copy(c.begin(), c.end(), colors);
    });
}

// ---- solvers ----------------------------------------------------------------------------------
// kind: 0 = VBD (vbd.cpp:12-195), 1 = XPBD parallel Gauss-Seidel (xpbd.cpp:16-131), 2 = XPBD serial,
// 3 = ImplicitEulerBaraffWitkin (implicitEulerBaraffWitkin.cpp:11-59), 4 = ImplicitEulerNewtonDx
// (implicitEulerNewtonDx.cpp:11-64), 5 = ImplicitEulerNewtonDv (implicitEulerNewtonDv.cpp:11-67),
// 6 = SimplecticEuler (simplecticEuler.cpp:12-62), 7 = BDF2 (bdf2.cpp:10-76), 8 = StaticNewton (staticNewton.cpp:10-51),
// 9 = QuasiStaticNewton (quasiStaticNewton.cpp:10-55), 10 = ImplicitEulerNewtonRobust (implicitEulerNewtonRobust.cpp:11-137),
// 11 = QuasiStaticNewtonRobust (quasiStaticNewtonRobust.cpp:11-151).
void *ref_solver_create(int kind, void *group, double dt, int substeps, int iterations, const double *gravity)
{
    auto *s = new RefSolver;
    auto *g = static_cast<RefGroup *>(group);
    if (kind == 0) {
        s->vbd = std::make_unique<OracleVBD>();
        s->vbd->setIterations(iterations);
    } else if (kind == 3) {
        makeImplicit<spg::solver::ImplicitEulerBaraffWitkin>(*s);
    } else if (kind == 4) {
        makeImplicit<spg::solver::ImplicitEulerNewtonDx>(*s);
    } else if (kind == 5) {
        makeImplicit<spg::solver::ImplicitEulerNewtonDv>(*s);
    } else if (kind == 6) {
        s->other = std::make_unique<spg::solver::SimplecticEuler>();
    } else if (kind == 7) {
        makeImplicit<spg::solver::BDF2>(*s);
    } else if (kind == 8) {
        makeImplicit<spg::solver::StaticNewton>(*s);
    } else if (kind == 9) {
        makeImplicit<spg::solver::QuasiStaticNewton>(*s);
    } else if (kind == 10) {
        makeImplicit<spg::solver::ImplicitEulerNewtonRobust>(*s);
    } else if (kind == 11) {
        makeImplicit<spg::solver::QuasiStaticNewtonRobust>(*s);
    } else {
        s->xpbd = std::make_unique<OracleXPBD>(kind == 1);
    }
    s->base()->addObject(g->pg);  // copies the group; energies are shared (baseSolver.cpp:10-14)
    s->base()->setDt(dt);
    s->base()->setNumSubsteps(substeps);
    if (gravity) s->base()->setGravity(spg::Vector3(gravity[0], gravity[1], gravity[2]));
    return s;
}

void ref_solver_destroy(void *h)
{
    delete static_cast<RefSolver *>(h);
}

int ref_solver_step(void *h, int nFrames)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        // VBD::computeParallelVertexGroups prints unconditionally (vbd.cpp:228); silence it.
        std::ostringstream sink;
        auto *old = std::cout.rdbuf(sink.rdbuf());
        if (s->linSolve) s->base()->setVerbosity(spg::Verbosity::Performance);  // the only way to learn the CG iteration count
        try {
            for (int f = 0; f < nFrames; ++f) s->base()->step();
        } catch (...) {
            std::cout.rdbuf(old);
            throw;
        }
        std::cout.rdbuf(old);
        s->base()->setVerbosity(spg::Verbosity::None);
        s->cgIterations.clear();
        if (s->linSolve) parseCgIterations(sink.str(), s->cgIterations);
    });
}

// ---- live topology edits (picking: apps/solverCompare.cpp:717-726, :1446-1455, :1471-1473, :1499-1502) ----------
// The application adds / removes an energy on the SOLVER's copy of the group and flags the solver
// (VBD::requirePrecomputationUpdate, vbd.h:26; XPBD has the same hook, xpbd.h:20).
int ref_solver_add_energy(void *h, int kind, int nStencils, const int32_t *stencils, const double *params)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        auto &pg = s->base()->particleGroups()[0];
        auto e = makeEnergy(kind, nStencils, stencils, params);
        pg.addEnergy(e);
        e->preparePrecomputations(pg);
        if (s->vbd) s->vbd->requirePrecomputationUpdate();
        if (s->xpbd) s->xpbd->requirePrecomputationUpdate();
    });
}

int ref_solver_remove_energy(void *h, int e)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        auto &pg = s->base()->particleGroups()[0];
        auto victim = pg.energies().at(e);  // (a copy: removeEnergy compares against it while erasing)
        pg.removeEnergy(victim);
        if (s->vbd) s->vbd->requirePrecomputationUpdate();
        if (s->xpbd) s->xpbd->requirePrecomputationUpdate();
    });
}

int ref_solver_update_anchor(void *h, int e, int stencil, const double *anchor3)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        auto a = std::dynamic_pointer_cast<spg::SpringAnchorEnergy>(s->base()->particleGroups()[0].energies().at(e));
        if (!a) throw std::runtime_error("energy is not a SpringAnchorEnergy");
        a->updateAnchor(stencil, spg::Vector3(anchor3[0], anchor3[1], anchor3[2]));
    });
}

// ---- implicit-Euler family probes (implicitEulerBase.cpp) ----------------------------------------
// CG iteration counts of the linear solves of the last ref_solver_step call (one per substep and frame)
int ref_solver_cg_iterations(void *h, int32_t *out, int capacity, int *n)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        *n = static_cast<int>(s->cgIterations.size());
        for (int i = 0; i < *n && i < capacity; ++i) out[i] = s->cgIterations[i];
    });
}
// getSystemForce (implicitEulerBase.cpp:102-143) on the solver's current state: 3V doubles
int ref_system_force(void *h, double *f)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        if (!s->sysForce) throw std::runtime_error("not an implicit-Euler solver");
        spg::VectorX fv(3 * s->base()->particleGroups()[0].size());
        s->sysForce(fv);
        std::copy(fv.data(), fv.data() + fv.size(), f);
    });
}
// getSystemStiffnessMatrix (:176-221) or getSystemMassMatrix (:145-174) after setFromTriplets, as COO sorted by
// (column, row) (the compressed column-major matrix walked in storage order). rows == NULL: only *nnz.
int ref_system_matrix(void *h, int which, int64_t *nnz, int32_t *rows, int32_t *cols, double *vals)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        if (!s->sysStiffness) throw std::runtime_error("not an implicit-Euler solver");
        const int n = 3 * s->base()->particleGroups()[0].size();
        spg::SparseMatrix A(n, n);
        if (which == 0)
            s->sysStiffness(A);
        else
            s->sysMass(A);
        int64_t k = 0;
        for (int c = 0; c < A.outerSize(); ++c)
            for (spg::SparseMatrix::InnerIterator it(A, c); it; ++it, ++k)
                if (rows) rows[k] = static_cast<int32_t>(it.row()), cols[k] = static_cast<int32_t>(it.col()), vals[k] = it.value();
        *nnz = k;
    });
}
// solveLinearSystem (:223-238): Eigen::ConjugateGradient<SparseMatrix, Lower | Upper>, Jacobi preconditioner,
// tolerance 1e-5, at most 300 iterations (implicitEulerBase.h:22-23), on a caller-supplied COO matrix
int ref_linear_solve(void *h, int n, int64_t nnz, const int32_t *rows, const int32_t *cols, const double *vals,
                     const double *b, double *x, int *iterations)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        if (!s->linSolve) throw std::runtime_error("not an implicit-Euler solver");
        std::vector<spg::Triplet> trip;
        trip.reserve(nnz);
        for (int64_t k = 0; k < nnz; ++k) trip.emplace_back(rows[k], cols[k], vals[k]);
        spg::SparseMatrix A(n, n);
        A.setFromTriplets(trip.begin(), trip.end());
        spg::VectorX bv = Eigen::Map<const spg::VectorX>(b, n), xv;
        std::ostringstream sink;
        auto *old = std::cout.rdbuf(sink.rdbuf());
        s->base()->setVerbosity(spg::Verbosity::Performance);
        try {
            s->linSolve(A, bv, xv);
        } catch (...) {
            std::cout.rdbuf(old);
            throw;
        }
        std::cout.rdbuf(old);
        s->base()->setVerbosity(spg::Verbosity::None);
        std::vector<int> it;
        parseCgIterations(sink.str(), it);
        if (iterations) *iterations = it.empty() ? -1 : it.back();
        std::copy(xv.data(), xv.data() + n, x);
    });
}

int ref_solver_reset(void *h)
{
    return guarded([&] { static_cast<RefSolver *>(h)->base()->reset(); });
}

void ref_solver_get_state(void *h, double *x, double *v)
{
    auto &pg = static_cast<RefSolver *>(h)->base()->particleGroups()[0];
    for (int i = 0; i < pg.size(); ++i) {
        for (int d = 0; d < 3; ++d) {
            if (x) x[3 * i + d] = pg.positions()[i][d];
            if (v) v[3 * i + d] = pg.velocities()[i][d];
        }
    }
}

void ref_solver_set_state(void *h, const double *x, const double *v)
{
    auto &pg = static_cast<RefSolver *>(h)->base()->particleGroups()[0];
    for (int i = 0; i < pg.size(); ++i) {
        for (int d = 0; d < 3; ++d) {
            if (x) pg.positions()[i][d] = x[3 * i + d];
            if (v) pg.velocities()[i][d] = v[3 * i + d];
        }
    }
}

// VBD colour of every vertex as the solver itself grouped them (vbd.cpp:203-229).
int ref_vbd_vertex_colors(void *h, int32_t *colors, int *nColors)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        std::ostringstream sink;
        auto *old = std::cout.rdbuf(sink.rdbuf());
        s->vbd->ensurePrecomputed();
        std::cout.rdbuf(old);
        const auto &groups = s->vbd->vertexGroups().at(0);
        *nColors = static_cast<int>(groups.size());
        for (int c = 0; c < static_cast<int>(groups.size()); ++c)
            for (int v : groups[c]) colors[v] = c;
    });
}

// Incidence table of the solver (vbd.cpp:231-268) flattened: for vertex v, energy e the entries
// [rowptr[v*nE+e], rowptr[v*nE+e+1]) of (stencilIdx, localIdx) pairs. Pass entries == nullptr to size.
int ref_vbd_incidence(void *h, int64_t *rowptr, int32_t *entries, int64_t *nEntries)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        std::ostringstream sink;
        auto *old = std::cout.rdbuf(sink.rdbuf());
        s->vbd->ensurePrecomputed();
        std::cout.rdbuf(old);
        const auto &epv = s->vbd->elementsPerVertex().at(0);
        int64_t n = 0;
        size_t row = 0;
        for (const auto &perEnergy : epv) {
            for (const auto &list : perEnergy) {
                if (rowptr) rowptr[row] = n;
                ++row;
                for (const auto &ent : list) {
                    if (entries) {
                        entries[2 * n] = ent[0];
                        entries[2 * n + 1] = ent[1];
                    }
                    ++n;
                }
            }
        }
        if (rowptr) rowptr[row] = n;
        *nEntries = n;
    });
}

int ref_vbd_inject_colors(void *h, const int32_t *colors, int nVerts)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        std::vector<std::vector<int>> groups;
        for (int i = 0; i < nVerts; ++i) {
            if (static_cast<int>(groups.size()) <= colors[i]) groups.resize(colors[i] + 1);
            groups[colors[i]].push_back(i);
        }
        s->vbd->injectVertexGroups(groups);
    });
}

int ref_xpbd_stencil_colors(void *h, int e, int32_t *colors, int *nColors)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        s->xpbd->ensurePrecomputed();
        auto &pg = s->xpbd->particleGroups()[0];
        const auto *groups = s->xpbd->groupsOf(pg.energies().at(e).get());
        if (!groups) throw std::runtime_error("no stencil groups for energy");
        *nColors = static_cast<int>(groups->size());
        for (int c = 0; c < static_cast<int>(groups->size()); ++c)
            for (int i : (*groups)[c]) colors[i] = c;
    });
}

int ref_xpbd_inject_colors(void *h, int e, const int32_t *colors, int nStencils)
{
    return guarded([&] {
        auto *s = static_cast<RefSolver *>(h);
        auto &pg = s->xpbd->particleGroups()[0];
        std::vector<std::vector<int>> groups;
        for (int i = 0; i < nStencils; ++i) {
            if (static_cast<int>(groups.size()) <= colors[i]) groups.resize(colors[i] + 1);
            groups[colors[i]].push_back(i);
        }
        s->xpbd->injectGroups(pg.energies().at(e).get(), groups);
    });
}

// ---- headless scene builders ---------------------------------------------------------------
// Re-statements of the solverCompare scene builders that only emit arrays (positions, rest positions,
// masses, and per-energy stencils + model parameters in insertion order). They exist to pin the
// numpy scene generators in spg_b200/scenes.py bit for bit. float arithmetic is kept where the
// reference uses float (apps/solverCompare.cpp:118-135, :332-347).
struct RefSceneEnergy {
    int kind;
    std::vector<int32_t> stencils;
    std::vector<double> params;
};
struct RefScene {
    std::vector<double> x, x0, mass;
    std::vector<RefSceneEnergy> energies;
};

void ref_scene_destroy(void *h)
{
    delete static_cast<RefScene *>(h);
}
int ref_scene_num_vertices(void *h)
{
    return static_cast<int>(static_cast<RefScene *>(h)->mass.size());
}
int ref_scene_num_energies(void *h)
{
    return static_cast<int>(static_cast<RefScene *>(h)->energies.size());
}
int ref_scene_energy_kind(void *h, int e)
{
    return static_cast<RefScene *>(h)->energies.at(e).kind;
}
int ref_scene_energy_num_stencils(void *h, int e)
{
    const auto &en = static_cast<RefScene *>(h)->energies.at(e);
    return static_cast<int>(en.stencils.size()) / kStencilSize[en.kind];
}
void ref_scene_get_vertices(void *h, double *x, double *x0, double *mass)
{
    auto *s = static_cast<RefScene *>(h);
    std::copy(s->x.begin(), s->x.end(), x);
    std::copy(s->x0.begin(), s->x0.end(), x0);
    std::copy(s->mass.begin(), s->mass.end(), mass);
}
void ref_scene_get_energy(void *h, int e, int32_t *stencils, double *params)
{
    const auto &en = static_cast<RefScene *>(h)->energies.at(e);
    std::copy(en.stencils.begin(), en.stencils.end(), stencils);
    std::copy(en.params.begin(), en.params.end(), params);
}

// Cloth (apps/solverCompare.cpp:332-498). membraneKind in {K_MEMBRANE_STVK, K_MEMBRANE_CHOI, K_MEMBRANE_BW}
// or K_SPRING / K_SPRING_CONTINUUM for the config-2 style scene (a spring of stiffness 1e2 on every mesh
// edge, rest length from the rest pose); bendingKind in {K_DISCRETE_BENDING, K_QUADRATIC_BENDING}.
void *ref_scene_cloth(int nRows, int nCols, int membraneKind, int bendingKind)
{
    const float mass = 1.0f, width = 1.0f, height = 1.0f;
    const float rowStep = height / (nRows - 1);
    const float colStep = width / (nCols - 1);
    std::vector<spg::Vector3> verts, verts0;
    std::vector<spg::Int3> faces;
    for (int r = 0; r < nRows; ++r) {
        for (int c = 0; c < nCols; ++c) {
            verts0.push_back({colStep * c, rowStep * r, 0});
            verts.push_back({colStep * c, 0, -rowStep * r});
            if (r == 0 || c == 0) continue;
            const int a = (r - 1) * nCols + c - 1, b = (r - 1) * nCols + c, d = r * nCols + c - 1, e = r * nCols + c;
            if ((r + c) % 2 == 0) {
                faces.push_back({d, a, b});
                faces.push_back({d, b, e});
            } else {
                faces.push_back({a, e, d});
                faces.push_back({a, b, e});
            }
        }
    }
    const spg::TriangleMesh mesh(verts, verts0, faces);
    auto *s = new RefScene;
    const int nV = static_cast<int>(verts.size());
    std::vector<double> areaPerVertex(nV, 0.0);
    double totalArea = 0;
    for (const auto &f : faces) {
        const spg::Vector3 e0 = verts0[f[1]] - verts0[f[0]];
        const spg::Vector3 e3 = verts0[f[2]] - verts0[f[1]];
        const double area = 0.5 * (e0.cross(e3).norm());
        totalArea += area;
        for (int k = 0; k < 3; ++k) areaPerVertex[f[k]] += area / 3.0;
    }
    const double density = mass / totalArea;
    for (int i = 0; i < nV; ++i) {
        for (int d = 0; d < 3; ++d) {
            s->x.push_back(verts[i][d]);
            s->x0.push_back(verts0[i][d]);
        }
        s->mass.push_back(areaPerVertex[i] * density);
    }
    RefSceneEnergy bend{bendingKind, {}, {}};
    const auto &edges = mesh.edges();
    const auto &opp = mesh.opposingVerticesPerEdge();
    for (size_t i = 0; i < edges.size(); ++i) {
        if (opp[i].size() != 2) continue;
        for (int v : {edges[i][0], edges[i][1], opp[i][0].vertexId, opp[i][1].vertexId}) bend.stencils.push_back(v);
        bend.params.push_back(0.0);
        bend.params.push_back(1e-2);
    }
    s->energies.push_back(std::move(bend));
    RefSceneEnergy mem{membraneKind, {}, {}};
    if (membraneKind == K_SPRING || membraneKind == K_SPRING_CONTINUUM) {
        for (const auto &ed : edges) {
            mem.stencils.push_back(ed[0]);
            mem.stencils.push_back(ed[1]);
            mem.params.push_back((verts0[ed[0]] - verts0[ed[1]]).norm());
            mem.params.push_back(1e2);
        }
    } else {
        for (const auto &f : faces) {
            for (int k = 0; k < 3; ++k) mem.stencils.push_back(f[k]);
            if (membraneKind == K_MEMBRANE_STVK) {
                mem.params.push_back(1e2);
                mem.params.push_back(0.0);
            } else {
                for (int k = 0; k < 3; ++k) mem.params.push_back(1e2);
            }
        }
    }
    s->energies.push_back(std::move(mem));
    RefSceneEnergy anchors{K_ANCHOR, {}, {}};
    for (int id : {0, nRows - 1}) {
        anchors.stencils.push_back(id);
        for (int d = 0; d < 3; ++d) anchors.params.push_back(verts[id][d]);
        anchors.params.push_back(1e5);
    }
    s->energies.push_back(std::move(anchors));
    return s;
}

// Tetrahedral beam (apps/solverCompare.cpp:118-189, :260-330) as createSceneObjects calls it
// (:679-690): mass 1, sides 4x1x1, young 1e3f, poisson 0.2f. tetKind in {K_TET_STVK, K_TET_SNH}.
void *ref_scene_beam(int nX, int nY, int nZ, int tetKind)
{
    const float mass = 1.0f, sideX = 4, sideY = 1, sideZ = 1, young = 1e3F, poisson = 0.2F;
    std::vector<spg::Vector3> verts;
    for (int i = 0; i < nX; ++i)
        for (int j = 0; j < nY; ++j)
            for (int k = 0; k < nZ; ++k)
                verts.push_back({-sideX * 0.5f + i * sideX / (nX - 1),
                                 -sideY * 0.5f + j * sideY / (nY - 1),
                                 -sideZ * 0.5f + k * sideZ / (nZ - 1)});
    auto volume = [&](int a, int b, int c, int d) {
        spg::Matrix3 m;
        m.col(0) = verts[b] - verts[a];
        m.col(1) = verts[c] - verts[a];
        m.col(2) = verts[d] - verts[a];
        return m.determinant() / 6.0;
    };
    std::vector<spg::Int4> tets;
    auto push = [&](int a, int b, int c, int d) {
        if (volume(a, b, c, d) > 0)
            tets.push_back({a, b, c, d});
        else
            tets.push_back({b, a, c, d});
    };
    for (int i = 0; i < nX - 1; ++i)
        for (int j = 0; j < nY - 1; ++j)
            for (int k = 0; k < nZ - 1; ++k) {
                int id[8];  // corner (di, dj, dk) -> id[di + 2*dk + 4*dj], the reference's id1..id8 numbering
                for (int dj = 0; dj < 2; ++dj)
                    for (int dk = 0; dk < 2; ++dk)
                        for (int di = 0; di < 2; ++di)
                            id[di + 2 * dk + 4 * dj] = (i + di) * nY * nZ + (j + dj) * nZ + k + dk;
                if (i % 2 == 1)
                    for (int q : {0, 2, 4, 6}) std::swap(id[q], id[q + 1]);
                if (j % 2 == 1)
                    for (int q : {0, 1, 2, 3}) std::swap(id[q], id[q + 4]);
                if (k % 2 == 1)
                    for (int q : {0, 1, 4, 5}) std::swap(id[q], id[q + 2]);
                push(id[0], id[2], id[3], id[6]);
                push(id[6], id[7], id[3], id[5]);
                push(id[3], id[1], id[0], id[5]);
                push(id[0], id[4], id[6], id[5]);
                push(id[6], id[3], id[5], id[0]);
            }
    const spg::TetrahedralMesh mesh(verts, verts, tets);
    auto *s = new RefScene;
    const int nV = static_cast<int>(verts.size());
    std::vector<double> volPerVertex(nV, 0.0);
    double totalVolume = 0;
    for (const auto &t : tets) {
        const double vol = volume(t[0], t[1], t[2], t[3]);
        totalVolume += vol;
        for (int k = 0; k < 4; ++k) volPerVertex[t[k]] += vol / 4.0;
    }
    const double density = mass / totalVolume;
    for (int i = 0; i < nV; ++i) {
        for (int d = 0; d < 3; ++d) {
            s->x.push_back(verts[i][d]);
            s->x0.push_back(verts[i][d]);
        }
        s->mass.push_back(volPerVertex[i] * density);
    }
    RefSceneEnergy anchors{K_ANCHOR, {}, {}};
    for (int j = 0; j < nY; ++j)
        for (int k = 0; k < nZ; ++k) {
            const int id = j * nZ + k;
            anchors.stencils.push_back(id);
            for (int d = 0; d < 3; ++d) anchors.params.push_back(verts[id][d]);
            anchors.params.push_back(1e4);
        }
    s->energies.push_back(std::move(anchors));
    RefSceneEnergy tet{tetKind, {}, {}};
    for (const auto &t : tets) {
        for (int k = 0; k < 4; ++k) tet.stencils.push_back(t[k]);
        tet.params.push_back(young);
        tet.params.push_back(poisson);
    }
    s->energies.push_back(std::move(tet));
    return s;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Rigid bodies: the reference's RigidBodyGroup and its four energies under the reference's own
// VBD / XPBD.  Kind numbering = enum spgx_rb_kind of include/spgx.h.
// ---------------------------------------------------------------------------------------------
namespace
{
struct RefRbGroup {
    spg::RigidBodyGroup rb;
    std::vector<std::shared_ptr<spg::RigidBodyGroup::EnergyT>> energies;
};
struct RefRbSolver {
    std::unique_ptr<OracleVBD> vbd;
    std::unique_ptr<OracleXPBD> xpbd;
    spg::solver::BaseSolver *base() { return vbd ? static_cast<spg::solver::BaseSolver *>(vbd.get()) : xpbd.get(); }
};
spg::Vector3 v3(const double *p) { return spg::Vector3(p[0], p[1], p[2]); }
}  // namespace

extern "C" {
void *ref_rb_group_create(int n, const double *x, const double *theta, const double *mass, const double *localInertia)
{
    auto *g = new RefRbGroup;
    // a visual mesh is carried per body (rigidBodyGroup.h:99) and never read by the solvers: one triangle
    const std::vector<spg::Vector3> verts{{0, 0, 0}, {1, 0, 0}, {0, 1, 0}};
    const spg::TriangleMesh mesh(verts, verts, {spg::Int3{0, 1, 2}});
    for (int i = 0; i < n; ++i) {
        spg::Matrix3 I;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) I(r, c) = localInertia[9 * i + 3 * r + c];
        g->rb.addBody(v3(x + 3 * i), v3(x + 3 * i), mass[i], v3(theta + 3 * i), v3(theta + 3 * i), I, mesh);
    }
    return g;
}
void ref_rb_group_destroy(void *h) { delete static_cast<RefRbGroup *>(h); }

int ref_rb_group_add_energy(void *h, int kind, int n, const int32_t *st, const double *p)
{
    return guarded([&] {
        auto *g = static_cast<RefRbGroup *>(h);
        std::shared_ptr<spg::RigidBodyGroup::EnergyT> out;
        if (kind == 0) {
            auto e = std::make_shared<spg::RBSpringEnergy>();
            for (int i = 0; i < n; ++i)
                e->addStencil({st[2 * i], st[2 * i + 1]}, {v3(p + 8 * i), v3(p + 8 * i + 3)}, p[8 * i + 6], p[8 * i + 7]);
            out = e;
        } else if (kind == 1) {
            auto e = std::make_shared<spg::RBSpringSquaredEnergy>();
            for (int i = 0; i < n; ++i)
                e->addStencil({st[2 * i], st[2 * i + 1]}, {v3(p + 8 * i), v3(p + 8 * i + 3)}, p[8 * i + 6], p[8 * i + 7]);
            out = e;
        } else if (kind == 2) {
            auto e = std::make_shared<spg::RBSpringAnchorEnergy>();
            for (int i = 0; i < n; ++i) e->addStencil({st[i]}, v3(p + 7 * i), v3(p + 7 * i + 3), p[7 * i + 6]);
            out = e;
        } else if (kind == 3) {
            auto e = std::make_shared<spg::RBOrientationAlignmentEnergy>();
            for (int i = 0; i < n; ++i) e->addStencil({st[2 * i], st[2 * i + 1]}, p[i]);
            out = e;
        } else {
            throw std::runtime_error("unknown rigid-body energy kind");
        }
        g->rb.addEnergy(out);
        g->energies.push_back(out);
    });
}

// state of a group (before it is handed to a solver) or of the solver's own copy
static void rbGetState(const spg::RigidBodyGroup &rb, double *x, double *theta, double *v, double *omega, double *R)
{
    for (int i = 0; i < rb.size(); ++i)
        for (int c = 0; c < 3; ++c) {
            if (x) x[3 * i + c] = rb.positions()[i][c];
            if (theta) theta[3 * i + c] = rb.thetas()[i][c];
            if (v) v[3 * i + c] = rb.velocities()[i][c];
            if (omega) omega[3 * i + c] = rb.omegas()[i][c];
            if (R)
                for (int d = 0; d < 3; ++d) R[9 * i + 3 * c + d] = rb.rotationMatrices()[i](c, d);
        }
}
static void rbSetState(spg::RigidBodyGroup &rb, const double *x, const double *theta, const double *v, const double *omega)
{
    const int n = rb.size();
    spg::VectorX pos(6 * n), vel(6 * n);
    rb.getPositions(pos, 0);
    rb.getVelocities(vel, 0);
    for (int i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) {
            if (x) pos[6 * i + c] = x[3 * i + c];
            if (theta) pos[6 * i + 3 + c] = theta[3 * i + c];
            if (v) vel[6 * i + c] = v[3 * i + c];
            if (omega) vel[6 * i + 3 + c] = omega[3 * i + c];
        }
    rb.setPositions(pos, 0);  // rebuilds rotation matrices and inertias (rigidBodyGroup.cpp:43-53)
    rb.setVelocities(vel, 0);
}
void ref_rb_group_get_state(void *h, double *x, double *theta, double *v, double *omega, double *R)
{
    rbGetState(static_cast<RefRbGroup *>(h)->rb, x, theta, v, omega, R);
}
void ref_rb_group_set_state(void *h, const double *x, const double *theta, const double *v, const double *omega)
{
    rbSetState(static_cast<RefRbGroup *>(h)->rb, x, theta, v, omega);
}

// constraintsAndGradient (C: 1, grad: 6 S) through the public virtual
extern "C++" template <class E>
bool tryRbConstraints(spg::RigidBodyGroup::EnergyT *en, spg::RigidBodyGroup &rb, int stencil, double *C, double *grad)
{
    auto *p = dynamic_cast<E *>(en);
    if (!p) return false;
    auto [c, gr] = p->constraintsAndGradient(stencil, rb);
    C[0] = c[0];
    for (int i = 0; i < gr.rows(); ++i) grad[i] = gr(i, 0);
    return true;
}
int ref_rb_constraint(void *h, int e, int stencil, double *C, double *grad)
{
    return guarded([&] {
        auto *g = static_cast<RefRbGroup *>(h);
        auto *en = g->energies.at(e).get();
        const bool ok = tryRbConstraints<spg::RBSpringEnergy>(en, g->rb, stencil, C, grad) ||
                        tryRbConstraints<spg::RBSpringSquaredEnergy>(en, g->rb, stencil, C, grad) ||
                        tryRbConstraints<spg::RBSpringAnchorEnergy>(en, g->rb, stencil, C, grad) ||
                        tryRbConstraints<spg::RBOrientationAlignmentEnergy>(en, g->rb, stencil, C, grad);
        if (!ok) throw std::runtime_error("unsupported energy type");
    });
}
// what VBD accumulates per incidence (vbd.cpp:135-136): force (6) and the 6 x 6 diagonal block, row-major
int ref_rb_vertex_force_hessian(void *h, int e, int stencil, int local, double *force6, double *hess36)
{
    return guarded([&] {
        auto *g = static_cast<RefRbGroup *>(h);
        auto &en = g->energies.at(e);
        spg::Vector<6> f = spg::Vector<6>::Zero();
        spg::Matrix<6, 6> H = spg::Matrix<6, 6>::Zero();
        en->accumulateVertexForce(stencil, local, g->rb, f);
        en->accumulateVertexHessian(stencil, local, g->rb, H);
        for (int i = 0; i < 6; ++i) {
            force6[i] = f[i];
            for (int j = 0; j < 6; ++j) hess36[6 * i + j] = H(i, j);
        }
    });
}

// kind: 0 VBD, 1 XPBD (parallel Gauss-Seidel), 2 XPBD (serial)
void *ref_rb_solver_create(int kind, void *group, double dt, int substeps, int iterations, const double *gravity)
{
    auto *s = new RefRbSolver;
    auto *g = static_cast<RefRbGroup *>(group);
    if (kind == 0) {
        s->vbd = std::make_unique<OracleVBD>();
        s->vbd->setIterations(iterations);
    } else {
        s->xpbd = std::make_unique<OracleXPBD>(kind == 1);
    }
    s->base()->addObject(g->rb);
    s->base()->setDt(dt);
    s->base()->setNumSubsteps(substeps);
    if (gravity) s->base()->setGravity(spg::Vector3(gravity[0], gravity[1], gravity[2]));
    return s;
}
void ref_rb_solver_destroy(void *h) { delete static_cast<RefRbSolver *>(h); }
int ref_rb_solver_step(void *h, int nFrames)
{
    return guarded([&] {
        auto *s = static_cast<RefRbSolver *>(h);
        std::ostringstream sink;  // VBD::computeParallelVertexGroups prints unconditionally (vbd.cpp:228)
        auto *old = std::cout.rdbuf(sink.rdbuf());
        try {
            for (int f = 0; f < nFrames; ++f) s->base()->step();
        } catch (...) {
            std::cout.rdbuf(old);
            throw;
        }
        std::cout.rdbuf(old);
    });
}
void ref_rb_solver_get_state(void *h, double *x, double *theta, double *v, double *omega, double *R)
{
    rbGetState(static_cast<RefRbSolver *>(h)->base()->rbGroups()[0], x, theta, v, omega, R);
}
void ref_rb_solver_set_state(void *h, const double *x, const double *theta, const double *v, const double *omega)
{
    rbSetState(static_cast<RefRbSolver *>(h)->base()->rbGroups()[0], x, theta, v, omega);
}
int ref_rb_solver_update_anchor(void *h, void *group, int e, int stencil, const double *anchor3)
{
    return guarded([&] {
        (void)h;  // energies are shared between the group and the solver's copy (rigidBodyGroup.h: shared_ptr)
        auto *a = dynamic_cast<spg::RBSpringAnchorEnergy *>(static_cast<RefRbGroup *>(group)->energies.at(e).get());
        if (!a) throw std::runtime_error("not an anchor energy");
        a->updateAnchor(stencil, v3(anchor3));
    });
}
}  // extern "C"
