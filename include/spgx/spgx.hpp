// spgx.hpp — C++ facade over the spgx C ABI (include/spgx.h) that re-creates the reference's class surface for
// the hot path, so that a caller of alexrodag/spg switches by changing the namespace:
//
//   spg::ParticleGroup                      -> spgx::ParticleGroup            (spg/include/spg/sim/simObject/particleGroup.h:14-60)
//   spg::SpringEnergy, ... (10 classes)     -> spgx::SpringEnergy, ...        (spg/include/spg/sim/energy/*.h: addStencil signatures)
//   spg::solver::VBD / spg::solver::XPBD    -> spgx::solver::VBD / XPBD       (spg/include/spg/sim/solver/{baseSolver,vbd,xpbd}.h)
//
// Semantics kept from the reference: addObject() copies the group into the solver (baseSolver.cpp:10-14); energies
// are shared_ptr and processed in insertion order (vbd.cpp:128, xpbd.cpp:84); step() is blocking; errors are
// std::runtime_error carrying the reference's message; requirePrecomputationUpdate() re-colours on the next step
// (vbd.h:26, xpbd.h:20); VBD::setIterations is the knob the reference leaves as a TODO (vbd.h:33).
// Difference that a GPU imposes: the solver's state lives in HBM. particleGroups()[i].positions() returns a host
// mirror that is refreshed lazily (download on first access after a step) and pushed back to the device before
// the next step if it was obtained through the non-const accessor (callers write positions, solverCompare.cpp:1388).
// Header only; link with spg_b200/lib/libspgx.so.  There is no CPU fallback.
#pragma once
#include <array>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../spgx.h"

namespace spgx
{
using Real = double;
struct Vector3 {
    Real x{0}, y{0}, z{0};
    Vector3() = default;
    Vector3(Real a, Real b, Real c) : x(a), y(b), z(c) {}
    Real &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const Real &operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
static_assert(sizeof(Vector3) == 3 * sizeof(double), "Vector3 must be three packed doubles");

enum class Verbosity { None, Performance, Debug };  // spg/include/spg/constants.h:5

inline void check(int status)
{
    if (status != SPGX_OK) throw std::runtime_error(spgx_last_error());
}

// ---- energies: host-side stencil + model-parameter lists with the reference's addStencil signatures -------------
class Energy
{
public:
    virtual ~Energy() = default;
    int kind() const { return m_kind; }
    int nStencils() const { return static_cast<int>(m_params.size()) / spgx_kind_num_model_params(m_kind); }
    int stencilSize() const { return spgx_kind_stencil_size(m_kind); }
    const std::vector<int> &flatStencils() const { return m_stencils; }  // energy.h:45
    const std::vector<Real> &modelParams() const { return m_params; }
    const std::string &name() const { return m_name; }
    // preparePrecomputations(obj) of the reference (energy.h:54-61) happens inside spgx_group_add_energy when
    // the group is handed to a solver; calling it here only validates, with the reference's messages.
    void preparePrecomputations(const class ParticleGroup &obj) const;
    unsigned long long revision() const { return m_revision; }

protected:
    Energy(int kind, const char *name) : m_kind(kind), m_name(name) {}
    template <size_t S, size_t P>
    void push(const std::array<int, S> &stencil, const std::array<Real, P> &params)
    {
        m_stencils.insert(m_stencils.end(), stencil.begin(), stencil.end());
        m_params.insert(m_params.end(), params.begin(), params.end());
        ++m_revision;
    }
    int m_kind;
    std::string m_name;
    std::vector<int> m_stencils;
    std::vector<Real> m_params;
    unsigned long long m_revision{0};
    friend class ParticleGroup;
};

#define SPGX_ENERGY_2(NAME, KIND, S, A, B)                                                  \
    class NAME : public Energy                                                              \
    {                                                                                       \
    public:                                                                                 \
        static constexpr int s_stencilSize = S;                                             \
        NAME() : Energy(KIND, #NAME) {}                                                     \
        void addStencil(const std::array<int, S> &stencil, Real A, Real B)                  \
        {                                                                                   \
            push<S, 2>(stencil, {A, B});                                                    \
        }                                                                                   \
    }
#define SPGX_ENERGY_3(NAME, KIND, S, A, B, C)                                               \
    class NAME : public Energy                                                              \
    {                                                                                       \
    public:                                                                                 \
        static constexpr int s_stencilSize = S;                                             \
        NAME() : Energy(KIND, #NAME) {}                                                     \
        void addStencil(const std::array<int, S> &stencil, Real A, Real B, Real C)          \
        {                                                                                   \
            push<S, 3>(stencil, {A, B, C});                                                 \
        }                                                                                   \
    }
SPGX_ENERGY_2(SpringEnergy, SPGX_SPRING, 2, restLength, stiffness);                    // springEnergy.h:16
SPGX_ENERGY_2(SpringContinuumEnergy, SPGX_SPRING_CONTINUUM, 2, restLength, stiffness); // springContinuumEnergy.h
SPGX_ENERGY_2(DiscreteBendingEnergy, SPGX_DISCRETE_BENDING, 4, restTheta, stiffness);  // discreteBendingEnergy.h:13
SPGX_ENERGY_2(QuadraticBendingEnergy, SPGX_QUADRATIC_BENDING, 4, restTheta, stiffness);
SPGX_ENERGY_2(MembraneStvkEnergy, SPGX_MEMBRANE_STVK, 3, young, poisson);              // membraneStvkEnergy.h:13
SPGX_ENERGY_3(MembraneChoiEnergy, SPGX_MEMBRANE_CHOI, 3, kWeft, kWarp, kBias);
SPGX_ENERGY_3(MembraneBaraffWitkinEnergy, SPGX_MEMBRANE_BW, 3, kWeft, kWarp, kBias);
SPGX_ENERGY_2(StvkEnergy, SPGX_TET_STVK, 4, young, poisson);
SPGX_ENERGY_2(StableNeoHookeanEnergy, SPGX_TET_SNH, 4, young, poisson);
SPGX_ENERGY_2(BaraffWitkinBendingEnergy, SPGX_BENDING_BW, 4, restTheta, stiffness);    // baraffWitkinBendingEnergy.h
SPGX_ENERGY_2(SpringSquaredContinuumEnergy, SPGX_SPRING_SQUARED_CONTINUUM, 2, restLength, stiffness);
SPGX_ENERGY_2(StableSquaredNeoHookeanEnergy, SPGX_TET_SSNH, 4, young, poisson);
#undef SPGX_ENERGY_2
#undef SPGX_ENERGY_3

class SpringAnchorEnergy : public Energy  // springAnchorEnergy.h:6-22
{
public:
    static constexpr int s_stencilSize = 1;
    SpringAnchorEnergy() : Energy(SPGX_ANCHOR, "SpringAnchorEnergy") {}
    void addStencil(std::array<int, 1> stencil, const Vector3 &anchor, Real stiffness)
    {
        push<1, 4>(stencil, {anchor.x, anchor.y, anchor.z, stiffness});
    }
    // live edit (solverCompare.cpp:1524): picked up by every solver holding this energy at its next step()
    void updateAnchor(int i, const Vector3 &anchor)
    {
        m_params[4 * i] = anchor.x, m_params[4 * i + 1] = anchor.y, m_params[4 * i + 2] = anchor.z;
        ++m_anchorRevision;
    }
    Vector3 anchor(int i) const { return {m_params[4 * i], m_params[4 * i + 1], m_params[4 * i + 2]}; }
    unsigned long long anchorRevision() const { return m_anchorRevision; }

protected:
    unsigned long long m_anchorRevision{0};  // solvers compare it with the revision they last uploaded
};

class StaticLoadEnergy : public Energy  // staticLoadEnergy.h:7-28 (energy only: XPBD::step throws, as in the reference)
{
public:
    static constexpr int s_stencilSize = 1;
    StaticLoadEnergy() : Energy(SPGX_STATIC_LOAD, "StaticLoadEnergy") {}
    void addStencil(std::array<int, 1> stencil, const Vector3 &loadDirection, Real stiffness)
    {
        push<1, 4>(stencil, {loadDirection.x, loadDirection.y, loadDirection.z, stiffness});
    }
};

// ---- ParticleGroup (particleGroup.h:14-60) ------------------------------------------------------------------------
class ParticleGroup
{
public:
    static constexpr int s_nDOFs = 3;
    using EnergyT = Energy;
    const std::vector<Vector3> &positions() const { return m_x; }
    std::vector<Vector3> &positions() { return m_x; }
    const std::vector<Vector3> &positions0() const { return m_x0; }
    const std::vector<Vector3> &initialPositions() const { return m_xInitial; }  // what reset() restores
    std::vector<Vector3> &velocities() { return m_v; }
    const std::vector<Vector3> &velocities() const { return m_v; }
    const std::vector<Real> &invMasses() const { return m_w; }
    const std::vector<Real> &masses() const { return m_m; }
    std::vector<std::shared_ptr<EnergyT>> &energies() { return m_energies; }
    const std::vector<std::shared_ptr<EnergyT>> &energies() const { return m_energies; }
    int nDOF() const { return static_cast<int>(m_x.size()) * 3; }
    int size() const { return static_cast<int>(m_x.size()); }
    void addParticle(const Vector3 &position, const Vector3 &position0, Real mass)  // particleGroup.cpp:74-82
    {
        m_x.push_back(position);
        m_xInitial.push_back(position);
        m_x0.push_back(position0);
        m_v.emplace_back(0, 0, 0);
        m_m.push_back(mass);
        m_w.push_back(mass == 0 ? 0 : 1 / mass);
    }
    void addEnergy(std::shared_ptr<EnergyT> energy) { m_energies.push_back(std::move(energy)); }
    void removeEnergy(const std::shared_ptr<EnergyT> &energy)  // particleGroup.cpp:89-95
    {
        for (auto it = m_energies.begin(); it != m_energies.end(); ++it)
            if (*it == energy) {
                m_energies.erase(it);
                return;
            }
    }
    void scaleMasses(Real f)  // particleGroup.cpp:105-113
    {
        for (size_t i = 0; i < m_m.size(); ++i) {
            m_m[i] *= f;
            m_w[i] = m_m[i] == 0 ? 0 : 1 / m_m[i];
        }
    }
    void setName(std::string name) { m_name = std::move(name); }
    void reset()  // particleGroup.cpp:97-103
    {
        m_x = m_xInitial;
        for (auto &v : m_v) v = Vector3(0, 0, 0);
    }

protected:
    std::string m_name;
    std::vector<Vector3> m_x, m_xInitial, m_x0, m_v;
    std::vector<Real> m_m, m_w;
    std::vector<std::shared_ptr<EnergyT>> m_energies;
};

inline void Energy::preparePrecomputations(const ParticleGroup &obj) const
{
    std::vector<Real> records(static_cast<size_t>(nStencils()) * spgx_kind_num_record_params(m_kind));
    check(spgx_prepare_energy(m_kind, nStencils(), m_stencils.data(), m_params.data(), obj.size(),
                              reinterpret_cast<const Real *>(obj.positions0().data()), records.data()));
}

// ---- rigid bodies (rigidBodyGroup.h:14-111 and the four energies of spg/include/spg/sim/energy/rigidBody*.h) --------------
using Matrix3 = std::array<Real, 9>;  // row-major 3 x 3

class RigidBodyEnergy  // Energy<RigidBodyGroup>: host-side stencil + model-parameter lists
{
public:
    virtual ~RigidBodyEnergy() = default;
    int kind() const { return m_kind; }
    int nStencils() const { return static_cast<int>(m_params.size()) / spgx_rb_kind_num_model_params(m_kind); }
    int stencilSize() const { return spgx_rb_kind_stencil_size(m_kind); }
    const std::vector<int> &flatStencils() const { return m_stencils; }
    const std::vector<Real> &modelParams() const { return m_params; }
    const std::string &name() const { return m_name; }
    unsigned long long revision() const { return m_revision; }

protected:
    RigidBodyEnergy(int kind, const char *name) : m_kind(kind), m_name(name) {}
    template <size_t S>
    void push(const std::array<int, S> &stencil, std::initializer_list<Real> params)
    {
        m_stencils.insert(m_stencils.end(), stencil.begin(), stencil.end());
        m_params.insert(m_params.end(), params.begin(), params.end());
        ++m_revision;
    }
    int m_kind;
    std::string m_name;
    std::vector<int> m_stencils;
    std::vector<Real> m_params;
    unsigned long long m_revision{0};
};

class RBSpringEnergy : public RigidBodyEnergy  // rigidBodySpringEnergy.h:8-22
{
public:
    static constexpr int s_stencilSize = 2;
    RBSpringEnergy() : RigidBodyEnergy(SPGX_RB_SPRING, "RBSpringEnergy") {}
    void addStencil(std::array<int, 2> stencil, const std::array<Vector3, 2> &localRigidBodyPoints, Real restLength, Real stiffness)
    {
        const auto &a = localRigidBodyPoints[0], &b = localRigidBodyPoints[1];
        push<2>(stencil, {a.x, a.y, a.z, b.x, b.y, b.z, restLength, stiffness});
    }
};
class RBSpringSquaredEnergy : public RigidBodyEnergy  // rigidBodySpringSquaredEnergy.h
{
public:
    static constexpr int s_stencilSize = 2;
    RBSpringSquaredEnergy() : RigidBodyEnergy(SPGX_RB_SPRING_SQUARED, "RBSpringSquaredEnergy") {}
    void addStencil(std::array<int, 2> stencil, const std::array<Vector3, 2> &localRigidBodyPoints, Real restLength, Real stiffness)
    {
        const auto &a = localRigidBodyPoints[0], &b = localRigidBodyPoints[1];
        push<2>(stencil, {a.x, a.y, a.z, b.x, b.y, b.z, restLength, stiffness});
    }
};
class RBSpringAnchorEnergy : public RigidBodyEnergy  // rigidBodySpringAnchorEnergy.h:8-30
{
public:
    static constexpr int s_stencilSize = 1;
    RBSpringAnchorEnergy() : RigidBodyEnergy(SPGX_RB_SPRING_ANCHOR, "RBSpringAnchorEnergy") {}
    void addStencil(std::array<int, 1> stencil, const Vector3 &localRigidBodyPoint, const Vector3 &anchor, Real stiffness)
    {
        push<1>(stencil, {localRigidBodyPoint.x, localRigidBodyPoint.y, localRigidBodyPoint.z, anchor.x, anchor.y, anchor.z, stiffness});
    }
    void updateAnchor(int i, const Vector3 &anchor)  // picked up by every solver holding this energy at its next step()
    {
        m_params[7 * i + 3] = anchor.x, m_params[7 * i + 4] = anchor.y, m_params[7 * i + 5] = anchor.z;
        ++m_anchorRevision;
    }
    Vector3 anchor(int i) const { return {m_params[7 * i + 3], m_params[7 * i + 4], m_params[7 * i + 5]}; }
    unsigned long long anchorRevision() const { return m_anchorRevision; }

protected:
    unsigned long long m_anchorRevision{0};
};
class RBOrientationAlignmentEnergy : public RigidBodyEnergy  // rigidBodyOrientationAlignmentEnergy.h
{
public:
    static constexpr int s_stencilSize = 2;
    RBOrientationAlignmentEnergy() : RigidBodyEnergy(SPGX_RB_ORIENTATION_ALIGNMENT, "RBOrientationAlignmentEnergy") {}
    void addStencil(std::array<int, 2> stencil, Real stiffness) { push<2>(stencil, {stiffness}); }
};

class RigidBodyGroup  // rigidBodyGroup.h:14-111 (host mirror; the solver owns the device copy)
{
public:
    static constexpr int s_nDOFs = 6;
    using EnergyT = RigidBodyEnergy;
    const std::vector<Vector3> &positions() const { return m_x; }
    std::vector<Vector3> &positions() { return m_x; }
    const std::vector<Vector3> &positions0() const { return m_x0; }
    const std::vector<Vector3> &thetas() const { return m_theta; }
    std::vector<Vector3> &thetas() { return m_theta; }
    const std::vector<Vector3> &thetas0() const { return m_theta0; }
    std::vector<Vector3> &velocities() { return m_v; }
    const std::vector<Vector3> &velocities() const { return m_v; }
    std::vector<Vector3> &omegas() { return m_omega; }
    const std::vector<Vector3> &omegas() const { return m_omega; }
    const std::vector<Real> &masses() const { return m_m; }
    const std::vector<Real> &invMasses() const { return m_w; }
    const std::vector<Matrix3> &localInertias() const { return m_localInertia; }
    // refreshed from the device together with the positions (the reference keeps them in step on the host)
    const std::vector<Matrix3> &rotationMatrices() const { return m_rotationMatrix; }
    std::vector<Matrix3> &rotationMatrices() { return m_rotationMatrix; }
    const std::vector<Vector3> &initialPositions() const { return m_xInitial; }
    const std::vector<Vector3> &initialThetas() const { return m_thetaInitial; }
    std::vector<std::shared_ptr<EnergyT>> &energies() { return m_energies; }
    const std::vector<std::shared_ptr<EnergyT>> &energies() const { return m_energies; }
    int nDOF() const { return static_cast<int>(m_x.size()) * 6; }
    int size() const { return static_cast<int>(m_x.size()); }
    // addBody (rigidBodyGroup.cpp:138-160); the visual mesh of the reference's signature is a rendering concern and not kept
    void addBody(const Vector3 &p, const Vector3 &p0, Real mass, const Vector3 &theta, const Vector3 &theta0, const Matrix3 &localInertia)
    {
        m_x.push_back(p), m_xInitial.push_back(p), m_x0.push_back(p0);
        m_v.emplace_back(0, 0, 0), m_omega.emplace_back(0, 0, 0);
        m_m.push_back(mass), m_w.push_back(mass == 0 ? 0 : 1 / mass);
        m_theta.push_back(theta), m_thetaInitial.push_back(theta), m_theta0.push_back(theta0);
        m_localInertia.push_back(localInertia);
        m_rotationMatrix.push_back(Matrix3{1, 0, 0, 0, 1, 0, 0, 0, 1});  // replaced by the solver's matrices at the first sync
    }
    void addEnergy(std::shared_ptr<EnergyT> energy) { m_energies.push_back(std::move(energy)); }
    void removeEnergy(const std::shared_ptr<EnergyT> &energy)  // rigidBodyGroup.cpp:167-178
    {
        for (auto it = m_energies.begin(); it != m_energies.end(); ++it)
            if (*it == energy) {
                m_energies.erase(it);
                return;
            }
    }
    void setName(std::string name) { m_name = std::move(name); }
    void reset()  // rigidBodyGroup.cpp:180-192
    {
        m_x = m_xInitial;
        m_theta = m_thetaInitial;
        for (auto &v : m_v) v = Vector3(0, 0, 0);
        for (auto &w : m_omega) w = Vector3(0, 0, 0);
    }

protected:
    std::string m_name;
    std::vector<Vector3> m_x, m_xInitial, m_x0, m_v, m_theta, m_thetaInitial, m_theta0, m_omega;
    std::vector<Matrix3> m_rotationMatrix, m_localInertia;
    std::vector<Real> m_m, m_w;
    std::vector<std::shared_ptr<EnergyT>> m_energies;
};

namespace solver
{
// BaseSolver (baseSolver.h:14-41) over device-resident copies of the particle groups
class BaseSolver
{
public:
    explicit BaseSolver(int device = 0) { check(spgx_ctx_create(device, &m_ctx)); }
    virtual ~BaseSolver()
    {
        for (auto &s : m_slots) spgx_group_destroy(s.handle);
        for (auto &s : m_rbSlots) spgx_rb_group_destroy(s.handle);
        spgx_ctx_destroy(m_ctx);
    }
    BaseSolver(const BaseSolver &) = delete;
    BaseSolver &operator=(const BaseSolver &) = delete;

    virtual void step() = 0;
    void addObject(const ParticleGroup &object)  // copies, baseSolver.cpp:10-14
    {
        m_groups.push_back(object);
        m_slots.emplace_back();
        m_slots.back().hostDirty = false;
    }
    void addObject(const RigidBodyGroup &object)  // baseSolver.cpp:35
    {
        m_rbGroups.push_back(object);
        m_rbSlots.emplace_back();
    }
    std::vector<RigidBodyGroup> &rbGroups()  // baseSolver.h:23
    {
        syncToHost();
        for (auto &s : m_rbSlots) s.hostDirty = true;
        return m_rbGroups;
    }
    const std::vector<RigidBodyGroup> &rbGroups() const
    {
        const_cast<BaseSolver *>(this)->syncToHost();
        return m_rbGroups;
    }
    // host mirrors; the non-const accessor marks them for upload before the next step
    std::vector<ParticleGroup> &particleGroups()
    {
        syncToHost();
        for (auto &s : m_slots) s.hostDirty = true;
        return m_groups;
    }
    const std::vector<ParticleGroup> &particleGroups() const
    {
        const_cast<BaseSolver *>(this)->syncToHost();
        return m_groups;
    }
    int numSimObjects() { return static_cast<int>(m_groups.size() + m_rbGroups.size()); }
    void setDt(Real dt) { m_dtStep = dt; }
    Real dt() { return m_dtStep; }
    void setNumSubsteps(int n) { m_nsubsteps = n; }
    int numSubsteps() { return m_nsubsteps; }
    virtual void reset()  // baseSolver.cpp:23-33
    {
        syncToHost();
        for (size_t i = 0; i < m_groups.size(); ++i) {
            m_groups[i].reset();
            m_slots[i].hostDirty = true;
        }
        for (size_t i = 0; i < m_rbGroups.size(); ++i) {
            m_rbGroups[i].reset();
            m_rbSlots[i].hostDirty = true;
        }
        m_time = 0;
    }
    void setVerbosity(Verbosity v) { m_verbosity = v; }
    void setGravity(const Vector3 &g) { m_gravity = g; }
    // the four phase times of the last step() (vbd.cpp:184-194 / xpbd.cpp:125-130), ms
    spgx_timings lastTimings(int group = 0)
    {
        spgx_timings t{};
        check(spgx_last_timings(m_slots.at(group).handle, &t));
        return t;
    }

protected:
    struct Slot {
        spgx_group *handle{nullptr};
        bool built{false};
        bool hostDirty{false};   // host mirror is newer than the device
        bool deviceDirty{false}; // device is newer than the host mirror
        std::vector<unsigned long long> revisions, anchorRevisions;
        std::vector<const Energy *> identities;
    };
    virtual int solverKind() const = 0;
    virtual int stepGroup(spgx_group *g, const Real *gravity) = 0;
    // rigid bodies run under VBD and XPBD (the implicit-Euler family's rigid-body branches are not built)
    virtual int rbIterations() const { return 1; }
    struct RbSlot {
        spgx_rb_group *handle{nullptr};
        bool built{false}, hostDirty{false}, deviceDirty{false};
        std::vector<unsigned long long> revisions, anchorRevisions;
        std::vector<const RigidBodyEnergy *> identities;
    };
    void buildRb(size_t i)
    {
        RigidBodyGroup &g = m_rbGroups[i];
        RbSlot &s = m_rbSlots[i];
        if (s.handle) {
            syncToHost();
            spgx_rb_group_destroy(s.handle);
            s.handle = nullptr;
        }
        check(spgx_rb_group_create(m_ctx, g.size(), reinterpret_cast<const Real *>(g.initialPositions().data()),
                                   reinterpret_cast<const Real *>(g.initialThetas().data()), g.masses().data(),
                                   reinterpret_cast<const Real *>(g.localInertias().data()), &s.handle));
        s.identities.clear(), s.revisions.clear(), s.anchorRevisions.clear();
        for (auto &e : g.energies()) {
            check(spgx_rb_group_add_energy(s.handle, e->kind(), e->nStencils(), e->flatStencils().data(), e->modelParams().data(), nullptr));
            s.identities.push_back(e.get());
            s.revisions.push_back(e->revision());
            auto *a = dynamic_cast<RBSpringAnchorEnergy *>(e.get());
            s.anchorRevisions.push_back(a ? a->anchorRevision() : 0);
        }
        check(spgx_rb_group_finalize(s.handle, solverKind()));
        check(spgx_rb_upload(s.handle, reinterpret_cast<const Real *>(g.positions().data()), reinterpret_cast<const Real *>(g.thetas().data()),
                             reinterpret_cast<const Real *>(g.velocities().data()), reinterpret_cast<const Real *>(g.omegas().data())));
        s.built = true;
        s.hostDirty = s.deviceDirty = false;
    }
    void stepRbGroups(const Real *gravity)
    {
        for (size_t i = 0; i < m_rbGroups.size(); ++i) {
            RigidBodyGroup &g = m_rbGroups[i];
            RbSlot &s = m_rbSlots[i];
            bool changed = g.energies().size() != s.identities.size();
            for (size_t e = 0; e < g.energies().size() && !changed; ++e)
                changed = g.energies()[e].get() != s.identities[e] || g.energies()[e]->revision() != s.revisions[e];
            if (!s.built || m_precomputationUpdateRequired || changed) buildRb(i);
            if (s.hostDirty) {
                check(spgx_rb_upload(s.handle, reinterpret_cast<const Real *>(g.positions().data()),
                                     reinterpret_cast<const Real *>(g.thetas().data()), reinterpret_cast<const Real *>(g.velocities().data()),
                                     reinterpret_cast<const Real *>(g.omegas().data())));
                s.hostDirty = false;
            }
            for (size_t e = 0; e < g.energies().size(); ++e)
                if (auto *a = dynamic_cast<RBSpringAnchorEnergy *>(g.energies()[e].get())) {
                    if (a->anchorRevision() == s.anchorRevisions[e]) continue;
                    for (int st = 0; st < a->nStencils(); ++st) {
                        const Vector3 p = a->anchor(st);
                        const Real v[3] = {p.x, p.y, p.z};
                        check(spgx_rb_energy_update_anchor(s.handle, static_cast<int>(e), st, v));
                    }
                    s.anchorRevisions[e] = a->anchorRevision();
                }
            check(spgx_rb_step(s.handle, m_dtStep, m_nsubsteps, rbIterations(), gravity, 1));
            s.deviceDirty = true;
        }
    }

    bool topologyChanged(const ParticleGroup &g, const Slot &s) const
    {
        if (g.energies().size() != s.identities.size()) return true;
        for (size_t e = 0; e < g.energies().size(); ++e)
            if (g.energies()[e].get() != s.identities[e] || g.energies()[e]->revision() != s.revisions[e]) return true;
        return false;
    }
    void build(size_t i)
    {
        ParticleGroup &g = m_groups[i];
        Slot &s = m_slots[i];
        if (s.handle) {
            syncToHost();
            spgx_group_destroy(s.handle);
            s.handle = nullptr;
        }
        check(spgx_group_create(m_ctx, g.size(), reinterpret_cast<const Real *>(g.initialPositions().data()),
                                reinterpret_cast<const Real *>(g.positions0().data()), g.masses().data(), &s.handle));
        s.identities.clear();
        s.revisions.clear();
        s.anchorRevisions.clear();
        for (auto &e : g.energies()) {
            int id = -1;
            check(spgx_group_add_energy(s.handle, e->kind(), e->nStencils(), e->flatStencils().data(), e->modelParams().data(),
                                        &id));
            s.identities.push_back(e.get());
            s.revisions.push_back(e->revision());
            auto *a = dynamic_cast<SpringAnchorEnergy *>(e.get());
            s.anchorRevisions.push_back(a ? a->anchorRevision() : 0);
        }
        check(spgx_group_finalize(s.handle, solverKind()));
        check(spgx_upload(s.handle, reinterpret_cast<const Real *>(g.positions().data()),
                          reinterpret_cast<const Real *>(g.velocities().data())));
        s.built = true;
        s.hostDirty = s.deviceDirty = false;
    }
    void stepAll()
    {
        const Real g[3] = {m_gravity.x, m_gravity.y, m_gravity.z};
        for (size_t i = 0; i < m_groups.size(); ++i) {
            Slot &s = m_slots[i];
            if (!s.built || m_precomputationUpdateRequired || topologyChanged(m_groups[i], s)) build(i);
            if (s.hostDirty) {
                check(spgx_upload(s.handle, reinterpret_cast<const Real *>(m_groups[i].positions().data()),
                                  reinterpret_cast<const Real *>(m_groups[i].velocities().data())));
                s.hostDirty = false;
            }
            // live anchor edits
            for (size_t e = 0; e < m_groups[i].energies().size(); ++e)
                if (auto *a = dynamic_cast<SpringAnchorEnergy *>(m_groups[i].energies()[e].get())) {
                    if (a->anchorRevision() == s.anchorRevisions[e]) continue;
                    for (int st = 0; st < a->nStencils(); ++st) {
                        const Vector3 p = a->anchor(st);
                        const Real v[3] = {p.x, p.y, p.z};
                        check(spgx_energy_update_record(s.handle, static_cast<int>(e), st, 0, v, 3));
                    }
                    s.anchorRevisions[e] = a->anchorRevision();
                }
            check(stepGroup(s.handle, g));
            s.deviceDirty = true;
            if (m_verbosity == Verbosity::Debug) {
                // a device-side NaN / Inf scan of the state after every step (the reference's Debug level prints; a GPU path
                // that goes non-finite would otherwise only show up at the next positions() read)
                int64_t bad[3] = {0, 0, -1};
                check(spgx_check_finite(s.handle, bad));
                if (bad[0] || bad[1])
                    throw std::runtime_error("non-finite state after step(): " + std::to_string(bad[0]) + " position and " +
                                             std::to_string(bad[1]) + " velocity components, first at vertex " + std::to_string(bad[2]));
            }
        }
        stepRbGroups(g);
        m_precomputationUpdateRequired = false;
        m_time += m_dtStep;
    }
    void syncToHost()
    {
        for (size_t i = 0; i < m_rbGroups.size(); ++i) {
            RbSlot &s = m_rbSlots[i];
            if (s.handle && s.deviceDirty) {
                RigidBodyGroup &g = m_rbGroups[i];
                check(spgx_rb_download(s.handle, reinterpret_cast<Real *>(g.positions().data()), reinterpret_cast<Real *>(g.thetas().data()),
                                       reinterpret_cast<Real *>(g.velocities().data()), reinterpret_cast<Real *>(g.omegas().data()),
                                       reinterpret_cast<Real *>(g.rotationMatrices().data())));
                s.deviceDirty = false;
            }
        }
        for (size_t i = 0; i < m_groups.size(); ++i) {
            Slot &s = m_slots[i];
            if (s.handle && s.deviceDirty) {
                check(spgx_download(s.handle, reinterpret_cast<Real *>(m_groups[i].positions().data()),
                                    reinterpret_cast<Real *>(m_groups[i].velocities().data())));
                s.deviceDirty = false;
            }
        }
    }

    spgx_ctx *m_ctx{nullptr};
    std::vector<ParticleGroup> m_groups;
    std::vector<Slot> m_slots;
    std::vector<RigidBodyGroup> m_rbGroups;
    std::vector<RbSlot> m_rbSlots;
    Vector3 m_gravity{0, -9.8, 0};
    Real m_time{0};
    Real m_dtStep{0.01};
    int m_nsubsteps{1};
    Verbosity m_verbosity{Verbosity::None};
    bool m_precomputationUpdateRequired{false};
};

class VBD : public BaseSolver  // vbd.h:15-41
{
public:
    enum class InitialGuessType { InertialWithAcceleration };
    using BaseSolver::BaseSolver;
    void step() override { stepAll(); }
    void setInitialGuessType(const InitialGuessType) {}
    void requirePrecomputationUpdate() { m_precomputationUpdateRequired = true; }
    void setIterations(int n) { m_iterations = n; }  // vbd.h:33 "TODO: Expose this in the solver API"
    int iterations() const { return m_iterations; }

protected:
    int solverKind() const override { return SPGX_VBD; }
    int rbIterations() const override { return m_iterations; }
    int stepGroup(spgx_group *g, const Real *gravity) override
    {
        return spgx_vbd_step(g, m_dtStep, m_nsubsteps, m_iterations, gravity, 1, 1);
    }
    int m_iterations{1};
};

class XPBD : public BaseSolver  // xpbd.h:15-29
{
public:
    explicit XPBD(bool useParallelGaussSeidel = true, int device = 0)
        : BaseSolver(device), m_useParallelGaussSeidel(useParallelGaussSeidel)
    {
    }
    void step() override { stepAll(); }
    void requirePrecomputationUpdate() { m_precomputationUpdateRequired = true; }
    void setIterations(int n) { m_iterations = n; }  // extension: lambda accumulated on device

protected:
    int solverKind() const override { return m_useParallelGaussSeidel ? SPGX_XPBD : SPGX_XPBD_SERIAL; }
    int stepGroup(spgx_group *g, const Real *gravity) override
    {
        return spgx_xpbd_step(g, m_dtStep, m_nsubsteps, m_iterations, gravity, 1, 1);
    }
    bool m_useParallelGaussSeidel;
    int m_iterations{1};
};

// The implicit-Euler family (implicitEulerBase.h:11-25): one linearised solve per substep, Eigen's Jacobi-preconditioned
// conjugate gradient with m_cgThreshold 1e-5 and m_cgMaxIterations 300, all on the device.
class ImplicitEulerBase : public BaseSolver
{
public:
    using BaseSolver::BaseSolver;
    void step() override { stepAll(); }
    // cg.iterations() of every linear solve of the last step() of group `i` (the reference only prints it,
    // implicitEulerBase.cpp:234-236)
    std::vector<int> lastCgIterations(size_t i = 0) const
    {
        std::vector<int32_t> it(4096);
        int n = 0;
        if (i < m_slots.size() && m_slots[i].handle) spgx_last_cg_iterations(m_slots[i].handle, it.data(), 4096, &n);
        return std::vector<int>(it.begin(), it.begin() + (n < 4096 ? n : 4096));
    }

protected:
    int solverKind() const override { return SPGX_NEWTON; }
    virtual int implicitKind() const = 0;
    int stepGroup(spgx_group *g, const Real *gravity) override
    {
        return spgx_implicit_step(g, implicitKind(), m_dtStep, m_nsubsteps, gravity, m_cgThreshold, m_cgMaxIterations, 1, 1);
    }
    const Real m_cgThreshold{1e-5};
    const int m_cgMaxIterations{300};
};
class ImplicitEulerBaraffWitkin : public ImplicitEulerBase  // implicitEulerBaraffWitkin.h
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_BARAFF_WITKIN; }
};
class ImplicitEulerNewtonDx : public ImplicitEulerBase  // implicitEulerNewtonDx.h
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_NEWTON_DX; }
};
class ImplicitEulerNewtonDv : public ImplicitEulerBase  // implicitEulerNewtonDv.h
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_NEWTON_DV; }
};
class BDF2 : public ImplicitEulerBase  // bdf2.h:12-21 (reset() forgets the previous state: spgx_reset does)
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_BDF2; }
};
class StaticNewton : public ImplicitEulerBase  // staticNewton.h:12
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_STATIC_NEWTON; }
};
class QuasiStaticNewton : public ImplicitEulerBase  // quasiStaticNewton.h:11
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_QUASI_STATIC_NEWTON; }
};
class ImplicitEulerNewtonRobust : public ImplicitEulerBase  // implicitEulerNewtonRobust.h:12-28
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_NEWTON_ROBUST; }
};
class QuasiStaticNewtonRobust : public ImplicitEulerBase  // quasiStaticNewtonRobust.h:12-31
{
public:
    using ImplicitEulerBase::ImplicitEulerBase;

protected:
    int implicitKind() const override { return SPGX_IE_QUASI_STATIC_NEWTON_ROBUST; }
};
class SimplecticEuler : public BaseSolver  // simplecticEuler.h:12-16 (the reference's spelling)
{
public:
    using BaseSolver::BaseSolver;
    void step() override { stepAll(); }

protected:
    int solverKind() const override { return SPGX_NEWTON; }
    int stepGroup(spgx_group *g, const Real *gravity) override
    {
        return spgx_symplectic_step(g, m_dtStep, m_nsubsteps, gravity, 1, 1);
    }
};
}  // namespace solver
}  // namespace spgx
