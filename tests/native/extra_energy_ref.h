// TEST INFRASTRUCTURE: the reference-side twin of tests/native/extra_kind.cuh, written against the reference's own
// authoring base class (StencilBlockEnergy<2>: spg/include/spg/sim/energy/energy.h:75-76) the way its energies are
// (compare spg/source/sim/energy/springContinuumEnergy.cpp:8-49).  Compiled into oracle/_ref/libspgref_extra.so
// (oracle/Makefile, target ref_extra) through the hook in oracle/ref_driver.cpp.
#pragma once
#include <spg/sim/energy/energy.h>
#include <spg/sim/simObject/particleGroup.h>

namespace spg
{
class SquaredDistanceEnergy : public StencilBlockEnergy<2>
{
public:
    SquaredDistanceEnergy() { m_name = "SquaredDistanceEnergy"; }
    void addStencil(const std::array<int, s_stencilSize> &stencil, const Real restLength, const Real stiffness)
    {
        m_stencils.push_back(stencil);
        if (restLength <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        m_restLength.push_back(restLength);
        m_modelStiffness.push_back(StiffnessMat{stiffness});
        m_effectiveStiffness.push_back(StiffnessMat{stiffness});
        m_modelCompliance.push_back(m_modelStiffness.back().inverse());
        m_effectiveCompliance.push_back(m_effectiveStiffness.back().inverse());
    }

protected:
    template <class TdC>
    void constraint(const int i, const ParticleGroup &pGroup, TdC &dC) const
    {
        const auto &x0{pGroup.positions()[m_stencils[i][0]]};
        const auto &x1{pGroup.positions()[m_stencils[i][1]]};
        using RealT = std::decay_t<decltype(dC[0])>;
        const Vector3T<RealT> x(RealT(x0.x(), 0), RealT(x0.y(), 1), RealT(x0.z(), 2));
        const Vector3T<RealT> y(RealT(x1.x(), 3), RealT(x1.y(), 4), RealT(x1.z(), 5));
        const Real L0 = m_restLength[i];
        dC[0] = (x - y).dot(x - y) - L0 * L0;
    }
    virtual void dConstraints(int i, const ParticleGroup &pGroup, ConstraintsAD1 &dC) const { constraint(i, pGroup, dC); }
    virtual void dConstraints(int i, const ParticleGroup &pGroup, ConstraintsAD2 &dC) const { constraint(i, pGroup, dC); }
    std::vector<Real> m_restLength;
};
}  // namespace spg

#define SPG_EXTRA_ENERGY_TYPE spg::SquaredDistanceEnergy
constexpr int K_TEST_SQUARED_DISTANCE = 14;
inline std::shared_ptr<spg::ParticleGroup::EnergyT> makeExtraEnergy(int kind, int n, const int32_t *st, const double *p)
{
    if (kind != K_TEST_SQUARED_DISTANCE) return nullptr;
    auto e = std::make_shared<spg::SquaredDistanceEnergy>();
    for (int i = 0; i < n; ++i) e->addStencil({st[2 * i], st[2 * i + 1]}, p[2 * i], p[2 * i + 1]);
    return e;
}
inline int extraKindStencilSize(int kind) { return kind == K_TEST_SQUARED_DISTANCE ? 2 : -1; }
inline int extraKindNumModelParams(int kind) { return kind == K_TEST_SQUARED_DISTANCE ? 2 : -1; }
