// A throw-away energy kind defined OUTSIDE the library sources (tests/test_extra_kind.py builds libspgx_extra.so with
//   -DSPGX_EXTRA_KINDS_DEF="…/extra_kind.def" -DSPGX_EXTRA_KINDS_IMPL="…/extra_kind.cuh").
// The authoring contract of the reference (StencilBlockEnergy<S, C>: supply C(x) as one generic lambda over the scalar
// type, spg/include/spg/sim/energy/energy.h:75-76, :305-347) restated for the device: one template body, instantiated with
// doubles, first-order duals (XPBD Jacobian) and second-order duals (VBD vertex block, Newton element Hessian).
// The reference-side twin of this energy, written against the reference's own base class, is extra_energy_ref.h.
//
//   C(x, y) = |x - y|^2 - L0^2        K_eff = stiffness        record p = {L0, k, 1/k}      model = {L0, k}
#pragma once
#include <stdexcept>
namespace spgx
{
template <>
struct EnergyKind<K_TEST_SQUARED_DISTANCE> {
    static constexpr int S = 2, NC = 1, NP = 3, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        const V3<T> d = x[0] - x[1];
        C[0] = dot(d, d) - p[0] * p[0];
    }
    SPGX_HD static double K(const double *p, int, int) { return p[1]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[2]; }
    // host: addStencil + preparePrecomputations of the twin class (model -> record)
    static void prepare(const int *, const double *model, const double *, double *p)
    {
        if (model[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        p[0] = model[0];
        p[1] = model[1];
        p[2] = 1.0 / model[1];
    }
};
}  // namespace spgx
