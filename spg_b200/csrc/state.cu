// Bandwidth-bound state passes fused from the reference's serial host loops:
//   predict : xOld <- x ; v += dt g ; x += v dt ; xInertial <- x        (vbd.cpp:49-75, xpbd.cpp:48-74,
//                                                                         particleGroup.cpp:7-14, :57-63)
//   finalize: v <- (x - xOld) * invdt                                    (particleGroup.cpp:65-72)
// Arrays are flat 3V doubles, so the kernels run over scalar components with fully coalesced accesses.
#include "kernels.h"

namespace spgx
{
namespace
{
__global__ void predictKernel(double *__restrict__ x, double *__restrict__ v, double *__restrict__ xOld,
                              double *__restrict__ xInertial, long long n3, double dt, double gx, double gy, double gz)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const int c = static_cast<int>(i % 3);
    const double dtg = dt * (c == 0 ? gx : (c == 1 ? gy : gz));
    const double xo = x[i];
    const double vn = v[i] + dtg;
    const double xn = xo + vn * dt;
    xOld[i] = xo;
    v[i] = vn;
    x[i] = xn;
    if (xInertial) xInertial[i] = xn;
}

__global__ void finalizeKernel(const double *__restrict__ x, const double *__restrict__ xOld, double *__restrict__ v,
                               long long n3, double invdt)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    v[i] = (x[i] - xOld[i]) * invdt;
}

// finalize of one substep + predict of the next in one launch (XPBD runs 20 substeps a frame: one kernel boundary less per
// substep).  The same roundings as the two kernels above: the velocity is rounded where finalizeKernel stores it (no
// contraction of its multiplication into the gravity add), the position update is the expression predictKernel uses.
__global__ void finalizePredictKernel(double *__restrict__ x, double *__restrict__ v, double *__restrict__ xOld, long long n3,
                                      double dt, double invdt, double gx, double gy, double gz)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const int c = static_cast<int>(i % 3);
    const double dtg = dt * (c == 0 ? gx : (c == 1 ? gy : gz));
    const double xo = x[i];
    const double vf = __dmul_rn(__dsub_rn(xo, xOld[i]), invdt);  // finalizeKernel: rounded here, as where it is stored
    const double vn = vf + dtg;                                  // predictKernel from here on (the same expressions, so the
                                                                 // same contractions: dt g + v is one fused multiply-add there)
    const double xn = xo + vn * dt;
    xOld[i] = xo;
    v[i] = vn;
    x[i] = xn;
}

__global__ void permute3Kernel(const double *__restrict__ src, double *__restrict__ dst,
                               const int32_t *__restrict__ perm, int n, bool gather)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t a = 3 * static_cast<size_t>(gather ? perm[i] : i);
    const size_t b = 3 * static_cast<size_t>(gather ? i : perm[i]);
    dst[b] = src[a];
    dst[b + 1] = src[a + 1];
    dst[b + 2] = src[a + 2];
}
// Verbosity::Debug check: counts the non-finite components of x and v over the first n slots and keeps the smallest
// offending vertex id (the caller's numbering).  out: [count in x, count in v, vertex + 1]
__global__ void checkFiniteKernel(const double *__restrict__ x, const double *__restrict__ v, const int32_t *__restrict__ slot2vert,
                                  long long n3, unsigned long long *__restrict__ out)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const bool bx = !isfinite(x[i]), bv = !isfinite(v[i]);
    if (bx) atomicAdd(out, 1ull);
    if (bv) atomicAdd(out + 1, 1ull);
    if (bx || bv) atomicMin(out + 2, static_cast<unsigned long long>(slot2vert[i / 3]) + 1ull);
}
}  // namespace

void launchCheckFinite(const double *x, const double *v, const int32_t *slot2vert, int nVerts, unsigned long long *out3,
                       cudaStream_t stream)
{
    const long long n3 = 3LL * nVerts;
    if (n3 == 0) return;
    const int threads = 256;
    checkFiniteKernel<<<static_cast<unsigned>((n3 + threads - 1) / threads), threads, 0, stream>>>(x, v, slot2vert, n3, out3);
}

namespace
{
__global__ void gatherScalarKernel(const double *__restrict__ src, const int32_t *__restrict__ idx, long long n, double *__restrict__ dst)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}
}  // namespace
void launchGatherScalar(const double *src, const int32_t *idx, long long n, double *dst, cudaStream_t stream)
{
    if (n == 0) return;
    gatherScalarKernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(src, idx, n, dst);
}

void launchPredict(double *x, double *v, double *xOld, double *xInertial, int nVerts, double dt, const double *g3,
                   cudaStream_t stream)
{
    const long long n3 = 3LL * nVerts;
    if (n3 == 0) return;
    const int threads = 256;
    predictKernel<<<static_cast<unsigned>((n3 + threads - 1) / threads), threads, 0, stream>>>(x, v, xOld, xInertial, n3,
                                                                                               dt, g3[0], g3[1], g3[2]);
}

void launchFinalize(const double *x, const double *xOld, double *v, int nVerts, double invdt, cudaStream_t stream)
{
    const long long n3 = 3LL * nVerts;
    if (n3 == 0) return;
    const int threads = 256;
    finalizeKernel<<<static_cast<unsigned>((n3 + threads - 1) / threads), threads, 0, stream>>>(x, xOld, v, n3, invdt);
}

void launchFinalizePredict(double *x, double *v, double *xOld, int nVerts, double dt, double invdt, const double *g3,
                           cudaStream_t stream)
{
    const long long n3 = 3LL * nVerts;
    if (n3 == 0) return;
    const int threads = 256;
    finalizePredictKernel<<<static_cast<unsigned>((n3 + threads - 1) / threads), threads, 0, stream>>>(x, v, xOld, n3, dt, invdt, g3[0],
                                                                                                       g3[1], g3[2]);
}

void launchPermute3(const double *src, double *dst, const int32_t *perm, int n, bool gather, cudaStream_t stream)
{
    if (n == 0) return;
    const int threads = 256;
    permute3Kernel<<<(n + threads - 1) / threads, threads, 0, stream>>>(src, dst, perm, n, gather);
}
}  // namespace spgx
