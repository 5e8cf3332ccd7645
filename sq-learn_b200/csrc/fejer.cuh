// O(1) exact sampler for the outcome law shared by amplitude estimation and
// phase estimation.
//
// Reference: amplitude_estimation (sklearn/QuantumUtility/Utility.py:494-508)
// and phase_estimation (Utility.py:637-653) build, for every grid point
// j in [0, M), the weight |sin(M*D*pi) / (M*sin(D*pi))|^2 with D the distance
// between j/M and the true phase w (circular for AE, and periodic by
// construction for PE), then draw one j with random.choices.  Since j is an
// integer, sin^2(pi*(M*w - j)) = sin^2(pi*M*w) does not depend on j, so
//
//     p(j) = sin^2(pi*t) / (M^2 * sin^2(pi*(j - x)/M)),   x = M*w, t = frac(x)
//
// and sum_j p(j) = 1 exactly (Fejer kernel identity).  We enumerate the four
// grid points around x exactly (>= 87% of the mass) and sample the two tails
// by rejection from a 1/z^2 envelope (sin(pi*y) >= 2y on [0, 1/2]).  For tiny
// M the O(M) inversion is used; it is also the truth the tests compare with.
#pragma once
#include "common.cuh"

namespace sq {

// exact pmf of offset m (grid point j = floor(x) + m), given s2 = sin^2(pi t)
__host__ __device__ inline double fejer_pmf(double M, double t, double s2, double m) {
    double y = (m - t) / M;
#ifdef __CUDA_ARCH__
    double s = sinpi(y);
#else
    double s = sin(3.14159265358979323846 * y);
#endif
    double ms = M * s;
    return s2 / (ms * ms);
}

// Draw j in [0, M) from p(j) above.  x = M*w must lie in [0, M].
__host__ __device__ inline double fejer_sample(double M, double x, Philox& rng) {
    double j0 = floor(x);
    double t = x - j0;
    if (j0 >= M) j0 -= M;
    if (t == 0.0) return j0;  // reference: weight 1 at distance 0, ~1e-32 elsewhere
#ifdef __CUDA_ARCH__
    double st = sinpi(t);
#else
    double st = sin(3.14159265358979323846 * t);
#endif
    double s2 = st * st;
    double u, v;
    rng.uniform2(u, v);

    if (M <= 16.0) {
        // brute-force inversion over offsets m = 0..M-1 (periodic, any window of M works)
        double acc = 0.0;
        int Mi = (int)M;
        for (int m = 0; m < Mi; ++m) {
            acc += fejer_pmf(M, t, s2, (double)m);
            if (u < acc || m == Mi - 1) {
                double j = j0 + m;
                return j >= M ? j - M : j;
            }
        }
    }

    // window m in {0, 1, -1, 2}, ordered by typical mass
    double m_sel;
    bool found = false;
    double acc = 0.0;
    const double wm[4] = {0.0, 1.0, -1.0, 2.0};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        acc += fejer_pmf(M, t, s2, wm[i]);
        if (!found && u < acc) { m_sel = wm[i]; found = true; }
    }
    if (!found) {
        // tails: right m in [3, mhi], left m in [mlo, -2], mhi - mlo + 1 = M,
        // chosen so |m - t| <= M/2 everywhere.
        double half = floor((M - 1.0) * 0.5);
        double mlo = -half;
        if (t > 0.5 && fmod(M, 2.0) != 0.0) mlo += 1.0;  // keeps t - mlo <= M/2 for odd M
        double mhi = mlo + M - 1.0;
        // envelope masses (common factor s2/4 dropped): integral of 1/z^2
        double eR = (mhi >= 3.0) ? (1.0 / (2.0 - t) - 1.0 / (mhi - t)) : 0.0;
        double eL = (mlo <= -2.0) ? (1.0 / (1.0 + t) - 1.0 / (t - mlo)) : 0.0;
        // v decides the side on the first try; later tries draw fresh numbers
        double side_u = v;
        for (int it = 0; it < 1000; ++it) {
            double a, b;
            rng.uniform2(a, b);
            bool right = side_u * (eR + eL) < eR;
            double zmin = right ? (2.0 - t) : (1.0 + t);
            double zmax = right ? (mhi - t) : (t - mlo);
            double inv = 1.0 / zmin - a * (1.0 / zmin - 1.0 / zmax);
            double z = 1.0 / inv;               // z in [zmin, zmax], density ~ 1/z^2
            double dist = right ? ceil(z + t) - t : ceil(z - t) + t;  // |m - t| of the cell holding z
            // cell is (dist-1, dist]; its envelope mass is 1/(dist-1) - 1/dist
            if (dist - 1.0 < zmin) dist = zmin + 1.0;
            if (dist > zmax) dist = zmax;
            double env = 0.25 * (1.0 / (dist - 1.0) - 1.0 / dist);
            double y = dist / M;
#ifdef __CUDA_ARCH__
            double s = sinpi(y);
#else
            double s = sin(3.14159265358979323846 * y);
#endif
            double p = 1.0 / (M * s * M * s);
            if (b * env <= p) {
                m_sel = right ? (dist + t) : (t - dist);
                m_sel = rint(m_sel);
                found = true;
                break;
            }
            side_u = rng.uniform();
        }
        if (!found) m_sel = 0.0;  // unreachable in practice (acceptance >= 4/pi^2 * 1/2)
    }
    double j = j0 + m_sel;
    if (j < 0.0) j += M;
    if (j >= M) j -= M;
    return j;
}

}  // namespace sq
