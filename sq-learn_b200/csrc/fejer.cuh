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
// and sum_j p(j) = 1 exactly (Fejer kernel identity).  The eight grid points
// around x are enumerated exactly (>= 95% of the mass) and the two tails are
// sampled by rejection from a 1/z^2 envelope (sin(pi*y) >= 2y on [0, 1/2]).
// The table is built once per (M, x) and reused for repeated draws (the
// median of 13 in inner-product estimation).  For tiny M the O(M) inversion
// is used; it is also the truth the tests compare with.
#pragma once
#include "common.cuh"

namespace sq {

__host__ __device__ inline double sq_sinpi(double y) {
#ifdef __CUDA_ARCH__
    return sinpi(y);
#else
    return sin(3.14159265358979323846 * y);
#endif
}

static constexpr int FEJER_W = 4;  // window offsets -3 .. 4

struct FejerTable {
    double M, j0, t, s2;
    double cum[2 * FEJER_W];  // cumulative mass of offsets 0, 1, -1, 2, -2, 3, -3, 4
    double eR, eL;            // envelope masses of the right / left tail (common factor dropped)
    double zminR, zmaxR, zminL, zmaxL;
    bool exact;               // x is a grid point: the draw is deterministic
    bool brute;               // M too small for window + tails
};

__host__ __device__ inline double fejer_offset(int i) {
    // 0, 1, -1, 2, -2, 3, -3, 4
    return (i & 1) ? (double)((i + 1) >> 1) : -(double)(i >> 1);
}

__host__ __device__ inline void fejer_setup(FejerTable& T, double M, double x) {
    T.M = M;
    double j0 = floor(x);
    T.t = x - j0;
    if (j0 >= M) j0 -= M;
    T.j0 = j0;
    T.exact = (T.t == 0.0);
    T.brute = (M <= 2.0 * FEJER_W + 1.0);
    double st = sq_sinpi(T.t);
    T.s2 = st * st;
    if (T.exact || T.brute) return;
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < 2 * FEJER_W; ++i) {
        double ms = M * sq_sinpi((fejer_offset(i) - T.t) / M);
        acc += T.s2 / (ms * ms);
        T.cum[i] = acc;
    }
    // tails: right m in [W+1, mhi], left m in [mlo, -W], mhi - mlo + 1 = M, |m - t| <= M/2 everywhere
    double half = floor((M - 1.0) * 0.5);
    double mlo = -half;
    if (T.t > 0.5 && fmod(M, 2.0) != 0.0) mlo += 1.0;
    double mhi = mlo + M - 1.0;
    T.zminR = FEJER_W - T.t;        T.zmaxR = mhi - T.t;
    T.zminL = FEJER_W - 1.0 + T.t;  T.zmaxL = T.t - mlo;
    T.eR = (T.zmaxR > T.zminR) ? (1.0 / T.zminR - 1.0 / T.zmaxR) : 0.0;
    T.eL = (T.zmaxL > T.zminL) ? (1.0 / T.zminL - 1.0 / T.zmaxL) : 0.0;
}

__host__ __device__ inline double fejer_wrap(const FejerTable& T, double m) {
    double j = T.j0 + m;
    if (j < 0.0) j += T.M;
    if (j >= T.M) j -= T.M;
    return j;
}

// window lookup for a uniform u; returns false when u falls in the tails
__host__ __device__ inline bool fejer_window(const FejerTable& T, double u, double& j) {
#pragma unroll
    for (int i = 0; i < 2 * FEJER_W; ++i) {
        if (u < T.cum[i]) { j = fejer_wrap(T, fejer_offset(i)); return true; }
    }
    return false;
}

// one draw from the tails (conditional law), by rejection
__host__ __device__ inline double fejer_tail(const FejerTable& T, Philox& rng) {
    if (!(T.eR + T.eL > 0.0)) return fejer_wrap(T, 0.0);  // no tail cells (rounding put u past the window)
    for (int it = 0; it < 1000; ++it) {
        double a, b;
        rng.uniform2(a, b);
        // a picks the side and the position inside it (one uniform, rescaled), b accepts
        double tot = T.eR + T.eL;
        bool right = a * tot < T.eR;
        double a2 = right ? (a * tot) / T.eR : (a * tot - T.eR) / T.eL;
        double zmin = right ? T.zminR : T.zminL, zmax = right ? T.zmaxR : T.zmaxL;
        double z = 1.0 / (1.0 / zmin - a2 * (1.0 / zmin - 1.0 / zmax));   // density ~ 1/z^2 on [zmin, zmax]
        double dist = right ? ceil(z + T.t) - T.t : ceil(z - T.t) + T.t;  // |m - t| of the cell (dist-1, dist] holding z
        if (dist - 1.0 < zmin) dist = zmin + 1.0;
        if (dist > zmax) dist = zmax;
        double env = 0.25 * (1.0 / (dist - 1.0) - 1.0 / dist);
        double ms = T.M * sq_sinpi(dist / T.M);
        if (b * env * ms * ms <= 1.0) return fejer_wrap(T, rint(right ? dist + T.t : T.t - dist));
    }
    return fejer_wrap(T, 0.0);  // unreachable in practice (acceptance >= 4/pi^2 * 3/4)
}

__host__ __device__ inline double fejer_brute(const FejerTable& T, double u) {
    double acc = 0.0;
    int Mi = (int)T.M;
    for (int m = 0; m < Mi; ++m) {
        double ms = T.M * sq_sinpi(((double)m - T.t) / T.M);
        acc += T.s2 / (ms * ms);
        if (u < acc || m == Mi - 1) return fejer_wrap(T, (double)m);
    }
    return T.j0;
}

// Draw j in [0, M) from p(j) above.  x = M*w must lie in [0, M].
__host__ __device__ inline double fejer_sample(double M, double x, Philox& rng) {
    FejerTable T;
    fejer_setup(T, M, x);
    if (T.exact) return T.j0;  // reference: weight 1 at distance 0, ~1e-32 elsewhere
    double u = rng.uniform();
    if (T.brute) return fejer_brute(T, u);
    double j;
    if (fejer_window(T, u, j)) return j;
    return fejer_tail(T, rng);
}

}  // namespace sq
