// Inner-product estimation for one (sample, centroid) pair: the median of the 13 amplitude-estimation outcomes
// sampled by ONE inversion.
//
// Reference: ipe -> median_evaluation(amplitude_estimation, gamma = 0.1 => Q = 13)
// (sklearn/QuantumUtility/Utility.py:697-737, :534-572, :442-531).  One outcome j has the law
//     p(j) = sin^2(pi t) f(|j - x|),   f(d) = 1 / (M sin(pi d / M))^2,   x = M theta_a / pi,  t = frac(x),
// and the estimate a~ = sin^2(pi j / M) depends on j through r = min(j, M - j) only, increasingly.  The median
// of 13 independent outcomes therefore has the CDF  G(F(r)),  F the CDF of ONE outcome in r-order and
// G(F) = P(Binomial(13, F) >= 7), so it can be drawn by inversion from a single uniform V: the smallest r with
// V <= G(F(r)).  No tail draws, no sort.
//
// F at the 8 grid points around x ("window") is  F_i = B + cum_i,  cum_i the running sum of the window weights
// (plus the weights of their mirror images M - r, negligible away from M/2) and B the mass of every r below the
// window.  In distance from x, B is ONE contiguous run:  B = sin^2(pi t) sum_{k=0}^{K-1} f(a + k),  a = t + 4,
// K = 2 floor(x) - 7 (the points left of the window down to 0, then their mirror images).  Every even derivative of
// f is positive on (0, M), so the Euler-Maclaurin expansion of the run envelops it (ipe_bracket_ends):
//     sum = int_a^b f + (f(a) + f(b))/2 + (f'(b) - f'(a))/12 - th (f'''(b) - f'''(a))/720,   0 <= th <= 1,  b = a + K - 1,
// all in closed form from (sin, cos) at the two ends (the antiderivative of csc^2 is -cot).  The gap is 3e-6 sin^2(pi t).
// V is compared with G at both ends of the bracket: if both agree the comparison is certain; otherwise (3e-5 of the
// pairs) ipe_tighten peels exact terms off both ends of the run, which shrinks the gap like 1/a^5, until every
// comparison is certain; a median outside the window (1e-6) is found by plain summation (ipe_brute).  All three use
// the same V, so together they ARE the inversion - exact in law, not an approximation.  ipe_lean declines
// what it is not built for before looking at V: small grids and windows touching 0 are summed plainly (IPE_BRUTE, a
// short sum there), windows touching M/2 on a large grid go to the general estimator of ipe.cuh with its own draws
// (IPE_GENERAL) - neither touches the law.
#pragma once
#include "common.cuh"
#include "fejer.cuh"

namespace sq {

enum { IPE_DONE = 0, IPE_GENERAL = 1, IPE_TIGHTEN = 2, IPE_BRUTE = 3 };

// The inversion's uniform V of the pair (row, col), in two parts: V = (H + L) / 65536.  H, 16 bits, comes from the block
// (seed, row, IPE_TAG_OCT, col / 8) that serves the eight columns 8c .. 8c + 7 of a row (32 bits per column pair); L, a
// uniform in (0, 1), from the pair's own stream (seed, row, IPE_TAG_LOW, col).  V is uniform in (0, 1) and V >= H / 65536,
// which is all the certificate below needs to know of it - so the E-step draws one block per eight pairs, and the
// second one only for the pairs it samples.  The general estimator of ipe.cuh draws from other tags.
static constexpr uint32_t IPE_TAG_OCT = 0x1bE1u, IPE_TAG_LOW = 0x1bE2u;
__host__ __device__ inline void ipe_oct_bits(uint64_t seed, uint64_t rowkey, uint32_t col, uint32_t out[4]) {
    Philox rng(seed, rowkey, IPE_TAG_OCT, col >> 3);
    rng.block(out);
}
__host__ __device__ inline uint32_t ipe_oct_high(const uint32_t bits[4], uint32_t col) {
    return (bits[(col & 7u) >> 1] >> ((col & 1u) * 16u)) & 0xffffu;
}
__host__ __device__ inline double ipe_pair_uniform(uint64_t seed, uint64_t rowkey, uint32_t col) {
    uint32_t bits[4];
    ipe_oct_bits(seed, rowkey, col, bits);
    Philox rng(seed, rowkey, IPE_TAG_LOW, col);
    return ((double)ipe_oct_high(bits, col) + rng.uniform()) * (1.0 / 65536.0);
}

struct IpeGrid {
    double S;      // |x|^2 + |c|^2
    double M;      // grid size of amplitude estimation (Utility.py:484)
    double x;      // M theta_a / pi
    double j0, t;  // floor and fraction of x
    int kind;      // 0: S <= 0 (estimate 0), 1: x on a grid point (deterministic), 2: to be sampled
    double invM;   // 1 / M
    double rS;     // 1 / S
    double c2t;    // cos(2 theta_a) = 1 - 2a
};

__host__ __device__ inline double sq_cospi(double y) {
#ifdef __CUDA_ARCH__
    return cospi(y);
#else
    return cos(3.14159265358979323846 * y);
#endif
}

__host__ __device__ inline void sq_sincospi(double y, double& sn, double& cs) {
#ifdef __CUDA_ARCH__
    sincospi(y, &sn, &cs);
#else
    sn = sin(3.14159265358979323846 * y);
    cs = cos(3.14159265358979323846 * y);
#endif
}

// a, epsilon_a and M of ipe / amplitude_estimation (Utility.py:725-733, :484).  Returns false for S = |x|^2 + |c|^2 <= 0
// (the estimate is 0; kind = 0).  `a` is handed to ipe_grid_tail.
__host__ __device__ inline bool ipe_grid_head(double dot, double xn, double cn, double eps, IpeGrid& g, double& a) {
    const double PI = 3.14159265358979323846;
    g.S = xn + cn;
    g.kind = 0;
    g.M = 16.0; g.invM = 1.0 / 16.0; g.rS = 0.0; g.x = 0.0; g.j0 = 0.0; g.t = 0.0; g.c2t = 1.0;
    a = 0.0;
    if (!(g.S > 0.0)) return false;
    // a = (S - 2 dot) / (2 S),  epsilon_a = eps max(1, |dot|) / S,  M = ceil(pi / (2 epsilon_a) (1 + sqrt(1 + 4 epsilon_a))):
    // the same quantities from ONE reciprocal, r = 1 / (S max(1, |dot|))  (1/S = r mx, S/mx = r S^2)
    const double mx = fmax(1.0, fabs(dot));
    const double r = 1.0 / (g.S * mx);
    const double rS = r * mx;
    g.rS = rS;
    a = fma(-dot, rS, 0.5);
    if (a < 1e-15) a = 0.0;  // math.isclose(a, 0, abs_tol=1e-15) (Utility.py:730); negatives are rounding noise
    if (a > 1.0) a = 1.0;
    const double eps_a = eps * mx * rS;
    g.M = ceil(((PI / (2.0 * eps)) * (g.S * (r * g.S))) * (1.0 + sqrt(1.0 + 4.0 * eps_a)));
    return true;
}

// theta_a and the position x = M theta_a / pi on the grid (Utility.py:493), after ipe_grid_head returned true
__host__ __device__ inline void ipe_grid_tail(double a, IpeGrid& g) {
    const double PI = 3.14159265358979323846;
    g.invM = 1.0 / g.M;
    // theta_a = asin(sqrt(a)) (Utility.py:493).  With c = 1 - 2a = cos(2 theta_a), theta_a = pi/4 - asin(c)/2; for the
    // usual pair (|x.c| a small fraction of |x|^2 + |c|^2) c is small and asin(c) is a short odd series - same
    // accuracy as the library call (both are within an ulp or two of the true angle) at a fifth of the instructions.
    double frac;                                            // theta_a / pi
    const double c = 1.0 - 2.0 * a;
    g.c2t = c;
    if (fabs(c) < 0.25) {
        // asin(c) = c * sum_k  q_k c^(2k),  q_k = (2k)! / (4^k (k!)^2 (2k+1)) (exact rationals); 14 terms, the next one is
        // < 0.25^28 * 0.02 = 3e-19.  Even and odd k as two chains in v = c^4, so that they run side by side.
        const double u = c * c, v = u * u;
        double E = 2704156.0 / (16777216.0 * 25.0);           // k = 12
        E = fma(E, v, 184756.0 / (1048576.0 * 21.0));       // 10
        E = fma(E, v, 12870.0 / (65536.0 * 17.0));          // 8
        E = fma(E, v, 924.0 / (4096.0 * 13.0));             // 6
        E = fma(E, v, 70.0 / (256.0 * 9.0));                // 4
        E = fma(E, v, 6.0 / (16.0 * 5.0));                  // 2
        E = fma(E, v, 1.0);                                 // 0
        double O = 10400600.0 / (67108864.0 * 27.0);          // k = 13
        O = fma(O, v, 705432.0 / (4194304.0 * 23.0));       // 11
        O = fma(O, v, 48620.0 / (262144.0 * 19.0));         // 9
        O = fma(O, v, 3432.0 / (16384.0 * 15.0));           // 7
        O = fma(O, v, 252.0 / (1024.0 * 11.0));             // 5
        O = fma(O, v, 20.0 / (64.0 * 7.0));                 // 3
        O = fma(O, v, 2.0 / (4.0 * 3.0));                   // 1
        frac = fma(-(0.5 / PI) * c, fma(O, u, E), 0.25);    // (exactly 1/4 at c = 0: x = M/4 lands on its grid point)
    } else {
        frac = asin(sqrt(a)) * (1.0 / PI);
    }
    g.x = g.M * frac;
    g.j0 = floor(g.x);
    g.t = g.x - g.j0;
    g.kind = (g.t == 0.0) ? 1 : 2;
}

__host__ __device__ inline void ipe_grid(double dot, double xn, double cn, double eps, IpeGrid& g) {
    double a;
    if (ipe_grid_head(dot, xn, cn, eps, g, a)) ipe_grid_tail(a, g);
}

// Can the pair be left out of the row's arg-min?  e_star = (some already sampled estimate of the same row) - |x|^2;
// the pair is out if its own estimate is certainly larger.  With a* = (e_star + |x|^2) / (2 S) and r* the grid position
// of a* (sin^2(pi r*/M) = a*), every outcome r <= r* lies at a circular distance >= x - r* >= d = M (a - a*) / pi from
// x (sin^2 has slope <= 1), p(j) <= 1 / (4 dist^2) there, so F(r*) <= (d + 1) / (2 d^2) <= 1 / (2 (d - 1)) and
// P(median <= r*) = G(F(r*)) <= 1716 F(r*)^7.  The median is the inversion of the pair's uniform V (ipe_lean /
// ipe_tighten / ipe_brute all are): V > 1716 / (2 (d - 1))^7 therefore certifies median > r*, i.e. estimate > e_star.
// No grid is needed for it: M >= pi / epsilon_a = pi S / (eps mx), mx = max(1, |x.c|), and a - a* =
// (|c|^2 - 2 x.c - e_star) / (2 S), so d >= D / (2 eps mx) with D = (true distance - |x|^2) - e_star: S cancels, no
// division, no square root.  Three grid points of margin (6 eps mx off D) cover the roundings and the clamps of a.
// Pairs the inversion does not serve (x within 4.5 points of M/2 on a grid of 64 or more, i.e. (1 - a) M^2 < 200:
// nearly opposite vectors) are never left out: (S + 2 x.c) S pi^2 > 500 (eps mx)^2 is asked for.
// Returns false if the pair must be sampled whatever its uniform; otherwise the pair is out iff V * lhs > rhs.
// Exact: the labels, distances and inertia are the ones the full computation gives.
__host__ __device__ inline bool ipe_out_possible(double dot, double xn, double cn, double eps, double e_star, double& lhs,
                                                 double& rhs) {
    const double PI = 3.14159265358979323846;
    const double S = xn + cn, em = eps * fmax(1.0, fabs(dot));
    const double D = fma(-2.0, dot, cn) - e_star - 6.0 * em;          // = 2 eps mx (d - 3)
    const double D2 = D * D, D4 = D2 * D2, e2 = em * em, e4 = e2 * e2;
    lhs = D4 * D2 * D;
    rhs = 1717.0 * (e4 * e2 * em);
    return D > 8.0 * em && fma(2.0, dot, S) * S * (PI * PI) > 500.0 * e2;
}

// P(Binomial(13, F) >= 7) = F^7 sum_{i=0}^{6} C(6 + i, i) (1 - F)^i
__host__ __device__ inline double ipe_G(double F) {
    F = fmin(fmax(F, 0.0), 1.0);
    const double q = 1.0 - F;
    double s = 924.0;
    s = fma(s, q, 462.0);
    s = fma(s, q, 210.0);
    s = fma(s, q, 84.0);
    s = fma(s, q, 28.0);
    s = fma(s, q, 7.0);
    s = fma(s, q, 1.0);
    const double F2 = F * F, F4 = F2 * F2;
    return F4 * F2 * F * s;
}

// the same without the clamp, for arguments already within rounding of [0, 1]
__host__ __device__ inline double ipe_G_raw(double F) {
    const double q = 1.0 - F;
    double s = 924.0;
    s = fma(s, q, 462.0);
    s = fma(s, q, 210.0);
    s = fma(s, q, 84.0);
    s = fma(s, q, 28.0);
    s = fma(s, q, 7.0);
    s = fma(s, q, 1.0);
    const double F2 = F * F, F4 = F2 * F2;
    return F4 * F2 * F * s;
}

__host__ __device__ inline double ipe_f(double M, double d) {
    const double ms = M * sq_sinpi(d / M);
    return 1.0 / (ms * ms);
}

// (M sin(pi z / M))^2 at the 8 window points z = i - 3 - t (grid points j0 - 3 .. j0 + 4), M >= 32, piM = pi / M
__host__ __device__ inline void ipe_window_sq(double piM, double t, double pw[8]) {
    const double PI = 3.14159265358979323846;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        // |pi z / M| < pi/8: M sin(pi z/M) = pi z P(u), u = (pi z/M)^2, Taylor to u^7 (next term < 2e-18)
        const double z = (double)(i - 3) - t, y = z * piM, u = y * y;
        double P = -1.0 / 1307674368000.0;
        P = fma(P, u, 1.0 / 6227020800.0);
        P = fma(P, u, -1.0 / 39916800.0);
        P = fma(P, u, 1.0 / 362880.0);
        P = fma(P, u, -1.0 / 5040.0);
        P = fma(P, u, 1.0 / 120.0);
        P = fma(P, u, -1.0 / 6.0);
        P = fma(P, u, 1.0);
        const double ms = PI * z * P;
        pw[i] = ms * ms;
    }
}

// w_i = s2 / pw_i from r = s2 / (pw_0 ... pw_7): prefix and suffix products, no further division
struct IpeProd8 {
    double a01, a23, a45, a67, lo4, hi4;
    __host__ __device__ inline void of(const double pw[8]) {
        a01 = pw[0] * pw[1]; a23 = pw[2] * pw[3]; a45 = pw[4] * pw[5]; a67 = pw[6] * pw[7];
        lo4 = a01 * a23; hi4 = a45 * a67;
    }
    __host__ __device__ inline void quotients(const double pw[8], double r, double w[8]) const {
        const double rl = r * hi4, rh = r * lo4;
        const double r01 = rl * a23, r23 = rl * a01, r45 = rh * a67, r67 = rh * a45;
        w[0] = r01 * pw[1]; w[1] = r01 * pw[0]; w[2] = r23 * pw[3]; w[3] = r23 * pw[2];
        w[4] = r45 * pw[5]; w[5] = r45 * pw[4]; w[6] = r67 * pw[7]; w[7] = r67 * pw[6];
    }
};

// window weights w_i = sin^2(pi t) f(|i - 3 - t|), i = 0..7, M >= 32
__host__ __device__ inline void ipe_window(double M, double t, double s2, double w[8]) {
    double pw[8];
    ipe_window_sq(3.14159265358979323846 / M, t, pw);
    IpeProd8 pr;
    pr.of(pw);
    pr.quotients(pw, s2 / (pr.lo4 * pr.hi4), w);
}

// sin and cos of a small angle (|y| <= 0.5), to full double precision
__host__ __device__ inline void ipe_sincos_small(double y, double& s, double& c) {
    const double u = y * y;
    double P = -1.0 / 1307674368000.0;
    P = fma(P, u, 1.0 / 6227020800.0);
    P = fma(P, u, -1.0 / 39916800.0);
    P = fma(P, u, 1.0 / 362880.0);
    P = fma(P, u, -1.0 / 5040.0);
    P = fma(P, u, 1.0 / 120.0);
    P = fma(P, u, -1.0 / 6.0);
    P = fma(P, u, 1.0);
    s = y * P;
    double Q = 1.0 / 20922789888000.0;
    Q = fma(Q, u, -1.0 / 87178291200.0);
    Q = fma(Q, u, 1.0 / 479001600.0);
    Q = fma(Q, u, -1.0 / 3628800.0);
    Q = fma(Q, u, 1.0 / 40320.0);
    Q = fma(Q, u, -1.0 / 720.0);
    Q = fma(Q, u, 1.0 / 24.0);
    Q = fma(Q, u, -0.5);
    c = fma(Q, u, 1.0);
}

// Bracket of  sum_{k=0}^{K-1} f(a + k)  (K >= 1, 0 < a, b = a + K - 1 < M) by the Euler-Maclaurin formula.  Every
// even derivative of csc^2 is positive on (0, pi), so the expansion envelops the sum:
//     sum = int_a^b f + (f(a) + f(b))/2 + (f'(b) - f'(a))/12 - th (f'''(b) - f'''(a))/720,   0 <= th <= 1,
// with f' = -(pi/M) 2c / (M^2 s^3), f''' = -(pi/M)^3 8c (2 + c^2) / (M^2 s^5), (s, c) = (sin, cos)(pi d / M), and
// int f = (cot y_a - cot y_b) / (pi M) = sin(y_b - y_a) / (pi M s_a s_b).  From the values at the two ends:
// (s3, c3) at a, (s4, c4) at b, q = 1 / (s3 s4).  The gap is 3e-6 at a = 4 (it falls like a^-5).
__host__ __device__ inline void ipe_bracket_ends(double invM, double piM, double s3, double c3, double s4, double c4,
                                                 double q, double& lo, double& hi) {
    const double r3 = q * s4, r4 = q * s3;                  // 1 / sin at the two ends
    const double r3_2 = r3 * r3, r4_2 = r4 * r4, invM2 = invM * invM;
    const double integral = (s4 * c3 - c4 * s3) * q * (invM * (1.0 / 3.14159265358979323846));
    const double ends = 0.5 * (r3_2 + r4_2) * invM2;
    const double e1 = (c3 * r3_2 * r3 - c4 * r4_2 * r4) * (piM * invM2 * (1.0 / 6.0));
    const double e3 = (c3 * (2.0 + c3 * c3) * (r3_2 * r3_2) * r3 - c4 * (2.0 + c4 * c4) * (r4_2 * r4_2) * r4) *
                      ((piM * piM) * (piM * invM2) * (1.0 / 90.0));
    hi = integral + ends + e1;
    lo = hi - e3;
    // floating-point slack: the brackets are sums of O(10) correctly rounded operations on positive terms
    hi = hi * (1.0 + 1e-12) + 1e-300;
    lo = lo * (1.0 - 1e-12);
    if (lo < 0.0) lo = 0.0;
}

// the general form (ipe_tighten): sines and cosines from the library
__host__ __device__ inline void ipe_run_bracket(double M, double a, double K, double& lo, double& hi) {
    const double invM = 1.0 / M;
    double s3, c3, s4, c4;
    sq_sincospi(a * invM, s3, c3);
    sq_sincospi((a + K - 1.0) * invM, s4, c4);
    ipe_bracket_ends(invM, 3.14159265358979323846 * invM, s3, c3, s4, c4, 1.0 / (s3 * s4), lo, hi);
}

// The common case.  Returns IPE_DONE with the median's r in `med`, or who should take the pair.
__host__ __device__ inline int ipe_lean(const IpeGrid& g, double V, double& med) {
    const double PI = 3.14159265358979323846;
    const double M = g.M, t = g.t, j0 = g.j0;
    // mirror images of the window points sit at distances t + 2 j0 - 3 .. t + 2 j0 + 4; sin(pi d/M) >= 2 min(d, M - d)/M
    const double m = fmin(t + 2.0 * j0 - 3.0, M - (t + 2.0 * j0 + 4.0));
    if (!(M >= 32.0 && j0 >= 4.0 && 2.0 * (j0 + 4.0) < M && m >= 4.0)) {
        // no room for the window.  On a small grid, or with x within a few points of 0 (nearly parallel vectors of
        // like norm), plain summation from r = 0 is short: the same inversion, the same V.  The rest (x near M/2 on a
        // large grid) goes to the general estimator.
        return (M < 64.0 || j0 < 8.0) ? IPE_BRUTE : IPE_GENERAL;
    }
    const double st = sq_sinpi(t), s2 = st * st;
    const double invM = g.invM, piM = PI * invM;
    // the window, and the two ends of the run below it: a = t + 4 (Taylor: pi a / M < 0.5) and b = 2x - a, where
    // pi b / M = 2 theta_a - pi a / M and cos 2 theta_a = 1 - 2a is known
    double pw[8], w[8];
    ipe_window_sq(piM, t, pw);
    IpeProd8 pr;
    pr.of(pw);
    double s3, c3;
    ipe_sincos_small((t + 4.0) * piM, s3, c3);
    const double s2t = sqrt(fmax(1.0 - g.c2t * g.c2t, 0.0));
    const double s4 = s2t * c3 - g.c2t * s3, c4 = g.c2t * c3 + s2t * s3;
    // one division for the eight window weights and the two 1 / sin of the bracket
    const double pwin = pr.lo4 * pr.hi4, pend = s3 * s4;
    const double inv = 1.0 / (pwin * pend);
    pr.quotients(pw, s2 * (inv * pend), w);
    double blo, bhi;
    ipe_bracket_ends(invM, piM, s3, c3, s4, c4, inv * pwin, blo, bhi);
    blo *= s2; bhi *= s2;
    double c0 = w[0], c1 = c0 + w[1], c2 = c1 + w[2], c3_ = c2 + w[3], c4_ = c3_ + w[4], c5 = c4_ + w[5],
           c6 = c5 + w[6], c7 = c6 + w[7];
    double up = bhi + 1e-15;                                // added to cum_i for the upper end of F_i
    const double dn = blo - 1e-15;
    double umu;                                             // what `up` carries for the mirror images
    if (2.0 * s2 > 1e-6 * (m * m)) {
        // sizeable mirror images (small grid, or x near 0 / M/2): their weights exactly.  They sit at the distances
        // b + 1 + i: rotate (sin, cos) of pi b / M by pi / M eight times; one division for the eight reciprocals.
        double sh, ch;
        ipe_sincos_small(piM, sh, ch);
        double sn = s4, cs = c4, pm[8], mw[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const double s_ = sn * ch + cs * sh, c_ = cs * ch - sn * sh;
            sn = s_; cs = c_;
            const double ms = M * sn;
            pm[i] = ms * ms;
        }
        IpeProd8 pq;
        pq.of(pm);
        pq.quotients(pm, s2 / (pq.lo4 * pq.hi4), mw);
        double acc = mw[0];
        c0 += acc; acc += mw[1]; c1 += acc; acc += mw[2]; c2 += acc; acc += mw[3]; c3_ += acc;
        acc += mw[4]; c4_ += acc; acc += mw[5]; c5 += acc; acc += mw[6]; c6 += acc; acc += mw[7]; c7 += acc;
        umu = 1e-14;                                        // rounding of the rotations
    } else {
        umu = 1e-6 * (1.0 + 1e-12);                         // >= 2 s2 / m^2 >= the 8 mirror weights together
    }
    up += umu;
    // V against G(F_i), F_i in [dn + c_i, up + c_i]: certain when both ends of the bracket say the same.
    // Smallest i in 0..7 with V <= G(F_i), taking the median to be in the window; the two edges are checked after.
    const bool a3 = V <= ipe_G_raw(dn + c3_), b3 = V > ipe_G_raw(up + c3_);
    if (a3 == b3) return IPE_TIGHTEN;
    const double cq = a3 ? c1 : c5;
    const bool a2 = V <= ipe_G_raw(dn + cq), b2 = V > ipe_G_raw(up + cq);
    if (a2 == b2) return IPE_TIGHTEN;
    const double cr = a3 ? (a2 ? c0 : c2) : (a2 ? c4_ : c6);
    const bool a1 = V <= ipe_G_raw(dn + cr), b1 = V > ipe_G_raw(up + cr);
    if (a1 == b1) return IPE_TIGHTEN;
    const int i = (a3 ? 0 : 4) + (a2 ? 0 : 2) + (a1 ? 0 : 1);
    if (i == 0) {
        // below the window?  B itself is in [dn, up - umu]
        const bool below = V <= ipe_G_raw(dn), above = V > ipe_G_raw(up - umu);
        if (below == above) return IPE_TIGHTEN;
        if (below) return IPE_BRUTE;
    } else if (i == 7) {
        const bool in7 = V <= ipe_G_raw(dn + c7), out7 = V > ipe_G_raw(up + c7);
        if (in7 == out7) return IPE_TIGHTEN;
        if (out7) return IPE_BRUTE;
    }
    med = j0 - 3.0 + (double)i;
    return IPE_DONE;
}

// The inversion by plain summation: smallest r with V <= G(F(r)).  O(r) - for the rare pairs only.
__host__ __device__ inline double ipe_brute(const IpeGrid& g, double V) {
    const double M = g.M, x = g.x, invM = g.invM;
    const double st = sq_sinpi(g.t), s2 = st * st;
    const double rmax = floor(M * 0.5);
    double F = 0.0;
    for (double r = 0.0; r <= rmax; r += 1.0) {
        // weight of r: the grid point at distance |r - x|, and (0 < r < M/2) its mirror image M - r at distance r + x
        const double ma = M * sq_sinpi(fabs(r - x) * invM), A = ma * ma;
        double P;
        if (r > 0.0 && 2.0 * r < M) {
            const double mb = M * sq_sinpi((r + x) * invM), B = mb * mb;
            P = (A + B) / (A * B);
        } else {
            P = 1.0 / A;
        }
        F += s2 * P;
        if (V <= ipe_G(F)) return r;
    }
    return rmax;
}

// Same decisions as ipe_lean with the mirror weights exact and the bracket of B tightened by peeling exact terms
// off both ends of its run (term k and term K-1-k together are the weight of r = j0 - 4 - k).  Returns IPE_DONE
// or IPE_BRUTE (the median left the window).
__host__ __device__ inline int ipe_tighten(const IpeGrid& g, double V, double& med) {
    const double M = g.M, t = g.t, j0 = g.j0;
    const double st = sq_sinpi(t), s2 = st * st;
    double w[8], c[8];
    ipe_window(M, t, s2, w);
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        acc += w[i] + s2 * ipe_f(M, t + 2.0 * j0 - 3.0 + (double)i);      // + the mirror image of the point
        c[i] = acc;
    }
    const double a = t + 4.0, K = 2.0 * j0 - 7.0;
    double peeled = 0.0, n = 0.0;                          // terms 0..n-1 and K-n..K-1 are in `peeled`
    double target = 8.0;
    while (true) {
        while (n < target && 2.0 * n < K) {
            const double lo_d = a + n, hi_d = a + K - 1.0 - n;
            peeled += ipe_f(M, lo_d);
            if (hi_d > lo_d) peeled += ipe_f(M, hi_d);
            n += 1.0;
        }
        const double rest = K - 2.0 * n;                   // terms n .. K-1-n are still bracketed
        double blo = 0.0, bhi = 0.0;
        if (rest >= 1.0) ipe_run_bracket(M, a + n, rest, blo, bhi);
        const double dn = (peeled + blo) * s2 * (1.0 - 1e-12) - 1e-16;
        const double up = (peeled + bhi) * s2 * (1.0 + 1e-12) + 1e-16;
        const bool exact = !(rest >= 1.0);
        if (V <= ipe_G(up) || V > ipe_G(dn + c[7])) {
            // possibly outside the window; certain once the bracket has closed
            if (exact || (V <= ipe_G(dn)) || (V > ipe_G(up + c[7]))) return IPE_BRUTE;
        } else {
            int i = 0;
            bool sure = true;
#pragma unroll
            for (int p = 0; p < 7; ++p) {                  // first i with V <= G(F_i); F_7 holds
                const bool le_lo = V <= ipe_G(dn + c[p]), le_hi = V <= ipe_G(up + c[p]);
                if (le_lo != le_hi) sure = false;
                if (!le_hi && i == p) i = p + 1;
            }
            if (sure || exact) { med = j0 - 3.0 + (double)i; return IPE_DONE; }
        }
        if (exact) { med = j0 - 3.0; return IPE_DONE; }    // not reached
        target *= 8.0;
    }
}

// the estimate |x|^2 + |c|^2 - 2 ipe(x, c) from the median's r (Utility.py:733-734, :520-531): a function of (S, M, r)
// alone, rounded the same way wherever it is inlined (no fused multiply-add across the call), so that equal outcomes
// give equal distances whichever pass of the kernel produced them
__host__ __device__ inline double ipe_value(const IpeGrid& g, double med) {
    const double sn = sq_sinpi(med * g.invM);
#ifdef __CUDA_ARCH__
    return __dmul_rn(__dmul_rn(2.0 * g.S, sn), sn);
#else
    return 2.0 * g.S * sn * sn;
#endif
}

// estimate - |x|^2, what the arg-min compares (one rounding, never fused with the product above)
__host__ __device__ inline double ipe_minus(double v, double xn) {
#ifdef __CUDA_ARCH__
    return __dsub_rn(v, xn);
#else
    return v - xn;
#endif
}

}  // namespace sq
