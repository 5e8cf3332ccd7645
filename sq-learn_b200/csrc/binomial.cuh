// Exact binomial sampler for counts up to 2^52, used to draw the multinomial
// measurement counts of the tomography kernel by conditional splitting.
//
// Reference behaviour being simulated: QuantumState.measure draws N outcomes
// with RandomState.choice(p=probabilities) (Utility.py:51-55) and
// estimate_wald counts them (Utility.py:61-64).  The vector of counts is
// Multinomial(N, p); splitting N down a binary tree of bucket masses with one
// Binomial(n_node, mass_left / mass_node) per node yields exactly that law at
// O(#buckets) cost instead of O(N).
//
// Algorithms (published, restated here): sequential inversion (BINV,
// Kachitvichyanukul & Schmeiser 1988) when n*min(p,q) < 10, otherwise
// Hormann's transformed rejection with squeeze (BTRS, 1993) with Stirling
// tail corrections.  Counts are carried as doubles (exact integers < 2^53).
#pragma once
#include "common.cuh"

namespace sq {

// Stirling series remainder: log(k!) - [ (k+1/2) log(k+1) - (k+1) + log(sqrt(2 pi)) ]
#ifdef __CUDACC__
__constant__ const double kStirlingTailDev[10] = {0.08106146679532726, 0.04134069595540929, 0.02767792568499834,
                                                  0.02079067210376509, 0.01664469118982119, 0.01387612882307075,
                                                  0.01189670994589177, 0.01041126526197209, 0.009255462182712733,
                                                  0.008330563433362871};
#endif
__host__ __device__ inline double stirling_tail(double k) {
    if (k <= 9.0) {
#ifdef __CUDA_ARCH__
        return kStirlingTailDev[(int)k];
#else
        const double tab[10] = {0.08106146679532726, 0.04134069595540929, 0.02767792568499834, 0.02079067210376509,
                                0.01664469118982119, 0.01387612882307075, 0.01189670994589177, 0.01041126526197209,
                                0.009255462182712733, 0.008330563433362871};
        return tab[(int)k];
#endif
    }
    // (1/12 - (1/360 - 1/1260/(k+1)^2)/(k+1)^2)/(k+1) with one division
    const double r = 1.0 / (k + 1.0), r2 = r * r;
    return r * (1.0 / 12.0 - r2 * (1.0 / 360.0 - r2 * (1.0 / 1260.0)));
}

// Binomial(n, p) with p = pl / (pl + pr).  pl, pr >= 0 are the two bucket
// masses, passed separately so that q = pr / (pl + pr) keeps full relative
// accuracy when p is close to 1.
__host__ __device__ inline double binomial_split(double n, double pl, double pr, Philox& rng) {
    if (n <= 0.0 || pl <= 0.0) return 0.0;
    if (pr <= 0.0) return n;
    double tot = pl + pr;
    bool flip = pl > pr;
    double p = (flip ? pr : pl) / tot;  // p <= 1/2
    double q = (flip ? pl : pr) / tot;
    double k;
    double np = n * p;
    if (np < 10.0) {
        // --- BINV: walk the pmf from 0 upwards ---
        double s = p / q;
        double a = (n + 1.0) * s;
        double r0 = exp(n * log1p(-p));
        double bound = fmin(n, np + 10.0 * sqrt(np * q + 1.0));
        while (true) {
            double u = rng.uniform();
            double r = r0;
            double x = 0.0;
            bool ok = true;
            while (u > r) {
                u -= r;
                x += 1.0;
                if (x > bound) { ok = false; break; }
                r *= (a / x - s);
            }
            if (ok) { k = x; break; }
        }
    } else {
        // --- BTRS ---
        double npq = np * q;
        double spq = sqrt(npq);
        double b = 1.15 + 2.53 * spq;
        double a = -0.0873 + 0.0248 * b + 0.01 * p;
        double c = np + 0.5;
        double vr = 0.92 - 4.2 / b;
        double alpha = (2.83 + 5.1 / b) * spq;
        double r = p / q;
        double m = floor((n + 1.0) * p);
        // the part of the acceptance bound that does not depend on the candidate k
        const double nm = n - m + 1.0;
        const double hm = (m + 0.5) * log((m + 1.0) / (r * nm)) + stirling_tail(m) + stirling_tail(n - m);
        while (true) {
            double u, v;
            rng.uniform2(u, v);
            u -= 0.5;
            double us = 0.5 - fabs(u);
            k = floor((2.0 * a / us + b) * u + c);
            if (us >= 0.07 && v <= vr) break;
            if (k < 0.0 || k > n) continue;
            v = log(v * alpha / (a / (us * us) + b));
            double nk = n - k + 1.0;
            double ub = hm + (n + 1.0) * log1p((k - m) / nk) + (k + 0.5) * log(r * nk / (k + 1.0)) -
                        stirling_tail(k) - stirling_tail(n - k);
            if (v <= ub) break;
        }
    }
    return flip ? n - k : k;
}

// First step of binomial_split, for callers that handle many nodes at once: settles the trivial cases and the
// (common) immediate acceptance of BTRS's first candidate, which needs one sqrt, four divisions and no logarithm.
// Returns BINOM_DONE with the count in `out`, or says which branch of binomial_split the node needs.  A node
// that is not done is handed, with a FRESH generator of the same stream, to binomial_split, which replays the
// first candidate and carries on - so head + split together return exactly what binomial_split alone returns.
// (tomography.cu gathers the unfinished nodes of a tree level and runs them densely, instead of letting one
// rejected lane hold its 31 neighbours through the logarithms of the full acceptance test.)
enum { BINOM_DONE = 0, BINOM_NEEDS_BINV = 1, BINOM_NEEDS_BTRS = 2 };
__host__ __device__ inline int binomial_head(double n, double pl, double pr, Philox& rng, double& out) {
    if (n <= 0.0 || pl <= 0.0) { out = 0.0; return BINOM_DONE; }
    if (pr <= 0.0) { out = n; return BINOM_DONE; }
    double tot = pl + pr;
    bool flip = pl > pr;
    double p = (flip ? pr : pl) / tot;
    double q = (flip ? pl : pr) / tot;
    double np = n * p;
    if (np < 10.0) return BINOM_NEEDS_BINV;
    double npq = np * q;
    double spq = sqrt(npq);
    double b = 1.15 + 2.53 * spq;
    double a = -0.0873 + 0.0248 * b + 0.01 * p;
    double c = np + 0.5;
    double vr = 0.92 - 4.2 / b;
    double u, v;
    rng.uniform2(u, v);
    u -= 0.5;
    double us = 0.5 - fabs(u);
    double k = floor((2.0 * a / us + b) * u + c);
    if (us >= 0.07 && v <= vr) { out = flip ? n - k : k; return BINOM_DONE; }
    return BINOM_NEEDS_BTRS;
}

}  // namespace sq
