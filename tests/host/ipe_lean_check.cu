// Host-side check of csrc/ipe_lean.cuh (the functions are __host__ __device__): for every uniform V the lean
// inversion (bracketed thresholds, then ipe_tighten, then plain summation) must return what plain summation alone
// returns - that is what "exact in law" means for the IPE E-step.  Prints one line per grid regime and the number of
// disagreements; exit status 0 iff there are none.
#include <cmath>
#include <cstdio>
#include <random>
#include "ipe_lean.cuh"
using namespace sq;

int main() {
    std::mt19937_64 gen(1);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    struct Cfg { double M, x; } cfgs[] = {{920, 230.37}, {920, 231.5}, {920, 100.001}, {920, 350.999}, {27001, 6750.25},
        {300, 75.3}, {300, 20.5}, {300, 140.2}, {64, 16.4}, {32, 8.5}, {5000, 4.2}, {5000, 8.9}, {5000, 2480.3}, {1000, 490.6},
        {401, 196.49}, {920, 455.2}, {100000, 50.5}, {2001, 1000.3}, {2001, 12.0001}, {33, 4.5}, {1000, 3.9}, {1000, 496.2}};
    long long total_bad = 0, lean_cfgs = 0;
    for (auto c : cfgs) {
        IpeGrid g;
        g.S = 1.0; g.M = c.M; g.invM = 1.0 / c.M; g.x = c.x; g.j0 = floor(c.x); g.t = c.x - g.j0; g.kind = 2;
        g.c2t = cos(2.0 * 3.14159265358979323846 * c.x / c.M);
        const long long n = (c.M > 20000) ? 4000 : 40000;
        long long cnt[4] = {0, 0, 0, 0}, bad = 0;
        for (long long i = 0; i < n; ++i) {
            double V = U(gen);
            if (i % 7 == 0) V = pow(V, 6.0) * 1e-3 + 1e-9;            // the far ends of the law too
            if (i % 11 == 0) V = 1.0 - pow(V, 6.0) * 1e-3 - 1e-9;
            double med = -1.0;
            int st = ipe_lean(g, V, med);
            cnt[st]++;
            if (st == IPE_GENERAL) continue;                         // declined before V was looked at
            if (st == IPE_TIGHTEN) st = ipe_tighten(g, V, med);
            if (st == IPE_BRUTE) med = ipe_brute(g, V);
            if (ipe_brute(g, V) != med) ++bad;
        }
        printf("M=%-7g x=%-10g lean %.4f declined %.4f tighten %.5f summed %.5f  disagreements %lld\n", c.M, c.x,
               (double)cnt[0] / n, (double)cnt[1] / n, (double)cnt[2] / n, (double)cnt[3] / n, bad);
        total_bad += bad;
        lean_cfgs += cnt[0] > 0;
    }
    printf("TOTAL disagreements %lld, regimes served by the lean path %lld\n", total_bad, lean_cfgs);

    // The certificate that lets the kernel leave a pair out (ipe_out_possible): whenever it holds, the full
    // computation with the same uniform must give an estimate above the reference estimate.  The uniform is put just
    // above the certificate's own threshold, where it is weakest.
    long long cert_n = 0, cert_bad = 0;
    for (int regime = 0; regime < 6; ++regime) {
        for (long long i = 0; i < 40000; ++i) {
            const double eps = (regime >= 4) ? 1e-3 : (regime == 3) ? 0.05 : 0.25;
            double xn = (regime == 0) ? 1.0 : 20.0 + 2000.0 * U(gen);
            double cn = (regime == 0) ? 1.0 : (regime == 1 ? 1.0 : xn * (0.5 + U(gen)));
            if (regime >= 4) cn = xn * (1.0 + 1e-3 * U(gen));
            double dot = (2.0 * U(gen) - 1.0) * sqrt(xn * cn);
            if (regime == 4) dot = -sqrt(xn * cn) * (1.0 - pow(10.0, -1.0 - 6.0 * U(gen)));   // nearly opposite: x next to M/2
            if (regime == 5) dot = sqrt(xn * cn) * (1.0 - pow(10.0, -1.0 - 6.0 * U(gen)));    // nearly equal: x next to 0
            IpeGrid g;
            double a;
            if (!ipe_grid_head(dot, xn, cn, eps, g, a)) continue;
            // a reference estimate somewhere below this pair's true distance
            const double e_star = (2.0 * g.S * a) * U(gen) - xn;
            double lhs, rhs;
            if (!ipe_out_possible(dot, xn, cn, eps, e_star, lhs, rhs)) continue;
            const double V = rhs / lhs * (1.0 + 1e-6 + U(gen) * ((i & 1) ? 1.0 : 1e-3));
            if (!(V < 1.0) || !(V * lhs > rhs)) continue;
            ++cert_n;
            ipe_grid_tail(a, g);
            double v;
            if (g.kind == 1) v = ipe_value(g, fmin(g.j0, g.M - g.j0));
            else {
                double med = -1.0;
                int st = ipe_lean(g, V, med);
                if (st == IPE_TIGHTEN) st = ipe_tighten(g, V, med);
                if (st == IPE_BRUTE) { med = ipe_brute(g, V); st = IPE_DONE; }
                if (st != IPE_DONE) { ++cert_bad; continue; }      // a pair of the general estimator must never be left out
                v = ipe_value(g, med);
            }
            if (!(ipe_minus(v, xn) > e_star)) ++cert_bad;
        }
    }
    printf("certificate: held for %lld pairs, violations %lld\n", cert_n, cert_bad);
    return (total_bad != 0 || lean_cfgs < 15 || cert_bad != 0 || cert_n < 10000) ? 1 : 0;
}
