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
    return (total_bad != 0 || lean_cfgs < 15) ? 1 : 0;
}
