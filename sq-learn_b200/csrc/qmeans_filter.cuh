// K2f - the E-step of the exact and delta-window modes as "filter in bf16, decide in fp64".
//
// Only the row's arg-min, its value and the columns within delta of it leave the E-step (labels_estimation,
// sklearn/cluster/_dmeans.py:732-777).  So the n x k distance matrix is first computed APPROXIMATELY on the bf16
// tensor cores (mma.sync.m16n8k16, fp32 accumulation) with a rigorous error bound, which tells, per row, which few
// columns can be the minimum or lie in the window; for exactly those columns the distance is then computed in fp64
// from the fp64 tile still sitting in shared memory, and every decision (arg-min, window, near-tie flag, uniform
// pick) is taken on the fp64 values by the same rules as qmeans_assign_kernel's epilogue.  Nothing approximate
// reaches an output: labels, distances and inertia are fp64 results, and rows the filter cannot narrow down (more
// than CAP columns inside the bound, or non-finite numbers) go to the exact fix-up kernels that already exist.
//
// Error bound.  bf16 has 8 significant bits: |bf16(u) - u| <= 2^-8 |u| (through fp32: + 2^-24), so
// |x~.c~ - x.c| <= (2 * 2^-8 + 2^-16 + d 2^-24 (fp32 accumulation)) |x||c| < 0.0079 |x||c|  (d <= 128).
// With v = |c|^2 - 2 x.c and E = 0.0158 |x| max_j|c_j| (+ the fp32 rounding of |c|^2): |v~ - v| <= E, hence every
// column with v <= min v + delta has v~ <= min v~ + delta + 2E.  Columns passing that test are the candidates.
//
// The kernel is HBM-bound by design: X is read once in fp64 (8 n d bytes) by bulk copies (cp.async.bulk, one per
// tile: the rows of a tile are contiguous) into a two-deep ring; the centroids stay in shared memory as bf16 for the
// whole launch.  One CTA of 512 threads per SM: two teams of eight warps, each on its own 32-row tiles, all k <= 256
// columns at once.  Measured at 10M x 128, k = 256: 6.8 ms = 0.24 of the HBM copy bandwidth (the fp64 tensor-core
// kernel: 20.6 ms); what separates it from the roofline is the instruction count per row, see DESIGN.md section 3c.
#pragma once
#include <cuda_bf16.h>

namespace sq {

struct FilterCfg {
    // TEAMS teams of TWARPS warps per CTA, each working on its own BM-row tiles (own ring, own barriers), sharing the
    // bf16 centroids: what several CTAs per SM would be if the centroids did not have to be there once per CTA
    // (measured: 2 teams x 8 warps x 32 rows 1.35 ms per 2M x 128, 4 x 4 x 16 rows 1.36 ms, one team of 16 warps x 64 rows 1.65 ms)
    static constexpr int TEAMS = 2, TWARPS = 8, TTHREADS = 32 * TWARPS, THREADS = TEAMS * TTHREADS;
    static constexpr int BM = 32, BN = 256, CAP = 8, DMAX = 128;
    static constexpr int MI = BM / 16, WN = BN / TWARPS, NI = WN / 8, NP = NI / 2;   // warp tile: BM rows x WN columns
};

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (unsigned)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar))
                 : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(unsigned& r0, unsigned& r1, unsigned& r2, unsigned& r3, const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"((unsigned)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0,
                                         unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// fp64 -> bf16, round to nearest (ties up), from the high word of the double alone: integer pipe only.  |error| <= 2^-8 |u|
// (+ 2^-21 |u| for the low word that is not looked at); magnitudes below 2^-126 become 0, above bf16's range +-inf
// (the row then has non-finite approximate distances and goes to the exact fix-up).
__device__ __forceinline__ unsigned f64_to_bf16_bits(double u) {
    const unsigned hi = (unsigned)__double2hiint(u);
    const unsigned r = (hi & 0x7fffffffu) + 0x1000u;               // round at bit 13 of the high word
    const unsigned em = r >> 13;                                   // 11-bit exponent : 7-bit mantissa
    unsigned out = em - (896u << 7);                               // re-bias 1023 -> 127
    if (em < (897u << 7)) out = 0u;
    if (em >= (1151u << 7)) out = 0x7f80u;
    return out | ((hi >> 16) & 0x8000u);
}

// dynamic shared memory of the kernel for feature count d (even, <= 128)
__host__ __device__ inline size_t filter_smem_bytes(int d) {
    const size_t cp = (size_t)((d + 15) & ~15) + 8;               // bf16 row pitch (d padded to 16): rows 16 bytes apart modulo 128
    const size_t T = FilterCfg::TEAMS, BM = FilterCfg::BM;
    return T * 2 * BM * d * 8                                      // xs: two fp64 tiles per team
           + (size_t)FilterCfg::BN * cp * 2                        // cb: centroids, bf16
           + T * BM * cp * 2                                       // xb: the teams' tiles, bf16
           + (size_t)FilterCfg::BN * 8                             // cn_s
           + T * BM * FilterCfg::TWARPS * 4                        // wmin
           + T * BM * 4 * 3                                        // thr_s, slack_s, ccnt
           + T * BM * FilterCfg::CAP * 4                           // ccol
           + T * BM * FilterCfg::CAP * 8                           // cval
           + 8                                                     // cmax
           + 8 * T * 2 + 64;                                       // mbarriers, padding
}

template <int MODE>
__global__ void __launch_bounds__(FilterCfg::THREADS, 1) qmeans_filter_kernel(AssignArgs a) {
    constexpr int BM = FilterCfg::BM, BN = FilterCfg::BN, CAP = FilterCfg::CAP, WARPS = FilterCfg::TWARPS;
    constexpr int TEAMS = FilterCfg::TEAMS, TT = FilterCfg::TTHREADS;
    constexpr int MI = FilterCfg::MI, NI = FilterCfg::NI, NP = FilterCfg::NP, WN = FilterCfg::WN;
    extern __shared__ __align__(1024) unsigned char fsm[];
    const int d = a.d, d16 = (d + 15) & ~15, cp = d16 + 8;        // (columns d .. d16 of the bf16 tiles stay zero)
    const int team = threadIdx.x / TT, tid = threadIdx.x % TT, lane = tid & 31, warp = tid >> 5;
    double* xs_all = reinterpret_cast<double*>(fsm);                                   // [TEAMS][2][BM*d]
    __nv_bfloat16* cb = reinterpret_cast<__nv_bfloat16*>(xs_all + (size_t)TEAMS * 2 * BM * d);   // [BN][cp]
    __nv_bfloat16* xb_all = cb + (size_t)BN * cp;                                      // [TEAMS][BM][cp]
    double* cn_s = reinterpret_cast<double*>(xb_all + (size_t)TEAMS * BM * cp);        // [BN]
    float* wmin_all = reinterpret_cast<float*>(cn_s + BN);                             // [TEAMS][BM][WARPS]
    float* thr_all = wmin_all + TEAMS * BM * WARPS;                                    // [TEAMS][BM]
    float* slack_all = thr_all + TEAMS * BM;                                           // [TEAMS][BM]
    int* ccnt_all = reinterpret_cast<int*>(slack_all + TEAMS * BM);                    // [TEAMS][BM]
    int* ccol_all = ccnt_all + TEAMS * BM;                                             // [TEAMS][BM][CAP]
    double* cval_all = reinterpret_cast<double*>(((uintptr_t)(ccol_all + TEAMS * BM * CAP) + 7) & ~(uintptr_t)7);   // [TEAMS][BM][CAP]
    double* cmax_s = cval_all + TEAMS * BM * CAP;                                      // [1] max_j |c_j|
    unsigned long long* bar_all = reinterpret_cast<unsigned long long*>(cmax_s + 1);   // [TEAMS][2]
    // this team's slices
    double* xs = xs_all + (size_t)team * 2 * BM * d;
    __nv_bfloat16* xb = xb_all + (size_t)team * BM * cp;
    float* wmin = wmin_all + team * BM * WARPS;
    float* thr_s = thr_all + team * BM;
    float* slack_s = slack_all + team * BM;
    int* ccnt = ccnt_all + team * BM;
    int* ccol = ccol_all + team * BM * CAP;
    double* cval = cval_all + team * BM * CAP;
    unsigned long long* bar = bar_all + team * 2;
    auto team_sync = [&]() { asm volatile("barrier.sync %0, %1;" ::"r"(1 + team), "r"(TT) : "memory"); };

    const int g = lane >> 2, t4 = lane & 3;
    const long long num_tiles = (a.n + BM - 1) / BM;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (d16 != d) {
        // the padding columns of the bf16 tiles are zero for the whole launch
        uint4* z = reinterpret_cast<uint4*>(cb);
        const int n16 = (int)(((size_t)BN * cp + (size_t)TEAMS * BM * cp) * 2 / 16);
        for (int i = threadIdx.x; i < n16; i += FilterCfg::THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
        __syncthreads();
    }
    // centroids: fp64 -> bf16 (zero rows past k), by the whole CTA
    // (pairs of features; a thread walks (row, pair) = (t / h, t % h), then + stride, without dividing again)
    const int h = d / 2;
    {
        const int t = threadIdx.x, sr = FilterCfg::THREADS / h, sc = FilterCfg::THREADS - sr * h;
        for (int j = t / h, c2 = t % h; j < BN;) {
            double2 v = make_double2(0.0, 0.0);
            if (j < a.k) v = *reinterpret_cast<const double2*>(a.C + (long long)j * d + 2 * c2);
            *reinterpret_cast<unsigned*>(cb + (size_t)j * cp + 2 * c2) = f64_to_bf16_bits(v.x) | (f64_to_bf16_bits(v.y) << 16);
            j += sr; c2 += sc;
            if (c2 >= h) { c2 -= h; ++j; }
        }
        for (int j = t; j < BN; j += FilterCfg::THREADS) cn_s[j] = (j < a.k) ? a.cnorm[j] : INFINITY;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        double m = 0.0;
        for (int j = lane; j < a.k; j += 32) m = fmax(m, cn_s[j]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) *cmax_s = sqrt(m);
    }
    __syncthreads();
    const double cmax = *cmax_s;
    const int step_r = TT / h, step_c = TT - step_r * h;
    const int r_first = tid / h, c_first = tid - r_first * h;

    auto issue = [&](long long tile, int stage) {
        // one thread: the whole tile is one contiguous block of X (rows of d doubles, no padding)
        const long long r0 = tile * BM;
        const long long rows = (a.n - r0 < BM) ? (a.n - r0) : BM;
        const unsigned bytes = (unsigned)(rows * d * 8);
        mbar_expect_tx(&bar[stage], bytes);
        bulk_load(xs + (size_t)stage * BM * d, a.X + r0 * d, bytes, &bar[stage]);
    };

    const long long tstride = (long long)gridDim.x * TEAMS;
    long long tile = (long long)blockIdx.x * TEAMS + team;
    if (tid == 0 && tile < num_tiles) issue(tile, 0);
    unsigned phases = 0u;                                  // bit s: parity to wait for on bar[s]
    int stage = 0;
    // relative error of a bf16 product sum (see the header) and of |c|^2 rounded to fp32
    const float EREL = 0.0158f;

    for (; tile < num_tiles; tile += tstride, stage ^= 1) {
        const long long r0 = tile * BM;
        const int rows = (int)((a.n - r0 < BM) ? (a.n - r0) : BM);
        // the team's next tile goes into its other buffer (whose readers finished before the barrier that ended the
        // previous iteration)
        if (tid == 0 && tile + tstride < num_tiles) issue(tile + tstride, stage ^ 1);
        mbar_wait(&bar[stage], (phases >> stage) & 1u);
        phases ^= 1u << stage;
        const double* xt = xs + (size_t)stage * BM * d;

        // ---- fp64 tile -> bf16 tile; per-row slack of the filter ----
        for (int r = r_first, c2 = c_first; r < BM;) {
            double2 v = *reinterpret_cast<const double2*>(xt + (size_t)r * d + 2 * c2);
            if (r >= rows) v = make_double2(0.0, 0.0);
            *reinterpret_cast<unsigned*>(xb + (size_t)r * cp + 2 * c2) = f64_to_bf16_bits(v.x) | (f64_to_bf16_bits(v.y) << 16);
            r += step_r; c2 += step_c;
            if (c2 >= h) { c2 -= h; ++r; }
        }
        if (tid < BM) {
            const double xn = (tid < rows) ? a.xnorm[r0 + tid] : 0.0;
            // E = 0.0158 |x| max|c| bounds |v~ - v|, plus the rounding of |c|^2 to fp32 and of the fp32 arithmetic on v~
            const double E = (double)EREL * sqrt(xn) * cmax + 2.4e-7 * (cmax * cmax + 2.0 * sqrt(xn) * cmax) + 1e-30;
            slack_s[tid] = __double2float_ru(2.0 * E);
            ccnt[tid] = 0;
        }
        team_sync();

        // ---- approximate x.c on the tensor cores: warp w of the team owns WN columns of the tile's rows ----
        float acc[MI][NI][4];
#pragma unroll
        for (int mi = 0; mi < MI; ++mi)
#pragma unroll
            for (int ni = 0; ni < NI; ++ni)
#pragma unroll
                for (int r = 0; r < 4; ++r) acc[mi][ni][r] = 0.0f;
        const int a_row = (lane & 7) + 8 * ((lane >> 3) & 1), a_k = 8 * (lane >> 4);
        const int b_row = (lane & 7) + 8 * (lane >> 4), b_k = 8 * ((lane >> 3) & 1);
        for (int k0 = 0; k0 < d16; k0 += 16) {
            unsigned af[MI][4], bf[NP][4];
#pragma unroll
            for (int mi = 0; mi < MI; ++mi)
                ldmatrix_x4(af[mi][0], af[mi][1], af[mi][2], af[mi][3], xb + (size_t)(mi * 16 + a_row) * cp + k0 + a_k);
#pragma unroll
            for (int np = 0; np < NP; ++np)
                ldmatrix_x4(bf[np][0], bf[np][1], bf[np][2], bf[np][3],
                            cb + (size_t)(warp * WN + np * 16 + b_row) * cp + k0 + b_k);
#pragma unroll
            for (int mi = 0; mi < MI; ++mi)
#pragma unroll
                for (int ni = 0; ni < NI; ++ni)
                    mma_bf16(acc[mi][ni], af[mi][0], af[mi][1], af[mi][2], af[mi][3], bf[ni >> 1][(ni & 1) * 2],
                             bf[ni >> 1][(ni & 1) * 2 + 1]);
        }
        // ---- v~ = |c|^2 - 2 x~.c~ in place; per-warp row minima ----
#pragma unroll
        for (int ni = 0; ni < NI; ++ni) {
            const int col = warp * WN + ni * 8 + 2 * t4;
            const float c0 = (float)cn_s[col], c1 = (float)cn_s[col + 1];
#pragma unroll
            for (int mi = 0; mi < MI; ++mi) {
                acc[mi][ni][0] = fmaf(-2.0f, acc[mi][ni][0], c0);
                acc[mi][ni][1] = fmaf(-2.0f, acc[mi][ni][1], c1);
                acc[mi][ni][2] = fmaf(-2.0f, acc[mi][ni][2], c0);
                acc[mi][ni][3] = fmaf(-2.0f, acc[mi][ni][3], c1);
            }
        }
#pragma unroll
        for (int mi = 0; mi < MI; ++mi)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                float m = INFINITY;
#pragma unroll
                for (int ni = 0; ni < NI; ++ni) m = fminf(m, fminf(acc[mi][ni][2 * hh], acc[mi][ni][2 * hh + 1]));
                m = fminf(m, __shfl_xor_sync(0xffffffffu, m, 1));
                m = fminf(m, __shfl_xor_sync(0xffffffffu, m, 2));
                if (t4 == 0) wmin[(mi * 16 + hh * 8 + g) * WARPS + warp] = m;
            }
        team_sync();
        if (tid < BM) {
            float m = wmin[tid * WARPS];
#pragma unroll
            for (int w = 1; w < WARPS; ++w) m = fminf(m, wmin[tid * WARPS + w]);
            // every column with v <= min v + delta has v~ <= min v~ + delta + 2E; a row without a finite minimum
            // (non-finite input) goes to the exact fix-up
            thr_s[tid] = (m + (MODE == MODE_WINDOW ? __double2float_ru(a.delta) : 0.0f) + slack_s[tid]) * (1.0f + 1e-6f) +
                         1e-30f;
            if (!(m < INFINITY)) ccnt[tid] = CAP + 1;
        }
        team_sync();
        // ---- candidates (columns past k carry +inf and never pass; rows past the end are never looked at again):
        // a bit per accumulator first, then a loop over the few hits ----
        {
            unsigned hits = 0u;
#pragma unroll
            for (int mi = 0; mi < MI; ++mi)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const float thr = thr_s[mi * 16 + hh * 8 + g];
#pragma unroll
                    for (int ni = 0; ni < NI; ++ni)
#pragma unroll
                        for (int e = 0; e < 2; ++e)
                            hits |= (!(acc[mi][ni][2 * hh + e] > thr) ? 1u : 0u) << (((mi * 2 + hh) * NI + ni) * 2 + e);
                }
            static_assert(MI * 2 * NI * 2 <= 32, "one bit per accumulator");
            while (hits) {
                const int b = __ffs(hits) - 1;
                hits &= hits - 1;
                const int e = b & 1, ni = (b >> 1) % NI, mh = (b >> 1) / NI;
                const int row = (mh >> 1) * 16 + (mh & 1) * 8 + g;
                const int slot = atomicAdd(&ccnt[row], 1);
                if (slot < CAP) ccol[row * CAP + slot] = warp * WN + ni * 8 + 2 * t4 + e;
            }
        }
        team_sync();
        // ---- fp64 for the candidates: a warp owns the rows warp, warp + 8, ...; its candidates, flattened, are dealt to
        // its eight groups of four lanes, so that eight fp64 dot products (their centroid rows come from L2) are in
        // flight at a time.  The summation order of a dot product depends on d alone. ----
        {
            constexpr int RPW = BM / WARPS;                    // rows per warp
            const int grp = lane >> 2, sub = lane & 3;
            int pre[RPW + 1];
            pre[0] = 0;
#pragma unroll
            for (int j = 0; j < RPW; ++j) {
                const int row = warp + WARPS * j;
                const int nc = (row < rows) ? ccnt[row] : 0;
                pre[j + 1] = pre[j] + ((nc >= 1 && nc <= CAP) ? nc : 0);
            }
            const int total = pre[RPW];
            for (int cbase = 0; cbase < total; cbase += 8) {       // (the same trip count for every lane: full-mask shuffles below)
                const bool live = cbase + grp < total;
                const int c = live ? cbase + grp : 0;
                int j = 0;
#pragma unroll
                for (int q = 1; q < RPW; ++q) j += (c >= pre[q]) ? 1 : 0;
                int base = 0;
#pragma unroll
                for (int q = 0; q < RPW; ++q) base = (q == j) ? pre[q] : base;
                const int row = warp + WARPS * j, slot = c - base;
                const int col = ccol[row * CAP + slot];
                const double* xr = xt + (size_t)row * d;
                const double* cr = a.C + (long long)col * d;
                double p0 = 0.0, p1 = 0.0;
                for (int f = 2 * sub; f < d; f += 8) {
                    const double2 xv = *reinterpret_cast<const double2*>(xr + f);
                    const double2 cv = __ldg(reinterpret_cast<const double2*>(cr + f));
                    p0 = fma(xv.x, cv.x, p0);
                    p1 = fma(xv.y, cv.y, p1);
                }
                double p = p0 + p1;
                p += __shfl_xor_sync(0xffffffffu, p, 2);
                p += __shfl_xor_sync(0xffffffffu, p, 1);
                if (live && sub == 0) cval[row * CAP + slot] = fma(-2.0, p, cn_s[col]);
            }
            __syncwarp();
            // ---- the decision, one lane per row ----
            if (lane < RPW && warp + WARPS * lane < rows) {
                const int row = warp + WARPS * lane;
                const long long gr = r0 + row;
                const int nc = ccnt[row];
                const double xn = a.xnorm[gr];
                if (nc > CAP || nc < 1) {
                    // too many columns inside the bound (or non-finite numbers): the exact fix-up kernels redo the row
                    const int slot = atomicAdd(a.flag_count, 1);
                    a.flag_rows[slot] = (int)gr;
                    a.labels[gr] = 0;
                    a.mindist[gr] = 0.0;
                } else {
                    double tmin = INFINITY;
                    int amin = 0x7fffffff;
                    for (int s2 = 0; s2 < nc; ++s2) {
                        const double v = cval[row * CAP + s2];
                        const int cc = ccol[row * CAP + s2];
                        if (v < tmin || (v == tmin && cc < amin)) { tmin = v; amin = cc; }
                    }
                    if (!(tmin < INFINITY)) {                  // NaN / inf in fp64 too
                        const int slot = atomicAdd(a.flag_count, 1);
                        a.flag_rows[slot] = (int)gr;
                        a.labels[gr] = 0;
                        a.mindist[gr] = 0.0;
                    } else if (MODE == MODE_WINDOW) {
                        const double thr = tmin + a.delta;
                        int c = 0, c0 = 0x7fffffff, c1 = 0x7fffffff, c2 = 0x7fffffff, c3 = 0x7fffffff;
                        for (int s2 = 0; s2 < nc; ++s2)
                            if (cval[row * CAP + s2] <= thr) {
                                // keep the four lowest columns, sorted
                                int x = ccol[row * CAP + s2], t;
                                if (x < c0) { t = c0; c0 = x; x = t; }
                                if (x < c1) { t = c1; c1 = x; x = t; }
                                if (x < c2) { t = c2; c2 = x; x = t; }
                                if (x < c3) { c3 = x; }
                                ++c;
                            }
                        int label = c0;
                        if (c > 4) {
                            const int slot = atomicAdd(a.flag_count, 1);
                            a.flag_rows[slot] = (int)gr;          // redone by qmeans_window_fixup_kernel
                        } else if (c > 1) {
                            // uniform pick among the candidates (reference: random.choice, _dmeans.py:2252-2257)
                            Philox rng(a.seed, a.offset + (unsigned long long)gr, 0x51CEu, 0u);
                            int r = (int)(rng.uniform() * c);
                            if (r >= c) r = c - 1;
                            label = r == 0 ? c0 : r == 1 ? c1 : r == 2 ? c2 : c3;
                        }
                        a.labels[gr] = label;
                        a.mindist[gr] = fmax(xn + tmin, 0.0);
                    } else {
                        const double thr = tmin + fmax(a.tie_tol, 0.0) * (xn + 2.0 * fabs(tmin));
                        int rc = 0;
                        for (int s2 = 0; s2 < nc; ++s2) rc += (cval[row * CAP + s2] <= thr) ? 1 : 0;
                        a.labels[gr] = amin;
                        a.mindist[gr] = fmax(xn + tmin, 0.0);
                        if (rc > 1 && a.tie_tol >= 0.0) {
                            const int slot = atomicAdd(a.flag_count, 1);
                            a.flag_rows[slot] = (int)gr;          // redone by qmeans_fixup_kernel in the reference's arithmetic
                        }
                    }
                }
            }
        }
        team_sync();   // xs[stage], xb, ccnt, ccol free for the team's next tiles
    }
}

}  // namespace sq
