// Shared device/host helpers for the sqb200 kernels (sm_100a only).
//
// Nothing here mirrors a reference file: the reference (sq-learn) is pure
// NumPy.  The helpers are the building blocks the kernels share: a
// counter-based Philox4x32-10 generator (so every random draw is a pure
// function of (seed, stream ids) and never of the launch geometry), warp
// reductions, and the error plumbing behind the C ABI in include/sqb200.h.
#pragma once

#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched at run time, no libcuda link)
#include <cuda_runtime.h>
#include <stdint.h>

#if !defined(__CUDA_ARCH__) || (__CUDA_ARCH__ >= 1000)
#define SQ_ARCH_OK 1
#endif

namespace sq {

// ---------------------------------------------------------------------------
// Error plumbing (host).  Every extern "C" entry returns 0 on success and a
// non-zero code otherwise; the message is kept per thread.
// ---------------------------------------------------------------------------
enum Status : int {
    SQ_OK = 0,
    SQ_ERR_INVALID = 1,    // bad argument
    SQ_ERR_CUDA = 2,       // a CUDA runtime call failed
    SQ_ERR_WORKSPACE = 3,  // caller-supplied workspace too small
    SQ_ERR_UNSUPPORTED = 4 // shape outside what this build supports
};

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define SQ_CUDA(call)                                        \
    do {                                                     \
        cudaError_t _e = (call);                             \
        if (_e != cudaSuccess) return ::sq::cuda_fail(_e, #call); \
    } while (0)

#define SQ_REQUIRE(cond, ...)                \
    do {                                     \
        if (!(cond)) {                       \
            ::sq::set_error(__VA_ARGS__);    \
            return ::sq::SQ_ERR_INVALID;     \
        }                                    \
    } while (0)

#define SQ_LAUNCH_CHECK(what) SQ_CUDA(cudaPeekAtLastError())

int device_sm_count();
// 2-D tensor map over a row-major fp64 matrix [rows, cols] with a box of {box_cols, box_rows}; 0 = ok
int make_tensor_map_2d(CUtensorMap* map, const double* base, unsigned long long cols, unsigned long long rows,
                       unsigned box_cols, unsigned box_rows);

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11).  Counter = 128 bit, key = 64 bit.
// ---------------------------------------------------------------------------
struct Philox {
    uint32_t k0, k1;      // key
    uint32_t c0, c1, c2, c3;  // counter; c0 is bumped per 128-bit block

    __host__ __device__ Philox() {}
    // (seed) keys the generator; (a, b, c) name the stream: e.g. vector id,
    // attempt/stage, tree node.  c0 counts blocks inside the stream.
    // The stream tag b enters the key additively (k1 = seed_hi + b * golden ratio): within one seed every
    // (a, b, c) is its own stream, which is all the kernels rely on; two DIFFERENT seeds whose high words differ
    // by (b - b') * 0x9E3779B9 mod 2^32 share streams between tags b and b' - seeds that are meant to be
    // independent should differ in their low word (sqlearn.QuantumUtility.seed() values do).
    __host__ __device__ Philox(uint64_t seed, uint64_t a, uint32_t b, uint32_t c) {
        k0 = (uint32_t)seed;
        k1 = (uint32_t)(seed >> 32) + b * 0x9E3779B9u;
        c0 = 0;
        c1 = (uint32_t)a;
        c2 = (uint32_t)(a >> 32);
        c3 = c;
    }

    __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#ifdef __CUDA_ARCH__
        lo = a * b;
        hi = __umulhi(a, b);
#else
        uint64_t p = (uint64_t)a * b;
        lo = (uint32_t)p;
        hi = (uint32_t)(p >> 32);
#endif
    }

    __host__ __device__ inline void block(uint32_t out[4]) {
        uint32_t x0 = c0, x1 = c1, x2 = c2, x3 = c3;
        uint32_t ka = k0, kb = k1;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t hi0, lo0, hi1, lo1;
            mulhilo(0xD2511F53u, x0, hi0, lo0);
            mulhilo(0xCD9E8D57u, x2, hi1, lo1);
            uint32_t y0 = hi1 ^ x1 ^ ka;
            uint32_t y1 = lo1;
            uint32_t y2 = hi0 ^ x3 ^ kb;
            uint32_t y3 = lo0;
            x0 = y0; x1 = y1; x2 = y2; x3 = y3;
            ka += 0x9E3779B9u;
            kb += 0xBB67AE85u;
        }
        out[0] = x0; out[1] = x1; out[2] = x2; out[3] = x3;
        ++c0;
    }

    // Two independent uniforms in the open interval (0,1), 53 bits each.
    __host__ __device__ inline void uniform2(double& u, double& v) {
        uint32_t r[4];
        block(r);
        u = to_unit(r[0], r[1]);
        v = to_unit(r[2], r[3]);
    }

    __host__ __device__ inline double uniform() {
        double u, v;
        uniform2(u, v);
        return u;
    }

    __host__ __device__ static inline double to_unit(uint32_t hi, uint32_t lo) {
        // 52 random bits + half-ulp offset: strictly inside (0,1)
        uint64_t bits = ((uint64_t)(hi >> 6) << 26) | (uint64_t)(lo >> 6);  // 52 bits
        return ((double)bits + 0.5) * (1.0 / 4503599627370496.0);
    }
};

#ifdef __CUDACC__
// The sign rule of real_tomography, `x > 0.4 * P_i[e] ** 2 * N` (Utility.py:350, :396), evaluated in the
// reference's own order of operations, ((0.4 * (P * P)) * N), each product rounded once.  It matters: P^2 = n/N, so
// the threshold is 0.4 n up to rounding, and whenever 0.4 n is an integer (n a multiple of 5) the count x can hit it
// exactly - the comparison is then decided by the last bit of the threshold.  With a few hundred components of
// 5..30 counts each that happens several times per vector, on components whose flipped sign IS the L-infinity
// error; another association of the products (or a fused multiply-add) changes the law of that error by ~1 %
// (seen against 10^4 reference draws at d = 784).
__device__ __forceinline__ double tomo_sign_threshold(double p, double N) {
    return __dmul_rn(__dmul_rn(0.4, __dmul_rn(p, p)), N);
}

// ---------------------------------------------------------------------------
// Warp helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    int bytes = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem, bool pred) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    int bytes = pred ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(s), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

// D(8x8) += A(8x4) * B(4x8), fp64 tensor core (DMMA.8x8x4 in SASS).
// lane l holds A[l/4][l%4], B[l%4][l/4], D[l/4][2*(l%4) + {0,1}].
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

// ---------------------------------------------------------------------------
// TMA / mbarrier helpers (sm_90+ PTX; SASS: UTMALDG, SYNCS)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}"
        ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity)
        : "memory");
}
// one 2-D box: coordinates (c0 = element along the row, c1 = row)
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(map), "r"(c0), "r"(c1),
          "r"((unsigned)__cvta_generic_to_shared(bar))
        : "memory");
}

#endif  // __CUDACC__

}  // namespace sq
