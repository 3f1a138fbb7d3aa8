// modarith.cuh — word-level modular arithmetic for the two device storage widths.
//
// Storage width T is chosen per context: uint32_t when every modulus is < 2^30,
// uint64_t otherwise (moduli < 2^62).  The bounds leave room for Harvey-style lazy
// butterflies that keep values in [0, 4q).
//
// Replaces the reference's `(unsigned __int128)a*b % mod` per product
// (cuda/src/matrix/MatrixUtils.cu:541-545) with Shoup multiplication for known
// twiddles and one Barrett reduction per lazily accumulated sum.
#pragma once
#include <stdint.h>
#include <string.h>
#include <cuda_runtime.h>

// Block-level code is written against GP_TID / GP_NT / GP_SYNC so that the very same
// templates also compile as host code, where they run as a ONE-thread block (every
// strided loop then covers all its work items in order and the barrier is a no-op).
// tests/ use that build (csrc/src/emu.cpp -> libgpupoly_emu.so) to check index math,
// twiddle addressing and lazy-reduction bounds against the oracle without a GPU.
// It is a test harness, not a CPU fallback: libgpupoly.so contains no host compute path.
#ifdef __CUDA_ARCH__
#define GP_TID threadIdx.x
#define GP_NT blockDim.x
#define GP_SYNC() __syncthreads()
#else
#define GP_TID 0u
#define GP_NT 1u
#define GP_SYNC() ((void)0)
#endif
#define GP_HD __host__ __device__ __forceinline__

namespace gp {

template <typename T> struct TwPair;
template <> struct __align__(8) TwPair<uint32_t> { uint32_t w, ws; };
template <> struct __align__(16) TwPair<uint64_t> { uint64_t w, ws; };

// ---- residues as doubles ---------------------------------------------------------------------------------
// F64 is the working type of the FP64-pipe NTT for moduli below 2^51 that are stored as u64 words: the value is
// an exact integer held in a double (|v| < 2^53), signed and lazily reduced.  In GLOBAL memory an F64 pointer still
// addresses the u64 residue words (the kernels convert on load / store); in registers and shared memory it is the
// double.  Why: the 64-bit Shoup butterfly costs ~22 passes of the integer multiply pipe, the same butterfly in
// doubles is 11-13 FP64 operations on a pipe that issues 64 per clock per SM on B200 (tools/fp64_bfly_lab.cu:
// 1.6e12 butterflies/s against 0.82e12).
struct F64 {
    double v;
};
template <> struct __align__(16) TwPair<F64> { double w, ws; };  // w in [0, q) and ws = w / q
template <typename T> struct is_f64 { static constexpr bool value = false; };
template <> struct is_f64<F64> { static constexpr bool value = true; };

// Per-limb constants, one entry per CRT limb, resident in global memory.
template <typename T> struct LimbConst {
    T q;             // modulus
    T q2;            // 2q
    T ninv, ninv_s;  // n^-1 mod q and its Shoup companion
    T nw, nw_s;      // n^-1 * psi^-bitrev(1): merged last inverse-stage twiddle
    T r64, r64_s;    // 2^64 mod q (+ Shoup), folds the high word of a 128-bit sum
    T one_s;         // floor(2^wordbits / q): Shoup companion of 1 (any word -> [0,2q))
    uint64_t mu64;   // floor(2^64 / q): Barrett constant for 64-bit inputs
    uint32_t bits;   // bit width of q
    uint32_t pad;
};

GP_HD uint32_t mulhi(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}
GP_HD uint64_t mulhi(uint64_t a, uint64_t b) {
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}
GP_HD uint32_t brev32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __brev(x);
#else
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
    x = ((x >> 8) & 0x00ff00ffu) | ((x & 0x00ff00ffu) << 8);
    return (x >> 16) | (x << 16);
#endif
}

// a*w mod q in [0, 2q) for ANY word a, w < q, ws = floor(w * 2^wordbits / q).
template <typename T> GP_HD T mul_shoup_lazy(T a, T w, T ws, T q) {
    T h = mulhi(a, ws);
    return a * w - h * q;
}
template <typename T> GP_HD T csub(T a, T q) { return a >= q ? a - q : a; }

// x mod q for a 64-bit x (both widths use the 64-bit Barrett constant).
template <typename T> GP_HD T red64(uint64_t x, const LimbConst<T> &c) {
    uint64_t qh = mulhi(x, c.mu64);
    uint64_t r = x - qh * (uint64_t)c.q;  // in [0, 2q)
    return (T)(r >= (uint64_t)c.q ? r - c.q : r);
}
// (hi*2^64 + lo) mod q, u64 limbs.
GP_HD uint64_t red128(uint64_t hi, uint64_t lo, const LimbConst<uint64_t> &c) {
    uint64_t h = red64<uint64_t>(hi, c);
    uint64_t a = mul_shoup_lazy<uint64_t>(h, c.r64, c.r64_s, c.q);  // [0,2q)
    uint64_t b = red64<uint64_t>(lo, c);                             // [0,q)
    uint64_t s = a + b;                                              // < 3q < 2^64
    s = csub(s, c.q2);
    return csub(s, c.q);
}

// 1.5 * 2^52: adding and subtracting it rounds |x| < 2^51 to the nearest integer
#define GP_F64_MAGIC 6755399441055744.0
GP_HD double f64_mul(double a, double b) {
#ifdef __CUDA_ARCH__
    return __dmul_rn(a, b);  // never contracted into a neighbouring add
#else
    return a * b;
#endif
}
GP_HD double f64_add(double a, double b) {
#ifdef __CUDA_ARCH__
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
}
// x - q * rint(x / q): any |x| < 2^53 -> |result| <= q/2 (+1 ulp of the quotient estimate); exact
GP_HD double f64_renorm(double x, double q, double qinv) {
    const double k = f64_add(fma(x, qinv, GP_F64_MAGIC), -GP_F64_MAGIC);
    return fma(-k, q, x);
}
// y * w mod q for |y| <= 2^51 / 1 (so that |y * w / q| < 2^51), w in [0, q), wq = w / q rounded:
// h + l = y * w exactly (l is the rounding error of the product); b = rint(y * wq) is within 1 of the true quotient;
// c = h - b q is exact (|c| < 2^53 and an integer multiple of ulp(h)); result = c + l, |result| <= (1/2 + |y| 2^-53) q.
GP_HD double f64_mulmod(double y, double w, double wq, double q) {
    const double h = f64_mul(y, w);
    const double l = fma(y, w, -h);
    const double b = f64_add(fma(y, wq, GP_F64_MAGIC), -GP_F64_MAGIC);
    const double c = fma(-b, q, h);
    return f64_add(c, l);
}
// Integer words <-> doubles without the slow 64-bit I2F/F2I conversions: for 0 <= u < 2^52 the double 2^52 + u has u
// as its mantissa field, so one OR / AND plus one exact add converts either way.
#define GP_F64_TWO52 4503599627370496.0
#define GP_F64_TWO52_BITS 0x4330000000000000ULL
GP_HD uint64_t f64_bits(double d) {
#ifdef __CUDA_ARCH__
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u;
    memcpy(&u, &d, 8);
    return u;
#endif
}
GP_HD double f64_from_bits(uint64_t u) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double d;
    memcpy(&d, &u, 8);
    return d;
#endif
}
GP_HD double f64_from_word(uint64_t u) { return f64_add(f64_from_bits(u | GP_F64_TWO52_BITS), -GP_F64_TWO52); }  // u < 2^52
GP_HD uint64_t f64_to_word(double r) { return f64_bits(f64_add(r, GP_F64_TWO52)) & 0x000FFFFFFFFFFFFFULL; }     // 0 <= r < 2^52
// signed lazy value -> canonical residue in [0, q), still a double
GP_HD double f64_canonical_f(double x, double q, double qinv) {
    double r = f64_renorm(x, q, qinv);
    if (r < 0.0) r += q;
    if (r >= q) r -= q;
    return r;
}
// ... and as an integer word
GP_HD uint64_t f64_canonical(double x, double q, double qinv) { return f64_to_word(f64_canonical_f(x, q, qinv)); }

// Lazy accumulator for sum_k a_k*b_k mod q.
template <typename T> struct Acc;
template <> struct Acc<uint32_t> {
    uint64_t v;
    GP_HD void clear() { v = 0; }
    GP_HD void mac(uint32_t a, uint32_t b) { v += (uint64_t)a * b; }
    // products < 2^60: 16 terms fit 64 bits for q < 2^30
    static constexpr int kTermsPerFold = 16;
    GP_HD void fold(const LimbConst<uint32_t> &c) { v = red64<uint32_t>(v, c); }
    GP_HD uint32_t result(const LimbConst<uint32_t> &c) const { return red64<uint32_t>(v, c); }
};
template <> struct Acc<uint64_t> {
    uint64_t lo, hi;
    GP_HD void clear() { lo = 0; hi = 0; }
    GP_HD void mac(uint64_t a, uint64_t b) {
        uint64_t pl = a * b, ph = mulhi(a, b);
        lo += pl;
        hi += ph + (lo < pl);
    }
    // products < 2^124: 16 terms fit 128 bits for q < 2^62
    static constexpr int kTermsPerFold = 16;
    GP_HD void fold(const LimbConst<uint64_t> &c) { lo = red128(hi, lo, c); hi = 0; }
    GP_HD uint64_t result(const LimbConst<uint64_t> &c) const { return red128(hi, lo, c); }
};

template <typename T> GP_HD T mul_mod(T a, T b, const LimbConst<T> &c) {
    Acc<T> acc;
    acc.clear();
    acc.mac(a, b);
    return acc.result(c);
}
template <typename T> GP_HD T add_mod(T a, T b, T q) { return csub<T>(a + b, q); }
template <typename T> GP_HD T sub_mod(T a, T b, T q) { return a >= b ? a - b : a + q - b; }

}  // namespace gp
