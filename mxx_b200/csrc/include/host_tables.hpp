// host_tables.hpp — host-side number theory and NTT table construction, shared by
// Runtime.cu (uploads the tables) and the host emulation harness emu.cu (tests only).
// Root definitions follow the reference so the EVAL layout at the boundary is identical:
// least primitive root g (cuda/src/Runtime.cu:121-165), psi = g^((p-1)/2n) (:167-190),
// n^-1 (:282-319).
#pragma once
#include <string.h>
#include <stdint.h>

#include <stdexcept>
#include <vector>

#include "modarith.cuh"

namespace gp {
namespace host {
typedef unsigned __int128 u128;

inline uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)(((u128)a * b) % q); }
inline uint64_t powmod(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q, c = b % q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, c, q);
        c = mulmod(c, c, q);
    }
    return r;
}
inline bool invmod(uint64_t a, uint64_t q, uint64_t &out) {
    __int128 t = 0, nt = 1, r = q, nr = a % q;
    while (nr != 0) {
        __int128 k = r / nr, tmp = t - k * nt;
        t = nt; nt = tmp;
        tmp = r - k * nr; r = nr; nr = tmp;
    }
    if (r != 1) return false;
    if (t < 0) t += q;
    out = (uint64_t)t;
    return true;
}
// distinct prime factors of m (trial division; 2 stripped first, odd candidates after)
inline std::vector<uint64_t> prime_factors(uint64_t m) {
    std::vector<uint64_t> f;
    if ((m & 1) == 0) {
        f.push_back(2);
        while ((m & 1) == 0) m >>= 1;
    }
    for (uint64_t d = 3; d * d <= m; d += 2) {
        if (m % d) continue;
        f.push_back(d);
        while (m % d == 0) m /= d;
    }
    if (m > 1) f.push_back(m);
    return f;
}
// least primitive root — same value the reference finds (Runtime.cu:121-165)
inline uint64_t least_primitive_root(uint64_t p) {
    const uint64_t phi = p - 1;
    const std::vector<uint64_t> fac = prime_factors(phi);
    for (uint64_t g = 2; g < p; ++g) {
        bool ok = true;
        for (uint64_t f : fac)
            if (powmod(g, phi / f, p) == 1) { ok = false; break; }
        if (ok) return g;
    }
    throw std::runtime_error("failed to find primitive root");
}
inline uint64_t primitive_2n_root(uint64_t p, uint64_t n) {  // Runtime.cu:167-190
    const uint64_t order = 2 * n;
    if (p % order != 1) throw std::runtime_error("modulus is not congruent to 1 mod 2N");
    const uint64_t root = powmod(least_primitive_root(p), (p - 1) / order, p);
    if (powmod(root, n, p) != p - 1) throw std::runtime_error("computed root is not a primitive 2N-th root");
    return root;
}
inline uint32_t bitrev(uint32_t x, int bits) { return bits == 0 ? 0 : (gp::brev32(x) >> (32 - bits)); }

template <typename T> inline T shoup(uint64_t w, uint64_t q) { return (T)((((u128)w) << (sizeof(T) * 8)) / q); }


template <typename T> struct Tables {
    std::vector<TwPair<T>> fwd, inv;  // [L][n]: psi^bitrev(i), psi^-bitrev(i) with Shoup companions
    std::vector<LimbConst<T>> limb;   // [L]
};

template <typename T>
Tables<T> build_tables(int logN, const std::vector<uint64_t> &moduli, const std::vector<uint64_t> &psis) {
    const size_t n = (size_t)1 << logN, L = moduli.size();
    Tables<T> t;
    t.fwd.resize(L * n);
    t.inv.resize(L * n);
    t.limb.resize(L);
    for (size_t l = 0; l < L; ++l) {
        const uint64_t q = moduli[l], psi = psis[l];
        uint64_t ipsi = 0, ninv = 0;
        if (!invmod(psi, q, ipsi) || !invmod(n % q, q, ninv)) throw std::runtime_error("failed to invert NTT constants");
        uint64_t w = 1, iw = 1;
        for (size_t i = 0; i < n; ++i) {
            const size_t r = bitrev((uint32_t)i, logN);
            t.fwd[l * n + r] = {(T)w, shoup<T>(w, q)};
            t.inv[l * n + r] = {(T)iw, shoup<T>(iw, q)};
            w = mulmod(w, psi, q);
            iw = mulmod(iw, ipsi, q);
        }
        LimbConst<T> &c = t.limb[l];
        c.q = (T)q;
        c.q2 = (T)(2 * q);
        c.ninv = (T)ninv;
        c.ninv_s = shoup<T>(ninv, q);
        const uint64_t nw = mulmod(ninv, n > 1 ? (uint64_t)t.inv[l * n + 1].w : 1, q);
        c.nw = (T)nw;
        c.nw_s = shoup<T>(nw, q);
        const uint64_t r64 = (uint64_t)((((u128)1) << 64) % q);
        c.r64 = (T)r64;
        c.r64_s = shoup<T>(r64, q);
        c.one_s = (T)((((u128)1) << (sizeof(T) * 8)) / q);
        c.mu64 = (uint64_t)((((u128)1) << 64) / q);
        c.bits = 64 - (uint32_t)__builtin_clzll(q);
        c.pad = 0;
    }
    return t;
}

// The same tables for the FP64-pipe kernels (modarith.cuh F64): twiddles as doubles with their quotients w/q;
// LimbConst<F64> reuses the generic field names — q2 carries 1/q (the F64 butterflies need no 2q), one_s too.
inline Tables<F64> f64_tables(const Tables<uint64_t> &t) {
    Tables<F64> r;
    const size_t L = t.limb.size(), per = L ? t.fwd.size() / L : 0;
    r.fwd.resize(t.fwd.size());
    r.inv.resize(t.inv.size());
    r.limb.resize(L);
    for (size_t l = 0; l < L; ++l) {
        const double q = (double)t.limb[l].q;
        for (size_t i = 0; i < per; ++i) {
            r.fwd[l * per + i] = {(double)t.fwd[l * per + i].w, (double)t.fwd[l * per + i].w / q};
            r.inv[l * per + i] = {(double)t.inv[l * per + i].w, (double)t.inv[l * per + i].w / q};
        }
        LimbConst<F64> &c = r.limb[l];
        memset(&c, 0, sizeof(c));
        c.q.v = q;
        c.q2.v = 1.0 / q;
        c.one_s.v = 1.0 / q;
        c.ninv.v = (double)t.limb[l].ninv;
        c.ninv_s.v = (double)t.limb[l].ninv / q;
        c.nw.v = (double)t.limb[l].nw;
        c.nw_s.v = (double)t.limb[l].nw / q;
        c.mu64 = t.limb[l].mu64;
        c.bits = t.limb[l].bits;
    }
    return r;
}
// FP64-pipe NTT: u64-word contexts whose moduli all stay below 2^51 (4q < 2^53 keeps every intermediate exact)
inline bool f64_ok(uint64_t qmax, bool wide) { return wide && qmax < ((uint64_t)1 << 51); }

// products that fit one lazy accumulator (2^64 for u32 words, 2^128 for u64) on top of a folded residue
// FP64 product (arith.cuh matmul_f64_thread): reduced terms of magnitude <= q on top of a renormalised
// accumulator (<= q/2, plus the rounding slack of the quotient estimate) must stay below 2^53
inline uint32_t f64_mac_fold_terms(uint64_t qmax) {
    const long double room = 9007199254740992.0L / (long double)qmax - 0.51L;
    uint64_t t = room < 1.0L ? 1 : (uint64_t)room;
    return (uint32_t)(t > 4096 ? 4096 : t);
}
inline uint32_t mac_fold_terms(uint64_t qmax, bool wide) {
    const u128 q2 = (u128)(qmax - 1) * (qmax - 1);
    const u128 room = wide ? (~(u128)0 - qmax) : ((u128)~0ull - qmax);
    u128 terms = room / q2;
    if (terms < 1) terms = 1;
    if (terms > (1u << 16)) terms = 1u << 16;
    return (uint32_t)terms;
}
// storage width rule: u32 words when every modulus is below 2^30 (lazy values stay below 4q < 2^32)
inline bool needs_wide(uint64_t qmax) { return qmax >= (1ull << 30); }
// forward transforms need no conditional subtraction when 36*q fits the word (ntt_fast.cuh)
inline bool nocsub_ok(uint64_t qmax, bool wide) { return wide ? qmax < ((uint64_t)1 << 58) : qmax * 36 < ((uint64_t)1 << 32); }

}  // namespace host
}  // namespace gp
