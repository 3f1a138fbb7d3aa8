// ntt_fast.cuh — compile-time-planned negacyclic NTT / INTT for the ring sizes that matter
// (one (poly, limb) block of 2^LOGB words resident in shared memory, LOGB = 12..15).
//
// Same mathematics and tables as ntt.cuh (merged-psi Cooley-Tukey forward, Gentleman-Sande
// inverse, Shoup twiddles, device EVAL order bit-reversed); what changes is the schedule:
//
//   forward  pass 1   global -> radix-2^R1 in registers -> smem      (coalesced 4/8-byte loads)
//            pass 2.. smem -> radix-16 -> smem                       (K passes)
//            last     smem (128-bit) -> radix-4 on 4 consecutive words -> global (128-bit)
//   inverse  the mirror image: 128-bit global loads -> radix-4 -> smem ... -> radix-2^R1 -> global
//
//   * every stride is a template constant, so shared-memory addresses are base + immediate:
//     no per-access index arithmetic (the v1 kernel spent more issue slots on LEA/SHF than on
//     butterflies, profiles/r01_ncu_v1_summary.md);
//   * padding of one 16-byte vector per 64 words keeps all passes bank-conflict free AND keeps
//     128-bit alignment for the vector pass; u64 residues live as two u32 planes (lo / hi) so
//     the u64 path has exactly the u32 access pattern;
//   * 256 threads and <= 80 registers: three CTAs per SM for a 64 KB block, so one CTA's
//     global loads overlap the other two CTAs' butterflies;
//   * NOCSUB: when 34*q < 2^wordbits (24..26-bit and 51-bit moduli) the forward transform needs
//     no conditional subtraction at all (values grow by 2q per stage and the Shoup multiply
//     accepts any word); one Shoup-by-one reduction at the final store brings them to [0, q).
#pragma once
#include <string.h>

#include "modarith.cuh"
#include "ntt.cuh"

namespace gp {

// ---- shared-memory view ---------------------------------------------------------------
GP_HD uint32_t padw(uint32_t a) { return a + ((a >> 6) << 2); }
template <int LOGB> struct FastGeom {
    static constexpr uint32_t B = 1u << LOGB;
    static constexpr uint32_t kPlane = B + ((B >> 6) << 2);  // padded words per plane
    // stage split: LOGB = R1 + 4*K + 2
    static constexpr int kRem = LOGB - 2;
    static constexpr int kR1 = (kRem % 4 == 0) ? 4 : ((kRem % 4 == 1 && kRem >= 5) ? 5 : kRem % 4);
    static constexpr int kK = (kRem - kR1) / 4;
    static_assert(kR1 >= 1 && kR1 <= 5 && kK >= 0 && LOGB - kR1 >= 6, "unsupported block size");
};
template <typename T, int LOGB> constexpr size_t fast_smem_bytes() {
    return (size_t)FastGeom<LOGB>::kPlane * sizeof(uint32_t) * (sizeof(T) / 4);
}

template <typename T, int LOGB> struct SmemView;
template <int LOGB> struct SmemView<uint32_t, LOGB> {
    uint32_t *p;
    GP_HD explicit SmemView(void *base) : p(reinterpret_cast<uint32_t *>(base)) {}
    GP_HD uint32_t ld(uint32_t i) const { return p[i]; }
    GP_HD void st(uint32_t i, uint32_t v) const { p[i] = v; }
    GP_HD void ld4(uint32_t i, uint32_t (&x)[4]) const {
        const uint4 v = *reinterpret_cast<const uint4 *>(p + i);
        x[0] = v.x, x[1] = v.y, x[2] = v.z, x[3] = v.w;
    }
    GP_HD void st4(uint32_t i, const uint32_t (&x)[4]) const { *reinterpret_cast<uint4 *>(p + i) = make_uint4(x[0], x[1], x[2], x[3]); }
};
template <int LOGB> struct SmemView<uint64_t, LOGB> {
    uint32_t *lo, *hi;
    GP_HD explicit SmemView(void *base)
        : lo(reinterpret_cast<uint32_t *>(base)), hi(reinterpret_cast<uint32_t *>(base) + FastGeom<LOGB>::kPlane) {}
    GP_HD uint64_t ld(uint32_t i) const { return ((uint64_t)hi[i] << 32) | lo[i]; }
    GP_HD void st(uint32_t i, uint64_t v) const { lo[i] = (uint32_t)v, hi[i] = (uint32_t)(v >> 32); }
    GP_HD void ld4(uint32_t i, uint64_t (&x)[4]) const {
        const uint4 a = *reinterpret_cast<const uint4 *>(lo + i), b = *reinterpret_cast<const uint4 *>(hi + i);
        x[0] = ((uint64_t)b.x << 32) | a.x, x[1] = ((uint64_t)b.y << 32) | a.y;
        x[2] = ((uint64_t)b.z << 32) | a.z, x[3] = ((uint64_t)b.w << 32) | a.w;
    }
    GP_HD void st4(uint32_t i, const uint64_t (&x)[4]) const {
        *reinterpret_cast<uint4 *>(lo + i) = make_uint4((uint32_t)x[0], (uint32_t)x[1], (uint32_t)x[2], (uint32_t)x[3]);
        *reinterpret_cast<uint4 *>(hi + i) = make_uint4((uint32_t)(x[0] >> 32), (uint32_t)(x[1] >> 32), (uint32_t)(x[2] >> 32), (uint32_t)(x[3] >> 32));
    }
};
// doubles travel through shared memory as their bit patterns in the u64 two-plane layout (f64_bits, modarith.cuh)
template <int LOGB> struct SmemView<F64, LOGB> {
    SmemView<uint64_t, LOGB> raw;
    GP_HD explicit SmemView(void *base) : raw(base) {}
    GP_HD F64 ld(uint32_t i) const { return F64{f64_from_bits(raw.ld(i))}; }
    GP_HD void st(uint32_t i, F64 v) const { raw.st(i, f64_bits(v.v)); }
    GP_HD void ld4(uint32_t i, F64 (&x)[4]) const {
        uint64_t u[4];
        raw.ld4(i, u);
#pragma unroll
        for (int j = 0; j < 4; j++) x[j].v = f64_from_bits(u[j]);
    }
    GP_HD void st4(uint32_t i, const F64 (&x)[4]) const {
        uint64_t u[4];
#pragma unroll
        for (int j = 0; j < 4; j++) u[j] = f64_bits(x[j].v);
        raw.st4(i, u);
    }
};

// launch shape per block size (needed by the passes: loop trip counts are compile-time constants)
template <typename T, int LOGB> struct FastCfg {
    static constexpr bool kSmall = fast_smem_bytes<T, LOGB>() <= 72 * 1024;  // three blocks per SM
#ifndef GP_FAST_THREADS
#define GP_FAST_THREADS 256
#endif
#ifndef GP_FAST_MINB
#define GP_FAST_MINB 3
#endif
    static constexpr int kThreads = kSmall ? GP_FAST_THREADS : 512;
#ifndef GP_FAST_MINB_WIDE
#define GP_FAST_MINB_WIDE 2
#endif
#ifndef GP_FAST_MINB_F64
#define GP_FAST_MINB_F64 3  // the FP64 butterflies fit 80 registers: a third resident CTA gives +10 % (tools/ntt_lab.cu)
#endif
    static constexpr int kMinBlocks = kSmall ? (sizeof(T) == 4 ? GP_FAST_MINB : (is_f64<T>::value ? GP_FAST_MINB_F64 : GP_FAST_MINB_WIDE)) : 1;
};

// ---- asynchronous global -> shared staging (LDGSTS) ---------------------------------------------
// A thread stages exactly the words it will consume itself, so cp.async.wait_group is the only
// synchronisation needed; all of a CTA's global loads are in flight within its first few hundred
// cycles and no register is tied up while they travel.
#ifndef GP_NTT_ASYNC
#define GP_NTT_ASYNC 1
#endif
#ifndef GP_NTT_ASYNC_FWD
#define GP_NTT_ASYNC_FWD 0
#endif
#ifndef GP_NTT_ASYNC_INV
#define GP_NTT_ASYNC_INV 1
#endif
#ifndef GP_FLAST_U
#define GP_FLAST_U 2
#endif
#if defined(__CUDA_ARCH__) && GP_NTT_ASYNC
#define GP_ASYNC_STAGE 1
__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait(int pending) {
    switch (pending) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
        case 4: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
        case 5: asm volatile("cp.async.wait_group 5;" ::: "memory"); break;
        case 6: asm volatile("cp.async.wait_group 6;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 7;" ::: "memory"); break;
    }
}
// warm the L2 with the block a CTA started one residency period later will read
__device__ __forceinline__ void prefetch_l2(const void *p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
#else
#define GP_ASYNC_STAGE 0
#endif

// 4 consecutive words from / to global memory (16 bytes for u32, 32 bytes for u64)
GP_HD void gld4(const uint32_t *p, uint32_t (&x)[4]) {
    const uint4 v = *reinterpret_cast<const uint4 *>(p);
    x[0] = v.x, x[1] = v.y, x[2] = v.z, x[3] = v.w;
}
GP_HD void gld4(const uint64_t *p, uint64_t (&x)[4]) {
    const ulonglong2 a = *reinterpret_cast<const ulonglong2 *>(p), b = *reinterpret_cast<const ulonglong2 *>(p + 2);
    x[0] = a.x, x[1] = a.y, x[2] = b.x, x[3] = b.y;
}
GP_HD void gst4(uint32_t *p, const uint32_t (&x)[4]) { *reinterpret_cast<uint4 *>(p) = make_uint4(x[0], x[1], x[2], x[3]); }
GP_HD void gst4(uint64_t *p, const uint64_t (&x)[4]) {
    *reinterpret_cast<ulonglong2 *>(p) = make_ulonglong2(x[0], x[1]);
    *reinterpret_cast<ulonglong2 *>(p + 2) = make_ulonglong2(x[2], x[3]);
}

// F64: the words in global memory are integers (canonical residues on the way in)
GP_HD void gld4(const F64 *p, F64 (&x)[4]) {
    uint64_t u[4];
    gld4(reinterpret_cast<const uint64_t *>(p), u);
#pragma unroll
    for (int j = 0; j < 4; j++) x[j].v = f64_from_word(u[j]);
}
GP_HD void gst4(F64 *p, const F64 (&x)[4]) {  // x already canonical, see f_last
    uint64_t u[4];
#pragma unroll
    for (int j = 0; j < 4; j++) u[j] = f64_to_word(x[j].v);
    gst4(reinterpret_cast<uint64_t *>(p), u);
}

// unsigned min trick: x - m wraps to a huge value when x < m
template <typename T> GP_HD T csub_min(T x, T m) {
    const T y = x - m;
    return y < x ? y : x;
}
// any word -> [0, 2q): Shoup multiplication by one, one_s = floor(2^wordbits / q)
template <typename T> GP_HD T reduce_lazy(T x, T one_s, T q) { return x - mulhi(x, one_s) * q; }


// COUNT consecutive twiddle pairs starting at an index that is a multiple of COUNT: 128-bit loads
#ifndef GP_TW_VEC
#define GP_TW_VEC 1
#endif
template <int COUNT> GP_HD void load_tw(const TwPair<uint32_t> *__restrict__ p, TwPair<uint32_t> (&w)[COUNT]) {
    if constexpr (COUNT >= 2 && GP_TW_VEC) {
#pragma unroll
        for (int i = 0; i < COUNT / 2; i++) {
            const uint4 v = reinterpret_cast<const uint4 *>(p)[i];
            w[2 * i].w = v.x, w[2 * i].ws = v.y, w[2 * i + 1].w = v.z, w[2 * i + 1].ws = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < COUNT; i++) w[i] = p[i];
    }
}
template <int COUNT> GP_HD void load_tw(const TwPair<uint64_t> *__restrict__ p, TwPair<uint64_t> (&w)[COUNT]) {
#pragma unroll
    for (int i = 0; i < COUNT; i++) w[i] = p[i];
}
template <int COUNT> GP_HD void load_tw(const TwPair<F64> *__restrict__ p, TwPair<F64> (&w)[COUNT]) {
#pragma unroll
    for (int i = 0; i < COUNT; i++) w[i] = p[i];
}

// ---- butterfly networks ---------------------------------------------------------------------
// Forward (Cooley-Tukey): stage s pairs j and j + (R >> (s+1)); twiddle index (twbase << s) + u.
// NOCSUB: inputs < B give outputs < B + 2q*LOGR; otherwise values stay in [0, 4q).
template <typename T, int LOGR, int S, bool NOCSUB>
GP_HD void fwd_stage(T (&x)[1 << LOGR], const TwPair<T> *__restrict__ tw, uint32_t twbase, T q, T q2) {
    constexpr int R = 1 << LOGR, half = R >> (S + 1), cnt = 1 << S;
    TwPair<T> w[cnt];
    load_tw<cnt>(tw + ((size_t)twbase << S), w);
#pragma unroll
    for (int u = 0; u < cnt; u++) {
#pragma unroll
        for (int v = 0; v < half; v++) {
            const int j0 = u * 2 * half + v, j1 = j0 + half;
            if constexpr (is_f64<T>::value) {
                // F64 (q2 carries 1/q): the multiplied leg is renormalised to |y| <= q/2, the other leg grows by
                // < 0.63 q per stage and is renormalised at the entry of every net (fwd_net): < 4q < 2^53 throughout
                const double t = f64_mulmod(f64_renorm(x[j1].v, q.v, q2.v), w[u].w, w[u].ws, q.v);
                const double X = x[j0].v;
                x[j0].v = f64_add(X, t);
                x[j1].v = f64_add(X, -t);
            } else {
                const T X = NOCSUB ? x[j0] : csub_min<T>(x[j0], q2);
                const T t = mul_shoup_lazy<T>(x[j1], w[u].w, w[u].ws, q);
                x[j0] = X + t;
                x[j1] = X - t + q2;
            }
        }
    }
    if constexpr (S + 1 < LOGR) fwd_stage<T, LOGR, S + 1, NOCSUB>(x, tw, twbase, q, q2);
}
template <typename T, int LOGR, bool NOCSUB>
GP_HD void fwd_net(T (&x)[1 << LOGR], const TwPair<T> *__restrict__ tw, uint32_t twbase, T q, T q2) {
    if constexpr (is_f64<T>::value) {
        static_assert(LOGR <= 5, "an F64 net may grow the unmultiplied leg for at most 5 stages");
        // the first stage's multiplied legs (upper half) are renormalised by their butterflies
#pragma unroll
        for (int j = 0; j < (1 << (LOGR - 1)); j++) x[j].v = f64_renorm(x[j].v, q.v, q2.v);
    }
    fwd_stage<T, LOGR, 0, NOCSUB>(x, tw, twbase, q, q2);
}
// Inverse (Gentleman-Sande): stage s pairs j and j + (1 << s); inputs and outputs in [0, 2q).
// FINAL merges n^-1 (and n^-1 * last twiddle) into the last stage.
template <typename T, int LOGR, int S, bool FINAL>
GP_HD void inv_stage(T (&x)[1 << LOGR], const TwPair<T> *__restrict__ tw, uint32_t twbase, const LimbConst<T> &c) {
    constexpr int R = 1 << LOGR, groups = R >> (S + 1);
    const T q = c.q, q2 = c.q2;
    if constexpr (FINAL && S == LOGR - 1) {
#pragma unroll
        for (int v = 0; v < (1 << S); v++) {
            const int j0 = v, j1 = j0 + (1 << S);
            const T X = x[j0], Y = x[j1];
            if constexpr (is_f64<T>::value) {
                x[j0].v = f64_mulmod(f64_renorm(f64_add(X.v, Y.v), q.v, q2.v), c.ninv.v, c.ninv_s.v, q.v);
                x[j1].v = f64_mulmod(f64_renorm(f64_add(X.v, -Y.v), q.v, q2.v), c.nw.v, c.nw_s.v, q.v);
            } else {
                x[j0] = mul_shoup_lazy<T>(X + Y, c.ninv, c.ninv_s, q);
                x[j1] = mul_shoup_lazy<T>(X - Y + q2, c.nw, c.nw_s, q);
            }
        }
    } else {
        TwPair<T> w[groups];
        load_tw<groups>(tw + (size_t)twbase * groups, w);
#pragma unroll
        for (int u = 0; u < groups; u++) {
#pragma unroll
            for (int v = 0; v < (1 << S); v++) {
                const int j0 = u * (2 << S) + v, j1 = j0 + (1 << S);
                const T X = x[j0], Y = x[j1];
                if constexpr (is_f64<T>::value) {
                    // the sum leg doubles per stage: renormalised on odd stages and at the end of the net, so that
                    // nets hand over |x| <= 0.63 q and nothing exceeds 4q < 2^53 in between
                    const double sum = f64_add(X.v, Y.v);
                    x[j0].v = ((S & 1) || S == LOGR - 1) ? f64_renorm(sum, q.v, q2.v) : sum;
                    x[j1].v = f64_mulmod(f64_renorm(f64_add(X.v, -Y.v), q.v, q2.v), w[u].w, w[u].ws, q.v);
                } else {
                    x[j0] = csub_min<T>(X + Y, q2);
                    x[j1] = mul_shoup_lazy<T>(X - Y + q2, w[u].w, w[u].ws, q);
                }
            }
        }
    }
    if constexpr (S + 1 < LOGR) inv_stage<T, LOGR, S + 1, FINAL>(x, tw, twbase, c);
}
template <typename T, int LOGR, bool FINAL>
GP_HD void inv_net(T (&x)[1 << LOGR], const TwPair<T> *__restrict__ tw, uint32_t twbase, const LimbConst<T> &c) {
    inv_stage<T, LOGR, 0, FINAL>(x, tw, twbase, c);
}

// ---- radix-4 networks on 4 consecutive words with preloaded twiddles -------------------------------
// The three twiddles of such a group are tw[tb], tw[2 tb], tw[2 tb + 1]: the last two are adjacent, so
// they come in one 128-bit load (u32) and all of a thread's twiddle loads can be issued ahead of the data.
template <typename T> struct Tw3 { TwPair<T> a, b0, b1; };
GP_HD Tw3<uint32_t> load_tw3(const TwPair<uint32_t> *__restrict__ tw, uint32_t tb) {
    Tw3<uint32_t> r;
    r.a = tw[tb];
    const uint4 v = *reinterpret_cast<const uint4 *>(tw + 2 * (size_t)tb);
    r.b0.w = v.x, r.b0.ws = v.y, r.b1.w = v.z, r.b1.ws = v.w;
    return r;
}
GP_HD Tw3<uint64_t> load_tw3(const TwPair<uint64_t> *__restrict__ tw, uint32_t tb) {
    Tw3<uint64_t> r;
    r.a = tw[tb];
    r.b0 = tw[2 * (size_t)tb];
    r.b1 = tw[2 * (size_t)tb + 1];
    return r;
}
GP_HD Tw3<F64> load_tw3(const TwPair<F64> *__restrict__ tw, uint32_t tb) {
    Tw3<F64> r;
    r.a = tw[tb];
    r.b0 = tw[2 * (size_t)tb];
    r.b1 = tw[2 * (size_t)tb + 1];
    return r;
}
template <typename T, bool NOCSUB> GP_HD void fwd_bfly(T &x0, T &x1, const TwPair<T> &w, T q, T q2) {
    if constexpr (is_f64<T>::value) {
        const double t = f64_mulmod(f64_renorm(x1.v, q.v, q2.v), w.w, w.ws, q.v);
        const double X = x0.v;
        x0.v = f64_add(X, t);
        x1.v = f64_add(X, -t);
    } else {
        const T X = NOCSUB ? x0 : csub_min<T>(x0, q2);
        const T t = mul_shoup_lazy<T>(x1, w.w, w.ws, q);
        x0 = X + t;
        x1 = X - t + q2;
    }
}
template <typename T, bool NOCSUB> GP_HD void fwd_net4(T (&x)[4], const Tw3<T> &t, T q, T q2) {
    if constexpr (is_f64<T>::value) {
#pragma unroll
        for (int j = 0; j < 2; j++) x[j].v = f64_renorm(x[j].v, q.v, q2.v);  // x[2], x[3]: by their butterflies
    }
    fwd_bfly<T, NOCSUB>(x[0], x[2], t.a, q, q2);
    fwd_bfly<T, NOCSUB>(x[1], x[3], t.a, q, q2);
    fwd_bfly<T, NOCSUB>(x[0], x[1], t.b0, q, q2);
    fwd_bfly<T, NOCSUB>(x[2], x[3], t.b1, q, q2);
}
template <typename T, bool RENORM = true> GP_HD void inv_bfly(T &x0, T &x1, const TwPair<T> &w, T q, T q2) {
    const T X = x0, Y = x1;
    if constexpr (is_f64<T>::value) {
        const double sum = f64_add(X.v, Y.v);
        x0.v = RENORM ? f64_renorm(sum, q.v, q2.v) : sum;
        x1.v = f64_mulmod(f64_renorm(f64_add(X.v, -Y.v), q.v, q2.v), w.w, w.ws, q.v);
    } else {
        x0 = csub_min<T>(X + Y, q2);
        x1 = mul_shoup_lazy<T>(X - Y + q2, w.w, w.ws, q);
    }
}
template <typename T> GP_HD void inv_net4(T (&x)[4], const Tw3<T> &t, T q, T q2) {
    inv_bfly<T, false>(x[0], x[1], t.b0, q, q2);
    inv_bfly<T, false>(x[2], x[3], t.b1, q, q2);
    inv_bfly<T, true>(x[0], x[2], t.a, q, q2);
    inv_bfly<T, true>(x[1], x[3], t.a, q, q2);
}

// ---- passes over one block of B = 2^LOGB words -------------------------------------------------
// Conventions: `nblk` = number of B-blocks in the whole transform (n / B), `sb` = this block's
// index; forward twiddle index of a span-block = (#span-blocks in the transform) + (its index).

// padded offset of j*stride for a stride that is a multiple of 64 words or smaller than 64
template <int LSTRIDE> GP_HD constexpr uint32_t joff(int j) {
    return LSTRIDE >= 6 ? padw((uint32_t)j << LSTRIDE) : ((uint32_t)j << LSTRIDE);
}

// forward pass 1: source functor -> radix-2^R1 -> smem
template <typename T, int LOGB, bool NOCSUB, class Src>
GP_HD void f_first(const Src &src, const SmemView<T, LOGB> &sm, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb,
                   T q, T q2) {
    constexpr int R1 = FastGeom<LOGB>::kR1, R = 1 << R1, LST = LOGB - R1;
    constexpr uint32_t stride = 1u << LST;
#if GP_ASYNC_STAGE && GP_NTT_ASYNC_FWD
    if constexpr (sizeof(T) == 4) {
        constexpr int NT = FastCfg<T, LOGB>::kThreads, ITER = (int)(stride / NT);
        static_assert(stride % NT == 0 && ITER >= 1 && ITER <= 8, "pass-1 staging plan");
        const T *raw = src.p;
#pragma unroll
        for (int it = 0; it < ITER; it++) {
            const uint32_t g = GP_TID + it * NT, base = padw(g);
#pragma unroll
            for (int j = 0; j < R; j++) cp_async4(sm.p + base + joff<LST>(j), raw + g + j * stride);
            cp_async_commit();
        }
#pragma unroll
        for (int it = 0; it < ITER; it++) {
            cp_async_wait(ITER - 1 - it);
            const uint32_t g = GP_TID + it * NT, base = padw(g);
            T x[R];
#pragma unroll
            for (int j = 0; j < R; j++) x[j] = src.xform(sm.ld(base + joff<LST>(j)));
            fwd_net<T, R1, NOCSUB>(x, tw, nblk + sb, q, q2);
#pragma unroll
            for (int j = 0; j < R; j++) sm.st(base + joff<LST>(j), x[j]);
        }
    } else
#endif
    {
#pragma unroll 2
        for (uint32_t g = GP_TID; g < stride; g += GP_NT) {
            T x[R];
#pragma unroll
            for (int j = 0; j < R; j++) x[j] = src(g + j * stride);
            fwd_net<T, R1, NOCSUB>(x, tw, nblk + sb, q, q2);
            const uint32_t base = padw(g);
#pragma unroll
            for (int j = 0; j < R; j++) sm.st(base + joff<LST>(j), x[j]);
        }
    }
}
// forward smem pass: radix 16 at stride 2^(LSPAN-4)
template <typename T, int LOGB, int LSPAN, bool NOCSUB>
GP_HD void f_mid(const SmemView<T, LOGB> &sm, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb, T q, T q2) {
    constexpr int LST = LSPAN - 4;
    constexpr uint32_t stride = 1u << LST, groups = 1u << (LOGB - 4);
    constexpr uint32_t per = LOGB - LSPAN;  // log2(span-blocks per B-block)
#pragma unroll 2
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        const uint32_t b = g >> LST, e = g & (stride - 1);
        const uint32_t base = padw((b << LSPAN) + e);
        T x[16];
#pragma unroll
        for (int j = 0; j < 16; j++) x[j] = sm.ld(base + joff<LST>(j));
        fwd_net<T, 4, NOCSUB>(x, tw, ((nblk + sb) << per) + b, q, q2);
#pragma unroll
        for (int j = 0; j < 16; j++) sm.st(base + joff<LST>(j), x[j]);
    }
}
// forward last pass: 4 consecutive words (one 128-bit smem vector) -> radix 4 -> [0,q) -> global
template <typename T, int LOGB, bool NOCSUB, int STAGES>
GP_HD void f_last(const SmemView<T, LOGB> &sm, T *__restrict__ dst, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb,
                  const LimbConst<T> &c) {
    constexpr uint32_t groups = 1u << (LOGB - 2);
    constexpr int U = sizeof(T) == 4 ? GP_FLAST_U : 2;  // groups whose twiddles are fetched ahead of the data
    const T q = c.q, q2 = c.q2;
    const uint32_t tb0 = (nblk + sb) << (LOGB - 2);
#pragma unroll 1
    for (uint32_t g0 = GP_TID; g0 < groups; g0 += U * GP_NT) {
        Tw3<T> t[U];
#pragma unroll
        for (int u = 0; u < U; u++) t[u] = load_tw3(tw, tb0 + g0 + u * GP_NT);
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t g = g0 + u * GP_NT;
            T x[4];
            sm.ld4(4 * g + ((g >> 4) << 2), x);
            fwd_net4<T, NOCSUB>(x, t[u], q, q2);
            if constexpr (is_f64<T>::value) {
#pragma unroll
                for (int j = 0; j < 4; j++) x[j].v = f64_canonical_f(x[j].v, q.v, q2.v);
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    T v = x[j];
                    if (NOCSUB) {
                        // < (1 + 2*STAGES) q: one Shoup-by-one brings it to [0, 2q)
                        v = reduce_lazy<T>(v, c.one_s, q);
                    } else {
                        v = csub_min<T>(v, q2);
                    }
                    x[j] = csub_min<T>(v, q);
                }
            }
            gst4(dst + 4 * g, x);
        }
    }
}

// inverse first pass: global (128-bit) -> radix 4 on 4 consecutive words -> smem
template <typename T, int LOGB>
GP_HD void i_first(const T *__restrict__ src, const SmemView<T, LOGB> &sm, const TwPair<T> *__restrict__ tw, uint32_t nblk,
                   uint32_t sb, const LimbConst<T> &c) {
    constexpr uint32_t groups = 1u << (LOGB - 2);
#if GP_ASYNC_STAGE && GP_NTT_ASYNC_INV
    if constexpr (sizeof(T) == 4) {
        constexpr int NT = FastCfg<T, LOGB>::kThreads, ITER = (int)(groups / NT), BATCH = 4, NB = ITER / BATCH;
        static_assert(groups % NT == 0 && ITER % BATCH == 0 && NB >= 1 && NB <= 8, "inverse pass-1 staging plan");
#pragma unroll
        for (int b = 0; b < NB; b++) {
#pragma unroll
            for (int i = 0; i < BATCH; i++) {
                const uint32_t g = GP_TID + (b * BATCH + i) * NT;
                cp_async16(sm.p + 4 * g + ((g >> 4) << 2), src + 4 * g);
            }
            cp_async_commit();
        }
#pragma unroll
        for (int b = 0; b < NB; b++) {
            Tw3<T> t[BATCH];
#pragma unroll
            for (int i = 0; i < BATCH; i++) t[i] = load_tw3(tw, ((nblk + sb) << (LOGB - 2)) + GP_TID + (b * BATCH + i) * NT);
            cp_async_wait(NB - 1 - b);
#pragma unroll
            for (int i = 0; i < BATCH; i++) {
                const uint32_t g = GP_TID + (b * BATCH + i) * NT, at = 4 * g + ((g >> 4) << 2);
                T x[4];
                sm.ld4(at, x);
                inv_net4<T>(x, t[i], c.q, c.q2);
                sm.st4(at, x);
            }
        }
    } else
#endif
    {
        constexpr int U = 2;
#pragma unroll 1
        for (uint32_t g0 = GP_TID; g0 < groups; g0 += U * GP_NT) {
            Tw3<T> t[U];
            T x[U][4];
#pragma unroll
            for (int u = 0; u < U; u++) {
                t[u] = load_tw3(tw, ((nblk + sb) << (LOGB - 2)) + g0 + u * GP_NT);
                gld4(src + 4 * (g0 + u * GP_NT), x[u]);
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const uint32_t g = g0 + u * GP_NT;
                inv_net4<T>(x[u], t[u], c.q, c.q2);
                sm.st4(4 * g + ((g >> 4) << 2), x[u]);
            }
        }
    }
}
// inverse smem pass: radix 16 at distance t0 = 2^LT0 (block span 2^(LT0+4))
template <typename T, int LOGB, int LT0>
GP_HD void i_mid(const SmemView<T, LOGB> &sm, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb, const LimbConst<T> &c) {
    constexpr int LSPAN = LT0 + 4;
    constexpr uint32_t t0 = 1u << LT0, groups = 1u << (LOGB - 4);
    constexpr uint32_t per = LOGB - LSPAN;
#pragma unroll 2
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        const uint32_t b = g >> LT0, e = g & (t0 - 1);
        const uint32_t base = padw((b << LSPAN) + e);
        T x[16];
#pragma unroll
        for (int j = 0; j < 16; j++) x[j] = sm.ld(base + joff<LT0>(j));
        inv_net<T, 4, false>(x, tw, ((nblk + sb) << per) + b, c);
#pragma unroll
        for (int j = 0; j < 16; j++) sm.st(base + joff<LT0>(j), x[j]);
    }
}
// inverse last pass: smem -> radix-2^R1 over the whole block -> global; FINAL merges n^-1
template <typename T, int LOGB, bool FINAL>
GP_HD void i_last(const SmemView<T, LOGB> &sm, T *__restrict__ dst, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb,
                  const LimbConst<T> &c) {
    constexpr int R1 = FastGeom<LOGB>::kR1, R = 1 << R1, LT0 = LOGB - R1;
    constexpr uint32_t t0 = 1u << LT0;
#pragma unroll 2
    for (uint32_t g = GP_TID; g < t0; g += GP_NT) {
        const uint32_t base = padw(g);
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = sm.ld(base + joff<LT0>(j));
        inv_net<T, R1, FINAL>(x, tw, nblk + sb, c);
#pragma unroll
        for (int j = 0; j < R; j++) {
            if constexpr (is_f64<T>::value)
                reinterpret_cast<uint64_t *>(dst)[g + j * t0] = f64_canonical(x[j].v, c.q.v, c.q2.v);
            else
                dst[g + j * t0] = csub_min<T>(x[j], c.q);
        }
    }
}

// ---- whole-block transforms ------------------------------------------------------------------------
// everything after pass 1 of a forward block: the radix-16 shared-memory passes and the radix-4 store pass
template <typename T, int LOGB, bool NOCSUB>
GP_HD void fast_fwd_tail(const SmemView<T, LOGB> &sm, T *__restrict__ dst, uint32_t nblk, uint32_t sb,
                         const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    using G = FastGeom<LOGB>;
    constexpr int S1 = LOGB - G::kR1;  // span after pass 1
    if constexpr (G::kK >= 1) {
        f_mid<T, LOGB, S1, NOCSUB>(sm, tw, nblk, sb, c.q, c.q2);
        GP_SYNC();
    }
    if constexpr (G::kK >= 2) {
        f_mid<T, LOGB, S1 - 4, NOCSUB>(sm, tw, nblk, sb, c.q, c.q2);
        GP_SYNC();
    }
    if constexpr (G::kK >= 3) {
        f_mid<T, LOGB, S1 - 8, NOCSUB>(sm, tw, nblk, sb, c.q, c.q2);
        GP_SYNC();
    }
    static_assert(G::kK <= 3, "add another f_mid");
    f_last<T, LOGB, NOCSUB, LOGB>(sm, dst, tw, nblk, sb, c);
}
template <typename T, int LOGB, bool NOCSUB, class Src>
GP_HD void fast_fwd_block(const Src &src, T *__restrict__ dst, void *smem, uint32_t logn, uint32_t sb,
                          const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    const SmemView<T, LOGB> sm(smem);
    const uint32_t nblk = 1u << (logn - LOGB);
    f_first<T, LOGB, NOCSUB>(src, sm, tw, nblk, sb, c.q, c.q2);
    GP_SYNC();
    fast_fwd_tail<T, LOGB, NOCSUB>(sm, dst, nblk, sb, tw, c);
}

// everything before the last pass of an inverse block: the radix-4 load pass and the radix-16 shared-memory passes
template <typename T, int LOGB>
GP_HD void fast_inv_head(const T *__restrict__ src, const SmemView<T, LOGB> &sm, uint32_t nblk, uint32_t sb,
                         const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    using G = FastGeom<LOGB>;
    i_first<T, LOGB>(src, sm, tw, nblk, sb, c);
    GP_SYNC();
    if constexpr (G::kK >= 1) {
        i_mid<T, LOGB, 2>(sm, tw, nblk, sb, c);
        GP_SYNC();
    }
    if constexpr (G::kK >= 2) {
        i_mid<T, LOGB, 6>(sm, tw, nblk, sb, c);
        GP_SYNC();
    }
    if constexpr (G::kK >= 3) {
        i_mid<T, LOGB, 10>(sm, tw, nblk, sb, c);
        GP_SYNC();
    }
}
template <typename T, int LOGB, bool FINAL>
GP_HD void fast_inv_block(const T *__restrict__ src, T *__restrict__ dst, void *smem, uint32_t logn, uint32_t sb,
                          const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    const SmemView<T, LOGB> sm(smem);
    const uint32_t nblk = 1u << (logn - LOGB);
    fast_inv_head<T, LOGB>(src, sm, nblk, sb, tw, c);
    i_last<T, LOGB, FINAL>(sm, dst, tw, nblk, sb, c);
}

// ---- kernels ------------------------------------------------------------------------------------------
// block sizes with a compile-time plan
template <typename T> inline bool fast_supported(int logb) {
    return sizeof(T) == 4 ? (logb >= 12 && logb <= 15) : (logb >= 12 && logb <= 14);
}
// block size used for a transform of 2^logn words: the whole transform when it fits shared memory,
// otherwise 64 KB blocks (three CTAs per SM) under radix-2^k global passes
template <typename T> inline int fast_block_log(int logn) {
    const int cap = sizeof(T) == 4 ? 15 : 14;
    return logn <= cap ? logn : cap - 1;
}
// ---- launch shape of the cluster kernels (ntt_cluster.cuh) ---------------------------------------------------------------------------------------------------
// (block, cluster) split of a ring of 2^logn words: 64 KB blocks (2^14 u32 / 2^13 u64 words), cluster of at most 8
template <typename T> inline int cluster_block_log() { return sizeof(T) == 4 ? 14 : 13; }
template <typename T> inline int cluster_log(int logn) {
    const int lc = logn - cluster_block_log<T>();
    return (lc >= 1 && lc <= 3) ? lc : 0;
}

template <typename T, int LOGB, bool NOCSUB>
GP_HD void fast_fwd_cta(uint32_t bid, void *smem, const T *src, T *dst, uint32_t limbs, uint32_t logn,
                        const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const uint32_t lsub = logn - LOGB;
    const size_t pl = bid >> lsub;
    const uint32_t sb = bid & ((1u << lsub) - 1), limb = (uint32_t)(pl % limbs);
    const size_t off = (pl << logn) + ((size_t)sb << LOGB);
    const LimbConst<T> c = lc[limb];
    if constexpr (is_f64<T>::value) {
        PlainSrc<T> s{src + off, 4 * (uint64_t)c.q.v};
        fast_fwd_block<T, LOGB, NOCSUB>(s, dst + off, smem, logn, sb, tw_all + ((size_t)limb << logn), c);
    } else {
        PlainSrc<T> s{src + off};
        fast_fwd_block<T, LOGB, NOCSUB>(s, dst + off, smem, logn, sb, tw_all + ((size_t)limb << logn), c);
    }
}
template <typename T, int LOGB, bool NOCSUB>
__global__ void __launch_bounds__(FastCfg<T, LOGB>::kThreads, FastCfg<T, LOGB>::kMinBlocks)
fast_ntt_fwd_kernel(const T *src, T *dst, uint32_t limbs, uint32_t logn, const TwPair<T> *__restrict__ tw_all,
                    const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    fast_fwd_cta<T, LOGB, NOCSUB>(blockIdx.x, smem_raw, src, dst, limbs, logn, tw_all, lc);
}

template <typename T, int LOGB>
GP_HD void fast_inv_cta(uint32_t bid, void *smem, const T *src, T *dst, uint32_t limbs, uint32_t logn,
                        const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const uint32_t lsub = logn - LOGB;
    const size_t pl = bid >> lsub;
    const uint32_t sb = bid & ((1u << lsub) - 1), limb = (uint32_t)(pl % limbs);
    const size_t off = (pl << logn) + ((size_t)sb << LOGB);
    const LimbConst<T> c = lc[limb];
    const TwPair<T> *tw = tw_all + ((size_t)limb << logn);
    if (lsub == 0)
        fast_inv_block<T, LOGB, true>(src + off, dst + off, smem, logn, sb, tw, c);
    else
        fast_inv_block<T, LOGB, false>(src + off, dst + off, smem, logn, sb, tw, c);
}
template <typename T, int LOGB>
__global__ void __launch_bounds__(FastCfg<T, LOGB>::kThreads, FastCfg<T, LOGB>::kMinBlocks)
fast_ntt_inv_kernel(const T *src, T *dst, uint32_t limbs, uint32_t logn, const TwPair<T> *__restrict__ tw_all,
                    const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    fast_inv_cta<T, LOGB>(blockIdx.x, smem_raw, src, dst, limbs, logn, tw_all, lc);
}

// gadget digit as a double: any digit below 2^51 is already an exact, small enough multiplicand (the nets renormalise)
struct F64DigitSrc {
    const uint64_t *p;
    uint32_t shift;
    uint64_t mask;
    GP_HD F64 operator()(uint32_t i) const { return F64{f64_from_word((p[i] >> shift) & mask)}; }  // source words < 2^51
};
GP_HD F64DigitSrc make_f64_digit_src(const F64 *__restrict__ src, const DigitSpec &d) {
    const uint64_t mask = d.dbits == 0 ? 0ULL : (d.dbits >= 64 ? ~0ULL : ((1ULL << d.dbits) - 1));
    return F64DigitSrc{reinterpret_cast<const uint64_t *>(src + d.soff), d.shift >= 64 ? 0u : d.shift, d.shift >= 64 ? 0ULL : mask};
}
// fused gadget decomposition + forward NTT (digit extraction is the load functor of pass 1);
// geometry as decompose_cta in ntt.cuh, block = the whole transform (logn == LOGB)
template <typename T, int LOGB, bool NOCSUB>
GP_HD void fast_decompose_cta(uint32_t bid, void *smem, const T *__restrict__ src, T *__restrict__ out, const DecompGeom &g,
                              const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const size_t pl = bid;  // out (poly, limb)
    const DigitSpec ds = digit_spec(g, pl, LOGB, lc);
    const uint32_t limb = ds.limb;
    const LimbConst<T> c = lc[limb];
    if constexpr (is_f64<T>::value) {
        const F64DigitSrc dsrc = make_f64_digit_src(src, ds);
        fast_fwd_block<T, LOGB, NOCSUB>(dsrc, out + (pl << LOGB), smem, LOGB, 0, tw_all + ((size_t)limb << LOGB), c);
    } else {
        const DigitSrc<T> dsrc = make_int_digit_src<T>(src, ds, c.q, c.one_s);
        fast_fwd_block<T, LOGB, NOCSUB>(dsrc, out + (pl << LOGB), smem, LOGB, 0, tw_all + ((size_t)limb << LOGB), c);
    }
}
template <typename T, int LOGB, bool NOCSUB>
__global__ void __launch_bounds__(FastCfg<T, LOGB>::kThreads, FastCfg<T, LOGB>::kMinBlocks)
fast_decompose_kernel(const T *__restrict__ src, T *__restrict__ out, DecompGeom g, const TwPair<T> *__restrict__ tw_all,
                      const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    fast_decompose_cta<T, LOGB, NOCSUB>(blockIdx.x, smem_raw, src, out, g, tw_all, lc);
}

}  // namespace gp
