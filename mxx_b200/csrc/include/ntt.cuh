// ntt.cuh — batched negacyclic NTT / INTT for sm_100a.
//
// Replaces cuda/src/matrix/MatrixNTT.cu (twist kernel :39, bit-reverse :117, one
// radix-2 global pass per stage :152, scale :79; host loop :655-769), i.e. log2(n)+2
// full HBM round trips with a pow_mod per butterfly, by ONE kernel per transform for
// n <= 2^15 (u32) / 2^14 (u64): one CTA owns one (poly, limb) in shared memory.
//
// Algorithm: Cooley-Tukey forward (natural -> bit-reversed) / Gentleman-Sande inverse
// (bit-reversed -> natural) with the psi-twist merged into the twiddles (table
// psi^bitrev(i), Shoup companions) and n^-1 merged into the last inverse stage.  No
// twist, bit-reverse or scale pass exists.  Butterflies are Harvey-lazy: forward values
// live in [0,4q), inverse in [0,2q); one correction at the final store.
//
// Schedule: log2(B) stages are split into register-resident radix-2^k passes
// (k = 5 for u32, 4 for u64: 32 data registers either way).  The forward first pass
// loads straight from global memory (stride >= 32 words: coalesced) and the inverse
// last pass stores straight to global; everything between goes through shared memory
// in a layout padded by one word per bank row, for which every pass is conflict-free
// (radix == words per bank row).
//
// Device EVAL order is bit-reversed: slot s holds X[bitrev(s)], X[k] = a(psi^(2k+1)).
#pragma once
#include "modarith.cuh"

namespace gp {

template <typename T> struct NttCfg {
    static constexpr int kLogRow = sizeof(T) == 4 ? 5 : 4;  // words per 128-byte bank row
    static constexpr int kLogR = kLogRow;                   // max radix (log2)
    static constexpr int kMaxLogB = sizeof(T) == 4 ? 15 : 14;  // largest block held in smem
    static constexpr int kThreads = 512;
};
template <typename T> GP_HD uint32_t spad(uint32_t a) { return a + (a >> NttCfg<T>::kLogRow); }
template <typename T> inline size_t ntt_smem_bytes(int logb) {
    size_t b = (size_t)1 << logb;
    return (b + (b >> NttCfg<T>::kLogRow) + 1) * sizeof(T);
}

// Source functors: logical coefficient index -> value in [0, 4q).
template <typename T> struct PlainSrc {
    const T *p;
    GP_HD T xform(T raw) const { return raw; }
    GP_HD T operator()(uint32_t i) const { return p[i]; }
};
// F64 working type (modarith.cuh): the words in memory are u64 integers below 8q (canonical, or lazily reduced by
// the integer global passes that precede the block kernel for rings above one CTA); two conditional subtractions
// bring the word below 2q < 2^52, where the mantissa trick converts it (modarith.cuh f64_from_word)
template <> struct PlainSrc<F64> {
    const F64 *p;
    uint64_t q4;  // 4q
    GP_HD F64 operator()(uint32_t i) const {
        uint64_t r = reinterpret_cast<const uint64_t *>(p)[i];
        if (r >= q4) r -= q4;
        if (r >= (q4 >> 1)) r -= q4 >> 1;
        return F64{f64_from_word(r)};
    }
};
// Gadget digit of a coefficient-form residue, the fused load of decompose + NTT:
// (x_t >> dg*bb) & mask, reduced into the target limb if needed
// (semantics of cuda/src/matrix/MatrixDecompose.cu:93-113).
template <typename T> struct DigitSrc {
    const T *p;
    uint32_t shift;
    T mask;
    T q;
    T one_s;  // floor(2^wordbits / q)
    // digit in [0, 2q): a digit can only exceed the target modulus when its mask does (a CTA-uniform
    // test), and then one Shoup-multiply-by-one brings it below 2q — the transforms accept that range
    GP_HD T xform(T raw) const {
        T d = (raw >> shift) & mask;
        if (mask >= q) d = d - mulhi(d, one_s) * q;
        return d;
    }
    GP_HD T operator()(uint32_t i) const { return xform(p[i]); }
    GP_HD T canonical(uint32_t i) const { return csub<T>(xform(p[i]), q); }  // [0, q): coefficient-form output
};

// ---- register butterfly networks ------------------------------------------------

// Forward: x[j] sits at position base + j*stride; stage s pairs j and j + (R >> (s+1)).
// twbase = (number of blocks before this pass) + (global block index).
template <typename T, int LOGR>
GP_HD void fwd_network(T (&x)[1 << LOGR], const TwPair<T> *__restrict__ tw, uint32_t twbase, T q, T q2) {
    constexpr int R = 1 << LOGR;
#pragma unroll
    for (int s = 0; s < LOGR; s++) {
        const int half = R >> (s + 1);
#pragma unroll
        for (int u = 0; u < (1 << s); u++) {
            const TwPair<T> w = tw[(twbase << s) + u];
#pragma unroll
            for (int v = 0; v < half; v++) {
                const int j0 = u * 2 * half + v, j1 = j0 + half;
                T X = csub<T>(x[j0], q2);
                T t = mul_shoup_lazy<T>(x[j1], w.w, w.ws, q);
                x[j0] = X + t;
                x[j1] = X - t + q2;
            }
        }
    }
}

// Inverse: x[j] sits at position base + j*t0; stage s pairs j and j + (1 << s).
// twbase = (number of blocks after this pass) + (global block index).
template <typename T, int LOGR, bool FINAL>
GP_HD void inv_network(T (&x)[1 << LOGR], const TwPair<T> *__restrict__ tw, uint32_t twbase,
                                            const LimbConst<T> &c) {
    constexpr int R = 1 << LOGR;
    const T q = c.q, q2 = c.q2;
#pragma unroll
    for (int s = 0; s < LOGR; s++) {
        const int groups = R >> (s + 1);
#pragma unroll
        for (int u = 0; u < groups; u++) {
            if (FINAL && s == LOGR - 1) {
#pragma unroll
                for (int v = 0; v < (1 << s); v++) {
                    const int j0 = u * (2 << s) + v, j1 = j0 + (1 << s);
                    T X = x[j0], Y = x[j1];
                    x[j0] = mul_shoup_lazy<T>(X + Y, c.ninv, c.ninv_s, q);
                    x[j1] = mul_shoup_lazy<T>(X - Y + q2, c.nw, c.nw_s, q);
                }
            } else {
                const TwPair<T> w = tw[twbase * groups + u];
#pragma unroll
                for (int v = 0; v < (1 << s); v++) {
                    const int j0 = u * (2 << s) + v, j1 = j0 + (1 << s);
                    T X = x[j0], Y = x[j1];
                    x[j0] = csub<T>(X + Y, q2);
                    x[j1] = mul_shoup_lazy<T>(X - Y + q2, w.w, w.ws, q);
                }
            }
        }
    }
}

// ---- one pass over a block of B = 2^logb words -----------------------------------
// span: block size at pass start (forward) / end (inverse); stride = span >> LOGR.
// blk0: global index of this CTA's first span-block; nblk: number of span-blocks in
// the whole transform (n / span).

template <typename T, int LOGR, class Src>
GP_HD void fwd_pass_from_src(const Src &src, T *sm, uint32_t logb, uint32_t nblk, uint32_t blk0,
                                                  const TwPair<T> *__restrict__ tw, T q, T q2) {
    constexpr int R = 1 << LOGR;
    const uint32_t lstride = logb - LOGR, stride = 1u << lstride, groups = 1u << lstride;
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = src(g + j * stride);
        fwd_network<T, LOGR>(x, tw, nblk + blk0, q, q2);
#pragma unroll
        for (int j = 0; j < R; j++) sm[spad<T>(g + j * stride)] = x[j];
    }
}

template <typename T, int LOGR>
GP_HD void fwd_pass_smem(T *sm, uint32_t logb, uint32_t lspan, uint32_t nblk, uint32_t blk0,
                                              const TwPair<T> *__restrict__ tw, T q, T q2) {
    constexpr int R = 1 << LOGR;
    const uint32_t lstride = lspan - LOGR, stride = 1u << lstride, groups = 1u << (logb - LOGR);
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        const uint32_t b = g >> lstride, e = g & (stride - 1), base = (b << lspan) + e;
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = sm[spad<T>(base + j * stride)];
        fwd_network<T, LOGR>(x, tw, nblk + blk0 + b, q, q2);
#pragma unroll
        for (int j = 0; j < R; j++) sm[spad<T>(base + j * stride)] = x[j];
    }
}

template <typename T, int LOGR>
GP_HD void inv_pass_smem(T *sm, uint32_t logb, uint32_t lt0, uint32_t nblk, uint32_t blk0,
                                              const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    constexpr int R = 1 << LOGR;
    const uint32_t t0 = 1u << lt0, lspan = lt0 + LOGR, groups = 1u << (logb - LOGR);
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        const uint32_t b = g >> lt0, e = g & (t0 - 1), base = (b << lspan) + e;
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = sm[spad<T>(base + j * t0)];
        inv_network<T, LOGR, false>(x, tw, nblk + blk0 + b, c);
#pragma unroll
        for (int j = 0; j < R; j++) sm[spad<T>(base + j * t0)] = x[j];
    }
}

// last inverse pass of a block: span == B, writes to global; FINAL merges n^-1.
template <typename T, int LOGR, bool FINAL>
GP_HD void inv_pass_to_dst(T *sm, T *__restrict__ dst, uint32_t logb, uint32_t nblk, uint32_t blk0,
                                                const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    constexpr int R = 1 << LOGR;
    const uint32_t lt0 = logb - LOGR, t0 = 1u << lt0, groups = t0;
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = sm[spad<T>(g + j * t0)];
        inv_network<T, LOGR, FINAL>(x, tw, nblk + blk0, c);
#pragma unroll
        for (int j = 0; j < R; j++) dst[g + j * t0] = csub<T>(x[j], c.q);
    }
}

// ---- whole-block transforms --------------------------------------------------------
// logn: full transform; logb: block per CTA; sb: block index within the transform.
// Forward expects `top = logn - logb` stages already done by global passes.

template <typename T, class Src>
GP_HD void fwd_block(const Src &src, T *__restrict__ dst, T *sm, uint32_t logn, uint32_t logb,
                                          uint32_t sb, const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    constexpr int LR = NttCfg<T>::kLogR;
    const T q = c.q, q2 = c.q2;
    const uint32_t r = logb % LR ? logb % LR : (logb < (uint32_t)LR ? logb : LR);  // first-pass radix
    uint32_t nblk = 1u << (logn - logb);  // blocks of size B in the transform
    switch (r) {
        case 1: fwd_pass_from_src<T, 1>(src, sm, logb, nblk, sb, tw, q, q2); break;
        case 2: fwd_pass_from_src<T, 2>(src, sm, logb, nblk, sb, tw, q, q2); break;
        case 3: fwd_pass_from_src<T, 3>(src, sm, logb, nblk, sb, tw, q, q2); break;
        case 4: fwd_pass_from_src<T, 4>(src, sm, logb, nblk, sb, tw, q, q2); break;
        default:
            if constexpr (LR == 5) fwd_pass_from_src<T, 5>(src, sm, logb, nblk, sb, tw, q, q2);
            break;
    }
    uint32_t lspan = logb - r;
    while (lspan > 0) {
        GP_SYNC();
        const uint32_t per = logb - lspan;  // log2(span-blocks per CTA block)
        fwd_pass_smem<T, LR>(sm, logb, lspan, nblk << per, sb << per, tw, q, q2);
        lspan -= LR;
    }
    GP_SYNC();
    const uint32_t B = 1u << logb;
    for (uint32_t i = GP_TID; i < B; i += GP_NT) {
        T v = sm[spad<T>(i)];
        v = csub<T>(v, q2);
        dst[i] = csub<T>(v, q);
    }
}

// Inverse: does the lowest logb stages of the block; FINAL (logb == logn) merges n^-1.
template <typename T, bool FINAL>
GP_HD void inv_block(const T *__restrict__ src, T *__restrict__ dst, T *sm, uint32_t logn,
                                          uint32_t logb, uint32_t sb, const TwPair<T> *__restrict__ tw,
                                          const LimbConst<T> &c) {
    constexpr int LR = NttCfg<T>::kLogR;
    const uint32_t B = 1u << logb;
    for (uint32_t i = GP_TID; i < B; i += GP_NT) sm[spad<T>(i)] = src[i];
    const uint32_t r = logb % LR ? logb % LR : (logb < (uint32_t)LR ? logb : LR);  // last-pass radix
    uint32_t lt0 = 0;
    while (lt0 + r < logb) {
        GP_SYNC();
        const uint32_t lspan = lt0 + LR;
        inv_pass_smem<T, LR>(sm, logb, lt0, 1u << (logn - lspan), sb << (logb - lspan), tw, c);
        lt0 += LR;
    }
    GP_SYNC();
    const uint32_t nblk = 1u << (logn - logb);
    switch (r) {
        case 1: inv_pass_to_dst<T, 1, FINAL>(sm, dst, logb, nblk, sb, tw, c); break;
        case 2: inv_pass_to_dst<T, 2, FINAL>(sm, dst, logb, nblk, sb, tw, c); break;
        case 3: inv_pass_to_dst<T, 3, FINAL>(sm, dst, logb, nblk, sb, tw, c); break;
        case 4: inv_pass_to_dst<T, 4, FINAL>(sm, dst, logb, nblk, sb, tw, c); break;
        default:
            if constexpr (LR == 5) inv_pass_to_dst<T, 5, FINAL>(sm, dst, logb, nblk, sb, tw, c);
            break;
    }
}

// ---- kernels -------------------------------------------------------------------------
// grid.x = polys * limbs * (n / B); data layout [poly][limb][n].

template <typename T>
GP_HD void ntt_fwd_cta(uint32_t bid, T *sm, const T *src, T *dst, uint32_t limbs, uint32_t logn, uint32_t logb,
                       const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const uint32_t lsub = logn - logb;
    const size_t pl = bid >> lsub;
    const uint32_t sb = bid & ((1u << lsub) - 1), limb = (uint32_t)(pl % limbs);
    const size_t off = (pl << logn) + ((size_t)sb << logb);
    const LimbConst<T> c = lc[limb];
    PlainSrc<T> s{src + off};
    fwd_block<T>(s, dst + off, sm, logn, logb, sb, tw_all + ((size_t)limb << logn), c);
}
template <typename T>
__global__ void __launch_bounds__(NttCfg<T>::kThreads)
ntt_fwd_kernel(const T *src, T *dst, uint32_t limbs, uint32_t logn, uint32_t logb, const TwPair<T> *__restrict__ tw_all,
               const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ntt_fwd_cta<T>(blockIdx.x, reinterpret_cast<T *>(smem_raw), src, dst, limbs, logn, logb, tw_all, lc);
}

template <typename T>
GP_HD void ntt_inv_cta(uint32_t bid, T *sm, const T *src, T *dst, uint32_t limbs, uint32_t logn, uint32_t logb,
                       const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const uint32_t lsub = logn - logb;
    const size_t pl = bid >> lsub;
    const uint32_t sb = bid & ((1u << lsub) - 1), limb = (uint32_t)(pl % limbs);
    const size_t off = (pl << logn) + ((size_t)sb << logb);
    const LimbConst<T> c = lc[limb];
    const TwPair<T> *tw = tw_all + ((size_t)limb << logn);
    if (lsub == 0)
        inv_block<T, true>(src + off, dst + off, sm, logn, logb, sb, tw, c);
    else
        inv_block<T, false>(src + off, dst + off, sm, logn, logb, sb, tw, c);
}
template <typename T>
__global__ void __launch_bounds__(NttCfg<T>::kThreads)
ntt_inv_kernel(const T *src, T *dst, uint32_t limbs, uint32_t logn, uint32_t logb, const TwPair<T> *__restrict__ tw_all,
               const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ntt_inv_cta<T>(blockIdx.x, reinterpret_cast<T *>(smem_raw), src, dst, limbs, logn, logb, tw_all, lc);
}

// Fused gadget decomposition + forward NTT: out row row*ell + t*dpt + dg, every limb
// l, is NTT_l(digit_{t,dg}(src[row][col])).  One CTA per (out poly, limb, block).
// Replaces memset + matrix_decompose_all_slots_kernel + a full NTT sweep
// (cuda/src/matrix/MatrixDecompose.cu:1155-1161, 1240-1328).
struct DecompGeom {
    uint32_t limbs;      // active limbs of the OUTPUT (the limbs of the tables the kernel is given)
    uint32_t src_cols;   // columns of the source matrix
    uint32_t src_col0;   // first source column of this chunk
    uint32_t cols;       // columns of the output (chunk width)
    uint32_t ell;        // digits per source poly in the output (dpt*towers, or dpt if small)
    uint32_t dpt;        // digits per tower
    uint32_t base_bits;
    // Source addressing.  An ordinary matrix stores [poly][tower][n] with as many towers as the output has limbs
    // (gathered == 0).  A LIMB-SHARDED caller (mxx_b200/dist.py, SURVEY 8e "Alternative (limbs)") hands over the
    // coefficient-form source as all-gathered per-rank blocks [group][poly][towers_per_group][n]: tower t of source
    // poly p then sits at (t / src_tpg) * src_group_stride + p * src_poly_stride + (t % src_tpg) * n words, and the
    // towers' bit widths (the FULL basis, not the output's limbs) travel in tower_bits.
    uint32_t gathered;
    uint32_t src_tpg;
    unsigned long long src_group_stride, src_poly_stride;
    unsigned long long tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7;  // 8 bits per tower, tower t in word t / 8 (see tower_bits_of)
};
// bit width of source tower t of a gathered source.  The word is selected by compares on named fields: a dynamic index
// into a kernel parameter makes the compiler copy the whole geometry to local memory (a 120-byte stack frame).
GP_HD uint32_t tower_bits_of(const DecompGeom &g, uint32_t t) {
    const uint32_t i = t >> 3;
    const unsigned long long lo = i < 2 ? (i == 0 ? g.tb0 : g.tb1) : (i == 2 ? g.tb2 : g.tb3);
    const unsigned long long hi = i < 6 ? (i == 4 ? g.tb4 : g.tb5) : (i == 6 ? g.tb6 : g.tb7);
    return (uint32_t)((i < 4 ? lo : hi) >> ((t & 7) * 8)) & 0xffu;
}
GP_HD void set_tower_bits(DecompGeom &g, uint32_t t, uint32_t bits) {
    unsigned long long *w[8] = {&g.tb0, &g.tb1, &g.tb2, &g.tb3, &g.tb4, &g.tb5, &g.tb6, &g.tb7};
    *w[(t >> 3) & 7] |= (unsigned long long)(bits & 0xffu) << ((t & 7) * 8);
}
// Where one output (poly, limb) of a decomposition reads its digits: source offset (words), shift and bit count of the
// digit (0 bits: the digit is zero), output limb.  Row order row*ell + t*dpt + dg with the narrowed last digit of a
// tower (cuda/src/matrix/MatrixDecompose.cu:93-113).
struct DigitSpec {
    size_t soff;
    uint32_t limb, shift, dbits;
};
template <typename LC>
GP_HD DigitSpec digit_spec(const DecompGeom &g, size_t pl, uint32_t logn, const LC *__restrict__ lc) {
    DigitSpec d;
    d.limb = (uint32_t)(pl % g.limbs);
    const size_t opoly = pl / g.limbs;
    const size_t orow = opoly / g.cols, col = opoly % g.cols;
    const size_t row = orow / g.ell;
    const uint32_t loc = (uint32_t)(orow % g.ell), t = loc / g.dpt, dg = loc % g.dpt;
    const size_t spoly = row * g.src_cols + g.src_col0 + col;
    uint32_t sbits;
    if (g.gathered) {
        sbits = tower_bits_of(g, t);
        d.soff = (size_t)(t / g.src_tpg) * g.src_group_stride + spoly * g.src_poly_stride + ((size_t)(t % g.src_tpg) << logn);
    } else {
        sbits = lc[t].bits;
        d.soff = (spoly * g.limbs + t) << logn;
    }
    d.shift = dg * g.base_bits;
    d.dbits = 0;
    if (d.shift < sbits) {
        const uint32_t rem = sbits - d.shift;
        d.dbits = g.base_bits < rem ? g.base_bits : rem;
    }
    return d;
}
template <typename T> GP_HD DigitSrc<T> make_int_digit_src(const T *__restrict__ src, const DigitSpec &d, T q, T one_s) {
    constexpr uint32_t W = sizeof(T) * 8;
    const T mask = d.dbits == 0 ? (T)0 : (d.dbits >= W ? ~(T)0 : (T)(((T)1 << d.dbits) - 1));
    return DigitSrc<T>{src + d.soff, d.shift >= W ? 0u : d.shift, d.shift >= W ? (T)0 : mask, q, one_s};
}
template <typename T, bool TO_EVAL>
GP_HD void decompose_cta(uint32_t bid, T *sm, const T *__restrict__ src, T *__restrict__ out, const DecompGeom &g,
                         uint32_t logn, uint32_t logb, const TwPair<T> *__restrict__ tw_all,
                         const LimbConst<T> *__restrict__ lc) {
    const uint32_t lsub = logn - logb;
    const size_t pl = bid >> lsub;  // out (poly, limb)
    const uint32_t sb = bid & ((1u << lsub) - 1);
    const DigitSpec ds = digit_spec(g, pl, logn, lc);
    const uint32_t limb = ds.limb;
    const LimbConst<T> c = lc[limb];
    T *dp = out + (pl << logn) + ((size_t)sb << logb);
    DigitSrc<T> dsrc = make_int_digit_src<T>(src, ds, c.q, c.one_s);
    dsrc.p += (size_t)sb << logb;
    if (TO_EVAL) {
        fwd_block<T>(dsrc, dp, sm, logn, logb, sb, tw_all + ((size_t)limb << logn), c);
    } else {
        const uint32_t B = 1u << logb;
        for (uint32_t i = GP_TID; i < B; i += GP_NT) dp[i] = dsrc.canonical(i);
    }
}
template <typename T, bool TO_EVAL>
__global__ void __launch_bounds__(NttCfg<T>::kThreads)
decompose_kernel(const T *__restrict__ src, T *__restrict__ out, DecompGeom g, uint32_t logn, uint32_t logb,
                 const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    decompose_cta<T, TO_EVAL>(blockIdx.x, reinterpret_cast<T *>(smem_raw), src, out, g, logn, logb, tw_all, lc);
}

// ---- global radix passes for n larger than one CTA's shared memory --------------------
// Forward top pass: stages with span in (B, n]; one thread per radix-R group.
template <typename T, int LOGR>
GP_HD void fwd_global_thread(size_t gid, T *data, uint32_t limbs, uint32_t logn, uint32_t lspan,
                             const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    constexpr int R = 1 << LOGR;
    const uint32_t lg = logn - LOGR;  // log2(groups per transform)
    const size_t pl = gid >> lg;
    const uint32_t g = (uint32_t)(gid & ((1u << lg) - 1)), limb = (uint32_t)(pl % limbs);
    const uint32_t lstride = lspan - LOGR, stride = 1u << lstride;
    const uint32_t b = g >> lstride, e = g & (stride - 1);
    T *base = data + (pl << logn) + ((size_t)b << lspan) + e;
    const LimbConst<T> c = lc[limb];
    T x[R];
#pragma unroll
    for (int j = 0; j < R; j++) x[j] = base[(size_t)j * stride];
    fwd_network<T, LOGR>(x, tw_all + ((size_t)limb << logn), (1u << (logn - lspan)) + b, c.q, c.q2);
#pragma unroll
    for (int j = 0; j < R; j++) base[(size_t)j * stride] = x[j];  // lazy [0,4q): the next pass accepts it
}
// Gadget decomposition fused into the FIRST forward top pass (span n): the digits are extracted while the radix-R
// groups are gathered, so the coefficient-form digit matrix is never written and re-read (rings above one CTA,
// e.g. n = 2^15 with u64 words).  Geometry as decompose_cta; out receives what decompose_kernel<false> followed by
// ntt_fwd_global_pass<LOGR>(lspan = logn) would leave.
// One CTA covers 2^kDecompTopChunkLog (or all, for small rings) radix-R groups of ONE output (poly, limb), so the
// digit geometry (a handful of 64-bit divisions) is worked out once per CTA and not per group.
constexpr uint32_t kDecompTopChunkLog = 11;
template <int LOGR> GP_HD uint32_t decompose_top_chunk_log(uint32_t logn) {
    const uint32_t lg = logn - LOGR;
    return lg < kDecompTopChunkLog ? lg : kDecompTopChunkLog;
}
template <typename T, int LOGR>
GP_HD void decompose_fwd_global_cta(size_t bid, const T *__restrict__ src, T *__restrict__ out, const DecompGeom &g,
                                    uint32_t logn, const TwPair<T> *__restrict__ tw_all,
                                    const LimbConst<T> *__restrict__ lc) {
    constexpr int R = 1 << LOGR;
    const uint32_t lg = logn - LOGR, stride = 1u << lg, lchunk = decompose_top_chunk_log<LOGR>(logn);
    const size_t pl = bid >> (lg - lchunk);  // out (poly, limb)
    const uint32_t e0 = (uint32_t)(bid & ((1u << (lg - lchunk)) - 1)) << lchunk;
    const DigitSpec ds = digit_spec(g, pl, logn, lc);
    const uint32_t limb = ds.limb;
    const LimbConst<T> c = lc[limb];
    DigitSrc<T> dsrc = make_int_digit_src<T>(src, ds, c.q, c.one_s);
    dsrc.p += e0;
    const TwPair<T> *tw = tw_all + ((size_t)limb << logn);
    T *base = out + (pl << logn) + e0;
    for (uint32_t i = GP_TID; i < (1u << lchunk); i += GP_NT) {
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = dsrc(i + (uint32_t)j * stride);
        fwd_network<T, LOGR>(x, tw, 1u, c.q, c.q2);
#pragma unroll
        for (int j = 0; j < R; j++) base[i + (size_t)j * stride] = x[j];
    }
}
template <typename T, int LOGR>
__global__ void __launch_bounds__(256)
decompose_fwd_global_pass(const T *__restrict__ src, T *__restrict__ out, DecompGeom g, uint32_t logn,
                          const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    decompose_fwd_global_cta<T, LOGR>(blockIdx.x, src, out, g, logn, tw_all, lc);
}
template <typename T, int LOGR>
__global__ void __launch_bounds__(256)
ntt_fwd_global_pass(T *data, size_t total, uint32_t limbs, uint32_t logn, uint32_t lspan,
                    const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid < total) fwd_global_thread<T, LOGR>(gid, data, limbs, logn, lspan, tw_all, lc);
}
// Inverse top pass: distances t0 .. t0*R/2 with t0 >= B; FINAL merges n^-1.
template <typename T, int LOGR, bool FINAL>
GP_HD void inv_global_thread(size_t gid, T *data, uint32_t limbs, uint32_t logn, uint32_t lt0,
                             const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    constexpr int R = 1 << LOGR;
    const uint32_t lg = logn - LOGR;
    const size_t pl = gid >> lg;
    const uint32_t g = (uint32_t)(gid & ((1u << lg) - 1)), limb = (uint32_t)(pl % limbs);
    const uint32_t t0 = 1u << lt0, lspan = lt0 + LOGR;
    const uint32_t b = g >> lt0, e = g & (t0 - 1);
    T *base = data + (pl << logn) + ((size_t)b << lspan) + e;
    const LimbConst<T> c = lc[limb];
    T x[R];
#pragma unroll
    for (int j = 0; j < R; j++) x[j] = base[(size_t)j * t0];
    inv_network<T, LOGR, FINAL>(x, tw_all + ((size_t)limb << logn), (1u << (logn - lspan)) + b, c);
#pragma unroll
    for (int j = 0; j < R; j++) base[(size_t)j * t0] = FINAL ? csub<T>(x[j], c.q) : x[j];
}
template <typename T, int LOGR, bool FINAL>
__global__ void __launch_bounds__(256)
ntt_inv_global_pass(T *data, size_t total, uint32_t limbs, uint32_t logn, uint32_t lt0,
                    const TwPair<T> *__restrict__ tw_all, const LimbConst<T> *__restrict__ lc) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid < total) inv_global_thread<T, LOGR, FINAL>(gid, data, limbs, logn, lt0, tw_all, lc);
}

// ---- host-side schedule shared by the launcher (MatrixNTT.inl) and the emulation harness ----
template <typename T> inline int ntt_block_log(int logn) { return logn < NttCfg<T>::kMaxLogB ? logn : NttCfg<T>::kMaxLogB; }
// Splits `top` stages into passes of at most 4 stages (16 strided words per thread keep
// every 32-byte sector of the global accesses fully used for stride >= 8 words).
inline void split_top(int top, int *passes, int *count) {
    *count = 0;
    while (top > 0) {
        int r = top >= 4 ? 4 : top;
        if (top > 4 && top < 8 && top - r < 2) r = top - 2;  // avoid a trailing radix-2 pass
        passes[(*count)++] = r;
        top -= r;
    }
}


}  // namespace gp
