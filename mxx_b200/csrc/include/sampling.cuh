// sampling.cuh — counter-based randomness and the integer samplers of the trapdoor path
// (SURVEY §8 f-1).  Per-element functions are GP_HD so the host emulation harness
// (src/emu.cu, tests only) can run the very same code without a GPU.
//
// Behaviour this file reproduces (so that a given GpuRngSeed yields the same matrix as the
// reference's CUDA library):
//   * random streams: ChaCha20 keyed by HChaCha20(seed, domain || stream2), block counter
//     starting at stream0, nonce = stream1, eight u64 per block (cuda/src/ChaCha.cu:109-166);
//   * uniform deviates u = (u64 >> 11) * 2^-53 clamped into (0,1)
//     (cuda/src/matrix/MatrixSampling.cu:6-19);
//   * the discrete Gaussian: Karney, "Sampling exactly from the normal distribution"
//     (ACM TOMS 2016), algorithms H, G, P, B and the integer sampler with the iteration
//     caps and the Box-Muller fall-back of MatrixSampling.cu:31-160;
//   * the G-lattice sampler for q != b^k (Genise-Micciancio, EUROCRYPT 2018, alg. 2-4) with
//     continuous perturbation, as in MatrixTrapdoor.cu:701-833.
// The order in which deviates are drawn is part of the contract; the code organisation
// (one register-resident stream object, one fused sample+scatter kernel per call, all RNS
// limbs written by the thread that drew the integer) is this library's own.
#pragma once
#include <math.h>
#include <stdint.h>

#include "modarith.cuh"

namespace gp {

struct Seed256 {
    uint64_t w[4];
};

// domain separators (ASCII tags) — one per distribution, MatrixSampling.cu:247-292,
// MatrixTrapdoor.cu:236, 770
constexpr uint64_t kDomUniform = 0x6f70656e66686531ULL;
constexpr uint64_t kDomGauss = 0x6f70656e66686532ULL;
constexpr uint64_t kDomBit = 0x6f70656e66686533ULL;
constexpr uint64_t kDomTernary = 0x6f70656e66686534ULL;
constexpr uint64_t kDomGadget = 0x6761646765746731ULL;
constexpr uint64_t kDomPerturb = 0x7065727475726231ULL;

constexpr int kMaxGaussDigits = 64;  // MatrixUtils.cu:1

#define GP_CC_QR(a, b, c, d)                       \
    x[a] += x[b]; x[d] ^= x[a]; x[d] = rotl(x[d], 16); \
    x[c] += x[d]; x[b] ^= x[c]; x[b] = rotl(x[b], 12); \
    x[a] += x[b]; x[d] ^= x[a]; x[d] = rotl(x[d], 8);  \
    x[c] += x[d]; x[b] ^= x[c]; x[b] = rotl(x[b], 7)

struct ChaChaStream {
    uint32_t key[8];   // HChaCha20 sub-key
    uint32_t ctr[4];   // words 12..15: 64-bit block counter (stream0), 64-bit nonce (stream1)
    uint32_t out[16];  // current key-stream block
    uint32_t pos;      // next u64 of `out`; 8 = exhausted

    static GP_HD uint32_t rotl(uint32_t v, int n) {
#ifdef __CUDA_ARCH__
        return __funnelshift_l(v, v, n);
#else
        return (v << n) | (v >> (32 - n));
#endif
    }
    // the 20-round ChaCha permutation (RFC 8439 §2.3, without the feed-forward).  Kept out of line (it is reached
    // from several draw sites) and run on a register copy: through the pointer every quarter-round step would be
    // a local-memory round trip (62 % of the instructions of the first version, profiles/r01_ncu_sampler.txt).
    static __host__ __device__ __noinline__ void permute(uint32_t *state) {
        uint32_t x[16];
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = state[i];
#pragma unroll
        for (int dr = 0; dr < 10; dr++) {
            GP_CC_QR(0, 4, 8, 12); GP_CC_QR(1, 5, 9, 13); GP_CC_QR(2, 6, 10, 14); GP_CC_QR(3, 7, 11, 15);
            GP_CC_QR(0, 5, 10, 15); GP_CC_QR(1, 6, 11, 12); GP_CC_QR(2, 7, 8, 13); GP_CC_QR(3, 4, 9, 14);
        }
#pragma unroll
        for (int i = 0; i < 16; i++) state[i] = x[i];
    }
    static GP_HD void sigma(uint32_t *x) {
        x[0] = 0x61707865u; x[1] = 0x3320646eu; x[2] = 0x79622d32u; x[3] = 0x6b206574u;
    }
    GP_HD void open(const Seed256 &seed, uint64_t stream0, uint64_t stream1, uint64_t stream2, uint64_t domain) {
        uint32_t x[16];
        sigma(x);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            x[4 + 2 * i] = (uint32_t)seed.w[i];
            x[5 + 2 * i] = (uint32_t)(seed.w[i] >> 32);
        }
        x[12] = (uint32_t)domain; x[13] = (uint32_t)(domain >> 32);
        x[14] = (uint32_t)stream2; x[15] = (uint32_t)(stream2 >> 32);
        permute(x);  // HChaCha20: first and last rows of the permuted state
#pragma unroll
        for (int i = 0; i < 4; i++) {
            key[i] = x[i];
            key[4 + i] = x[12 + i];
        }
        ctr[0] = (uint32_t)stream0; ctr[1] = (uint32_t)(stream0 >> 32);
        ctr[2] = (uint32_t)stream1; ctr[3] = (uint32_t)(stream1 >> 32);
        pos = 8;
    }
    GP_HD void refill() {
        uint32_t x[16];
        sigma(x);
#pragma unroll
        for (int i = 0; i < 8; i++) x[4 + i] = key[i];
#pragma unroll
        for (int i = 0; i < 4; i++) x[12 + i] = ctr[i];
        permute(x);
        out[0] = x[0] + 0x61707865u; out[1] = x[1] + 0x3320646eu; out[2] = x[2] + 0x79622d32u; out[3] = x[3] + 0x6b206574u;
#pragma unroll
        for (int i = 0; i < 8; i++) out[4 + i] = x[4 + i] + key[i];
#pragma unroll
        for (int i = 0; i < 4; i++) out[12 + i] = x[12 + i] + ctr[i];
        if (++ctr[0] == 0) ++ctr[1];
        pos = 0;
    }
    GP_HD uint64_t next() {
        if (pos >= 8) refill();
        const uint64_t v = (uint64_t)out[2 * pos] | ((uint64_t)out[2 * pos + 1] << 32);
        pos++;
        return v;
    }
    // stream words handed out so far by a stream opened at block counter stream0 (fewer than 2^35)
    GP_HD uint32_t consumed(uint64_t stream0) const { return pos == 8 && ctr[0] == (uint32_t)stream0 ? 0u : (ctr[0] - (uint32_t)stream0 - 1u) * 8u + pos; }
    // uniform deviate in the open interval (0,1), 53 bits
    static GP_HD double to_unit(uint64_t w) {
        const double scale = 1.0 / 9007199254740992.0;
        double u = (double)(w >> 11) * scale;
        if (u <= 0.0) u = scale;
        else if (u >= 1.0) u = 1.0 - scale;
        return u;
    }
    GP_HD double unit() { return to_unit(next()); }
    // uniform residue in [0, q): rejection below the largest multiple of q
    GP_HD uint64_t below(uint64_t q) {
        if (q == 0) return 0;
        const uint64_t lim = ~0ULL - (~0ULL % q);
        for (;;) {
            const uint64_t v = next();
            if (v < lim) return v % q;
        }
    }
    // Box-Muller (only the fall-back of the integer sampler and the continuous perturbation)
    GP_HD double normal() {
        const double u1 = unit();
        const double u2 = unit();
        const double r = sqrt(-2.0 * log(u1));
        const double theta = 6.283185307179586476925286766559 * u2;
        return r * cos(theta);
    }
};

// z mod q for a signed 64-bit z
GP_HD uint64_t signed_residue(int64_t z, uint64_t q) {
    if (z >= 0) return (uint64_t)z % q;
    const uint64_t mag = (uint64_t)(-(z + 1)) + 1;
    const uint64_t r = mag % q;
    return r == 0 ? 0 : q - r;
}

// centred representative of x mod q in (-q/2, q/2]
GP_HD int64_t centred(uint64_t x, uint64_t q) {
    const uint64_t r = x % q;
    return r <= (q >> 1) ? (int64_t)r : -(int64_t)(q - r);
}

// ---- Karney's exact discrete Gaussian --------------------------------------------------
// Algorithms H (true w.p. e^{-1/2}), G (k >= 0 w.p. e^{-k/2}(1 - e^{-1/2})), P (true w.p. e^{-n/2}),
// B (true w.p. e^{-x(2k+x)/(2k+2)}) and the integer sampler built on them, with the reference's iteration caps and
// Box-Muller fall-back (MatrixSampling.cu:31-160) — written as a state machine that consumes exactly ONE stream
// word per step().  Why: as nested rejection loops the lanes of a warp sit at different program points almost all
// the time (3.6 of 32 lanes active per instruction and a ChaCha permutation per lane per block, ncu
// profiles/r01_ncu_sampler.txt).  With one word per step every lane of a warp draws at the same instruction,
// key-stream blocks are refilled by all lanes together, and only the few-instruction case bodies diverge.
// The sequence of words each decision consumes is unchanged, so samples are identical to the nested-loop form.
// What a draw needs from its deviation alone: callers that draw many integers with the same deviation (every task of a
// sample_distribution launch; the first / the later draws of every G-lattice task of a tower) set it up once, with all lanes
// active, instead of dividing 2^64 by the span inside the rejection loop.
struct KarneyScale {
    double sd;
    int64_t span;     // ceil(sd)
    uint64_t span_m;  // floor((2^64 - 1) / span): `word % span` without a 64-bit division per attempt
    bool valid;       // false: the draw degenerates to llround(mean)
    GP_HD void init(double sd_) {
        sd = sd_, span = 0, span_m = 0;
        valid = sd > 0.0 && isfinite(sd);
        if (valid) {
            span = (int64_t)ceil(sd);
            valid = span > 0;
        }
        if (valid) span_m = ~0ULL / (uint64_t)span;
    }
};

struct KarneyMachine {
    enum : uint32_t { kGH0, kGH1, kGH2, kPH0, kPH1, kPH2, kSign, kJ, kB0, kB1, kFb1, kFb2, kDone };
    uint32_t st;
    int32_t k, n, h, bn, tries;
    double mean, sd, hv, x, x0, y, z, bthr;  // bthr = (2k + x) / (2k + 2), fixed for all B tests of one attempt
    int64_t i0, j, sgn, span, result;
    uint64_t span_m;

    GP_HD bool done() const { return st == kDone; }
    GP_HD void finish(int64_t r) { result = r, st = kDone; }
    GP_HD void begin(double mean_, const KarneyScale &sc) {
        mean = mean_, sd = sc.sd, span = sc.span, span_m = sc.span_m, tries = 0;
        if (!sc.valid || !isfinite(mean)) return finish((int64_t)llround(mean));
        restart();
    }
    GP_HD void begin(double mean_, double sd_) {
        KarneyScale sc;
        sc.init(sd_);
        begin(mean_, sc);
    }
    // top of the rejection loop (65536 attempts, then the continuous fall-back)
    GP_HD void restart() {
        if (tries >= (1 << 16)) { st = kFb1; return; }
        ++tries, k = 0, st = kGH0;
    }
    // P: `while (n-- && H()) {}` accepted when n ran out
    GP_HD void p_check() {
        if (n == 0) st = kSign;
        else --n, st = kPH0;
    }
    // `h = k + 1; while (h-- > 0 && B()) {}` accepted when h ran out
    GP_HD void b_check() {
        if (h > 0) --h, y = x, bn = 0, st = kB0;
        else finish(sgn * (i0 + j));
    }
    GP_HD void h_result(bool r, bool in_g) {
        if (in_g) {
            if (r && ++k <= 1024) { st = kGH0; return; }
            n = k * (k - 1);
            p_check();
        } else if (r) p_check();
        else restart();
    }
    GP_HD void b_result() {
        if ((bn & 1) == 0) b_check();
        else restart();
    }
    // The six H states share one body and so do the two B states: lanes of a warp that sit in different states of the
    // same test then execute the same instructions (ncu: with one case body per state, half of the warp instructions
    // of the sampler kernels ran with fewer than 4 active lanes, profiles/r01_ncu_sampler.txt).
    GP_HD void step(uint64_t w) {
        const double u = ChaChaStream::to_unit(w);
        if (st <= kPH2) {
            // H: a = U, !(a < 1/2) -> true; b = U, !(b < a) -> false; a = U, !(a < b) -> true; ...
            const bool in_g = st < kPH0;
            const uint32_t sub = in_g ? st : st - kPH0;
            const double thr = sub == 0 ? 0.5 : hv;
            if (u < thr) hv = u, st = st - sub + (sub == 1 ? 2u : 1u);
            else h_result(sub != 1, in_g);
            return;
        }
        if (st == kB0 || st == kB1) {
            // B: z = U, !(z < y) ends; r = U, !(r < (2k + x) / (2k + 2)) ends; y = z; ...
            const bool first = st == kB0;
            const double thr = first ? y : bthr;
            if (first) z = u;
            if (!(u < thr)) b_result();
            else if (first) st = kB1;
            else {
                y = z;
                if (bn > 4096) b_result();
                else ++bn, st = kB0;
            }
            return;
        }
        switch (st) {
            case kSign:
                // only the sign is taken here; everything that follows from it waits for kJ, so that the divisions of an
                // attempt sit in ONE case body (with a dozen busy lanes most trips find some lane in kSign and some lane in
                // kJ, and a warp pays for every body any of its lanes needs)
                sgn = (w & 1ULL) ? 1 : -1;
                st = kJ;
                break;
            case kJ: {
                const double di0 = sd * (double)k + (double)sgn * mean;
                i0 = (int64_t)ceil(di0);
                x0 = ((double)i0 - di0) / sd;
                // j = w % span through the precomputed floor((2^64 - 1) / span): the estimate never exceeds the quotient
                const uint64_t us = (uint64_t)span;
                uint64_t r = w - mulhi(w, span_m) * us;
                while (r >= us) r -= us;
                j = (int64_t)r;
                x = x0 + (double)j / sd;
                if (!(x < 1.0) || (x == 0.0 && sgn < 0 && k == 0)) restart();
                else h = k + 1, bthr = (2.0 * (double)k + x) / (double)(2 * k + 2), b_check();
                break;
            }
            case kFb1:
                hv = u, st = kFb2;
                break;
            case kFb2: {
                const double r = sqrt(-2.0 * log(hv));
                const double theta = 6.283185307179586476925286766559 * u;
                finish((int64_t)llround(mean + sd * (r * cos(theta))));
                break;
            }
            default:
                break;
        }
    }
};

// integer z ~ D_{Z, stddev, mean}: ONE draw from a stream.  Callers that draw several integers from the same stream
// (gq_sample_digits, p1_sample_column) do not call this once per integer: after the first draw the lanes of a warp
// would sit at different positions of their key-stream blocks, and from then on every step of the following draws
// would find 1/8 of the lanes at a block boundary — a ~1000-instruction ChaCha permutation executed per STEP with
// 4 of 32 lanes active (ncu on the round-1 kernel: 67 % of the warp instructions with 4-7 lanes,
// profiles/r01_ncu_sampler.txt).  They run one loop over the whole task instead, in which every lane takes exactly
// one word per trip and a finished draw hands over to the next one inside the trip: positions stay equal across
// the warp and all lanes refill together.  The word sequence each draw sees is unchanged.
GP_HD int64_t karney_integer(ChaChaStream &g, double mean, double stddev) {
    KarneyMachine km;
    km.begin(mean, stddev);
    while (!km.done()) km.step(g.next());
    return km.result;
}

// ---- one coefficient of gpu_matrix_sample_distribution ----------------------------------
enum { kDistUniform = 0, kDistGauss = 1, kDistBit = 2, kDistTernary = 3 };  // Matrix.cuh:21-27

// The integer behind a Gauss / bit / ternary coefficient is the same in every RNS limb (the
// stream is keyed by (poly, coeff) only), so it is drawn once and reduced per limb.
GP_HD int64_t draw_small_integer(const Seed256 &seed, int dist, double sigma, uint64_t global_poly, uint64_t coeff) {
    ChaChaStream g;
    if (dist == kDistGauss) {
        g.open(seed, global_poly + 1, coeff + 1, 0, kDomGauss);
        return karney_integer(g, 0.0, sigma);
    }
    if (dist == kDistBit) {
        g.open(seed, global_poly + 1, coeff + 1, 0, kDomBit);
        return (int64_t)(g.next() & 1ULL);
    }
    g.open(seed, global_poly + 1, coeff + 1, 0, kDomTernary);
    const uint64_t pick = g.next() % 3ULL;
    return pick == 0 ? 0 : (pick == 1 ? 1 : -1);
}
GP_HD uint64_t draw_uniform_residue(const Seed256 &seed, uint64_t global_poly, uint64_t coeff, uint32_t limb, uint64_t q) {
    ChaChaStream g;
    g.open(seed, global_poly + 1, coeff + 1, (uint64_t)limb + 1, kDomUniform);
    return g.below(q);
}

// ---- G-lattice sampler, one (poly, coeff, tower) --------------------------------------
// Writes the dpt signed digits x with  sum_i x_i b^i == value (mod q)  and x ~ D_{Lambda, c}.
// DMAX bounds the local arrays (8 covers every base >= ceil(crt_bits/8); 64 is the ABI limit).
template <int DMAX> struct GqTower {  // what depends on (q, base, dpt, c) only
    uint32_t base_bits, dpt, last;
    uint64_t base;
    double bf, sigma, kf;
    int64_t qd[DMAX];
    double cv[DMAX];
    KarneyScale first, later;  // deviations of z[last] (sigma / cv[last]) and of the other draws (sigma)
    GP_HD void init(uint64_t q, uint32_t base_bits_, uint32_t dpt_, double c) {
        base_bits = base_bits_, dpt = dpt_, last = dpt_ - 1;
        base = 1ULL << base_bits;
        bf = (double)base;
        sigma = c / (bf + 1.0);
        kf = (double)dpt;
        uint64_t qq = q;
        for (uint32_t i = 0; i < dpt; i++) {
            qd[i] = (int64_t)(qq & (base - 1));
            qq >>= base_bits;
        }
        cv[0] = (double)qd[0] / bf;
        for (uint32_t i = 1; i < dpt; i++) cv[i] = (cv[i - 1] + (double)qd[i]) / bf;
        first.init(sigma / cv[last]);
        later.init(sigma);
    }
    GP_HD int64_t value_digit(uint64_t value, uint32_t i) const { return (int64_t)((value >> (base_bits * i)) & (base - 1)); }
    // the running target a[] of one task: dpt continuous normals (the first 2 dpt words of the stream), the perturbation
    // p = L-factor of the covariance times zf, folded in
    GP_HD void target(ChaChaStream &g, uint64_t value, double *a) const {
        double zf[DMAX];
        for (uint32_t i = 0; i < dpt; i++) zf[i] = sigma * g.normal();
        double prev = 0.0;
        for (uint32_t t = 0; t < dpt; t++) {
            const double lt = t == 0 ? sqrt(bf * (1.0 + 1.0 / kf) + 1.0) : sqrt(bf * (1.0 + 1.0 / (kf - (double)t)));
            double p;
            if (t < last) {
                const double hn = sqrt(bf * (1.0 - 1.0 / (kf - (double)t)));  // h[t+1]
                p = lt * zf[t] + hn * zf[t + 1];
            } else {
                const double hl = t == 0 ? 0.0 : sqrt(bf * (1.0 - 1.0 / (kf - (double)(t - 1))));  // h[last]
                p = hl * zf[t];
            }
            const double vdt = (double)value_digit(value, t);
            a[t] = t == 0 ? (vdt - p) / bf : (prev + vdt - p) / bf;
            prev = a[t];
        }
    }
    // a finished draw of the chain z[last], z[0], .., z[last-1] (`drawn` = draws finished before this one); returns false
    // when it was the last one, otherwise the mean of the next draw
    GP_HD bool chain(uint32_t drawn, int64_t r, double *a, int64_t *z, double &mean) const {
        if (drawn == 0) {
            z[last] = r;
            for (uint32_t i = 0; i < dpt; i++) a[i] += (double)r * cv[i];
        } else {
            z[drawn - 1] = r;
        }
        if (drawn == last) return false;
        mean = -a[drawn];  // deviation: sigma (`later`)
        return true;
    }
    GP_HD void digits(uint64_t value, const int64_t *z, int64_t *out) const {
        const int64_t sb = (int64_t)base;
        if (dpt == 1) {
            out[0] = sb * z[0] + qd[0] * z[0] + value_digit(value, 0);
            return;
        }
        out[0] = sb * z[0] + qd[0] * z[last] + value_digit(value, 0);
        for (uint32_t i = 1; i < last; i++) out[i] = sb * z[i] - z[i - 1] + qd[i] * z[last] + value_digit(value, i);
        out[last] = qd[last] * z[last] - z[last - 1] + value_digit(value, last);
    }
};

template <int DMAX>
GP_HD void gq_sample_digits(const Seed256 &seed, uint64_t value, uint64_t q, uint32_t base_bits, uint32_t dpt, double c,
                            uint32_t tower, uint64_t poly, uint64_t coeff, int64_t *digits_out, uint32_t *words_used = nullptr) {
    GqTower<DMAX> tw;
    tw.init(q, base_bits, dpt, c);
    int64_t z[DMAX];
    double a[DMAX];
    ChaChaStream g;
    g.open(seed, (uint64_t)tower + 1, poly + 1, coeff + 1, kDomGadget);
    tw.target(g, value, a);
    // z[last] ~ D(-a[last]/cv[last], sigma/cv[last]), then a += z[last] cv and z[i] ~ D(-a[i], sigma) for i < last —
    // as one word-per-trip loop (see karney_integer)
    {
        KarneyMachine km;
        uint32_t drawn = 0;
        km.begin(-a[tw.last] / tw.cv[tw.last], tw.first);
        for (;;) {
            bool fin = false;
            while (km.done()) {  // also a draw that needs no word at all (degenerate deviation)
                double mean;
                if (!tw.chain(drawn, km.result, a, z, mean)) {
                    fin = true;
                    break;
                }
                km.begin(mean, tw.later);
                ++drawn;
            }
            if (fin) break;
            km.step(g.next());
        }
    }
    if (words_used) *words_used = g.consumed((uint64_t)tower + 1);  // host-side lane-utilisation model only (src/emu.cu)
    tw.digits(value, z, digits_out);
}

// ---- lane-balanced form: a thread runs K tasks through ONE word-per-trip loop --------------------------------------------
// The number of words a draw consumes is long-tailed (one Gaussian integer: mean 19 words; a G-lattice task with two digits:
// mean 42, 99th percentile 124), so a warp whose lanes own one task each runs for the longest of 32 tasks: 12 of 32 lanes
// in the loop on average (warp model over the word counts of real streams, tests/test_samplers.py).  Here a lane owns K
// tasks and moves to its next one as soon as the current one is done, so the warp runs for the longest SUM of K: 18 of 32
// lanes for K = 4, 21 for K = 8.  What keeps this in lock step:
//   * everything that is not the rejection loop (HChaCha20 key derivation, RNS reduction, stores) runs for all K tasks up
//     front or at the end, with all lanes active;
//   * a lane takes its next task only on a trip at which the whole warp is at a key-stream block boundary (and would first
//     skip the words a task's up-front phase had used), so the position inside the block stays equal across the warp and all
//     lanes run the ChaCha permutation together (cost: 3.5 idle trips per task on average).
// Each task reads exactly the words the one-task form reads, so the samples are identical.
// Measured on B200 (ncu, profiles/r02_sampler_lab.txt): the Gaussian matrix of the C5 preimage (16.4 M draws) 12.2 ms with one
// task per thread, 10.9 ms for K = 4, 8.7 ms for K = 8 (7.5 G instead of 10.1 G warp instructions).  The same driver for the
// G-lattice sampler was built and measured and is NOT kept: its task switches (mean -a/cv, a += z cv, the next draw's set-up)
// are per-lane work inside the loop and K per-task arrays live in local memory — 17.9 ms (K = 8) and 20.8 ms (K = 4) against
// 13.4 ms for gq_sample_digits, with more executed instructions (10.2 G vs 9.2 G) at a lower issue rate.
struct BatchLane {
    ChaChaStream g;
    KarneyMachine km;
    uint32_t cur, skip, drawn;  // task index; words still to skip; finished draws of the current task
    bool active;                // inside a task (false: waiting for a block boundary, or out of tasks)
    GP_HD void start() { cur = 0, skip = 0, drawn = 0, active = false, g.pos = 8; }
    GP_HD bool at_boundary() const { return g.pos >= 8; }
    // point the stream at word `used` of the stream (key, stream0, stream1); only on a block-boundary trip, so that the
    // refill inside the next g.next() produces the block that word lives in
    GP_HD void resume(const uint32_t *key, uint64_t stream0, uint64_t stream1, uint32_t used) {
#pragma unroll
        for (int i = 0; i < 8; i++) g.key[i] = key[i];
        const uint64_t blk = stream0 + (used >> 3);
        g.ctr[0] = (uint32_t)blk, g.ctr[1] = (uint32_t)(blk >> 32);
        g.ctr[2] = (uint32_t)stream1, g.ctr[3] = (uint32_t)(stream1 >> 32);
        skip = used & 7u, drawn = 0, active = true;
    }
};

// K Gaussian integers D(0, sigma) from the streams (poly + 1, coeff_k + 1, 0, kDomGauss), coeff_k = coeff0 + k * stride.
// The coefficient is the nonce of these streams (stream2 = 0), so all K share one HChaCha20 sub-key.
template <int K>
GP_HD void gauss_integers_batch(const Seed256 &seed, double sigma, uint64_t global_poly, uint64_t coeff0, uint64_t stride, int64_t *z) {
    uint32_t key[8];
    {
        ChaChaStream s;
        s.open(seed, global_poly + 1, coeff0 + 1, 0, kDomGauss);
#pragma unroll
        for (int i = 0; i < 8; i++) key[i] = s.key[i];
    }
    KarneyScale sc;
    sc.init(sigma);
    BatchLane ln;
    ln.start();
    for (;;) {
        // transitions that need no stream word: a finished draw ends its task; a lane between tasks takes the next one
        // when the warp is at a block boundary.  Every lane that stays in the loop reaches g.next() on every trip.
        for (;;) {
            if (ln.active) {
                if (!ln.km.done()) break;
                z[ln.cur++] = ln.km.result, ln.active = false;
            } else {
                if (ln.cur >= (uint32_t)K || !ln.at_boundary()) break;
                ln.resume(key, global_poly + 1, coeff0 + (uint64_t)ln.cur * stride + 1, 0);
                ln.km.begin(0.0, sc);
            }
        }
        if (!ln.active && ln.cur >= (uint32_t)K) break;
        const uint64_t w = ln.g.next();
        if (ln.active) ln.km.step(w);  // else: waiting for the boundary, the word is dropped
    }
}

// ---- perturbation p1: per-coefficient 2d x 2d conditional Gaussian ----------------------
// Tables are coefficient-minor ([entry][n]) so that neighbouring threads read neighbouring
// words (the reference keeps them coefficient-major, MatrixTrapdoor.cu:92-93).
//   sd  [m][n]    : conditional standard deviation of row t given rows > t
//   upd [m*m][n]  : upd[t*m + i] = cov(i, t | rows > t) / var(t | rows > t),  i < t
// `cov` is an [m*m][n] scratch.  centred_abd(which, i, j) returns the centred limb-0
// coefficient of A (0), B (1) or D (2) at (i, j).
template <typename LoadABD>
GP_HD void p1_factor_coeff(LoadABD centred_abd, uint32_t d, size_t n, size_t coeff, double sigma, double s, double dgg_stddev,
                           double *cov, double *sd, double *upd) {
    const uint32_t m = 2 * d;
    const double sigma2 = sigma * sigma;
    const double s2 = s * s;
    const double fallback_var = dgg_stddev * dgg_stddev;
    const double eps = 1e-9;
#define GP_COV(i, j) cov[((size_t)(i) * m + (j)) * n + coeff]
#define GP_UPD(t, i) upd[((size_t)(t) * m + (i)) * n + coeff]
    for (uint32_t i = 0; i < d; i++)
        for (uint32_t j = 0; j < d; j++) {
            const double a_ij = (double)centred_abd(0, i, j);
            const double d_ij = (double)centred_abd(2, i, j);
            const double b_ij = (double)centred_abd(1, i, j);
            const double b_ji = (double)centred_abd(1, j, i);
            GP_COV(i, j) = -sigma2 * a_ij + (i == j ? s2 : 0.0);
            GP_COV(i + d, j + d) = -sigma2 * d_ij + (i == j ? s2 : 0.0);
            GP_COV(i, j + d) = -sigma2 * b_ij;
            GP_COV(i + d, j) = -sigma2 * b_ji;
        }
    for (int t = (int)m - 1; t >= 0; --t) {
        double var = GP_COV(t, t);
        if (!(var > eps)) var = fallback_var;
        sd[(size_t)t * n + coeff] = sqrt(var);
        for (int i = 0; i < t; i++) GP_UPD(t, i) = GP_COV(i, t) / var;
        // Schur complement of row/column t
        for (int i = 0; i < t; i++) {
            const double ci = GP_UPD(t, i);
            for (int j = 0; j <= i; j++) {
                const double cj = GP_UPD(t, j) * var;
                const double v = GP_COV(i, j) - ci * cj;
                GP_COV(i, j) = v;
                GP_COV(j, i) = v;
            }
        }
    }
#undef GP_COV
}

// one (column, coefficient) of the perturbation: rows sampled from the last to the first,
// each conditioned on the ones already drawn.  `mean` (m doubles) and `zs` (m integers) are
// caller-provided scratch with element stride `st` (1 for thread-local arrays).
template <typename LoadTp2>
GP_HD void p1_sample_column(LoadTp2 centred_tp2, const Seed256 &seed, uint32_t m, size_t n, size_t col, size_t coeff,
                            double c_scale, const double *sd, const double *upd, double *mean, int64_t *zs, size_t st) {
    ChaChaStream g;
    g.open(seed, (uint64_t)col + 1, (uint64_t)coeff + 1, 0, kDomPerturb);
    for (uint32_t r = 0; r < m; r++) mean[r * st] = c_scale * (double)centred_tp2(r);
    // rows m-1 .. 0 as one word-per-trip loop (see karney_integer)
    if (m == 0) return;
    KarneyMachine km;
    int t = (int)m - 1;
    double mu = mean[(size_t)t * st];
    km.begin(mu, sd[(size_t)t * n + coeff]);
    for (;;) {
        bool fin = false;
        while (km.done()) {
            const int64_t zt = km.result;
            zs[(size_t)t * st] = zt;
            const double delta = (double)zt - mu;
            for (int i = 0; i < t; i++) mean[(size_t)i * st] += GP_UPD(t, i) * delta;
            if (t == 0) {
                fin = true;
                break;
            }
            --t;
            mu = mean[(size_t)t * st];
            km.begin(mu, sd[(size_t)t * n + coeff]);
        }
        if (fin) break;
        km.step(g.next());
    }
#undef GP_UPD
}

// ---- kernels ----------------------------------------------------------------------------
template <typename T> GP_HD T small_to_residue(int64_t z, const LimbConst<T> &c) {
    if (z >= 0) return red64<T>((uint64_t)z, c);
    const T r = red64<T>((uint64_t)(-(z + 1)) + 1, c);
    return r == 0 ? (T)0 : (T)(c.q - r);
}

struct SampleGeom {
    size_t polys;       // rows * local columns
    size_t local_cols;  // columns of `out`
    size_t full_cols;   // columns of the matrix `out` is a slice of (stream indexing)
    size_t col_offset;
    uint32_t n;
    uint32_t limbs;
};

#ifdef __CUDACC__
// Gauss / bit / ternary: one thread per (poly, coeff) draws the integer once and writes every
// limb (the reference re-runs the sampler per limb, MatrixSampling.cu:260-293).
template <typename T>
__global__ void __launch_bounds__(256) sample_small_kernel(T *__restrict__ dst, SampleGeom g, int dist, double sigma, Seed256 seed,
                                                           const LimbConst<T> *__restrict__ lc) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= g.polys * g.n) return;
    const size_t poly = idx / g.n, coeff = idx - poly * g.n;
    const size_t row = poly / g.local_cols, col = poly - row * g.local_cols;
    const int64_t z = draw_small_integer(seed, dist, sigma, row * g.full_cols + g.col_offset + col, coeff);
    T *p = dst + poly * g.limbs * (size_t)g.n + coeff;
    for (uint32_t l = 0; l < g.limbs; l++) p[(size_t)l * g.n] = small_to_residue<T>(z, lc[l]);
}

// uniform: one stream per stored word, storage order
template <typename T>
__global__ void __launch_bounds__(256) sample_uniform_kernel(T *__restrict__ dst, SampleGeom g, Seed256 seed,
                                                             const LimbConst<T> *__restrict__ lc) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= g.polys * g.limbs * g.n) return;
    const size_t pl = idx / g.n, coeff = idx - pl * g.n;
    const size_t poly = pl / g.limbs;
    const uint32_t limb = (uint32_t)(pl - poly * g.limbs);
    const size_t row = poly / g.local_cols, col = poly - row * g.local_cols;
    dst[idx] = (T)draw_uniform_residue(seed, row * g.full_cols + g.col_offset + col, coeff, limb, (uint64_t)lc[limb].q);
}

// G-lattice sampling fused with the scatter: thread (src poly, tower, coeff) reads one residue and
// writes dpt digit coefficients into all limbs of output rows row*ell + tower*dpt + digit
// (replaces sample kernel + int64 staging + scatter kernel per tower, MatrixTrapdoor.cu:701-897).
template <typename T, int DMAX>
__global__ void __launch_bounds__(128) gauss_samp_gq_kernel(const T *__restrict__ src, T *__restrict__ out, size_t src_polys,
                                                            size_t src_cols, uint32_t n, uint32_t limbs, uint32_t base_bits,
                                                            uint32_t dpt, double c, Seed256 seed,
                                                            const LimbConst<T> *__restrict__ lc) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= src_polys * limbs * n) return;
    const size_t pl = idx / n, coeff = idx - pl * n;
    const size_t poly = pl / limbs;
    const uint32_t tower = (uint32_t)(pl - poly * limbs);
    int64_t digits[DMAX];
    gq_sample_digits<DMAX>(seed, (uint64_t)src[idx], (uint64_t)lc[tower].q, base_bits, dpt, c, tower, poly, coeff, digits);
    const size_t row = poly / src_cols, col = poly - row * src_cols;
    const size_t ell = (size_t)dpt * limbs;
    for (uint32_t dg = 0; dg < dpt; dg++) {
        const size_t out_poly = (row * ell + (size_t)tower * dpt + dg) * src_cols + col;
        T *p = out + out_poly * limbs * (size_t)n + coeff;
        for (uint32_t l = 0; l < limbs; l++) p[(size_t)l * n] = small_to_residue<T>(digits[dg], lc[l]);
    }
}

// ---- lane-balanced Gaussian kernel: a warp owns 32 K consecutive coefficients of one polynomial (n % (32 K) == 0); lane l
// runs coefficients c0 + l + 32 k through gauss_integers_batch, so for every k the warp writes 32 consecutive words.  The body
// is GP_HD so the host harness walks the same index arithmetic.
template <typename T, int K>
GP_HD void sample_gauss_batch_thread(size_t tid, T *__restrict__ dst, const SampleGeom &g, double sigma, const Seed256 &seed,
                                     const LimbConst<T> *__restrict__ lc) {
    const size_t base = (tid >> 5) * (size_t)(32 * K);
    if (base >= g.polys * g.n) return;
    const size_t poly = base / g.n, c0 = base - poly * g.n + (tid & 31);
    const size_t row = poly / g.local_cols, col = poly - row * g.local_cols;
    int64_t z[K];
    gauss_integers_batch<K>(seed, sigma, row * g.full_cols + g.col_offset + col, c0, 32, z);
    for (int k = 0; k < K; k++) {
        T *p = dst + poly * g.limbs * (size_t)g.n + c0 + 32 * (size_t)k;
        for (uint32_t l = 0; l < g.limbs; l++) p[(size_t)l * g.n] = small_to_residue<T>(z[k], lc[l]);
    }
}
template <typename T, int K>
__global__ void __launch_bounds__(128) sample_gauss_batch_kernel(T *__restrict__ dst, SampleGeom g, double sigma, Seed256 seed,
                                                                 const LimbConst<T> *__restrict__ lc) {
    sample_gauss_batch_thread<T, K>((size_t)blockIdx.x * blockDim.x + threadIdx.x, dst, g, sigma, seed, lc);
}
// covariance factorisation: one thread per coefficient (limb 0 of A, B, D in coefficient form)
template <typename T>
__global__ void __launch_bounds__(128) p1_factor_kernel(const T *__restrict__ a, const T *__restrict__ b, const T *__restrict__ dm,
                                                        uint32_t d, uint32_t n, uint32_t limbs, uint64_t q0, double sigma, double s,
                                                        double dgg_stddev, double *cov, double *sd, double *upd) {
    const size_t coeff = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (coeff >= n) return;
    auto load = [&](int which, uint32_t i, uint32_t j) -> int64_t {
        const T *base = which == 0 ? a : (which == 1 ? b : dm);
        return centred((uint64_t)base[((size_t)i * d + j) * limbs * (size_t)n + coeff], q0);
    };
    p1_factor_coeff(load, d, (size_t)n, coeff, sigma, s, dgg_stddev, cov, sd, upd);
}

// perturbation sampling fused with the reduction into every limb: thread (column, coeff).
// MLOC > 0: scratch in registers/local memory; MLOC == 0: [m][cols*n] global scratch.
template <typename T, int MLOC>
__global__ void __launch_bounds__(128) p1_sample_kernel(const T *__restrict__ tp2, T *__restrict__ out, uint32_t m, size_t cols,
                                                        uint32_t n, uint32_t limbs, uint64_t q0, double c_scale, Seed256 seed,
                                                        const double *__restrict__ sd, const double *__restrict__ upd,
                                                        double *mean_ws, int64_t *z_ws, const LimbConst<T> *__restrict__ lc) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t total = cols * n;
    if (idx >= total) return;
    const size_t col = idx / n, coeff = idx - col * n;
    auto load = [&](uint32_t r) -> int64_t {
        return centred((uint64_t)tp2[((size_t)r * cols + col) * limbs * (size_t)n + coeff], q0);
    };
    double mean_loc[MLOC > 0 ? MLOC : 1];
    int64_t z_loc[MLOC > 0 ? MLOC : 1];
    double *mean = MLOC > 0 ? mean_loc : mean_ws + idx;
    int64_t *zs = MLOC > 0 ? z_loc : z_ws + idx;
    const size_t st = MLOC > 0 ? 1 : total;
    p1_sample_column(load, seed, m, (size_t)n, col, coeff, c_scale, sd, upd, mean, zs, st);
    for (uint32_t r = 0; r < m; r++) {
        const int64_t z = zs[(size_t)r * st];
        T *p = out + ((size_t)r * cols + col) * limbs * (size_t)n + coeff;
        for (uint32_t l = 0; l < limbs; l++) p[(size_t)l * n] = small_to_residue<T>(z, lc[l]);
    }
}
#endif  // __CUDACC__

}  // namespace gp
