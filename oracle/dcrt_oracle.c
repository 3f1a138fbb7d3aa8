/*
 * dcrt_oracle.c — CPU oracle (TEST INFRASTRUCTURE ONLY; see dcrt_oracle.h).
 *
 * Every function cites the reference file:line whose semantics it restates.
 * Written from the algorithm descriptions, not from the reference sources.
 */
#include "dcrt_oracle.h"

#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------------- */
/* scalar number theory                                                       */
/* ------------------------------------------------------------------------- */

/* (a*b) mod q through a 128-bit product — the reference's modular multiply
 * (cuda/src/matrix/MatrixUtils.cu:541-545, cuda/src/Runtime.cu:98-102). */
uint64_t orc_mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)(((u128)a * b) % q); }

/* square-and-multiply, Runtime.cu:104-119 */
uint64_t orc_powmod(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q, c = b % q;
    while (e) {
        if (e & 1) r = orc_mulmod(r, c, q);
        c = orc_mulmod(c, c, q);
        e >>= 1;
    }
    return r;
}

/* extended Euclid, Runtime.cu:40-74 */
uint64_t orc_invmod(uint64_t a, uint64_t q) {
    __int128 t = 0, nt = 1, r = q, nr = a % q;
    while (nr != 0) {
        __int128 k = r / nr, tmp;
        tmp = t - k * nt; t = nt; nt = tmp;
        tmp = r - k * nr; r = nr; nr = tmp;
    }
    if (r != 1) return 0;
    if (t < 0) t += q;
    return (uint64_t)t;
}

/* deterministic Miller-Rabin for 64-bit integers */
int orc_is_prime(uint64_t n) {
    static const uint64_t small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return 0;
    for (size_t i = 0; i < sizeof small / sizeof *small; i++) {
        if (n == small[i]) return 1;
        if (n % small[i] == 0) return 0;
    }
    uint64_t d = n - 1;
    int s = 0;
    while ((d & 1) == 0) { d >>= 1; s++; }
    for (size_t i = 0; i < sizeof small / sizeof *small; i++) {
        uint64_t x = orc_powmod(small[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) {
            x = orc_mulmod(x, x, n);
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* least primitive root: factor p-1 by trial division, test candidates 2,3,...
 * (Runtime.cu:121-165) */
uint64_t orc_least_primitive_root(uint64_t p) {
    uint64_t phi = p - 1, m = phi, fac[64];
    int nf = 0;
    for (uint64_t d = 2; d * d <= m; d++) {
        if (m % d) continue;
        fac[nf++] = d;
        while (m % d == 0) m /= d;
    }
    if (m > 1) fac[nf++] = m;
    for (uint64_t g = 2; g < p; g++) {
        int ok = 1;
        for (int i = 0; i < nf; i++)
            if (orc_powmod(g, phi / fac[i], p) == 1) { ok = 0; break; }
        if (ok) return g;
    }
    return 0;
}

/* primitive 2n-th root of unity: psi = g^((p-1)/2n) (Runtime.cu:167-190) */
uint64_t orc_psi(uint64_t p, uint64_t n) {
    uint64_t order = 2 * n;
    if (p % order != 1) return 0;
    uint64_t g = orc_least_primitive_root(p);
    uint64_t psi = orc_powmod(g, (p - 1) / order, p);
    if (orc_powmod(psi, n, p) != p - 1) return 0;
    return psi;
}

/* Standalone CRT basis: the moduli are an INPUT to the GPU path
 * (src/poly/dcrt/gpu.rs:559-594); for standalone runs use the largest primes
 * below 2^bits with p = 1 (mod 2n), descending.  Reproduces the one literal basis
 * in the reference tree, [131041, 131009] for (n=4, 17 bits)
 * (src/slot_transfer/bgg_poly_encoding_gpu.rs:1501). */
int orc_prime_chain(uint64_t n, uint32_t bits, uint32_t depth, uint64_t *out) {
    if (bits < 2 || bits > 62) return 0;
    uint64_t step = 2 * n, top = (uint64_t)1 << bits;
    uint64_t c = ((top - 2) / step) * step + 1; /* largest c <= top-1 with c = 1 mod 2n */
    uint32_t k = 0;
    while (k < depth && c > step) {
        if (orc_is_prime(c)) out[k++] = c;
        c -= step;
    }
    return (int)k;
}

/* ------------------------------------------------------------------------- */
/* seeded synthetic inputs                                                    */
/* ------------------------------------------------------------------------- */
static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

uint64_t orc_rand_residue(uint64_t seed, uint64_t idx, uint64_t q) {
    uint64_t x = splitmix64(splitmix64(seed) ^ (idx * 0xD1342543DE82EF95ull + 1));
    return (uint64_t)(((u128)x * q) >> 64);
}

void orc_fill_uniform(uint64_t *m, size_t polys, uint32_t L, size_t n,
                      const uint64_t *moduli, uint64_t seed) {
#pragma omp parallel for schedule(static)
    for (size_t pl = 0; pl < polys * L; pl++) {
        uint64_t q = moduli[pl % L];
        for (size_t i = 0; i < n; i++) m[pl * n + i] = orc_rand_residue(seed, pl * n + i, q);
    }
}

/* ------------------------------------------------------------------------- */
/* single-limb transforms                                                     */
/* ------------------------------------------------------------------------- */
static inline uint32_t ilog2(size_t n) {
    uint32_t l = 0;
    while (((size_t)1 << l) < n) l++;
    return l;
}
static inline size_t bitrev(size_t x, uint32_t bits) {
    size_t r = 0;
    for (uint32_t i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}
static inline uint64_t addmod(uint64_t a, uint64_t b, uint64_t q) {
    uint64_t s = a + b;
    return (s >= q || s < a) ? s - q : s;
}
static inline uint64_t submod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : q - (b - a); }

/* the eval-layout DEFINITION at the boundary: slot k = a(psi^(2k+1)) */
void orc_ntt_naive(const uint64_t *a, uint64_t *out, size_t n, uint64_t q, uint64_t psi) {
    for (size_t k = 0; k < n; k++) {
        uint64_t x = orc_powmod(psi, 2 * k + 1, q), xp = 1, acc = 0;
        for (size_t i = 0; i < n; i++) {
            acc = addmod(acc, orc_mulmod(a[i], xp, q), q);
            xp = orc_mulmod(xp, x, q);
        }
        out[k] = acc;
    }
}

/* radix-2 DIT stages on bit-reversed input with w = wlen^j per butterfly, wlen =
 * omega^(n/len) (MatrixNTT.cu:152-206; tables Runtime.cu:630-639) */
static void ref_stages(uint64_t *a, size_t n, uint64_t q, uint64_t omega) {
    for (size_t len = 2; len <= n; len <<= 1) {
        uint64_t wlen = orc_powmod(omega, n / len, q);
        size_t half = len >> 1;
        for (size_t g = 0; g < n; g += len) {
            uint64_t w = 1;
            for (size_t j = 0; j < half; j++) {
                uint64_t u = a[g + j], v = orc_mulmod(a[g + j + half], w, q);
                a[g + j] = addmod(u, v, q);
                a[g + j + half] = submod(u, v, q);
                w = orc_mulmod(w, wlen, q);
            }
        }
    }
}
static void ref_bitrev(uint64_t *a, size_t n) { /* MatrixNTT.cu:117-150 */
    uint32_t lg = ilog2(n);
    for (size_t i = 0; i < n; i++) {
        size_t r = bitrev(i, lg);
        if (i < r) { uint64_t t = a[i]; a[i] = a[r]; a[r] = t; }
    }
}
static void ref_twist(uint64_t *a, size_t n, uint64_t q, uint64_t base) { /* MatrixNTT.cu:39-77 */
    uint64_t w = 1;
    for (size_t i = 0; i < n; i++) { a[i] = orc_mulmod(a[i], w, q); w = orc_mulmod(w, base, q); }
}

void orc_ntt_ref(uint64_t *a, size_t n, uint64_t q, uint64_t psi) { /* MatrixNTT.cu:655-705 */
    ref_twist(a, n, q, psi);
    ref_bitrev(a, n);
    ref_stages(a, n, q, orc_mulmod(psi, psi, q));
}
void orc_intt_ref(uint64_t *a, size_t n, uint64_t q, uint64_t psi) { /* MatrixNTT.cu:706-769 */
    uint64_t ipsi = orc_invmod(psi, q);
    ref_bitrev(a, n);
    ref_stages(a, n, q, orc_mulmod(ipsi, ipsi, q));
    uint64_t ninv = orc_invmod(n % q, q);
    for (size_t i = 0; i < n; i++) a[i] = orc_mulmod(a[i], ninv, q); /* MatrixNTT.cu:79-115 */
    ref_twist(a, n, q, ipsi);
}

void orc_negacyclic_schoolbook(const uint64_t *a, const uint64_t *b, uint64_t *c, size_t n, uint64_t q) {
    memset(c, 0, n * sizeof *c);
    for (size_t i = 0; i < n; i++)
        for (size_t j = 0; j < n; j++) {
            uint64_t p = orc_mulmod(a[i], b[j], q);
            size_t k = i + j;
            if (k < n) c[k] = addmod(c[k], p, q);
            else c[k - n] = submod(c[k - n], p, q);
        }
}

/* ------------------------------------------------------------------------- */
/* plan: per-limb tables for the fast transforms                              */
/* ------------------------------------------------------------------------- */
struct OrcPlan {
    size_t n;
    uint32_t logn, L, crt_bits;
    uint64_t *q;        /* [L] */
    uint64_t *mu;       /* [L] Barrett floor(2^(2k)/q), k = bit width of q */
    uint32_t *kbits;    /* [L] */
    uint64_t *psi;      /* [L] */
    uint64_t *ninv;     /* [L] */
    uint64_t *fw;       /* [L][n]   psi^bitrev(i) */
    uint64_t *fws;      /* Shoup companions floor(w*2^64/q) */
    uint64_t *iw;       /* [L][n]   psi^-bitrev(i) */
    uint64_t *iws;
    uint32_t *rev;      /* [n] */
};

static inline uint64_t shoup(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / q); }
/* w*a mod q with precomputed ws = floor(w*2^64/q); result in [0,q) */
static inline uint64_t mul_shoup(uint64_t a, uint64_t w, uint64_t ws, uint64_t q) {
    uint64_t h = (uint64_t)(((u128)a * ws) >> 64);
    uint64_t r = a * w - h * q;
    return r >= q ? r - q : r;
}

OrcPlan *orc_plan_create(size_t n, const uint64_t *moduli, uint32_t L) {
    OrcPlan *p = (OrcPlan *)calloc(1, sizeof *p);
    p->n = n; p->logn = ilog2(n); p->L = L;
    p->mu = (uint64_t *)malloc(L * 8); p->kbits = (uint32_t *)malloc(L * 4);
    p->q = (uint64_t *)malloc(L * 8); p->psi = (uint64_t *)malloc(L * 8); p->ninv = (uint64_t *)malloc(L * 8);
    p->fw = (uint64_t *)malloc(L * n * 8); p->fws = (uint64_t *)malloc(L * n * 8);
    p->iw = (uint64_t *)malloc(L * n * 8); p->iws = (uint64_t *)malloc(L * n * 8);
    p->rev = (uint32_t *)malloc(n * 4);
    for (size_t i = 0; i < n; i++) p->rev[i] = (uint32_t)bitrev(i, p->logn);
    p->crt_bits = 0;
    for (uint32_t l = 0; l < L; l++) {
        uint64_t q = moduli[l];
        p->q[l] = q;
        uint32_t bits = 64 - (uint32_t)__builtin_clzll(q);
        if (bits > p->crt_bits) p->crt_bits = bits;
        p->kbits[l] = bits;
        p->mu[l] = (uint64_t)((((u128)1) << (2 * bits)) / q); /* bits <= 62 */
        uint64_t psi = orc_psi(q, n);
        if (!psi) { orc_plan_destroy(p); return NULL; }
        p->psi[l] = psi;
        p->ninv[l] = orc_invmod(n % q, q);
        uint64_t ipsi = orc_invmod(psi, q), w = 1, iw = 1;
        for (size_t i = 0; i < n; i++) { /* power i stored at bitrev(i) */
            size_t r = p->rev[i];
            p->fw[l * n + r] = w; p->fws[l * n + r] = shoup(w, q);
            p->iw[l * n + r] = iw; p->iws[l * n + r] = shoup(iw, q);
            w = orc_mulmod(w, psi, q); iw = orc_mulmod(iw, ipsi, q);
        }
    }
    return p;
}
void orc_plan_destroy(OrcPlan *p) {
    if (!p) return;
    free(p->mu); free(p->kbits);
    free(p->q); free(p->psi); free(p->ninv); free(p->fw); free(p->fws); free(p->iw); free(p->iws); free(p->rev);
    free(p);
}
int orc_get_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_threads(int t) {
#ifdef _OPENMP
    if (t > 0) omp_set_num_threads(t);
#else
    (void)t;
#endif
}
uint32_t orc_crt_bits(const OrcPlan *p) { return p->crt_bits; }
uint32_t orc_digits_per_tower(const OrcPlan *p, uint32_t bb) { return (p->crt_bits + bb - 1) / bb; }

/* Cooley-Tukey, merged psi powers, natural in -> bit-reversed out; then permuted
 * to the natural order the boundary defines.  An algorithm independent of
 * orc_ntt_ref (tests require the two and orc_ntt_naive to agree). */
static void fast_ntt(const OrcPlan *p, uint32_t l, uint64_t *a, uint64_t *tmp) {
    size_t n = p->n, t = n;
    uint64_t q = p->q[l];
    const uint64_t *w = p->fw + l * n, *ws = p->fws + l * n;
    for (size_t m = 1; m < n; m <<= 1) {
        t >>= 1;
        for (size_t i = 0; i < m; i++) {
            uint64_t s = w[m + i], ss = ws[m + i];
            uint64_t *x = a + 2 * i * t, *y = x + t;
            for (size_t j = 0; j < t; j++) {
                uint64_t u = x[j], v = mul_shoup(y[j], s, ss, q);
                x[j] = addmod(u, v, q);
                y[j] = submod(u, v, q);
            }
        }
    }
    for (size_t i = 0; i < n; i++) tmp[p->rev[i]] = a[i];
    memcpy(a, tmp, n * 8);
}
/* Gentleman-Sande inverse: natural -> (permute to bit-reversed) -> natural coefficients */
static void fast_intt(const OrcPlan *p, uint32_t l, uint64_t *a, uint64_t *tmp) {
    size_t n = p->n, t = 1;
    uint64_t q = p->q[l];
    const uint64_t *w = p->iw + l * n, *ws = p->iws + l * n;
    for (size_t i = 0; i < n; i++) tmp[i] = a[p->rev[i]];
    memcpy(a, tmp, n * 8);
    for (size_t m = n; m > 1; m >>= 1) {
        size_t h = m >> 1;
        for (size_t i = 0; i < h; i++) {
            uint64_t s = w[h + i], ss = ws[h + i];
            uint64_t *x = a + 2 * i * t, *y = x + t;
            for (size_t j = 0; j < t; j++) {
                uint64_t u = x[j], v = y[j];
                x[j] = addmod(u, v, q);
                y[j] = mul_shoup(submod(u, v, q), s, ss, q);
            }
        }
        t <<= 1;
    }
    uint64_t ni = p->ninv[l], nis = shoup(ni, q);
    for (size_t i = 0; i < n; i++) a[i] = mul_shoup(a[i], ni, nis, q);
}

void orc_matrix_ntt(const OrcPlan *p, uint64_t *m, size_t polys) {
#pragma omp parallel
    {
        uint64_t *tmp = (uint64_t *)malloc(p->n * 8);
#pragma omp for schedule(static)
        for (size_t pl = 0; pl < polys * p->L; pl++) fast_ntt(p, (uint32_t)(pl % p->L), m + pl * p->n, tmp);
        free(tmp);
    }
}
void orc_matrix_intt(const OrcPlan *p, uint64_t *m, size_t polys) {
#pragma omp parallel
    {
        uint64_t *tmp = (uint64_t *)malloc(p->n * 8);
#pragma omp for schedule(static)
        for (size_t pl = 0; pl < polys * p->L; pl++) fast_intt(p, (uint32_t)(pl % p->L), m + pl * p->n, tmp);
        free(tmp);
    }
}

/* ------------------------------------------------------------------------- */
/* matrix arithmetic                                                          */
/* ------------------------------------------------------------------------- */

/* out[i][j] = sum_k lhs[i][k] * rhs[k][j]: parallel over output entries (i,j),
 * serial over k, one modular multiply + one modular add per term
 * (src/matrix/base/memory.rs:589-605: rayon over (i,j), serial k;
 *  DCRTPolyMul/DCRTPolyAdd per term: src/poly/dcrt/poly.rs:389-395).  Operands are
 * in EVALUATION format, so a ring product is a pointwise product per limb. */
/* Barrett reduction of the 128-bit product (q < 2^62): same value as `% q`. */
static inline uint64_t barrett_mulmod(uint64_t a, uint64_t b, uint64_t q, uint64_t mu, uint32_t k) {
    u128 x = (u128)a * b;
    uint64_t t = (uint64_t)(x >> (k - 1));
    uint64_t qh = (uint64_t)(((u128)t * mu) >> (k + 1));
    uint64_t r = (uint64_t)x - qh * q;
    if (r >= q) r -= q;
    if (r >= q) r -= q;
    return r;
}
uint64_t orc_barrett_mulmod_check(uint64_t a, uint64_t b, uint64_t q) {
    uint32_t k = 64 - (uint32_t)__builtin_clzll(q);
    uint64_t mu = (uint64_t)((((u128)1) << (2 * k)) / q);
    return barrett_mulmod(a, b, q, mu, k);
}

void orc_matrix_mul(const OrcPlan *p, uint64_t *out, const uint64_t *lhs, const uint64_t *rhs,
                    size_t r, size_t k, size_t c) {
    size_t n = p->n, L = p->L, ps = L * n;
#pragma omp parallel for collapse(2) schedule(dynamic)
    for (size_t i = 0; i < r; i++)
        for (size_t j = 0; j < c; j++) {
            uint64_t *o = out + (i * c + j) * ps;
            memset(o, 0, ps * 8);
            for (size_t kk = 0; kk < k; kk++) {
                const uint64_t *a = lhs + (i * k + kk) * ps, *b = rhs + (kk * c + j) * ps;
                for (size_t l = 0; l < L; l++) {
                    uint64_t q = p->q[l], mu = p->mu[l];
                    uint32_t kb = p->kbits[l];
                    const uint64_t *al = a + l * n, *bl = b + l * n;
                    uint64_t *ol = o + l * n;
                    for (size_t s = 0; s < n; s++) ol[s] = addmod(ol[s], barrett_mulmod(al[s], bl[s], q, mu, kb), q);
                }
            }
        }
}

void orc_matrix_add(const OrcPlan *p, uint64_t *out, const uint64_t *a, const uint64_t *b, size_t polys) {
    size_t n = p->n;
#pragma omp parallel for schedule(static)
    for (size_t pl = 0; pl < polys * p->L; pl++) {
        uint64_t q = p->q[pl % p->L];
        for (size_t s = 0; s < n; s++) out[pl * n + s] = addmod(a[pl * n + s], b[pl * n + s], q);
    }
}
void orc_matrix_sub(const OrcPlan *p, uint64_t *out, const uint64_t *a, const uint64_t *b, size_t polys) {
    size_t n = p->n;
#pragma omp parallel for schedule(static)
    for (size_t pl = 0; pl < polys * p->L; pl++) {
        uint64_t q = p->q[pl % p->L];
        for (size_t s = 0; s < n; s++) out[pl * n + s] = submod(a[pl * n + s], b[pl * n + s], q);
    }
}
void orc_matrix_mul_scalar(const OrcPlan *p, uint64_t *out, const uint64_t *a, const uint64_t *sc, size_t polys) {
    size_t n = p->n, L = p->L;
#pragma omp parallel for schedule(static)
    for (size_t pl = 0; pl < polys * L; pl++) {
        size_t l = pl % L;
        uint64_t q = p->q[l];
        for (size_t s = 0; s < n; s++) out[pl * n + s] = barrett_mulmod(a[pl * n + s], sc[l * n + s], q, p->mu[l], p->kbits[l]);
    }
}

/* ------------------------------------------------------------------------- */
/* gadget decomposition                                                        */
/* ------------------------------------------------------------------------- */

/* For source limb t and digit dg:
 *   digit = (x_t >> dg*bb) & mask(min(bb, bits(q_t) - dg*bb))   (0 if dg*bb >= bits(q_t))
 * reduced mod q_l if needed and stored in EVERY limb l of output row
 * row*ell + t*dpt + dg  (MatrixDecompose.cu:77-113; src_bits = bit width of q_t,
 * :1268; dpt from the widest modulus, :924-946; small: only limb 0, :1069, 1270).
 * CPU twin: OpenFHE Decompose + last-digit mask (src/matrix/dcrt_poly.rs:453-493,
 * src/poly/dcrt/params.rs:81-89). */
void orc_matrix_decompose(const OrcPlan *p, uint64_t *out, const uint64_t *src, size_t rows, size_t cols,
                          uint32_t bb, int small) {
    size_t n = p->n, L = p->L, ps = L * n;
    uint32_t dpt = orc_digits_per_tower(p, bb);
    size_t ell = small ? dpt : (size_t)dpt * L;
    size_t tl = small ? 1 : L;
#pragma omp parallel for collapse(2) schedule(static)
    for (size_t row = 0; row < rows; row++)
        for (size_t col = 0; col < cols; col++) {
            const uint64_t *sp = src + (row * cols + col) * ps;
            for (size_t t = 0; t < tl; t++) {
                uint32_t sbits = 64 - (uint32_t)__builtin_clzll(p->q[t]);
                for (uint32_t dg = 0; dg < dpt; dg++) {
                    size_t orow = row * ell + t * dpt + dg;
                    uint64_t *op = out + (orow * cols + col) * ps;
                    uint32_t shift = dg * bb;
                    uint64_t mask = 0;
                    if (shift < sbits) {
                        uint32_t rem = sbits - shift, db = bb < rem ? bb : rem;
                        mask = db >= 64 ? ~0ull : (((uint64_t)1 << db) - 1);
                    }
                    for (size_t s = 0; s < n; s++) {
                        uint64_t d = shift >= 64 ? 0 : ((sp[t * n + s] >> shift) & mask);
                        for (size_t l = 0; l < L; l++) op[l * n + s] = d >= p->q[l] ? d % p->q[l] : d;
                    }
                }
            }
        }
}

/* G = I_rows (x) g: entry (row, row*ell + t*dpt + dg) is the constant polynomial
 * whose limb t is 2^(bb*dg) mod q_t and whose other limbs are 0; small variant:
 * every limb holds 2^(bb*dg) mod q_l, ell = dpt
 * (MatrixDecompose.cu:212-252, 346-365, 411). COEFF form. */
void orc_matrix_gadget(const OrcPlan *p, uint64_t *out, size_t rows, uint32_t bb, int small) {
    size_t n = p->n, L = p->L, ps = L * n;
    uint32_t dpt = orc_digits_per_tower(p, bb);
    size_t ell = small ? dpt : (size_t)dpt * L, cols = rows * ell;
    memset(out, 0, rows * cols * ps * 8);
    for (size_t row = 0; row < rows; row++)
        for (size_t loc = 0; loc < ell; loc++) {
            size_t tower = loc / dpt;
            uint32_t dg = (uint32_t)(loc % dpt);
            uint64_t *op = out + (row * cols + row * ell + loc) * ps;
            for (size_t l = 0; l < L; l++)
                if (small || l == tower) op[l * n] = orc_powmod((uint64_t)1 << bb, dg, p->q[l]);
        }
}

/* S * G^-1(other): INTT(other) -> digits -> NTT -> matrix product
 * (src/matrix/gpu_dcrt_poly.rs:1414-1493; CPU twin src/matrix/dcrt_poly.rs:248-255) */
void orc_matrix_mul_decompose(const OrcPlan *p, uint64_t *out, const uint64_t *S, const uint64_t *other,
                              size_t r, size_t d, size_t c, uint32_t bb, int small) {
    size_t ps = p->L * p->n;
    uint32_t dpt = orc_digits_per_tower(p, bb);
    size_t ell = small ? dpt : (size_t)dpt * p->L, m = d * ell;
    uint64_t *oc = (uint64_t *)malloc(d * c * ps * 8);
    uint64_t *dec = (uint64_t *)malloc(m * c * ps * 8);
    memcpy(oc, other, d * c * ps * 8);
    orc_matrix_intt(p, oc, d * c);
    orc_matrix_decompose(p, dec, oc, d, c, bb, small);
    orc_matrix_ntt(p, dec, m * c);
    orc_matrix_mul(p, out, S, dec, r, m, c);
    free(oc);
    free(dec);
}
