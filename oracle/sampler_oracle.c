/*
 * sampler_oracle.c — CPU oracle for the samplers of the trapdoor path.
 *
 * TEST INFRASTRUCTURE ONLY (see dcrt_oracle.h): used by tests/ as the checker, never by
 * the product path.
 *
 * Restates, in plain C:
 *   - ChaCha20 block / HChaCha20 (RFC 8439 §2.3; draft-irtf-cfrg-xchacha §2.2) and the
 *     stream derivation of cuda/src/ChaCha.cu:109-166 (key = HChaCha20(seed, domain||stream2),
 *     block counter = stream0, nonce = stream1, 8 u64 per block)
 *   - uniform (0,1) deviates, rejection-uniform residues       MatrixSampling.cu:6-19, 187-203
 *   - Karney's discrete Gaussian (algorithms H, G, P, B)       MatrixSampling.cu:31-160
 *   - gpu_matrix_sample_distribution(_columns)                 MatrixSampling.cu:221-306, 412-470
 *   - the G-lattice sampler gpu_matrix_gauss_samp_gq_arb_base  MatrixTrapdoor.cu:701-897
 *   - the perturbation p1 (covariance factorisation + conditional sampling)
 *                                                              MatrixTrapdoor.cu:95-277, 674-699
 *
 * Parity status: pinned for everything integer-exact — the ChaCha block against the RFC 8439
 * §2.3.2 known answer, and the distributions (uniform / bit / ternary / Gauss with mean 0,
 * whose floating-point steps are exactly rounded single operations) against outputs of the
 * reference's CUDA library run on a B200 (tests/golden/ref_cuda_sampler_vectors.npz).
 * The G-lattice sampler and p1 evaluate log/cos and fused multiply-adds on the device; this C
 * restatement uses libm and unfused arithmetic, so for those two it is a checker of the
 * RELATIONS the reference tests (G z = v mod q; the conditional-Gaussian statistics), not of
 * individual samples — sample-for-sample equality for them is checked GPU-vs-golden.
 *
 * Matrices: host layout [poly row-major][limb][coeff] u64, coefficient domain.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "dcrt_oracle.h"

#define ROTL32(v, n) (((v) << (n)) | ((v) >> (32 - (n))))
#define QR(a, b, c, d)                 \
    do {                               \
        a += b; d ^= a; d = ROTL32(d, 16); \
        c += d; b ^= c; b = ROTL32(b, 12); \
        a += b; d ^= a; d = ROTL32(d, 8);  \
        c += d; b ^= c; b = ROTL32(b, 7);  \
    } while (0)

static const uint32_t kSigma[4] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u};

static void twenty_rounds(uint32_t x[16]) {
    for (int i = 0; i < 10; i++) {
        QR(x[0], x[4], x[8], x[12]);
        QR(x[1], x[5], x[9], x[13]);
        QR(x[2], x[6], x[10], x[14]);
        QR(x[3], x[7], x[11], x[15]);
        QR(x[0], x[5], x[10], x[15]);
        QR(x[1], x[6], x[11], x[12]);
        QR(x[2], x[7], x[8], x[13]);
        QR(x[3], x[4], x[9], x[14]);
    }
}

/* RFC 8439 §2.3: out = rounds(state) + state, state = sigma | key | words 12..15 */
void orc_chacha20_block(const uint32_t key[8], const uint32_t w12_15[4], uint32_t out[16]) {
    uint32_t st[16], x[16];
    memcpy(st, kSigma, 16);
    memcpy(st + 4, key, 32);
    memcpy(st + 12, w12_15, 16);
    memcpy(x, st, 64);
    twenty_rounds(x);
    for (int i = 0; i < 16; i++) out[i] = x[i] + st[i];
}

/* HChaCha20: rows 0 and 3 of the permuted state, no feed-forward */
void orc_hchacha20(const uint32_t key[8], const uint32_t nonce[4], uint32_t out[8]) {
    uint32_t x[16];
    memcpy(x, kSigma, 16);
    memcpy(x + 4, key, 32);
    memcpy(x + 12, nonce, 16);
    twenty_rounds(x);
    memcpy(out, x, 16);
    memcpy(out + 4, x + 12, 16);
}

typedef struct {
    uint32_t key[8];
    uint32_t w[4];
    uint32_t block[16];
    unsigned idx; /* next u64 of block, 8 = empty */
} Stream;

/* ChaCha.cu:109-150 */
static void stream_init(Stream *s, const uint64_t seed[4], uint64_t stream0, uint64_t stream1, uint64_t stream2,
                        uint64_t domain) {
    uint32_t base_key[8], nonce[4];
    for (int i = 0; i < 4; i++) {
        base_key[2 * i] = (uint32_t)seed[i];
        base_key[2 * i + 1] = (uint32_t)(seed[i] >> 32);
    }
    nonce[0] = (uint32_t)domain;
    nonce[1] = (uint32_t)(domain >> 32);
    nonce[2] = (uint32_t)stream2;
    nonce[3] = (uint32_t)(stream2 >> 32);
    orc_hchacha20(base_key, nonce, s->key);
    s->w[0] = (uint32_t)stream0;
    s->w[1] = (uint32_t)(stream0 >> 32);
    s->w[2] = (uint32_t)stream1;
    s->w[3] = (uint32_t)(stream1 >> 32);
    s->idx = 8;
}

/* ChaCha.cu:152-166 */
static uint64_t stream_u64(Stream *s) {
    if (s->idx >= 8) {
        orc_chacha20_block(s->key, s->w, s->block);
        s->w[0] += 1;
        if (s->w[0] == 0) s->w[1] += 1;
        s->idx = 0;
    }
    uint64_t v = (uint64_t)s->block[2 * s->idx] | ((uint64_t)s->block[2 * s->idx + 1] << 32);
    s->idx++;
    return v;
}

void orc_stream_words(const uint64_t seed[4], uint64_t stream0, uint64_t stream1, uint64_t stream2, uint64_t domain,
                      size_t count, uint64_t *out) {
    Stream s;
    stream_init(&s, seed, stream0, stream1, stream2, domain);
    for (size_t i = 0; i < count; i++) out[i] = stream_u64(&s);
}

/* MatrixSampling.cu:6-19 */
static double uniform_open01(Stream *s) {
    const double scale = 1.0 / 9007199254740992.0;
    double u = (double)(stream_u64(s) >> 11) * scale;
    if (u <= 0.0) u = scale;
    else if (u >= 1.0) u = 1.0 - scale;
    return u;
}

/* MatrixSampling.cu:21-29 */
static double standard_normal(Stream *s) {
    double u1 = uniform_open01(s);
    double u2 = uniform_open01(s);
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925286766559 * u2);
}

/* Karney 2016, algorithm H; MatrixSampling.cu:31-50 */
static int alg_h(Stream *s) {
    double a = uniform_open01(s);
    if (!(a < 0.5)) return 1;
    for (;;) {
        double b = uniform_open01(s);
        if (!(b < a)) return 0;
        a = uniform_open01(s);
        if (!(a < b)) return 1;
    }
}
/* MatrixSampling.cu:52-64 */
static int32_t alg_g(Stream *s) {
    int32_t n = 0;
    while (alg_h(s)) {
        ++n;
        if (n > 1024) break;
    }
    return n;
}
/* MatrixSampling.cu:66-72 */
static int alg_p(Stream *s, int32_t n) {
    while (n-- && alg_h(s)) {
    }
    return n < 0;
}
/* MatrixSampling.cu:74-98 */
static int alg_b(Stream *s, int32_t k, double x) {
    double y = x;
    int32_t n = 0;
    double m = (double)(2 * k + 2);
    for (;; ++n) {
        double z = uniform_open01(s);
        if (!(z < y)) break;
        double r = uniform_open01(s);
        if (!(r < (2.0 * (double)k + x) / m)) break;
        y = z;
        if (n > 4096) break;
    }
    return (n % 2) == 0;
}
/* MatrixSampling.cu:100-160 */
static int64_t integer_karney(Stream *s, double mean, double stddev) {
    if (!(stddev > 0.0) || !isfinite(mean) || !isfinite(stddev)) return (int64_t)llround(mean);
    int64_t ceil_std = (int64_t)ceil(stddev);
    if (ceil_std <= 0) return (int64_t)llround(mean);
    for (int iter = 0; iter < (1 << 16); ++iter) {
        int32_t k = alg_g(s);
        if (!alg_p(s, k * (k - 1))) continue;
        int64_t sg = (stream_u64(s) & 1ULL) ? 1 : -1;
        double di0 = stddev * (double)k + (double)sg * mean;
        int64_t i0 = (int64_t)ceil(di0);
        double x0 = ((double)i0 - di0) / stddev;
        int64_t j = (int64_t)(stream_u64(s) % (uint64_t)ceil_std);
        double x = x0 + (double)j / stddev;
        if (!(x < 1.0) || (x == 0.0 && sg < 0 && k == 0)) continue;
        int32_t h = k + 1;
        while (h-- > 0 && alg_b(s, k, x)) {
        }
        if (h >= 0) continue;
        return sg * (i0 + j);
    }
    return (int64_t)llround(mean + stddev * standard_normal(s));
}

void orc_karney_samples(const uint64_t seed[4], double mean, double stddev, size_t count, int64_t *out) {
    for (size_t i = 0; i < count; i++) {
        Stream s;
        stream_init(&s, seed, i + 1, 1, 0, 0x6f70656e66686532ULL);
        out[i] = integer_karney(&s, mean, stddev);
    }
}

/* MatrixSampling.cu:175-185 */
static uint64_t signed_mod(int64_t v, uint64_t q) {
    if (v >= 0) return (uint64_t)v % q;
    uint64_t mag = (uint64_t)(-(v + 1)) + 1;
    uint64_t r = mag % q;
    return r ? q - r : 0;
}
/* MatrixSampling.cu:205-219 */
static int64_t centered(uint64_t v, uint64_t q) {
    uint64_t r = v % q, half = q >> 1;
    return r <= half ? (int64_t)r : -(int64_t)(q - r);
}
/* MatrixSampling.cu:187-203 */
static uint64_t uniform_mod(Stream *s, uint64_t q) {
    const uint64_t max = ~(uint64_t)0;
    const uint64_t threshold = max - (max % q);
    for (;;) {
        uint64_t x = stream_u64(s);
        if (x < threshold) return x % q;
    }
}

/* gpu_matrix_sample_distribution(_columns) before its final NTT: MatrixSampling.cu:221-306.
 * dist: 0 uniform, 1 gauss, 2 bit, 3 ternary. out: [rows*local_cols][L][n]. */
void orc_sample_distribution(int dist, double sigma, const uint64_t seed[4], size_t rows, size_t local_cols,
                             size_t full_cols, size_t col_offset, size_t n, uint32_t L, const uint64_t *moduli,
                             uint64_t *out) {
#pragma omp parallel for schedule(static)
    for (size_t poly = 0; poly < rows * local_cols; poly++) {
        size_t row = poly / local_cols, col = poly % local_cols;
        uint64_t gpoly = row * full_cols + col_offset + col;
        for (uint32_t l = 0; l < L; l++) {
            uint64_t q = moduli[l];
            uint64_t *dst = out + (poly * L + l) * n;
            for (size_t i = 0; i < n; i++) {
                Stream s;
                if (dist == 0) {
                    stream_init(&s, seed, gpoly + 1, i + 1, (uint64_t)l + 1, 0x6f70656e66686531ULL);
                    dst[i] = uniform_mod(&s, q);
                } else if (dist == 1) {
                    stream_init(&s, seed, gpoly + 1, i + 1, 0, 0x6f70656e66686532ULL);
                    dst[i] = signed_mod(integer_karney(&s, 0.0, sigma), q);
                } else if (dist == 2) {
                    stream_init(&s, seed, gpoly + 1, i + 1, 0, 0x6f70656e66686533ULL);
                    dst[i] = (stream_u64(&s) & 1ULL) % q;
                } else {
                    stream_init(&s, seed, gpoly + 1, i + 1, 0, 0x6f70656e66686534ULL);
                    uint64_t pick = stream_u64(&s) % 3ULL;
                    dst[i] = signed_mod(pick == 0 ? 0 : (pick == 1 ? 1 : -1), q);
                }
            }
        }
    }
}

/* ---- G-lattice sampler: MatrixTrapdoor.cu:701-833 (sample) + 835-897 (scatter) ----------
 * src: [rows*cols][L][n] coefficient domain; out: [(rows*dpt*L)*cols][L][n] coefficient domain,
 * out row = row*dpt*L + tower*dpt + digit.  dpt = ceil(crt_bits / base_bits). */
#define ORC_MAX_DIGITS 64
static void gq_digits(Stream *rng, uint64_t value, uint64_t q, uint32_t bb, uint32_t k, double c, int64_t *out) {
    uint64_t base = (uint64_t)1 << bb;
    double base_f = (double)base, sigma = c / (base_f + 1.0), kf = (double)k;
    int64_t m_digits[ORC_MAX_DIGITS], v_digits[ORC_MAX_DIGITS], z[ORC_MAX_DIGITS];
    double l[ORC_MAX_DIGITS], h[ORC_MAX_DIGITS], c_vec[ORC_MAX_DIGITS], p[ORC_MAX_DIGITS], a[ORC_MAX_DIGITS],
        zf[ORC_MAX_DIGITS];
    uint64_t t = q, v = value % q;
    for (uint32_t i = 0; i < k; i++) {
        m_digits[i] = (int64_t)(t % base);
        t /= base;
        v_digits[i] = (int64_t)(v % base);
        v /= base;
    }
    l[0] = sqrt(base_f * (1.0 + 1.0 / kf) + 1.0);
    for (uint32_t i = 1; i < k; i++) l[i] = sqrt(base_f * (1.0 + 1.0 / (kf - (double)i)));
    h[0] = 0.0;
    for (uint32_t i = 1; i < k; i++) h[i] = sqrt(base_f * (1.0 - 1.0 / (kf - (double)(i - 1))));
    c_vec[0] = (double)m_digits[0] / base_f;
    for (uint32_t i = 1; i < k; i++) c_vec[i] = (c_vec[i - 1] + (double)m_digits[i]) / base_f;
    for (uint32_t i = 0; i < k; i++) zf[i] = sigma * standard_normal(rng);
    for (uint32_t i = 0; i + 1 < k; i++) p[i] = l[i] * zf[i] + h[i + 1] * zf[i + 1];
    p[k - 1] = h[k - 1] * zf[k - 1];
    a[0] = ((double)v_digits[0] - p[0]) / base_f;
    for (uint32_t i = 1; i < k; i++) a[i] = (a[i - 1] + (double)v_digits[i] - p[i]) / base_f;
    uint32_t last = k - 1;
    z[last] = integer_karney(rng, -a[last] / c_vec[last], sigma / c_vec[last]);
    for (uint32_t i = 0; i < k; i++) a[i] += (double)z[last] * c_vec[i];
    for (uint32_t i = 0; i < last; i++) z[i] = integer_karney(rng, -a[i], sigma);
    for (uint32_t i = 0; i < k; i++) {
        int64_t b = (int64_t)base;
        if (k == 1) out[i] = b * z[0] + m_digits[0] * z[0] + v_digits[0];
        else if (i == 0) out[i] = b * z[0] + m_digits[0] * z[last] + v_digits[0];
        else if (i < last) out[i] = b * z[i] - z[i - 1] + m_digits[i] * z[last] + v_digits[i];
        else out[i] = m_digits[last] * z[last] - z[last - 1] + v_digits[last];
    }
}

void orc_gauss_samp_gq(const uint64_t *src, size_t rows, size_t cols, size_t n, uint32_t L, const uint64_t *moduli,
                       uint32_t crt_bits, uint32_t base_bits, double c, const uint64_t seed[4], uint64_t *out) {
    const uint32_t k = (crt_bits + base_bits - 1) / base_bits;
    const size_t ell = (size_t)k * L;
#pragma omp parallel for schedule(static)
    for (size_t poly = 0; poly < rows * cols; poly++) {
        size_t row = poly / cols, col = poly % cols;
        for (uint32_t tower = 0; tower < L; tower++)
            for (size_t i = 0; i < n; i++) {
                Stream rng;
                int64_t digits[ORC_MAX_DIGITS];
                stream_init(&rng, seed, (uint64_t)tower + 1, poly + 1, i + 1, 0x6761646765746731ULL);
                gq_digits(&rng, src[(poly * L + tower) * n + i], moduli[tower], base_bits, k, c, digits);
                for (uint32_t dg = 0; dg < k; dg++) {
                    size_t out_poly = (row * ell + (size_t)tower * k + dg) * cols + col;
                    for (uint32_t l = 0; l < L; l++) out[(out_poly * L + l) * n + i] = signed_mod(digits[dg], moduli[l]);
                }
            }
    }
}

/* ---- perturbation p1 ------------------------------------------------------------------
 * MatrixTrapdoor.cu:95-198 (factorisation, per coefficient, limb 0 of A, B, D centred) and
 * :200-277 (conditional sampling), :674-699 (reduction into every limb).
 * a, b, dm: [d*d][L][n] coefficient domain; tp2: [2d*cols][L][n]; out likewise.
 * `sigma` here is the sampler's c (src/sampler/trapdoor/gpu.rs:145-169). */
void orc_sample_p1(const uint64_t *a, const uint64_t *b, const uint64_t *dm, const uint64_t *tp2, size_t d, size_t cols,
                   size_t n, uint32_t L, const uint64_t *moduli, double sigma, double s, double dgg_stddev,
                   const uint64_t seed[4], uint64_t *out) {
    const size_t m = 2 * d;
    const uint64_t q0 = moduli[0];
    const double sigma2 = sigma * sigma, s2 = s * s, fallback_var = dgg_stddev * dgg_stddev, eps = 1e-9;
    const double c_scale = -(sigma * sigma) / (s * s - sigma * sigma);
#pragma omp parallel
    {
        double *cov = malloc(m * m * sizeof(double)), *upd = calloc(m * m, sizeof(double)), *sd = malloc(m * sizeof(double)),
               *mean = malloc(m * sizeof(double));
        int64_t *zs = malloc(m * sizeof(int64_t));
#pragma omp for schedule(static)
        for (size_t ci = 0; ci < n; ci++) {
#define AT(M, i, j) ((double)centered((M)[(((i) * d + (j)) * L) * n + ci], q0))
            for (size_t i = 0; i < d; i++)
                for (size_t j = 0; j < d; j++) {
                    cov[i * m + j] = -sigma2 * AT(a, i, j) + (i == j ? s2 : 0.0);
                    cov[(i + d) * m + (j + d)] = -sigma2 * AT(dm, i, j) + (i == j ? s2 : 0.0);
                    cov[i * m + (j + d)] = -sigma2 * AT(b, i, j);
                    cov[(i + d) * m + j] = -sigma2 * AT(b, j, i);
                }
#undef AT
            for (long t = (long)m - 1; t >= 0; --t) {
                double var = cov[t * m + t];
                if (!(var > eps)) var = fallback_var;
                sd[t] = sqrt(var);
                for (long i = 0; i < t; i++) upd[t * m + i] = cov[i * m + t] / var;
                for (long i = 0; i < t; i++)
                    for (long j = 0; j <= i; j++) {
                        double u = cov[i * m + j] - upd[t * m + i] * (upd[t * m + j] * var);
                        cov[i * m + j] = u;
                        cov[j * m + i] = u;
                    }
            }
            for (size_t col = 0; col < cols; col++) {
                Stream rng;
                stream_init(&rng, seed, col + 1, ci + 1, 0, 0x7065727475726231ULL);
                for (size_t r = 0; r < m; r++) mean[r] = c_scale * (double)centered(tp2[((r * cols + col) * L) * n + ci], q0);
                for (long t = (long)m - 1; t >= 0; --t) {
                    double mu = mean[t];
                    zs[t] = integer_karney(&rng, mu, sd[t]);
                    double delta = (double)zs[t] - mu;
                    for (long i = 0; i < t; i++) mean[i] += upd[t * m + i] * delta;
                }
                for (size_t r = 0; r < m; r++)
                    for (uint32_t l = 0; l < L; l++) out[((r * cols + col) * L + l) * n + ci] = signed_mod(zs[r], moduli[l]);
            }
        }
        free(cov); free(upd); free(sd); free(mean); free(zs);
    }
}

/* how many u64 words of its stream each Karney sample consumed (sizing of the device-side key-stream prefetch) */
void orc_karney_draw_counts(const uint64_t seed[4], double mean, double stddev, size_t count, uint32_t *out) {
    for (size_t i = 0; i < count; i++) {
        Stream s;
        stream_init(&s, seed, i + 1, 1, 0, 0x6f70656e66686532ULL);
        (void)integer_karney(&s, mean, stddev);
        uint64_t blocks = ((uint64_t)s.w[0] | ((uint64_t)s.w[1] << 32)) - (i + 1);
        out[i] = (uint32_t)((blocks - 1) * 8 + s.idx);
    }
}
