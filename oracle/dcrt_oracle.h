/*
 * dcrt_oracle.h — CPU oracle for the mxx DCRT polynomial-matrix hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked, imported or
 * executed by the product path (mxx_b200/).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may use it, and there only
 * as the checker / reported CPU baseline.
 *
 * It restates, in plain C, the reference semantics of the path:
 *   - CRT basis / roots            cuda/src/Runtime.cu:121-190, 282-319
 *   - negacyclic NTT / INTT        cuda/src/matrix/MatrixNTT.cu:39-206, 655-769
 *   - pointwise matrix product     src/matrix/base/memory.rs:450-479, 589-605 over
 *                                  src/poly/dcrt/poly.rs:389-415
 *   - gadget decomposition         cuda/src/matrix/MatrixDecompose.cu:77-113, 924-946, 1268-1270
 *   - gadget matrix                cuda/src/matrix/MatrixDecompose.cu:235-252, 346-365
 *   - mul_decompose                src/matrix/dcrt_poly.rs:248-255, src/matrix/gpu_dcrt_poly.rs:1414-1493
 *
 * Parity status: the reference's CPU arithmetic lives in OpenFHE (crate `openfhe`
 * 0.3.2 @ MachinaIO/openfhe-rs#9c9d81ce, un-vendored, absent here) and the
 * reference tests hold no golden vectors for this path (test_data/ is empty).
 * The oracle is therefore pinned against (1) the single literal CRT basis in the
 * tree ([131041, 131009] for n=4, 17 bits: src/slot_transfer/bgg_poly_encoding_gpu.rs:1501),
 * (2) the reference's own relational tests, and (3) outputs of the reference's
 * CUDA library itself (oracle/_ref/libgpupoly_ref.so, built from
 * /root/reference/cuda by oracle/Makefile) run on the GPU box, with golden
 * vectors from that run committed under tests/golden/.
 *
 * Host matrix layout everywhere: [poly row-major][limb][coeff] little-endian u64
 * (src/matrix/gpu_dcrt_poly.rs:769-817, 1916-1920).
 */
#ifndef DCRT_ORACLE_H
#define DCRT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- scalar number theory ------------------------------------------------ */
uint64_t orc_mulmod(uint64_t a, uint64_t b, uint64_t q);
uint64_t orc_powmod(uint64_t b, uint64_t e, uint64_t q);
uint64_t orc_invmod(uint64_t a, uint64_t q);                 /* 0 if not invertible */
uint64_t orc_barrett_mulmod_check(uint64_t a, uint64_t b, uint64_t q); /* == orc_mulmod, q < 2^62 */
int orc_is_prime(uint64_t n);
/* least primitive root of prime p, Runtime.cu:121-165 */
uint64_t orc_least_primitive_root(uint64_t p);
/* psi = g^((p-1)/2n), Runtime.cu:167-190; returns 0 if p != 1 mod 2n */
uint64_t orc_psi(uint64_t p, uint64_t n);
/* descending chain of the `depth` largest primes < 2^bits with p = 1 mod 2n.
 * Returns number written (== depth on success). */
int orc_prime_chain(uint64_t n, uint32_t bits, uint32_t depth, uint64_t *out);

/* ---- seeded synthetic inputs --------------------------------------------- */
/* counter-based: value(seed, idx) uniform in [0,q) (SplitMix64 + mul-high). */
uint64_t orc_rand_residue(uint64_t seed, uint64_t idx, uint64_t q);
/* fills [polys][limb][coeff]; idx = (poly*L + limb)*n + coeff */
void orc_fill_uniform(uint64_t *m, size_t polys, uint32_t L, size_t n,
                      const uint64_t *moduli, uint64_t seed);

/* ---- single-limb transforms ---------------------------------------------- */
/* O(n^2) definition: X[k] = sum_i a_i psi^{i(2k+1)} mod q, natural order. */
void orc_ntt_naive(const uint64_t *a, uint64_t *out, size_t n, uint64_t q, uint64_t psi);
/* Step-for-step restatement of the reference pipeline:
 * forward = twist(psi^i) -> bit-reverse -> log2(n) radix-2 DIT stages (omega = psi^2)
 * inverse = bit-reverse -> stages (omega^-1) -> scale n^-1 -> twist(psi^-i)
 * (MatrixNTT.cu:655-769).  In place, natural order in and out. */
void orc_ntt_ref(uint64_t *a, size_t n, uint64_t q, uint64_t psi);
void orc_intt_ref(uint64_t *a, size_t n, uint64_t q, uint64_t psi);
/* O(n^2) negacyclic product c = a*b mod (x^n+1, q). */
void orc_negacyclic_schoolbook(const uint64_t *a, const uint64_t *b, uint64_t *c,
                               size_t n, uint64_t q);

/* ---- matrix-level operations (OpenMP) ------------------------------------- */
typedef struct OrcPlan OrcPlan; /* precomputed tables for (n, moduli) */
OrcPlan *orc_plan_create(size_t n, const uint64_t *moduli, uint32_t L);
void orc_plan_destroy(OrcPlan *p);
int orc_get_threads(void);
void orc_set_threads(int t);

/* fast in-place transforms over [polys][L][n]; natural order both domains;
 * independent algorithm (merged-psi CT/GS + permutation) from orc_ntt_ref. */
void orc_matrix_ntt(const OrcPlan *p, uint64_t *m, size_t polys);
void orc_matrix_intt(const OrcPlan *p, uint64_t *m, size_t polys);

/* out(r x c) = lhs(r x k) * rhs(k x c), pointwise per (limb, slot): both EVAL. */
void orc_matrix_mul(const OrcPlan *p, uint64_t *out, const uint64_t *lhs,
                    const uint64_t *rhs, size_t r, size_t k, size_t c);
void orc_matrix_add(const OrcPlan *p, uint64_t *out, const uint64_t *a,
                    const uint64_t *b, size_t polys);
void orc_matrix_sub(const OrcPlan *p, uint64_t *out, const uint64_t *a,
                    const uint64_t *b, size_t polys);
/* out[i] = a[i] * s (s = one poly, EVAL), pointwise */
void orc_matrix_mul_scalar(const OrcPlan *p, uint64_t *out, const uint64_t *a,
                           const uint64_t *s, size_t polys);

uint32_t orc_crt_bits(const OrcPlan *p);
uint32_t orc_digits_per_tower(const OrcPlan *p, uint32_t base_bits);

/* src (rows x cols) in COEFF form -> out ((rows*ell) x cols) in COEFF form,
 * ell = dpt*L (or dpt when small).  Digit (t,dg) of source limb t is written to
 * every limb of out row row*ell + t*dpt + dg. */
void orc_matrix_decompose(const OrcPlan *p, uint64_t *out, const uint64_t *src,
                          size_t rows, size_t cols, uint32_t base_bits, int small);
/* G (rows x rows*ell), COEFF form (constant polynomials). */
void orc_matrix_gadget(const OrcPlan *p, uint64_t *out, size_t rows,
                       uint32_t base_bits, int small);
/* out(r x c) = S(r x d*ell) * NTT(decompose(INTT(other(d x c)))) — all EVAL at the
 * interface (natural order). */
void orc_matrix_mul_decompose(const OrcPlan *p, uint64_t *out, const uint64_t *S,
                              const uint64_t *other, size_t r, size_t d, size_t c,
                              uint32_t base_bits, int small);

/* ---- samplers of the trapdoor path (sampler_oracle.c) ------------------ */
void orc_chacha20_block(const uint32_t key[8], const uint32_t w12_15[4], uint32_t out[16]); /* RFC 8439 §2.3 */
void orc_hchacha20(const uint32_t key[8], const uint32_t nonce[4], uint32_t out[8]);
/* `count` u64 of the stream cuda/src/ChaCha.cu:109-166 derives from (seed, streams, domain) */
void orc_stream_words(const uint64_t seed[4], uint64_t stream0, uint64_t stream1, uint64_t stream2, uint64_t domain,
                      size_t count, uint64_t *out);
/* sample i drawn from stream (i+1, 1, 0) of the Gauss domain: D_{Z, stddev, mean} */
void orc_karney_samples(const uint64_t seed[4], double mean, double stddev, size_t count, int64_t *out);
/* coefficient-domain result of gpu_matrix_sample_distribution(_columns), MatrixSampling.cu:221-306 */
void orc_sample_distribution(int dist, double sigma, const uint64_t seed[4], size_t rows, size_t local_cols,
                             size_t full_cols, size_t col_offset, size_t n, uint32_t L, const uint64_t *moduli,
                             uint64_t *out);
/* coefficient-domain result of gpu_matrix_gauss_samp_gq_arb_base, MatrixTrapdoor.cu:701-897 */
void orc_gauss_samp_gq(const uint64_t *src, size_t rows, size_t cols, size_t n, uint32_t L, const uint64_t *moduli,
                       uint32_t crt_bits, uint32_t base_bits, double c, const uint64_t seed[4], uint64_t *out);
/* coefficient-domain result of gpu_matrix_sample_p1_full_cached, MatrixTrapdoor.cu:95-277, 674-699 */
void orc_sample_p1(const uint64_t *a, const uint64_t *b, const uint64_t *dm, const uint64_t *tp2, size_t d, size_t cols,
                   size_t n, uint32_t L, const uint64_t *moduli, double sigma, double s, double dgg_stddev,
                   const uint64_t seed[4], uint64_t *out);

#ifdef __cplusplus
}
#endif
#endif
