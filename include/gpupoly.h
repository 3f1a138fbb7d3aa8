/*
 * gpupoly.h — C ABI of the B200-native DCRT polynomial-matrix library (libgpupoly).
 *
 * This is the drop-in boundary for MachinaIO/mxx's `gpu` feature: the symbols below
 * are exactly those the reference's Rust side binds in its `unsafe extern "C"` block
 * (src/poly/dcrt/gpu.rs:69-240, plus gpu_device_mem_info from
 * src/bench_estimator/gpu.rs:10-14).  The reference C prototypes they replace live in
 * cuda/include/Runtime.cuh:20-44 and cuda/include/matrix/Matrix{Data,Arith,NTT,
 * Decompose,Serde,Sampling,Trapdoor}.cuh.  Plain pointers and sizes only.
 *
 * Conventions (identical to the reference):
 *   - every int-returning call returns 0 on success, non-zero on error; the message is
 *     in a thread-local string returned by gpu_last_error() (Runtime.cu:16-22, 817-820);
 *     no exception crosses the ABI;
 *   - handles are opaque, created by *_create and freed by the matching *_destroy;
 *   - host RNS layout is [poly row-major][limb][coeff] little-endian u64
 *     (src/matrix/gpu_dcrt_poly.rs:769-817, 1916-1920);
 *   - EVAL layout at the boundary is natural order, slot k = a(psi^(2k+1)),
 *     psi = (least primitive root)^((p-1)/2n)  (Runtime.cu:121-190, MatrixNTT.cu:655-705);
 *   - `level` = active limbs - 1; binary ops require equal ctx and level.
 *
 * The section "extensions" at the end adds entry points the reference does not have
 * (fused mul_decompose, stream access for timing).  They are optional for a drop-in:
 * the reference call sequences work without them.  See INTEGRATION.md.
 */
#ifndef GPUPOLY_H
#define GPUPOLY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct GpuContext GpuContext;               /* gpu.rs:27-31  GpuContextOpaque */
typedef struct GpuMatrix GpuMatrix;                 /* gpu.rs:33-37  GpuMatrixOpaque */
typedef struct GpuEventSet GpuEventSet;             /* gpu.rs:63-67  GpuEventSetOpaque */
typedef struct GpuP1CovarianceCache GpuP1CovarianceCache; /* gpu.rs:39-43 */

/* gpu.rs:45-49: #[repr(C)] struct GpuRngSeed { words: [u64; 4] }, passed by value */
typedef struct GpuRngSeed { uint64_t words[4]; } GpuRngSeed;

/* gpu.rs:242-247 */
enum { GPU_POLY_FORMAT_COEFF = 0, GPU_POLY_FORMAT_EVAL = 1 };
enum { GPU_MATRIX_DIST_UNIFORM = 0, GPU_MATRIX_DIST_GAUSS = 1, GPU_MATRIX_DIST_BIT = 2, GPU_MATRIX_DIST_TERNARY = 3 };

/* ---- context / runtime: replaces cuda/src/Runtime.cu (prototypes Runtime.cuh:20-44) ---- */
/* gpu.rs:70-79; Runtime.cu:527-693.  L = limbs-1, moduli_len must be L+1, every q = 1 mod 2N. */
int gpu_context_create(uint32_t logN, uint32_t L, uint32_t dnum, const uint64_t *moduli, size_t moduli_len,
                       const int *gpu_ids, size_t gpu_ids_len, GpuContext **out_ctx);
void gpu_context_destroy(GpuContext *ctx);                              /* gpu.rs:80 */
int gpu_context_get_N(const GpuContext *ctx, int *out_N);               /* gpu.rs:81 */
int gpu_event_set_wait(GpuEventSet *events);                            /* gpu.rs:83 */
void gpu_event_set_destroy(GpuEventSet *events);                        /* gpu.rs:84 */
int gpu_device_synchronize(void);                                       /* gpu.rs:230 */
int gpu_device_count(int *out_count);                                   /* gpu.rs:231 */
int gpu_device_mem_info(int device, size_t *out_free, size_t *out_total); /* bench_estimator/gpu.rs:10-14 */
int gpu_device_reset(void);                                             /* Runtime.cuh:39 */
const char *gpu_last_error(void);                                       /* gpu.rs:234 */
void *gpu_pinned_alloc(size_t bytes);                                   /* gpu.rs:236 */
void gpu_pinned_free(void *ptr);                                        /* gpu.rs:237 */

/* ---- matrix storage: replaces cuda/src/matrix/MatrixData.cu (MatrixData.cuh) ---- */
int gpu_matrix_create(GpuContext *ctx, int level, size_t rows, size_t cols, int format, GpuMatrix **out); /* gpu.rs:86-93 */
void gpu_matrix_destroy(GpuMatrix *mat);                                /* gpu.rs:94 */
int gpu_matrix_copy(GpuMatrix *dst, const GpuMatrix *src);              /* gpu.rs:95 */
int gpu_matrix_copy_block(GpuMatrix *out, const GpuMatrix *src, size_t dst_row, size_t dst_col,
                          size_t src_row, size_t src_col, size_t rows, size_t cols); /* gpu.rs:159-168 */

/* ---- host <-> device: replaces cuda/src/matrix/MatrixSerde.cu (MatrixSerde.cuh) ---- */
int gpu_matrix_load_rns_batch(GpuMatrix *mat, const uint8_t *bytes, size_t bytes_per_poly, int format,
                              GpuEventSet **out_events);                /* gpu.rs:96-102; MatrixSerde.cu:566-741 */
int gpu_matrix_store_rns_batch(const GpuMatrix *mat, uint8_t *bytes_out, size_t bytes_per_poly, int format,
                               GpuEventSet **out_events);               /* gpu.rs:103-109; MatrixSerde.cu:743-924 */
int gpu_matrix_store_const_coeff_batch(const GpuMatrix *mat, uint64_t *words_out, size_t words_per_poly,
                                       GpuEventSet **out_events);       /* gpu.rs:110-115; MatrixSerde.cu:926-1064 */
int gpu_matrix_store_compact_bytes(GpuMatrix *mat, uint8_t *payload_out, size_t payload_capacity,
                                   uint16_t *out_max_coeff_bits, uint16_t *out_bytes_per_coeff,
                                   size_t *out_payload_len);            /* gpu.rs:116-123 */
int gpu_matrix_load_compact_bytes(GpuMatrix *mat, const uint8_t *payload, size_t payload_len,
                                  uint16_t max_coeff_bits);             /* gpu.rs:124-129 */

/* ---- arithmetic: replaces cuda/src/matrix/MatrixArith.cu (MatrixArith.cuh) ---- */
int gpu_matrix_add(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs);   /* gpu.rs:130-134 */
int gpu_matrix_add_block(GpuMatrix *out, const GpuMatrix *src, size_t dst_row, size_t dst_col,
                         size_t src_row, size_t src_col, size_t rows, size_t cols); /* gpu.rs:135-144 */
int gpu_matrix_sub(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs);   /* gpu.rs:145-149 */
int gpu_matrix_mul(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs);   /* gpu.rs:150-154; MatrixArith.cu:2840-2907 */
int gpu_matrix_equal(const GpuMatrix *lhs, const GpuMatrix *rhs, int *out_equal); /* gpu.rs:155-159 */
int gpu_matrix_mul_scalar(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *scalar); /* gpu.rs:160-164 */

/* ---- transforms: replaces cuda/src/matrix/MatrixNTT.cu (MatrixNTT.cuh) ---- */
int gpu_matrix_ntt_all(GpuMatrix *mat);                                 /* gpu.rs:228; MatrixNTT.cu:787-797 */
int gpu_matrix_intt_all(GpuMatrix *mat);                                /* gpu.rs:229; MatrixNTT.cu:800-811 */

/* ---- gadget: replaces cuda/src/matrix/MatrixDecompose.cu (MatrixDecompose.cuh) ---- */
int gpu_matrix_fill_gadget(GpuMatrix *out, uint32_t base_bits);         /* gpu.rs:169; MatrixDecompose.cu:307-446 */
int gpu_matrix_fill_small_gadget(GpuMatrix *out, uint32_t base_bits);   /* gpu.rs:170 */
int gpu_matrix_fill_small_decomposed_identity_chunk(GpuMatrix *out, const GpuMatrix *scalar_by_digit,
                                                    size_t chunk_idx);  /* gpu.rs:171-175; MatrixDecompose.cu:507-890 */
int gpu_matrix_decompose_base(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out);       /* gpu.rs:176-180; MatrixDecompose.cu:892-1342 */
int gpu_matrix_decompose_base_small(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out); /* gpu.rs:181-185 */

/* ---- samplers / trapdoor: replace cuda/src/matrix/MatrixSampling.cu, MatrixTrapdoor.cu, ChaCha.cu ----
 * Every integer is drawn on the device from ChaCha20 streams keyed by the GpuRngSeed (ChaCha.cu:109-166); for a given
 * seed the result is, residue for residue, the matrix the reference library returns (tests/test_samplers.py).
 * gauss_samp: out is (rows*ell) x cols created in the format wanted back; an EVAL src is INTT'd in place
 *   (MatrixTrapdoor.cu:1739-1746).  sample_distribution(_columns): dist_type = GPU_MATRIX_DIST_*, result EVAL;
 *   the _columns form fills columns [col_offset, col_offset + out->cols) of the full_ncol-wide matrix the seed defines.
 * p1 cache: A, B, D are d x d coefficient-form matrices (limb 0 is read, centred); sample_p1_full_cached: tp2 is
 *   2d x cols in coefficient form, result EVAL (MatrixTrapdoor.cu:2159-2474, 2476-2918). */
int gpu_matrix_gauss_samp_gq_arb_base(GpuMatrix *src, uint32_t base_bits, double c, double dgg_stddev,
                                      GpuRngSeed seed, GpuMatrix *out); /* gpu.rs:186-193 */
int gpu_matrix_create_p1_covariance_cache(const GpuMatrix *a_mat, const GpuMatrix *b_mat, const GpuMatrix *d_mat,
                                          double sigma, double s, double dgg_stddev,
                                          GpuP1CovarianceCache **out_cache); /* gpu.rs:194-202 */
void gpu_matrix_destroy_p1_covariance_cache(GpuP1CovarianceCache *cache);   /* gpu.rs:203 */
int gpu_matrix_sample_p1_full_cached(const GpuP1CovarianceCache *cache, const GpuMatrix *tp2, GpuRngSeed seed,
                                     GpuMatrix *out);                   /* gpu.rs:204-209 */
int gpu_matrix_sample_distribution(GpuMatrix *out, int dist_type, double sigma, GpuRngSeed seed); /* gpu.rs:210-215 */
int gpu_matrix_sample_distribution_columns(GpuMatrix *out, int dist_type, double sigma, GpuRngSeed seed,
                                           size_t full_ncol, size_t col_offset); /* gpu.rs:216-223 */

/* ---- extensions (not in the reference ABI) -------------------------------------------- */
/* out(r x c) = lhs(r x d*ell) * G^-1(rhs(d x c)) in ONE call: the fused replacement for the
 * slice -> decompose_owned -> mul -> copy_block loop of GpuDCRTPolyMatrix::mul_decompose
 * (src/matrix/gpu_dcrt_poly.rs:1414-1493).  lhs EVAL; rhs EVAL or COEFF; out EVAL.
 * small != 0 selects mul_decompose_small (gpu_dcrt_poly.rs:1495-1574). */
int gpu_matrix_mul_decompose(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs, uint32_t base_bits, int small);
/* The same product for a LIMB-SHARDED job (SURVEY.md 8e "Alternative (limbs)"; the reference has only the scaffold,
 * cuda/src/Runtime.cu:254-280, and rejects it at run time, cuda/src/matrix/MatrixNTT.cu:495-498).  out and lhs live in a
 * context whose moduli are this GPU's subset of the CRT basis; other_coeff is the coefficient-form right operand (d x c)
 * with ALL `towers` towers of the basis as an all-gather of the ranks' [poly][towers_per_group][n] blocks leaves it:
 * tower t of polynomial p at (t / towers_per_group) * group_stride_words + p * poly_stride_words + (t % towers_per_group) * n
 * words of this context's storage width.  out(r x out_cols) = lhs(r x d*ell) * G^-1(other[:, col0 : col0 + out_cols]) on this
 * context's limbs; ell follows the FULL basis (tower_moduli). */
int gpu_matrix_mul_decompose_gathered(GpuMatrix *out, const GpuMatrix *lhs, const void *other_coeff, size_t d, size_t c, size_t col0,
                                      const uint64_t *tower_moduli, size_t towers, size_t towers_per_group,
                                      size_t group_stride_words, size_t poly_stride_words, uint32_t base_bits, int small);
/* rows [dst_row, dst_row + src->rows) of dst <- coefficient form of src (src EVAL or COEFF, untouched; dst tagged COEFF): the
 * inverse transform written straight into this rank's block of the all-gather buffer of a limb-sharded mul_decompose. */
int gpu_matrix_intt_to(GpuMatrix *dst, size_t dst_row, const GpuMatrix *src);
/* out[dst block] += src[src block] * scalar (scalar 1x1; all EVAL): the `+ c_y * x` tail of a BGG encoding product
 * (src/bgg/encoding.rs:135) applied in place to the vector rows of a stacked product, one launch. */
int gpu_matrix_add_scaled_block(GpuMatrix *out, const GpuMatrix *src, const GpuMatrix *scalar, size_t dst_row, size_t dst_col,
                                size_t src_row, size_t src_col, size_t rows, size_t cols);
/* gpu_matrix_decompose_base for callers that must not allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments, validation
 * and result as gpu_matrix_decompose_base; the coefficient form of an EVAL input goes to a static device array of
 * GPU_NOALLOC_SCRATCH_BYTES instead of pool memory, and an EVAL input larger than that is rejected with an error.  The call
 * makes no cudaMalloc / cudaFree / cudaMallocAsync.  All calls on a device share the array: calls whose kernels overlap on
 * the GPU (two contexts or streams on one device) race on it; calls on one context are stream-ordered. */
#define GPU_NOALLOC_SCRATCH_BYTES (16u << 20)
int gpu_matrix_decompose_base_noalloc(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out);
/* Big-integer coefficients <-> residues ON THE DEVICE: the conversion step either side of the path
 * (SURVEY 8f-4).  words[(poly * n + j) * words_per_coeff + w], little-endian u64 words of the coefficient.
 * load: every active limb receives the coefficient mod q_limb — the host BigUint `%` loop of
 *   GpuDCRTPolyMatrix::from_cpu_matrix (src/matrix/gpu_dcrt_poly.rs:793-803) and GpuDCRTPoly::from_coeffs
 *   (src/poly/dcrt/gpu.rs:903-920); `format` tags the data (EVAL: natural slot order, as load_rns_batch).
 * store: CRT reconstruction to [0, Q), Q = product of the active moduli — the host loop of
 *   GpuDCRTPoly::coeffs (gpu.rs:959-994); requires coefficient form. */
int gpu_matrix_load_bigint_batch(GpuMatrix *mat, const uint64_t *words, size_t words_per_coeff, int format, GpuEventSet **out_events);
int gpu_matrix_store_bigint_batch(const GpuMatrix *mat, uint64_t *words_out, size_t words_per_coeff, GpuEventSet **out_events);
/* The CUDA stream (cudaStream_t) all work of this context is ordered on; for CUDA-event timing. */
void *gpu_context_stream(const GpuContext *ctx);
/* Stream override for the CALLING THREAD: this library's launches go to `stream` (a cudaStream_t) instead of each context's
 * stream until it is called with NULL.  Ordering against the context stream is the caller's business (events); allocation
 * stays on the context stream, so only operations on existing buffers belong on a side stream. */
int gpu_thread_stream_override(void *stream);
/* Device pointer / word size (4 or 8) / format of a matrix's storage, [poly][limb][coeff]. */
int gpu_matrix_device_info(const GpuMatrix *mat, void **out_ptr, int *out_word_bytes, int *out_format);
/* Number of kernels launched by this library since process start (bench.py's gpu_launches). */
uint64_t gpu_kernel_launch_count(void);
/* CUDA-event stopwatch on the context stream: record(stop=0), work, record(stop=1), elapsed. */
typedef struct GpuTimer GpuTimer;
int gpu_timer_create(const GpuContext *ctx, GpuTimer **out);
int gpu_timer_record(const GpuContext *ctx, GpuTimer *timer, int stop);
int gpu_timer_elapsed_ms(GpuTimer *timer, float *out_ms);
void gpu_timer_destroy(GpuTimer *timer);
/* Waits for the context stream only (gpu_device_synchronize waits for the whole device). */
int gpu_context_synchronize(const GpuContext *ctx);

#ifdef __cplusplus
}
#endif
#endif
