// arith.cuh — element-wise, block and matrix-product kernels over [poly][limb][coeff].
//
// Replaces cuda/src/matrix/MatrixArith.cu: block_elementwise_all_limbs_kernel (:26),
// block_add_rect_all_limbs_kernel (:137), block_matmul_kernel (:191, one launch per
// limb, stride-bytes_per_poly loads, `__int128 %` per product and per add) and
// block_equal_kernel (:602).  Here every access is a 128-bit vector along the
// coefficient axis, all limbs go in one launch, and products are accumulated lazily
// with one Barrett reduction per `fold` terms.
#pragma once
#include "modarith.cuh"

namespace gp {

template <typename T, int V> struct alignas(sizeof(T) * V) Vec { T v[V]; };
template <typename T, int V> GP_HD Vec<T, V> ldv(const T *p) {
    return *reinterpret_cast<const Vec<T, V> *>(p);
}
template <typename T, int V> GP_HD void stv(T *p, const Vec<T, V> &x) {
    *reinterpret_cast<Vec<T, V> *>(p) = x;
}

// streaming variants (read-once / write-once data: evict-first, so that the operand every CTA re-reads
// — the left matrix of a product — stays resident in L2 while the right matrix streams through)
template <typename T, int V> GP_HD Vec<T, V> ldv_stream(const T *p) {
#ifdef __CUDA_ARCH__
    if constexpr (sizeof(T) * V == 16) {
        const uint4 r = __ldcs(reinterpret_cast<const uint4 *>(p));
        Vec<T, V> out;
        *reinterpret_cast<uint4 *>(&out) = r;
        return out;
    } else
#endif
    {
        return ldv<T, V>(p);
    }
}
template <typename T, int V> GP_HD void stv_stream(T *p, const Vec<T, V> &x) {
#ifdef __CUDA_ARCH__
    if constexpr (sizeof(T) * V == 16) {
        __stcs(reinterpret_cast<uint4 *>(p), *reinterpret_cast<const uint4 *>(&x));
    } else
#endif
    {
        stv<T, V>(p, x);
    }
}

enum class EwOp { Add, Sub, MulScalar };

// out[p] = a[p] (+|-) b[p], or a[p] * s (s one poly).  total = words / V.
template <typename T, int V, EwOp OP>
__global__ void elementwise_kernel(T *__restrict__ out, const T *__restrict__ a, const T *__restrict__ b, size_t total,
                                   uint32_t limbs, uint32_t logn, const LimbConst<T> *__restrict__ lc) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    const size_t poly_words = (size_t)limbs << logn;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total; i += step) {
        const size_t off = i * V;
        const uint32_t limb = (uint32_t)((off >> logn) % limbs);
        const LimbConst<T> c = lc[limb];
        Vec<T, V> x = ldv<T, V>(a + off), y, r;
        if (OP == EwOp::MulScalar)
            y = ldv<T, V>(b + (off % poly_words));
        else
            y = ldv<T, V>(b + off);
#pragma unroll
        for (int k = 0; k < V; k++) {
            if (OP == EwOp::Add) r.v[k] = add_mod<T>(x.v[k], y.v[k], c.q);
            if (OP == EwOp::Sub) r.v[k] = sub_mod<T>(x.v[k], y.v[k], c.q);
            if (OP == EwOp::MulScalar) r.v[k] = mul_mod<T>(x.v[k], y.v[k], c);
        }
        stv<T, V>(out + off, r);
    }
}

struct RectGeom {
    size_t dst_cols, src_cols, dst_row, dst_col, src_row, src_col, rows, cols;
};
// dst block += src block (mod q); total = rows*cols*poly_words / V
template <typename T, int V>
__global__ void add_rect_kernel(T *__restrict__ dst, const T *__restrict__ src, RectGeom g, size_t total, uint32_t limbs,
                                uint32_t logn, const LimbConst<T> *__restrict__ lc) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    const size_t poly_words = (size_t)limbs << logn;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total; i += step) {
        const size_t off = i * V, rp = off / poly_words, w = off % poly_words;
        const size_t r = rp / g.cols, cc = rp % g.cols;
        const uint32_t limb = (uint32_t)(w >> logn);
        const T q = lc[limb].q;
        T *d = dst + ((g.dst_row + r) * g.dst_cols + g.dst_col + cc) * poly_words + w;
        const T *s = src + ((g.src_row + r) * g.src_cols + g.src_col + cc) * poly_words + w;
        Vec<T, V> x = ldv<T, V>(d), y = ldv<T, V>(s);
#pragma unroll
        for (int k = 0; k < V; k++) x.v[k] = add_mod<T>(x.v[k], y.v[k], q);
        stv<T, V>(d, x);
    }
}

// dst block += src block * scalar (one polynomial, EVAL); the tail of a BGG encoding product (src/bgg/encoding.rs:135)
// applied in place to the rows of the stacked product that hold the vector part
template <typename T, int V>
__global__ void fma_rect_kernel(T *__restrict__ dst, const T *__restrict__ src, const T *__restrict__ scalar, RectGeom g, size_t total,
                                uint32_t limbs, uint32_t logn, const LimbConst<T> *__restrict__ lc) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    const size_t poly_words = (size_t)limbs << logn;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total; i += step) {
        const size_t off = i * V, rp = off / poly_words, w = off % poly_words;
        const size_t r = rp / g.cols, cc = rp % g.cols;
        const LimbConst<T> c = lc[(uint32_t)(w >> logn)];
        T *d = dst + ((g.dst_row + r) * g.dst_cols + g.dst_col + cc) * poly_words + w;
        const T *s = src + ((g.src_row + r) * g.src_cols + g.src_col + cc) * poly_words + w;
        Vec<T, V> x = ldv<T, V>(d), y = ldv<T, V>(s), z = ldv<T, V>(scalar + w);
#pragma unroll
        for (int k = 0; k < V; k++) x.v[k] = add_mod<T>(x.v[k], mul_mod<T>(y.v[k], z.v[k], c), c.q);
        stv<T, V>(d, x);
    }
}

// dst block = src block (backs copy, copy_block, slice, concat, entry; replaces
// block_copy_rect_all_limbs_kernel, MatrixArith.cu:88)
template <typename T, int V>
__global__ void copy_rect_kernel(T *__restrict__ dst, const T *__restrict__ src, RectGeom g, size_t total,
                                 size_t poly_words) {
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
        const size_t off = i * V, rp = off / poly_words, w = off % poly_words;
        const size_t r = rp / g.cols, cc = rp % g.cols;
        T *d = dst + ((g.dst_row + r) * g.dst_cols + g.dst_col + cc) * poly_words + w;
        const T *s = src + ((g.src_row + r) * g.src_cols + g.src_col + cc) * poly_words + w;
        stv<T, V>(d, ldv<T, V>(s));
    }
}

// out (size x size) = rows [chunk*size, (chunk+1)*size) of identity(size, scalar).small_decompose():
// global row gr = src_row*digits + digit holds scalar_by_digit[digit] at column src_row, zero elsewhere
// (replaces matrix_fill_small_decomposed_identity_chunk_all_limbs_kernel, MatrixDecompose.cu:455-505).
template <typename T, int V>
__global__ void small_identity_chunk_kernel(T *__restrict__ out, const T *__restrict__ digits_row, size_t size,
                                            size_t chunk_idx, size_t digits, size_t total, size_t poly_words) {
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
        const size_t off = i * V, poly = off / poly_words, w = off % poly_words;
        const size_t lr = poly / size, col = poly % size;
        const size_t gr = chunk_idx * size + lr, src_row = gr / digits, digit = gr % digits;
        Vec<T, V> x;
        if (col == src_row) {
            x = ldv<T, V>(digits_row + digit * poly_words + w);
        } else {
#pragma unroll
            for (int k = 0; k < V; k++) x.v[k] = 0;
        }
        stv<T, V>(out + off, x);
    }
}

template <typename T, int V>
__global__ void equal_kernel(const T *__restrict__ a, const T *__restrict__ b, size_t total, int *__restrict__ diff) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    bool d = false;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total; i += step) {
        Vec<T, V> x = ldv<T, V>(a + i * V), y = ldv<T, V>(b + i * V);
#pragma unroll
        for (int k = 0; k < V; k++) d |= x.v[k] != y.v[k];
    }
    if (__any_sync(0xffffffffu, d) && (GP_TID & 31) == 0) atomicOr(diff, 1);
}

// scalar fill of whole polys: out[poly][limb][*] = vals[poly*limbs + limb] (gadget matrix in EVAL form:
// a constant polynomial evaluates to the constant at every slot).
template <typename T>
__global__ void fill_const_polys_kernel(T *__restrict__ out, const T *__restrict__ vals, size_t total_words, uint32_t logn) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total_words; i += step) out[i] = vals[i >> logn];
}

// Gadget matrix G = I_rows (x) g in EVAL form (a constant polynomial evaluates to the constant at
// every slot): entry (row, row*ell + t*dpt + dg) holds pow[limb][dg] = 2^(bb*dg) mod q_limb in limb t
// only (every limb when small), everything else is zero.  Replaces
// matrix_fill_gadget_multi_limb_kernel + the NTT of a constant (MatrixDecompose.cu:212-252, 431-437).
template <typename T>
__global__ void gadget_fill_kernel(T *__restrict__ out, const T *__restrict__ pow, size_t total_words, uint32_t rows,
                                   uint32_t ell, uint32_t dpt, uint32_t limbs, uint32_t logn, int small) {
    const size_t step = (size_t)gridDim.x * blockDim.x;
    const size_t cols = (size_t)rows * ell;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_words; i += step) {
        const size_t pl = i >> logn, poly = pl / limbs;
        const uint32_t limb = (uint32_t)(pl % limbs);
        const size_t row = poly / cols, col = poly % cols;
        T v = 0;
        if (col / ell == row) {
            const uint32_t loc = (uint32_t)(col % ell), t = loc / dpt, dg = loc % dpt;
            if (small || t == limb) v = pow[limb * dpt + dg];
        }
        out[i] = v;
    }
}

// ---- matrix product ----------------------------------------------------------------
// out(r x c) = A(r x k) * B(k x c), independently per (limb, slot).  A thread owns a
// TR x TC tile of output polynomials at V consecutive slots of one limb.
// grid: x = limbs * ceil(n / (V*GP_NT)), y = ceil(c/TC), z = ceil(r/TR).
// The product lands in columns [out_col0, out_col0 + cols) of an output with out_cols columns.
struct MatmulGeom {
    uint32_t rows, inner, cols;   // product shape (cols = columns computed by this launch)
    uint32_t b_cols, b_col0;      // B has b_cols columns; the product uses [b_col0, b_col0 + cols)
    uint32_t out_cols, out_col0;  // the product lands in columns [out_col0, out_col0 + cols) of out
    uint32_t limbs, logn, fold;
};
template <typename T, int V, int TR, int TC, int UNROLL = 2>
GP_HD void matmul_thread(uint32_t bx, uint32_t by, uint32_t bz, uint32_t tid, uint32_t ntid, T *__restrict__ out,
                         const T *__restrict__ A, const T *__restrict__ B, const MatmulGeom &g,
                         const LimbConst<T> *__restrict__ lc) {
    const uint32_t rows = g.rows, inner = g.inner, cols = g.cols, limbs = g.limbs, logn = g.logn, fold = g.fold;
    const uint32_t chunks = ((1u << logn) / V + ntid - 1) / ntid;
    const uint32_t limb = bx / chunks, chunk = bx % chunks;
    const uint32_t slot = (chunk * ntid + tid) * V;
    if (slot >= (1u << logn)) return;
    const size_t poly_words = (size_t)limbs << logn;
    const size_t in_poly = ((size_t)limb << logn) + slot;
    const uint32_t r0 = bz * TR, c0 = by * TC;
    const LimbConst<T> c = lc[limb];

    Acc<T> acc[TR][TC][V];
#pragma unroll
    for (int i = 0; i < TR; i++)
#pragma unroll
        for (int j = 0; j < TC; j++)
#pragma unroll
            for (int v = 0; v < V; v++) acc[i][j][v].clear();

    // clamp out-of-range tile members onto valid polynomials; their results are not stored
    size_t arow[TR], bcol[TC];
#pragma unroll
    for (int i = 0; i < TR; i++) arow[i] = (size_t)(r0 + i < rows ? r0 + i : rows - 1) * inner * poly_words + in_poly;
#pragma unroll
    for (int j = 0; j < TC; j++) bcol[j] = (size_t)(g.b_col0 + (c0 + j < cols ? c0 + j : cols - 1)) * poly_words + in_poly;
    const size_t bstep = (size_t)g.b_cols * poly_words;

    uint32_t since = 0;
    constexpr int kUnroll = UNROLL;
#pragma unroll kUnroll
    for (uint32_t k = 0; k < inner; k++) {
        Vec<T, V> a[TR], b[TC];
#pragma unroll
        for (int i = 0; i < TR; i++) a[i] = ldv<T, V>(A + arow[i] + (size_t)k * poly_words);
#pragma unroll
        for (int j = 0; j < TC; j++) b[j] = ldv_stream<T, V>(B + bcol[j] + (size_t)k * bstep);
#pragma unroll
        for (int i = 0; i < TR; i++)
#pragma unroll
            for (int j = 0; j < TC; j++)
#pragma unroll
                for (int v = 0; v < V; v++) acc[i][j][v].mac(a[i].v[v], b[j].v[v]);
        if (++since == fold) {
            since = 0;
#pragma unroll
            for (int i = 0; i < TR; i++)
#pragma unroll
                for (int j = 0; j < TC; j++)
#pragma unroll
                    for (int v = 0; v < V; v++) acc[i][j][v].fold(c);
        }
    }
#pragma unroll
    for (int i = 0; i < TR; i++)
#pragma unroll
        for (int j = 0; j < TC; j++) {
            if (r0 + i < rows && c0 + j < cols) {
                Vec<T, V> o;
#pragma unroll
                for (int v = 0; v < V; v++) o.v[v] = acc[i][j][v].result(c);
                stv_stream<T, V>(out + ((size_t)(r0 + i) * g.out_cols + g.out_col0 + c0 + j) * poly_words + in_poly, o);
            }
        }
}
// UNROLL (inner-dimension unroll) and NT (CTA size) are tuned per tile shape on the C1-R product
// (tools/matmul_lab.cu): <1,6> unroll 2 / 64 threads and <1,1> unroll 6 / 128 threads stream the right
// operand at the measured HBM copy bandwidth; see DESIGN.md 4.
// COLS_FAST: the column tiles are the fastest-varying CTA index (grid x = column tiles, y = limb x slot chunk), so the
// CTAs resident at one time share a few (limb, slot chunk) slices of A and every A word is fetched from HBM once.
// With the slot chunks fastest, one sweep over x touches ALL of A before any of it is used again: a left operand
// larger than the L2 (C3-b: 4 x 120 polynomials = 472 MB) is streamed once per column tile.
template <typename T, int V, int TR, int TC, int UNROLL = 2, int NT = 128, bool COLS_FAST = false>
__global__ void __launch_bounds__(NT)
matmul_kernel(T *__restrict__ out, const T *__restrict__ A, const T *__restrict__ B, MatmulGeom g,
              const LimbConst<T> *__restrict__ lc) {
    matmul_thread<T, V, TR, TC, UNROLL>(COLS_FAST ? blockIdx.y : blockIdx.x, COLS_FAST ? blockIdx.x : blockIdx.y, blockIdx.z,
                                        threadIdx.x, blockDim.x, out, A, B, g, lc);
}
// Launch bounds: `__launch_bounds__(NT)` alone.  `__launch_bounds__(NT, 1)` is NOT the same to ptxas: it took the one-row C1-R
// tile from 164 to 214 registers (0.545 -> 0.625 ms) and the wide tiles from 162-168 to 254 (profiles/r02_matmul_rows_lab.txt).

// ---- the same product on the FP64 pipe (u64 words, every modulus below 2^51; modarith.cuh f64_*) --------------
// A 64-bit lazy MAC is a 128-bit multiply-add (measured 0.78e12 MAC/s); on exact doubles a term costs 7 FP64
// operations: b/q is estimated as b * (1/q) (one DMUL per loaded B word, relative error 2^-52, so the rounded quotient
// of a*b is within 1 of the true one and the reduced term |a*b - k q| <= q), then f64_mulmod's exact h/l split.
// The accumulator is an integer-valued double: q/2 + fold*q < 2^53 terms between renormalisations
// (fold = f64_mac_fold_terms, 3 for 51-bit moduli).  Operands are canonical residues (what every kernel stores);
// the result is the canonical residue, bit-identical with matmul_kernel<uint64_t>.
template <int V, int TR, int TC, int UNROLL = 2>
GP_HD void matmul_f64_thread(uint32_t bx, uint32_t by, uint32_t bz, uint32_t tid, uint32_t ntid, uint64_t *__restrict__ out,
                             const uint64_t *__restrict__ A, const uint64_t *__restrict__ B, const MatmulGeom &g,
                             const LimbConst<F64> *__restrict__ lc) {
    typedef uint64_t T;
    const uint32_t rows = g.rows, inner = g.inner, cols = g.cols, limbs = g.limbs, logn = g.logn, fold = g.fold;
    const uint32_t chunks = ((1u << logn) / V + ntid - 1) / ntid;
    const uint32_t limb = bx / chunks, chunk = bx % chunks;
    const uint32_t slot = (chunk * ntid + tid) * V;
    if (slot >= (1u << logn)) return;
    const size_t poly_words = (size_t)limbs << logn;
    const size_t in_poly = ((size_t)limb << logn) + slot;
    const uint32_t r0 = bz * TR, c0 = by * TC;
    const double q = lc[limb].q.v, qinv = lc[limb].q2.v;  // LimbConst<F64>::q2 carries 1/q (host_tables.hpp)

    double acc[TR][TC][V];
#pragma unroll
    for (int i = 0; i < TR; i++)
#pragma unroll
        for (int j = 0; j < TC; j++)
#pragma unroll
            for (int v = 0; v < V; v++) acc[i][j][v] = 0.0;

    size_t arow[TR], bcol[TC];
#pragma unroll
    for (int i = 0; i < TR; i++) arow[i] = (size_t)(r0 + i < rows ? r0 + i : rows - 1) * inner * poly_words + in_poly;
#pragma unroll
    for (int j = 0; j < TC; j++) bcol[j] = (size_t)(g.b_col0 + (c0 + j < cols ? c0 + j : cols - 1)) * poly_words + in_poly;
    const size_t bstep = (size_t)g.b_cols * poly_words;

    uint32_t since = 0;
    constexpr int kUnroll = UNROLL;
#pragma unroll kUnroll
    for (uint32_t k = 0; k < inner; k++) {
        Vec<T, V> a[TR], b[TC];
#pragma unroll
        for (int i = 0; i < TR; i++) a[i] = ldv<T, V>(A + arow[i] + (size_t)k * poly_words);
#pragma unroll
        for (int j = 0; j < TC; j++) b[j] = ldv_stream<T, V>(B + bcol[j] + (size_t)k * bstep);
        double af[TR][V], bf[TC][V], bq[TC][V];
#pragma unroll
        for (int i = 0; i < TR; i++)
#pragma unroll
            for (int v = 0; v < V; v++) af[i][v] = f64_from_word(a[i].v[v]);
#pragma unroll
        for (int j = 0; j < TC; j++)
#pragma unroll
            for (int v = 0; v < V; v++) {
                bf[j][v] = f64_from_word(b[j].v[v]);
                bq[j][v] = f64_mul(bf[j][v], qinv);
            }
#pragma unroll
        for (int i = 0; i < TR; i++)
#pragma unroll
            for (int j = 0; j < TC; j++)
#pragma unroll
                for (int v = 0; v < V; v++) acc[i][j][v] = f64_add(acc[i][j][v], f64_mulmod(af[i][v], bf[j][v], bq[j][v], q));
        if (++since == fold) {
            since = 0;
#pragma unroll
            for (int i = 0; i < TR; i++)
#pragma unroll
                for (int j = 0; j < TC; j++)
#pragma unroll
                    for (int v = 0; v < V; v++) acc[i][j][v] = f64_renorm(acc[i][j][v], q, qinv);
        }
    }
#pragma unroll
    for (int i = 0; i < TR; i++)
#pragma unroll
        for (int j = 0; j < TC; j++) {
            if (r0 + i < rows && c0 + j < cols) {
                Vec<T, V> o;
#pragma unroll
                for (int v = 0; v < V; v++) o.v[v] = f64_canonical(acc[i][j][v], q, qinv);
                stv_stream<T, V>(out + ((size_t)(r0 + i) * g.out_cols + g.out_col0 + c0 + j) * poly_words + in_poly, o);
            }
        }
}
// MINB caps the registers (ptxas took 222 for the 2 x 4 tile uncapped: 8 warps per SM, 3.7 TB/s)
template <int V, int TR, int TC, int UNROLL = 2, int NT = 128, bool COLS_FAST = false, int MINB = 512 / NT>
__global__ void __launch_bounds__(NT, MINB)
matmul_f64_kernel(uint64_t *__restrict__ out, const uint64_t *__restrict__ A, const uint64_t *__restrict__ B, MatmulGeom g,
                  const LimbConst<F64> *__restrict__ lc) {
    matmul_f64_thread<V, TR, TC, UNROLL>(COLS_FAST ? blockIdx.y : blockIdx.x, COLS_FAST ? blockIdx.x : blockIdx.y, blockIdx.z,
                                         threadIdx.x, blockDim.x, out, A, B, g, lc);
}

}  // namespace gp
