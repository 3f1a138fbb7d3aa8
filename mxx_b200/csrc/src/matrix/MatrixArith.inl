// MatrixArith.inl — add / sub / add_block / mul / mul_scalar / equal.
// Replaces cuda/src/matrix/MatrixArith.cu (gpu_matrix_add :2632, add_block :2698, sub
// :2774, mul :2840-2907, equal :2909, mul_scalar :2982): every op is ONE launch over all
// limbs on the context stream; nothing is uploaded per call.

namespace {

template <typename T, gp::EwOp OP>
int launch_elementwise(GpuContext *ctx, T *out, const T *a, const T *b, size_t words, uint32_t limbs) {
    if (words == 0) return 0;
    const int V = vec_width<T>(ctx);
    const size_t total = words / V;
    const unsigned threads = 256, grid = stride_grid(ctx, total, threads);
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    const uint32_t logn = (uint32_t)ctx->logN;
    switch (V) {
        case 4: gp::elementwise_kernel<T, 4, OP><<<grid, threads, 0, gp::cur(ctx)>>>(out, a, b, total, limbs, logn, lc); break;
        case 2: gp::elementwise_kernel<T, 2, OP><<<grid, threads, 0, gp::cur(ctx)>>>(out, a, b, total, limbs, logn, lc); break;
        default: gp::elementwise_kernel<T, 1, OP><<<grid, threads, 0, gp::cur(ctx)>>>(out, a, b, total, limbs, logn, lc); break;
    }
    gp::count_launch();
    GP_CHECK_LAUNCH("elementwise_kernel");
    return 0;
}

// grid.y / grid.z hold the column / row tiles and are limited to 65535: wider products go out in slabs
// (the reference accepts any size, cuda/src/matrix/MatrixArith.cu:2886-2904)
constexpr uint32_t kMaxGridYZ = 65535;
template <class Launch> int for_each_matmul_slab(const gp::MatmulGeom &g, int TR, int TC, size_t poly_words, Launch launch) {
    const uint32_t col_slab = kMaxGridYZ * (uint32_t)TC, row_slab = kMaxGridYZ * (uint32_t)TR;
    for (uint32_t r0 = 0; r0 < g.rows; r0 += row_slab)
        for (uint32_t c0 = 0; c0 < g.cols; c0 += col_slab) {
            gp::MatmulGeom s = g;
            s.rows = std::min(row_slab, g.rows - r0);
            s.cols = std::min(col_slab, g.cols - c0);
            s.b_col0 = g.b_col0 + c0;
            s.out_col0 = g.out_col0 + c0;
            if (launch(s, (size_t)r0 * g.inner * poly_words, (size_t)r0 * g.out_cols * poly_words)) return 1;
        }
    return 0;
}

// Launch order of the row-tiled products (TR >= 2): column tiles fastest, so that the CTAs in flight share their slices of
// the left operand (arith.cuh matmul_kernel COLS_FAST).  MXX_MATMUL_COLS_FAST=0 keeps the slot chunks fastest (A/B runs).
bool matmul_cols_fast(unsigned slot_ctas) {
    static const int knob = getenv("MXX_MATMUL_COLS_FAST") ? atoi(getenv("MXX_MATMUL_COLS_FAST")) : 1;
    return knob != 0 && slot_ctas <= kMaxGridYZ;
}

template <typename T, int V, int TR, int TC, int UNROLL = 2, int NT = 128>
int launch_matmul_tile(GpuContext *ctx, T *out, const T *A, const T *B, const gp::MatmulGeom &g) {
    const unsigned threads = NT;
    const unsigned per_limb = (unsigned)(((size_t)ctx->N / V + threads - 1) / threads);
    return for_each_matmul_slab(g, TR, TC, (size_t)g.limbs << g.logn, [&](const gp::MatmulGeom &s, size_t a_off, size_t o_off) {
        const unsigned slot_ctas = s.limbs * per_limb, col_tiles = (s.cols + TC - 1) / TC, row_tiles = (s.rows + TR - 1) / TR;
        const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
        if (TR >= 2 && col_tiles > 1 && matmul_cols_fast(slot_ctas))
            gp::matmul_kernel<T, V, TR, TC, UNROLL, NT, true><<<dim3(col_tiles, slot_ctas, row_tiles), threads, 0, gp::cur(ctx)>>>(out + o_off, A + a_off, B, s, lc);
        else
            gp::matmul_kernel<T, V, TR, TC, UNROLL, NT, false><<<dim3(slot_ctas, col_tiles, row_tiles), threads, 0, gp::cur(ctx)>>>(out + o_off, A + a_off, B, s, lc);
        gp::count_launch();
        GP_CHECK_LAUNCH("matmul_kernel");
        return 0;
    });
}

template <int V, int TR, int TC, int UNROLL = 2, int NT = 128, int MINB = 512 / NT>
int launch_matmul_f64_tile(GpuContext *ctx, uint64_t *out, const uint64_t *A, const uint64_t *B, const gp::MatmulGeom &g) {
    const unsigned threads = NT;
    const unsigned per_limb = (unsigned)(((size_t)ctx->N / V + threads - 1) / threads);
    return for_each_matmul_slab(g, TR, TC, (size_t)g.limbs << g.logn, [&](const gp::MatmulGeom &s, size_t a_off, size_t o_off) {
        const unsigned slot_ctas = s.limbs * per_limb, col_tiles = (s.cols + TC - 1) / TC, row_tiles = (s.rows + TR - 1) / TR;
        const auto *lc = (const gp::LimbConst<gp::F64> *)ctx->d_limb_f64;
        if (TR >= 2 && col_tiles > 1 && matmul_cols_fast(slot_ctas))
            gp::matmul_f64_kernel<V, TR, TC, UNROLL, NT, true, MINB><<<dim3(col_tiles, slot_ctas, row_tiles), threads, 0, gp::cur(ctx)>>>(out + o_off, A + a_off, B, s, lc);
        else
            gp::matmul_f64_kernel<V, TR, TC, UNROLL, NT, false, MINB><<<dim3(slot_ctas, col_tiles, row_tiles), threads, 0, gp::cur(ctx)>>>(out + o_off, A + a_off, B, s, lc);
        gp::count_launch();
        GP_CHECK_LAUNCH("matmul_f64_kernel");
        return 0;
    });
}

}  // namespace

// out[:, out_col0 : out_col0+cols] = A(rows x inner) * B[:, b_col0 : b_col0+cols]
template <typename T> int gpm::run_matmul(GpuContext *ctx, T *out, const T *A, const T *B, gp::MatmulGeom g) {
    if (g.rows == 0 || g.cols == 0 || g.limbs == 0) return 0;
    g.logn = (uint32_t)ctx->logN;
    g.fold = ctx->mac_fold;
    constexpr int VMAX = 16 / (int)sizeof(T);
    const int V = vec_width<T>(ctx);
    if (g.inner == 0) {  // empty sum: the zero matrix
        for (uint32_t r = 0; r < g.rows; r++)
            GP_CUDA(cudaMemsetAsync(out + ((size_t)r * g.out_cols + g.out_col0) * ((size_t)g.limbs << g.logn), 0,
                                    (size_t)g.cols * ((size_t)g.limbs << g.logn) * sizeof(T), gp::cur(ctx)),
                    "cudaMemsetAsync(matmul)");
        return 0;
    }
    if (V != VMAX) return launch_matmul_tile<T, 1, 1, 1>(ctx, out, A, B, g);  // tiny rings (n < 16 bytes of words)
    if constexpr (sizeof(T) == 8) {
        if (ctx->f64_ntt && ctx->f64_matmul) {  // moduli < 2^51: the product on the FP64 pipe (arith.cuh)
            g.fold = ctx->f64_mac_fold;
            if (g.rows >= 4) return launch_matmul_f64_tile<VMAX / 2, 4, 4, 1>(ctx, out, A, B, g);
            // two rows (the C4 gate): five resident CTAs (94 registers) — 0.265 ms against 0.294 ms at four (108 registers) and
            // 0.357 ms at six on the gate's product (tools/matmul_f64_lab.cu, profiles/r02_matmul_rows_lab.txt)
            if (g.rows >= 2) return launch_matmul_f64_tile<VMAX, 2, 4, 1, 128, 5>(ctx, out, A, B, g);
            if (g.cols % 6 == 0) return launch_matmul_f64_tile<VMAX, 1, 6, 2, 64>(ctx, out, A, B, g);
            return launch_matmul_f64_tile<VMAX, 1, 1, 6, 128>(ctx, out, A, B, g);
        }
    }
    // Register tile: the wider the tile the fewer operand loads per product.  rows == 1 (the
    // BGG / bench_matrix_mul shape) streams B once and re-reads A from L2.
    // four rows: for 32-bit words a 4 x 2 tile of full 128-bit vectors (162 registers) — 0.77 ms against 1.75 ms for the 4 x 4
    // tile of half vectors on one column chunk of C3-b (tools/matmul_rows_lab.cu, profiles/r02_matmul_rows_lab.txt)
    if constexpr (sizeof(T) == 4) {
        static const bool wide_rows = !(getenv("MXX_MATMUL_ROWS4") && getenv("MXX_MATMUL_ROWS4")[0] == '0');
        if (g.rows >= 4 && wide_rows) return launch_matmul_tile<T, VMAX, 4, 2>(ctx, out, A, B, g);
    }
    if (g.rows >= 4) return launch_matmul_tile<T, VMAX / 2, 4, 4>(ctx, out, A, B, g);
    if (g.rows >= 2) return launch_matmul_tile<T, VMAX, 2, 4>(ctx, out, A, B, g);
    // one output row: the product is a stream over B.  Tile shapes measured on C1-R (tools/matmul_lab.cu):
    // six columns per thread when the width allows, else one column with a deep unroll.
    if (g.cols % 6 == 0) return launch_matmul_tile<T, VMAX, 1, 6, 2, 64>(ctx, out, A, B, g);
    return launch_matmul_tile<T, VMAX, 1, 1, 6, 128>(ctx, out, A, B, g);
}

template int gpm::run_matmul<uint32_t>(GpuContext *, uint32_t *, const uint32_t *, const uint32_t *, gp::MatmulGeom);
template int gpm::run_matmul<uint64_t>(GpuContext *, uint64_t *, const uint64_t *, const uint64_t *, gp::MatmulGeom);

int gpm::binary_check(const char *name, GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs) {
    if (!out || !lhs || !rhs) return gp::set_error(std::string("invalid ") + name + " arguments");
    if (!same_ctx(out, lhs) || !same_ctx(out, rhs)) return gp::set_error(std::string("context mismatch in ") + name);
    return 0;
}

extern "C" int gpu_matrix_add(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs) {
    if (binary_check("gpu_matrix_add", out, lhs, rhs)) return 1;
    if (!same_shape(lhs, rhs) || !same_shape(out, lhs)) return gp::set_error("size mismatch in gpu_matrix_add");
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const size_t words = out->polys() * out->poly_words();
    int st = GP_DISPATCH_WIDTH(ctx, return (launch_elementwise<T, gp::EwOp::Add>(ctx, (T *)out->data, (const T *)lhs->data,
                                                                                 (const T *)rhs->data, words, (uint32_t)out->limbs())););
    // the reference marks the result EVAL unconditionally (MatrixArith.cu:2694) while its Rust
    // side tracks the domain itself; propagating the operands' domain is a superset of that
    if (!st) out->format = lhs->format;
    return st;
}

extern "C" int gpu_matrix_sub(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs) {
    if (binary_check("gpu_matrix_sub", out, lhs, rhs)) return 1;
    if (!same_shape(lhs, rhs) || !same_shape(out, lhs)) return gp::set_error("size mismatch in gpu_matrix_sub");
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const size_t words = out->polys() * out->poly_words();
    int st = GP_DISPATCH_WIDTH(ctx, return (launch_elementwise<T, gp::EwOp::Sub>(ctx, (T *)out->data, (const T *)lhs->data,
                                                                                 (const T *)rhs->data, words, (uint32_t)out->limbs())););
    if (!st) out->format = lhs->format;
    return st;
}

extern "C" int gpu_matrix_mul_scalar(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *scalar) {
    if (binary_check("gpu_matrix_mul_scalar", out, lhs, scalar)) return 1;
    if (scalar->rows != 1 || scalar->cols != 1) return gp::set_error("scalar must be 1x1 in gpu_matrix_mul_scalar");
    if (!same_shape(out, lhs)) return gp::set_error("size mismatch in gpu_matrix_mul_scalar");
    if (lhs->format != GPU_POLY_FORMAT_EVAL || scalar->format != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("gpu_matrix_mul_scalar requires EVAL format");  // MatrixArith.cu:3011-3014
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const size_t words = out->polys() * out->poly_words();
    int st = GP_DISPATCH_WIDTH(ctx, return (launch_elementwise<T, gp::EwOp::MulScalar>(
                                        ctx, (T *)out->data, (const T *)lhs->data, (const T *)scalar->data, words, (uint32_t)out->limbs())););
    if (!st) out->format = GPU_POLY_FORMAT_EVAL;
    return st;
}

extern "C" int gpu_matrix_add_block(GpuMatrix *out, const GpuMatrix *src, size_t dst_row, size_t dst_col, size_t src_row,
                                    size_t src_col, size_t rows, size_t cols) {
    if (!out || !src) return gp::set_error("invalid gpu_matrix_add_block arguments");
    if (!same_ctx(out, src)) return gp::set_error("context mismatch in gpu_matrix_add_block");
    if (src_row + rows > src->rows || src_col + cols > src->cols || dst_row + rows > out->rows ||
        dst_col + cols > out->cols)
        return gp::set_error("block out of range in gpu_matrix_add_block");
    if (rows == 0 || cols == 0 || out->level < 0) return 0;
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const gp::RectGeom g{out->cols, src->cols, dst_row, dst_col, src_row, src_col, rows, cols};
    return GP_DISPATCH_WIDTH(
        ctx, const int V = vec_width<T>(ctx); const size_t total = rows * cols * out->poly_words() / V;
        const unsigned threads = 256, grid = stride_grid(ctx, total, threads);
        const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb; const uint32_t limbs = (uint32_t)out->limbs(), logn = (uint32_t)ctx->logN;
        T *d = (T *)out->data; const T *s = (const T *)src->data;
        switch (V) {
            case 4: gp::add_rect_kernel<T, 4><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, g, total, limbs, logn, lc); break;
            case 2: gp::add_rect_kernel<T, 2><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, g, total, limbs, logn, lc); break;
            default: gp::add_rect_kernel<T, 1><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, g, total, limbs, logn, lc); break;
        } gp::count_launch();
        GP_CHECK_LAUNCH("add_rect_kernel"); return 0;);
}

// extension: out block += src block * scalar (scalar 1x1, everything EVAL) — one launch instead of mul_scalar + add + the
// slice/concat copies around them (the vector rows of a stacked BGG encoding product, mxx_b200/bgg.py)
extern "C" int gpu_matrix_add_scaled_block(GpuMatrix *out, const GpuMatrix *src, const GpuMatrix *scalar, size_t dst_row, size_t dst_col,
                                           size_t src_row, size_t src_col, size_t rows, size_t cols) {
    if (binary_check("gpu_matrix_add_scaled_block", out, src, scalar)) return 1;
    if (scalar->rows != 1 || scalar->cols != 1) return gp::set_error("scalar must be 1x1 in gpu_matrix_add_scaled_block");
    if (out->format != GPU_POLY_FORMAT_EVAL || src->format != GPU_POLY_FORMAT_EVAL || scalar->format != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("gpu_matrix_add_scaled_block requires EVAL format");
    if (src_row + rows > src->rows || src_col + cols > src->cols || dst_row + rows > out->rows || dst_col + cols > out->cols)
        return gp::set_error("block out of range in gpu_matrix_add_scaled_block");
    if (rows == 0 || cols == 0 || out->level < 0) return 0;
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const gp::RectGeom g{out->cols, src->cols, dst_row, dst_col, src_row, src_col, rows, cols};
    return GP_DISPATCH_WIDTH(
        ctx, const int V = vec_width<T>(ctx); const size_t total = rows * cols * out->poly_words() / V;
        const unsigned threads = 256, grid = stride_grid(ctx, total, threads);
        const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb; const uint32_t limbs = (uint32_t)out->limbs(), logn = (uint32_t)ctx->logN;
        T *d = (T *)out->data; const T *s = (const T *)src->data; const T *z = (const T *)scalar->data;
        switch (V) {
            case 4: gp::fma_rect_kernel<T, 4><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, z, g, total, limbs, logn, lc); break;
            case 2: gp::fma_rect_kernel<T, 2><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, z, g, total, limbs, logn, lc); break;
            default: gp::fma_rect_kernel<T, 1><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, z, g, total, limbs, logn, lc); break;
        } gp::count_launch();
        GP_CHECK_LAUNCH("fma_rect_kernel"); return 0;);
}

extern "C" int gpu_matrix_mul(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs) {
    if (binary_check("gpu_matrix_mul", out, lhs, rhs)) return 1;
    if (lhs->cols != rhs->rows || out->rows != lhs->rows || out->cols != rhs->cols)
        return gp::set_error("size mismatch in gpu_matrix_mul");
    if (lhs->format != GPU_POLY_FORMAT_EVAL || rhs->format != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("gpu_matrix_mul requires EVAL format");  // MatrixArith.cu:2863-2866
    if (out == lhs || out == rhs || (out->data && (out->data == lhs->data || out->data == rhs->data)))
        return gp::set_error("gpu_matrix_mul output must not alias an input");
    GpuContext *ctx = out->ctx;
    out->format = GPU_POLY_FORMAT_EVAL;
    if (out->polys() == 0 || out->level < 0) return 0;
    if (bind_device(ctx)) return 1;
    gp::MatmulGeom g{};
    g.rows = (uint32_t)lhs->rows;
    g.inner = (uint32_t)lhs->cols;
    g.cols = (uint32_t)rhs->cols;
    g.b_cols = (uint32_t)rhs->cols;
    g.b_col0 = 0;
    g.out_cols = (uint32_t)out->cols;
    g.out_col0 = 0;
    g.limbs = (uint32_t)out->limbs();
    return GP_DISPATCH_WIDTH(ctx, return run_matmul<T>(ctx, (T *)out->data, (const T *)lhs->data, (const T *)rhs->data, g););
}

extern "C" int gpu_matrix_equal(const GpuMatrix *lhs, const GpuMatrix *rhs, int *out_equal) {
    if (!lhs || !rhs || !out_equal) return gp::set_error("invalid gpu_matrix_equal arguments");
    *out_equal = 0;
    if (lhs->ctx != rhs->ctx || lhs->level != rhs->level || !same_shape(lhs, rhs)) return 0;
    if (lhs->format != rhs->format) return 0;  // MatrixArith.cu:2934-2937: not equal, no error
    GpuContext *ctx = lhs->ctx;
    const size_t words = lhs->polys() * lhs->poly_words();
    if (words == 0 || lhs->data == rhs->data) {
        *out_equal = 1;
        return 0;
    }
    if (bind_device(ctx)) return 1;
    Scratch flag;
    if (flag.alloc(sizeof(int), ctx->stream)) return 1;
    GP_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int), ctx->stream), "cudaMemsetAsync(equal)");
    int st = GP_DISPATCH_WIDTH(
        ctx, const int V = vec_width<T>(ctx); const size_t total = words / V;
        const unsigned threads = 256, grid = stride_grid(ctx, total, threads);
        const T *a = (const T *)lhs->data; const T *b = (const T *)rhs->data;
        switch (V) {
            case 4: gp::equal_kernel<T, 4><<<grid, threads, 0, gp::cur(ctx)>>>(a, b, total, (int *)flag.p); break;
            case 2: gp::equal_kernel<T, 2><<<grid, threads, 0, gp::cur(ctx)>>>(a, b, total, (int *)flag.p); break;
            default: gp::equal_kernel<T, 1><<<grid, threads, 0, gp::cur(ctx)>>>(a, b, total, (int *)flag.p); break;
        } gp::count_launch();
        GP_CHECK_LAUNCH("equal_kernel"); return 0;);
    if (st) return st;
    int diff = 0;
    GP_CUDA(cudaMemcpyAsync(&diff, flag.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "cudaMemcpyAsync(equal)");
    GP_CUDA(cudaStreamSynchronize(ctx->stream), "cudaStreamSynchronize(equal)");  // the result is a host value
    *out_equal = diff ? 0 : 1;
    return 0;
}
