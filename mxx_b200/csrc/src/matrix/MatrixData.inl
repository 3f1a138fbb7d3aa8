// MatrixData.inl — matrix storage: create / destroy / copy / copy_block.
// Replaces cuda/src/matrix/MatrixData.cu (gpu_matrix_create :191-461, gpu_matrix_copy
// :473-488, gpu_matrix_copy_block :490-560).  One stream-ordered allocation per matrix
// from the device's default memory pool (no per-limb streams/events, no aux pointer
// workspace, no 2x over-allocation).

namespace {
std::once_flag g_pool_once[16];

// keep freed blocks cached in the pool instead of returning them to the driver at
// every synchronisation (the reference exposes the same knob through
// MXX_CUDA_MEMPOOL_RELEASE_THRESHOLD_BYTES, Runtime.cu:337-361)
void tune_pool(int device) {
    if (device < 0 || device >= 16) return;
    std::call_once(g_pool_once[device], [device]() {
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, device) != cudaSuccess) return;
        uint64_t threshold = ~0ull;
        if (const char *env = getenv("MXX_CUDA_MEMPOOL_RELEASE_THRESHOLD_BYTES")) {
            char *end = nullptr;
            unsigned long long v = strtoull(env, &end, 10);
            if (end && end != env) threshold = v;
        }
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
    });
}

template <typename T>
int launch_copy_rect(GpuContext *ctx, T *dst, const T *src, const gp::RectGeom &g, size_t poly_words) {
    const size_t words = g.rows * g.cols * poly_words;
    if (words == 0) return 0;
    const int V = vec_width<T>(ctx);
    const size_t total = words / V;
    const unsigned threads = 256, grid = stride_grid(ctx, total, threads);
    switch (V) {
        case 4: gp::copy_rect_kernel<T, 4><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, g, total, poly_words); break;
        case 2: gp::copy_rect_kernel<T, 2><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, g, total, poly_words); break;
        default: gp::copy_rect_kernel<T, 1><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, g, total, poly_words); break;
    }
    gp::count_launch();
    GP_CHECK_LAUNCH("copy_rect_kernel");
    return 0;
}
}  // namespace

extern "C" int gpu_matrix_create(GpuContext *ctx, int level, size_t rows, size_t cols, int format, GpuMatrix **out) {
    if (!ctx || !out) return gp::set_error("invalid gpu_matrix_create arguments");
    *out = nullptr;
    if (level < -1 || level > ctx->level) return gp::set_error("invalid level in gpu_matrix_create");  // MatrixData.cu:209
    if (format != GPU_POLY_FORMAT_COEFF && format != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("invalid format in gpu_matrix_create");
    if (rows != 0 && cols > SIZE_MAX / rows) return gp::set_error("matrix size overflow in gpu_matrix_create");
    if (bind_device(ctx)) return 1;
    tune_pool(ctx->device);
    GpuMatrix *m = new (std::nothrow) GpuMatrix();
    if (!m) return gp::set_error("out of host memory in gpu_matrix_create");
    m->ctx = ctx;
    m->rows = rows;
    m->cols = cols;
    m->level = level;
    m->format = format;
    const size_t bytes = m->bytes();
    if (bytes) {
        cudaError_t err = cudaMallocAsync(&m->data, bytes, ctx->stream);
        if (err != cudaSuccess) {
            delete m;
            return gp::set_cuda_error(err, "cudaMallocAsync(gpu_matrix_create)");
        }
    }
    *out = m;
    return 0;
}

extern "C" void gpu_matrix_destroy(GpuMatrix *mat) {
    if (!mat) return;
    if (mat->data && mat->ctx && gp::context_alive(mat->ctx)) {
        cudaSetDevice(mat->ctx->device);
        cudaFreeAsync(mat->data, mat->ctx->stream);  // ordered after every use: all work is on this stream
    }
    delete mat;
}

extern "C" int gpu_matrix_copy_block(GpuMatrix *out, const GpuMatrix *src, size_t dst_row, size_t dst_col, size_t src_row,
                                     size_t src_col, size_t rows, size_t cols) {
    if (!out || !src) return gp::set_error("invalid gpu_matrix_copy_block arguments");
    if (!same_ctx(out, src)) return gp::set_error("context mismatch in gpu_matrix_copy_block");
    if (src_row + rows > src->rows || src_col + cols > src->cols || dst_row + rows > out->rows ||
        dst_col + cols > out->cols)
        return gp::set_error("block out of range in gpu_matrix_copy_block");  // MatrixData.cu:504-540
    out->format = src->format;  // MatrixData.cu:558
    if (rows == 0 || cols == 0 || out->level < 0) return 0;
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const gp::RectGeom g{out->cols, src->cols, dst_row, dst_col, src_row, src_col, rows, cols};
    // whole contiguous matrices: a plain device copy
    if (rows == src->rows && cols == src->cols && same_shape(out, src)) {
        if (out->data != src->data)
            GP_CUDA(cudaMemcpyAsync(out->data, src->data, src->bytes(), cudaMemcpyDeviceToDevice, ctx->stream),
                    "cudaMemcpyAsync(gpu_matrix_copy)");
        return 0;
    }
    return GP_DISPATCH_WIDTH(ctx, return launch_copy_rect<T>(ctx, (T *)out->data, (const T *)src->data, g, src->poly_words()););
}

extern "C" int gpu_matrix_copy(GpuMatrix *dst, const GpuMatrix *src) {
    if (!dst || !src) return gp::set_error("invalid gpu_matrix_copy arguments");
    if (!same_shape(dst, src)) return gp::set_error("size mismatch in gpu_matrix_copy");
    if (!same_ctx(dst, src)) return gp::set_error("context mismatch in gpu_matrix_copy");
    return gpu_matrix_copy_block(dst, src, 0, 0, 0, 0, src->rows, src->cols);
}

extern "C" int gpu_matrix_device_info(const GpuMatrix *mat, void **out_ptr, int *out_word_bytes, int *out_format) {
    if (!mat || !mat->ctx) return gp::set_error("invalid gpu_matrix_device_info arguments");
    if (out_ptr) *out_ptr = mat->data;
    if (out_word_bytes) *out_word_bytes = (int)mat->ctx->word_bytes();
    if (out_format) *out_format = mat->format;
    return 0;
}
