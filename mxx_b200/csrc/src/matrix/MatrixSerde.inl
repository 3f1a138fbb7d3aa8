// MatrixSerde.inl — host RNS words <-> device storage.
// Replaces gpu_matrix_load_rns_batch / store_rns_batch / store_const_coeff_batch of
// cuda/src/matrix/MatrixSerde.cu (:566-741, :743-924, :926-1064): ONE DMA of the whole
// host buffer into a stream-ordered staging area and ONE pack kernel for all limbs,
// instead of a staging allocation + strided 2D copy + kernel per limb.  The host layout
// is the reference's: [poly row-major][limb][coeff] little-endian u64 with a caller
// supplied per-poly stride (src/matrix/gpu_dcrt_poly.rs:629-663, 1916-1920).

namespace {

int make_event_set(GpuContext *ctx, GpuEventSet **out_events) {
    GpuEventSet *ev = new (std::nothrow) GpuEventSet();
    if (!ev) return gp::set_error("out of host memory for event set");
    ev->device = ctx->device;
    cudaError_t err = cudaEventCreateWithFlags(&ev->event, cudaEventDisableTiming);
    if (err == cudaSuccess) err = cudaEventRecord(ev->event, ctx->stream);
    if (err != cudaSuccess) {
        if (ev->event) cudaEventDestroy(ev->event);
        delete ev;
        return gp::set_cuda_error(err, "cudaEventRecord");
    }
    *out_events = ev;
    return 0;
}

template <typename T>
int launch_pack(GpuContext *ctx, T *dst, const uint64_t *src, size_t words, bool bitrev, uint32_t limbs) {
    const unsigned threads = 256, grid = stride_grid(ctx, words, threads);
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    if (bitrev)
        gp::pack_kernel<T, uint64_t, true><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, words, (uint32_t)ctx->logN, 0, limbs, lc);
    else
        gp::pack_kernel<T, uint64_t, false><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, words, (uint32_t)ctx->logN, 0, limbs, lc);
    gp::count_launch();
    GP_CHECK_LAUNCH("pack_kernel");
    return 0;
}
template <typename T>
int launch_unpack(GpuContext *ctx, uint64_t *dst, const T *src, size_t words, bool bitrev) {
    const unsigned threads = 256, grid = stride_grid(ctx, words, threads);
    if (bitrev)
        gp::unpack_kernel<T, true><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, words, (uint32_t)ctx->logN);
    else
        gp::unpack_kernel<T, false><<<grid, threads, 0, gp::cur(ctx)>>>(dst, src, words, (uint32_t)ctx->logN);
    gp::count_launch();
    GP_CHECK_LAUNCH("unpack_kernel");
    return 0;
}
}  // namespace

extern "C" int gpu_matrix_load_rns_batch(GpuMatrix *mat, const uint8_t *bytes, size_t bytes_per_poly, int format,
                                         GpuEventSet **out_events) {
    if (!mat || !out_events) return gp::set_error("invalid gpu_matrix_load_rns_batch arguments");
    *out_events = nullptr;
    if (format != GPU_POLY_FORMAT_COEFF && format != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("invalid format in gpu_matrix_load_rns_batch");
    GpuContext *ctx = mat->ctx;
    const size_t polys = mat->polys();
    if (polys == 0 || mat->level < 0) {
        mat->format = format;
        return 0;
    }
    if (!bytes) return gp::set_error("null bytes in gpu_matrix_load_rns_batch");
    const size_t tight = mat->poly_words() * sizeof(uint64_t);
    if (bytes_per_poly % sizeof(uint64_t) != 0)
        return gp::set_error("bytes_per_poly must be a multiple of 8 in gpu_matrix_load_rns_batch");  // MatrixSerde.cu:618
    if (bytes_per_poly < tight)
        return gp::set_error("bytes_per_poly too small in gpu_matrix_load_rns_batch");  // MatrixSerde.cu:622
    if (bind_device(ctx)) return 1;
    if (bytes_per_poly == tight) {  // large u32-context loads: chunked two-lane pipeline (MatrixHostLoad.inl)
        const int hs = load_hybrid(ctx, mat, (const uint64_t *)bytes, polys * mat->poly_words(), format == GPU_POLY_FORMAT_EVAL);
        if (hs > 0) return hs;
        if (hs == 0) {
            mat->format = format;
            return make_event_set(ctx, out_events);
        }
    }
    Scratch stage;
    if (stage.alloc(polys * tight, ctx->stream)) return 1;
    if (bytes_per_poly == tight)
        GP_CUDA(cudaMemcpyAsync(stage.p, bytes, polys * tight, cudaMemcpyHostToDevice, ctx->stream), "cudaMemcpyAsync(H2D)");
    else
        GP_CUDA(cudaMemcpy2DAsync(stage.p, tight, bytes, bytes_per_poly, tight, polys, cudaMemcpyHostToDevice, ctx->stream),
                "cudaMemcpy2DAsync(H2D)");
    const size_t words = polys * mat->poly_words();
    const bool bitrev = format == GPU_POLY_FORMAT_EVAL;
    int st = GP_DISPATCH_WIDTH(ctx, return launch_pack<T>(ctx, (T *)mat->data, (const uint64_t *)stage.p, words, bitrev, (uint32_t)mat->limbs()););
    if (st) return st;
    mat->format = format;  // MatrixSerde.cu:739
    return make_event_set(ctx, out_events);
}

extern "C" int gpu_matrix_store_rns_batch(const GpuMatrix *mat, uint8_t *bytes_out, size_t bytes_per_poly, int format,
                                          GpuEventSet **out_events) {
    if (!mat || !out_events) return gp::set_error("invalid gpu_matrix_store_rns_batch arguments");
    *out_events = nullptr;
    if (format != GPU_POLY_FORMAT_COEFF && format != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("invalid format in gpu_matrix_store_rns_batch");
    GpuContext *ctx = mat->ctx;
    const size_t polys = mat->polys();
    if (polys == 0 || mat->level < 0) return 0;
    if (format != mat->format)
        return gp::set_error("format mismatch in gpu_matrix_store_rns_batch");  // MatrixSerde.cu:765-768
    if (!bytes_out) return gp::set_error("null bytes_out in gpu_matrix_store_rns_batch");
    const size_t tight = mat->poly_words() * sizeof(uint64_t);
    if (bytes_per_poly % sizeof(uint64_t) != 0)
        return gp::set_error("bytes_per_poly must be a multiple of 8 in gpu_matrix_store_rns_batch");
    if (bytes_per_poly < tight) return gp::set_error("bytes_per_poly too small in gpu_matrix_store_rns_batch");
    if (bind_device(ctx)) return 1;
    Scratch stage;
    if (stage.alloc(polys * tight, ctx->stream)) return 1;
    const size_t words = polys * mat->poly_words();
    const bool bitrev = format == GPU_POLY_FORMAT_EVAL;
    int st = GP_DISPATCH_WIDTH(ctx, return launch_unpack<T>(ctx, (uint64_t *)stage.p, (const T *)mat->data, words, bitrev););
    if (st) return st;
    if (bytes_per_poly == tight)
        GP_CUDA(cudaMemcpyAsync(bytes_out, stage.p, polys * tight, cudaMemcpyDeviceToHost, ctx->stream), "cudaMemcpyAsync(D2H)");
    else
        GP_CUDA(cudaMemcpy2DAsync(bytes_out, bytes_per_poly, stage.p, tight, tight, polys, cudaMemcpyDeviceToHost, ctx->stream),
                "cudaMemcpy2DAsync(D2H)");
    return make_event_set(ctx, out_events);
}

extern "C" int gpu_matrix_store_const_coeff_batch(const GpuMatrix *mat, uint64_t *words_out, size_t words_per_poly,
                                                  GpuEventSet **out_events) {
    if (!mat || !out_events) return gp::set_error("invalid gpu_matrix_store_const_coeff_batch arguments");
    *out_events = nullptr;
    if (!mat->ctx) return gp::set_error("invalid context in gpu_matrix_store_const_coeff_batch");
    if (mat->format != GPU_POLY_FORMAT_COEFF)
        return gp::set_error("gpu_matrix_store_const_coeff_batch requires coeff format");  // MatrixSerde.cu:943
    GpuContext *ctx = mat->ctx;
    const size_t polys = mat->polys();
    if (polys && !words_out) return gp::set_error("null words_out in gpu_matrix_store_const_coeff_batch");
    if (polys == 0) return 0;
    if (mat->level < 0) return gp::set_error("invalid level in gpu_matrix_store_const_coeff_batch");
    const size_t limbs = mat->limbs();
    if (words_per_poly < limbs) return gp::set_error("words_per_poly too small in gpu_matrix_store_const_coeff_batch");
    if (bind_device(ctx)) return 1;
    if (words_per_poly > limbs) std::fill_n(words_out, polys * words_per_poly, (uint64_t)0);  // MatrixSerde.cu:981
    Scratch stage;
    const size_t count = polys * limbs;
    if (stage.alloc(count * sizeof(uint64_t), ctx->stream)) return 1;
    const unsigned threads = 256, grid = (unsigned)((count + threads - 1) / threads);
    if (ctx->wide)
        gp::const_coeff_kernel<uint64_t><<<grid, threads, 0, gp::cur(ctx)>>>((uint64_t *)stage.p, (const uint64_t *)mat->data, count,
                                                                            (uint32_t)ctx->logN);
    else
        gp::const_coeff_kernel<uint32_t><<<grid, threads, 0, gp::cur(ctx)>>>((uint64_t *)stage.p, (const uint32_t *)mat->data, count,
                                                                            (uint32_t)ctx->logN);
    gp::count_launch();
    GP_CHECK_LAUNCH("const_coeff_kernel");
    GP_CUDA(cudaMemcpy2DAsync(words_out, words_per_poly * 8, stage.p, limbs * 8, limbs * 8, polys, cudaMemcpyDeviceToHost,
                              ctx->stream),
            "cudaMemcpy2DAsync(const coeff)");
    return make_event_set(ctx, out_events);
}
