// MatrixNTT.inl — host side of the batched negacyclic NTT / INTT.
// Replaces gpu_matrix_ntt_all / gpu_matrix_intt_all and the host loop of
// cuda/src/matrix/MatrixNTT.cu:402-811 (twist + bit-reverse + log2 n stage launches
// + scale): one kernel per transform when a (poly, limb) fits a CTA's shared memory,
// otherwise radix-2^k global passes for the top stages plus one shared-memory kernel.

namespace {

template <typename T, int LOGB, bool NOCSUB>
int launch_fast_fwd(GpuContext *ctx, const T *src, T *dst, unsigned grid, uint32_t limbs) {
    static std::atomic<uint32_t> done{0};
    constexpr size_t smem = gp::fast_smem_bytes<T, LOGB>();
    auto kernel = gp::fast_ntt_fwd_kernel<T, LOGB, NOCSUB>;
    if (optin_smem(kernel, smem, ctx->device, done)) return 1;
    kernel<<<grid, gp::FastCfg<T, LOGB>::kThreads, smem, gp::cur(ctx)>>>(src, dst, limbs, (uint32_t)ctx->logN, fwd_tab<T>(ctx),
                                                                       limb_tab<T>(ctx));
    gp::count_launch();
    GP_CHECK_LAUNCH("fast_ntt_fwd_kernel");
    return 0;
}
template <typename T, int LOGB>
int launch_fast_inv(GpuContext *ctx, const T *src, T *dst, unsigned grid, uint32_t limbs) {
    static std::atomic<uint32_t> done{0};
    constexpr size_t smem = gp::fast_smem_bytes<T, LOGB>();
    auto kernel = gp::fast_ntt_inv_kernel<T, LOGB>;
    if (optin_smem(kernel, smem, ctx->device, done)) return 1;
    kernel<<<grid, gp::FastCfg<T, LOGB>::kThreads, smem, gp::cur(ctx)>>>(src, dst, limbs, (uint32_t)ctx->logN, inv_tab<T>(ctx),
                                                                       limb_tab<T>(ctx));
    gp::count_launch();
    GP_CHECK_LAUNCH("fast_ntt_inv_kernel");
    return 0;
}
template <typename T> int dispatch_fast_fwd(GpuContext *ctx, const T *src, T *dst, unsigned grid, uint32_t limbs, int logb) {
    if constexpr (gp::is_f64<T>::value) {
#define GP_CALL(LB) launch_fast_fwd<T, LB, false>(ctx, src, dst, grid, limbs)
        GP_FAST_LOGB_SWITCH(logb, GP_CALL)
#undef GP_CALL
    } else {
#define GP_CALL(LB) (ctx->nocsub ? launch_fast_fwd<T, LB, true>(ctx, src, dst, grid, limbs) : launch_fast_fwd<T, LB, false>(ctx, src, dst, grid, limbs))
        GP_FAST_LOGB_SWITCH(logb, GP_CALL)
#undef GP_CALL
    }
}
template <typename T> int dispatch_fast_inv(GpuContext *ctx, const T *src, T *dst, unsigned grid, uint32_t limbs, int logb) {
#define GP_CALL(LB) launch_fast_inv<T, LB>(ctx, src, dst, grid, limbs)
    GP_FAST_LOGB_SWITCH(logb, GP_CALL)
#undef GP_CALL
}

// opt in to > 48 KB dynamic shared memory for the runtime-planned kernels of ntt.cuh (once per device; the
// decomposition kernels of the same family are configured by their own translation unit, MatrixDecompose.inl)
template <typename T> int ensure_ntt_smem(GpuContext *ctx, size_t bytes) {
    static std::atomic<uint32_t> done_f{0}, done_i{0};
    if (bytes <= 48 * 1024) return 0;
    if ((int)bytes > ctx->max_smem_optin) return gp::set_error("NTT block does not fit shared memory on this device");
    if (optin_smem(gp::ntt_fwd_kernel<T>, (size_t)ctx->max_smem_optin, ctx->device, done_f)) return 1;
    return optin_smem(gp::ntt_inv_kernel<T>, (size_t)ctx->max_smem_optin, ctx->device, done_i);
}

template <typename T, int LOGR>
int launch_fwd_global(GpuContext *ctx, T *data, size_t pl, uint32_t limbs, uint32_t lspan) {
    const size_t total = pl << (ctx->logN - LOGR);
    const unsigned threads = 256, grid = (unsigned)((total + threads - 1) / threads);
    gp::ntt_fwd_global_pass<T, LOGR><<<grid, threads, 0, gp::cur(ctx)>>>(data, total, limbs, (uint32_t)ctx->logN, lspan,
                                                                         (const gp::TwPair<T> *)ctx->d_fwd,
                                                                         (const gp::LimbConst<T> *)ctx->d_limb);
    gp::count_launch();
    GP_CHECK_LAUNCH("ntt_fwd_global_pass");
    return 0;
}
template <typename T, int LOGR>
int launch_inv_global(GpuContext *ctx, T *data, size_t pl, uint32_t limbs, uint32_t lt0, bool final_pass) {
    const size_t total = pl << (ctx->logN - LOGR);
    const unsigned threads = 256, grid = (unsigned)((total + threads - 1) / threads);
    const auto *tw = (const gp::TwPair<T> *)ctx->d_inv;
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    if (final_pass)
        gp::ntt_inv_global_pass<T, LOGR, true><<<grid, threads, 0, gp::cur(ctx)>>>(data, total, limbs, (uint32_t)ctx->logN, lt0, tw, lc);
    else
        gp::ntt_inv_global_pass<T, LOGR, false><<<grid, threads, 0, gp::cur(ctx)>>>(data, total, limbs, (uint32_t)ctx->logN, lt0, tw, lc);
    gp::count_launch();
    GP_CHECK_LAUNCH("ntt_inv_global_pass");
    return 0;
}

template <typename T> int dispatch_fwd_global(GpuContext *ctx, T *d, size_t pl, uint32_t limbs, uint32_t lspan, int r) {
    switch (r) {
        case 1: return launch_fwd_global<T, 1>(ctx, d, pl, limbs, lspan);
        case 2: return launch_fwd_global<T, 2>(ctx, d, pl, limbs, lspan);
        case 3: return launch_fwd_global<T, 3>(ctx, d, pl, limbs, lspan);
        default: return launch_fwd_global<T, 4>(ctx, d, pl, limbs, lspan);
    }
}
template <typename T>
int dispatch_inv_global(GpuContext *ctx, T *d, size_t pl, uint32_t limbs, uint32_t lt0, int r, bool fin) {
    switch (r) {
        case 1: return launch_inv_global<T, 1>(ctx, d, pl, limbs, lt0, fin);
        case 2: return launch_inv_global<T, 2>(ctx, d, pl, limbs, lt0, fin);
        case 3: return launch_inv_global<T, 3>(ctx, d, pl, limbs, lt0, fin);
        default: return launch_inv_global<T, 4>(ctx, d, pl, limbs, lt0, fin);
    }
}

}  // namespace

// dst = NTT(src) for `polys` polynomials of `limbs` limbs each (src may equal dst).
// first_top_pass_done: the caller already ran the first global pass (fused with its producer, MatrixDecompose.inl).
template <typename T>
int gpm::run_ntt_fwd(GpuContext *ctx, const T *src, T *dst, size_t polys, uint32_t limbs, bool first_top_pass_done) {
    const size_t pl = polys * limbs;
    if (pl == 0) return 0;
    if (const int lc = ntt_cluster_log<T>(ctx); lc > 0 && !first_top_pass_done) {
        return gpm::run_cluster_fwd<T>(ctx, src, dst, pl, limbs, lc);
    }
    const int logn = ctx->logN, logb = ntt_block_log<T>(ctx), top = logn - logb;
    const bool fast = use_fast<T>(ctx, logb);
    const size_t smem = gp::ntt_smem_bytes<T>(logb);
    if (!fast && ensure_ntt_smem<T>(ctx, smem)) return 1;
    if (top > 0) {
        if (src != dst)
            GP_CUDA(cudaMemcpyAsync(dst, src, (pl << logn) * sizeof(T), cudaMemcpyDeviceToDevice, gp::cur(ctx)), "cudaMemcpyAsync(ntt)");
        src = dst;
        int passes[8], np = 0;
        split_top(top, passes, &np);
        uint32_t lspan = (uint32_t)logn;
        for (int i = 0; i < np; i++) {
            if (!(i == 0 && first_top_pass_done) && dispatch_fwd_global<T>(ctx, dst, pl, limbs, lspan, passes[i])) return 1;
            lspan -= passes[i];
        }
    }
    const unsigned grid = (unsigned)(pl << top);
    if (fast) {
        if constexpr (sizeof(T) == 8) {
            if (ctx->f64_ntt) return dispatch_fast_fwd<gp::F64>(ctx, (const gp::F64 *)src, (gp::F64 *)dst, grid, limbs, logb);
        }
        return dispatch_fast_fwd<T>(ctx, src, dst, grid, limbs, logb);
    }
    gp::ntt_fwd_kernel<T><<<grid, gp::NttCfg<T>::kThreads, smem, gp::cur(ctx)>>>(
        src, dst, limbs, (uint32_t)logn, (uint32_t)logb, (const gp::TwPair<T> *)ctx->d_fwd, (const gp::LimbConst<T> *)ctx->d_limb);
    gp::count_launch();
    GP_CHECK_LAUNCH("ntt_fwd_kernel");
    return 0;
}

template <typename T> int gpm::run_ntt_inv(GpuContext *ctx, const T *src, T *dst, size_t polys, uint32_t limbs) {
    const size_t pl = polys * limbs;
    if (pl == 0) return 0;
    if (const int lc = ntt_cluster_log<T>(ctx); lc > 0) {
        return gpm::run_cluster_inv<T>(ctx, src, dst, pl, limbs, lc);
    }
    const int logn = ctx->logN, logb = ntt_block_log<T>(ctx), top = logn - logb;
    const bool fast = use_fast<T>(ctx, logb);
    const size_t smem = gp::ntt_smem_bytes<T>(logb);
    const unsigned grid = (unsigned)(pl << top);
    if (fast) {
        bool f64 = false;
        if constexpr (sizeof(T) == 8) f64 = ctx->f64_ntt;
        if (f64) {
            if (dispatch_fast_inv<gp::F64>(ctx, (const gp::F64 *)src, (gp::F64 *)dst, grid, limbs, logb)) return 1;
        } else if (dispatch_fast_inv<T>(ctx, src, dst, grid, limbs, logb)) return 1;
    } else {
        if (ensure_ntt_smem<T>(ctx, smem)) return 1;
        gp::ntt_inv_kernel<T><<<grid, gp::NttCfg<T>::kThreads, smem, gp::cur(ctx)>>>(
            src, dst, limbs, (uint32_t)logn, (uint32_t)logb, (const gp::TwPair<T> *)ctx->d_inv, (const gp::LimbConst<T> *)ctx->d_limb);
        gp::count_launch();
        GP_CHECK_LAUNCH("ntt_inv_kernel");
    }
    if (top > 0) {
        int passes[8], np = 0;
        split_top(top, passes, &np);
        uint32_t lt0 = (uint32_t)logb;
        // smallest distances first; the radices are applied in reverse so the forward and
        // inverse transforms use the same pass boundaries
        for (int i = np - 1; i >= 0; i--) {
            if (dispatch_inv_global<T>(ctx, dst, pl, limbs, lt0, passes[i], i == 0)) return 1;
            lt0 += passes[i];
        }
    }
    return 0;
}

template int gpm::run_ntt_fwd<uint32_t>(GpuContext *, const uint32_t *, uint32_t *, size_t, uint32_t, bool);
template int gpm::run_ntt_fwd<uint64_t>(GpuContext *, const uint64_t *, uint64_t *, size_t, uint32_t, bool);
template int gpm::run_ntt_inv<uint32_t>(GpuContext *, const uint32_t *, uint32_t *, size_t, uint32_t);
template int gpm::run_ntt_inv<uint64_t>(GpuContext *, const uint64_t *, uint64_t *, size_t, uint32_t);

int gpm::matrix_transform(GpuMatrix *mat, bool inverse) {
    GpuContext *ctx = mat->ctx;
    if (mat->polys() == 0 || mat->level < 0) {
        mat->format = inverse ? GPU_POLY_FORMAT_COEFF : GPU_POLY_FORMAT_EVAL;
        return 0;
    }
    if (bind_device(ctx)) return 1;
    const uint32_t limbs = (uint32_t)mat->limbs();
    int st = GP_DISPATCH_WIDTH(ctx, T *d = (T *)mat->data;
                               return inverse ? run_ntt_inv<T>(ctx, d, d, mat->polys(), limbs)
                                              : run_ntt_fwd<T>(ctx, d, d, mat->polys(), limbs););
    if (st) return st;
    mat->format = inverse ? GPU_POLY_FORMAT_COEFF : GPU_POLY_FORMAT_EVAL;
    return 0;
}

extern "C" int gpu_matrix_ntt_all(GpuMatrix *mat) {
    if (!mat || !mat->ctx) return gp::set_error("invalid gpu_matrix_ntt_all arguments");
    if (mat->format == GPU_POLY_FORMAT_EVAL) return 0;  // MatrixNTT.cu:793-796
    return matrix_transform(mat, false);
}

extern "C" int gpu_matrix_intt_all(GpuMatrix *mat) {
    if (!mat || !mat->ctx) return gp::set_error("invalid gpu_matrix_intt_all arguments");
    if (mat->format == GPU_POLY_FORMAT_COEFF) return 0;  // MatrixNTT.cu:806-809
    return matrix_transform(mat, true);
}

// extension: rows [dst_row, dst_row + src->rows) of dst <- coefficient form of src (src untouched): the inverse transform
// written straight into its block of an all-gather buffer (limb-sharded mul_decompose, mxx_b200/dist.py gather_towers),
// instead of clone + intt_all + copy_block.  dst is tagged COEFF.
extern "C" int gpu_matrix_intt_to(GpuMatrix *dst, size_t dst_row, const GpuMatrix *src) {
    if (!dst || !src || !dst->ctx) return gp::set_error("invalid gpu_matrix_intt_to arguments");
    if (dst->ctx != src->ctx || dst->level != src->level) return gp::set_error("context mismatch in gpu_matrix_intt_to");
    if (dst->cols != src->cols || dst_row + src->rows > dst->rows) return gp::set_error("size mismatch in gpu_matrix_intt_to");
    if (dst == src || (dst->data && dst->data == src->data)) return gp::set_error("gpu_matrix_intt_to output must not alias the input");
    dst->format = GPU_POLY_FORMAT_COEFF;
    if (src->polys() == 0 || src->level < 0) return 0;
    GpuContext *ctx = dst->ctx;
    if (bind_device(ctx)) return 1;
    const uint32_t limbs = (uint32_t)src->limbs();
    const size_t off = dst_row * dst->cols * src->poly_words();
    return GP_DISPATCH_WIDTH(
        ctx, T *d = (T *)dst->data + off; const T *sp = (const T *)src->data;
        if (src->format == GPU_POLY_FORMAT_COEFF) {
            GP_CUDA(cudaMemcpyAsync(d, sp, src->bytes(), cudaMemcpyDeviceToDevice, gp::cur(ctx)), "cudaMemcpyAsync(intt_to)");
            return 0;
        } return run_ntt_inv<T>(ctx, sp, d, src->polys(), limbs););
}
