// MatrixSampling.inl — samplers of the trapdoor path: random matrices from a
// distribution, the G-lattice Gaussian sampler and the perturbation p1.
// Replaces cuda/src/matrix/MatrixSampling.cu:412-491 and MatrixTrapdoor.cu:1672-2157,
// 2159-2474, 2476-2918.  Same validation and result formats; the work is one fused kernel
// per call on the context stream (the reference: per-limb launches, an int64 staging
// buffer, a scatter kernel per limb and event bookkeeping between limb streams).
// There is no host sampling path: every integer is drawn on the device.

struct GpuP1CovarianceCache {
    GpuContext *ctx = nullptr;
    int level = -1;
    size_t d = 0;         // A, B, D are d x d
    double sigma = 0.0;   // the `c` of the preimage sampler
    double s = 0.0;
    double *sd = nullptr;   // [2d][n]
    double *upd = nullptr;  // [2d*2d][n]
};

namespace {

inline gp::Seed256 to_seed(const GpuRngSeed &s) {
    gp::Seed256 r;
    for (int i = 0; i < 4; i++) r.w[i] = s.words[i];
    return r;
}

inline unsigned blocks_for(size_t items, unsigned threads) { return (unsigned)((items + threads - 1) / threads); }

// Tasks per lane of the lane-balanced Gaussian kernel (sampling.cuh gauss_integers_batch): MXX_SAMPLER_BATCH = 1 (one task per
// thread), 2, 4 or 8 (default, 12.2 -> 8.7 ms on the C5 matrix, profiles/r02_sampler_lab.txt).  A warp owns 32 K consecutive
// coefficients of one polynomial, so K is halved until 32 K divides the ring dimension.
inline int batch_for(int knob, uint32_t n) {
    int k = knob >= 8 ? 8 : (knob >= 4 ? 4 : (knob >= 2 ? 2 : 1));
    while (k > 1 && n % (32u * (unsigned)k)) k >>= 1;
    return k;
}
inline int sampler_batch(uint32_t n) {
    static const int knob = getenv("MXX_SAMPLER_BATCH") ? atoi(getenv("MXX_SAMPLER_BATCH")) : 8;
    return batch_for(knob, n);
}
template <typename T>
int run_sample_distribution(GpuContext *ctx, GpuMatrix *out, int dist, double sigma, const gp::Seed256 &seed, size_t full_cols,
                            size_t col_offset) {
    gp::SampleGeom g;
    g.polys = out->polys();
    g.local_cols = out->cols;
    g.full_cols = full_cols;
    g.col_offset = col_offset;
    g.n = (uint32_t)ctx->N;
    g.limbs = (uint32_t)out->limbs();
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    if (dist == GPU_MATRIX_DIST_UNIFORM) {
        const size_t items = g.polys * g.limbs * g.n;
        if (items > (size_t)0x7fffffff * 256) return gp::set_error("matrix too large in gpu_matrix_sample_distribution");
        gp::sample_uniform_kernel<T><<<blocks_for(items, 256), 256, 0, gp::cur(ctx)>>>((T *)out->data, g, seed, lc);
    } else {
        const size_t items = g.polys * g.n;
        if (items > (size_t)0x7fffffff * 256) return gp::set_error("matrix too large in gpu_matrix_sample_distribution");
        const int batch = dist == GPU_MATRIX_DIST_GAUSS ? sampler_batch(g.n) : 1;
        switch (batch) {
#define GP_SGB(K)                                                                                                              \
    case K:                                                                                                                    \
        gp::sample_gauss_batch_kernel<T, K><<<blocks_for(items / K, 128), 128, 0, gp::cur(ctx)>>>((T *)out->data, g, sigma, seed, lc); \
        break;
            GP_SGB(2)
            GP_SGB(4)
            GP_SGB(8)
#undef GP_SGB
            default:
                gp::sample_small_kernel<T><<<blocks_for(items, 256), 256, 0, gp::cur(ctx)>>>((T *)out->data, g, dist, sigma, seed, lc);
        }
    }
    gp::count_launch();
    GP_CHECK_LAUNCH("sample_distribution kernel");
    return run_ntt_fwd<T>(ctx, (T *)out->data, (T *)out->data, out->polys(), g.limbs);
}

int sample_distribution_common(GpuMatrix *out, int dist, double sigma, GpuRngSeed seed, size_t full_cols, size_t col_offset) {
    if (!out || !out->ctx) return gp::set_error("invalid gpu_matrix_sample_distribution arguments");
    if (dist < GPU_MATRIX_DIST_UNIFORM || dist > GPU_MATRIX_DIST_TERNARY)
        return gp::set_error("invalid dist_type in gpu_matrix_sample_distribution");
    if (dist == GPU_MATRIX_DIST_GAUSS && !(sigma > 0.0))
        return gp::set_error("sigma must be positive in gpu_matrix_sample_distribution");
    if (col_offset > full_cols || out->cols > full_cols - col_offset)
        return gp::set_error("column range out of bounds in gpu_matrix_sample_distribution");
    if (out->polys() == 0) {
        out->format = GPU_POLY_FORMAT_EVAL;
        return 0;
    }
    if (out->level < 0) return gp::set_error("invalid level in gpu_matrix_sample_distribution");
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const gp::Seed256 s = to_seed(seed);
    int st = GP_DISPATCH_WIDTH(ctx, return run_sample_distribution<T>(ctx, out, dist, sigma, s, full_cols, col_offset););
    if (!st) out->format = GPU_POLY_FORMAT_EVAL;  // MatrixSampling.cu:462-469: sampled in COEFF, returned in EVAL
    return st;
}

template <typename T>
int run_gauss_samp(GpuContext *ctx, const GpuMatrix *src, GpuMatrix *out, uint32_t base_bits, uint32_t dpt, double c,
                   const gp::Seed256 &seed, bool to_eval) {
    const uint32_t limbs = (uint32_t)src->limbs(), n = (uint32_t)ctx->N;
    const size_t items = src->polys() * limbs * n;
    if (items > (size_t)0x7fffffff * 128) return gp::set_error("matrix too large in gpu_matrix_gauss_samp_gq_arb_base");
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    const unsigned grid = blocks_for(items, 128);
    if (dpt <= 8)
        gp::gauss_samp_gq_kernel<T, 8><<<grid, 128, 0, gp::cur(ctx)>>>((const T *)src->data, (T *)out->data, src->polys(), src->cols,
                                                                       n, limbs, base_bits, dpt, c, seed, lc);
    else
        gp::gauss_samp_gq_kernel<T, gp::kMaxGaussDigits><<<grid, 128, 0, gp::cur(ctx)>>>(
            (const T *)src->data, (T *)out->data, src->polys(), src->cols, n, limbs, base_bits, dpt, c, seed, lc);
    gp::count_launch();
    GP_CHECK_LAUNCH("gauss_samp_gq_kernel");
    if (to_eval) return run_ntt_fwd<T>(ctx, (T *)out->data, (T *)out->data, out->polys(), limbs);
    return 0;
}

}  // namespace

extern "C" int gpu_matrix_sample_distribution(GpuMatrix *out, int dist_type, double sigma, GpuRngSeed seed) {
    return sample_distribution_common(out, dist_type, sigma, seed, out ? out->cols : 0, 0);
}

extern "C" int gpu_matrix_sample_distribution_columns(GpuMatrix *out, int dist_type, double sigma, GpuRngSeed seed,
                                                      size_t full_ncol, size_t col_offset) {
    return sample_distribution_common(out, dist_type, sigma, seed, full_ncol, col_offset);
}

extern "C" int gpu_matrix_gauss_samp_gq_arb_base(GpuMatrix *src, uint32_t base_bits, double c, double dgg_stddev, GpuRngSeed seed,
                                                 GpuMatrix *out) {
    (void)dgg_stddev;  // unused by the reference too (MatrixTrapdoor.cu:1680)
    if (!src || !out || !src->ctx) return gp::set_error("invalid gpu_matrix_gauss_samp_gq_arb_base arguments");
    if (base_bits == 0 || base_bits >= 63) return gp::set_error("invalid base_bits in gpu_matrix_gauss_samp_gq_arb_base");
    if (!(c > 0.0)) return gp::set_error("c must be positive in gpu_matrix_gauss_samp_gq_arb_base");
    if (!same_ctx(src, out)) return gp::set_error("context mismatch in gpu_matrix_gauss_samp_gq_arb_base");
    const int want = out->format;
    if (want != GPU_POLY_FORMAT_COEFF && want != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("invalid output format in gpu_matrix_gauss_samp_gq_arb_base");
    if (src->level < 0) return gp::set_error("invalid level in gpu_matrix_gauss_samp_gq_arb_base");
    if (out == src || (out->data && out->data == src->data))
        return gp::set_error("gpu_matrix_gauss_samp_gq_arb_base output must not alias the input");
    GpuContext *ctx = src->ctx;
    const uint32_t dpt = digits_per_tower(ctx, base_bits);  // widest modulus of the context, :1712-1722
    if (dpt == 0 || dpt > (uint32_t)gp::kMaxGaussDigits)
        return gp::set_error("invalid digits_per_tower in gpu_matrix_gauss_samp_gq_arb_base");
    const size_t ell = (size_t)dpt * src->limbs();
    if (out->rows != src->rows * ell || out->cols != src->cols)
        return gp::set_error("output size mismatch in gpu_matrix_gauss_samp_gq_arb_base");
    if (src->polys() == 0) {
        out->format = GPU_POLY_FORMAT_EVAL;
        return 0;
    }
    // like the reference (:1739-1746) an EVAL source is brought to coefficient form in place
    if (src->format == GPU_POLY_FORMAT_EVAL && matrix_transform(src, true)) return 1;
    if (bind_device(ctx)) return 1;
    const gp::Seed256 s = to_seed(seed);
    int st = GP_DISPATCH_WIDTH(ctx, return run_gauss_samp<T>(ctx, src, out, base_bits, dpt, c, s, want == GPU_POLY_FORMAT_EVAL););
    if (!st) out->format = want;
    return st;
}

extern "C" int gpu_matrix_create_p1_covariance_cache(const GpuMatrix *a_mat, const GpuMatrix *b_mat, const GpuMatrix *d_mat,
                                                     double sigma, double s, double dgg_stddev, GpuP1CovarianceCache **out_cache) {
    if (!out_cache) return gp::set_error("null output in gpu_matrix_create_p1_covariance_cache");
    *out_cache = nullptr;
    if (!a_mat || !b_mat || !d_mat || !a_mat->ctx) return gp::set_error("invalid gpu_matrix_create_p1_covariance_cache arguments");
    if (!(sigma > 0.0) || !(s > sigma) || !(dgg_stddev > 0.0))
        return gp::set_error("invalid Gaussian parameters in gpu_matrix_create_p1_covariance_cache");
    if (a_mat->ctx != b_mat->ctx || a_mat->ctx != d_mat->ctx)
        return gp::set_error("context mismatch in gpu_matrix_create_p1_covariance_cache");
    if (a_mat->level != b_mat->level || a_mat->level != d_mat->level)
        return gp::set_error("level mismatch in gpu_matrix_create_p1_covariance_cache");
    if (a_mat->format != GPU_POLY_FORMAT_COEFF || b_mat->format != GPU_POLY_FORMAT_COEFF || d_mat->format != GPU_POLY_FORMAT_COEFF)
        return gp::set_error("p1 covariance cache inputs must be coefficient matrices");
    const size_t d = a_mat->rows;
    if (a_mat->cols != d || b_mat->rows != d || b_mat->cols != d || d_mat->rows != d || d_mat->cols != d)
        return gp::set_error("A/B/D must be dxd in gpu_matrix_create_p1_covariance_cache");
    GpuContext *ctx = a_mat->ctx;
    auto *cache = new (std::nothrow) GpuP1CovarianceCache();
    if (!cache) return gp::set_error("out of host memory in gpu_matrix_create_p1_covariance_cache");
    cache->ctx = ctx;
    cache->level = a_mat->level;
    if (d == 0) {
        *out_cache = cache;
        return 0;
    }
    if (a_mat->level < 0) {
        delete cache;
        return gp::set_error("invalid level in gpu_matrix_create_p1_covariance_cache");
    }
    if (d > 0xffffu) {
        delete cache;
        return gp::set_error("invalid dimension in gpu_matrix_create_p1_covariance_cache");
    }
    cache->d = d;
    cache->sigma = sigma;
    cache->s = s;
    const size_t n = (size_t)ctx->N, m = 2 * d;
    auto fail = [&](int) {
        gpu_matrix_destroy_p1_covariance_cache(cache);
        return 1;
    };
    if (bind_device(ctx)) return fail(1);
    cudaStream_t stream = gp::cur(ctx);
    if (cudaMallocAsync((void **)&cache->sd, m * n * sizeof(double), stream) != cudaSuccess ||
        cudaMallocAsync((void **)&cache->upd, m * m * n * sizeof(double), stream) != cudaSuccess) {
        gp::set_cuda_error(cudaGetLastError(), "cudaMallocAsync(p1 covariance cache)");
        return fail(1);
    }
    Scratch cov;
    if (cov.alloc(m * m * n * sizeof(double), stream)) return fail(1);
    const uint32_t limbs = (uint32_t)a_mat->limbs();
    const uint64_t q0 = ctx->moduli[0];
    int st = GP_DISPATCH_WIDTH(
        ctx, gp::p1_factor_kernel<T><<<blocks_for(n, 128), 128, 0, stream>>>(
                 (const T *)a_mat->data, (const T *)b_mat->data, (const T *)d_mat->data, (uint32_t)d, (uint32_t)n, limbs, q0, sigma, s,
                 dgg_stddev, (double *)cov.p, cache->sd, cache->upd);
        gp::count_launch(); GP_CHECK_LAUNCH("p1_factor_kernel"); return 0;);
    if (st) return fail(1);
    *out_cache = cache;
    return 0;
}

extern "C" void gpu_matrix_destroy_p1_covariance_cache(GpuP1CovarianceCache *cache) {
    if (!cache) return;
    if (cache->ctx && (cache->sd || cache->upd) && gp::context_alive(cache->ctx)) {
        cudaSetDevice(cache->ctx->device);
        if (cache->sd) cudaFreeAsync(cache->sd, cache->ctx->stream);
        if (cache->upd) cudaFreeAsync(cache->upd, cache->ctx->stream);
    }
    delete cache;
}

extern "C" int gpu_matrix_sample_p1_full_cached(const GpuP1CovarianceCache *cache, const GpuMatrix *tp2, GpuRngSeed seed,
                                                GpuMatrix *out) {
    if (!cache || !tp2 || !out) return gp::set_error("invalid gpu_matrix_sample_p1_full_cached arguments");
    if (cache->ctx != tp2->ctx || cache->ctx != out->ctx) return gp::set_error("context mismatch in gpu_matrix_sample_p1_full_cached");
    if (cache->level != tp2->level || cache->level != out->level)
        return gp::set_error("level mismatch in gpu_matrix_sample_p1_full_cached");
    if (tp2->format != GPU_POLY_FORMAT_COEFF)
        return gp::set_error("tp2 must be coefficient format in gpu_matrix_sample_p1_full_cached");
    if (tp2->rows != 2 * cache->d || out->rows != 2 * cache->d || out->cols != tp2->cols)
        return gp::set_error("tp2/out shape mismatch in gpu_matrix_sample_p1_full_cached");
    if (out == tp2 || (out->data && out->data == tp2->data))
        return gp::set_error("gpu_matrix_sample_p1_full_cached output must not alias tp2");
    const size_t cols = tp2->cols;
    if (cols == 0 || cache->d == 0) {
        out->format = GPU_POLY_FORMAT_EVAL;
        return 0;
    }
    if (cache->level < 0) return gp::set_error("invalid level in gpu_matrix_sample_p1_full_cached");
    const double denom = cache->s * cache->s - cache->sigma * cache->sigma;
    if (!(denom > 0.0)) return gp::set_error("invalid cached Gaussian denominator");
    const double c_scale = -(cache->sigma * cache->sigma) / denom;  // MatrixTrapdoor.cu:1455-1460
    GpuContext *ctx = cache->ctx;
    if (bind_device(ctx)) return 1;
    cudaStream_t stream = gp::cur(ctx);
    const uint32_t m = (uint32_t)(2 * cache->d), n = (uint32_t)ctx->N, limbs = (uint32_t)tp2->limbs();
    const size_t items = cols * n;
    if (items > (size_t)0x7fffffff * 128) return gp::set_error("matrix too large in gpu_matrix_sample_p1_full_cached");
    const gp::Seed256 s = to_seed(seed);
    const uint64_t q0 = ctx->moduli[0];
    Scratch mean_ws, z_ws;
    if (m > 8) {
        if (mean_ws.alloc((size_t)m * items * sizeof(double), stream)) return 1;
        if (z_ws.alloc((size_t)m * items * sizeof(int64_t), stream)) return 1;
    }
    int st = GP_DISPATCH_WIDTH(
        ctx, const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb; const unsigned grid = blocks_for(items, 128);
        if (m <= 8) gp::p1_sample_kernel<T, 8><<<grid, 128, 0, stream>>>((const T *)tp2->data, (T *)out->data, m, cols, n, limbs, q0, c_scale, s,
                                                                       cache->sd, cache->upd, nullptr, nullptr, lc);
        else gp::p1_sample_kernel<T, 0><<<grid, 128, 0, stream>>>((const T *)tp2->data, (T *)out->data, m, cols, n, limbs, q0, c_scale, s,
                                                                cache->sd, cache->upd, (double *)mean_ws.p, (int64_t *)z_ws.p, lc);
        gp::count_launch(); GP_CHECK_LAUNCH("p1_sample_kernel");
        return run_ntt_fwd<T>(ctx, (T *)out->data, (T *)out->data, out->polys(), limbs););
    if (!st) out->format = GPU_POLY_FORMAT_EVAL;
    return st;
}
