// MatrixDecompose.inl — gadget matrix, gadget decomposition, fused mul_decompose.
// Replaces cuda/src/matrix/MatrixDecompose.cu (gpu_matrix_fill_gadget :307-453,
// fill_small_decomposed_identity_chunk :507-890, gpu_matrix_decompose_base :892-1342).
// Decomposition + forward NTT is ONE kernel (digit extraction is the load functor of the
// NTT, ntt.cuh decompose_kernel): no memset, no coefficient-form digit matrix in HBM,
// no separate NTT sweep.

namespace {

uint64_t host_powmod(uint64_t b, uint64_t e, uint64_t q) {
    unsigned __int128 r = 1 % q, c = b % q;
    for (; e; e >>= 1) {
        if (e & 1) r = r * c % q;
        c = c * c % q;
    }
    return (uint64_t)r;
}

// gpu_matrix_fill_gadget takes the widest modulus among the matrix's ACTIVE limbs (cuda/src/matrix/MatrixDecompose.cu:346-355);
// decompose and gauss_samp use all moduli of the context (:924-935) — digits_per_tower(ctx, ...)
uint32_t gadget_digits_per_tower(const GpuMatrix *m, uint32_t base_bits) {
    uint32_t bits = 0;
    for (size_t i = 0; i < m->limbs() && i < m->ctx->moduli.size(); i++)
        bits = std::max(bits, 64u - (uint32_t)__builtin_clzll(m->ctx->moduli[i]));
    return (bits + base_bits - 1) / base_bits;
}

template <typename T>
int run_fill_gadget(GpuContext *ctx, GpuMatrix *out, uint32_t base_bits, bool small) {
    const uint32_t dpt = gadget_digits_per_tower(out, base_bits), limbs = (uint32_t)out->limbs();
    const uint32_t ell = small ? dpt : dpt * limbs;
    std::vector<T> pow((size_t)limbs * dpt);
    for (uint32_t l = 0; l < limbs; l++)
        for (uint32_t dg = 0; dg < dpt; dg++)  // 2^(bb*dg) mod q_l (MatrixDecompose.cu:235-252)
            pow[(size_t)l * dpt + dg] = (T)host_powmod(2, (uint64_t)base_bits * dg, ctx->moduli[l]);
    Scratch tab;
    if (tab.alloc(pow.size() * sizeof(T), ctx->stream)) return 1;
    GP_CUDA(cudaMemcpyAsync(tab.p, pow.data(), pow.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream),
            "cudaMemcpyAsync(gadget table)");
    const size_t words = out->polys() * out->poly_words();
    const unsigned threads = 256, grid = stride_grid(ctx, words, threads);
    gp::gadget_fill_kernel<T><<<grid, threads, 0, gp::cur(ctx)>>>((T *)out->data, (const T *)tab.p, words, (uint32_t)out->rows, ell,
                                                               dpt, limbs, (uint32_t)ctx->logN, small ? 1 : 0);
    gp::count_launch();
    GP_CHECK_LAUNCH("gadget_fill_kernel");
    return 0;
}

int fill_gadget_common(GpuMatrix *out, uint32_t base_bits, bool small, const char *name) {
    if (!out || !out->ctx) return gp::set_error(std::string("invalid ") + name + " arguments");
    if (base_bits == 0 || base_bits >= 63) return gp::set_error(std::string("invalid base_bits in ") + name);  // :316
    if (out->level < 0) return gp::set_error(std::string("invalid level in ") + name);
    GpuContext *ctx = out->ctx;
    const size_t dpt = gadget_digits_per_tower(out, base_bits);
    const size_t ell = small ? dpt : dpt * out->limbs();
    if (out->cols != out->rows * ell) return gp::set_error(std::string("output size mismatch in ") + name);  // :363
    out->format = GPU_POLY_FORMAT_EVAL;  // :431-437
    if (out->polys() == 0) return 0;
    if (bind_device(ctx)) return 1;
    return GP_DISPATCH_WIDTH(ctx, return run_fill_gadget<T>(ctx, out, base_bits, small););
}

template <typename T, int LOGB, bool NOCSUB>
int launch_fast_decompose(GpuContext *ctx, const T *src, T *out, const gp::DecompGeom &g, unsigned grid) {
    static std::atomic<uint32_t> done{0};
    constexpr size_t smem = gp::fast_smem_bytes<T, LOGB>();
    auto kernel = gp::fast_decompose_kernel<T, LOGB, NOCSUB>;
    if (optin_smem(kernel, smem, ctx->device, done)) return 1;
    kernel<<<grid, gp::FastCfg<T, LOGB>::kThreads, smem, gp::cur(ctx)>>>(src, out, g, fwd_tab<T>(ctx), limb_tab<T>(ctx));
    gp::count_launch();
    GP_CHECK_LAUNCH("fast_decompose_kernel");
    return 0;
}
template <typename T> int dispatch_fast_decompose(GpuContext *ctx, const T *src, T *out, const gp::DecompGeom &g, unsigned grid, int logb) {
    if constexpr (gp::is_f64<T>::value) {
#define GP_CALL(LB) launch_fast_decompose<T, LB, false>(ctx, src, out, g, grid)
        GP_FAST_LOGB_SWITCH(logb, GP_CALL)
#undef GP_CALL
    } else {
#define GP_CALL(LB) (ctx->nocsub ? launch_fast_decompose<T, LB, true>(ctx, src, out, g, grid) : launch_fast_decompose<T, LB, false>(ctx, src, out, g, grid))
        GP_FAST_LOGB_SWITCH(logb, GP_CALL)
#undef GP_CALL
    }
}

// A coefficient-form source that is NOT an ordinary matrix of this context: the all-gathered per-rank blocks of a
// limb-sharded operand (gp::DecompGeom, ntt.cuh), carrying ALL towers of the full CRT basis
struct GatheredSrc {
    uint32_t towers = 0, tpg = 0, crt_bits = 0;
    size_t group_stride = 0, poly_stride = 0;
    unsigned char bits[64] = {};
};

// out (rows*ell x cols) <- digits of src[:, src_col0 : src_col0 + cols] (src in COEFF form)
template <typename T>
int run_decompose(GpuContext *ctx, const T *src_coeff, T *out, size_t rows, size_t src_cols, size_t src_col0, size_t cols,
                  uint32_t limbs, uint32_t base_bits, bool small, bool to_eval, const GatheredSrc *gs = nullptr) {
    const uint32_t dpt = gs ? (gs->crt_bits + base_bits - 1) / base_bits : digits_per_tower(ctx, base_bits);
    gp::DecompGeom g{};
    g.limbs = limbs;
    g.src_cols = (uint32_t)src_cols;
    g.src_col0 = (uint32_t)src_col0;
    g.cols = (uint32_t)cols;
    g.dpt = dpt;
    g.ell = small ? dpt : dpt * (gs ? gs->towers : limbs);
    g.base_bits = base_bits;
    if (gs) {
        g.gathered = 1;
        g.src_tpg = gs->tpg;
        g.src_group_stride = gs->group_stride;
        g.src_poly_stride = gs->poly_stride;
        for (uint32_t t = 0; t < gs->towers && t < 64; t++) gp::set_tower_bits(g, t, gs->bits[t]);
    }
    const size_t opl = rows * g.ell * cols * limbs;
    if (opl == 0) return 0;
    if (const int lc = ntt_cluster_log<T>(ctx); to_eval && lc > 0) {  // rings of 2..8 blocks: digits extracted by the cluster's push phase
        return gpm::run_cluster_decompose<T>(ctx, src_coeff, out, g, opl, lc);
    }
    const int logn = ctx->logN, logb = ntt_block_log<T>(ctx), top = logn - logb;
    if (to_eval && top == 0 && use_fast<T>(ctx, logb)) {
        if constexpr (sizeof(T) == 8) {
            if (ctx->f64_ntt) return dispatch_fast_decompose<gp::F64>(ctx, (const gp::F64 *)src_coeff, (gp::F64 *)out, g, (unsigned)opl, logb);
        }
        return dispatch_fast_decompose<T>(ctx, src_coeff, out, g, (unsigned)opl, logb);
    }
    const size_t smem = gp::ntt_smem_bytes<T>(logb);
    {  // > 48 KB dynamic shared memory for this translation unit's runtime-planned decomposition kernels
        static std::atomic<uint32_t> done_e{0}, done_c{0};
        if ((int)smem > ctx->max_smem_optin) return gp::set_error("NTT block does not fit shared memory on this device");
        if (smem > 48 * 1024 && (optin_smem(gp::decompose_kernel<T, true>, (size_t)ctx->max_smem_optin, ctx->device, done_e) ||
                                 optin_smem(gp::decompose_kernel<T, false>, (size_t)ctx->max_smem_optin, ctx->device, done_c)))
            return 1;
    }
    const auto *tw = (const gp::TwPair<T> *)ctx->d_fwd;
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    const unsigned grid = (unsigned)(opl << top), threads = gp::NttCfg<T>::kThreads;
    if (to_eval && top == 0) {
        gp::decompose_kernel<T, true><<<grid, threads, smem, gp::cur(ctx)>>>(src_coeff, out, g, (uint32_t)logn, (uint32_t)logb, tw, lc);
        gp::count_launch();
        GP_CHECK_LAUNCH("decompose_kernel");
        return 0;
    }
    if (to_eval && ctx->fuse_decompose_top) {
        // rings above one CTA: the digits are extracted inside the first global pass of their forward transform
        int passes[8], np = 0;
        gp::split_top(top, passes, &np);
        const unsigned gthreads = 256;
        switch (passes[0]) {
#define GP_DFG(R)                                                                                                    \
    case R: {                                                                                                        \
        const unsigned ggrid = (unsigned)(opl << ((logn - R) - gp::decompose_top_chunk_log<R>((uint32_t)logn)));     \
        gp::decompose_fwd_global_pass<T, R><<<ggrid, gthreads, 0, gp::cur(ctx)>>>(src_coeff, out, g, (uint32_t)logn, tw, lc); \
    } break;
            GP_DFG(1)
            GP_DFG(2)
            GP_DFG(3)
            default:
                GP_DFG(4)
#undef GP_DFG
        }
        gp::count_launch();
        GP_CHECK_LAUNCH("decompose_fwd_global_pass");
        return run_ntt_fwd<T>(ctx, out, out, rows * g.ell * cols, limbs, true);
    }
    gp::decompose_kernel<T, false><<<grid, threads, 0, gp::cur(ctx)>>>(src_coeff, out, g, (uint32_t)logn, (uint32_t)logb, tw, lc);
    gp::count_launch();
    GP_CHECK_LAUNCH("decompose_kernel");
    if (to_eval) return run_ntt_fwd<T>(ctx, out, out, rows * g.ell * cols, limbs);
    return 0;
}

int decompose_common(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out, bool small) {
    if (!src || !out) return gp::set_error("invalid gpu_matrix_decompose_base arguments");
    if (base_bits == 0) return gp::set_error("base_bits must be non-zero in gpu_matrix_decompose_base");
    if (base_bits >= 63) return gp::set_error("base_bits too large in gpu_matrix_decompose_base");
    if (!same_ctx(src, out)) return gp::set_error("context mismatch in gpu_matrix_decompose_base");
    if (src->level < 0) return gp::set_error("invalid level in gpu_matrix_decompose_base");
    if (out == src || (out->data && out->data == src->data))
        return gp::set_error("gpu_matrix_decompose_base output must not alias the input");
    GpuContext *ctx = src->ctx;
    const int want = out->format;  // the caller creates `out` in the format it wants back (MatrixDecompose.cu:910-914)
    if (want != GPU_POLY_FORMAT_COEFF && want != GPU_POLY_FORMAT_EVAL)
        return gp::set_error("invalid output format in gpu_matrix_decompose_base");
    const size_t dpt = digits_per_tower(ctx, base_bits);
    const size_t ell = small ? dpt : dpt * src->limbs();
    if (out->rows != src->rows * ell || out->cols != src->cols)
        return gp::set_error("output size mismatch in gpu_matrix_decompose_base");  // :937-946
    if (src->polys() == 0) {
        out->format = GPU_POLY_FORMAT_EVAL;
        return 0;
    }
    if (bind_device(ctx)) return 1;
    const uint32_t limbs = (uint32_t)src->limbs();
    int st = GP_DISPATCH_WIDTH(
        ctx, const T *coeff = (const T *)src->data; Scratch tmp;
        if (src->format == GPU_POLY_FORMAT_EVAL) {  // coefficient form of the input, without touching it (:965-988)
            if (tmp.alloc(src->bytes(), ctx->stream)) return 1;
            if (run_ntt_inv<T>(ctx, (const T *)src->data, (T *)tmp.p, src->polys(), limbs)) return 1;
            coeff = (const T *)tmp.p;
        } return run_decompose<T>(ctx, coeff, (T *)out->data, src->rows, src->cols, 0, src->cols, limbs, base_bits, small,
                                  want == GPU_POLY_FORMAT_EVAL););
    if (!st) out->format = want;
    return st;
}
}  // namespace

extern "C" int gpu_matrix_fill_gadget(GpuMatrix *out, uint32_t base_bits) {
    return fill_gadget_common(out, base_bits, false, "gpu_matrix_fill_gadget");
}
extern "C" int gpu_matrix_fill_small_gadget(GpuMatrix *out, uint32_t base_bits) {
    return fill_gadget_common(out, base_bits, true, "gpu_matrix_fill_small_gadget");
}
extern "C" int gpu_matrix_decompose_base(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out) {
    return decompose_common(src, base_bits, out, false);
}
extern "C" int gpu_matrix_decompose_base_small(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out) {
    return decompose_common(src, base_bits, out, true);
}

extern "C" int gpu_matrix_fill_small_decomposed_identity_chunk(GpuMatrix *out, const GpuMatrix *scalar_by_digit, size_t chunk_idx) {
    const char *fn = "gpu_matrix_fill_small_decomposed_identity_chunk";
    if (!out || !scalar_by_digit) return gp::set_error(std::string("invalid ") + fn + " arguments");
    if (!same_ctx(out, scalar_by_digit)) return gp::set_error(std::string("context mismatch in ") + fn);
    if (out->rows != out->cols) return gp::set_error(std::string("output must be square in ") + fn);
    if (scalar_by_digit->rows != 1 || scalar_by_digit->cols == 0)
        return gp::set_error(std::string("scalar_by_digit must be 1 x chunk_count in ") + fn);
    if (out->format != scalar_by_digit->format) return gp::set_error(std::string("format mismatch in ") + fn);
    const size_t size = out->rows, digits = scalar_by_digit->cols;
    if (chunk_idx >= digits) return gp::set_error(std::string("chunk_idx out of range in ") + fn);
    if (size == 0) return 0;
    if (out->level < 0) return gp::set_error(std::string("invalid level in ") + fn);
    GpuContext *ctx = out->ctx;
    if (bind_device(ctx)) return 1;
    const size_t pw = out->poly_words();
    return GP_DISPATCH_WIDTH(
        ctx, const int V = vec_width<T>(ctx); const size_t total = size * size * pw / V;
        const unsigned threads = 256, grid = stride_grid(ctx, total, threads);
        T *d = (T *)out->data; const T *s = (const T *)scalar_by_digit->data;
        switch (V) {
            case 4: gp::small_identity_chunk_kernel<T, 4><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, size, chunk_idx, digits, total, pw); break;
            case 2: gp::small_identity_chunk_kernel<T, 2><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, size, chunk_idx, digits, total, pw); break;
            default: gp::small_identity_chunk_kernel<T, 1><<<grid, threads, 0, gp::cur(ctx)>>>(d, s, size, chunk_idx, digits, total, pw); break;
        } gp::count_launch();
        GP_CHECK_LAUNCH("small_identity_chunk_kernel"); return 0;);
}

// ---- extension: S * G^-1(other) in one call ------------------------------------------
// Replaces the slice -> decompose_owned -> mul -> copy_block loop of
// GpuDCRTPolyMatrix::mul_decompose (src/matrix/gpu_dcrt_poly.rs:1414-1493; ~5 matrix
// creations and log2(n)+5 launches per column there).  Here: one INTT of `other`, then per
// column chunk one fused digit+NTT kernel and one product kernel writing straight into the
// output columns (MXX_MUL_DECOMPOSE_SCRATCH_BYTES bounds the digit scratch; the reference's
// knob is MXX_MUL_DECOMPOSE_COLUMN_CHUNK_WIDTH, src/env.rs:148).
namespace {
constexpr size_t kOneChunkMinCtas = 8192;  // digit-NTT CTAs (64 KB blocks) from which the product runs as ONE chunk
// out = lhs * G^-1(source columns [col0, col0 + out->cols)): per column chunk one fused digit+NTT kernel and one product
// kernel writing straight into the output columns.  `coeff` is the coefficient-form source: an ordinary d x c matrix of
// this context, or (gs != nullptr) the gathered blocks of a limb-sharded operand.
template <typename T>
int mul_decompose_core(GpuContext *ctx, GpuMatrix *out, const GpuMatrix *lhs, const T *coeff, size_t d, size_t c, size_t col0,
                       size_t m, uint32_t base_bits, bool small, const GatheredSrc *gs) {
    // digit polynomials in flight: two lanes of at most this many bytes each (they stream through L2/HBM while
    // the issue-bound digit-NTT kernel runs, so the traffic is hidden; the cap only bounds scratch memory)
    size_t budget = 2ull << 30;
    if (const char *env = getenv("MXX_MUL_DECOMPOSE_SCRATCH_BYTES")) {
        char *end = nullptr;
        unsigned long long v = strtoull(env, &end, 10);
        if (end && end != env && v) budget = v;
    }
    const uint32_t limbs = (uint32_t)out->limbs();
    const size_t oc = out->cols;
    // Column chunks are independent.  Chunk k runs on auxiliary stream k % 2 with its own
    // digit buffer, so the partial last wave of one chunk's digit-NTT kernel is filled by the next
    // chunk's CTAs and the HBM/L2-bound product kernel overlaps the issue-bound NTT kernel.
    const size_t poly_bytes = out->poly_words() * sizeof(T);
    // Wide chunks: a digit-NTT launch of w columns has w*m*limbs CTAs, and only a launch of many waves keeps
    // the last, partial wave cheap (C3-a: 450 CTAs per column on 444 resident slots).  Two chunks run
    // concurrently on the two auxiliary streams, so one chunk's product overlaps the other's NTTs.
    size_t width = std::max<size_t>(1, (budget / 2) / std::max<size_t>(1, m * poly_bytes));
    // One chunk or two lanes?  Measured on B200 (tools/md_lanes_ab.py, profiles/r02_md_lanes_ab.txt): launches of many waves are
    // 6-8 % faster as ONE chunk on the context stream (C3-a 0.90 -> 0.83 ms, C3-b 20.3 -> 19.1 ms: no product kernel competing
    // with the digit-NTT kernel, fewer partial waves), small ones 2-7 % faster as two chunks on the auxiliary streams (the second
    // chunk's CTAs fill the first one's tail; a limb-sharded C4 gate on 8 GPUs is 2300 CTAs).  MXX_MUL_DECOMPOSE_LANES=1|2 forces either.
    static const int lanes_knob = getenv("MXX_MUL_DECOMPOSE_LANES") ? atoi(getenv("MXX_MUL_DECOMPOSE_LANES")) : 0;
    const size_t ctas_all = ((m * oc * limbs) << ctx->logN) >> (sizeof(T) == 4 ? 14 : 13);
    const bool one_lane = lanes_knob == 1 || (lanes_knob == 0 && ctas_all >= kOneChunkMinCtas);
    width = std::min(width * (one_lane ? 2 : 1), one_lane ? oc : (oc >= 2 ? (oc + 1) / 2 : oc));
    const size_t chunks = (oc + width - 1) / width;
    const int lanes = chunks > 1 ? 2 : 1;  // a job too large for one chunk of the scratch budget alternates wide chunks over the two lanes
    Scratch digits[2];
    for (int i = 0; i < lanes; i++)
        if (digits[i].alloc(m * width * poly_bytes, ctx->stream)) return 1;
    if (lanes > 1) {
        GP_CUDA(cudaEventRecord(ctx->ev_fork, ctx->stream), "cudaEventRecord(fork)");
        for (int i = 0; i < lanes; i++) GP_CUDA(cudaStreamWaitEvent(ctx->aux[i], ctx->ev_fork, 0), "cudaStreamWaitEvent(fork)");
    }
    int st = 0;
    for (size_t k = 0; k < chunks && !st; k++) {
        const size_t c0 = k * width, w = std::min(width, oc - c0);
        gp::StreamScope scope(lanes > 1 ? ctx->aux[k & 1] : ctx->stream);
        T *dg = (T *)digits[lanes > 1 ? (k & 1) : 0].p;
        st = run_decompose<T>(ctx, coeff, dg, d, c, col0 + c0, w, limbs, base_bits, small, true, gs);
        if (st) break;
        gp::MatmulGeom g{};
        g.rows = (uint32_t)lhs->rows; g.inner = (uint32_t)m; g.cols = (uint32_t)w; g.b_cols = (uint32_t)w; g.b_col0 = 0;
        g.out_cols = (uint32_t)oc; g.out_col0 = (uint32_t)c0; g.limbs = limbs;
        st = run_matmul<T>(ctx, (T *)out->data, (const T *)lhs->data, dg, g);
    }
    if (lanes > 1)  // join even after an error so the scratch buffers are not freed under running kernels
        for (int i = 0; i < lanes; i++) {
            cudaEventRecord(ctx->ev_join[i], ctx->aux[i]);
            cudaStreamWaitEvent(ctx->stream, ctx->ev_join[i], 0);
        }
    return st;
}
}  // namespace

extern "C" int gpu_matrix_mul_decompose(GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs, uint32_t base_bits, int small) {
    if (binary_check("gpu_matrix_mul_decompose", out, lhs, rhs)) return 1;
    if (base_bits == 0 || base_bits >= 63) return gp::set_error("invalid base_bits in gpu_matrix_mul_decompose");
    if (out->level < 0) return gp::set_error("invalid level in gpu_matrix_mul_decompose");
    GpuContext *ctx = out->ctx;
    const size_t dpt = digits_per_tower(ctx, base_bits);
    const size_t ell = small ? dpt : dpt * out->limbs();
    const size_t d = rhs->rows, c = rhs->cols, m = d * ell;
    if (lhs->cols != m || out->rows != lhs->rows || out->cols != c)
        return gp::set_error("size mismatch in gpu_matrix_mul_decompose");
    if (lhs->format != GPU_POLY_FORMAT_EVAL) return gp::set_error("gpu_matrix_mul_decompose requires an EVAL left operand");
    if (out == lhs || out == rhs || (out->data && (out->data == lhs->data || out->data == rhs->data)))
        return gp::set_error("gpu_matrix_mul_decompose output must not alias an input");
    out->format = GPU_POLY_FORMAT_EVAL;
    if (out->polys() == 0) return 0;
    if (bind_device(ctx)) return 1;
    const uint32_t limbs = (uint32_t)out->limbs();
    return GP_DISPATCH_WIDTH(
        ctx, const T *coeff = (const T *)rhs->data; Scratch tmp;
        if (rhs->format == GPU_POLY_FORMAT_EVAL) {
            if (tmp.alloc(rhs->bytes(), ctx->stream)) return 1;
            if (run_ntt_inv<T>(ctx, (const T *)rhs->data, (T *)tmp.p, rhs->polys(), limbs)) return 1;
            coeff = (const T *)tmp.p;
        } return mul_decompose_core<T>(ctx, out, lhs, coeff, d, c, 0, m, base_bits, small != 0, nullptr););
}

// ---- extension: the same for a LIMB-SHARDED job (SURVEY 8e "Alternative (limbs)", mxx_b200/dist.py) -----------------
// `out` and `lhs` live in this rank's context, whose moduli are a subset of the job's CRT basis; `other_coeff` is the
// coefficient-form right operand (d x c polynomials) with ALL `towers` towers of the basis, as the all-gather of the ranks'
// [poly][towers_per_group][n] blocks left it: tower t of polynomial p at
//     (t / towers_per_group) * group_stride_words + p * poly_stride_words + (t % towers_per_group) * n   words
// (an ordinary full-basis matrix is towers_per_group = towers, poly_stride_words = towers * n).  Words have the width of
// this context's storage.  out (lhs->rows x cols) = lhs * G^-1(other[:, col0 : col0 + cols]) restricted to this rank's
// limbs; digit rows follow the FULL basis (ell = digits_per_tower(max tower bits) * towers, or digits_per_tower if small).
extern "C" int gpu_matrix_mul_decompose_gathered(GpuMatrix *out, const GpuMatrix *lhs, const void *other_coeff, size_t d, size_t c,
                                                 size_t col0, const uint64_t *tower_moduli, size_t towers, size_t towers_per_group,
                                                 size_t group_stride_words, size_t poly_stride_words, uint32_t base_bits, int small) {
    const char *fn = "gpu_matrix_mul_decompose_gathered";
    if (!out || !lhs || !tower_moduli || (!other_coeff && d * c)) return gp::set_error(std::string("invalid ") + fn + " arguments");
    if (!same_ctx(out, lhs)) return gp::set_error(std::string("context mismatch in ") + fn);
    if (base_bits == 0 || base_bits >= 63) return gp::set_error(std::string("invalid base_bits in ") + fn);
    if (out->level < 0) return gp::set_error(std::string("invalid level in ") + fn);
    if (towers == 0 || towers > 64 || towers_per_group == 0 || towers % towers_per_group)
        return gp::set_error(std::string("invalid tower grouping in ") + fn);
    GpuContext *ctx = out->ctx;
    GatheredSrc gs;
    gs.towers = (uint32_t)towers, gs.tpg = (uint32_t)towers_per_group;
    gs.group_stride = group_stride_words, gs.poly_stride = poly_stride_words;
    for (size_t t = 0; t < towers; t++) {
        if (tower_moduli[t] < 2) return gp::set_error(std::string("invalid tower modulus in ") + fn);
        gs.bits[t] = (unsigned char)(64 - __builtin_clzll(tower_moduli[t]));
        gs.crt_bits = std::max<uint32_t>(gs.crt_bits, gs.bits[t]);
        if (!ctx->wide && gs.bits[t] > 30) return gp::set_error(std::string("tower does not fit this context's 32-bit words in ") + fn);
    }
    const size_t dpt = (gs.crt_bits + base_bits - 1) / base_bits;
    const size_t ell = small ? dpt : dpt * towers, m = d * ell;
    if (lhs->cols != m || out->rows != lhs->rows || col0 + out->cols > c) return gp::set_error(std::string("size mismatch in ") + fn);
    if (lhs->format != GPU_POLY_FORMAT_EVAL) return gp::set_error(std::string(fn) + " requires an EVAL left operand");
    if (out == lhs || (out->data && out->data == lhs->data)) return gp::set_error(std::string(fn) + " output must not alias an input");
    out->format = GPU_POLY_FORMAT_EVAL;
    if (out->polys() == 0) return 0;
    if (bind_device(ctx)) return 1;
    return GP_DISPATCH_WIDTH(ctx, return mul_decompose_core<T>(ctx, out, lhs, (const T *)other_coeff, d, c, col0, m, base_bits, small != 0, &gs););
}
