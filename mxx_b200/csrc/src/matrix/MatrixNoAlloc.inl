// MatrixNoAlloc.inl — gpu_matrix_decompose_base_noalloc: gadget decomposition for callers that must not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Part of the MatrixDecompose.cu translation unit.
//
// gpu_matrix_decompose_base needs scratch when its input is in EVAL form: the coefficient form of the input
// (src->bytes()), taken from the stream-ordered pool (decompose_common, MatrixDecompose.inl).  Here that scratch is a
// file-scope device array of GPU_NOALLOC_SCRATCH_BYTES (16 MiB), which every GPU that loads the library holds until
// exit.  The entry point calls no cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine; inputs
// larger than the array are rejected.  Every call on a device shares the array: two calls whose kernels overlap on the
// GPU (different contexts / streams on one device) race on it; calls on one context are stream-ordered and safe.
// Nothing has to start at zero: the inverse transform writes every word of the region before the digit kernel reads it.

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[GPU_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address kernels may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
int noalloc_scratch(const GpuContext *ctx, void **out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static void *ptrs[kMaxDevices] = {};
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return gp::set_error("device index out of range in gpu_matrix_decompose_base_noalloc");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[ctx->device]) {
        void *p = nullptr;  // bind_device(ctx) has made ctx->device current
        GP_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(noalloc scratch)");
        if (!p) return gp::set_error("cudaGetSymbolAddress returned a null scratch address");
        ptrs[ctx->device] = p;
    }
    *out = ptrs[ctx->device];
    return 0;
}

}  // namespace

extern "C" int gpu_matrix_decompose_base_noalloc(const GpuMatrix *src, uint32_t base_bits, GpuMatrix *out) {
    const char *fn = "gpu_matrix_decompose_base_noalloc";
    // validation: that of gpu_matrix_decompose_base (decompose_common)
    if (!src || !out) return gp::set_error(std::string("invalid ") + fn + " arguments");
    if (base_bits == 0) return gp::set_error(std::string("base_bits must be non-zero in ") + fn);
    if (base_bits >= 63) return gp::set_error(std::string("base_bits too large in ") + fn);
    if (!same_ctx(src, out)) return gp::set_error(std::string("context mismatch in ") + fn);
    if (src->level < 0) return gp::set_error(std::string("invalid level in ") + fn);
    if (out == src || (out->data && out->data == src->data)) return gp::set_error(std::string(fn) + " output must not alias the input");
    GpuContext *ctx = src->ctx;
    const int want = out->format;
    if (want != GPU_POLY_FORMAT_COEFF && want != GPU_POLY_FORMAT_EVAL) return gp::set_error(std::string("invalid output format in ") + fn);
    const size_t ell = (size_t)digits_per_tower(ctx, base_bits) * src->limbs();
    if (out->rows != src->rows * ell || out->cols != src->cols) return gp::set_error(std::string("output size mismatch in ") + fn);
    if (src->polys() == 0) {
        out->format = GPU_POLY_FORMAT_EVAL;
        return 0;
    }
    const bool need_scratch = src->format == GPU_POLY_FORMAT_EVAL;
    if (need_scratch && src->bytes() > (size_t)GPU_NOALLOC_SCRATCH_BYTES)
        return gp::set_error(std::string(fn) + ": an EVAL input of " + std::to_string(src->bytes()) + " bytes exceeds the static scratch of " +
                             std::to_string((size_t)GPU_NOALLOC_SCRATCH_BYTES) + " bytes (GPU_NOALLOC_SCRATCH_BYTES); use gpu_matrix_decompose_base");
    if (bind_device(ctx)) return 1;
    void *scratch = nullptr;
    if (need_scratch && noalloc_scratch(ctx, &scratch)) return 1;
    const uint32_t limbs = (uint32_t)src->limbs();
    int st = GP_DISPATCH_WIDTH(
        ctx, const T *coeff = (const T *)src->data;
        if (need_scratch) {  // coefficient form of the input in the static array, without touching the input
            if (run_ntt_inv<T>(ctx, (const T *)src->data, (T *)scratch, src->polys(), limbs)) return 1;
            coeff = (const T *)scratch;
        } return run_decompose<T>(ctx, coeff, (T *)out->data, src->rows, src->cols, 0, src->cols, limbs, base_bits, false,
                                  want == GPU_POLY_FORMAT_EVAL););
    if (!st) out->format = want;
    return st;
}
