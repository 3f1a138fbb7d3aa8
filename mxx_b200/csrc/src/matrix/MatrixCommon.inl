// MatrixCommon.inl — shared host helpers: width dispatch, launch sizing, stream-ordered
// scratch.  Replaces the reference's per-op pointer/stride table uploads and per-limb
// stream/event bookkeeping (cuda/src/matrix/MatrixUtils.cu:259-405): with one typed,
// contiguous buffer per matrix and one stream per context none of that is needed.

namespace {

#define GP_CHECK_LAUNCH(where)                                          \
    do {                                                                \
        cudaError_t _e = cudaGetLastError();                            \
        if (_e != cudaSuccess) return gp::set_cuda_error(_e, where);    \
    } while (0)

// Calls fn(T{}) with T = uint64_t or uint32_t according to the context's storage width.
#define GP_DISPATCH_WIDTH(ctx, ...)                         \
    ((ctx)->wide ? ([&]() -> int { using T = uint64_t; __VA_ARGS__ })() \
                 : ([&]() -> int { using T = uint32_t; __VA_ARGS__ })())

inline int bind_device(const GpuContext *ctx) {
    GP_CUDA(cudaSetDevice(ctx->device), "cudaSetDevice");
    return 0;
}

// grid for a grid-stride kernel: enough CTAs to fill the chip a few times over, a whole
// number of waves (148 SMs on B200; sm_count read from the device).
inline unsigned stride_grid(const GpuContext *ctx, size_t items, unsigned threads, unsigned ctas_per_sm = 8) {
    size_t need = (items + threads - 1) / threads;
    size_t cap = (size_t)std::max(ctx->sm_count, 1) * ctas_per_sm;
    return (unsigned)std::max<size_t>(1, std::min(need, cap));
}

// stream-ordered scratch buffer; freed (stream-ordered) by the destructor
struct Scratch {
    void *p = nullptr;
    cudaStream_t s = nullptr;
    int alloc(size_t bytes, cudaStream_t stream) {
        s = stream;
        if (bytes == 0) return 0;
        GP_CUDA(cudaMallocAsync(&p, bytes, stream), "cudaMallocAsync(scratch)");
        return 0;
    }
    ~Scratch() {
        if (p) cudaFreeAsync(p, s);
    }
};

inline bool same_shape(const GpuMatrix *a, const GpuMatrix *b) { return a->rows == b->rows && a->cols == b->cols; }
inline bool same_ctx(const GpuMatrix *a, const GpuMatrix *b) { return a->ctx == b->ctx && a->level == b->level; }

// largest vector width (in words, up to 16 bytes) that divides n
template <typename T> inline int vec_width(const GpuContext *ctx) {
    int v = 16 / (int)sizeof(T);
    while (v > 1 && (ctx->N % v) != 0) v >>= 1;
    return v;
}

inline uint32_t digits_per_tower(const GpuContext *ctx, uint32_t base_bits) {
    return (ctx->crt_bits + base_bits - 1) / base_bits;  // MatrixDecompose.cu:924-935
}

}  // namespace
