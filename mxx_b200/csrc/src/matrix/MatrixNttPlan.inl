// MatrixNttPlan.inl — planning helpers shared by the translation units that launch NTT-family kernels
// (MatrixNTT.cu, MatrixNTTCluster.cu, MatrixDecompose.cu, MatrixDecomposeCluster.cu): block / cluster size per ring,
// device tables by working type, the opt-in to > 48 KB of dynamic shared memory, the LOGB / cluster-size switches.

namespace {

// block (per-CTA) size: the compile-time-planned kernels of ntt_fast.cuh when one exists for this
// ring, else the generic kernel of ntt.cuh
template <typename T> inline int ntt_block_log(const GpuContext *ctx) {
    const int fb = gp::fast_block_log<T>(ctx->logN);
    if (!ctx->force_generic_ntt && ctx->ntt_block_override >= 12 && ctx->ntt_block_override < ctx->logN &&
        gp::fast_supported<T>(ctx->ntt_block_override))
        return ctx->ntt_block_override;  // MXX_NTT_BLOCK_LOG: tuning knob for rings above one CTA's shared memory
    if (!ctx->force_generic_ntt && gp::fast_supported<T>(fb)) return fb;
    return gp::ntt_block_log<T>(ctx->logN);
}
template <typename T> inline bool use_fast(const GpuContext *ctx, int logb) {
    return !ctx->force_generic_ntt && gp::fast_supported<T>(logb) && logb == ntt_block_log<T>(ctx);
}
using gp::split_top;

// device tables by working type: the F64 kernels have their own (Runtime.cu upload_tables)
template <typename T> inline const gp::TwPair<T> *fwd_tab(const GpuContext *c) {
    return (const gp::TwPair<T> *)(gp::is_f64<T>::value ? c->d_fwd_f64 : c->d_fwd);
}
template <typename T> inline const gp::TwPair<T> *inv_tab(const GpuContext *c) {
    return (const gp::TwPair<T> *)(gp::is_f64<T>::value ? c->d_inv_f64 : c->d_inv);
}
template <typename T> inline const gp::LimbConst<T> *limb_tab(const GpuContext *c) {
    return (const gp::LimbConst<T> *)(gp::is_f64<T>::value ? c->d_limb_f64 : c->d_limb);
}

// opt-in to > 48 KB of dynamic shared memory, once per kernel instantiation and device
template <typename K> int optin_smem(K kernel, size_t bytes, int device, std::atomic<uint32_t> &done) {
    if (bytes <= 48 * 1024) return 0;
    const uint32_t bit = 1u << (device & 31);
    if (done.load(std::memory_order_acquire) & bit) return 0;
    GP_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes), "cudaFuncSetAttribute");
    done.fetch_or(bit, std::memory_order_release);
    return 0;
}

#define GP_FAST_LOGB_SWITCH(logb, CALL)                                   \
    switch (logb) {                                                       \
        case 12: return CALL(12);                                         \
        case 13: return CALL(13);                                         \
        case 14: return CALL(14);                                         \
        case 15:                                                          \
            if constexpr (sizeof(T) == 4) return CALL(15);                \
            [[fallthrough]];                                              \
        default: return gp::set_error("no compile-time NTT plan for this block size"); \
    }

// ---- cluster / DSMEM kernels (ntt_cluster.cuh): rings of 2, 4 or 8 shared-memory blocks in ONE pass over HBM ----
// log2(cluster size) for this context's ring, 0 = not applicable (ring fits one block, knobs, ring too large)
template <typename T> inline int ntt_cluster_log(const GpuContext *ctx) {
    if (!ctx->ntt_cluster || ctx->force_generic_ntt || ctx->ntt_block_override) return 0;
    return gp::cluster_log<T>(ctx->logN);
}
#define GP_CLUSTER_SWITCH(lc, CALL)                                                     \
    switch (lc) {                                                                       \
        case 1: return CALL(1);                                                         \
        case 2: return CALL(2);                                                         \
        case 3: return CALL(3);                                                         \
        default: return gp::set_error("no cluster NTT plan for this ring size");        \
    }

}  // namespace
