// MatrixNTTCluster.inl — host side of the cluster / DSMEM transforms (ntt_cluster.cuh): rings of 2, 4 or 8
// shared-memory blocks in ONE pass over HBM.  Own translation unit (MatrixNTTCluster.cu) so that the kernel families
// compile in parallel.

namespace {
template <typename K, int LOGB, int LC, bool NOCSUB>
int launch_cluster_fwd(GpuContext *ctx, const K *src, K *dst, size_t pl, uint32_t limbs) {
    static std::atomic<uint32_t> done{0};
    constexpr size_t smem = gp::fast_smem_bytes<K, LOGB>();
    auto kernel = gp::fast_ntt_fwd_cluster_kernel<K, LOGB, LC, NOCSUB>;
    if (optin_smem(kernel, smem, ctx->device, done)) return 1;
    kernel<<<(unsigned)(pl << LC), gp::FastCfg<K, LOGB>::kThreads, smem, gp::cur(ctx)>>>(src, dst, limbs, fwd_tab<K>(ctx), limb_tab<K>(ctx));
    gp::count_launch();
    GP_CHECK_LAUNCH("fast_ntt_fwd_cluster_kernel");
    return 0;
}
template <typename K, int LOGB, int LC>
int launch_cluster_inv(GpuContext *ctx, const K *src, K *dst, size_t pl, uint32_t limbs) {
    static std::atomic<uint32_t> done{0};
    constexpr size_t smem = gp::fast_smem_bytes<K, LOGB>();
    auto kernel = gp::fast_ntt_inv_cluster_kernel<K, LOGB, LC>;
    if (optin_smem(kernel, smem, ctx->device, done)) return 1;
    kernel<<<(unsigned)(pl << LC), gp::FastCfg<K, LOGB>::kThreads, smem, gp::cur(ctx)>>>(src, dst, limbs, inv_tab<K>(ctx), limb_tab<K>(ctx));
    gp::count_launch();
    GP_CHECK_LAUNCH("fast_ntt_inv_cluster_kernel");
    return 0;
}
// K = the kernels' working type: T, or gp::F64 over u64 words
template <typename K> int dispatch_cluster_fwd(GpuContext *ctx, const K *src, K *dst, size_t pl, uint32_t limbs, int lc) {
    constexpr int LB = sizeof(K) == 4 ? 14 : 13;
    if constexpr (gp::is_f64<K>::value) {
#define GP_CALL(LC) launch_cluster_fwd<K, LB, LC, false>(ctx, src, dst, pl, limbs)
        GP_CLUSTER_SWITCH(lc, GP_CALL)
#undef GP_CALL
    } else {
#define GP_CALL(LC) (ctx->nocsub ? launch_cluster_fwd<K, LB, LC, true>(ctx, src, dst, pl, limbs) : launch_cluster_fwd<K, LB, LC, false>(ctx, src, dst, pl, limbs))
        GP_CLUSTER_SWITCH(lc, GP_CALL)
#undef GP_CALL
    }
}
template <typename K> int dispatch_cluster_inv(GpuContext *ctx, const K *src, K *dst, size_t pl, uint32_t limbs, int lc) {
    constexpr int LB = sizeof(K) == 4 ? 14 : 13;
#define GP_CALL(LC) launch_cluster_inv<K, LB, LC>(ctx, src, dst, pl, limbs)
    GP_CLUSTER_SWITCH(lc, GP_CALL)
#undef GP_CALL
}

}  // namespace

// T = storage word; u64 contexts with moduli < 2^51 run the FP64-pipe kernels
template <typename T> int gpm::run_cluster_fwd(GpuContext *ctx, const T *src, T *dst, size_t pl, uint32_t limbs, int lc) {
    if constexpr (sizeof(T) == 8) {
        if (ctx->f64_ntt) return dispatch_cluster_fwd<gp::F64>(ctx, (const gp::F64 *)src, (gp::F64 *)dst, pl, limbs, lc);
    }
    return dispatch_cluster_fwd<T>(ctx, src, dst, pl, limbs, lc);
}
template <typename T> int gpm::run_cluster_inv(GpuContext *ctx, const T *src, T *dst, size_t pl, uint32_t limbs, int lc) {
    if constexpr (sizeof(T) == 8) {
        if (ctx->f64_ntt) return dispatch_cluster_inv<gp::F64>(ctx, (const gp::F64 *)src, (gp::F64 *)dst, pl, limbs, lc);
    }
    return dispatch_cluster_inv<T>(ctx, src, dst, pl, limbs, lc);
}
template int gpm::run_cluster_fwd<uint32_t>(GpuContext *, const uint32_t *, uint32_t *, size_t, uint32_t, int);
template int gpm::run_cluster_fwd<uint64_t>(GpuContext *, const uint64_t *, uint64_t *, size_t, uint32_t, int);
template int gpm::run_cluster_inv<uint32_t>(GpuContext *, const uint32_t *, uint32_t *, size_t, uint32_t, int);
template int gpm::run_cluster_inv<uint64_t>(GpuContext *, const uint64_t *, uint64_t *, size_t, uint32_t, int);
